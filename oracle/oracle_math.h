// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product; nothing under
// ray_tracer_b200/ may include, link or call this.
//
// Minimal float vector/matrix algebra with the *evaluation order* of GLM 0.9.9.7
// (the version vendored by the reference under third_party/glm). The reference's
// host code uses GLM directly (src/vk_engine.h:16-19) and SURVEY.md §8c2 pins the
// shader's built-ins (reflect/refract/mix/smoothstep/inverse/normalize/dot/cross)
// to the same GLM semantics. Every function cites the GLM file:line whose
// operation order it follows; tests/test_oracle_math.py checks them bit-for-bit
// against vectors produced by the real GLM headers (oracle/glm_kat_gen.cpp).
//
// Build with -ffp-contract=off: each operator below must be one IEEE-754 single op.
#pragma once
#include <cmath>
#include <cstdint>

namespace om {

struct vec2 { float x, y; };

struct vec3 {
    float x, y, z;
    vec3() : x(0.f), y(0.f), z(0.f) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

struct vec4 {
    float x, y, z, w;
    vec4() : x(0.f), y(0.f), z(0.f), w(0.f) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(const vec3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
    vec3 xyz() const { return vec3(x, y, z); }
};

// glm/detail/type_vec3.inl — component-wise operators
inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }  // type_vec3.inl:715
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3& operator+=(vec3& a, const vec3& b) { a = a + b; return a; }
inline vec3& operator*=(vec3& a, const vec3& b) { a = a * b; return a; }
inline vec3& operator*=(vec3& a, float s) { a = a * s; return a; }
inline vec3& operator/=(vec3& a, float s) { a = a / s; return a; }

inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(const vec4& a, const vec4& b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator*(const vec4& a, const vec4& b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }

// glm/detail/func_common.inl:17-30 — select form, NOT fminf/fmaxf (NaN behaviour, SURVEY A-Q20)
inline float gmin(float x, float y) { return (y < x) ? y : x; }
inline float gmax(float x, float y) { return (x < y) ? y : x; }
inline vec3 gmin(const vec3& a, const vec3& b) { return vec3(gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)); }
inline vec3 gmax(const vec3& a, const vec3& b) { return vec3(gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)); }
inline float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }  // func_common.inl:505

// glm/detail/func_geometric.inl:49-66
inline float dot(const vec3& a, const vec3& b) { vec3 t = a * b; return t.x + t.y + t.z; }
inline float dot(const vec4& a, const vec4& b) { vec4 t = a * b; return (t.x + t.y) + (t.z + t.w); }
// func_geometric.inl:69-80
inline vec3 cross(const vec3& x, const vec3& y) {
    return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
// func_exponential.inl:136-139
inline float inversesqrt(float x) { return 1.f / std::sqrt(x); }
// func_geometric.inl:83-92
inline vec3 normalize(const vec3& v) { return v * inversesqrt(dot(v, v)); }
inline vec4 normalize(const vec4& v) { return v * inversesqrt(dot(v, v)); }
// func_geometric.inl:106-113
inline vec3 reflect(const vec3& I, const vec3& N) { return I - N * dot(N, I) * 2.f; }
// func_geometric.inl:116-127
inline vec3 refract(const vec3& I, const vec3& N, float eta) {
    float const d = dot(N, I);
    float const k = 1.f - eta * eta * (1.f - d * d);
    return (k >= 0.f) ? (eta * I - (eta * d + std::sqrt(k)) * N) : vec3(0.f);
}
// func_common.inl:123-132 (scalar) and :104-112 (vector with scalar interpolant)
inline float mix(float x, float y, float a) { return x * (1.f - a) + y * a; }
inline vec3 mix(const vec3& x, const vec3& y, float a) { return x * (1.f - a) + y * a; }
// func_common.inl:564-570
inline float smoothstep(float e0, float e1, float x) {
    float const t = gclamp((x - e0) / (e1 - e0), 0.f, 1.f);
    return t * t * (3.f - 2.f * t);
}
// func_common.inl:144-150  (0<x) - (x<0)
inline float sign(float x) { return float(0.f < x) - float(x < 0.f); }
// func_trigonometric.inl:9-14
inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }

struct mat3 {
    vec3 c[3];
    vec3& operator[](int i) { return c[i]; }
    const vec3& operator[](int i) const { return c[i]; }
};
// type_mat3x3.inl:468-474
inline vec3 operator*(const mat3& m, const vec3& v) {
    return vec3(m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z,
                m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z,
                m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z);
}
// type_mat3x3.inl:486-520
inline mat3 operator*(const mat3& a, const mat3& b) {
    mat3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
            r[j][i] = a[0][i] * b[j][0] + a[1][i] * b[j][1] + a[2][i] * b[j][2];
    return r;
}

struct mat4 {
    vec4 c[4];
    mat4() {}
    explicit mat4(float d) { c[0] = vec4(d, 0, 0, 0); c[1] = vec4(0, d, 0, 0); c[2] = vec4(0, 0, d, 0); c[3] = vec4(0, 0, 0, d); }
    explicit mat4(const mat3& m) {  // type_mat4x4.inl mat3->mat4 conversion: pad with identity
        c[0] = vec4(m[0], 0.f); c[1] = vec4(m[1], 0.f); c[2] = vec4(m[2], 0.f); c[3] = vec4(0.f, 0.f, 0.f, 1.f);
    }
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};
// type_mat4x4.inl mat*vec: (m0*x + m1*y) + (m2*z + m3*w)
inline vec4 operator*(const mat4& m, const vec4& v) {
    vec4 const Mul0 = m[0] * vec4(v.x);
    vec4 const Mul1 = m[1] * vec4(v.y);
    vec4 const Add0 = Mul0 + Mul1;
    vec4 const Mul2 = m[2] * vec4(v.z);
    vec4 const Mul3 = m[3] * vec4(v.w);
    vec4 const Add1 = Mul2 + Mul3;
    return Add0 + Add1;
}
// type_mat4x4.inl mat*mat: ((A0*b0 + A1*b1) + A2*b2) + A3*b3 per column
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r;
    for (int j = 0; j < 4; ++j)
        r[j] = a[0] * b[j][0] + a[1] * b[j][1] + a[2] * b[j][2] + a[3] * b[j][3];
    return r;
}
// func_matrix.inl:294-351
inline mat4 inverse(const mat4& m) {
    float Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    float Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
    float Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    float Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    float Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
    float Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    float Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    float Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
    float Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    float Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    float Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
    float Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    float Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    float Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
    float Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    float Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    float Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
    float Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];

    vec4 Fac0(Coef00, Coef00, Coef02, Coef03);
    vec4 Fac1(Coef04, Coef04, Coef06, Coef07);
    vec4 Fac2(Coef08, Coef08, Coef10, Coef11);
    vec4 Fac3(Coef12, Coef12, Coef14, Coef15);
    vec4 Fac4(Coef16, Coef16, Coef18, Coef19);
    vec4 Fac5(Coef20, Coef20, Coef22, Coef23);

    vec4 Vec0(m[1][0], m[0][0], m[0][0], m[0][0]);
    vec4 Vec1(m[1][1], m[0][1], m[0][1], m[0][1]);
    vec4 Vec2(m[1][2], m[0][2], m[0][2], m[0][2]);
    vec4 Vec3(m[1][3], m[0][3], m[0][3], m[0][3]);

    vec4 Inv0(Vec1 * Fac0 - Vec2 * Fac1 + Vec3 * Fac2);
    vec4 Inv1(Vec0 * Fac0 - Vec2 * Fac3 + Vec3 * Fac4);
    vec4 Inv2(Vec0 * Fac1 - Vec1 * Fac3 + Vec3 * Fac5);
    vec4 Inv3(Vec0 * Fac2 - Vec1 * Fac4 + Vec2 * Fac5);

    vec4 SignA(+1.f, -1.f, +1.f, -1.f);
    vec4 SignB(-1.f, +1.f, -1.f, +1.f);
    mat4 Inverse;
    Inverse[0] = Inv0 * SignA; Inverse[1] = Inv1 * SignB; Inverse[2] = Inv2 * SignA; Inverse[3] = Inv3 * SignB;

    vec4 Row0(Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0]);
    vec4 Dot0(m[0] * Row0);
    float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
    float OneOverDeterminant = 1.f / Dot1;
    mat4 r;
    for (int j = 0; j < 4; ++j) r[j] = Inverse[j] * OneOverDeterminant;
    return r;
}
// ext/matrix_transform.inl:10-15 with m = identity (gtx/transform.hpp translate(v))
inline mat4 translate(const vec3& v) {
    mat4 m(1.f), r(1.f);
    r[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3];
    return r;
}
// ext/matrix_transform.inl:18-46 with m = identity (gtx/transform.hpp rotate(angle, v))
inline mat4 rotate(float angle, const vec3& v) {
    mat4 m(1.f);
    float const c = std::cos(angle);
    float const s = std::sin(angle);
    vec3 axis(normalize(v));
    vec3 temp((1.f - c) * axis);
    float R[3][3];
    R[0][0] = c + temp[0] * axis[0];
    R[0][1] = temp[0] * axis[1] + s * axis[2];
    R[0][2] = temp[0] * axis[2] - s * axis[1];
    R[1][0] = temp[1] * axis[0] - s * axis[2];
    R[1][1] = c + temp[1] * axis[1];
    R[1][2] = temp[1] * axis[2] + s * axis[0];
    R[2][0] = temp[2] * axis[0] + s * axis[1];
    R[2][1] = temp[2] * axis[1] - s * axis[0];
    R[2][2] = c + temp[2] * axis[2];
    mat4 r;
    r[0] = m[0] * R[0][0] + m[1] * R[0][1] + m[2] * R[0][2];
    r[1] = m[0] * R[1][0] + m[1] * R[1][1] + m[2] * R[1][2];
    r[2] = m[0] * R[2][0] + m[1] * R[2][1] + m[2] * R[2][2];
    r[3] = m[3];
    return r;
}
// ext/matrix_transform.inl:78-86 with m = identity (gtx/transform.hpp scale(v))
inline mat4 scale(const vec3& v) {
    mat4 m(1.f), r;
    r[0] = m[0] * v[0]; r[1] = m[1] * v[1]; r[2] = m[2] * v[2]; r[3] = m[3];
    return r;
}

inline bool isnan3(const vec3& v) { return std::isnan(v.x) || std::isnan(v.y) || std::isnan(v.z); }
inline bool isinf3(const vec3& v) { return std::isinf(v.x) || std::isinf(v.y) || std::isinf(v.z); }

}  // namespace om
