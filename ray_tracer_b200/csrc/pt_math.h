// Float vector algebra for the kernels. Every operator is ONE IEEE-754 single-precision
// operation, written in the order the reference's arithmetic evaluates (GLSL built-ins with
// GLM 0.9.9.7 semantics: third_party/glm/glm/detail/func_geometric.inl, func_common.inl,
// type_mat4x4.inl). The translation units that include this are compiled with -fmad=false,
// so nvcc never contracts a*b+c; where a fused multiply-add is wanted (conservative box
// tests only) it is written explicitly as pt_fma().
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define PT_HD __host__ __device__ __forceinline__
#define PT_HD_NOINLINE __host__ __device__ inline
#else
#define PT_HD inline
#define PT_HD_NOINLINE inline
#endif

namespace pt {

struct v3 { float x, y, z; };

PT_HD v3 mk3(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
PT_HD v3 splat3(float s) { return mk3(s, s, s); }
PT_HD v3 operator+(v3 a, v3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
PT_HD v3 operator-(v3 a, v3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
PT_HD v3 operator*(v3 a, v3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
PT_HD v3 operator*(v3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
PT_HD v3 operator*(float s, v3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
PT_HD v3 operator/(v3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
PT_HD v3 operator-(v3 a) { return mk3(-a.x, -a.y, -a.z); }

// select forms of min/max (func_common.inl:17-30): NaN behaviour must match, so no fminf/fmaxf here
PT_HD float gmin(float x, float y) { return (y < x) ? y : x; }
PT_HD float gmax(float x, float y) { return (x < y) ? y : x; }
PT_HD float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }

PT_HD float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }   // (x+y)+z, func_geometric.inl:49-66
PT_HD v3 cross(v3 x, v3 y) {                                                 // func_geometric.inl:69-80
    return mk3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
PT_HD float inversesqrt(float x) { return 1.f / sqrtf(x); }                  // func_exponential.inl:136-139
PT_HD v3 normalize(v3 v) { return v * inversesqrt(dot(v, v)); }              // func_geometric.inl:83-92
PT_HD v3 reflect(v3 I, v3 N) { return I - N * dot(N, I) * 2.f; }             // func_geometric.inl:106-113
PT_HD v3 refract(v3 I, v3 N, float eta) {                                    // func_geometric.inl:116-127
    float d = dot(N, I);
    float k = 1.f - eta * eta * (1.f - d * d);
    if (!(k >= 0.f)) return splat3(0.f);
    return eta * I - (eta * d + sqrtf(k)) * N;
}
PT_HD float mix(float x, float y, float a) { return x * (1.f - a) + y * a; } // func_common.inl:123-132
PT_HD v3 mix(v3 x, v3 y, float a) { return x * (1.f - a) + y * a; }
PT_HD float smoothstep(float e0, float e1, float x) {                        // func_common.inl:564-570
    float t = gclamp((x - e0) / (e1 - e0), 0.f, 1.f);
    return t * t * (3.f - 2.f * t);
}
PT_HD float sign(float x) { return float(0.f < x) - float(x < 0.f); }        // func_common.inl:144-150
PT_HD float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }
PT_HD bool isnan3(v3 v) { return (v.x != v.x) || (v.y != v.y) || (v.z != v.z); }
PT_HD bool isinf1(float f) { return f == INFINITY || f == -INFINITY; }
PT_HD bool isinf3(v3 v) { return isinf1(v.x) || isinf1(v.y) || isinf1(v.z); }

// Column-major 4x4 as the reference stores it. mat*vec4 is (m0*x + m1*y) + (m2*z + m3*w),
// type_mat4x4.inl; only the xyz rows are ever consumed by the shader.
struct m4 { float c[16]; };
PT_HD v3 mul_point(const float* m, v3 p) {    // (M * vec4(p, 1)).xyz
    return mk3((m[0] * p.x + m[4] * p.y) + (m[8] * p.z + m[12] * 1.f),
               (m[1] * p.x + m[5] * p.y) + (m[9] * p.z + m[13] * 1.f),
               (m[2] * p.x + m[6] * p.y) + (m[10] * p.z + m[14] * 1.f));
}
PT_HD v3 mul_dir(const float* m, v3 d) {      // (M * vec4(d, 0)).xyz
    return mk3((m[0] * d.x + m[4] * d.y) + (m[8] * d.z + m[12] * 0.f),
               (m[1] * d.x + m[5] * d.y) + (m[9] * d.z + m[13] * 0.f),
               (m[2] * d.x + m[6] * d.y) + (m[10] * d.z + m[14] * 0.f));
}
// normalize(M * vec4(n, 0)).xyz — the normalisation is over all FOUR components
// (raytrace.comp:318), dot4 = (x*x + y*y) + (z*z + w*w)
PT_HD v3 mul_dir_normalize4(const float* m, v3 d) {
    float x = (m[0] * d.x + m[4] * d.y) + (m[8] * d.z + m[12] * 0.f);
    float y = (m[1] * d.x + m[5] * d.y) + (m[9] * d.z + m[13] * 0.f);
    float z = (m[2] * d.x + m[6] * d.y) + (m[10] * d.z + m[14] * 0.f);
    float w = (m[3] * d.x + m[7] * d.y) + (m[11] * d.z + m[15] * 0.f);
    float s = inversesqrt((x * x + y * y) + (z * z + w * w));
    return mk3(x * s, y * s, z * s);
}

// Cofactor inverse in GLM's operation order (func_matrix.inl:294-351); the shader's
// inverse(object.transformMatrix) is hoisted to upload time (a pure function of the matrix).
PT_HD_NOINLINE void inverse4(const float* a, float* out) {
#define M(c, r) a[4 * (c) + (r)]
    float c00 = M(2,2) * M(3,3) - M(3,2) * M(2,3), c02 = M(1,2) * M(3,3) - M(3,2) * M(1,3), c03 = M(1,2) * M(2,3) - M(2,2) * M(1,3);
    float c04 = M(2,1) * M(3,3) - M(3,1) * M(2,3), c06 = M(1,1) * M(3,3) - M(3,1) * M(1,3), c07 = M(1,1) * M(2,3) - M(2,1) * M(1,3);
    float c08 = M(2,1) * M(3,2) - M(3,1) * M(2,2), c10 = M(1,1) * M(3,2) - M(3,1) * M(1,2), c11 = M(1,1) * M(2,2) - M(2,1) * M(1,2);
    float c12 = M(2,0) * M(3,3) - M(3,0) * M(2,3), c14 = M(1,0) * M(3,3) - M(3,0) * M(1,3), c15 = M(1,0) * M(2,3) - M(2,0) * M(1,3);
    float c16 = M(2,0) * M(3,2) - M(3,0) * M(2,2), c18 = M(1,0) * M(3,2) - M(3,0) * M(1,2), c19 = M(1,0) * M(2,2) - M(2,0) * M(1,2);
    float c20 = M(2,0) * M(3,1) - M(3,0) * M(2,1), c22 = M(1,0) * M(3,1) - M(3,0) * M(1,1), c23 = M(1,0) * M(2,1) - M(2,0) * M(1,1);
    float f0[4] = {c00, c00, c02, c03}, f1[4] = {c04, c04, c06, c07}, f2[4] = {c08, c08, c10, c11};
    float f3[4] = {c12, c12, c14, c15}, f4[4] = {c16, c16, c18, c19}, f5[4] = {c20, c20, c22, c23};
    float v0[4] = {M(1,0), M(0,0), M(0,0), M(0,0)}, v1[4] = {M(1,1), M(0,1), M(0,1), M(0,1)};
    float v2[4] = {M(1,2), M(0,2), M(0,2), M(0,2)}, v3_[4] = {M(1,3), M(0,3), M(0,3), M(0,3)};
    const float sa[4] = {1.f, -1.f, 1.f, -1.f}, sb[4] = {-1.f, 1.f, -1.f, 1.f};
    float inv[16];
    for (int i = 0; i < 4; ++i) {
        inv[0 + i] = (v1[i] * f0[i] - v2[i] * f1[i] + v3_[i] * f2[i]) * sa[i];
        inv[4 + i] = (v0[i] * f0[i] - v2[i] * f3[i] + v3_[i] * f4[i]) * sb[i];
        inv[8 + i] = (v0[i] * f1[i] - v1[i] * f3[i] + v3_[i] * f5[i]) * sa[i];
        inv[12 + i] = (v0[i] * f2[i] - v1[i] * f4[i] + v2[i] * f5[i]) * sb[i];
    }
    float d0 = M(0,0) * inv[0], d1 = M(0,1) * inv[4], d2 = M(0,2) * inv[8], d3 = M(0,3) * inv[12];
    float det = (d0 + d1) + (d2 + d3);
    float ood = 1.f / det;
    for (int i = 0; i < 16; ++i) out[i] = inv[i] * ood;
#undef M
}

PT_HD float pt_fma(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
PT_HD float as_float(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { uint32_t u; float f; } c; c.u = u; return c.f;
#endif
}
PT_HD uint32_t as_uint(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { uint32_t u; float f; } c; c.f = f; return c.u;
#endif
}

}  // namespace pt
