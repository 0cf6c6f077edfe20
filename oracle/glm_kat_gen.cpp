// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Known-answer generator that evaluates the REAL GLM 0.9.9.7 vendored by the reference
// (include path /root/reference/third_party/glm, headers used where they lie, nothing
// copied). Built into oracle/_ref/glm_kat_gen by `make -C oracle ref`; its output is the
// committed fixture tests/golden/glm_kat.json, which pins oracle_math.h bit-for-bit.
//
// Output: one JSON object {"cases":[{"fn":..., "in":[u32 bit patterns], "out":[u32 bit patterns]}, ...]}
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <vector>
#include <string>

#include <glm/glm.hpp>
#include <glm/gtx/transform.hpp>

static uint64_t s = 0x9E3779B97F4A7C15ull;
static uint32_t rnd() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return uint32_t(s >> 16); }
static float rf(float lo, float hi) { return lo + (hi - lo) * (rnd() / 4294967296.0f); }
static uint32_t bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

static bool first = true;
static void emit(const char* fn, const std::vector<float>& in, const std::vector<float>& out) {
    std::printf("%s{\"fn\":\"%s\",\"in\":[", first ? "" : ",\n", fn);
    first = false;
    for (size_t i = 0; i < in.size(); ++i) std::printf("%s%u", i ? "," : "", bits(in[i]));
    std::printf("],\"out\":[");
    for (size_t i = 0; i < out.size(); ++i) std::printf("%s%u", i ? "," : "", bits(out[i]));
    std::printf("]}");
}
static std::vector<float> V(const glm::vec3& v) { return {v.x, v.y, v.z}; }
static std::vector<float> V(const glm::vec4& v) { return {v.x, v.y, v.z, v.w}; }
static std::vector<float> V(const glm::mat4& m) {
    std::vector<float> o;
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) o.push_back(m[c][r]);
    return o;
}
static std::vector<float> cat(std::vector<float> a, const std::vector<float>& b) { a.insert(a.end(), b.begin(), b.end()); return a; }
static glm::vec3 rv3(float lo = -2.f, float hi = 2.f) { return glm::vec3(rf(lo, hi), rf(lo, hi), rf(lo, hi)); }
static glm::vec4 rv4(float lo = -2.f, float hi = 2.f) { return glm::vec4(rf(lo, hi), rf(lo, hi), rf(lo, hi), rf(lo, hi)); }
static glm::mat4 rm4() { return glm::mat4(rv4(), rv4(), rv4(), rv4()); }

int main() {
    std::printf("{\"glm_version\":%d,\"cases\":[\n", GLM_VERSION);
    for (int i = 0; i < 40; ++i) {
        glm::vec3 a = rv3(), b = rv3();
        emit("dot3", cat(V(a), V(b)), {glm::dot(a, b)});
        emit("cross", cat(V(a), V(b)), V(glm::cross(a, b)));
        emit("normalize3", V(a), V(glm::normalize(a)));
        glm::vec3 n = glm::normalize(b);
        emit("reflect", cat(V(a), V(n)), V(glm::reflect(a, n)));
        glm::vec3 I = glm::normalize(a);
        float eta = rf(0.3f, 2.5f);
        emit("refract", cat(cat(V(I), V(n)), {eta}), V(glm::refract(I, n, eta)));
        glm::vec4 p = rv4(), q = rv4();
        emit("dot4", cat(V(p), V(q)), {glm::dot(p, q)});
        emit("normalize4", V(p), V(glm::normalize(p)));
        float t = rf(-0.2f, 1.2f);
        emit("mix1", {a.x, b.x, t}, {glm::mix(a.x, b.x, t)});
        emit("mix3", cat(cat(V(a), V(b)), {t}), V(glm::mix(a, b, t)));
        float e0 = rf(-1.f, 0.f), e1 = rf(0.1f, 1.f), x = rf(-1.5f, 1.5f);
        emit("smoothstep", {e0, e1, x}, {glm::smoothstep(e0, e1, x)});
        emit("sign", {x}, {glm::sign(x)});
        emit("radians", {x * 100.f}, {glm::radians(x * 100.f)});
        emit("min", {a.x, b.x}, {glm::min(a.x, b.x)});
        emit("max", {a.x, b.x}, {glm::max(a.x, b.x)});
        glm::mat4 m = rm4(), m2 = rm4();
        emit("inverse", V(m), V(glm::inverse(m)));
        emit("mat4vec4", cat(V(m), V(p)), V(m * p));
        emit("mat4mat4", cat(V(m), V(m2)), V(m * m2));
        emit("translate", V(a), V(glm::translate(a)));
        float ang = rf(-3.2f, 3.2f);
        glm::vec3 axes[3] = {glm::vec3(1.f, 0.f, 0.f), glm::vec3(0.f, 1.f, 0.f), glm::vec3(0.f, 0.f, 1.f)};
        glm::vec3 ax = (i % 4 == 3) ? b : axes[i % 3];
        emit("rotate", cat({ang}, V(ax)), V(glm::rotate(ang, ax)));
        emit("scale", V(a), V(glm::scale(a)));
        // the reference's object matrix, src/vk_engine.cpp:807-811
        glm::vec3 pos = rv3(), rot = rv3(-180.f, 180.f), scl = rv3(0.1f, 2.f);
        if (i % 5 == 0) rot = glm::vec3(90.f, 0.f, -90.f);
        if (i % 5 == 1) { rot = glm::vec3(0.f); scl = glm::vec3(1.f); pos = glm::vec3(0.f); }
        glm::mat4 om = glm::translate(pos) * glm::rotate(glm::radians(rot.x), glm::vec3(1.f, 0.f, 0.f)) *
                       glm::rotate(glm::radians(rot.y), glm::vec3(0.f, 1.f, 0.f)) *
                       glm::rotate(glm::radians(rot.z), glm::vec3(0.f, 0.f, 1.f)) * glm::scale(scl);
        emit("object_matrix", cat(cat(V(pos), V(rot)), V(scl)), V(om));
        emit("inverse", V(om), V(glm::inverse(om)));
        // the reference's camera matrix, src/vk_engine.cpp:1633-1653
        float angs[3] = {rf(-90.f, 90.f), rf(-180.f, 180.f), rf(-30.f, 30.f)};
        if (i == 0) { angs[0] = 4.f; angs[1] = 0.f; angs[2] = 0.f; }
        float thetaX = glm::radians(angs[0]), thetaY = glm::radians(angs[1]), thetaZ = glm::radians(angs[2]);
        glm::mat3 rotX = glm::mat3(glm::vec3(1, 0, 0), glm::vec3(0, std::cos(thetaX), -std::sin(thetaX)), glm::vec3(0, std::sin(thetaX), std::cos(thetaX)));
        glm::mat3 rotY = glm::mat3(glm::vec3(std::cos(thetaY), 0, std::sin(thetaY)), glm::vec3(0, 1, 0), glm::vec3(-std::sin(thetaY), 0, std::cos(thetaY)));
        glm::mat3 rotZ = glm::mat3(glm::vec3(std::cos(thetaZ), -std::sin(thetaZ), 0), glm::vec3(std::sin(thetaZ), std::cos(thetaZ), 0), glm::vec3(0, 0, 1));
        glm::mat4 cam = glm::mat4(rotY * rotX * rotZ);
        emit("camera_rotation", {angs[0], angs[1], angs[2]}, V(cam));
        glm::mat3 m3(a, b, n);
        glm::vec3 r3 = m3 * I;
        emit("mat3vec3", cat(cat(cat(V(a), V(b)), V(n)), V(I)), V(r3));
    }
    // NaN / signed-zero behaviour of the select-form min/max (SURVEY A-Q20)
    float nan = std::nanf(""), inf = INFINITY;
    float sp[][2] = {{nan, 1.f}, {1.f, nan}, {0.f, -0.f}, {-0.f, 0.f}, {inf, -inf}, {nan, nan}};
    for (auto& pr : sp) {
        emit("min", {pr[0], pr[1]}, {glm::min(pr[0], pr[1])});
        emit("max", {pr[0], pr[1]}, {glm::max(pr[0], pr[1])});
    }
    std::printf("\n]}\n");
    return 0;
}
