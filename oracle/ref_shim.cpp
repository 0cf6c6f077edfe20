// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Harness around the REFERENCE'S OWN SHADER TEXT. oracle/ref_from_glsl.py rewrites
// /root/reference/shaders/raytrace.comp syntactically into oracle/_ref/raytrace_comp.inc (git-ignored);
// this file supplies what a GLSL compiler would: vector types and built-ins (the reference's vendored GLM
// 0.9.9.7, used in place from /root/reference/third_party/glm), the storage image, gl_GlobalInvocationID, the
// std140 decode of the six storage buffers and of the 208-byte push-constant block, and a plain-C API with the
// same shapes as oracle_api.cpp. Built by `make -C oracle ref` into oracle/_ref/libraytrace_comp.so.
// Nothing here restates shader arithmetic: every float result comes from the included text + GLM.
//
// Built-ins are declared with float parameters so that GLSL's implicit int -> float conversions
// (`max(0, x)`, `pow(x, 5)`, `smoothstep(0, 0.4, t)`, `vec3 * int`, `1 / vec3`) resolve as GLSL resolves them.
// Transcendentals are the host libm's (glm::sin -> std::sin(float) ...): the reference ran on a Metal driver whose
// last-ulp behaviour nothing in the repository pins ("parity unpinned" at that boundary).
#define GLM_FORCE_SWIZZLE
#include <glm/glm.hpp>

#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace glsl {

using glm::ivec2;
using glm::mat3;
using glm::mat4;
using glm::uvec2;
using glm::uvec3;
using glm::vec2;
using glm::vec3;
using glm::vec4;
typedef uint32_t uint;

// ---- scalar built-ins (GLSL.std.450) ----
inline float sqrt(float x) { return glm::sqrt(x); }
inline float log(float x) { return glm::log(x); }
inline float cos(float x) { return glm::cos(x); }
inline float sin(float x) { return glm::sin(x); }
inline float tan(float x) { return glm::tan(x); }
inline float pow(float x, float y) { return glm::pow(x, y); }
inline float abs(float x) { return glm::abs(x); }
inline float sign(float x) { return glm::sign(x); }
inline float radians(float x) { return glm::radians(x); }
inline float min(float a, float b) { return glm::min(a, b); }
inline float max(float a, float b) { return glm::max(a, b); }
inline float mix(float a, float b, float t) { return glm::mix(a, b, t); }
inline float smoothstep(float e0, float e1, float x) { return glm::smoothstep(e0, e1, x); }
inline bool isnan(float x) { return glm::isnan(x); }
// ---- vector built-ins ----
inline vec3 normalize(const vec3& v) { return glm::normalize(v); }
inline vec4 normalize(const vec4& v) { return glm::normalize(v); }
inline float dot(const vec3& a, const vec3& b) { return glm::dot(a, b); }
inline vec3 cross(const vec3& a, const vec3& b) { return glm::cross(a, b); }
inline vec3 reflect(const vec3& i, const vec3& n) { return glm::reflect(i, n); }
inline vec3 refract(const vec3& i, const vec3& n, float eta) { return glm::refract(i, n, eta); }
inline vec3 min(const vec3& a, const vec3& b) { return glm::min(a, b); }
inline vec3 max(const vec3& a, const vec3& b) { return glm::max(a, b); }
inline vec3 mix(const vec3& a, const vec3& b, float t) { return glm::mix(a, b, t); }
inline mat4 inverse(const mat4& m) { return glm::inverse(m); }
inline glm::bvec3 isnan(const vec3& v) { return glm::isnan(v); }
inline glm::bvec3 isinf(const vec3& v) { return glm::isinf(v); }
inline bool any(const glm::bvec3& b) { return glm::any(b); }
// ---- implicit-conversion operators GLSL has and C++ templates do not deduce ----
inline vec3 operator*(const vec3& v, int s) { return v * float(s); }
inline vec3 operator/(int s, const vec3& v) { return float(s) / v; }
inline vec3 operator/(const vec3& v, uint s) { return v / float(s); }
inline vec2 operator/(const vec2& a, const ivec2& b) { return a / vec2(b); }
// unary minus of a swizzle (`-env.lightDir.rgb`): GLM's swizzle proxies only define the binary operators
template <int N, typename T, glm::qualifier Q, int E0, int E1, int E2, int E3>
inline glm::vec<N, T, Q> operator-(const glm::detail::_swizzle<N, T, Q, E0, E1, E2, E3>& s) { return -s(); }

// ---- storage image + invocation id: one pixel per call of glsl_main(), state per host thread ----
struct image2D {};
struct InvocationID { uint x, y, z; uvec2 xy; };
static thread_local InvocationID gl_GlobalInvocationID;
static thread_local ivec2 tl_imageSize;
static thread_local vec4 tl_imageOld, tl_imageNew;
inline ivec2 imageSize(const image2D&) { return tl_imageSize; }
inline vec4 imageLoad(const image2D&, const ivec2&) { return tl_imageOld; }
inline void imageStore(const image2D&, const ivec2&, const vec4& v) { tl_imageNew = v; }

#include "_ref/raytrace_comp.inc"

}  // namespace glsl

// ------------------------------------------------------------------------------------------------------
namespace {

using namespace glsl;

float rdf(const uint8_t* p) { float f; std::memcpy(&f, p, 4); return f; }
uint32_t rdu(const uint8_t* p) { uint32_t u; std::memcpy(&u, p, 4); return u; }
int32_t rdi(const uint8_t* p) { int32_t u; std::memcpy(&u, p, 4); return u; }
vec2 rd2(const uint8_t* p) { return vec2(rdf(p), rdf(p + 4)); }
vec3 rd3(const uint8_t* p) { return vec3(rdf(p), rdf(p + 4), rdf(p + 8)); }
vec4 rd4(const uint8_t* p) { return vec4(rdf(p), rdf(p + 4), rdf(p + 8), rdf(p + 12)); }
mat4 rdm(const uint8_t* p) { mat4 m; for (int c = 0; c < 4; ++c) m[c] = rd4(p + 16 * c); return m; }

// std140 decode of the six storage buffers (offsets: SURVEY Appendix A-Q1, src/vk_engine.h:49-123,185-189)
struct RefScene {
    std::vector<Sphere> spheres;
    std::vector<Material> materials;
    std::vector<TrianglePoint> points;
    std::vector<Triangle> triangles;
    std::vector<RenderObject> objects;
    std::vector<BVHNode> nodes;
};

void bind_scene(const RefScene& s) {
    glsl::spheres = s.spheres.data();
    glsl::materials = s.materials.data();
    glsl::trianglePoints = s.points.data();
    glsl::triangles = s.triangles.data();
    glsl::objects = s.objects.data();
    glsl::bvhNodes = s.nodes.data();
}

// 208-byte push-constant block (src/vk_engine.h:145-178). bools: the host writes a 1-byte C++ bool into a 4-byte
// slot; read as byte != 0 (same pin as the oracle, A-Q1).
void set_push_constants(const uint8_t* b) {
    auto& P = glsl::PushConstants;
    P.camInfo.cameraRotation = rdm(b);
    P.camInfo.pos = rd3(b + 64);
    P.camInfo.nearPlane = rdf(b + 76);
    P.camInfo.aspectRatio = rdf(b + 80);
    P.camInfo.fov = rdf(b + 84);
    P.environment.horizonColor = rd4(b + 96);
    P.environment.zenithColor = rd4(b + 112);
    P.environment.groundColor = rd3(b + 128);
    P.environment.lightDir = rd4(b + 144);
    const uint8_t* r = b + 160;
    P.rayTracerParams.progressive = r[0] != 0;
    P.rayTracerParams.singleRender = r[4] != 0;
    P.rayTracerParams.debugMode = rdi(r + 8);
    P.rayTracerParams.raysPerPixel = rdu(r + 12);
    P.rayTracerParams.bounceLimit = rdu(r + 16);
    P.rayTracerParams.sphereCount = rdu(r + 20);
    P.rayTracerParams.objectCount = rdu(r + 24);
    P.rayTracerParams.triCap = rdu(r + 28);
    P.rayTracerParams.boxCap = rdu(r + 32);
    P.rayTracerParams.sampleLimit = rdu(r + 36);
    P.frameCount = rdu(b + 200);
}

// one invocation of the shader's main() for pixel (x, y) of a W x H storage image
inline vec4 invoke(uint x, uint y, int W, int H, const vec4& oldColor) {
    glsl::gl_GlobalInvocationID.x = x;
    glsl::gl_GlobalInvocationID.y = y;
    glsl::gl_GlobalInvocationID.z = 0;
    glsl::gl_GlobalInvocationID.xy = uvec2(x, y);
    glsl::tl_imageSize = ivec2(W, H);
    glsl::tl_imageOld = oldColor;
    glsl::glsl_main();
    return glsl::tl_imageNew;
}

struct HitOut {  // same record as orc::HitOut (oracle_shader.h)
    float dst;
    uint32_t kind, objectHitIndex, triHitIndex, materialIndex, frontFace;
    float normal[3], hitPoint[3];
};
static_assert(sizeof(HitOut) == 48, "HitOut");

}  // namespace

extern "C" {

void* refsh_scene_create(const void* sph, uint32_t nS, const void* mat, uint32_t nM, const void* pts, uint32_t nP,
                         const void* tri, uint32_t nT, const void* obj, uint32_t nO, const void* nod, uint32_t nN) {
    RefScene* s = new RefScene();
    const uint8_t* p = (const uint8_t*)sph;
    for (uint32_t i = 0; i < nS; ++i, p += 32) {
        Sphere e{}; e.position = rd3(p); e.radius = rdf(p + 12); e.materialIndex = rdu(p + 16);
        s->spheres.push_back(e);
    }
    p = (const uint8_t*)mat;
    for (uint32_t i = 0; i < nM; ++i, p += 64) {
        Material e{}; e.albedo = rd3(p); e.emissionColor = rd3(p + 16); e.emissionStrength = rdf(p + 28);
        e.reflectance = rdf(p + 32); e.ior = rdf(p + 36); e.albedoIndex = rdi(p + 40); e.metalnessIndex = rdi(p + 44);
        e.alphaIndex = rdi(p + 48); e.bumpIndex = rdi(p + 52);
        s->materials.push_back(e);
    }
    p = (const uint8_t*)pts;
    for (uint32_t i = 0; i < nP; ++i, p += 32) {
        TrianglePoint e{}; e.position = rd4(p); e.normal = rd4(p + 16);
        s->points.push_back(e);
    }
    p = (const uint8_t*)tri;
    for (uint32_t i = 0; i < nT; ++i, p += 48) {
        Triangle e{}; e.v0 = rdu(p); e.v1 = rdu(p + 4); e.v2 = rdu(p + 8); e.frontOnly = rdu(p + 12);
        e.binormal = rd3(p + 16); e.tangent = rd3(p + 32);
        s->triangles.push_back(e);
    }
    p = (const uint8_t*)obj;
    for (uint32_t i = 0; i < nO; ++i, p += 80) {
        RenderObject e{}; e.transformMatrix = rdm(p); e.smoothShade = rdu(p + 64); e.bvhIndex = rdu(p + 68);
        e.materialIndex = rdu(p + 72); e.samplerIndex = rdu(p + 76);
        s->objects.push_back(e);
    }
    p = (const uint8_t*)nod;
    for (uint32_t i = 0; i < nN; ++i, p += 32) {
        BVHNode e{}; e.boundsX = rd2(p); e.boundsY = rd2(p + 8); e.boundsZ = rd2(p + 16); e.index = rdu(p + 24);
        e.triCount = rdu(p + 28);
        s->nodes.push_back(e);
    }
    return s;
}
void refsh_scene_destroy(void* s) { delete (RefScene*)s; }

// one literal dispatch: image (W*H*4 floats) is the storage image, read and written as the shader does
int refsh_dispatch(void* scene, const void* pc208, int W, int H, float* image, int threads) {
    bind_scene(*(RefScene*)scene);
    set_push_constants((const uint8_t*)pc208);
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float* px = image + 4 * (size_t(y) * W + x);
            vec4 o = invoke(uint(x), uint(y), W, H, vec4(px[0], px[1], px[2], px[3]));
            px[0] = o.x; px[1] = o.y; px[2] = o.z; px[3] = o.w;
        }
    return 0;
}

// pixels (x0 + i*sx, y0 + j*sy) of the W x H image, frames first..first+n-1: accum.rgb += what main() stores
// with progressive = false and debugMode = -1 (= outColor, magenta when non-finite), accum.a += 1.
// Same contract as orc_render_window.
int refsh_render_window(void* scene, const void* pc208, int W, int H, uint32_t first, uint32_t n, int x0, int y0, int w,
                        int h, int sx, int sy, float* accum, int threads) {
    if (w < 0 || h < 0 || sx < 1 || sy < 1 || x0 < 0 || y0 < 0) return 1;
    bind_scene(*(RefScene*)scene);
    uint8_t pc[208];
    std::memcpy(pc, pc208, 208);
    pc[160] = 0;                       // progressive = false
    int32_t dbg = -1;
    std::memcpy(pc + 168, &dbg, 4);    // debugMode = -1
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    for (uint32_t f = first; f < first + n; ++f) {
        std::memcpy(pc + 200, &f, 4);
        set_push_constants(pc);
#pragma omp parallel for schedule(dynamic, 1)
        for (int j = 0; j < h; ++j)
            for (int i = 0; i < w; ++i) {
                vec4 o = invoke(uint(x0 + i * sx), uint(y0 + j * sy), W, H, vec4(0.f));
                float* px = accum + 4 * (size_t(j) * w + i);
                px[0] += o.x; px[1] += o.y; px[2] += o.z; px[3] += 1.f;
            }
    }
    return 0;
}

// calculateIntersections() on caller-supplied rays (8 floats each: origin.xyz, pad, dir.xyz, pad);
// stats2: optional 2 floats per ray = the shader's stats[0] (box tests), stats[1] (triangle tests).
// kind is 1 for any hit (the shader's HitInfo does not say sphere or triangle); objectHitIndex / triHitIndex are
// what the shader leaves in HitInfo (value-initialised 0 unless a triangle was hit).
int refsh_trace_batch(void* scene, const void* pc208, const float* rays, uint64_t n, void* hits48, float* stats2,
                      int threads) {
    bind_scene(*(RefScene*)scene);
    set_push_constants((const uint8_t*)pc208);
    HitOut* out = (HitOut*)hits48;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
        Ray r{};
        r.origin = vec3(rays[8 * i], rays[8 * i + 1], rays[8 * i + 2]);
        r.dir = vec3(rays[8 * i + 4], rays[8 * i + 5], rays[8 * i + 6]);
        float stats[2] = {0.f, 0.f};
        HitInfo h = calculateIntersections(r, stats);
        HitOut& o = out[i];
        std::memset(&o, 0, sizeof(o));
        o.dst = h.dst;
        o.kind = h.didHit ? 1u : 0u;
        o.objectHitIndex = h.objectHitIndex;
        o.triHitIndex = h.triHitIndex;
        o.materialIndex = h.materialIndex;
        o.frontFace = h.didHit ? (h.frontFace ? 1u : 0u) : 0u;
        if (h.didHit) {
            o.normal[0] = h.normal.x; o.normal[1] = h.normal.y; o.normal[2] = h.normal.z;
            o.hitPoint[0] = h.hitPoint.x; o.hitPoint[1] = h.hitPoint.y; o.hitPoint[2] = h.hitPoint.z;
        }
        if (stats2) { stats2[2 * i] = stats[0]; stats2[2 * i + 1] = stats[1]; }
    }
    return 0;
}

float refsh_random(uint32_t* state) { return glsl::random(*state); }

}  // extern "C"
