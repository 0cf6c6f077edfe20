// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// CPU restatement of the reference compute shader shaders/raytrace.comp.
// PARITY STATUS: the device-side built-ins of the reference (SPIR-V GLSL.std.450 as
// implemented by the author's Metal driver) are not pinned by any reference test or
// golden vector ("parity unpinned" for last-ulp behaviour). What IS
// pinned and checked in tests/: the RNG/seed stream (Appendix B), the struct layouts
// (A-Q1), GLM 0.9.9.7 operation order (tests/golden/glm_kat.json), the BVH builder.
#pragma once
#include <cstdint>
#include <vector>
#include "oracle_layout.h"
#include "oracle_math.h"

namespace orc {

// Algorithmic-work counters of SURVEY.md §8(d4), extended to ALL rays
// (primary, NEE and both pdf rays), in reference-layout units.
struct Work {
    uint64_t n_box = 0;     // 2 per inner node visited   (raytrace.comp:338)
    uint64_t n_tri = 0;     // triCount per leaf visited  (raytrace.comp:310)
    uint64_t n_obj = 0;     // object-loop iterations     (raytrace.comp:289)
    uint64_t n_sph = 0;     // sphere tests               (raytrace.comp:282)
    uint64_t n_hit = 0;     // rays that returned a hit (one material fetch)
    uint64_t n_ray = 0;     // calculateIntersections calls
    uint64_t n_bounce = 0;  // bounce-loop iterations that found a hit
    uint64_t n_path = 0;    // trace() calls
    void add(const Work& o) {
        n_box += o.n_box; n_tri += o.n_tri; n_obj += o.n_obj; n_sph += o.n_sph;
        n_hit += o.n_hit; n_ray += o.n_ray; n_bounce += o.n_bounce; n_path += o.n_path;
    }
};

struct Scene {
    std::vector<ol::Sphere> spheres;
    std::vector<ol::RayMaterial> materials;
    std::vector<ol::TrianglePoint> points;
    std::vector<ol::Triangle> triangles;
    std::vector<ol::RenderObject> objects;
    std::vector<ol::BVHNode> nodes;
    // hoisted out of the ray loop; bit-identical to evaluating inverse() per ray
    // (raytrace.comp:292-293) because inverse() is a pure function of the matrix.
    std::vector<om::mat4> M, Minv;
    void finalize();
    // TEXTURE EXTENSION — NOT in the HEAD shader, which declares TextureBuffer[64] and computes
    // hit.uv but never samples (raytrace.comp:122,148,249-256). Semantics defined by this project ("parity
    // unpinned" against the reference): see sample_texture() in oracle_shader.cpp. Off unless useTextures is set.
    struct Texture { std::vector<uint8_t> rgba; uint32_t w = 0, h = 0; };
    Texture textures[64];      // TextureBuffer[slot], RGBA8 with sRGB-encoded colour (src/vk_textures.cpp:103-200)
    bool useTextures = false;
};

struct HitOut {  // first-hit record for the fixed-ray-batch integer parity test
    float dst;
    uint32_t kind;            // 0 miss, 1 sphere, 2 triangle
    uint32_t objectHitIndex;  // 0xffffffff for spheres / miss (uninitialised in the reference)
    uint32_t triHitIndex;     // triangle index, or sphere index for kind 1
    uint32_t materialIndex;   // pinned to 0 on miss (A-Q9)
    uint32_t frontFace;
    float normal[3];
    float hitPoint[3];
};
static_assert(sizeof(HitOut) == 48, "HitOut");

// One literal dispatch of main() (raytrace.comp:539-594) over a W×H float image.
//   image  : W*H*4 floats, in/out — the storage image in float instead of rgba8
//   rgba8  : optional W*H*4 bytes, in/out — literal 8-bit read-modify-write (A-Q3, unorm RN)
//   outcol : optional W*H*3 floats — outColor (mean over in-dispatch samples) before the blend
//   stats  : optional W*H*2 floats — the shader's stats[0], stats[1]
void render_dispatch(const Scene& sc, const ol::PushConstants& pc, int W, int H, float* image,
                     uint8_t* rgba8, float* outcol, float* stats, Work* work, int threads);

// Throughput form: frames first..first+n-1 with frameCount=f; accum.rgb += outColor_f
// (non-finite outColor contributes magenta, mirroring raytrace.comp:579-581); accum.a += 1.
void render_window(const Scene& sc, const ol::PushConstants& pc, int W, int H, uint32_t first, uint32_t n, int x0, int y0,
                   int w, int h, int sx, int sy, float* accum, Work* work = nullptr, int threads = 0, Work* workExt = nullptr);
void render_frames(const Scene& sc, const ol::PushConstants& pc, int W, int H, uint32_t first, uint32_t n,
                   float* accum, Work* work, int threads, Work* workExt = nullptr);

// calculateIntersections (raytrace.comp:276-353) on caller-supplied rays (8 floats each:
// origin.xyz, pad, dir.xyz, pad).
void trace_batch(const Scene& sc, const ol::PushConstants& pc, const float* rays, size_t n, HitOut* out,
                 Work* work, int threads);

// tie detector: see oracle_shader.cpp
void pixel_ties(const Scene& sc, const ol::PushConstants& pc, int W, int H, uint32_t frame, const uint32_t* xy, size_t n,
                uint32_t* ties);

size_t pixel_rays(const Scene& sc, const ol::PushConstants& pc, int W, int H, uint32_t frame, uint32_t x, uint32_t y, size_t cap,
                  float* rays8, HitOut* hits);

// 0 = libm transcendentals (default), 1 = the backend's portable double-precision polynomials
void set_math_mode(int mode);
int math_mode();

// exposed for known-answer tests
float random(uint32_t& state);
uint32_t starting_seed(uint32_t frameCount);

}  // namespace orc
