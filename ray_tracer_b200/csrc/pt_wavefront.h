// Wavefront form of trace()/main() (shaders/raytrace.comp:483-594): the bounce loop is turned
// inside out — all paths of a batch of progressive frames advance one bounce per stage launch,
// survivors are compacted into queues. Per-item stage bodies live here as plain functions so
// that the CUDA kernels (wavefront.cu) stay thin wrappers; the arithmetic per path is the
// shader's, in the shader's order (pt_shade.h).
//
//   generate : main() :541-566       primary ray + per-pixel RNG seed for every (frame, pixel)
//   extend   : calculateIntersections :276-353   closest hit (spheres, then the wide BVH)
//   shade    : trace() :492-535      emission / deferred direct light, BSDF sample, roulette
//   connect  : diffuseBRDF :441-460  NEE + the two lightSamplePDF queries -> directLight, MIS weight
//   accumulate : main() :568-581     frames summed per pixel in frame order
//
// Result-preserving restructurings: (1) the NEE ray and lightSamplePDF(lightSample)
// are the same ray (A-Q10) and are traced once; (2) lightSamplePDF only needs "is the closest hit
// emissive, and its distance": closest hit among EMISSIVE geometry (tiny) + a bounded occlusion
// query against the rest, instead of two more full closest-hit searches.
#pragma once
#include "pt_traverse.h"
#include "pt_reftrace.h"

namespace pt {

#if !defined(__CUDACC__)
struct float4 { float x, y, z, w; };
inline float4 make_float4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
#endif

struct PathArrays {
    float4* rayO;      // origin.xyz | misWeight
    float4* rayD;      // dir.xyz    | rng state (bits)
    float4* atten;     // attenuation.xyz
    float4* total;     // totalColor.xyz
    float4* direct;    // directLight.xyz (x == -1: specular sentinel, raytrace.comp:468,480,500)
    Hit* hit;          // extend result
    float4* qO;        // NEE origin.xyz (hit + n*0.01) | cosinePDF(lightSample)
    float4* qL;        // lightSample.xyz               | realCosinePDF
    float4* qA;        // albedo*INV_PI*max(0, n.lightSample) | material of the emissive light-sample hit
    float4* qE;        // emissive candidate distances and loop-order keys of the two pdf rays (connect)
};

struct WfParams {
    ptl::PushConstants pc;
    CameraPlane cp;
    int W, H;
    uint32_t row0, rows;            // rows rendered: row0 + k * rowStride, k < rows
    uint32_t rowStride;             // 1 = a contiguous band; N = every N-th row (interleaved split of one frame over N GPUs)
    uint32_t firstFrame, stride;    // slot k renders frameCount = firstFrame + k*stride
    uint32_t nSlots, nPaths;        // nPaths = nSlots * rows * W
    uint32_t triVote32;             // persistent kernels: run a triangle step when >= triVote32/32 of the busy lanes hold triangles
    uint32_t fetchMinIdle;          // ... refill a warp once this many lanes are idle
    uint32_t triVoteOccl32, fetchMinIdleOccl;   // the same two for k_occlude (any-hit walks: a hit ends the ray)
    uint32_t sharePrimary;          // bounce 0: the hit of path p is stored at p % (rows * W) (one primary ray per pixel)
    uint32_t connectAfterExtend;    // connect of bounce j runs after extend of bounce j+1 (A.hit is the NEXT hit)
    uint32_t useTextures;           // texture extension on (option "textures"); 0 = the HEAD shader's behaviour
    uint32_t continueSample;        // sample s > 0 of a multi-sample dispatch: generate continues the RNG stream in rayD[p].w
};

PT_HD v3 xyz(const float4& f) { return mk3(f.x, f.y, f.z); }
PT_HD float4 f4(v3 v, float w) { return make_float4(v.x, v.y, v.z, w); }

// ---- generate -------------------------------------------------------------------------------
PT_HD void stage_generate(const WfParams& P, const PathArrays& A, uint32_t p) {
    uint32_t perSlot = P.rows * (uint32_t)P.W;
    uint32_t k = p / perSlot, rem = p - k * perSlot;
    uint32_t gy = P.row0 + (rem / (uint32_t)P.W) * P.rowStride, gx = rem % (uint32_t)P.W;
    uint32_t frame = P.firstFrame + k * P.stride;
    Ray r = primary_ray(P.pc.cam, P.cp, gx, gy, P.W, P.H);
    // samples of one dispatch share the pixel's RNG stream (raytrace.comp:562-570): sample s > 0 continues from the state
    // the path of sample s-1 ended with, which every exit of stage_shade leaves in rayD[p].w
    uint32_t state = P.continueSample ? as_uint(A.rayD[p].w) : gy * (uint32_t)P.W + gx + frame_seed(frame);
    A.rayO[p] = f4(r.o, 1.f);                       // misWeight = 1
    A.rayD[p] = f4(r.d, as_float(state));
    A.atten[p] = make_float4(1.f, 1.f, 1.f, 0.f);
    A.total[p] = make_float4(0.f, 0.f, 0.f, 0.f);
    A.direct[p] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ---- extend ---------------------------------------------------------------------------------
PT_HD int lowest_bit(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
PT_HD uint32_t live_sphere_mask(uint32_t sphereCount) { return sphereCount >= 32u ? 0xffffffffu : ((1u << sphereCount) - 1u); }
// raytrace.comp:282-287: closest sphere, first of equal distances (see WideScene for the slot masks)
PT_HD void spheres_closest(const WideScene& sc, const Ray& ray, uint32_t sphereCount, Hit& h, uint32_t& key) {
    h.dst = PT_FAR; h.id = PT_HIT_MISS; h.u = 0.f; h.v = 0.f; key = 0xffffffffu;
    if (sc.nSpheres <= 32u) {
        for (uint32_t mk = sc.uniqSphereMask & live_sphere_mask(sphereCount); mk; mk &= mk - 1u) {
            uint32_t i = (uint32_t)lowest_bit(mk);
            float dst; bool front;
            if (sphere_test(sc.spheres[i], ray, dst, front) && dst < h.dst) {
                h.dst = dst; h.id = PT_HIT_SPHERE | (front ? PT_HIT_FRONT : 0u) | i; key = i;
            }
        }
    } else {
        for (uint32_t i = 0; i < sphereCount; i++) {
            float dst; bool front;
            if (sphere_test(sc.spheres[i], ray, dst, front) && dst < h.dst) {
                h.dst = dst; h.id = PT_HIT_SPHERE | (front ? PT_HIT_FRONT : 0u) | i; key = i;
            }
        }
    }
}
PT_HD Hit wide_closest_hit(const WideScene& sc, const Ray& ray, uint32_t sphereCount, uint32_t objectCount, uint32_t& contested) {
    Hit h; uint32_t key;
    spheres_closest(sc, ray, sphereCount, h, key);
    bvh8_closest(sc.all, sc.xforms, sc.up, sc.nSpheres, ray, objectCount, sc.extent, h, key, contested);
    return h;
}
// The hit of a CONTESTED ray (pt_traverse.h): calculateIntersections in the reference's own order over the reference's own
// buffers (pt_reftrace.h), expressed as the wavefront's Hit (record index of the triangle in the wide layout).
PT_HD Hit exact_closest_hit(const RefScene& ref, const WideScene& sc, const ptl::RayTracerData& td, const Ray& ray, float wideDst) {
    // wideDst: the distance the wide traversal found (the smallest candidate): phantom bound of the re-trace (pt_reftrace.h)
    // wideDst >= PT_FAR: the wide traversal's own hit was not reachable in the reference's BVH2 (k_verify_extend): no bound
    const RefHit r = ref_closest_hit(ref, td, ray, nullptr, wideDst >= PT_FAR ? PT_FAR : wideDst + 3.f * contest_window(wideDst, sc.extent));
    Hit h; h.dst = r.dst; h.u = 0.f; h.v = 0.f; h.id = PT_HIT_MISS;
    if (r.kind == 1u) h.id = PT_HIT_SPHERE | (r.front ? PT_HIT_FRONT : 0u) | r.triHitIndex;
    else if (r.kind == 2u) {
        h.id = sc.primOf[sc.instBase[r.objectHitIndex] + r.triHitIndex - sc.triFirst[r.objectHitIndex]] | (r.front ? PT_HIT_FRONT : 0u);
        h.u = r.u; h.v = r.v;
    }
    return h;
}
// Would the reference have tested the triangle of this (optimistically accepted, pt_traverse.h) closest hit? Spheres and
// misses: nothing to confirm.
// In two halves, for a kernel pair: the quick accept at full lane count, the rare walk from a compacted queue
PT_HD bool hit_reachable_quick(const WideScene& sc, const Ray& ray, const Hit& h) {
    if (h.id == PT_HIT_MISS || (h.id & PT_HIT_SPHERE)) return true;
    const PrimRecord p = load_prim(sc.all.prims, h.id & PT_HIT_IDMASK & ~PT_HIT_CONTESTED);
    ObjRayCache oc; oc.xform = 0xffffffffu;
    object_ray(oc, sc.xforms, p.xformId, ray);
    return ref_reachable_quick(p, oc, h.dst);
}
PT_HD bool hit_reachable_walk(const WideScene& sc, const Ray& ray, const Hit& h) {
    const PrimRecord p = load_prim(sc.all.prims, h.id & PT_HIT_IDMASK & ~PT_HIT_CONTESTED);
    ObjRayCache oc; oc.xform = 0xffffffffu;
    object_ray(oc, sc.xforms, p.xformId, ray);
    return ref_reachable_walk(sc.up, p, oc, h.dst);
}
PT_HD void stage_extend(const WfParams& P, const WideScene& sc, const RefScene& ref, const PathArrays& A, uint32_t p) {
    Ray r; r.o = xyz(A.rayO[p]); r.d = xyz(A.rayD[p]);
    uint32_t contested = 0u;
    Hit h = wide_closest_hit(sc, r, P.pc.params.sphereCount, P.pc.params.objectCount, contested);
    if (contested) h = exact_closest_hit(ref, sc, P.pc.params, r, h.dst);
    A.hit[p] = h;
}

// World-space hit point, shading normal and material of a Hit (raytrace.comp:218-221 for spheres;
// :247,:259,:318-319 for triangles).
PT_HD void hit_surface(const WideScene& sc, const Ray& ray, const Hit& h, v3& hp, v3& n, uint32_t& material, bool& front,
                       uint32_t& objectIndex, uint32_t& primIndex) {
    front = (h.id & PT_HIT_FRONT) != 0;
    uint32_t idx = h.id & PT_HIT_IDMASK;
    if (h.id & PT_HIT_SPHERE) {
        const ptl::Sphere s = sc.spheres[idx];
        sphere_surface(s, ray, h.dst, front, hp, n);
        material = s.materialIndex; objectIndex = 0xffffffffu; primIndex = idx;
    } else {
        const ShadeRecord s = sc.shade[idx];
        const Xform& x = sc.xforms[sc.objXform[s.objectIndex]];
        Ray orr; orr.d = mul_dir(x.Minv, ray.d); orr.o = mul_point(x.Minv, ray.o);
        tri_surface(x.M, orr, h.dst, h.u, h.v, front, mk3(s.n0x, s.n0y, s.n0z), mk3(s.n1x, s.n1y, s.n1z),
                    mk3(s.n2x, s.n2y, s.n2z), hp, n);
        material = s.materialIndex; objectIndex = s.objectIndex; primIndex = s.triIndex;
    }
}

// ---- texture extension ------------------------------------------------------------------------
// NOT in the HEAD shader (it binds TextureBuffer[64] and computes hit.uv, raytrace.comp:122,148,249-256, but never
// samples). Semantics of this project (the test checker restates them independently):
// the reference's samplers are NEAREST without mips, REPEAT (0) or CLAMP_TO_EDGE (1) (src/vk_engine.cpp:525-531),
// images are RGBA8 sRGB (src/vk_textures.cpp:212): texel = floor(uv * size) wrapped/clamped, bytes decoded to linear
// through a 256-entry table. The fetch goes through a CUDA texture object (cudaArray: 2-D tiled, texture cache)
// with point filtering and unnormalised coordinates, so the texel chosen is exactly that index.
#if defined(__CUDA_ARCH__)
PT_HD int texel_index(float coord, uint32_t size, uint32_t samplerIndex) {
    float f = floorf(coord * float(size));
    if (!(fabsf(f) < 1.0e9f)) f = 0.f;
    int i = int(f), n = int(size);
    if (samplerIndex == 0u) { i %= n; if (i < 0) i += n; }
    else { i = i < 0 ? 0 : (i > n - 1 ? n - 1 : i); }
    return i;
}
__device__ __forceinline__ bool sample_texture(const WideScene& sc, int32_t slot, float tu, float tv, uint32_t samplerIndex, v3& rgb) {
    if (slot < 0 || slot >= 64) return false;
    const TexEntry t = sc.tex[slot];
    if (t.object == 0ull) return false;
    int x = texel_index(tu, t.w, samplerIndex), y = texel_index(tv, t.h, samplerIndex);
    uchar4 px = tex2D<uchar4>((cudaTextureObject_t)t.object, float(x) + 0.5f, float(y) + 0.5f);
    rgb = mk3(sc.srgbLut[px.x], sc.srgbLut[px.y], sc.srgbLut[px.z]);
    return true;
}
// albedo map: factor on the material's albedo; metalness map: mirror BSDF where its red channel >= 0.5
__device__ __forceinline__ void apply_textures(const WideScene& sc, const ptl::Material& mat, uint32_t primIdx, float u, float v,
                                               v3& albedoTexel, float& reflectance) {
    if (mat.albedoIndex < 0 && mat.metalnessIndex < 0) return;
    const UvRecord r = sc.uv[primIdx];
    float w = 1.f - u - v;
    float tu = w * r.u0 + u * r.u1 + v * r.u2;          // hit.uv, raytrace.comp:252
    float tv = w * r.v0 + u * r.v1 + v * r.v2;
    if (r.degenerate) { tu = 0.5f; tv = 0.5f; }         // :254-256
    v3 t;
    if (sample_texture(sc, mat.albedoIndex, tu, tv, r.samplerIndex, t)) albedoTexel = t;
    if (sample_texture(sc, mat.metalnessIndex, tu, tv, r.samplerIndex, t)) reflectance = t.x >= 0.5f ? 1.f : 0.f;
}
#endif

// ---- shade ----------------------------------------------------------------------------------
#define PT_SHADE_CONTINUE 1u      /* path goes on to bounce j+1 */
#define PT_SHADE_CONNECT 2u       /* diffuse bounce: NEE/pdf queries pending for this path */

// Light-sample ray whose result cannot matter: with Afac = albedo/pi * max(0, n.l) == 0 in all three components
// (light below the surface's horizon, or a black surface) directLight = E * (Afac * g) is a zero whatever the ray
// finds, as long as g = w1/pdf and E are finite: g is (pdf == 0 ? 0 : w1/pdf) with w1 <= 1, which is finite unless
// pdf = dst^2 / (cos * 0.444) is NaN, i.e. cos == 0 — excluded by l.y != 0 (cos = -l.y, raytrace.comp:399) — and
// E is finite when every material's emission is (WideScene::emissionFinite). The sign of that zero can differ
// from the traced result; it is only ever added to totalColor, which is never -0 (it starts at +0).
// About half of the NEE rays of an interior scene are of this kind; they are neither searched nor traced.
PT_HD bool nee_cannot_contribute(const WideScene& sc, const float4& qa, v3 lightSample) {
    return sc.emissionFinite && qa.x == 0.f && qa.y == 0.f && qa.z == 0.f && lightSample.y != 0.f;
}
// The second lightSamplePDF call of diffuseBRDF (raytrace.comp:453, the BSDF-sampled direction) only feeds
// misWeight, and misWeight is only read at the NEXT hit as emission / misWeight, which trace() uses only when
// directLight.x == -1 (raytrace.comp:499-500; at j == 0 misWeight is still 1). After a diffuse bounce directLight is
// the NEE result, so that pdf ray is dead unless the NEE result is EXACTLY -1.0f. It is therefore evaluated lazily:
// connect_job_finish reports that rare case and connect_mis_weight_exact() then runs both parts of the query.
PT_HD uint32_t stage_shade(const WfParams& P, const WideScene& sc, const PathArrays& A, uint32_t p, uint32_t j) {
    const ptl::RayTracerData& td = P.pc.params;
    float4 ro = A.rayO[p], rd = A.rayD[p];
    Ray ray; ray.o = xyz(ro); ray.d = xyz(rd);
    float misWeight = ro.w;
    uint32_t state = as_uint(rd.w);
    v3 atten = xyz(A.atten[p]), direct = xyz(A.direct[p]);
    const Hit h = A.hit[(j == 0 && P.sharePrimary) ? p % (P.rows * (uint32_t)P.W) : p];

    if (h.id == PT_HIT_MISS) {                                   // raytrace.comp:531-534
        v3 total = xyz(A.total[p]) + atten * environment_light(P.pc.env, ray.d);
        A.total[p] = f4(total, 0.f);
        return 0u;
    }
    v3 hp, n; uint32_t matIdx, objIdx, primIdx; bool front;
    hit_surface(sc, ray, h, hp, n, matIdx, front, objIdx, primIdx);
    const ptl::Material mat = sc.materials[matIdx];

    v3 emission = mk3(mat.emissionColor[0], mat.emissionColor[1], mat.emissionColor[2]) * mat.emissionStrength / misWeight;
    v3 finalLight = direct.x == -1.f ? emission : direct;        // :499-500
    // totalColor is read and written only when this hit changes it. Adding a zero leaves it as it is (it is never -0:
    // it starts at +0 and only grows), and it already passed the guard of :505 at the bounce before; a NaN term is not
    // a zero and takes the full path. Most bounces of an interior add nothing (no direct light, no emission).
    const v3 gain = finalLight * atten;
    if (j == 0 || !(gain.x == 0.f && gain.y == 0.f && gain.z == 0.f)) {
        v3 total = xyz(A.total[p]) + gain;
        if (j == 0) total = total + emission;
        if (isnan3(total) || total.x < 0.f || total.y < 0.f || total.z < 0.f) {   // :505 — the whole path returns 0
            A.total[p] = make_float4(0.f, 0.f, 0.f, 0.f);
            return 0u;
        }
        A.total[p] = f4(total, 0.f);
    }

    v3 dirOut, radiance;
    float originSign = 1.f, nextMis = 1.f;
    uint32_t flags = 0u;
    float reflectance = mat.reflectance;
    v3 albedoTexel = splat3(1.f);
#if defined(__CUDA_ARCH__)
    if (P.useTextures && !(h.id & PT_HIT_SPHERE)) apply_textures(sc, mat, h.id & PT_HIT_IDMASK, h.u, h.v, albedoTexel, reflectance);
#endif
    if (reflectance != 0.f) {                                    // :509-510, specularBRDF :466-469
        dirOut = reflect(ray.d, n); radiance = splat3(1.f); direct = splat3(-1.f);
    } else if (mat.ior != -1.f) {                                // :511-512, dielectricBTDF :471-481
        dirOut = dielectric_dir(ray.d, n, front, mat.ior, state, originSign);
        radiance = splat3(1.f); direct = splat3(-1.f);
    } else {                                                     // :513-514, diffuseBRDF :430-464 (sampling half)
        v3 albedo = mk3(mat.albedo[0], mat.albedo[1], mat.albedo[2]);
        if (P.useTextures) albedo = albedo * albedoTexel;
        v3 origin = hp + n * 0.01f;
        v3 lightSample = light_sample_dir(origin, state);
        v3 cosineSample = cosine_hemisphere_dir(n, state);
        float cosinePDF = cosine_hemisphere_pdf(n, lightSample);
        float realCosinePDF = cosine_hemisphere_pdf(n, cosineSample);
        radiance = albedo * PT_INV_PI * dot(n, cosineSample) / realCosinePDF;
        v3 Afac = albedo * PT_INV_PI * gmax(0.f, dot(n, lightSample));
        dirOut = cosineSample;
        const float4 qa = f4(Afac, 0.f);
        if (nee_cannot_contribute(sc, qa, lightSample)) {
            direct = splat3(0.f) * (Afac * 0.f);                 // what connect would store: E = 0 times (Afac * 0); no query
        } else {
            A.qO[p] = f4(origin, cosinePDF);
            A.qL[p] = f4(lightSample, realCosinePDF);
            A.qA[p] = qa;
            // what diffuseBRDF yields when the light-sample ray's closest hit is not emissive (E = 0, pdf = 0,
            // raytrace.comp:459-460); connect overwrites it only for an unoccluded emissive hit
            direct = splat3(0.f) * (Afac * 0.f);
            flags |= PT_SHADE_CONNECT;
        }
    }
    atten = atten * radiance;

    float rr = gmax(gmax(atten.x, atten.y), atten.z);            // :519-524
    rr = gmin(rr, 0.95f);
    rr = j <= 5 ? 1.f : rr;
    const bool stop = rng_next(state) > rr;                      // break: pending direct light is dropped
    if (stop || j >= td.bounceLimit) {                           // or the loop ends: nothing computed now is consumed,
        reinterpret_cast<float*>(&A.rayD[p])[3] = as_float(state);   // except the RNG state (next sample of the dispatch)
        return 0u;
    }
    atten = atten * (1.f / rr);

    A.atten[p] = f4(atten, 0.f);
    A.direct[p] = f4(direct, 0.f);
    A.rayO[p] = f4(hp + n * originSign * 0.00001f, nextMis);     // :527-529 (misWeight overwritten by connect for diffuse)
    A.rayD[p] = f4(dirOut, as_float(state));
    return flags | PT_SHADE_CONTINUE;
}

// ---- connect --------------------------------------------------------------------------------
// lightSamplePDF() (raytrace.comp:389-403) for (origin, dir) is nonzero only when the CLOSEST hit of
// the ray is emissive. It is evaluated in three parts:
//   prep    : closest hit among EMISSIVE geometry only (spheres + the small emissive BVH), and the
//             non-emissive spheres that could hide it -> candidate distance tE and loop-order key
//   occlude : is there a non-emissive triangle the reference would have preferred? (bounded any-hit
//             walk of the full BVH, the expensive part: its own persistent kernel, wavefront.cu)
//   finish  : pdfs, MIS weights, direct light (raytrace.comp:447-460)
// qE[p] = { tE(lightSample), material of that emissive hit, loop-order key of it, scale of the contest window }, tE < 0: no emissive closest hit
// (none found, or occluded). Those paths keep the directLight shade stored (a zero: E = 0, pdf = 0); finish only reads
// their qE. (Finishing the unoccluded ones inside the occlusion kernel, by the lane that ends the walk, was measured:
// k_connect_finish 3.7 -> 0 ms but k_occlude 35.3 -> 39.7 ms per step — four dependent loads in a converged warp.)
#define PT_NO_EMISSIVE (-1.f)

// Conservative slab test against the bounding box of all emissive triangles (the padded world boxes the small BVH
// was built from, grown once more by 1e-4 relative at upload: far above this test's rounding). Most BSDF-sampled
// rays miss it and skip the call into the traversal altogether.
PT_HD bool ray_meets_box(const Ray& r, const float* lo, const float* hi) {
    BoxRay b = make_box_ray(r);
    float t0x = (lo[0] - b.o.x) * b.rd.x, t1x = (hi[0] - b.o.x) * b.rd.x;
    float t0y = (lo[1] - b.o.y) * b.rd.y, t1y = (hi[1] - b.o.y) * b.rd.y;
    float t0z = (lo[2] - b.o.z) * b.rd.z, t1z = (hi[2] - b.o.z) * b.rd.z;
    float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), 0.f));
    float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    return tn <= tf;
}
// Sphere slots are split at upload into emissive / non-emissive bit masks (WideScene), so the two
// loops below touch only the spheres that matter; with more than 32 slots every slot is visited.
// contested: see pt_traverse.h ("exactness against the reference's traversal order") — two emissive candidates, or
// the emissive candidate and a sphere, closer together than the slack: the caller asks the reference-order search.
PT_HD float emissive_candidate(const WideScene& sc, v3 origin, v3 dir, uint32_t sphereCount, uint32_t objectCount, uint32_t& keyE,
                               uint32_t& matOut, uint32_t& contested, float& windowScale) {
    matOut = 0u; keyE = 0xffffffffu; windowScale = 1.f;
    if (!sc.anyEmissive) return PT_NO_EMISSIVE;
    Ray r; r.o = origin; r.d = dir;
    Hit e; e.dst = PT_FAR; e.id = PT_HIT_MISS; e.u = 0.f; e.v = 0.f;
    const bool masks = sc.nSpheres <= 32u;
    const uint32_t live = live_sphere_mask(sphereCount);
    if (masks) {
        for (uint32_t mk = sc.emisSphereMask & live; mk; mk &= mk - 1u) {
            uint32_t i = (uint32_t)lowest_bit(mk);
            const ptl::Sphere s = sc.spheres[i];
            float dst; bool front;
            if (sphere_test(s, r, dst, front) && dst < e.dst) { e.dst = dst; e.id = PT_HIT_SPHERE | i; keyE = i; matOut = s.materialIndex; }
        }
    } else {
        for (uint32_t i = 0; i < sphereCount; i++) {
            const ptl::Sphere s = sc.spheres[i];
            if (sc.materials[s.materialIndex].emissionStrength == 0.f) continue;
            float dst; bool front;
            if (sphere_test(s, r, dst, front) && dst < e.dst) { e.dst = dst; e.id = PT_HIT_SPHERE | i; keyE = i; matOut = s.materialIndex; }
        }
    }
    uint32_t before = e.id;
    if (sc.emis.nPrims && ray_meets_box(r, sc.emisLo, sc.emisHi)) {
        float g = 0.f;
        bvh8_closest(sc.emis, sc.xforms, sc.up, sc.nSpheres, r, objectCount, sc.extent, e, keyE, contested, &g);
        if (e.id != before) windowScale = fmaxf(1.f, g * (1.f / 16.f));     // the window grows with 1 / cos of the emissive hit
    }
    if (e.id == PT_HIT_MISS) return PT_NO_EMISSIVE;
    if (e.id != before) matOut = sc.shadeEmis[e.id & PT_HIT_IDMASK].materialIndex;
    const float slackE = (e.id & PT_HIT_SPHERE) ? -1.f : contest_window(e.dst, sc.extent) * windowScale;      // a sphere is never pruned by the reference
    if (masks) {
        for (uint32_t mk = sc.plainSphereMask & live; mk; mk &= mk - 1u) {
            uint32_t i = (uint32_t)lowest_bit(mk);
            float dst; bool front;
            if (sphere_test(sc.spheres[i], r, dst, front)) {
                if (fabsf(dst - e.dst) <= slackE) contested = 1u;     // the reference tests spheres first: the triangle may be pruned against it
                if (dst < e.dst || (dst == e.dst && i < keyE)) return PT_NO_EMISSIVE;
            }
        }
    } else {
        for (uint32_t i = 0; i < sphereCount; i++) {
            const ptl::Sphere s = sc.spheres[i];
            if (!(sc.materials[s.materialIndex].emissionStrength == 0.f)) continue;
            float dst; bool front;
            if (sphere_test(s, r, dst, front)) {
                if (fabsf(dst - e.dst) <= slackE) contested = 1u;
                if (dst < e.dst || (dst == e.dst && i < keyE)) return PT_NO_EMISSIVE;
            }
        }
    }
    return e.dst;
}

#define PT_CONNECT_JOB_LIGHT 1u     /* occlusion query pending for the light-sample ray */
#define PT_CONNECT_JOB_EXACT 2u     /* contested: the reference-order search answers (connect_exact) */
PT_HD uint32_t stage_connect_prep(const WfParams& P, const WideScene& sc, const PathArrays& A, uint32_t p) {
    const ptl::RayTracerData& td = P.pc.params;
    uint32_t keyL = 0xffffffffu, matL = 0u, contested = 0u;
    float tL = PT_NO_EMISSIVE, scaleL = 1.f;
    // connectAfterExtend: bounce j+1 was already extended; directLight of bounce j is dropped when it missed (:531-534)
    bool dropped = P.connectAfterExtend && A.hit[p].id == PT_HIT_MISS;
    if (!dropped)            // (rays that cannot contribute never get here: stage_shade)
        tL = emissive_candidate(sc, xyz(A.qO[p]), xyz(A.qL[p]), td.sphereCount, td.objectCount, keyL, matL, contested, scaleL);  // :443-447
    A.qE[p] = make_float4(tL, as_float(matL), as_float(keyL), scaleL);
    if (contested) return PT_CONNECT_JOB_EXACT;
    return tL >= 0.f ? PT_CONNECT_JOB_LIGHT : 0u;
}
// the occlusion query of path p's light-sample ray
PT_HD void connect_job_ray(const PathArrays& A, uint32_t p, Ray& r, float& tE, uint32_t& keyE, float& windowScale) {
    float4 e = A.qE[p];
    r.o = xyz(A.qO[p]);
    r.d = xyz(A.qL[p]);
    tE = e.x;
    keyE = as_uint(e.z);
    windowScale = e.w;
}
// A contested light-sample ray: lightSamplePDF's closest-hit search (raytrace.comp:396-397) in the reference's own order.
PT_HD void connect_exact(const RefScene& ref, const WideScene& sc, const ptl::RayTracerData& td, const PathArrays& A, uint32_t p) {
    Ray r; r.o = xyz(A.qO[p]); r.d = xyz(A.qL[p]);
    // phantom bound (pt_reftrace.h): nothing beyond the emissive candidate's window can be the closest hit of this ray
    const float4 qe = A.qE[p];
    const float bound = qe.x >= 0.f ? qe.x + 3.f * contest_window(qe.x, sc.extent) * fmaxf(qe.w, 1.f) : PT_FAR;
    const RefHit h = ref_closest_hit(ref, td, r, nullptr, bound);
    const bool emissive = h.kind != 0u && !(ref.materials[h.materialIndex].emissionStrength == 0.f);
    A.qE[p] = make_float4(emissive ? h.dst : PT_NO_EMISSIVE, as_float(h.materialIndex), 0.f, 0.f);
}
PT_HD void connect_job_occluded(const PathArrays& A, uint32_t p) {
    reinterpret_cast<float*>(&A.qE[p])[0] = PT_NO_EMISSIVE;
}
// k_occlude<*, true>: occluded by a triangle that is not yet confirmed reachable (pt_traverse.h) — the record index and its
// distance ride in qE.y / qE.z (dead once qE.x says "no emissive hit") until k_verify_occluded has looked at them
PT_HD void connect_job_occluded_by(const PathArrays& A, uint32_t p, uint32_t prim, float dst) {
    float4 e = A.qE[p];
    e.x = PT_NO_EMISSIVE; e.y = as_float(prim); e.z = dst;
    A.qE[p] = e;
}
// -> false when the reference would not have tested that occluder: the query has to be answered by connect_exact
PT_HD bool connect_job_occluder_reachable(const WideScene& sc, const PathArrays& A, uint32_t p) {
    const float4 e = A.qE[p];
    if (e.x >= 0.f) return true;                       // not occluded (or contested: k_exact_nee answers)
    const PrimRecord pr = load_prim(sc.all.prims, as_uint(e.y));
    Ray r; r.o = xyz(A.qO[p]); r.d = xyz(A.qL[p]);
    ObjRayCache oc; oc.xform = 0xffffffffu;
    object_ray(oc, sc.xforms, pr.xformId, r);
    return ref_reachable(sc.up, pr, oc, e.z);
}
// The light-sample ray of path p: when its closest hit is the emissive one at distance qE.x (found by prep, not occluded)
// — pdf, power-heuristic weight and direct light of diffuseBRDF (raytrace.comp:447-460); otherwise nothing (see above).
// Returns true when misWeight is live (directLight.x == -1 exactly, see stage_shade) and has to be computed.
PT_HD bool connect_job_finish(const WideScene& sc, const PathArrays& A, uint32_t p) {
    const float4 qe = A.qE[p];
    if (!(qe.x >= 0.f)) return false;
    float4 qo = A.qO[p], ql = A.qL[p], qa = A.qA[p];
    v3 lightSample = xyz(ql);
    float cosinePDF = qo.w;
    float realLightPDF = light_pdf_from_dst(qe.x, lightSample);                                            // :447
    float w1 = mis_weight(realLightPDF, cosinePDF);                                                        // :449-450
    const ptl::Material lm = sc.materials[as_uint(qe.y)];
    v3 E = mk3(lm.emissionColor[0], lm.emissionColor[1], lm.emissionColor[2]) * lm.emissionStrength;
    v3 direct = E * (xyz(qa) * (realLightPDF == 0.f ? 0.f : w1 / realLightPDF));                            // :459-460
    A.direct[p] = f4(direct, 0.f);
    return direct.x == -1.f;
}
PT_HD_NOINLINE void connect_mis_weight_exact(const RefScene& ref, const ptl::RayTracerData& td, const PathArrays& A, uint32_t p) {
    v3 origin = xyz(A.qO[p]), cosineSample = xyz(A.rayD[p]);
    float realCosinePDF = A.qL[p].w;
    float lightPDF = ref_light_pdf(ref, td, origin, cosineSample);                                          // :453, in the reference's order
    float4 ro = A.rayO[p];
    ro.w = mis_weight(realCosinePDF, lightPDF);                                                            // :455-456
    A.rayO[p] = ro;
}
// all parts for one path (host emulation; the kernels run them as separate launches)
PT_HD void stage_connect(const WfParams& P, const WideScene& sc, const RefScene& ref, const PathArrays& A, uint32_t p) {
    uint32_t job = stage_connect_prep(P, sc, A, p);
    if (job == PT_CONNECT_JOB_LIGHT) {
        Ray r; float tE, ws; uint32_t keyE;
        connect_job_ray(A, p, r, tE, keyE, ws);
        const int o = bvh8_occluded(sc.all, sc.xforms, sc.up, sc.nSpheres, r, P.pc.params.objectCount, sc.extent, ws, tE, keyE);
        if (o == PT_OCCL_YES) connect_job_occluded(A, p);
        else if (o == PT_OCCL_CONTESTED) job = PT_CONNECT_JOB_EXACT;
    }
    if (job == PT_CONNECT_JOB_EXACT) connect_exact(ref, sc, P.pc.params, A, p);
    if (connect_job_finish(sc, A, p)) connect_mis_weight_exact(ref, P.pc.params, A, p);
}

// ---- accumulate -----------------------------------------------------------------------------
// outColor of each frame slot added in frame order (raytrace.comp:568-581: mean over 1 sample,
// NaN/Inf -> magenta); accum.a counts frames.
PT_HD void stage_accumulate(const WfParams& P, const PathArrays& A, float4* accum, uint32_t pix) {
    uint32_t perSlot = P.rows * (uint32_t)P.W;
    uint32_t gy = P.row0 + (pix / (uint32_t)P.W) * P.rowStride, gx = pix % (uint32_t)P.W;
    size_t g = (size_t)gy * P.W + gx;
    float4 a = accum[g];
    for (uint32_t k = 0; k < P.nSlots; ++k) {
        v3 c = xyz(A.total[k * perSlot + pix]);
        if (isnan3(c) || isinf3(c)) c = mk3(1.f, 0.f, 1.f);
        a.x += c.x; a.y += c.y; a.z += c.z; a.w += 1.f;
    }
    accum[g] = a;
}

}  // namespace pt
