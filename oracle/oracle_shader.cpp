// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Function-for-function CPU restatement of shaders/raytrace.comp. Names, statement
// order and float literal types follow the shader; each function cites its lines.
// Compile with -O2 -ffp-contract=off (no FMA contraction, no fast-math).
#include "oracle_shader.h"

// Optional "portable" transcendental mode (orc::set_math_mode(1)): sin/cos/pow/tan come from the
// backend's own double-precision polynomials instead of libm, so that a bit-for-bit image
// comparison isolates everything else (traversal, RNG stream, BSDF operation order). The default
// (mode 0) is libm, and the tolerance tests use that. The header is plain C++, no CUDA.
#include "../ray_tracer_b200/csrc/pt_portable_math.h"

#include <cmath>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {
using namespace om;

static int g_math_mode = 0;
void set_math_mode(int mode) { g_math_mode = mode; }
int math_mode() { return g_math_mode; }

namespace {

inline float o_pow(float x, float y) { return g_math_mode ? pt::pt_powf(x, y) : std::pow(x, y); }
inline float o_sin(float x) { return g_math_mode ? pt::pt_sinf(x) : std::sin(x); }
inline float o_cos(float x) { return g_math_mode ? pt::pt_cosf(x) : std::cos(x); }
inline float o_tan(float x) { return g_math_mode ? pt::pt_tanf(x) : std::tan(x); }

const float PI = 3.1415926535897932384f;   // raytrace.comp:6
const float INV_PI = 0.3183098862f;        // raytrace.comp:7

struct Ray {            // raytrace.comp:17-22 (dimSign is written but never read)
    vec3 origin, dir, invDir;
};

struct HitInfo {        // raytrace.comp:72-82
    vec3 hitPoint, normal;
    vec2 uv{0.f, 0.f};
    float dst = 0.f;
    uint32_t objectHitIndex = 0xffffffffu;
    uint32_t triHitIndex = 0;
    uint32_t materialIndex = 0;   // A-Q9: uninitialised in the shader, pinned to 0
    bool didHit = false;
    bool frontFace = false;
    uint32_t kind = 0;            // oracle bookkeeping only: 1 sphere, 2 triangle
    vec3 albedoTexel{1.f, 1.f, 1.f};   // texture extension: factor on the material's albedo (1 without a map)
};

struct BxDFResult {     // raytrace.comp:112-118
    vec3 sampledDir, radiance, directLight;
    float originSign, cosineMisWeight;
};

struct Ctx {
    const Scene& sc;
    const ol::PushConstants& pc;
    Work work;      // all rays
    Work ext;       // extension rays only (the calls from trace(), raytrace.comp:493): roofline of the extend stage
    // Tie detector (tests only, pixel_ties()): counts rays whose closest hit was met at EXACTLY the same distance on two
    // triangles of one object — the one documented class where a traversal in another order may keep the other triangle
    // (the reference keeps the first in its own loop order, raytrace.comp:316). Does not change any result.
    bool detectTies = false;
    uint32_t ties = 0;
    // Ray recorder (tests only, pixel_rays()): every calculateIntersections call of one pixel's path with its result.
    std::vector<float>* recRays = nullptr;      // 8 floats per ray
    std::vector<HitOut>* recHits = nullptr;
};

vec3 v3(const float* p) { return vec3(p[0], p[1], p[2]); }

}  // namespace

// raytrace.comp:158-163. 4294967295.f is not representable; the literal is 2^32.
float random(uint32_t& state) {
    state = state * 747796405u + 2891336453u;
    uint32_t result = ((state >> ((state >> 28) + 4u)) ^ state) * 277803737u;
    result = (result >> 22) ^ result;
    return float(result) / 4294967295.f;
}

// raytrace.comp:562-563. The int literal 23892183 converts to the float 23892184.
uint32_t starting_seed(uint32_t frameCount) {
    uint32_t lol = frameCount;
    return uint32_t(random(lol) * float(23892183));
}

namespace {

// raytrace.comp:177-181
float schlick(float cosine, float refraction_index) {
    float r0 = (1.f - refraction_index) / (1.f + refraction_index);
    r0 = r0 * r0;
    return r0 + (1.f - r0) * o_pow((1.f - cosine), 5.f);
}

// raytrace.comp:195-224
HitInfo sphereIntersection(const ol::Sphere& sphere, const Ray& ray) {
    HitInfo hitInfo;
    hitInfo.didHit = false;

    vec3 pos = v3(sphere.position);
    vec3 oc = pos - ray.origin;
    float a = dot(ray.dir, ray.dir);
    float b = dot(oc, ray.dir);
    float c = dot(oc, oc) - sphere.radius * sphere.radius;
    float discriminant = b * b - a * c;

    if (discriminant >= 0.f) {
        float sqrtd = std::sqrt(discriminant);
        float dst = (b - sqrtd) / a;
        hitInfo.frontFace = true;
        if (dst < 0.f) {
            dst = (b + sqrtd) / a;
            hitInfo.frontFace = false;
            if (dst < 0.f) {
                return hitInfo;
            }
        }

        hitInfo.didHit = true;
        hitInfo.dst = dst;
        hitInfo.hitPoint = ray.origin + ray.dir * dst;
        hitInfo.normal = normalize(hitInfo.hitPoint - pos) * (hitInfo.frontFace ? 1.f : -1.f);
        hitInfo.materialIndex = sphere.materialIndex;
    }
    return hitInfo;
}

// raytrace.comp:227-261. smoothShade is ignored by the shader (`rachit` is always true).
HitInfo triangleIntersection(const Ray& ray, const ol::TrianglePoint& v0, const ol::TrianglePoint& v1,
                             const ol::TrianglePoint& v2, bool /*smoothShade*/, bool frontOnly) {
    vec3 p0 = v3(v0.position), p1 = v3(v1.position), p2 = v3(v2.position);
    vec3 v1v0 = p1 - p0;
    vec3 v2v0 = p2 - p0;
    vec3 rov0 = ray.origin - p0;
    vec3 n = cross(v1v0, v2v0);

    vec3 q = cross(rov0, ray.dir);
    float d0 = -dot(ray.dir, n);
    float d = 1.f / d0;

    float dst = dot(rov0, n) * d;
    float u = dot(v2v0, q) * d;
    float v = -dot(v1v0, q) * d;
    float w = 1.f - u - v;

    HitInfo hit;
    hit.frontFace = d0 >= 0.00000001f;
    hit.didHit = dst >= 0.f && u >= 0.f && v >= 0.f && w >= 0.f && !(!hit.frontFace && frontOnly);
    hit.hitPoint = ray.origin + ray.dir * dst;
    hit.dst = dst;

    vec2 v0uv{v0.position[3], v0.normal[3]};
    vec2 v1uv{v1.position[3], v1.normal[3]};
    vec2 v2uv{v2.position[3], v2.normal[3]};
    hit.uv = vec2{w * v0uv.x + u * v1uv.x + v * v2uv.x, w * v0uv.y + u * v1uv.y + v * v2uv.y};
    auto eq = [](const vec2& a, const vec2& b) { return a.x == b.x && a.y == b.y; };
    if (eq(v0uv, v1uv) || eq(v1uv, v2uv) || eq(v2uv, v0uv)) {
        hit.uv = vec2{0.5f, 0.5f};
    }

    hit.normal = (w * v3(v0.normal) + u * v3(v1.normal) + v * v3(v2.normal)) * (hit.frontFace ? 1.f : -1.f);
    return hit;
}

// raytrace.comp:263-274
float boxIntersection(const vec3& bmin, const vec3& bmax, const Ray& ray) {
    vec3 tMin = (bmin - ray.origin) * ray.invDir;
    vec3 tMax = (bmax - ray.origin) * ray.invDir;
    vec3 t1 = gmin(tMin, tMax);
    vec3 t2 = gmax(tMin, tMax);
    float tNear = gmax(gmax(t1.x, t1.y), t1.z);
    float tFar = gmin(gmin(t2.x, t2.y), t2.z);

    bool hit = tFar >= tNear && tFar > 0.f;
    float dst = hit ? (tNear > 0.f ? tNear : 0.f) : 99999999.f;
    return dst;
}

// raytrace.comp:276-353
HitInfo calculateIntersections(Ctx& cx, const Ray& ray, float* stats) {
    const Scene& sc = cx.sc;
    HitInfo closestHit;
    closestHit.didHit = false;
    closestHit.dst = 99999999.f;
    const ol::RayTracerData& traceData = cx.pc.rayTraceParams;
    const Work before = cx.work;
    cx.work.n_ray++;
    float tieDst = -1.f;

    for (uint32_t i = 0; i < traceData.sphereCount; i++) {
        cx.work.n_sph++;
        HitInfo hitInfo = sphereIntersection(sc.spheres[i], ray);
        if (hitInfo.didHit && hitInfo.dst < closestHit.dst) {
            closestHit = hitInfo;
            closestHit.kind = 1;
            closestHit.triHitIndex = i;
        }
    }

    for (uint32_t i = 0; i < traceData.objectCount; i++) {
        cx.work.n_obj++;
        const ol::RenderObject& object = sc.objects[i];
        Ray transformRay;
        transformRay.dir = (sc.Minv[i] * vec4(ray.dir, 0.f)).xyz();
        transformRay.origin = (sc.Minv[i] * vec4(ray.origin, 1.f)).xyz();
        transformRay.invDir = vec3(1.f / transformRay.dir.x, 1.f / transformRay.dir.y, 1.f / transformRay.dir.z);

        // bvh traversal (A-Q16: the shader's stack has 64 slots; the oracle sizes it for depth 64)
        uint32_t stack[72];
        uint32_t stackIndex = 1;
        stack[0] = object.bvhIndex;
        while (stackIndex > 0) {
            const ol::BVHNode& currentNode = sc.nodes[stack[--stackIndex]];

            if (currentNode.triCount != 0) {
                if (stats) stats[1] += float(currentNode.triCount);
                cx.work.n_tri += currentNode.triCount;
                for (uint32_t j = currentNode.index; j < currentNode.index + currentNode.triCount; j++) {
                    const ol::Triangle& tri = sc.triangles[j];
                    HitInfo hitInfo = triangleIntersection(transformRay, sc.points[tri.v0], sc.points[tri.v1],
                                                           sc.points[tri.v2], object.smoothShade != 0,
                                                           tri.frontOnly != 0);
                    hitInfo.materialIndex = object.materialIndex;
                    if (cx.detectTies && hitInfo.didHit && closestHit.didHit && closestHit.kind == 2 && closestHit.objectHitIndex == i &&
                        hitInfo.dst == closestHit.dst)
                        tieDst = hitInfo.dst;

                    if (hitInfo.didHit && hitInfo.dst < closestHit.dst) {
                        closestHit = hitInfo;
                        closestHit.normal = normalize(sc.M[i] * vec4(closestHit.normal, 0.f)).xyz();
                        closestHit.hitPoint = (sc.M[i] * vec4(closestHit.hitPoint, 1.f)).xyz();
                        closestHit.triHitIndex = j;
                        closestHit.objectHitIndex = i;
                        closestHit.kind = 2;
                    }
                }
            } else {
                const ol::BVHNode& child1 = sc.nodes[currentNode.index];
                const ol::BVHNode& child2 = sc.nodes[currentNode.index + 1];

                float dst1 = boxIntersection(vec3(child1.boundsX[0], child1.boundsY[0], child1.boundsZ[0]),
                                             vec3(child1.boundsX[1], child1.boundsY[1], child1.boundsZ[1]),
                                             transformRay);
                float dst2 = boxIntersection(vec3(child2.boundsX[0], child2.boundsY[0], child2.boundsZ[0]),
                                             vec3(child2.boundsX[1], child2.boundsY[1], child2.boundsZ[1]),
                                             transformRay);
                if (stats) stats[0] += 2.f;
                cx.work.n_box += 2;

                bool isNearestA = dst1 <= dst2;
                float dstNear = isNearestA ? dst1 : dst2;
                float dstFar = isNearestA ? dst2 : dst1;
                uint32_t childIndexNear = isNearestA ? currentNode.index : currentNode.index + 1;
                uint32_t childIndexFar = isNearestA ? currentNode.index + 1 : currentNode.index;

                if (dstFar < closestHit.dst) stack[stackIndex++] = childIndexFar;
                if (dstNear < closestHit.dst) stack[stackIndex++] = childIndexNear;
            }
        }
    }

    if (closestHit.didHit) cx.work.n_hit++;
    if (cx.detectTies && closestHit.didHit && closestHit.kind == 2 && tieDst == closestHit.dst) cx.ties++;
    if (cx.recRays) {
        const float r8[8] = {ray.origin.x, ray.origin.y, ray.origin.z, 0.f, ray.dir.x, ray.dir.y, ray.dir.z, 0.f};
        cx.recRays->insert(cx.recRays->end(), r8, r8 + 8);
        HitOut o;
        std::memset(&o, 0, sizeof(o));
        o.dst = closestHit.dst;
        o.kind = closestHit.didHit ? closestHit.kind : 0;
        o.objectHitIndex = closestHit.didHit ? closestHit.objectHitIndex : 0xffffffffu;
        o.triHitIndex = closestHit.didHit ? closestHit.triHitIndex : 0;
        o.materialIndex = closestHit.materialIndex;
        o.frontFace = closestHit.didHit ? (closestHit.frontFace ? 1u : 0u) : 0u;
        cx.recHits->push_back(o);
    }
    if (stats) {
        cx.ext.n_box += cx.work.n_box - before.n_box; cx.ext.n_tri += cx.work.n_tri - before.n_tri;
        cx.ext.n_obj += cx.work.n_obj - before.n_obj; cx.ext.n_sph += cx.work.n_sph - before.n_sph;
        cx.ext.n_hit += cx.work.n_hit - before.n_hit; cx.ext.n_ray += 1;
    }
    return closestHit;
}

// raytrace.comp:356-365
vec3 getEnvironmentLight(Ctx& cx, const Ray& ray) {
    const ol::EnvironmentData& env = cx.pc.environment;
    vec3 horizon = v3(env.horizonColor), zenith = v3(env.zenithColor), ground = v3(env.groundColor);
    vec3 lightDir = v3(env.lightDir);
    float skyGradientT = o_pow(smoothstep(0.f, 0.4f, -ray.dir.y), 0.35f);
    vec3 skyGradient = mix(horizon, zenith, skyGradientT);
    float sun = o_pow(gmax(0.f, dot(ray.dir, -lightDir)), env.horizonColor[3]) * env.zenithColor[3];

    float groundToSkyT = smoothstep(-0.01f, 0.f, -ray.dir.y);
    float sunMask = float(groundToSkyT >= 1.f);
    if (env.lightDir[3] == 1.f) {
        float s = sun * sunMask;
        vec3 m = mix(ground, skyGradient, groundToSkyT);
        return vec3(m.x + s, m.y + s, m.z + s);
    }
    return vec3(0.f);
}

// raytrace.comp:368-387
vec3 lightSampleDir(const vec3& rayOrigin, uint32_t& state) {
    float x = random(state);
    float z = random(state);
    float randomX = mix(-0.33333f, 0.33333f, x);
    float randomZ = mix(-0.33333f, 0.33333f, z);
    vec3 randomPoint = vec3(randomX, -1.5f, randomZ);

    vec3 pointVector = randomPoint - rayOrigin;
    return normalize(pointVector);
}

// raytrace.comp:389-403
float lightSamplePDF(Ctx& cx, const vec3& rayOrigin, const vec3& /*rayNormal*/, const vec3& direction) {
    Ray ray;
    ray.origin = rayOrigin;
    ray.dir = direction;
    HitInfo hit = calculateIntersections(cx, ray, nullptr);
    if (!hit.didHit || cx.sc.materials[hit.materialIndex].emissionStrength == 0.f) return 0.f;

    float sqRadius = hit.dst * hit.dst;
    float cosTheta = dot(vec3(0.f, -1.f, 0.f), direction);
    float pdf = sqRadius / (cosTheta * 0.4444444f);
    return pdf;
}

// raytrace.comp:405-424
vec3 cosineHemisphereDir(const vec3& rayNormal, uint32_t& state) {
    float r1 = random(state);
    float r2 = random(state);

    float phi = 2.f * PI * r1;
    float sqrtR2 = std::sqrt(r2);
    float x = o_cos(phi) * sqrtR2;
    float y = o_sin(phi) * sqrtR2;
    float z = std::sqrt(1.f - r2);

    vec3 zAlignedDir = vec3(x, y, z);
    vec3 nonParallelAxis = std::fabs(dot(rayNormal, vec3(1.f, 0.f, 0.f))) < 1.f ? vec3(1.f, 0.f, 0.f) : vec3(0.f, 0.f, 1.f);
    vec3 t = normalize(cross(rayNormal, nonParallelAxis));
    vec3 b = cross(rayNormal, t);
    mat3 changeOfBasis;
    changeOfBasis[0] = t; changeOfBasis[1] = b; changeOfBasis[2] = rayNormal;

    return changeOfBasis * zAlignedDir;
}

// raytrace.comp:426-428
float cosineHemispherePDF(const vec3& rayNormal, const vec3& direction) {
    return gmax(0.f, dot(direction, rayNormal) * INV_PI);
}

// raytrace.comp:430-464
BxDFResult diffuseBRDF(Ctx& cx, const vec3& /*incomingDir*/, const HitInfo& prevHit, uint32_t& state) {
    const ol::RayMaterial& hitMaterial = cx.sc.materials[prevHit.materialIndex];
    vec3 albedo = v3(hitMaterial.albedo) * prevHit.albedoTexel;      // albedoTexel is 1 unless the texture extension is on
    vec3 origin = prevHit.hitPoint + prevHit.normal * 0.01f;

    vec3 lightSample = lightSampleDir(origin, state);
    vec3 cosineSample = cosineHemisphereDir(prevHit.normal, state);

    // next event estimation
    Ray lightRay;
    lightRay.origin = origin;
    lightRay.dir = lightSample;
    HitInfo lightHit = calculateIntersections(cx, lightRay, nullptr);
    const ol::RayMaterial& lightMaterial = cx.sc.materials[lightHit.materialIndex];

    // light MIS weight
    float realLightPDF = lightSamplePDF(cx, origin, prevHit.normal, lightSample);
    float cosinePDF = cosineHemispherePDF(prevHit.normal, lightSample);
    float misWeight1 = realLightPDF * realLightPDF / (realLightPDF * realLightPDF + cosinePDF * cosinePDF);
    if (std::isnan(misWeight1)) misWeight1 = 0.f;

    // cosine MIS weight
    float lightPDF = lightSamplePDF(cx, origin, prevHit.normal, cosineSample);
    float realCosinePDF = cosineHemispherePDF(prevHit.normal, cosineSample);
    float misWeight2 = realCosinePDF * realCosinePDF / (lightPDF * lightPDF + realCosinePDF * realCosinePDF);
    if (std::isnan(misWeight2)) misWeight2 = 0.f;

    vec3 radiance = albedo * INV_PI * dot(prevHit.normal, cosineSample) / realCosinePDF;
    vec3 directLight = v3(lightMaterial.emissionColor) * lightMaterial.emissionStrength;
    directLight *= albedo * INV_PI * gmax(0.f, dot(prevHit.normal, lightSample)) *
                   (realLightPDF == 0.f ? 0.f : misWeight1 / realLightPDF);

    BxDFResult result = {cosineSample, radiance, directLight, 1.f, misWeight2};
    return result;
}

// ---- TEXTURE EXTENSION (not in the HEAD shader; semantics of this project) -------------------------------------
// Texel of TextureBuffer[slot] at uv with the reference's two samplers (src/vk_engine.cpp:525-531: both NEAREST, no
// mips; sampler 0 REPEAT, sampler 1 CLAMP_TO_EDGE), colour decoded sRGB -> linear (images are R8G8B8A8_SRGB,
// src/vk_textures.cpp:212): texel index = floor(uv * size), wrapped or clamped. All float operations are single.
float srgb_to_linear(uint8_t c) {
    double x = double(c) / 255.0;
    return float(x <= 0.04045 ? x / 12.92 : std::pow((x + 0.055) / 1.055, 2.4));
}
int texel_index(float coord, uint32_t size, uint32_t samplerIndex) {
    float f = std::floor(coord * float(size));
    if (!(std::fabs(f) < 1.0e9f)) f = 0.f;          // non-finite or absurd coordinates: texel 0
    int i = int(f), n = int(size);
    if (samplerIndex == 0u) { i %= n; if (i < 0) i += n; }
    else { i = i < 0 ? 0 : (i > n - 1 ? n - 1 : i); }
    return i;
}
bool sample_texture(const Scene& sc, int32_t slot, vec2 uv, uint32_t samplerIndex, vec3& rgb) {
    if (slot < 0 || slot >= 64 || sc.textures[slot].w == 0) return false;
    const Scene::Texture& t = sc.textures[slot];
    int x = texel_index(uv.x, t.w, samplerIndex), y = texel_index(uv.y, t.h, samplerIndex);
    const uint8_t* px = &t.rgba[(size_t(y) * t.w + x) * 4];
    rgb = vec3(srgb_to_linear(px[0]), srgb_to_linear(px[1]), srgb_to_linear(px[2]));
    return true;
}
// Applied to a triangle hit before the BSDF is chosen: albedo map multiplies the material's albedo (read_mtl puts
// map_Kd in albedoIndex, :1109-1117); metalness map (map_Ks, :1119-1125) switches the mirror BSDF on where its red
// channel is >= 0.5 and off elsewhere. Spheres carry no uv.
void apply_textures(const Scene& sc, HitInfo& hit, float& reflectance) {
    if (!sc.useTextures || hit.kind != 2u) return;
    const ol::RayMaterial& m = sc.materials[hit.materialIndex];
    uint32_t sampler = sc.objects[hit.objectHitIndex].samplerIndex;
    vec3 t;
    if (sample_texture(sc, m.albedoIndex, hit.uv, sampler, t)) hit.albedoTexel = t;
    if (sample_texture(sc, m.metalnessIndex, hit.uv, sampler, t)) reflectance = t.x >= 0.5f ? 1.f : 0.f;
}

// raytrace.comp:466-469
BxDFResult specularBRDF(const vec3& incomingDir, const HitInfo& prevHit, uint32_t& /*state*/) {
    BxDFResult result = {reflect(incomingDir, prevHit.normal), vec3(1.f), vec3(-1.f), 1.f, 1.f};
    return result;
}

// raytrace.comp:471-481. `||` short-circuits: random() is not drawn under total internal reflection.
BxDFResult dielectricBTDF(Ctx& cx, const vec3& incomingDir, const HitInfo& prevHit, uint32_t& state) {
    const ol::RayMaterial& hitMaterial = cx.sc.materials[prevHit.materialIndex];
    float ior = !prevHit.frontFace ? hitMaterial.ior : 1.f / hitMaterial.ior;

    float cosine = dot(-incomingDir, prevHit.normal);
    float sine = std::sqrt(1.f - cosine * cosine);
    bool solution = (ior * sine) > 1.f || schlick(cosine, ior) > random(state);
    vec3 dir = solution ? reflect(incomingDir, prevHit.normal) : refract(incomingDir, prevHit.normal, ior);
    BxDFResult result = {dir, vec3(1.f), vec3(-1.f), (solution ? 1.f : sign(dot(prevHit.normal, incomingDir))), 1.f};
    return result;
}

// raytrace.comp:483-537
vec3 trace(Ctx& cx, const Ray& ray, uint32_t& state, float* stats) {
    vec3 totalColor = vec3(0.f);
    vec3 attenuation = vec3(1.f);
    vec3 directLight = vec3(0.f);
    float misWeight = 1.f;
    HitInfo hit;
    Ray newRay = ray;
    const ol::RayTracerData& traceData = cx.pc.rayTraceParams;
    cx.work.n_path++;

    for (uint32_t j = 0; j <= traceData.bounceLimit; j++) {
        hit = calculateIntersections(cx, newRay, stats);
        if (hit.didHit) {
            cx.work.n_bounce++;
            const ol::RayMaterial& hitMaterial = cx.sc.materials[hit.materialIndex];

            // 0-1 NEE
            vec3 emission = v3(hitMaterial.emissionColor) * hitMaterial.emissionStrength / misWeight;
            vec3 finalLight = directLight.x == -1.f ? emission : directLight;
            totalColor += finalLight * attenuation;
            if (j == 0) totalColor += emission;
            if (isnan3(totalColor) || totalColor.x < 0.f || totalColor.y < 0.f || totalColor.z < 0.f) return vec3(0.f);

            // BxDF
            BxDFResult bxdf;
            float reflectance = hitMaterial.reflectance;
            apply_textures(cx.sc, hit, reflectance);          // no-op unless the texture extension is on
            if (reflectance != 0.f) {
                bxdf = specularBRDF(newRay.dir, hit, state);
            } else if (hitMaterial.ior != -1.f) {
                bxdf = dielectricBTDF(cx, newRay.dir, hit, state);
            } else {
                bxdf = diffuseBRDF(cx, newRay.dir, hit, state);
            }
            attenuation *= bxdf.radiance;
            directLight = bxdf.directLight;

            // russian roulette
            float rrProb = gmax(gmax(attenuation.x, attenuation.y), attenuation.z);
            rrProb = gmin(rrProb, 0.95f);
            rrProb = j <= 5 ? 1.f : rrProb;
            if (random(state) > rrProb) break;
            attenuation *= 1.f / rrProb;

            // prep new bounce
            misWeight = bxdf.cosineMisWeight;
            newRay.origin = hit.hitPoint + hit.normal * bxdf.originSign * 0.00001f;
            newRay.dir = bxdf.sampledDir;
        } else {
            totalColor += attenuation * getEnvironmentLight(cx, newRay);
            break;
        }
    }
    return totalColor;
}

inline float unorm8_load(uint8_t b) { return float(b) / 255.f; }
inline uint8_t unorm8_store(float f) {  // Vulkan unorm conversion: clamp, scale, round-to-nearest-even
    if (!(f > 0.f)) return 0;           // also maps NaN to 0
    if (f >= 1.f) return 255;
    return uint8_t(std::nearbyint(f * 255.f));
}

struct PixelOut {
    vec3 outColor;     // mean over in-dispatch samples, raytrace.comp:569-574
    float stats[2];
};

// raytrace.comp:539-574 — everything before the blend
PixelOut shade_pixel(Ctx& cx, uint32_t gx, uint32_t gy, int W, int H) {
    const ol::PushConstants& pc = cx.pc;
    float uvx = float(gx) / float(W);
    float uvy = float(gy) / float(H);

    const ol::CameraInfo& cam = pc.camInfo;
    float planeHeight = cam.nearPlane * o_tan(radians(cam.fov * 0.5f)) * 2.f;
    float planeWidth = planeHeight * cam.aspectRatio;

    vec3 bottomLeft = vec3(-planeWidth / 2.f, -planeHeight / 2.f, 0.1f);
    vec3 point = bottomLeft + vec3(planeWidth * uvx, planeHeight * uvy, 0.f);
    vec3 dir = normalize(point);

    mat4 rot;
    for (int c = 0; c < 4; ++c) rot[c] = vec4(cam.cameraRotation[4 * c], cam.cameraRotation[4 * c + 1],
                                              cam.cameraRotation[4 * c + 2], cam.cameraRotation[4 * c + 3]);
    Ray ray;
    ray.dir = (rot * vec4(dir, 1.f)).xyz();
    ray.origin = v3(cam.pos);
    ray.invDir = vec3(1.f / ray.dir.x, 1.f / ray.dir.y, 1.f / ray.dir.z);

    uint32_t startingSeed = starting_seed(pc.frameCount);
    uint32_t state = gy * uint32_t(W) + gx + startingSeed;

    const ol::RayTracerData& traceData = pc.rayTraceParams;
    PixelOut po;
    po.stats[0] = 0.f; po.stats[1] = 0.f;
    vec3 outColor = vec3(0.f);
    uint32_t samples = traceData.singleRender ? traceData.sampleLimit : traceData.raysPerPixel;
    for (uint32_t i = 0; i < samples; i++) {
        outColor += trace(cx, ray, state, po.stats);
    }
    outColor /= float(samples);
    po.outColor = outColor;
    return po;
}

}  // namespace

void Scene::finalize() {
    M.resize(objects.size());
    Minv.resize(objects.size());
    for (size_t i = 0; i < objects.size(); ++i) {
        const float* t = objects[i].transformMatrix;
        for (int c = 0; c < 4; ++c) M[i][c] = vec4(t[4 * c], t[4 * c + 1], t[4 * c + 2], t[4 * c + 3]);
        Minv[i] = inverse(M[i]);
    }
}

void render_dispatch(const Scene& sc, const ol::PushConstants& pc, int W, int H, float* image, uint8_t* rgba8,
                     float* outcol, float* stats, Work* work, int threads) {
    Work total;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel
    {
        Ctx cx{sc, pc, Work(), Work()};
#pragma omp for schedule(dynamic, 1)
        for (int y = 0; y < H; ++y) {
            for (int x = 0; x < W; ++x) {
                size_t p = size_t(y) * W + x;
                PixelOut po = shade_pixel(cx, uint32_t(x), uint32_t(y), W, H);
                const ol::RayTracerData& traceData = pc.rayTraceParams;
                if (outcol) { outcol[3 * p] = po.outColor.x; outcol[3 * p + 1] = po.outColor.y; outcol[3 * p + 2] = po.outColor.z; }
                if (stats) { stats[2 * p] = po.stats[0]; stats[2 * p + 1] = po.stats[1]; }

                // raytrace.comp:576-593, once per storage image flavour
                auto finish = [&](const vec3& oldColor) {
                    float weight = 1.f / (float(pc.frameCount) + 1.f);
                    vec3 finalColor = oldColor * (1.f - weight) + po.outColor * weight;
                    finalColor = traceData.progressive ? finalColor : po.outColor;
                    if (isnan3(finalColor) || isinf3(finalColor)) {
                        finalColor = vec3(1.f, 0.f, 1.f);
                    }
                    if (traceData.debug == 0) {
                        finalColor = po.stats[0] > float(traceData.boxCap) ? vec3(1.f, 0.f, 0.f)
                                                                            : vec3(po.stats[0]) / float(traceData.boxCap);
                    } else if (traceData.debug == 1) {
                        finalColor = po.stats[1] > float(traceData.triangleCap) ? vec3(1.f, 0.f, 0.f)
                                                                                 : vec3(po.stats[1]) / float(traceData.triangleCap);
                    } else if (traceData.debug == 2) {
                        finalColor.x = po.stats[0] / float(traceData.boxCap);
                        finalColor.y = 0.f;
                        finalColor.z = po.stats[1] / float(traceData.triangleCap);
                    }
                    return finalColor;
                };
                if (image) {
                    vec3 f = finish(vec3(image[4 * p], image[4 * p + 1], image[4 * p + 2]));
                    image[4 * p] = f.x; image[4 * p + 1] = f.y; image[4 * p + 2] = f.z; image[4 * p + 3] = 1.f;
                }
                if (rgba8) {
                    vec3 f = finish(vec3(unorm8_load(rgba8[4 * p]), unorm8_load(rgba8[4 * p + 1]), unorm8_load(rgba8[4 * p + 2])));
                    rgba8[4 * p] = unorm8_store(f.x); rgba8[4 * p + 1] = unorm8_store(f.y);
                    rgba8[4 * p + 2] = unorm8_store(f.z); rgba8[4 * p + 3] = 255;
                }
            }
        }
#pragma omp critical
        total.add(cx.work);
    }
    if (work) work->add(total);
}

void render_frames(const Scene& sc, const ol::PushConstants& pc_in, int W, int H, uint32_t first, uint32_t n,
                   float* accum, Work* work, int threads, Work* workExt) {
    Work total, totalExt;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    for (uint32_t f = first; f < first + n; ++f) {
        ol::PushConstants pc = pc_in;
        pc.frameCount = f;
#pragma omp parallel
        {
            Ctx cx{sc, pc, Work(), Work()};
#pragma omp for schedule(dynamic, 1)
            for (int y = 0; y < H; ++y) {
                for (int x = 0; x < W; ++x) {
                    size_t p = size_t(y) * W + x;
                    PixelOut po = shade_pixel(cx, uint32_t(x), uint32_t(y), W, H);
                    vec3 c = po.outColor;
                    if (isnan3(c) || isinf3(c)) c = vec3(1.f, 0.f, 1.f);
                    accum[4 * p] += c.x; accum[4 * p + 1] += c.y; accum[4 * p + 2] += c.z; accum[4 * p + 3] += 1.f;
                }
            }
#pragma omp critical
            { total.add(cx.work); totalExt.add(cx.ext); }
        }
    }
    if (work) work->add(total);
    if (workExt) workExt->add(totalExt);
}

// Pixels (x0 + i*sx, y0 + j*sy), i < w, j < h, of the W x H image (seeds and rays are those of the full
// image: raytrace.comp:541,562-564), frames first .. first+n-1, accumulated into a w x h x 4 buffer. Lets the
// tests and the timed CPU baseline look at full-resolution configurations through a window or a pixel lattice.
void render_window(const Scene& sc, const ol::PushConstants& pc_in, int W, int H, uint32_t first, uint32_t n, int x0, int y0,
                   int w, int h, int sx, int sy, float* accum, Work* work, int threads, Work* workExt) {
    Work total, totalExt;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    for (uint32_t f = first; f < first + n; ++f) {
        ol::PushConstants pc = pc_in;
        pc.frameCount = f;
#pragma omp parallel
        {
            Ctx cx{sc, pc, Work(), Work()};
#pragma omp for schedule(dynamic, 1)
            for (int j = 0; j < h; ++j) {
                for (int i = 0; i < w; ++i) {
                    size_t p = size_t(j) * w + i;
                    PixelOut po = shade_pixel(cx, uint32_t(x0 + i * sx), uint32_t(y0 + j * sy), W, H);
                    vec3 c = po.outColor;
                    if (isnan3(c) || isinf3(c)) c = vec3(1.f, 0.f, 1.f);
                    accum[4 * p] += c.x; accum[4 * p + 1] += c.y; accum[4 * p + 2] += c.z; accum[4 * p + 3] += 1.f;
                }
            }
#pragma omp critical
            { total.add(cx.work); totalExt.add(cx.ext); }
        }
    }
    if (work) work->add(total);
    if (workExt) workExt->add(totalExt);
}

// For each listed pixel (x, y) of frame `frame`: how many rays of its path (extension, NEE and pdf rays alike) met their
// closest hit at exactly the same distance on two triangles of one object. A pixel whose image value differs between two
// correct traversals in different orders must have at least one (tests/test_gpu_parity.py: every differing pixel is
// explained, none is merely budgeted).
void pixel_ties(const Scene& sc, const ol::PushConstants& pc_in, int W, int H, uint32_t frame, const uint32_t* xy, size_t n,
                uint32_t* ties) {
    ol::PushConstants pc = pc_in;
    pc.frameCount = frame;
#pragma omp parallel for schedule(dynamic, 16)
    for (long long i = 0; i < (long long)n; ++i) {
        Ctx cx{sc, pc, Work(), Work()};
        cx.detectTies = true;
        shade_pixel(cx, xy[2 * i], xy[2 * i + 1], W, H);
        ties[i] = cx.ties;
    }
}

// Every calculateIntersections call on the path of pixel (x, y) of frame `frame`, in call order (extension ray, then for
// a diffuse bounce the NEE ray and the two pdf rays), with the hit the oracle found. Returns the number of rays; fills up
// to `cap`. Lets a test replay the oracle's own rays through another traversal and find the first ray that differs.
size_t pixel_rays(const Scene& sc, const ol::PushConstants& pc_in, int W, int H, uint32_t frame, uint32_t x, uint32_t y, size_t cap,
                  float* rays8, HitOut* hits) {
    ol::PushConstants pc = pc_in;
    pc.frameCount = frame;
    std::vector<float> rr; std::vector<HitOut> hh;
    Ctx cx{sc, pc, Work(), Work()};
    cx.recRays = &rr; cx.recHits = &hh;
    shade_pixel(cx, x, y, W, H);
    const size_t n = hh.size(), m = n < cap ? n : cap;
    if (m) { std::memcpy(rays8, rr.data(), m * 32); std::memcpy(hits, hh.data(), m * sizeof(HitOut)); }
    return n;
}

void trace_batch(const Scene& sc, const ol::PushConstants& pc, const float* rays, size_t n, HitOut* out, Work* work,
                 int threads) {
    Work total;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel
    {
        Ctx cx{sc, pc, Work(), Work()};
#pragma omp for schedule(dynamic, 256)
        for (long long i = 0; i < (long long)n; ++i) {
            Ray r;
            r.origin = vec3(rays[8 * i], rays[8 * i + 1], rays[8 * i + 2]);
            r.dir = vec3(rays[8 * i + 4], rays[8 * i + 5], rays[8 * i + 6]);
            HitInfo h = calculateIntersections(cx, r, nullptr);
            HitOut& o = out[i];
            std::memset(&o, 0, sizeof(o));
            o.dst = h.dst;
            o.kind = h.didHit ? h.kind : 0;
            o.objectHitIndex = h.didHit ? h.objectHitIndex : 0xffffffffu;
            o.triHitIndex = h.didHit ? h.triHitIndex : 0;
            o.materialIndex = h.materialIndex;
            o.frontFace = h.didHit ? (h.frontFace ? 1u : 0u) : 0u;
            if (h.didHit) {
                o.normal[0] = h.normal.x; o.normal[1] = h.normal.y; o.normal[2] = h.normal.z;
                o.hitPoint[0] = h.hitPoint.x; o.hitPoint[1] = h.hitPoint.y; o.hitPoint[2] = h.hitPoint.z;
            }
        }
#pragma omp critical
        total.add(cx.work);
    }
    if (work) work->add(total);
}

}  // namespace orc
