// Per-path arithmetic of the compute pass, shared by the reference-layout kernel
// (megakernel.cu) and the wavefront stages (wavefront.cu). Each function states the lines
// of shaders/raytrace.comp whose results it reproduces; float operations appear in the
// shader's evaluation order (see pt_math.h for why that matters).
#pragma once
#include "pt_layout.h"
#include "pt_math.h"
#include "pt_portable_math.h"

namespace pt {

#define PT_PI 3.1415926535897932384f      /* raytrace.comp:6 */
#define PT_INV_PI 0.3183098862f           /* raytrace.comp:7 */
#define PT_FAR 99999999.f                 /* raytrace.comp:273,279 (rounds to 1e8f) */

struct Ray { v3 o, d; };

// random(): PCG-style LCG + output hash, raytrace.comp:158-163. The divisor literal
// 4294967295.f rounds to 2^32, and the result can be exactly 1.0 (A-Q6).
PT_HD float rng_next(uint32_t& state) {
    state = state * 747796405u + 2891336453u;
    uint32_t r = ((state >> ((state >> 28) + 4u)) ^ state) * 277803737u;
    r = (r >> 22) ^ r;
    return float(r) / 4294967296.f;
}
// raytrace.comp:562-564; int literal 23892183 converts to the float 23892184
PT_HD uint32_t frame_seed(uint32_t frameCount) {
    uint32_t s = frameCount;
    return uint32_t(rng_next(s) * 23892184.f);
}

// Camera constants of main() that do not depend on the pixel (raytrace.comp:547-550)
struct CameraPlane { float planeW, planeH, blx, bly; };
PT_HD CameraPlane camera_plane(const ptl::CameraInfo& cam) {
    CameraPlane c;
    c.planeH = cam.nearPlane * pt_tanf(radians(cam.fov * 0.5f)) * 2.f;
    c.planeW = c.planeH * cam.aspectRatio;
    c.blx = -c.planeW / 2.f;
    c.bly = -c.planeH / 2.f;
    return c;
}
// raytrace.comp:541,550-556: uv = gid/dim (no pixel-centre offset, no jitter: A-Q4)
PT_HD Ray primary_ray(const ptl::CameraInfo& cam, const CameraPlane& cp, uint32_t gx, uint32_t gy, int W, int H) {
    float uvx = float(gx) / float(W);
    float uvy = float(gy) / float(H);
    v3 point = mk3(cp.blx + cp.planeW * uvx, cp.bly + cp.planeH * uvy, 0.1f + 0.f);
    v3 dir = normalize(point);
    Ray r;
    r.d = mul_point(cam.rot, dir);          // cameraRotation * vec4(dir, 1)
    r.o = mk3(cam.pos[0], cam.pos[1], cam.pos[2]);
    return r;
}

// schlick(), raytrace.comp:177-181
PT_HD float schlick(float cosine, float ri) {
    float r0 = (1.f - ri) / (1.f + ri);
    r0 = r0 * r0;
    return r0 + (1.f - r0) * pt_powf(1.f - cosine, 5.f);
}

// sphereIntersection(), raytrace.comp:195-224. Returns hit?; dst and frontFace by reference.
// hitPoint/normal are pure functions of (ray, sphere, dst, frontFace): see sphere_surface().
PT_HD bool sphere_test(const ptl::Sphere& s, const Ray& ray, float& dst, bool& front) {
    v3 oc = mk3(s.px, s.py, s.pz) - ray.o;
    float a = dot(ray.d, ray.d);
    float b = dot(oc, ray.d);
    float c = dot(oc, oc) - s.radius * s.radius;
    float disc = b * b - a * c;
    if (!(disc >= 0.f)) return false;
    float sq = sqrtf(disc);
    dst = (b - sq) / a;
    front = true;
    if (dst < 0.f) {
        dst = (b + sq) / a;
        front = false;
        if (dst < 0.f) return false;
    }
    return true;
}
PT_HD void sphere_surface(const ptl::Sphere& s, const Ray& ray, float dst, bool front, v3& hp, v3& n) {
    hp = ray.o + ray.d * dst;
    n = normalize(hp - mk3(s.px, s.py, s.pz)) * (front ? 1.f : -1.f);
}

// The edge vectors and face normal of triangleIntersection() (raytrace.comp:229-232) depend
// only on the vertices; they are evaluated once at upload with the same operations.
struct TriEdges { v3 p0, e1, e2, n; };      // p0, v1v0, v2v0, cross(v1v0, v2v0)
PT_HD TriEdges tri_edges(v3 p0, v3 p1, v3 p2) {
    TriEdges t;
    t.p0 = p0;
    t.e1 = p1 - p0;
    t.e2 = p2 - p0;
    t.n = cross(t.e1, t.e2);
    return t;
}
// triangleIntersection(), raytrace.comp:227-248 (the ray is in object space). u, v, dst are the
// shader's values; didHit as at :246.
PT_HD bool tri_test_r(const TriEdges& t, const Ray& ray, bool frontOnly, float& dst, float& u, float& v, bool& front, float& invD0) {
    v3 rov0 = ray.o - t.p0;
    v3 q = cross(rov0, ray.d);
    float d0 = -dot(ray.d, t.n);
    float d = 1.f / d0;
    invD0 = d;                  // for the distance slack of an accepted hit (pt_traverse.h: tri_dst_slack)
    dst = dot(rov0, t.n) * d;
    u = dot(t.e2, q) * d;
    v = -dot(t.e1, q) * d;
    float w = 1.f - u - v;
    front = d0 >= 0.00000001f;
    return dst >= 0.f && u >= 0.f && v >= 0.f && w >= 0.f && !(!front && frontOnly);
}
PT_HD bool tri_test(const TriEdges& t, const Ray& ray, bool frontOnly, float& dst, float& u, float& v, bool& front) {
    float invD0;
    return tri_test_r(t, ray, frontOnly, dst, u, v, front, invD0);
}
// raytrace.comp:247,259 then :318-319 — object-space hit point / interpolated normal, moved to
// world space with the object's matrix (normal by M, not inverse-transpose: A-Q13).
PT_HD void tri_surface(const float* M, const Ray& objRay, float dst, float u, float v, bool front,
                       v3 n0, v3 n1, v3 n2, v3& hp, v3& n) {
    float w = 1.f - u - v;
    v3 hpo = objRay.o + objRay.d * dst;
    v3 no = (w * n0 + u * n1 + v * n2) * (front ? 1.f : -1.f);
    n = mul_dir_normalize4(M, no);
    hp = mul_point(M, hpo);
}

// boxIntersection(), raytrace.comp:263-274 (reference-layout traversal only)
PT_HD float box_test_ref(v3 bmin, v3 bmax, v3 o, v3 invDir) {
    v3 tMin = (bmin - o) * invDir;
    v3 tMax = (bmax - o) * invDir;
    v3 t1 = mk3(gmin(tMin.x, tMax.x), gmin(tMin.y, tMax.y), gmin(tMin.z, tMax.z));
    v3 t2 = mk3(gmax(tMin.x, tMax.x), gmax(tMin.y, tMax.y), gmax(tMin.z, tMax.z));
    float tNear = gmax(gmax(t1.x, t1.y), t1.z);
    float tFar = gmin(gmin(t2.x, t2.y), t2.z);
    bool hit = tFar >= tNear && tFar > 0.f;
    return hit ? (tNear > 0.f ? tNear : 0.f) : PT_FAR;
}

// getEnvironmentLight(), raytrace.comp:356-365
PT_HD v3 environment_light(const ptl::EnvironmentData& env, v3 dir) {
    if (!(env.lightDir[3] == 1.f)) return splat3(0.f);
    v3 horizon = mk3(env.horizon[0], env.horizon[1], env.horizon[2]);
    v3 zenith = mk3(env.zenith[0], env.zenith[1], env.zenith[2]);
    v3 ground = mk3(env.ground[0], env.ground[1], env.ground[2]);
    v3 ld = mk3(env.lightDir[0], env.lightDir[1], env.lightDir[2]);
    float skyT = pt_powf(smoothstep(0.f, 0.4f, -dir.y), 0.35f);
    v3 sky = mix(horizon, zenith, skyT);
    float sun = pt_powf(gmax(0.f, dot(dir, -ld)), env.horizon[3]) * env.zenith[3];
    float g2s = smoothstep(-0.01f, 0.f, -dir.y);
    float sunMask = float(g2s >= 1.f);
    float s = sun * sunMask;
    v3 m = mix(ground, sky, g2s);
    return mk3(m.x + s, m.y + s, m.z + s);
}

// lightSampleDir(), raytrace.comp:368-387: uniform point on the hard-wired ceiling quad
PT_HD v3 light_sample_dir(v3 origin, uint32_t& state) {
    float x = rng_next(state);
    float z = rng_next(state);
    float rx = mix(-0.33333f, 0.33333f, x);
    float rz = mix(-0.33333f, 0.33333f, z);
    return normalize(mk3(rx, -1.5f, rz) - origin);
}
// tail of lightSamplePDF(), raytrace.comp:398-401, given the closest emissive hit distance
PT_HD float light_pdf_from_dst(float dst, v3 direction) {
    float sq = dst * dst;
    float cosTheta = dot(mk3(0.f, -1.f, 0.f), direction);
    return sq / (cosTheta * 0.4444444f);
}
// cosineHemisphereDir(), raytrace.comp:405-424
PT_HD v3 cosine_hemisphere_dir(v3 n, uint32_t& state) {
    float r1 = rng_next(state);
    float r2 = rng_next(state);
    float phi = 2.f * PT_PI * r1;
    float sqrtR2 = sqrtf(r2);
    float sn, cs;
    pt_sincosf(phi, &sn, &cs);
    float x = cs * sqrtR2;
    float y = sn * sqrtR2;
    float z = sqrtf(1.f - r2);
    v3 axis = fabsf(dot(n, mk3(1.f, 0.f, 0.f))) < 1.f ? mk3(1.f, 0.f, 0.f) : mk3(0.f, 0.f, 1.f);
    v3 t = normalize(cross(n, axis));
    v3 b = cross(n, t);
    // mat3(t, b, n) * (x, y, z): type_mat3x3.inl:468-474
    return mk3(t.x * x + b.x * y + n.x * z, t.y * x + b.y * y + n.y * z, t.z * x + b.z * y + n.z * z);
}
// cosineHemispherePDF(), raytrace.comp:426-428
PT_HD float cosine_hemisphere_pdf(v3 n, v3 d) { return gmax(0.f, dot(d, n) * PT_INV_PI); }

// power-heuristic weight with the shader's NaN -> 0 rule (raytrace.comp:449-450,455-456)
PT_HD float mis_weight(float num, float other) {
    float w = num * num / (other * other + num * num);
    return (w != w) ? 0.f : w;
}

// dielectricBTDF(), raytrace.comp:471-481. `||` short-circuits: no random number under TIR.
PT_HD v3 dielectric_dir(v3 in, v3 n, bool frontFace, float matIor, uint32_t& state, float& originSign) {
    float ior = !frontFace ? matIor : 1.f / matIor;
    float cosine = dot(-in, n);
    float sine = sqrtf(1.f - cosine * cosine);
    bool solution = (ior * sine) > 1.f || schlick(cosine, ior) > rng_next(state);
    originSign = solution ? 1.f : sign(dot(n, in));
    return solution ? reflect(in, n) : refract(in, n, ior);
}

}  // namespace pt
