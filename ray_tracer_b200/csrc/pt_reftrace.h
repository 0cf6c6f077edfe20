// The compute pass on the REFERENCE buffer layout: a literal per-pixel evaluation of
// calculateIntersections()/trace()/main() (shaders/raytrace.comp:276-353,483-594) that reads the
// six storage buffers exactly as uploaded. It serves the single-dispatch entry point
// (b200pt_render: every RayTracerData switch, debug heat-maps, in-dispatch multi-sampling) and
// is the correctness anchor the wavefront pipeline is compared with. The throughput path is
// wavefront.cu, not this.
#pragma once
#include "pt_shade.h"

namespace pt {

struct RefScene {
    const ptl::Sphere* spheres;
    const ptl::Material* materials;
    const ptl::TrianglePoint* points;
    const ptl::Triangle* tris;
    const ptl::RenderObject* objects;
    const ptl::BVHNode* nodes;
    const float* minv;          // 16 floats per object: inverse(transformMatrix), hoisted (A-Q13)
    uint32_t nSpheres, nMaterials, nPoints, nTris, nObjects, nNodes;
};

struct RefHit {
    v3 hp, n;
    float dst;
    float u, v;                 // barycentrics of a triangle hit (raytrace.comp:239-240)
    uint32_t objectHitIndex, triHitIndex, materialIndex;
    uint32_t kind;              // 0 miss, 1 sphere, 2 triangle
    bool front;
};

PT_HD v3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }

// calculateIntersections(), raytrace.comp:276-353. stats may be null (NEE / pdf rays use
// throw-away arrays in the shader, :391,:442).
// bound0 (default: none): a PHANTOM upper bound for the re-trace of contested rays (pt_traverse.h). The wide traversal
// already knows the smallest candidate distance m; every candidate that can win lies within the contest window w of m and
// its enclosing boxes start before m + 2w. Starting the search with closestHit.dst = m + 3w instead of 99999999 prunes, from
// the first node on, only subtrees whose boxes start beyond it: their triangles cannot win, and the only other thing they
// could do — lower closestHit.dst to some value still above m + 2w — changes no pruning decision for a candidate that can
// win (its boxes start before that value either way). Same winner, a tenth of the node visits. A search that ends without a
// hit is repeated without the bound.
PT_HD_NOINLINE RefHit ref_closest_hit(const RefScene& sc, const ptl::RayTracerData& td, const Ray& ray, float* stats, float bound0 = PT_FAR) {
    RefHit best;
    best.kind = 0;
    best.dst = bound0;
    best.materialIndex = 0;                     // uninitialised in the shader on a miss; pinned (A-Q9)
    best.objectHitIndex = 0xffffffffu;
    best.triHitIndex = 0;
    best.front = false;
    best.hp = splat3(0.f);
    best.n = splat3(0.f);
    best.u = 0.f; best.v = 0.f;

    for (uint32_t i = 0; i < td.sphereCount; i++) {
        float dst; bool front;
        if (sphere_test(sc.spheres[i], ray, dst, front) && dst < best.dst) {
            best.kind = 1; best.dst = dst; best.front = front; best.triHitIndex = i;
            best.materialIndex = sc.spheres[i].materialIndex;
            sphere_surface(sc.spheres[i], ray, dst, front, best.hp, best.n);
        }
    }

    for (uint32_t i = 0; i < td.objectCount; i++) {
        const ptl::RenderObject& obj = sc.objects[i];
        const float* Minv = sc.minv + 16 * i;
        Ray tr;
        tr.d = mul_dir(Minv, ray.d);
        tr.o = mul_point(Minv, ray.o);
        v3 inv = mk3(1.f / tr.d.x, 1.f / tr.d.y, 1.f / tr.d.z);

        // Without counters to keep (the re-trace of contested rays): an object whose ROOT box the ray misses, or enters beyond
        // the closest hit so far, contributes nothing — the shader never tests that box (:301-307), but both children's boxes
        // lie inside it float for float (update_bvh_bounds takes min / max over subsets of the same vertices) and
        // boxIntersection is monotone in the box (rounding of (b - o) * inv, min, max), so each child's distance is >= the
        // root's and the loop below would push neither. Exactly the same result, minus two node loads and box tests.
        if (!stats) {
            const ptl::BVHNode root = sc.nodes[obj.bvhIndex];
            if (root.triCount == 0 && tr.d.x != 0.f && tr.d.y != 0.f && tr.d.z != 0.f &&      // finite 1 / d: no 0 * inf = NaN in either test, the order argument holds
                !(box_test_ref(mk3(root.minX, root.minY, root.minZ), mk3(root.maxX, root.maxY, root.maxZ), tr.o, inv) < best.dst)) continue;
        }
        uint32_t stack[72];                     // the shader has 64 slots for a depth-64 tree (A-Q16)
        uint32_t sp = 1;
        stack[0] = obj.bvhIndex;
        while (sp > 0) {
            const ptl::BVHNode nd = sc.nodes[stack[--sp]];
            if (nd.triCount != 0) {
                if (stats) stats[1] += float(nd.triCount);
                for (uint32_t j = nd.index; j < nd.index + nd.triCount; j++) {
                    const ptl::Triangle& tri = sc.tris[j];
                    const ptl::TrianglePoint& a = sc.points[tri.v0];
                    const ptl::TrianglePoint& b = sc.points[tri.v1];
                    const ptl::TrianglePoint& c = sc.points[tri.v2];
                    TriEdges te = tri_edges(mk3(a.px, a.py, a.pz), mk3(b.px, b.py, b.pz), mk3(c.px, c.py, c.pz));
                    float dst, u, v; bool front;
                    if (tri_test(te, tr, tri.frontOnly != 0, dst, u, v, front) && dst < best.dst) {
                        best.kind = 2; best.dst = dst; best.front = front; best.u = u; best.v = v;
                        best.triHitIndex = j; best.objectHitIndex = i; best.materialIndex = obj.materialIndex;
                        tri_surface(obj.m, tr, dst, u, v, front, mk3(a.nx, a.ny, a.nz), mk3(b.nx, b.ny, b.nz),
                                    mk3(c.nx, c.ny, c.nz), best.hp, best.n);
                    }
                }
            } else {
                const ptl::BVHNode c1 = sc.nodes[nd.index];
                const ptl::BVHNode c2 = sc.nodes[nd.index + 1];
                float d1 = box_test_ref(mk3(c1.minX, c1.minY, c1.minZ), mk3(c1.maxX, c1.maxY, c1.maxZ), tr.o, inv);
                float d2 = box_test_ref(mk3(c2.minX, c2.minY, c2.minZ), mk3(c2.maxX, c2.maxY, c2.maxZ), tr.o, inv);
                if (stats) stats[0] += 2.f;
                bool nearA = d1 <= d2;
                float dNear = nearA ? d1 : d2, dFar = nearA ? d2 : d1;
                uint32_t iNear = nearA ? nd.index : nd.index + 1, iFar = nearA ? nd.index + 1 : nd.index;
                if (dFar < best.dst && sp < 72) stack[sp++] = iFar;
                if (dNear < best.dst && sp < 72) stack[sp++] = iNear;
            }
        }
    }
    if (best.kind == 0u) {
        if (bound0 < PT_FAR) return ref_closest_hit(sc, td, ray, stats, PT_FAR);
        best.dst = PT_FAR;
    }
    return best;
}

// lightSamplePDF(), raytrace.comp:389-403
PT_HD float ref_light_pdf(const RefScene& sc, const ptl::RayTracerData& td, v3 origin, v3 dir) {
    Ray r; r.o = origin; r.d = dir;
    RefHit h = ref_closest_hit(sc, td, r, nullptr);
    if (h.kind == 0 || sc.materials[h.materialIndex].emissionStrength == 0.f) return 0.f;
    return light_pdf_from_dst(h.dst, dir);
}

// trace(), raytrace.comp:483-537 with diffuseBRDF (:430-464) / specularBRDF (:466-469) /
// dielectricBTDF (:471-481) in line.
PT_HD_NOINLINE v3 ref_trace(const RefScene& sc, const ptl::PushConstants& pc, Ray ray, uint32_t& state, float* stats) {
    const ptl::RayTracerData& td = pc.params;
    v3 total = splat3(0.f), atten = splat3(1.f), direct = splat3(0.f);
    float misWeight = 1.f;
    for (uint32_t j = 0; j <= td.bounceLimit; j++) {
        RefHit hit = ref_closest_hit(sc, td, ray, stats);
        if (hit.kind == 0) {
            total = total + atten * environment_light(pc.env, ray.d);
            break;
        }
        const ptl::Material& mat = sc.materials[hit.materialIndex];
        v3 emission = ld3(mat.emissionColor) * mat.emissionStrength / misWeight;
        v3 finalLight = direct.x == -1.f ? emission : direct;
        total = total + finalLight * atten;
        if (j == 0) total = total + emission;
        if (isnan3(total) || total.x < 0.f || total.y < 0.f || total.z < 0.f) return splat3(0.f);

        v3 dirOut, radiance;
        float originSign = 1.f, nextMis = 1.f;
        if (mat.reflectance != 0.f) {
            dirOut = reflect(ray.d, hit.n); radiance = splat3(1.f); direct = splat3(-1.f);
        } else if (mat.ior != -1.f) {
            dirOut = dielectric_dir(ray.d, hit.n, hit.front, mat.ior, state, originSign);
            radiance = splat3(1.f); direct = splat3(-1.f);
        } else {
            v3 albedo = ld3(mat.albedo);
            v3 origin = hit.hp + hit.n * 0.01f;
            v3 lightSample = light_sample_dir(origin, state);
            v3 cosineSample = cosine_hemisphere_dir(hit.n, state);
            Ray lr; lr.o = origin; lr.d = lightSample;
            RefHit lightHit = ref_closest_hit(sc, td, lr, nullptr);
            const ptl::Material& lm = sc.materials[lightHit.materialIndex];
            float realLightPDF = ref_light_pdf(sc, td, origin, lightSample);
            float cosinePDF = cosine_hemisphere_pdf(hit.n, lightSample);
            float w1 = mis_weight(realLightPDF, cosinePDF);
            float lightPDF = ref_light_pdf(sc, td, origin, cosineSample);
            float realCosinePDF = cosine_hemisphere_pdf(hit.n, cosineSample);
            float w2 = mis_weight(realCosinePDF, lightPDF);
            radiance = albedo * PT_INV_PI * dot(hit.n, cosineSample) / realCosinePDF;
            direct = ld3(lm.emissionColor) * lm.emissionStrength;
            direct = direct * (albedo * PT_INV_PI * gmax(0.f, dot(hit.n, lightSample)) *
                               (realLightPDF == 0.f ? 0.f : w1 / realLightPDF));
            dirOut = cosineSample; nextMis = w2;
        }
        atten = atten * radiance;

        float rr = gmax(gmax(atten.x, atten.y), atten.z);
        rr = gmin(rr, 0.95f);
        rr = j <= 5 ? 1.f : rr;
        if (rng_next(state) > rr) break;
        atten = atten * (1.f / rr);

        misWeight = nextMis;
        ray.o = hit.hp + hit.n * originSign * 0.00001f;
        ray.d = dirOut;
    }
    return total;
}

struct RefPixel { v3 outColor; float stats0, stats1; };

// main() up to the blend, raytrace.comp:539-574
PT_HD RefPixel ref_pixel(const RefScene& sc, const ptl::PushConstants& pc, const CameraPlane& cp, uint32_t gx, uint32_t gy, int W, int H) {
    Ray ray = primary_ray(pc.cam, cp, gx, gy, W, H);
    uint32_t state = gy * uint32_t(W) + gx + frame_seed(pc.frameCount);
    float stats[2] = {0.f, 0.f};
    v3 out = splat3(0.f);
    uint32_t samples = pc.params.singleRender ? pc.params.sampleLimit : pc.params.raysPerPixel;
    for (uint32_t i = 0; i < samples; i++) out = out + ref_trace(sc, pc, ray, state, stats);
    out = out / float(samples);
    RefPixel r; r.outColor = out; r.stats0 = stats[0]; r.stats1 = stats[1];
    return r;
}

// option "storage_rgba8": the reference's storage image is rgba8 (raytrace.comp:120,593) — what the next dispatch reads
// back is the 8-bit value. Vulkan unorm conversion: clamp, scale, round to nearest even; no transfer function (A-Q3:
// whether MoltenVK applies sRGB on a storage image is driver-defined — parity unpinned).
PT_HD float unorm8_round(float f) {
    if (!(f > 0.f)) return 0.f;
    if (f >= 1.f) return 1.f;
#if defined(__CUDA_ARCH__)
    return float(__float2int_rn(f * 255.f)) / 255.f;
#else
    return float((int)nearbyintf(f * 255.f)) / 255.f;
#endif
}

// main() after the sample loop, raytrace.comp:576-591: progressive blend, NaN/Inf -> magenta,
// debug heat-maps. `old` is the stored colour of this pixel.
PT_HD v3 ref_resolve(const ptl::PushConstants& pc, const RefPixel& px, v3 old) {
    const ptl::RayTracerData& td = pc.params;
    float weight = 1.f / (float(pc.frameCount) + 1.f);
    v3 fin = old * (1.f - weight) + px.outColor * weight;
    fin = td.progressive ? fin : px.outColor;
    if (isnan3(fin) || isinf3(fin)) fin = mk3(1.f, 0.f, 1.f);
    if (td.debug == 0) {
        fin = px.stats0 > float(td.boxCap) ? mk3(1.f, 0.f, 0.f) : splat3(px.stats0) / float(td.boxCap);
    } else if (td.debug == 1) {
        fin = px.stats1 > float(td.triangleCap) ? mk3(1.f, 0.f, 0.f) : splat3(px.stats1) / float(td.triangleCap);
    } else if (td.debug == 2) {
        fin.x = px.stats0 / float(td.boxCap);
        fin.y = 0.f;
        fin.z = px.stats1 / float(td.triangleCap);
    }
    return fin;
}

}  // namespace pt
