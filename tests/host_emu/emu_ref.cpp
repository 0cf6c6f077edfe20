// TEST INFRASTRUCTURE — host-side emulation of the backend's kernels.
// Compiles the SAME headers the CUDA translation units include (PT_HD functions are plain
// inline functions under g++) and runs them in loops on the CPU, so that the kernels' logic can
// be checked bit-for-bit against the oracle in the CPU-only container. Nothing here is linked
// into libb200pt.so; the product has no CPU path.
#include <cstring>
#include <vector>

#include "../../ray_tracer_b200/csrc/pt_reftrace.h"

using namespace pt;

extern "C" {

// one dispatch of the reference-layout kernel: outcol (W*H*3), stats (W*H*2, optional),
// image (W*H*4 in/out, optional)
int emu_ref_dispatch(const void* spheres, uint32_t nS, const void* materials, uint32_t nM, const void* points,
                     uint32_t nP, const void* tris, uint32_t nT, const void* objects, uint32_t nO, const void* nodes,
                     uint32_t nN, const void* pc208, int W, int H, float* outcol, float* stats, float* image) {
    RefScene sc;
    sc.spheres = (const ptl::Sphere*)spheres; sc.materials = (const ptl::Material*)materials;
    sc.points = (const ptl::TrianglePoint*)points; sc.tris = (const ptl::Triangle*)tris;
    sc.objects = (const ptl::RenderObject*)objects; sc.nodes = (const ptl::BVHNode*)nodes;
    sc.nSpheres = nS; sc.nMaterials = nM; sc.nPoints = nP; sc.nTris = nT; sc.nObjects = nO; sc.nNodes = nN;
    std::vector<float> minv(16 * (size_t)nO);
    for (uint32_t i = 0; i < nO; ++i) inverse4(sc.objects[i].m, &minv[16 * i]);
    sc.minv = minv.data();
    ptl::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    CameraPlane cp = camera_plane(pc.cam);
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            size_t p = (size_t)y * W + x;
            RefPixel px = ref_pixel(sc, pc, cp, (uint32_t)x, (uint32_t)y, W, H);
            if (outcol) { outcol[3 * p] = px.outColor.x; outcol[3 * p + 1] = px.outColor.y; outcol[3 * p + 2] = px.outColor.z; }
            if (stats) { stats[2 * p] = px.stats0; stats[2 * p + 1] = px.stats1; }
            if (image) {
                v3 f = ref_resolve(pc, px, mk3(image[4 * p], image[4 * p + 1], image[4 * p + 2]));
                image[4 * p] = f.x; image[4 * p + 1] = f.y; image[4 * p + 2] = f.z; image[4 * p + 3] = 1.f;
            }
        }
    return 0;
}

}  // extern "C"
