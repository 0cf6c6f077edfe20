// C entry points of the host-side scene builder (libb200pt_host.so), for the Python bench/tests and
// for any C caller that wants reference-layout buffers from OBJ/MTL files or procedural meshes.
#include <cstring>
#include <string>

#include "scene_builder.h"

using pth::Placement;
using pth::SceneBuilder;

static Placement mk(const float* pos, const float* rot, const float* scl, uint32_t sampler, int frontOnly) {
    Placement p;
    for (int i = 0; i < 3; ++i) { p.position[i] = pos[i]; p.rotation[i] = rot[i]; p.scale[i] = scl[i]; }
    p.samplerIndex = sampler; p.frontOnly = frontOnly != 0;
    return p;
}

extern "C" {

void* b200host_builder_create() { return new SceneBuilder(); }
// fn: address of b200pt_build_bvh (or any function with that signature); NULL restores the host build
void b200host_set_bvh_builder(void* b, void* fn, int device, uint32_t minTriangles) {
    ((SceneBuilder*)b)->set_bvh_builder((SceneBuilder::BvhBuildFn)fn, device, minTriangles);
}
void b200host_bvh_builder_stats(void* b, double* deviceMs, uint32_t* builds) {
    *deviceMs = ((SceneBuilder*)b)->bvhDeviceMs; *builds = ((SceneBuilder*)b)->bvhPluggedBuilds;
}
void b200host_builder_destroy(void* b) { delete (SceneBuilder*)b; }

// 0 ok, 1 file not found (the reference silently skips it), 2 malformed input
int b200host_read_obj(void* b, const char* path, const float* pos, const float* rot, const float* scl, uint32_t sampler,
                      int frontOnly, int material) {
    try { return ((SceneBuilder*)b)->read_obj(path, mk(pos, rot, scl, sampler, frontOnly), material) ? 0 : 1; }
    catch (const std::exception&) { return 2; }
}
int b200host_add_mesh(void* b, const void* faceVertices, uint64_t nTris, const float* pos, const float* rot, const float* scl,
                      int frontOnly, int material) {
    try { ((SceneBuilder*)b)->add_mesh((const ptl::TrianglePoint*)faceVertices, (size_t)nTris, mk(pos, rot, scl, 0, frontOnly), material); return 0; }
    catch (const std::exception&) { return 2; }
}
int b200host_default_scene(void* b, const char* assetDir) {
    try { ((SceneBuilder*)b)->default_scene(assetDir); return 0; } catch (const std::exception&) { return 2; }
}
int b200host_cornell_box(void* b, const char* assetDir) {
    try { ((SceneBuilder*)b)->cornell_box(assetDir); return 0; } catch (const std::exception&) { return 2; }
}
int b200host_read_mtl(void* b, const char* path) {
    try { ((SceneBuilder*)b)->read_mtl(path); return 0; } catch (const std::exception&) { return 2; }
}
void b200host_push_default_materials(void* b) { ((SceneBuilder*)b)->push_default_materials(); }
void b200host_add_material(void* b, const void* mat64) {
    ptl::Material m; std::memcpy(&m, mat64, 64); ((SceneBuilder*)b)->materials.push_back(m);
}
void b200host_set_spheres(void* b, const void* sph, uint32_t n) {
    ((SceneBuilder*)b)->spheres.assign((const ptl::Sphere*)sph, (const ptl::Sphere*)sph + n);
}
// spheres, materials, points, triangles, objects, nodes, textures
void b200host_counts(void* b, uint32_t* out7) {
    SceneBuilder* s = (SceneBuilder*)b;
    out7[0] = (uint32_t)s->spheres.size(); out7[1] = (uint32_t)s->materials.size(); out7[2] = (uint32_t)s->points.size();
    out7[3] = (uint32_t)s->triangles.size(); out7[4] = (uint32_t)s->objects.size(); out7[5] = (uint32_t)s->nodes.size();
    out7[6] = (uint32_t)s->texturePaths.size();
}
void b200host_copy(void* b, int which, void* dst) {
    SceneBuilder* s = (SceneBuilder*)b;
    switch (which) {
        case 0: std::memcpy(dst, s->spheres.data(), s->spheres.size() * 32); break;
        case 1: std::memcpy(dst, s->materials.data(), s->materials.size() * 64); break;
        case 2: std::memcpy(dst, s->points.data(), s->points.size() * 32); break;
        case 3: std::memcpy(dst, s->triangles.data(), s->triangles.size() * 48); break;
        case 4: std::memcpy(dst, s->objects.data(), s->objects.size() * 80); break;
        case 5: std::memcpy(dst, s->nodes.data(), s->nodes.size() * 32); break;
    }
}
int b200host_texture_path(void* b, uint32_t slot, char* dst, int cap) {
    SceneBuilder* s = (SceneBuilder*)b;
    if (slot >= s->texturePaths.size()) return -1;
    const std::string& p = s->texturePaths[slot];
    if (dst && cap > 0) { std::strncpy(dst, p.c_str(), cap - 1); dst[cap - 1] = 0; }
    return (int)p.size();
}
void b200host_object_matrix(const float* pos, const float* rot, const float* scl, float* out16) {
    SceneBuilder::object_matrix(mk(pos, rot, scl, 0, 0), out16);
}

}  // extern "C"
