// C entry points of the host-side scene builder (libb200pt_host.so), for the Python bench/tests and
// for any C caller that wants reference-layout buffers from OBJ/MTL files or procedural meshes.
#include <cstdio>
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

// ---- on-disk scene cache: the six storage buffers of prepare_storage_buffers() (src/vk_engine.cpp:680-758)
// exactly as they are uploaded, so a large OBJ is parsed and its BVH built once. Layout, little endian:
//   0   char[8]  "B2PTSCN1"
//   8   u32      version (1), u32 reserved
//   16  u64[6]   record counts: spheres, materials, triPoints, triangles, objects, bvhNodes
//   64  u32[6]   record sizes: 32, 64, 32, 48, 80, 32
//   88  u64      FNV-1a 64 of the payload
//   96  ... 128  zero
//   128 payload: the six arrays in that order, each padded to a multiple of 64 bytes
static const uint32_t kRecord[6] = {32, 64, 32, 48, 80, 32};
static uint64_t fnv1a(uint64_t h, const unsigned char* p, size_t n) {
    for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 1099511628211ull; }
    return h;
}
static size_t pad64(size_t n) { return (n + 63) & ~(size_t)63; }

// 0 ok, 1 cannot open/write
int b200host_scene_write(const char* path, const void* const* blobs6, const uint64_t* counts6) {
    unsigned char head[128] = {0};
    std::memcpy(head, "B2PTSCN1", 8);
    uint32_t version = 1;
    std::memcpy(head + 8, &version, 4);
    std::memcpy(head + 16, counts6, 48);
    std::memcpy(head + 64, kRecord, 24);
    uint64_t h = 14695981039346656037ull;
    for (int i = 0; i < 6; ++i) h = fnv1a(h, (const unsigned char*)blobs6[i], (size_t)counts6[i] * kRecord[i]);
    std::memcpy(head + 88, &h, 8);
    FILE* f = std::fopen(path, "wb");
    if (!f) return 1;
    bool ok = std::fwrite(head, 1, 128, f) == 128;
    static const unsigned char zeros[64] = {0};
    for (int i = 0; i < 6 && ok; ++i) {
        size_t n = (size_t)counts6[i] * kRecord[i];
        ok = (n == 0 || std::fwrite(blobs6[i], 1, n, f) == n) && std::fwrite(zeros, 1, pad64(n) - n, f) == pad64(n) - n;
    }
    ok = std::fclose(f) == 0 && ok;
    return ok ? 0 : 1;
}
// Reads the header: 0 ok, 1 cannot open, 2 not a scene file / unsupported version / record sizes differ, 3 truncated
int b200host_scene_probe(const char* path, uint64_t* counts6) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return 1;
    unsigned char head[128];
    size_t got = std::fread(head, 1, 128, f);
    int rc = 0;
    uint32_t version = 0;
    if (got != 128 || std::memcmp(head, "B2PTSCN1", 8) != 0) rc = 2;
    else {
        std::memcpy(&version, head + 8, 4);
        if (version != 1 || std::memcmp(head + 64, kRecord, 24) != 0) rc = 2;
    }
    if (!rc) {
        std::memcpy(counts6, head + 16, 48);
        uint64_t need = 128;
        for (int i = 0; i < 6 && !rc; ++i) {
            if (counts6[i] > ((uint64_t)1 << 40) / kRecord[i]) rc = 2;         // a byte size that cannot be a real scene (and cannot wrap below)
            else need += pad64((size_t)counts6[i] * kRecord[i]);
        }
        long end = -1;
        if (!rc && (std::fseek(f, 0, SEEK_END) != 0 || (end = std::ftell(f)) < 0 || (uint64_t)end < need)) rc = 3;
    }
    std::fclose(f);
    return rc;
}
// Fills six caller-allocated arrays whose capacities (records) are `counts6` as b200host_scene_probe returned them; the
// file's header is read again and must still say the same — a file replaced between the two calls is refused before
// anything is written: 0 ok, 1 cannot open, 2 not a scene file, 3 truncated, 4 checksum mismatch, 5 counts changed
int b200host_scene_read_checked(const char* path, void* const* blobs6, const uint64_t* counts6) {
    uint64_t counts[6];
    if (int rc = b200host_scene_probe(path, counts)) return rc;
    if (counts6 && std::memcmp(counts, counts6, sizeof(counts)) != 0) return 5;
    FILE* f = std::fopen(path, "rb");
    if (!f) return 1;
    unsigned char head[128];
    int rc = std::fread(head, 1, 128, f) == 128 ? 0 : 3;
    if (!rc && std::memcmp(head + 16, counts, 48) != 0) rc = 5;      // this open sees another file than the probe did
    uint64_t h = 14695981039346656037ull, want = 0;
    std::memcpy(&want, head + 88, 8);
    for (int i = 0; i < 6 && !rc; ++i) {
        size_t n = (size_t)counts[i] * kRecord[i];
        if (n && std::fread(blobs6[i], 1, n, f) != n) rc = 3;
        else if (std::fseek(f, (long)(pad64(n) - n), SEEK_CUR) != 0) rc = 3;
        else h = fnv1a(h, (const unsigned char*)blobs6[i], n);
    }
    std::fclose(f);
    if (!rc && h != want) rc = 4;
    return rc;
}
int b200host_scene_read(const char* path, void* const* blobs6) { return b200host_scene_read_checked(path, blobs6, nullptr); }

}  // extern "C"
