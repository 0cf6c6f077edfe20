// C ABI of the backend (include/b200pt.h). Host-side plumbing only: device buffers, the upload
// boundary that replaces prepare_storage_buffers/copy_buffer/update_buffer
// (src/vk_engine.cpp:680-758,1401-1475) and the dispatch boundary that replaces run_compute
// (src/vk_engine.cpp:1623-1676). All arithmetic of the pass runs in the CUDA kernels; there is no
// CPU path.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>

#include "../../include/b200pt.h"
#include <cmath>

#include "pt_bvh8.h"
#include "pt_context.h"
#include "wavefront.h"

using namespace pt;

static thread_local std::string g_error;
namespace pt { void set_global_error(const char* what) { g_error = what; } }     // bvh2_build.cu

static int fail(b200pt_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->error = buf; else g_error = buf;
    return code;
}
#define CU(c, call)                                                                                  \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) return fail(c, B200PT_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

static cudaError_t dev_resize(DeviceBuffer& b, size_t bytes) {
    if (bytes <= b.bytes && b.p) return cudaSuccess;
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.bytes = 0;
    if (bytes == 0) return cudaSuccess;
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e == cudaSuccess) b.bytes = bytes;
    return e;
}
static void dev_free(DeviceBuffer& b) { if (b.p) cudaFree(b.p); b.p = nullptr; b.bytes = 0; }

static int alloc_outputs(b200pt_ctx* c) {
    size_t px = (size_t)c->W * c->H;
    CU(c, dev_resize(c->dAccumOwn, px * 16));
    CU(c, dev_resize(c->dImage, px * 16));
    CU(c, dev_resize(c->dStats, px * 8));
    c->accum = (float4*)c->dAccumOwn.p;
    CU(c, cudaMemsetAsync(c->dAccumOwn.p, 0, px * 16, c->stream));
    CU(c, cudaMemsetAsync(c->dImage.p, 0, px * 16, c->stream));
    CU(c, cudaMemsetAsync(c->dStats.p, 0, px * 8, c->stream));
    return 0;
}

// inverse(object.transformMatrix) for every object, evaluated once per upload with the same
// operation order the shader's inverse() has in the oracle (raytrace.comp:292-293, A-Q13)
static int upload_objects(b200pt_ctx* c) {
    size_t n = c->hObjects.size();
    std::vector<float> minv(16 * n);
    for (size_t i = 0; i < n; ++i) inverse4(c->hObjects[i].m, &minv[16 * i]);
    CU(c, dev_resize(c->dObjects, n * sizeof(ptl::RenderObject)));
    CU(c, dev_resize(c->dMinv, n * 64));
    if (n) {
        CU(c, cudaMemcpyAsync(c->dObjects.p, c->hObjects.data(), n * sizeof(ptl::RenderObject), cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(c->dMinv.p, minv.data(), n * 64, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));      // minv is a local
    }
    c->ref.objects = (const ptl::RenderObject*)c->dObjects.p;
    c->ref.minv = (const float*)c->dMinv.p;
    c->ref.nObjects = (uint32_t)n;
    return 0;
}

template <class T>
static int upload_vec(b200pt_ctx* c, DeviceBuffer& d, const std::vector<T>& h) {
    CU(c, dev_resize(d, h.size() * sizeof(T)));
    if (!h.empty()) CU(c, cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return 0;
}

// indices the kernels dereference are validated once here, so a malformed scene is an error
// return instead of an illegal address on the device
static const char* validate_scene(const b200pt_ctx* c) {
    size_t nM = c->hMaterials.size(), nP = c->hPoints.size(), nT = c->hTris.size(), nN = c->hNodes.size();
    if (nM == 0) return "scene has no materials (materials[0] is read on a miss)";
    for (const auto& s : c->hSpheres) if (s.materialIndex >= nM) return "sphere materialIndex out of range";
    for (const auto& t : c->hTris) if (t.v0 >= nP || t.v1 >= nP || t.v2 >= nP) return "triangle vertex index out of range";
    for (const auto& o : c->hObjects) {
        if (o.materialIndex >= nM) return "object materialIndex out of range";
        if (o.bvhIndex >= nN) return "object bvhIndex out of range";
    }
    for (const auto& n : c->hNodes) {
        if (n.triCount) { if ((size_t)n.index + n.triCount > nT) return "BVH leaf triangle range out of range"; }
        else if ((size_t)n.index + 1 >= nN) return "BVH child index out of range";
    }
    return nullptr;
}

extern "C" {

int b200pt_version(void) { return 100; }
const char* b200pt_global_error(void) { return g_error.c_str(); }
const char* b200pt_last_error(b200pt_ctx* c) { return c ? c->error.c_str() : g_error.c_str(); }

int b200pt_create(int device, int width, int height, b200pt_ctx** out) {
    if (!out) return fail(nullptr, B200PT_ERR_ARG, "b200pt_create: out is NULL");
    *out = nullptr;
    if (width <= 0 || height <= 0 || (int64_t)width * height > (int64_t)1 << 28)
        return fail(nullptr, B200PT_ERR_ARG, "b200pt_create: bad size %dx%d", width, height);
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, B200PT_ERR_CUDA, "b200pt_create: no CUDA device (%s); this backend has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (device < 0 || device >= count) return fail(nullptr, B200PT_ERR_ARG, "b200pt_create: device %d of %d", device, count);
    b200pt_ctx* c = new (std::nothrow) b200pt_ctx();
    if (!c) return fail(nullptr, B200PT_ERR_STATE, "out of host memory");
    c->device = device; c->W = width; c->H = height;
#define CUC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(nullptr, B200PT_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); delete c; return B200PT_ERR_CUDA; } } while (0)
    CUC(cudaSetDevice(device));
    CUC(cudaStreamCreateWithFlags(&c->ownStream, cudaStreamNonBlocking));
    c->stream = c->ownStream;
    CUC(cudaEventCreate(&c->evStart));
    CUC(cudaEventCreate(&c->evStop));
#undef CUC
    if (int rc = alloc_outputs(c)) { g_error = c->error; b200pt_destroy(c); return rc; }
    *out = c;
    return 0;
}

static void free_texture(b200pt_ctx::Texture& t);

int b200pt_destroy(b200pt_ctx* c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    b200pt_comm_destroy(c);
    wavefront_destroy(c);
    dev_free(c->dSpheres); dev_free(c->dMaterials); dev_free(c->dPoints); dev_free(c->dTris);
    dev_free(c->dObjects); dev_free(c->dNodes); dev_free(c->dMinv);
    dev_free(c->dAccumOwn); dev_free(c->dImage); dev_free(c->dStats);
    for (auto& t : c->textures) free_texture(t);
    dev_free(c->dTexTable); dev_free(c->dSrgbLut);
    if (c->evStart) cudaEventDestroy(c->evStart);
    if (c->evStop) cudaEventDestroy(c->evStop);
    if (c->ownStream) cudaStreamDestroy(c->ownStream);
    delete c;
    return 0;
}

int b200pt_resize(b200pt_ctx* c, int width, int height) {
    if (!c) return B200PT_ERR_ARG;
    if (width <= 0 || height <= 0 || (int64_t)width * height > (int64_t)1 << 28) return fail(c, B200PT_ERR_ARG, "bad size %dx%d", width, height);
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    bool attached = c->accum != (float4*)c->dAccumOwn.p;
    if (attached) return fail(c, B200PT_ERR_STATE, "detach the external accumulator before resizing");
    c->W = width; c->H = height;
    wavefront_invalidate_frame_state(c);
    return alloc_outputs(c);
}

int b200pt_set_stream(b200pt_ctx* c, void* s) {
    if (!c) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    c->stream = s ? (cudaStream_t)s : c->ownStream;
    return 0;
}

int b200pt_sync(b200pt_ctx* c) {
    if (!c) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int b200pt_upload_scene(b200pt_ctx* c, const void* spheres, uint32_t nS, const void* materials, uint32_t nM,
                        const void* points, uint32_t nP, const void* tris, uint32_t nT, const void* objects, uint32_t nO,
                        const void* nodes, uint32_t nN) {
    if (!c) return B200PT_ERR_ARG;
    if ((nS && !spheres) || (nM && !materials) || (nP && !points) || (nT && !tris) || (nO && !objects) || (nN && !nodes))
        return fail(c, B200PT_ERR_ARG, "b200pt_upload_scene: NULL buffer with non-zero count");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    c->sceneLoaded = false;
    c->hSpheres.assign((const ptl::Sphere*)spheres, (const ptl::Sphere*)spheres + nS);
    c->hMaterials.assign((const ptl::Material*)materials, (const ptl::Material*)materials + nM);
    c->hPoints.assign((const ptl::TrianglePoint*)points, (const ptl::TrianglePoint*)points + nP);
    c->hTris.assign((const ptl::Triangle*)tris, (const ptl::Triangle*)tris + nT);
    c->hObjects.assign((const ptl::RenderObject*)objects, (const ptl::RenderObject*)objects + nO);
    c->hNodes.assign((const ptl::BVHNode*)nodes, (const ptl::BVHNode*)nodes + nN);
    if (const char* why = validate_scene(c)) return fail(c, B200PT_ERR_ARG, "b200pt_upload_scene: %s", why);
    if (int rc = upload_vec(c, c->dSpheres, c->hSpheres)) return rc;
    if (int rc = upload_vec(c, c->dMaterials, c->hMaterials)) return rc;
    if (int rc = upload_vec(c, c->dPoints, c->hPoints)) return rc;
    if (int rc = upload_vec(c, c->dTris, c->hTris)) return rc;
    if (int rc = upload_vec(c, c->dNodes, c->hNodes)) return rc;
    if (int rc = upload_objects(c)) return rc;
    c->ref.spheres = (const ptl::Sphere*)c->dSpheres.p; c->ref.nSpheres = nS;
    c->ref.materials = (const ptl::Material*)c->dMaterials.p; c->ref.nMaterials = nM;
    c->ref.points = (const ptl::TrianglePoint*)c->dPoints.p; c->ref.nPoints = nP;
    c->ref.tris = (const ptl::Triangle*)c->dTris.p; c->ref.nTris = nT;
    c->ref.nodes = (const ptl::BVHNode*)c->dNodes.p; c->ref.nNodes = nN;
    CU(c, cudaStreamSynchronize(c->stream));
    if (int rc = wavefront_build_scene(c)) return rc;
    c->sceneLoaded = true;
    return 0;
}

int b200pt_update_materials(b200pt_ctx* c, const void* materials, uint32_t n) {
    if (!c || !materials) return fail(c, B200PT_ERR_ARG, "b200pt_update_materials: NULL");
    if (!c->sceneLoaded || n != c->hMaterials.size()) return fail(c, B200PT_ERR_STATE, "update_materials: count %u differs from the uploaded %zu", n, c->hMaterials.size());
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    std::vector<ptl::Material> old = c->hMaterials;
    c->hMaterials.assign((const ptl::Material*)materials, (const ptl::Material*)materials + n);
    if (const char* why = validate_scene(c)) { c->hMaterials = old; return fail(c, B200PT_ERR_ARG, "update_materials: %s", why); }
    if (int rc = upload_vec(c, c->dMaterials, c->hMaterials)) return rc;
    CU(c, cudaStreamSynchronize(c->stream));
    return wavefront_build_scene(c, pt::WF_CHANGED_MATERIALS);      // emissive/non-emissive split depends on the materials
}

int b200pt_update_spheres(b200pt_ctx* c, const void* spheres, uint32_t n) {
    if (!c || !spheres) return fail(c, B200PT_ERR_ARG, "b200pt_update_spheres: NULL");
    if (!c->sceneLoaded || n != c->hSpheres.size()) return fail(c, B200PT_ERR_STATE, "update_spheres: count %u differs from the uploaded %zu", n, c->hSpheres.size());
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    std::vector<ptl::Sphere> old = c->hSpheres;
    c->hSpheres.assign((const ptl::Sphere*)spheres, (const ptl::Sphere*)spheres + n);
    if (const char* why = validate_scene(c)) { c->hSpheres = old; return fail(c, B200PT_ERR_ARG, "update_spheres: %s", why); }
    if (int rc = upload_vec(c, c->dSpheres, c->hSpheres)) return rc;
    CU(c, cudaStreamSynchronize(c->stream));
    return wavefront_build_scene(c, pt::WF_CHANGED_SPHERES);
}

int b200pt_update_objects(b200pt_ctx* c, const void* objects, uint32_t n) {
    if (!c || !objects) return fail(c, B200PT_ERR_ARG, "b200pt_update_objects: NULL");
    if (!c->sceneLoaded || n != c->hObjects.size()) return fail(c, B200PT_ERR_STATE, "update_objects: count %u differs from the uploaded %zu", n, c->hObjects.size());
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    std::vector<ptl::RenderObject> old = c->hObjects;
    c->hObjects.assign((const ptl::RenderObject*)objects, (const ptl::RenderObject*)objects + n);
    if (const char* why = validate_scene(c)) { c->hObjects = old; return fail(c, B200PT_ERR_ARG, "update_objects: %s", why); }
    if (int rc = upload_objects(c)) return rc;
    // world-space boxes move with the transforms. When nothing but matrices changed the device-built tree is refitted
    // (same topology, boxes recomputed bottom-up); anything else re-lays the scene out.
    bool onlyMatrices = true;
    for (size_t i = 0; i < old.size() && onlyMatrices; ++i) {
        const ptl::RenderObject &a = old[i], &b = c->hObjects[i];
        onlyMatrices = a.smoothShade == b.smoothShade && a.bvhIndex == b.bvhIndex && a.materialIndex == b.materialIndex && a.samplerIndex == b.samplerIndex;
    }
    return wavefront_build_scene(c, onlyMatrices ? pt::WF_CHANGED_TRANSFORMS : pt::WF_CHANGED_ALL);
}

static int upload_texture_table(b200pt_ctx* c) {
    pt::TexEntry table[B200PT_MAX_TEXTURES];
    for (int i = 0; i < B200PT_MAX_TEXTURES; ++i) {
        table[i].object = (unsigned long long)c->textures[i].object; table[i].w = c->textures[i].w; table[i].h = c->textures[i].h;
    }
    CU(c, dev_resize(c->dTexTable, sizeof(table)));
    CU(c, cudaMemcpyAsync(c->dTexTable.p, table, sizeof(table), cudaMemcpyHostToDevice, c->stream));
    if (!c->dSrgbLut.p) {
        float lut[256];
        for (int i = 0; i < 256; ++i) {       // IEC 61966-2-1 decoding, evaluated in double and rounded once
            double x = double(i) / 255.0;
            lut[i] = float(x <= 0.04045 ? x / 12.92 : std::pow((x + 0.055) / 1.055, 2.4));
        }
        CU(c, dev_resize(c->dSrgbLut, sizeof(lut)));
        CU(c, cudaMemcpyAsync(c->dSrgbLut.p, lut, sizeof(lut), cudaMemcpyHostToDevice, c->stream));
    }
    CU(c, cudaStreamSynchronize(c->stream));
    wavefront_set_textures(c);
    return 0;
}
static void free_texture(b200pt_ctx::Texture& t) {
    if (t.object) cudaDestroyTextureObject(t.object);
    if (t.array) cudaFreeArray(t.array);
    t.object = 0; t.array = nullptr; t.w = t.h = 0;
}

int b200pt_upload_texture(b200pt_ctx* c, uint32_t slot, const void* rgba8, uint32_t w, uint32_t h) {
    if (!c || !rgba8 || slot >= B200PT_MAX_TEXTURES || w == 0 || h == 0 || w > 32768 || h > 32768)
        return fail(c, B200PT_ERR_ARG, "b200pt_upload_texture: bad argument (slot %u, %ux%u)", slot, w, h);
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    auto& t = c->textures[slot];
    free_texture(t);
    cudaChannelFormatDesc fmt = cudaCreateChannelDesc<uchar4>();
    CU(c, cudaMallocArray(&t.array, &fmt, w, h));
    CU(c, cudaMemcpy2DToArray(t.array, 0, 0, rgba8, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyHostToDevice));
    cudaResourceDesc res;
    std::memset(&res, 0, sizeof(res));
    res.resType = cudaResourceTypeArray; res.res.array.array = t.array;
    cudaTextureDesc td;
    std::memset(&td, 0, sizeof(td));
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;     // wrap/clamp is applied to the texel index by the kernel
    td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    CU(c, cudaCreateTextureObject(&t.object, &res, &td, nullptr));
    t.w = w; t.h = h;
    return upload_texture_table(c);
}

static int check_render(b200pt_ctx* c, const void* pc208, ptl::PushConstants& pc) {
    if (!c || !pc208) return fail(c, B200PT_ERR_ARG, "render: NULL argument");
    if (!c->sceneLoaded) return fail(c, B200PT_ERR_STATE, "render: no scene uploaded");
    std::memcpy(&pc, pc208, sizeof(pc));
    if (pc.params.sphereCount > c->hSpheres.size()) return fail(c, B200PT_ERR_ARG, "render: sphereCount %u > uploaded %zu", pc.params.sphereCount, c->hSpheres.size());
    if (pc.params.objectCount > c->hObjects.size()) return fail(c, B200PT_ERR_ARG, "render: objectCount %u > uploaded %zu", pc.params.objectCount, c->hObjects.size());
    uint32_t samples = pc.params.singleRender ? pc.params.sampleLimit : pc.params.raysPerPixel;
    if (samples > (1u << 20)) return fail(c, B200PT_ERR_ARG, "render: %u samples per dispatch", samples);
    if (pc.params.bounceLimit > 4096) return fail(c, B200PT_ERR_ARG, "render: bounceLimit %u", pc.params.bounceLimit);
    return 0;
}

int b200pt_render(b200pt_ctx* c, const void* pc208) {
    ptl::PushConstants pc;
    if (int rc = check_render(c, pc208, pc)) return rc;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaEventRecord(c->evStart, c->stream));
    // dispatches without a debug heat-map go through the wavefront kernels (same image; the box/triangle test
    // counters of b200pt_read_stats exist only on the reference-order kernel: pipeline 0 or a debug mode)
    const uint64_t samples = pc.params.singleRender ? pc.params.sampleLimit : pc.params.raysPerPixel;
    const bool heatmap = pc.params.debug >= 0 && pc.params.debug <= 2;
    if (c->pipeline == 1 && samples >= 1 && !heatmap) {
        CU(c, cudaMemsetAsync(c->dStats.p, 0, (size_t)c->W * c->H * 8, c->stream));
        if (int rc = wavefront_render_frames(c, pc, pc.frameCount, 1, 1, 0, (uint32_t)c->H, (float4*)c->dImage.p)) return rc;
    } else {
        CU(c, launch_ref_dispatch(c, pc));
    }
    CU(c, cudaEventRecord(c->evStop, c->stream));
    return 0;
}

int b200pt_render_frames_rows_strided(b200pt_ctx* c, const void* pc208, uint32_t first, uint32_t n, uint32_t stride,
                                      uint32_t row0, uint32_t row1, uint32_t rowStride) {
    ptl::PushConstants pc;
    if (int rc = check_render(c, pc208, pc)) return rc;
    if (stride == 0) stride = 1;
    if (rowStride == 0) rowStride = 1;
    if (row1 > (uint32_t)c->H || row0 > row1) return fail(c, B200PT_ERR_ARG, "render_frames: rows [%u,%u) of %d", row0, row1, c->H);
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaEventRecord(c->evStart, c->stream));
    if (c->pipeline == 0) {
        if (rowStride != 1) return fail(c, B200PT_ERR_STATE, "render_frames: interleaved rows need the wavefront pipeline");
        CU(c, launch_ref_frames(c, pc, first, n, stride, row0, row1));
    } else {
        if (int rc = wavefront_render_frames(c, pc, first, n, stride, row0, row1, nullptr, rowStride)) return rc;
    }
    CU(c, cudaEventRecord(c->evStop, c->stream));
    return 0;
}
int b200pt_render_frames_rows(b200pt_ctx* c, const void* pc208, uint32_t first, uint32_t n, uint32_t stride,
                              uint32_t row0, uint32_t row1) {
    return b200pt_render_frames_rows_strided(c, pc208, first, n, stride, row0, row1, 1);
}

int b200pt_render_frames(b200pt_ctx* c, const void* pc208, uint32_t first, uint32_t n, uint32_t stride) {
    if (!c) return B200PT_ERR_ARG;
    return b200pt_render_frames_rows(c, pc208, first, n, stride, 0, (uint32_t)c->H);
}

int b200pt_clear(b200pt_ctx* c) {
    if (!c) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    size_t px = (size_t)c->W * c->H;
    CU(c, cudaMemsetAsync(c->accum, 0, px * 16, c->stream));
    CU(c, cudaMemsetAsync(c->dImage.p, 0, px * 16, c->stream));
    CU(c, cudaMemsetAsync(c->dStats.p, 0, px * 8, c->stream));
    std::memset(c->counters, 0, sizeof(c->counters));
    return 0;
}

static int read_back(b200pt_ctx* c, void* host, const void* dev, size_t bytes) {
    if (!c || !host) return fail(c, B200PT_ERR_ARG, "read: NULL");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return 0;
}
int b200pt_read_accum_f32(b200pt_ctx* c, float* host) { return c ? read_back(c, host, c->accum, (size_t)c->W * c->H * 16) : B200PT_ERR_ARG; }
int b200pt_read_image_f32(b200pt_ctx* c, float* host) { return c ? read_back(c, host, c->dImage.p, (size_t)c->W * c->H * 16) : B200PT_ERR_ARG; }
int b200pt_read_stats(b200pt_ctx* c, float* host) { return c ? read_back(c, host, c->dStats.p, (size_t)c->W * c->H * 8) : B200PT_ERR_ARG; }
int b200pt_write_image_f32(b200pt_ctx* c, const float* host) {
    if (!c || !host) return fail(c, B200PT_ERR_ARG, "write_image: NULL");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(c->dImage.p, host, (size_t)c->W * c->H * 16, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return 0;
}
int b200pt_read_rgba8(b200pt_ctx* c, uint8_t* host) {
    if (!c || !host) return fail(c, B200PT_ERR_ARG, "read_rgba8: NULL");
    CU(c, cudaSetDevice(c->device));
    size_t px = (size_t)c->W * c->H;
    DeviceBuffer tmp;
    CU(c, dev_resize(tmp, px * 4));
    cudaError_t e = launch_image_to_rgba8(c, (const float4*)c->dImage.p, (uchar4*)tmp.p, px);
    if (e == cudaSuccess) e = cudaMemcpyAsync(host, tmp.p, px * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    dev_free(tmp);
    CU(c, e);
    return 0;
}
int b200pt_device_accum_ptr(b200pt_ctx* c, void** p) { if (!c || !p) return B200PT_ERR_ARG; *p = c->accum; return 0; }
int b200pt_device_image_ptr(b200pt_ctx* c, void** p) { if (!c || !p) return B200PT_ERR_ARG; *p = c->dImage.p; return 0; }
int b200pt_attach_accum(b200pt_ctx* c, void* p) {
    if (!c) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    c->accum = p ? (float4*)p : (float4*)c->dAccumOwn.p;
    return 0;
}

int b200pt_trace_batch(b200pt_ctx* c, const void* pc208, const float* rays, uint64_t n, void* hits) {
    ptl::PushConstants pc;
    if (int rc = check_render(c, pc208, pc)) return rc;
    if (n && (!rays || !hits)) return fail(c, B200PT_ERR_ARG, "trace_batch: NULL buffer");
    if (n == 0) return 0;
    CU(c, cudaSetDevice(c->device));
    DeviceBuffer dR, dH;
    cudaError_t e = dev_resize(dR, n * 32);
    if (e == cudaSuccess) e = dev_resize(dH, n * 48);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dR.p, rays, n * 32, cudaMemcpyHostToDevice, c->stream);
    int rc = 0;
    if (e == cudaSuccess) {
        if (c->pipeline == 0) e = launch_ref_trace_batch(c, pc, (const float*)dR.p, n, (HitRecord48*)dH.p);
        else rc = wavefront_trace_batch(c, pc, (const float*)dR.p, n, (HitRecord48*)dH.p);
    }
    if (e == cudaSuccess && rc == 0) e = cudaMemcpyAsync(hits, dH.p, n * 48, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    dev_free(dR); dev_free(dH);
    if (rc) return rc;
    CU(c, e);
    return 0;
}

int b200pt_trace_batch_device(b200pt_ctx* c, const void* pc208, const void* dRays, uint64_t n, void* dHits) {
    ptl::PushConstants pc;
    if (int rc = check_render(c, pc208, pc)) return rc;
    if (n && (!dRays || !dHits)) return fail(c, B200PT_ERR_ARG, "trace_batch_device: NULL buffer");
    if (n == 0) return 0;
    CU(c, cudaSetDevice(c->device));
    if (c->pipeline == 0) { CU(c, launch_ref_trace_batch(c, pc, (const float*)dRays, n, (HitRecord48*)dHits)); return 0; }
    return wavefront_trace_batch(c, pc, (const float*)dRays, n, (HitRecord48*)dHits);
}

int b200pt_trace_batch_device_static(b200pt_ctx* c, const void* pc208, const void* dRays, uint64_t n, void* dHits) {
    ptl::PushConstants pc;
    if (int rc = check_render(c, pc208, pc)) return rc;
    if (n && (!dRays || !dHits)) return fail(c, B200PT_ERR_ARG, "trace_batch_device_static: NULL buffer");
    if (n > B200PT_STATIC_MAX_RAYS)
        return fail(c, B200PT_ERR_ARG, "trace_batch_device_static: %llu rays exceed the static scratch (B200PT_STATIC_MAX_RAYS = %u); split the batch or use b200pt_trace_batch_device",
                    (unsigned long long)n, (unsigned)B200PT_STATIC_MAX_RAYS);
    if (n == 0) return 0;
    CU(c, cudaSetDevice(c->device));
    return wavefront_trace_batch_static(c, pc, (const float*)dRays, n, (HitRecord48*)dHits);
}

int b200pt_set_option(b200pt_ctx* c, const char* key, int64_t v) {
    if (!c || !key) return B200PT_ERR_ARG;
    std::string k(key);
    if (k == "pipeline") { if (v < 0 || v > 1) return fail(c, B200PT_ERR_ARG, "pipeline must be 0 or 1"); c->pipeline = (int)v; return 0; }
    if (k == "frames_in_flight") { if (v < 1 || v > 64) return fail(c, B200PT_ERR_ARG, "frames_in_flight must be 1..64"); c->framesInFlight = (int)v; return 0; }
    if (k == "profile_stages") { c->profileStages = v != 0; return 0; }
    if (k == "count_work") { c->countWork = v != 0; return 0; }
    if (k == "textures") { c->useTextures = v != 0; return 0; }
    if (k == "storage_rgba8") { c->storageRgba8 = v != 0; return 0; }
    if (k == "trav_tri_vote") { if (v < 1 || v > 32) return fail(c, B200PT_ERR_ARG, "trav_tri_vote must be 1..32"); c->triVote32 = (int)v; return 0; }
    if (k == "trav_fetch_idle") { if (v < 1 || v > 32) return fail(c, B200PT_ERR_ARG, "trav_fetch_idle must be 1..32"); c->fetchMinIdle = (int)v; return 0; }
    if (k == "occl_tri_vote") { if (v < 1 || v > 32) return fail(c, B200PT_ERR_ARG, "occl_tri_vote must be 1..32"); c->triVoteOccl32 = (int)v; return 0; }
    if (k == "occl_fetch_idle") { if (v < 1 || v > 32) return fail(c, B200PT_ERR_ARG, "occl_fetch_idle must be 1..32"); c->fetchMinIdleOccl = (int)v; return 0; }
    if (k == "share_primary") { c->sharePrimary = v != 0; return 0; }
    if (k == "overlap_connect") { if (v < 0 || v > 2) return fail(c, B200PT_ERR_ARG, "overlap_connect must be 0..2"); c->overlapConnect = (int)v; return 0; }
    if (k == "wide_build") { if (v < 0 || v > 1) return fail(c, B200PT_ERR_ARG, "wide_build must be 0 (host) or 1 (device)"); c->wideBuild = (int)v; return 0; }
    if (k == "refit_on_update") { c->refitOnUpdate = v != 0; return 0; }
    if (k == "exact_order") { c->exactOrder = v != 0; return 0; }
    return fail(c, B200PT_ERR_ARG, "unknown option '%s'", key);
}
int b200pt_get_option(b200pt_ctx* c, const char* key, int64_t* v) {
    if (!c || !key || !v) return B200PT_ERR_ARG;
    std::string k(key);
    if (k == "pipeline") { *v = c->pipeline; return 0; }
    if (k == "frames_in_flight") { *v = c->framesInFlight; return 0; }
    if (k == "profile_stages") { *v = c->profileStages; return 0; }
    if (k == "count_work") { *v = c->countWork; return 0; }
    if (k == "textures") { *v = c->useTextures; return 0; }
    if (k == "storage_rgba8") { *v = c->storageRgba8; return 0; }
    if (k == "trav_tri_vote") { *v = c->triVote32; return 0; }
    if (k == "trav_fetch_idle") { *v = c->fetchMinIdle; return 0; }
    if (k == "occl_tri_vote") { *v = c->triVoteOccl32; return 0; }
    if (k == "occl_fetch_idle") { *v = c->fetchMinIdleOccl; return 0; }
    if (k == "share_primary") { *v = c->sharePrimary; return 0; }
    if (k == "overlap_connect") { *v = c->overlapConnect; return 0; }
    if (k == "wide_build") { *v = c->wideBuild; return 0; }
    if (k == "refit_on_update") { *v = c->refitOnUpdate; return 0; }
    if (k == "exact_order") { *v = c->exactOrder; return 0; }
    if (k == "last_update_was_refit") { *v = c->lastRefit; return 0; }                               // read-only
    if (k == "last_scene_build_us") { *v = (int64_t)(c->lastSceneBuildMs * 1000.0); return 0; }      // read-only
    if (k == "last_wide_device_us") { *v = (int64_t)(c->lastWideDeviceMs * 1000.0); return 0; }      // read-only
    return fail(c, B200PT_ERR_ARG, "unknown option '%s'", key);
}

int b200pt_read_counters(b200pt_ctx* c, uint64_t* out16) {
    if (!c || !out16) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    wavefront_fold_counters(c);
    std::memcpy(out16, c->counters, sizeof(c->counters));
    return 0;
}
int b200pt_reset_counters(b200pt_ctx* c) {
    if (!c) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    wavefront_fold_counters(c);
    std::memset(c->counters, 0, sizeof(c->counters));
    return 0;
}
int b200pt_read_stage_times(b200pt_ctx* c, double* ms8, uint64_t* launches8, int reset) {
    if (!c || !ms8 || !launches8) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    wavefront_resolve_stage_times(c);
    wavefront_stage_times(c, ms8, launches8, reset != 0);
    return 0;
}
int b200pt_last_render_ms(b200pt_ctx* c, float* ms) {
    if (!c || !ms) return B200PT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaEventSynchronize(c->evStop));
    CU(c, cudaEventElapsedTime(ms, c->evStart, c->evStop));
    return 0;
}

}  // extern "C"
