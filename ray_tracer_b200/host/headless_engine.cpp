// See headless_engine.h. Compile with -ffp-contract=off: the camera matrix must be the one GLM builds on the host.
#include "headless_engine.h"

#include <cmath>
#include <cstdio>
#include <cstring>

#include "../../include/b200pt.h"

extern "C" {   // on-disk scene cache, host_api.cpp
int b200host_scene_probe(const char* path, uint64_t* counts6);
int b200host_scene_read_checked(const char* path, void* const* blobs6, const uint64_t* counts6);
}

namespace pth {

namespace {
struct M3 { float c[3][3]; };      // columns, as glm::mat3
// glm::mat3 * glm::mat3 (GLM 0.9.9.7 type_mat3x3.inl operator*): each element is ((a0*b0 + a1*b1) + a2*b2)
M3 mul(const M3& a, const M3& b) {
    M3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
    return r;
}
float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }   // glm::radians
}  // namespace

HeadlessEngine::HeadlessEngine() {
    std::memset(&cameraInfo, 0, sizeof cameraInfo);
    std::memset(&environment, 0, sizeof environment);
    std::memset(&rayTracerParams, 0, sizeof rayTracerParams);
    std::memset(&constants, 0, sizeof constants);
    cameraInfo.pos[0] = 0.f; cameraInfo.pos[1] = -0.5f; cameraInfo.pos[2] = -3.5f;
    cameraInfo.nearPlane = 0.1f; cameraInfo.fov = 50.f;
    const float h[4] = {0.986f, 1.f, 0.902f, 1000.f}, z[4] = {0.265f, 0.595f, 0.887f, 10.f};
    std::memcpy(environment.horizon, h, sizeof h);
    std::memcpy(environment.zenith, z, sizeof z);
    environment.ground[0] = environment.ground[1] = environment.ground[2] = 0.431f;
    const float d[3] = {2.f, 0.8f, -3.f};
    const float inv = 1.f / std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);   // glm::normalize = v * inversesqrt(dot(v, v))
    for (int i = 0; i < 3; ++i) environment.lightDir[i] = d[i] * inv;
    environment.lightDir[3] = 0.f;
    rayTracerParams.debug = -1;
    rayTracerParams.raysPerPixel = 1; rayTracerParams.bounceLimit = 8;
    rayTracerParams.triangleCap = 50; rayTracerParams.boxCap = 200; rayTracerParams.sampleLimit = 10;
}

bool HeadlessEngine::fail(const char* what) {
    const std::string msg = ctx_ ? b200pt_last_error(ctx_) : b200pt_global_error();
    error_ = msg.compare(0, std::strlen(what), what) == 0 ? msg : std::string(what) + ": " + msg;
    return false;
}

bool HeadlessEngine::init_image(int device) {
    cleanup();
    if (b200pt_create(device, (int)width, (int)height, &ctx_)) { ctx_ = nullptr; return fail("b200pt_create"); }
    return true;
}

bool HeadlessEngine::load_scene_file(const std::string& path) {
    uint64_t n[6];
    if (int rc = b200host_scene_probe(path.c_str(), n)) { error_ = path + ": scene file rejected (code " + std::to_string(rc) + ")"; return false; }
    scene.spheres.resize(n[0]); scene.materials.resize(n[1]); scene.points.resize(n[2]);
    scene.triangles.resize(n[3]); scene.objects.resize(n[4]); scene.nodes.resize(n[5]);
    void* blobs[6] = {scene.spheres.data(), scene.materials.data(), scene.points.data(), scene.triangles.data(), scene.objects.data(), scene.nodes.data()};
    if (int rc = b200host_scene_read_checked(path.c_str(), blobs, n)) { error_ = path + ": scene file unreadable (code " + std::to_string(rc) + ")"; return false; }
    return true;
}

bool HeadlessEngine::prepare_storage_buffers() {
    if (!ctx_) { error_ = "prepare_storage_buffers before init_image"; return false; }
    if (b200pt_upload_scene(ctx_, scene.spheres.data(), (uint32_t)scene.spheres.size(), scene.materials.data(), (uint32_t)scene.materials.size(),
                            scene.points.data(), (uint32_t)scene.points.size(), scene.triangles.data(), (uint32_t)scene.triangles.size(),
                            scene.objects.data(), (uint32_t)scene.objects.size(), scene.nodes.data(), (uint32_t)scene.nodes.size()))
        return fail("b200pt_upload_scene");
    return true;
}

bool HeadlessEngine::update_buffer(Buffer which) {
    int rc = 0;
    switch (which) {
        case MATERIALS: rc = b200pt_update_materials(ctx_, scene.materials.data(), (uint32_t)scene.materials.size()); break;
        case SPHERES: rc = b200pt_update_spheres(ctx_, scene.spheres.data(), (uint32_t)scene.spheres.size()); break;
        case OBJECTS: rc = b200pt_update_objects(ctx_, scene.objects.data(), (uint32_t)scene.objects.size()); break;
    }
    if (rc) return fail("b200pt_update");
    return true;          // like the UI handlers (src/vk_engine.cpp:1545, 1572, 1603): the accumulation is not restarted
}

void HeadlessEngine::fill_constants() {
    cameraInfo.aspectRatio = width / (float)height;
    const float tx = radians(cameraAngles[0]), ty = radians(cameraAngles[1]), tz = radians(cameraAngles[2]);
    const M3 rotX = {{{1.f, 0.f, 0.f}, {0.f, std::cos(tx), -std::sin(tx)}, {0.f, std::sin(tx), std::cos(tx)}}};
    const M3 rotY = {{{std::cos(ty), 0.f, std::sin(ty)}, {0.f, 1.f, 0.f}, {-std::sin(ty), 0.f, std::cos(ty)}}};
    const M3 rotZ = {{{std::cos(tz), -std::sin(tz), 0.f}, {std::sin(tz), std::cos(tz), 0.f}, {0.f, 0.f, 1.f}}};
    const M3 r = mul(mul(rotY, rotX), rotZ);
    std::memset(cameraInfo.rot, 0, sizeof cameraInfo.rot);            // mat4(mat3): upper-left block + identity corner
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) cameraInfo.rot[4 * j + i] = r.c[j][i];
    cameraInfo.rot[15] = 1.f;
    rayTracerParams.sphereCount = (uint32_t)scene.spheres.size();
    rayTracerParams.objectCount = (uint32_t)scene.objects.size();
    constants.cam = cameraInfo;
    constants.env = environment;
    constants.params = rayTracerParams;
    constants.frameCount = (uint32_t)frameNumber;
}

bool HeadlessEngine::run_compute() {
    fill_constants();
    if (b200pt_render(ctx_, &constants)) return fail("b200pt_render");
    return true;
}

bool HeadlessEngine::draw() {
    const bool active = totalSamples < rayTracerParams.sampleLimit;
    if (active && !run_compute()) return false;
    if (b200pt_sync(ctx_)) return fail("b200pt_sync");               // vkQueueWaitIdle, :1808
    frameNumber = rayTracerParams.progressive ? frameNumber + 1 : 0;
    totalSamples += active ? rayTracerParams.sampleLimit : 0;
    if (!rayTracerParams.singleRender) totalSamples = 0;
    return true;
}

bool HeadlessEngine::render_batch(uint32_t firstFrame, uint32_t nFrames) {
    fill_constants();
    if (b200pt_render_frames(ctx_, &constants, firstFrame, nFrames, 1)) return fail("b200pt_render_frames");
    return true;
}

bool HeadlessEngine::comm_init(int nranks, int rank, const void* uniqueId128) {
    if (b200pt_comm_init_rank(ctx_, nranks, rank, uniqueId128)) return fail("b200pt_comm_init_rank");
    return true;
}
bool HeadlessEngine::render_batch_split(uint32_t firstFrame, uint32_t nFrames) {
    fill_constants();
    if (b200pt_render_frames_split(ctx_, &constants, firstFrame, nFrames)) return fail("b200pt_render_frames_split");
    return true;
}
bool HeadlessEngine::reduce_accum(int root) {
    if (b200pt_reduce_accum(ctx_, root, 1)) return fail("b200pt_reduce_accum");
    return true;
}
bool HeadlessEngine::render_frame_interleaved(uint32_t frame, int root) {
    fill_constants();
    if (b200pt_render_frame_interleaved(ctx_, &constants, frame, root)) return fail("b200pt_render_frame_interleaved");
    return true;
}

bool HeadlessEngine::set_option(const char* key, long long value) {
    if (b200pt_set_option(ctx_, key, (int64_t)value)) return fail("b200pt_set_option");
    return true;
}

bool HeadlessEngine::read_image(std::vector<float>& rgba) {
    rgba.resize((size_t)width * height * 4);
    if (b200pt_read_image_f32(ctx_, rgba.data())) return fail("b200pt_read_image_f32");
    return true;
}

bool HeadlessEngine::read_accum_mean(std::vector<float>& rgba) {
    rgba.resize((size_t)width * height * 4);
    if (b200pt_read_accum_f32(ctx_, rgba.data())) return fail("b200pt_read_accum_f32");
    for (size_t i = 0; i < rgba.size(); i += 4) {
        const float n = rgba[i + 3];
        if (n > 0.f) { rgba[i] /= n; rgba[i + 1] /= n; rgba[i + 2] /= n; }
        rgba[i + 3] = 1.f;
    }
    return true;
}

bool HeadlessEngine::read_rgba8(std::vector<uint8_t>& rgba8) {
    rgba8.resize((size_t)width * height * 4);
    if (b200pt_read_rgba8(ctx_, rgba8.data())) return fail("b200pt_read_rgba8");
    return true;
}

bool HeadlessEngine::last_render_ms(float& ms) {
    if (b200pt_last_render_ms(ctx_, &ms)) return fail("b200pt_last_render_ms");
    return true;
}

void HeadlessEngine::cleanup() {
    if (ctx_) { b200pt_destroy(ctx_); ctx_ = nullptr; }
}

bool HeadlessEngine::write_ppm(const std::string& path, const uint8_t* rgba8, uint32_t w, uint32_t h) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "P6\n%u %u\n255\n", w, h);
    std::vector<uint8_t> row((size_t)w * 3);
    bool ok = true;
    for (uint32_t y = 0; y < h && ok; ++y) {          // image row 0 is the top of the screen (A-Q2)
        for (uint32_t x = 0; x < w; ++x) std::memcpy(&row[3 * x], rgba8 + 4 * ((size_t)y * w + x), 3);
        ok = std::fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    return std::fclose(f) == 0 && ok;
}

bool HeadlessEngine::write_pfm(const std::string& path, const float* rgba, uint32_t w, uint32_t h) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "PF\n%u %u\n-1.0\n", w, h);        // negative scale: little endian
    std::vector<float> row((size_t)w * 3);
    bool ok = true;
    for (uint32_t yy = 0; yy < h && ok; ++yy) {        // PFM stores the bottom row first
        const uint32_t y = h - 1 - yy;
        for (uint32_t x = 0; x < w; ++x) std::memcpy(&row[3 * x], rgba + 4 * ((size_t)y * w + x), 12);
        ok = std::fwrite(row.data(), 4, row.size(), f) == row.size();
    }
    return std::fclose(f) == 0 && ok;
}

}  // namespace pth
