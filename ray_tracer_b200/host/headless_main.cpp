// b200pt_headless — the reference application's main() (src/main.cpp: init, run, cleanup) without a window:
// builds or loads a scene, runs the compute pass through the C ABI and writes the image to a file.
//   --mode loop  : the engine's own loop, one run_compute() = one b200pt_render per draw() (progressive blend)
//   --mode batch : progressive frames in flight together (b200pt_render_frames), the throughput path
//   --gpus N     : one engine + host thread per GPU, joined by the library's NCCL communicator. batch: the frames are
//                  split (rank r renders f = r mod N) and the accumulators summed onto GPU 0 at the end; --mode interleaved:
//                  every frame is split by interleaved rows and gathered on GPU 0 (1-spp latency mode)
// Prints one JSON line with the timing. No CPU fallback: without a CUDA device the first call fails with its message.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/b200pt.h"
#include "headless_engine.h"

using pth::HeadlessEngine;

static void usage() {
    std::puts(
        "usage: b200pt_headless (--scene FILE | --default-scene ASSETDIR | --cornell ASSETDIR | --obj FILE [--mtl FILE] [--material I])\n"
        "         [--width W] [--height H] [--frames N] [--mode loop|batch|interleaved] [--gpus N] [--in-flight K] [--bounces D] [--rays-per-pixel R]\n"
        "         [--env 0|1] [--angles AX AY AZ] [--pos X Y Z] [--fov F] [--device I] [--gpu-bvh] [--option KEY=VALUE]...\n"
        "         [--out-ppm FILE] [--out-pfm FILE] [--dump-constants]");
}

int main(int argc, char** argv) {
    HeadlessEngine e;
    std::string sceneFile, assetDir, objFile, mtlFile, outPpm, outPfm, mode = "batch";
    bool cornellOnly = false, dump = false, gpuBvh = false;
    int device = 0, material = 0, gpus = 1;
    uint32_t frames = 16, inFlight = 8;
    std::vector<std::pair<std::string, long long>> options;
    e.width = 512; e.height = 512;
    e.rayTracerParams.progressive = 1;
    auto need = [&](int& i, int n) { if (i + n >= argc) { usage(); std::exit(2); } };
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--scene") { need(i, 1); sceneFile = argv[++i]; }
        else if (a == "--default-scene") { need(i, 1); assetDir = argv[++i]; }
        else if (a == "--cornell") { need(i, 1); assetDir = argv[++i]; cornellOnly = true; }
        else if (a == "--obj") { need(i, 1); objFile = argv[++i]; }
        else if (a == "--mtl") { need(i, 1); mtlFile = argv[++i]; }
        else if (a == "--material") { need(i, 1); material = std::atoi(argv[++i]); }
        else if (a == "--width") { need(i, 1); e.width = (uint32_t)std::atoi(argv[++i]); }
        else if (a == "--height") { need(i, 1); e.height = (uint32_t)std::atoi(argv[++i]); }
        else if (a == "--frames") { need(i, 1); frames = (uint32_t)std::atoi(argv[++i]); }
        else if (a == "--mode") { need(i, 1); mode = argv[++i]; }
        else if (a == "--in-flight") { need(i, 1); inFlight = (uint32_t)std::atoi(argv[++i]); }
        else if (a == "--bounces") { need(i, 1); e.rayTracerParams.bounceLimit = (uint32_t)std::atoi(argv[++i]); }
        else if (a == "--rays-per-pixel") { need(i, 1); e.rayTracerParams.raysPerPixel = (uint32_t)std::atoi(argv[++i]); }
        else if (a == "--env") { need(i, 1); e.environment.lightDir[3] = (float)std::atoi(argv[++i]); }
        else if (a == "--angles") { need(i, 3); for (int k = 0; k < 3; ++k) e.cameraAngles[k] = (float)std::atof(argv[++i]); }
        else if (a == "--pos") { need(i, 3); for (int k = 0; k < 3; ++k) e.cameraInfo.pos[k] = (float)std::atof(argv[++i]); }
        else if (a == "--fov") { need(i, 1); e.cameraInfo.fov = (float)std::atof(argv[++i]); }
        else if (a == "--device") { need(i, 1); device = std::atoi(argv[++i]); }
        else if (a == "--gpus") { need(i, 1); gpus = std::atoi(argv[++i]); }
        else if (a == "--gpu-bvh") gpuBvh = true;
        else if (a == "--option") {
            need(i, 1);
            const std::string kv = argv[++i];
            const size_t eq = kv.find('=');
            if (eq == std::string::npos) { usage(); return 2; }
            options.emplace_back(kv.substr(0, eq), std::atoll(kv.c_str() + eq + 1));
        }
        else if (a == "--out-ppm") { need(i, 1); outPpm = argv[++i]; }
        else if (a == "--out-pfm") { need(i, 1); outPfm = argv[++i]; }
        else if (a == "--dump-constants") dump = true;
        else { usage(); return a == "--help" || a == "-h" ? 0 : 2; }
    }
    if (mode != "loop" && mode != "batch" && mode != "interleaved") { usage(); return 2; }
    if (e.width == 0 || e.height == 0 || inFlight == 0 || gpus < 1) { usage(); return 2; }
    if (gpus > 1 && mode == "loop") { std::fprintf(stderr, "--gpus needs --mode batch or interleaved\n"); return 2; }
    if (mode == "interleaved" && e.rayTracerParams.raysPerPixel != 1) { std::fprintf(stderr, "--mode interleaved renders 1-spp frames\n"); return 2; }

    if (dump) {           // the 208-byte block run_compute() would push for frame 0, as hex (no device needed)
        e.scene.spheres.resize(10);
        e.fill_constants();
        const unsigned char* p = (const unsigned char*)&e.constants;
        for (size_t i = 0; i < sizeof e.constants; ++i) std::printf("%02x", p[i]);
        std::printf("\n");
        return 0;
    }

    // ---- init(): scene on the host ----
    const auto t0 = std::chrono::steady_clock::now();
    if (gpuBvh) e.scene.set_bvh_builder(&b200pt_build_bvh, device, 4096);
    try {
        if (!sceneFile.empty()) {
            if (!e.load_scene_file(sceneFile)) { std::fprintf(stderr, "%s\n", e.error().c_str()); return 1; }
        } else if (!assetDir.empty()) {
            if (assetDir.back() != '/') assetDir += '/';
            if (cornellOnly) {
                e.scene.spheres.resize(10);
                std::memset(e.scene.spheres.data(), 0, 10 * sizeof(ptl::Sphere));
                e.scene.push_default_materials();
                e.scene.cornell_box(assetDir);
            } else {
                e.scene.default_scene(assetDir);
            }
        } else if (!objFile.empty()) {
            e.scene.spheres.resize(10);
            std::memset(e.scene.spheres.data(), 0, 10 * sizeof(ptl::Sphere));
            e.scene.push_default_materials();
            if (!mtlFile.empty()) e.scene.read_mtl(mtlFile);
            if (!e.scene.read_obj(objFile, pth::Placement(), material)) { std::fprintf(stderr, "%s: cannot read\n", objFile.c_str()); return 1; }
        } else {
            usage();
            return 2;
        }
    } catch (const std::exception& ex) {
        std::fprintf(stderr, "scene: %s\n", ex.what());
        return 1;
    }
    const double buildS = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    if (gpus > 1 || mode == "interleaved") {
        // ---- one engine per GPU (devices device .. device + gpus - 1), one host thread each; GPU `device` is the root ----
        unsigned char uid[B200PT_UNIQUE_ID_BYTES];
        if (b200pt_comm_unique_id(uid)) { std::fprintf(stderr, "b200pt_comm_unique_id: %s\n", b200pt_global_error()); return 1; }
        std::vector<std::unique_ptr<HeadlessEngine>> eng((size_t)gpus);
        std::vector<std::string> errs((size_t)gpus);
        std::vector<double> runS((size_t)gpus, 0.0), upS((size_t)gpus, 0.0);
        std::vector<float> img;
        auto worker = [&](int r) {
            HeadlessEngine* g = r == 0 ? &e : eng[(size_t)r].get();
            auto bad = [&](const std::string& w) { errs[(size_t)r] = w.empty() ? "failed" : w; };
            const auto tu = std::chrono::steady_clock::now();
            if (!g->init_image(device + r)) return bad(g->error());
            for (const auto& kv : options) if (!g->set_option(kv.first.c_str(), kv.second)) return bad(g->error());
            if (!g->prepare_storage_buffers()) return bad(g->error());
            if (!g->comm_init(gpus, r, uid)) return bad(g->error());
            if (b200pt_sync(g->context())) return bad("b200pt_sync");
            upS[(size_t)r] = std::chrono::duration<double>(std::chrono::steady_clock::now() - tu).count();
            const auto tr = std::chrono::steady_clock::now();
            if (mode == "interleaved") {
                for (uint32_t f = 0; f < frames; ++f) {
                    if (!g->render_frame_interleaved(f, 0)) return bad(g->error());
                    if (b200pt_sync(g->context())) return bad("b200pt_sync");           // one frame at a time: latency mode
                }
            } else {
                const uint32_t per = inFlight * (uint32_t)gpus;
                for (uint32_t f = 0; f < frames; f += per)
                    if (!g->render_batch_split(f, std::min(per, frames - f))) return bad(g->error());
                if (!g->reduce_accum(0)) return bad(g->error());
            }
            if (r == 0) { if (!g->read_accum_mean(img)) return bad(g->error()); }       // D2H of the finished image ends the clock
            else if (b200pt_sync(g->context())) return bad("b200pt_sync");
            runS[(size_t)r] = std::chrono::duration<double>(std::chrono::steady_clock::now() - tr).count();
        };
        for (int r = 1; r < gpus; ++r) {
            eng[(size_t)r].reset(new HeadlessEngine());
            HeadlessEngine& g = *eng[(size_t)r];
            g.width = e.width; g.height = e.height; g.scene = e.scene;
            std::memcpy(g.cameraAngles, e.cameraAngles, sizeof e.cameraAngles);
            g.cameraInfo = e.cameraInfo; g.environment = e.environment; g.rayTracerParams = e.rayTracerParams;
        }
        std::vector<std::thread> th;
        for (int r = 0; r < gpus; ++r) th.emplace_back(worker, r);
        for (auto& t : th) t.join();
        for (int r = 0; r < gpus; ++r) if (!errs[(size_t)r].empty()) { std::fprintf(stderr, "gpu %d: %s\n", device + r, errs[(size_t)r].c_str()); return 1; }
        double mean = 0.0;
        for (size_t i = 0; i < img.size(); i += 4) mean += (double)img[i] + img[i + 1] + img[i + 2];
        mean /= 3.0 * (double)(img.size() / 4);
        if (!outPfm.empty() && !HeadlessEngine::write_pfm(outPfm, img.data(), e.width, e.height)) { std::fprintf(stderr, "cannot write %s\n", outPfm.c_str()); return 1; }
        const double run = *std::max_element(runS.begin(), runS.end());
        const double paths = (double)e.width * e.height * frames;
        std::printf("{\"mode\": \"%s\", \"gpus\": %d, \"width\": %u, \"height\": %u, \"frames\": %u, \"bounce_limit\": %u, \"triangles\": %zu, "
                    "\"scene_build_s\": %.4f, \"scene_upload_s\": %.4f, \"run_s\": %.6f, \"ms_per_frame\": %.4f, \"mpaths_per_s\": %.3f, "
                    "\"mean_radiance\": %.9g}\n",
                    mode.c_str(), gpus, e.width, e.height, frames, e.rayTracerParams.bounceLimit, e.scene.triangles.size(), buildS,
                    *std::max_element(upS.begin(), upS.end()), run, run * 1e3 / frames, paths / run / 1e6, mean);
        for (int r = 1; r < gpus; ++r) eng[(size_t)r]->cleanup();
        e.cleanup();
        return 0;
    }

    // ---- init(): device ----
    if (!e.init_image(device)) { std::fprintf(stderr, "%s\n", e.error().c_str()); return 1; }
    for (const auto& kv : options)
        if (!e.set_option(kv.first.c_str(), kv.second)) { std::fprintf(stderr, "%s\n", e.error().c_str()); return 1; }
    const auto t1 = std::chrono::steady_clock::now();
    if (!e.prepare_storage_buffers()) { std::fprintf(stderr, "%s\n", e.error().c_str()); return 1; }
    const double uploadS = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();

    // ---- run() ----
    const auto t2 = std::chrono::steady_clock::now();
    bool ok = true;
    if (mode == "loop") {
        for (uint32_t f = 0; f < frames && ok; ++f) ok = e.draw();
    } else {
        for (uint32_t f = 0; f < frames && ok; f += inFlight) ok = e.render_batch(f, std::min(inFlight, frames - f));
    }
    if (ok && b200pt_sync(e.context())) { std::fprintf(stderr, "b200pt_sync: %s\n", b200pt_last_error(e.context())); return 1; }
    if (!ok) { std::fprintf(stderr, "%s\n", e.error().c_str()); return 1; }
    const double runS = std::chrono::duration<double>(std::chrono::steady_clock::now() - t2).count();

    // ---- the image ----
    std::vector<float> img;
    ok = mode == "loop" ? e.read_image(img) : e.read_accum_mean(img);
    if (!ok) { std::fprintf(stderr, "%s\n", e.error().c_str()); return 1; }
    double mean = 0.0;
    for (size_t i = 0; i < img.size(); i += 4) mean += (double)img[i] + img[i + 1] + img[i + 2];
    mean /= 3.0 * (double)(img.size() / 4);
    if (!outPfm.empty() && !HeadlessEngine::write_pfm(outPfm, img.data(), e.width, e.height)) { std::fprintf(stderr, "cannot write %s\n", outPfm.c_str()); return 1; }
    if (!outPpm.empty()) {
        std::vector<uint8_t> px;
        if (mode == "loop") {
            if (!e.read_rgba8(px)) { std::fprintf(stderr, "%s\n", e.error().c_str()); return 1; }
        } else {          // unorm8, round to nearest, of the mean image
            px.resize(img.size());
            for (size_t i = 0; i < img.size(); ++i) {
                float v = img[i];
                v = v < 0.f ? 0.f : (v > 1.f ? 1.f : v);
                px[i] = (uint8_t)std::floor(v * 255.f + 0.5f);
            }
        }
        if (!HeadlessEngine::write_ppm(outPpm, px.data(), e.width, e.height)) { std::fprintf(stderr, "cannot write %s\n", outPpm.c_str()); return 1; }
    }
    const double paths = (double)e.width * e.height * frames * e.rayTracerParams.raysPerPixel;
    std::printf("{\"mode\": \"%s\", \"width\": %u, \"height\": %u, \"frames\": %u, \"rays_per_pixel\": %u, \"bounce_limit\": %u, "
                "\"triangles\": %zu, \"objects\": %zu, \"bvh_nodes\": %zu, \"scene_build_s\": %.4f, \"scene_upload_s\": %.4f, "
                "\"run_s\": %.6f, \"ms_per_frame\": %.4f, \"mpaths_per_s\": %.3f, \"mean_radiance\": %.9g}\n",
                mode.c_str(), e.width, e.height, frames, e.rayTracerParams.raysPerPixel, e.rayTracerParams.bounceLimit,
                e.scene.triangles.size(), e.scene.objects.size(), e.scene.nodes.size(), buildS, uploadS, runS,
                runS * 1e3 / frames, paths / runS / 1e6, mean);
    e.cleanup();
    return 0;
}
