#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 path-tracing backend (contract: see DESIGN.md §7).

Metric (BASELINE.json): Mpaths/s on textured-Sponza-class geometry at 1920x1080, depth 8, NEE+MIS;
one path = one trace() call of shaders/raytrace.comp (one pixel of one progressive frame).
`assets/sponza.obj` is not in the reference checkout and there is no network, so the workload is the
labelled procedural stand-in of ray_tracer_b200/scenes.py (264 796 triangles, 25 material groups +
the light box = 26 RenderObjects, environment light on) unless --obj names a real file.

A step = `--frames-per-step` progressive frames per GPU (frameCount = f, raysPerPixel = 1) of the
full 1920x1080 image through b200pt_render_frames. N GPUs split the frames (rank r renders
f = r mod N: independent seeds, raytrace.comp:562-564), each step ends with an NCCL reduce of the
step's float accumulators onto rank 0 (weak scaling: per-GPU work fixed).

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 10 --warmup 3
    python bench.py --impl reference ...     # the CPU restatement (oracle/) on the host cores

The oracle (tests/oracle_binding.py -> oracle/liboracle.so) is imported here ONLY for the
cpu_baseline leg, the --impl reference arm and the algorithmic work counters of the roofline.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mpaths_per_s_sponza_1080p_depth8_nee_mis"
UNIT = "Mpaths/s"
# reference-layout bytes touched per unit of counted work
BYTES = dict(n_box=32, n_tri=144, n_obj=80, n_sph=32, n_hit=64)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames-per-step", type=int, default=24)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--bounce-limit", type=int, default=8)
    ap.add_argument("--workload", default="atrium", choices=["atrium", "cornell", "soup", "statue"])
    ap.add_argument("--soup-triangles", type=int, default=10_000_000)
    ap.add_argument("--obj", default=None, help="path of a real sponza.obj (optional; replaces the stand-in)")
    ap.add_argument("--gpu-bvh", action="store_true", help="build the reference BVH2 of the procedural scenes with b200pt_build_bvh")
    ap.add_argument("--pipeline", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample", default="960x540x12", help="WxHxframes of the CPU baseline sample")
    ap.add_argument("--opt", action="append", default=[], help="backend option key=value (b200pt_set_option), repeatable")
    ap.add_argument("--dump", default=None, help="write the final mean image as .npy (rank 0)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
def make_workload(args):
    """-> (blobs, PushConstants, config dict)."""
    from ray_tracer_b200 import PushConstants, scenes
    W, H = args.width, args.height
    if args.gpu_bvh and args.impl == "b200":
        scenes.GPU_BVH = (int(os.environ.get("LOCAL_RANK", "0")), 4096)
    if args.workload == "atrium":
        if args.obj:
            from ray_tracer_b200 import scene as hs
            b = hs.SceneBuilder()
            b.set_spheres(bytes(320))
            b.add_material(hs.material())
            if not b.read_obj(args.obj, material=0):
                raise SystemExit("cannot read " + args.obj)
            blobs = b.blobs()
            info = dict(triangles=blobs["triangles"].size // 48, objects=blobs["objects"].size // 80, stand_in_for=None)
            name = "C4 Sponza (%s)" % os.path.basename(args.obj)
        else:
            blobs, info = scenes.atrium()
            name = "C4 Sponza STAND-IN: procedural atrium (sponza.obj absent from the reference checkout)"
        pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
        scenes.atrium_camera(pc)
    elif args.workload == "statue":
        blobs, info = scenes.statue_room(800_000)
        pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
        name = "C3 dragon STAND-IN: mirror statue, ~800k triangles, in a Cornell-style room (dragon.obj absent from the reference checkout)"
    elif args.workload == "soup":
        blobs, info = scenes.soup(args.soup_triangles)
        pc = PushConstants(W, H, 0, 1)
        pc["lightDir"][3] = 1.0
        name = "soup: %d random triangles (traversal roofline case)" % args.soup_triangles
    else:
        import numpy as np
        z = np.load(os.path.join(ROOT, "tests", "golden", "scene_cornell.npz"))
        blobs = {k: np.ascontiguousarray(z[k]) for k in ("spheres", "materials", "points", "triangles", "objects", "nodes")}
        info = dict(triangles=blobs["triangles"].size // 48, objects=blobs["objects"].size // 80)
        pc = PushConstants(W, H, 10, info["objects"])
        name = "C1 Cornell box + spheres"
    pc["bounceLimit"] = args.bounce_limit
    scene_mb = sum(v.size for v in blobs.values()) / 1e6
    cfg = dict(workload=name, width=W, height=H, bounce_limit=args.bounce_limit, rays_per_pixel=1,
               frames_per_step_per_gpu=args.frames_per_step, triangles=int(info["triangles"]), objects=int(info["objects"]),
               scene_reference_layout_mb=round(scene_mb, 1), sampling="NEE + MIS (diffuseBRDF, raytrace.comp:430-464)")
    return blobs, pc, cfg


def algorithmic_bytes(work, paths):
    """B per path = 32 N_box + 144 N_tri + 80 N_obj + 32 N_sph + 64 N_hit (+32 accumulator RMW)."""
    return sum(BYTES[k] * work[k] for k in BYTES) / max(paths, 1)


def oracle_sample(blobs, pc, args, sample, threads=0):
    """Runs the CPU restatement on a bounded sample of the SAME workload: every (W/w)-th pixel in x and (H/h)-th in y
    of the real W x H image (the full image's rays and seeds), `n` frames; sample = "wxhxn".
    -> dict(mpaths, seconds, paths, work_all, work_ext, cores)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    w, h, n = (int(x) for x in sample.lower().split("x"))
    W, H = args.width, args.height
    sx, sy = max(1, W // w), max(1, H // h)
    w, h = min(w, (W + sx - 1) // sx), min(h, (H + sy - 1) // sy)
    # torchrun exports OMP_NUM_THREADS=1: ask for the cores this process may run on explicitly
    threads = threads or len(os.sched_getaffinity(0))
    sc = ob.OracleScene(blobs)
    t0 = time.perf_counter()
    _, work_all, work_ext = sc.render_window(pc.buffer, W, H, 0, 0, 0, w, h, n=n, sx=sx, sy=sy, threads=threads, want_work=True)
    dt = time.perf_counter() - t0
    paths = w * h * n
    return dict(mpaths=paths / dt / 1e6, seconds=dt, paths=paths, work_all=work_all, work_ext=work_ext,
                cores=threads, lattice="%dx%d pixels (every %d-th column, %d-th row of the %dx%d image) x %d frames" % (w, h, sx, sy, W, H, n))


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], 0, [], set()
        for t, ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7 or not (t0 - 0.05 <= t <= t1 + 0.15):
                continue
            try:
                sm.append(float(f[0])); smax = max(smax, float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(self.NAMES, f[3:7]):
                if v == "Active":
                    reasons.add(name)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=smax, reasons=sorted(reasons), samples=len(sm),
                    power_w_max=max(power) if power else None)


# ---------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the reference's path on the host cores. The path exists in the reference only as a
    GLSL shader (no CPU code, no graphics stack here), so this times the CPU restatement (oracle/, kind "port")."""
    if rank != 0:
        return
    blobs, pc, cfg = make_workload(args)
    sample = "480x270x4"
    for _ in range(min(args.warmup, 1)):
        oracle_sample(blobs, pc, args, "240x135x1")
    paths, secs, cores = 0, 0.0, 1
    for _ in range(args.steps):
        r = oracle_sample(blobs, pc, args, sample)
        paths += r["paths"]; secs += r["seconds"]; cores = r["cores"]
    v = paths / secs / 1e6
    cfg["frames_per_step_per_gpu"] = None
    out = dict(impl="reference", metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
               ms_per_step=secs / args.steps * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
               data="synthetic", config=cfg,
               cpu_baseline=dict(value=v, unit=UNIT, cores=cores, kind="port",
                                 sample="%s per step, OpenMP over rows, all host cores" % r["lattice"]),
               e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(out), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the backend has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        from ray_tracer_b200 import build
        build.build()
    if world > 1:
        dist.barrier()
    import ray_tracer_b200 as rt

    t_build = time.time()
    blobs, pc, cfg = make_workload(args)
    cfg["scene_build_s"] = round(time.time() - t_build, 3)
    cfg["bvh2_builder"] = "b200pt_build_bvh (GPU)" if args.gpu_bvh else "host"
    W, H, F, N = args.width, args.height, args.frames_per_step, world
    cp = rt.ComputePass(W, H, device=local)
    stream = torch.cuda.Stream(dev)              # a real (non-default) stream: handle 0 would mean "the context's own"
    torch.cuda.set_stream(stream)
    cp.set_stream(stream.cuda_stream)            # kernels run on torch's current stream: torch events time them
    cp.set_option("pipeline", args.pipeline)
    cp.set_option("frames_in_flight", max(1, min(64, F)))
    for kv in args.opt:
        k, v = kv.split("=")
        cp.set_option(k, int(v))
    t_up = time.perf_counter()
    cp.copy_buffers(blobs)
    cp.sync()
    upload_s = time.perf_counter() - t_up
    wide_build = dict(where="device" if cp.get_option("wide_build") else "host", relayout_ms=cp.get_option("last_scene_build_us") / 1e3,
                      device_ms=cp.get_option("last_wide_device_us") / 1e3)

    # rank-local step buffer (attached as the context's accumulator) and, on rank 0, the running total
    stepbuf = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    cp.attach_accum(stepbuf.data_ptr())
    total = torch.zeros_like(stepbuf) if (N > 1 and rank == 0) else None
    pinned = torch.empty(H, W, 4, dtype=torch.float32, pin_memory=True)
    pinned_np = pinned.numpy()
    pcbuf = pc.buffer
    frame_cursor = [0]

    from ray_tracer_b200 import multi_gpu

    def step(e2e=False):
        f0 = frame_cursor[0]
        frame_cursor[0] += F * N
        if N > 1:
            # rank r renders frames f0 + r, f0 + r + N, ... (F of them); NCCL sum over NVLink onto rank 0's running total
            multi_gpu.render_frames_split(cp, pcbuf, f0, F * N, stepbuf, total)
        else:
            cp.render_frames(pcbuf, f0, F)                         # host push constants -> one batch of F frames
        if e2e and rank == 0:
            if N > 1:
                pinned.copy_(total, non_blocking=True)
                stream.synchronize()
            else:
                cp.read_accum(out=pinned_np)                       # D2H of the progressive float image + sync

    def fence():
        torch.cuda.synchronize()
        if N > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(k, e2e):
        fence()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(k):
            step(e2e)
        e1.record(stream)
        fence()
        t1 = time.perf_counter()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if N > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), t0, t1

    for _ in range(max(args.warmup, 0)):
        step()
    fence()
    cp.reset_counters()
    cp.set_option("profile_stages", 1)
    cp.stage_times(reset=True)
    sampler = ClockSampler(local) if rank == 0 else None
    ms, t0, t1 = timed(args.steps, e2e=False)
    clocks = sampler.stop(t0, t1) if sampler else None
    counters = cp.counters()
    stages = cp.stage_times(reset=True)
    cp.set_option("profile_stages", 0)

    for _ in range(2):
        step(e2e=True)
    ms_e2e, _, _ = timed(args.steps, e2e=True)

    # interactive mode (BASELINE.json: "ms/frame @1spp", frames split by image tiles across GPUs): single progressive
    # frames, each split into row bands over the ranks and gathered on rank 0 (multi_gpu.render_frame_bands)
    band_img = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    cp.attach_accum(band_img.data_ptr())
    cp.set_option("frames_in_flight", 1)
    n_inter = max(4, min(32, 4 * args.steps))

    def interactive(k):
        for f in range(k):
            if N > 1:
                multi_gpu.render_frame_bands(cp, pcbuf, 100000 + f, band_img)
            else:
                cp.render_frames(pcbuf, 100000 + f, 1)

    interactive(3)
    fence()
    i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    i0.record(stream)
    interactive(n_inter)
    i1.record(stream)
    fence()
    ms_inter = torch.tensor([i0.elapsed_time(i1)], dtype=torch.float64, device=dev)
    if N > 1:
        dist.all_reduce(ms_inter, op=dist.ReduceOp.MAX)
    ms_inter = float(ms_inter.item()) / n_inter
    cp.attach_accum(stepbuf.data_ptr())

    # the engine's own per-frame call: b200pt_render = one literal dispatch (progressive blend into the storage image).
    # 1-sample dispatches are served by the wavefront kernels; pipeline 0 is the per-pixel reference-order kernel.
    def dispatches(k, first):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fence()
        e0.record(stream)
        for f in range(k):
            pc["frameCount"] = first + f
            cp.run_compute(pc.buffer)
        e1.record(stream)
        fence()
        return e0.elapsed_time(e1) / k
    keep_progressive = int(pc["progressive"])
    pc["progressive"] = 1
    dispatches(2, 0)
    ms_dispatch = dispatches(n_inter, 2)
    cp.set_option("pipeline", 0)
    ms_dispatch_ref = dispatches(2, 2 + n_inter)
    cp.set_option("pipeline", args.pipeline)
    pc["progressive"] = keep_progressive
    pc["frameCount"] = 0

    paths_per_step = W * H * F * N
    value = paths_per_step * args.steps / (ms * 1e-3) / 1e6
    e2e_value = paths_per_step * args.steps / (ms_e2e * 1e-3) / 1e6

    if rank == 0:
        frames_done = frame_cursor[0]
        img = (total if N > 1 else stepbuf)
        mean_rgb = (img[..., :3].double().mean() / max(img[..., 3].double().mean().item(), 1.0)).item()
        if args.dump:
            a = img.cpu().numpy()
            np.save(args.dump, a[..., :3] / np.maximum(a[..., 3:4], 1))

        # ---- algorithmic work (oracle counters) + CPU baseline ----
        cpu = None
        if N == 1 and not args.no_cpu_baseline:
            r = oracle_sample(blobs, pc, args, args.cpu_sample)
            cpu = dict(value=r["mpaths"], unit=UNIT, cores=r["cores"], kind="port",
                       sample="%s; %.1f s, OpenMP over rows, all host cores" % (r["lattice"], r["seconds"]))
        else:
            r = oracle_sample(blobs, pc, args, "240x135x1")
        b_all = algorithmic_bytes(r["work_all"], r["paths"]) + 32.0
        b_ext = algorithmic_bytes(r["work_ext"], r["paths"])
        b_con = b_all - b_ext - 32.0
        stage_bytes = dict(extend=b_ext, occlude=b_con)      # NEE/pdf rays: their traversal is k_occlude
        dom = max(("extend", "occlude"), key=lambda s: stages[s][0])
        dom_ms, dom_launches = stages[dom]
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        paths_rank = W * H * F * args.steps                      # stage times are this rank's
        achieved = stage_bytes[dom] * paths_rank / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if (args.width, args.height, F, args.obj) == (1920, 1080, tj.get("frames_in_flight"), None):      # the configuration the capture was taken on
                traffic = tj.get(args.workload, {}).get(dom)
        except (OSError, ValueError):
            pass
        # second candidate bound of SURVEY.md §8(d3): FP32 issue. Flops per path of the extension rays in the
        # reference's formulation: 24/box test, 64/triangle test, 20/sphere test, 150/object visit.
        wx = r["work_ext"]
        flops_ext = (24.0 * wx["n_box"] + 64.0 * wx["n_tri"] + 20.0 * wx["n_sph"] + 150.0 * wx["n_obj"]) / max(r["paths"], 1)
        ext_ms = stages["extend"][0]
        fp32 = dict(achieved=flops_ext * paths_rank / (ext_ms * 1e-3) / 1e12 if ext_ms > 0 else None, peak=74.4, unit="TFLOP/s",
                    kernel="k_extend", peak_source="nominal 148 SM x 128 lanes x 2 x 1.965 GHz (not measured)",
                    algorithmic_flops_per_path=flops_ext)
        fp32["frac"] = fp32["achieved"] / fp32["peak"] if fp32["achieved"] else None
        roofline = dict(bound="hbm", kernel="k_" + dom, fp32_issue=fp32, achieved=achieved, peak=peak, unit="GB/s",
                        frac=(achieved / peak) if achieved else None, traffic=traffic,
                        peak_source="MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                        algorithmic_bytes_per_path=dict(all_rays=b_all, extension_rays=b_ext, nee_and_pdf_rays=b_con),
                        launches=dom_launches, avg_launch_ms=dom_ms / max(dom_launches, 1),
                        bytes_per_launch=stage_bytes[dom] * paths_rank / max(dom_launches, 1),
                        stage_ms_per_step={k: v[0] / args.steps for k, v in stages.items()},
                        note="algorithmic bytes are the REFERENCE layout's (32 B/box test, 144 B/triangle test, 80 B/object visit); "
                             "the scene is L2-resident, so achieved may exceed the HBM figure")
        cfg.update(parallelism="spp-split x%d (frames f = r mod N) + NCCL reduce per step" % N if N > 1 else "single GPU",
                   l2="path state per step %.0f MB >> 126 MB L2; scene (%.0f MB) is re-read by every frame by design; no flush"
                      % (W * H * F * 160 / 1e6, cfg["scene_reference_layout_mb"]),
                   frames_rendered=frames_done, mean_radiance=mean_rgb, scene_upload_s=round(upload_s, 3), wide_bvh_build=wide_build,
                   rays_per_path=(counters["extension_rays"] + counters["shadow_rays"]) / max(counters["paths"], 1))
        out = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=N, steps=args.steps, warmup=args.warmup,
                   ms_per_step=ms / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                   data="synthetic", config=cfg, clocks=clocks,
                   e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=208 * N, d2h_bytes_per_step=W * H * 16,
                            ms_per_step=ms_e2e / args.steps),
                   gpu_launches=counters["launches"], roofline=roofline, cpu_baseline=cpu,
                   ms_per_frame_1spp=ms / args.steps / F,
                   interactive=dict(ms_per_frame=ms_inter, frames=n_inter, run_compute_ms=ms_dispatch, run_compute_pipeline0_ms=ms_dispatch_ref, split="row bands over %d GPU(s), gathered on rank 0" % N,
                                    note="one 1-spp progressive frame at a time (latency), vs ms_per_frame_1spp = throughput mode / frames"))
        print(json.dumps(out), flush=True)
    cp.attach_accum(0)
    cp.close()
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
