#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 path-tracing backend (contract: see DESIGN.md §7).

Metric (BASELINE.json): Mpaths/s on textured-Sponza-class geometry at 1920x1080, depth 8, NEE+MIS;
one path = one trace() call of shaders/raytrace.comp (one pixel of one progressive frame).
`assets/sponza.obj` is not in the reference checkout and there is no network, so the workload is the
labelled procedural stand-in of ray_tracer_b200/scenes.py (264 796 triangles, 25 material groups +
the light box = 26 RenderObjects, environment light on) unless --obj names a real file.

A step = `--frames-per-step` progressive frames per GPU (frameCount = f, raysPerPixel = 1) of the
full 1920x1080 image through b200pt_render_frames. N GPUs split the frames (rank r renders
f = r mod N: independent seeds, raytrace.comp:562-564) and each step ends with the library's NCCL
reduce of the float accumulators onto rank 0 (b200pt_reduce_accum; weak scaling: per-GPU work fixed).

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 10 --warmup 3
    python bench.py --impl reference ...     # the reference's own shader text compiled for the host cores

Besides the contract's line the run measures (all printed in the same JSON object):
  strong_scaling   time to the finished, reduced, host-resident image of C4 at --strong-spp (1024) frames TOTAL
  c5               C5 of BASELINE.json (atrium + statue stand-in, 3840x2160, depth 16): Mpaths/s, reduce time
  interactive      one 1-spp frame at a time: split by interleaved rows over the GPUs (b200pt_render_frame_interleaved)
  parity           relRMSE of this run's own image against the CPU reference on the cpu_baseline lattice

The checker libraries (oracle/liboracle.so, oracle/_ref/libraytrace_comp.so) are loaded here ONLY for the
cpu_baseline leg, the --impl reference arm, the parity block and the "work avoided" counters of the roofline.
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
NODE_BYTES, PRIM_BYTES = 80, 64          # wide layout: pt_bvh8.h Bvh8Node / PrimRecord


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
    ap.add_argument("--workload", default="atrium", choices=["atrium", "cornell", "soup", "statue", "c5"])
    ap.add_argument("--soup-triangles", type=int, default=10_000_000)
    ap.add_argument("--soup-rays", type=int, default=1 << 24)
    ap.add_argument("--obj", default=None, help="path of a real sponza.obj (optional; replaces the stand-in)")
    ap.add_argument("--gpu-bvh", action="store_true", help="build the reference BVH2 of the procedural scenes with b200pt_build_bvh")
    ap.add_argument("--pipeline", type=int, default=1)
    ap.add_argument("--textures", type=int, default=0, help="1: texture extension on, procedural albedo maps bound to the atrium's materials")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample", default="480x270x12", help="WxHxframes of the CPU baseline sample")
    ap.add_argument("--parity-sample", default="64x36x1024", help="WxHxframes of the parity lattice: C4's own 1024 spp on every 30th pixel")
    ap.add_argument("--strong-spp", type=int, default=1024, help="frames of the strong-scaling leg (0 = skip)")
    ap.add_argument("--c5-frames", type=int, default=24, help="frames per GPU of the C5 leg (0 = skip)")
    ap.add_argument("--only-value", action="store_true", help="skip every leg but the timed value (short runs under ncu)")
    ap.add_argument("--opt", action="append", default=[], help="backend option key=value (b200pt_set_option), repeatable")
    ap.add_argument("--dump", default=None, help="write the final mean image as .npy (rank 0)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
def make_workload(args, make_builder=None, workload=None, size=None, bounce=None):
    """-> (blobs, PushConstants, config dict). make_builder: scene builder factory (the reference arm passes the
    oracle's so that it loads no product library)."""
    from ray_tracer_b200 import scenes
    from ray_tracer_b200.params import PushConstants
    workload = workload or args.workload
    W, H = size or (args.width, args.height)
    bounce = args.bounce_limit if bounce is None else bounce
    if args.gpu_bvh and args.impl == "b200" and make_builder is None:
        scenes.GPU_BVH = (int(os.environ.get("LOCAL_RANK", "0")), 4096)
    kw = dict(make_builder=make_builder) if make_builder else {}
    if workload == "atrium":
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
            blobs, info = scenes.atrium(textured=bool(args.textures), **kw)
            name = "C4 Sponza STAND-IN: procedural atrium (sponza.obj absent from the reference checkout)"
        pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
        scenes.atrium_camera(pc)
    elif workload == "c5":
        blobs, info = scenes.atrium_with_statue(**kw)
        pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
        scenes.atrium_camera(pc)
        name = "C5 Sponza + dragon STAND-IN: procedural atrium + mirror statue (both assets absent from the reference checkout)"
    elif workload == "statue":
        blobs, info = scenes.statue_room(800_000, **kw)
        pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
        name = "C3 dragon STAND-IN: mirror statue, ~800k triangles, in a Cornell-style room (dragon.obj absent from the reference checkout)"
    elif workload == "soup":
        blobs, info = scenes.soup(args.soup_triangles, **kw)
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
    pc["bounceLimit"] = bounce
    scene_mb = sum(v.size for v in blobs.values()) / 1e6
    cfg = dict(workload=name, width=W, height=H, bounce_limit=bounce, rays_per_pixel=1,
               frames_per_step_per_gpu=args.frames_per_step, triangles=int(info["triangles"]), objects=int(info["objects"]),
               scene_reference_layout_mb=round(scene_mb, 1), sampling="NEE + MIS (diffuseBRDF, raytrace.comp:430-464)")
    if info.get("textures"):
        cfg["textures"] = info["textures"]
    return blobs, pc, cfg, info


def algorithmic_bytes(work, paths):
    """B per path = 32 N_box + 144 N_tri + 80 N_obj + 32 N_sph + 64 N_hit (+32 accumulator RMW)."""
    return sum(BYTES[k] * work[k] for k in BYTES) / max(paths, 1)


def _lattice(W, H, sample):
    w, h, n = (int(x) for x in sample.lower().split("x"))
    sx, sy = max(1, W // w), max(1, H // h)
    w, h = min(w, (W + sx - 1) // sx), min(h, (H + sy - 1) // sy)
    return w, h, n, sx, sy


def cpu_reference_sample(blobs, pc, W, H, sample, want_work=False, threads=0, force_port=False, textures=None):
    """The reference's path on the host cores, on a bounded sample of the SAME workload: every (W/w)-th column and
    (H/h)-th row of the real W x H image (the full image's rays and seeds), n frames; sample = "wxhxn".
    kind "reference": oracle/_ref/libraytrace_comp.so = the reference's own shaders/raytrace.comp compiled for the CPU
    (oracle/ref_from_glsl.py); kind "port": the hand-written restatement (oracle/liboracle.so), used when that library
    is not in the tree, and always for the work counters (the shader text has none for NEE / pdf rays)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    import ref_shader_binding as rb
    w, h, n, sx, sy = _lattice(W, H, sample)
    threads = threads or len(os.sched_getaffinity(0))      # torchrun exports OMP_NUM_THREADS=1: ask explicitly
    # textures: the HEAD shader never samples, so the texture extension is checked against the port's restatement
    use_ref = os.path.exists(rb.LIB_PATH) and not force_port and not want_work and not textures
    work_all = work_ext = None
    if use_ref:
        sc = rb.RefShaderScene(blobs)
        t0 = time.perf_counter()
        img = sc.render_window(pc.buffer, W, H, 0, 0, 0, w, h, n=n, sx=sx, sy=sy, threads=threads)
    else:
        sc = ob.OracleScene(blobs)
        for slot, tex in (textures or {}).items():
            sc.set_texture(slot, tex)
        if textures:
            sc.use_textures(True)
        t0 = time.perf_counter()
        img, work_all, work_ext = sc.render_window(pc.buffer, W, H, 0, 0, 0, w, h, n=n, sx=sx, sy=sy, threads=threads, want_work=True)
    dt = time.perf_counter() - t0
    paths = w * h * n
    return dict(mpaths=paths / dt / 1e6, seconds=dt, paths=paths, work_all=work_all, work_ext=work_ext, cores=threads, image=img,
                kind="reference" if use_ref else "port", lattice_shape=(w, h, n, sx, sy),
                lattice="%dx%d pixels (every %d-th column, %d-th row of the %dx%d image) x %d frames" % (w, h, sx, sy, W, H, n))


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
    """--impl reference: the reference's own implementation of the path on the host cores. The path exists in the
    reference only as a GLSL shader; oracle/_ref/libraytrace_comp.so is that shader's text compiled for the CPU
    (kind "reference"); without it the hand-written port is timed (kind "port"). The scene is built by the oracle's
    builder: this arm loads no library of the product."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    blobs, pc, cfg, _ = make_workload(args, make_builder=ob.OracleBuilder)
    W, H = args.width, args.height
    sample = "320x180x4"
    for _ in range(min(args.warmup, 1)):
        cpu_reference_sample(blobs, pc, W, H, "240x135x1")
    paths, secs, cores, r = 0, 0.0, 1, None
    for _ in range(args.steps):
        r = cpu_reference_sample(blobs, pc, W, H, sample)
        paths += r["paths"]; secs += r["seconds"]; cores = r["cores"]
    v = paths / secs / 1e6
    cfg["frames_per_step_per_gpu"] = None
    out = dict(impl="reference", metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
               ms_per_step=secs / args.steps * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
               data="synthetic", config=cfg,
               cpu_baseline=dict(value=v, unit=UNIT, cores=cores, kind=r["kind"],
                                 sample="%s per step, OpenMP over rows, all host cores" % r["lattice"],
                                 what="shaders/raytrace.comp compiled for the CPU by oracle/ref_from_glsl.py (oracle/_ref/libraytrace_comp.so)"
                                      if r["kind"] == "reference" else "oracle/liboracle.so (hand-written restatement)"),
               e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    emit(out)


def load_json(*path):
    try:
        return json.load(open(os.path.join(ROOT, *path)))
    except (OSError, ValueError):
        return {}


def nearest_gather_peak(pm, prefix, mb):
    """measured random-gather bandwidth (profiles/peaks_micro.json) for records `prefix` over an array of `mb` MB:
    the entry with the smallest footprint that is at least as large (the ceiling falls with the footprint)."""
    best = None
    for k, v in pm.get("l2_gather", {}).items():
        if not k.startswith(prefix):
            continue
        size = k.split("_")[-1]
        smb = float(size[:-2]) * (1024.0 if size.endswith("GB") else 1.0)
        if smb >= mb and (best is None or smb < best[0]):
            best = (smb, k, v["gbs"])
    return best


def build_roofline(args, cfg, F, steps, stages, counted, peaks, pm, tj, avoided, scene_sizes):
    """Three candidate bounds for the dominant kernel, each a TIME for the launches of the timed region:
      hbm    DRAM bytes the kernel really moves (ncu dram__bytes of the same command, profiles/traffic.json) / the
             driver's measured HBM copy bandwidth
      l2     the kernel's own algorithmic bytes in the WIDE layout — node visits x 80 B + triangle tests x 64 B, counted
             by the kernel (option count_work) — / the measured random-gather bandwidth of L2-resident records
      issue  warp instructions (ncu smsp__inst_executed of the same command) / the measured issue rate
    bound = the largest; frac = that time / the measured time (CUDA events) <= ~1."""
    W, H = args.width, args.height
    dom = max(("extend", "occlude"), key=lambda s: stages[s][0])
    dom_ms, dom_launches = stages[dom]
    paths_rank = W * H * F * steps
    kname = "k_" + dom
    cand = {}
    # --- l2: counted work
    nv, tt = counted.get(dom + "_node_visits_per_path"), counted.get(dom + "_tri_tests_per_path")
    if nv is not None:
        node_mb, prim_mb = scene_sizes
        pn = nearest_gather_peak(pm, "node80", node_mb)
        pp = nearest_gather_peak(pm, "prim64", prim_mb)
        bytes_l2 = (nv * NODE_BYTES + tt * PRIM_BYTES) * paths_rank
        c = dict(bytes=bytes_l2, node_visits_per_path=nv, tri_tests_per_path=tt)
        if pn and pp:
            t_ms = (nv * NODE_BYTES * paths_rank / pn[2] + tt * PRIM_BYTES * paths_rank / pp[2]) / 1e6
            c.update(time_ms=t_ms, peak_gbs=dict(nodes=pn[2], prims=pp[2]), peak_source="profiles/peaks_micro.json l2_gather %s / %s (measured)" % (pn[1], pp[1]))
        cand["l2"] = c
    # --- hbm / issue: ncu figures of the same command, per path
    tk = tj.get(args.workload, {})
    same_cfg = tk.get("config") == dict(width=W, height=H, frames_in_flight=F, bounce_limit=args.bounce_limit) and not args.obj
    kk = tk.get("kernels", {}).get(kname) if same_cfg else None
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None
    if kk:
        per_path_dram = kk["dram_bytes"] / tk["paths"]
        per_path_inst = kk["warp_inst"] / tk["paths"]
        traffic = per_path_dram * paths_rank / max(dom_launches, 1)
        cand["hbm"] = dict(bytes=per_path_dram * paths_rank, time_ms=per_path_dram * paths_rank / hbm / 1e6, peak_gbs=hbm,
                           peak_source="MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)")
        issue_peak = pm.get("issue_ffma_lop3_ginst")
        c = dict(warp_instructions=per_path_inst * paths_rank)
        if issue_peak:
            c.update(time_ms=per_path_inst * paths_rank / issue_peak / 1e6, peak_ginst=issue_peak,
                     peak_source="profiles/peaks_micro.json issue_ffma_lop3_ginst (measured)")
        cand["issue"] = c
    timed = {k: v for k, v in cand.items() if "time_ms" in v}
    bound = max(timed, key=lambda k: timed[k]["time_ms"]) if timed else None
    out = dict(kernel=kname, bound=bound, launches=dom_launches, measured_ms=dom_ms, avg_launch_ms=dom_ms / max(dom_launches, 1),
               candidates=cand, traffic=traffic,
               stage_ms_per_step={k: v[0] / steps for k, v in stages.items()},
               work_avoided=avoided)
    if bound:
        b = timed[bound]
        if bound == "issue":
            out.update(achieved=b["warp_instructions"] / dom_ms / 1e6, peak=b["peak_ginst"], unit="G warp-instructions/s")
        elif bound == "hbm":
            out.update(achieved=b["bytes"] / dom_ms / 1e6, peak=b["peak_gbs"], unit="GB/s")
        else:
            out.update(achieved=b["bytes"] / dom_ms / 1e6, peak=b["bytes"] / b["time_ms"] / 1e6, unit="GB/s")
        out["frac"] = b["time_ms"] / dom_ms
        out["peak_source"] = b["peak_source"]
    else:
        out.update(achieved=None, peak=None, unit=None, frac=None,
                   note="no measured ceiling available for this configuration (profiles/peaks_micro.json / traffic.json)")
    out["fractions"] = {k: v["time_ms"] / dom_ms for k, v in timed.items()}
    return out


_REAL_STDOUT = None


def quiet_stdout():
    """From here on whatever native libraries write to file descriptor 1 (NCCL's version banner, for one) goes to stderr; the
    contract's one JSON line is written by emit() to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = json.dumps(obj) + "\n"
    if _REAL_STDOUT is None:
        sys.stdout.write(line)
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, line.encode())


def main():
    args = parse_args()
    quiet_stdout()
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
    # NCCL writes its debug output — the "NCCL version ..." banner of its first communicator included — to stdout, in front of
    # the JSON line: send it to stderr instead (warnings stay visible, stdout holds the one line the contract asks for)
    # (NCCL honours NCCL_DEBUG_FILE only above the VERSION level, and prints the banner at VERSION and WARN)
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        from ray_tracer_b200 import build
        build.build()
    if world > 1:
        dist.barrier()
    import ray_tracer_b200 as rt
    from ray_tracer_b200 import backend

    N = world
    stream = torch.cuda.Stream(dev)              # a real (non-default) stream: handle 0 would mean "the context's own"
    torch.cuda.set_stream(stream)

    def fence():
        torch.cuda.synchronize()
        if N > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if N > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def new_pass(W, H, blobs, frames_in_flight, textures=None):
        cp = rt.ComputePass(W, H, device=local)
        cp.set_stream(stream.cuda_stream)        # kernels run on torch's current stream: torch events time them
        cp.set_option("pipeline", args.pipeline)
        cp.set_option("frames_in_flight", max(1, min(64, frames_in_flight)))
        for kv in args.opt:
            k, v = kv.split("=")
            cp.set_option(k, int(v))
        cp.copy_buffers(blobs)
        if textures:
            for slot, tex in textures.items():
                cp.upload_texture(slot, tex)
            cp.set_option("textures", 1)
        if N > 1:                                # the library's own communicator (csrc/comm.cu); torch only carries the id
            uid = torch.from_numpy(backend.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).to(dev)
            dist.broadcast(uid, src=0)
            cp.comm_init(N, rank, uid.cpu().numpy())
        cp.sync()
        return cp

    def device_ms(fn):
        """fn() between two events on the stream, fenced on both sides; max over ranks."""
        fence()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        fn()
        e1.record(stream)
        fence()
        t1 = time.perf_counter()
        return max_over_ranks(e0.elapsed_time(e1)), t0, t1

    # ------------------------------------------------------------------------------------------
    if args.workload == "soup":
        return bench_soup(args, rank, N, dev, stream, new_pass, device_ms)

    t_build = time.time()
    blobs, pc, cfg, info = make_workload(args)
    cfg["scene_build_s"] = round(time.time() - t_build, 3)
    cfg["bvh2_builder"] = "b200pt_build_bvh (GPU)" if args.gpu_bvh else "host"
    W, H, F = args.width, args.height, args.frames_per_step
    t_up = time.perf_counter()
    cp = new_pass(W, H, blobs, F, textures=info.get("texture_images"))
    upload_s = time.perf_counter() - t_up
    wide_build = dict(where="device" if cp.get_option("wide_build") else "host", relayout_ms=cp.get_option("last_scene_build_us") / 1e3,
                      device_ms=cp.get_option("last_wide_device_us") / 1e3)

    pinned = torch.empty(H, W, 4, dtype=torch.float32, pin_memory=True)
    pinned_np = pinned.numpy()
    pcbuf = pc.buffer
    frame_cursor = [0]

    def step(e2e=False):
        """One step: F frames on every rank (frames f0 + r, f0 + r + N, ...), the step's sum on rank 0."""
        f0 = frame_cursor[0]
        frame_cursor[0] += F * N
        cp.render_frames_split(pcbuf, f0, F * N)      # host push constants -> this rank's F frames in one batch
        cp.reduce_accum(0, zero_others=True)          # ncclReduce over NVLink onto rank 0's running total (a no-op on one GPU)
        if e2e and rank == 0:
            cp.read_accum(out=pinned_np)              # D2H of the progressive float image + sync

    for _ in range(max(args.warmup, 0)):
        step()
    fence()
    cp.reset_counters()
    sampler = ClockSampler(local) if rank == 0 else None
    ms, t0, t1 = device_ms(lambda: [step() for _ in range(args.steps)])       # the library's default path (no per-stage events)
    clocks = sampler.stop(t0, t1) if sampler else None
    counters = cp.counters()
    # per-stage device times: a second, shorter pass with an event between consecutive launches (serialises the streams)
    cp.set_option("profile_stages", 1)
    cp.stage_times(reset=True)
    stage_steps = max(1, min(3, args.steps))
    ms_staged, _, _ = device_ms(lambda: [step() for _ in range(stage_steps)])
    stages = {k: (v[0] * args.steps / stage_steps, v[1] * args.steps // stage_steps) for k, v in cp.stage_times(reset=True).items()}
    cp.set_option("profile_stages", 0)
    paths_per_step = W * H * F * N
    value = paths_per_step * args.steps / (ms * 1e-3) / 1e6
    out = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=N, steps=args.steps, warmup=args.warmup, ms_per_step=ms / args.steps,
               higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic", config=cfg, clocks=clocks,
               gpu_launches=counters["launches"], ms_per_frame_1spp=ms / args.steps / F, ms_per_step_with_stage_events=ms_staged / stage_steps)
    out["exact_retrace"] = dict(extension_rays_per_path=counters["exact_extension_rays"] / max(counters["paths"], 1),
                                light_queries_per_path=counters["exact_light_queries"] / max(counters["paths"], 1),
                                note="contested rays (two candidates closer together than the rounding slack of their distances: coincident "
                                     "surfaces) re-traced in the reference's own order so that results equal the reference's bit for bit")
    if args.only_value:
        out["stage_ms_per_step"] = {k: v[0] / args.steps for k, v in stages.items()}
        if rank == 0:
            emit(out)
        cp.close()
        if N > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    for _ in range(2):
        step(e2e=True)
    ms_e2e, _, _ = device_ms(lambda: [step(e2e=True) for _ in range(args.steps)])
    out["e2e"] = dict(value=paths_per_step * args.steps / (ms_e2e * 1e-3) / 1e6, unit=UNIT, h2d_bytes_per_step=208 * N,
                      d2h_bytes_per_step=W * H * 16, ms_per_step=ms_e2e / args.steps)
    mean_rgb = float(pinned_np[..., :3].astype(np.float64).mean() / max(float(pinned_np[..., 3].mean()), 1.0)) if rank == 0 else None

    # ---- counted pass: the traversal kernels count their node visits / triangle tests (one step, not timed)
    cp.set_option("count_work", 1)
    cp.reset_counters()
    step()
    cp.sync()
    cw = cp.counters()
    cp.set_option("count_work", 0)
    pr = max(cw["paths"], 1)
    counted = dict(extend_node_visits_per_path=cw["extend_node_visits"] / pr, extend_tri_tests_per_path=cw["extend_tri_tests"] / pr,
                   occlude_node_visits_per_path=cw["occlude_node_visits"] / pr, occlude_tri_tests_per_path=cw["occlude_tri_tests"] / pr,
                   extension_rays_traced_per_path=cw["extension_rays_traced"] / pr, occlusion_rays_traced_per_path=cw["occlusion_rays_traced"] / pr,
                   pdf_rays_traced_per_path=cw["pdf_rays_traced"] / pr,
                   extension_rays_algorithmic_per_path=cw["extension_rays"] / pr, connect_items_per_path=cw["connect_items"] / pr)

    # ---- strong scaling: a FIXED number of frames split over the ranks, timed to the reduced image in host memory
    strong = None
    if args.strong_spp > 0:
        spp = args.strong_spp

        def strong_run():
            cp.clear()
            per = F * N
            for f in range(0, spp, per):
                cp.render_frames_split(pcbuf, f, min(per, spp - f))
            cp.reduce_accum(0, zero_others=True)
            if rank == 0:
                cp.read_accum(out=pinned_np)
        fence()
        tw0 = time.perf_counter()
        ms_strong, _, _ = device_ms(strong_run)
        wall = max_over_ranks(time.perf_counter() - tw0)
        strong = dict(spp=spp, width=W, height=H, seconds=ms_strong / 1e3, wall_seconds=wall, mpaths_per_s=W * H * spp / ms_strong / 1e3,
                      note="C4 at %d spp TOTAL (frames f = r mod N per rank), one ncclReduce of the %.1f MB accumulators, D2H of the image on rank 0; "
                           "time-to-image ratio across N = strong scaling" % (spp, W * H * 16 / 1e6))
        if rank == 0:
            strong["mean_radiance"] = float(pinned_np[..., :3].astype(np.float64).mean() / spp)

    # ---- interactive mode (BASELINE.json: "ms/frame @1spp"): single frames, interleaved rows over the GPUs
    cp.set_option("frames_in_flight", 1)
    n_inter = max(4, min(32, 4 * args.steps))

    def interactive(k):
        for f in range(k):
            cp.render_frame_interleaved(pcbuf, 100000 + f, 0)
    interactive(3)
    ms_inter, _, _ = device_ms(lambda: interactive(n_inter))
    ms_inter /= n_inter
    cp.set_option("exact_order", 0)         # without the reference-order re-trace of contested rays (not bit-exact: see exact_retrace)
    interactive(2)
    ms_inter_fast = device_ms(lambda: interactive(n_inter))[0] / n_inter
    cp.set_option("exact_order", 1)

    # the engine's own per-frame call: b200pt_render = one literal dispatch (progressive blend into the storage image)
    def dispatches(k, first):
        for f in range(k):
            pc["frameCount"] = first + f
            cp.run_compute(pc.buffer)
    keep_progressive = int(pc["progressive"])
    pc["progressive"] = 1
    dispatches(2, 0)
    ms_dispatch = device_ms(lambda: dispatches(n_inter, 2))[0] / n_inter
    cp.set_option("pipeline", 0)
    ms_dispatch_ref = device_ms(lambda: dispatches(2, 2 + n_inter))[0] / 2
    cp.set_option("pipeline", args.pipeline)
    pc["progressive"] = keep_progressive
    pc["frameCount"] = 0
    cp.set_option("frames_in_flight", max(1, min(64, F)))

    # ---- UI "update objects" (src/vk_engine.cpp:1594-1603): one object's transform edited; wall time until the scene is usable
    upd = None
    if blobs["objects"].size >= 80:
        moved = blobs["objects"].copy()
        moved[:64].view(np.float32)[12] += 0.25                    # object 0: translate x
        times = []
        for k in range(6):
            fence()
            tu = time.perf_counter()
            cp.update_buffer("objects", moved if k % 2 == 0 else blobs["objects"])
            cp.sync()
            times.append((time.perf_counter() - tu) * 1e3)
        upd = dict(ms=min(times), ms_all=[round(t, 3) for t in times], refit=bool(cp.get_option("last_update_was_refit")),
                   device_ms=cp.get_option("last_wide_device_us") / 1e3,
                   note="b200pt_update_objects of all %d objects with one matrix changed (wall clock incl. upload and sync); the scene is back to the original afterwards" % info["objects"])

    # ---- CPU baseline (timing) and parity of this run's own output at the configuration's OWN sample count: the lattice
    # pixels of the full image, frames 0 .. n-1, rendered here (whole rows, every sy-th) and by the CPU reference
    cpu = parity = None
    if rank == 0 and N == 1 and not args.no_cpu_baseline:
        r = cpu_reference_sample(blobs, pc, W, H, args.cpu_sample, textures=info.get("texture_images"))
        cpu = dict(value=r["mpaths"], unit=UNIT, cores=r["cores"], kind=r["kind"],
                   sample="%s; %.1f s, OpenMP over rows, all host cores" % (r["lattice"], r["seconds"]),
                   what="shaders/raytrace.comp compiled for the CPU (oracle/_ref/libraytrace_comp.so)" if r["kind"] == "reference"
                        else "oracle/liboracle.so (hand-written restatement)")
        w, h, n, sx, sy = _lattice(W, H, args.parity_sample)
        cp.clear()
        cp.render_frames(pcbuf, 0, n, rows=(0, H), row_stride=sy)
        got = cp.read_accum()[0:h * sy:sy, 0:w * sx:sx][:h, :w]
        rp = cpu_reference_sample(blobs, pc, W, H, args.parity_sample, textures=info.get("texture_images"))
        ref = rp["image"]
        a, b = got[..., :3].astype(np.float64), ref[..., :3].astype(np.float64)
        rel = float(np.sqrt(np.mean((a - b) ** 2)) / np.mean(b))
        err = np.abs(a - b).max(axis=2) / n
        parity = dict(rel_rmse=rel, tolerance=1e-3, ok=bool(rel <= 1e-3), spp=n, pixels=w * h,
                      against="%s (host libm transcendentals), %s" % (
                          "shaders/raytrace.comp compiled for the CPU (oracle/_ref/libraytrace_comp.so)" if rp["kind"] == "reference" else "oracle/liboracle.so", rp["lattice"]),
                      pixels_bit_equal=float((got.view(np.uint32) == ref.view(np.uint32)).all(axis=2).mean()),
                      max_abs_error_of_the_mean=float(err.max()), mean_radiance=float(b.mean() / n), cpu_seconds=rp["seconds"],
                      note="GPU kernels use their own sin/cos/pow/tan (<= 1 ulp from libm); a last-bit difference in a sampled direction can send a "
                           "path to another triangle after a few bounces on curved geometry, so individual paths differ while the estimate agrees")

    # ---- C5 of BASELINE.json: atrium + statue, 3840x2160, depth 16 (frames split over the ranks + reduce)
    c5 = None
    if args.c5_frames > 0 and args.workload == "atrium" and not args.obj:
        cp.close()
        cp = None
        blobs5, pc5, cfg5, _ = make_workload(args, workload="c5", size=(3840, 2160), bounce=16)
        F5 = args.c5_frames
        cp5 = new_pass(3840, 2160, blobs5, min(F5, 12))
        pin5 = torch.empty(2160, 3840, 4, dtype=torch.float32, pin_memory=True).numpy() if rank == 0 else None

        def c5_run(first):
            cp5.clear()
            cp5.render_frames_split(pc5.buffer, first, F5 * N)
        c5_run(0)
        cp5.reduce_accum(0, zero_others=True)          # first collective of this communicator: NCCL sets its channels up here
        ms_c5, _, _ = device_ms(lambda: c5_run(F5 * N))
        ms_red, _, _ = device_ms(lambda: cp5.reduce_accum(0, zero_others=True))
        t_d2h = time.perf_counter()
        if rank == 0:
            cp5.read_accum(out=pin5)
        t_d2h = time.perf_counter() - t_d2h
        c5 = dict(workload=cfg5["workload"], width=3840, height=2160, bounce_limit=16, triangles=cfg5["triangles"], frames_per_gpu=F5,
                  frames_total=F5 * N, mpaths_per_s=3840 * 2160 * F5 * N / ms_c5 / 1e3, render_ms=ms_c5, reduce_ms=ms_red,
                  reduce_bytes=3840 * 2160 * 16, d2h_s=t_d2h,
                  note="4096 spp of BASELINE.json configs[4] = %.0f s at this rate on %d GPU(s)" % (3840 * 2160 * 4096 / (3840 * 2160 * F5 * N / ms_c5 * 1e3), N))
        if rank == 0:
            c5["mean_radiance"] = float(pin5[..., :3].astype(np.float64).mean() / max(float(pin5[..., 3].mean()), 1.0))
        cp5.close()

    if rank == 0:
        # ---- "work avoided": the reference's algorithmic bytes / flops for the same paths (oracle counters, §8 d4)
        rw = cpu_reference_sample(blobs, pc, W, H, "240x135x1", want_work=True)
        b_all = algorithmic_bytes(rw["work_all"], rw["paths"]) + 32.0
        b_ext = algorithmic_bytes(rw["work_ext"], rw["paths"])
        wx = rw["work_ext"]
        flops_ext = (24.0 * wx["n_box"] + 64.0 * wx["n_tri"] + 20.0 * wx["n_sph"] + 150.0 * wx["n_obj"]) / max(rw["paths"], 1)
        avoided = dict(reference_layout_bytes_per_path=dict(all_rays=b_all, extension_rays=b_ext, nee_and_pdf_rays=b_all - b_ext - 32.0),
                       reference_flops_per_path_extension_rays=flops_ext,
                       reference_rays_per_path=rw["work_all"]["n_ray"] / max(rw["paths"], 1),
                       note="what shaders/raytrace.comp would touch for the same paths in its own layout (32 B/box test, 144 B/triangle "
                            "test, 80 B/object visit; four traversals per diffuse bounce) — the work this design removes, not a roofline")
        peaks = load_json("MEASURED_PEAKS.json")
        pm = load_json("profiles", "peaks_micro.json")
        tj = load_json("profiles", "traffic.json")
        roofline = build_roofline(args, cfg, F, args.steps, stages, counted, peaks, pm, tj, avoided, wide_sizes(info))
        cfg.update(parallelism="spp-split x%d (frames f = r mod N) + ncclReduce per step (b200pt_reduce_accum)" % N if N > 1 else "single GPU",
                   l2="path state per step %.0f MB >> 126 MB L2; scene (%.0f MB) is re-read by every frame by design; no flush"
                      % (W * H * F * 160 / 1e6, cfg["scene_reference_layout_mb"]),
                   frames_rendered=frame_cursor[0], mean_radiance=mean_rgb, scene_upload_s=round(upload_s, 3), wide_bvh_build=wide_build,
                   primary_rays="traced once per pixel and shared by the %d frame slots of a batch (no jitter: raytrace.comp:541)" % F,
                   rays_per_path=dict(traced=counted["extension_rays_traced_per_path"] + counted["occlusion_rays_traced_per_path"] + counted["pdf_rays_traced_per_path"],
                                      reference=avoided["reference_rays_per_path"]))
        out.update(roofline=roofline, cpu_baseline=cpu, parity=parity, counted_work=counted, strong_scaling=strong, c5=c5, update_objects=upd,
                   interactive=dict(ms_per_frame=ms_inter, ms_per_frame_without_exact_retrace=ms_inter_fast, frames=n_inter, run_compute_ms=ms_dispatch, run_compute_pipeline0_ms=ms_dispatch_ref,
                                    split="interleaved rows over %d GPU(s), gathered on rank 0 inside the library (ncclSend/Recv)" % N,
                                    note="one 1-spp progressive frame at a time (latency), vs ms_per_frame_1spp = throughput mode / frames"))
        if args.dump:
            np.save(args.dump, pinned_np[..., :3] / np.maximum(pinned_np[..., 3:4], 1))
        emit(out)
    if cp is not None:
        cp.close()
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()


def wide_sizes(info):
    """(node array MB, triangle record MB) of the internal wide layout: ~0.17 nodes of 80 B and one 64-byte record per triangle."""
    t = float(info["triangles"])
    return 0.17 * t * NODE_BYTES / 1e6, t * PRIM_BYTES / 1e6


def bench_soup(args, rank, N, dev, stream, new_pass, device_ms):
    """The HBM-bound traversal case (north_star: "synthetic 10M-triangle soup for traversal roofline"): closest hits of
    --soup-rays rays with uniform-random origins in the cube and uniform-random directions, through the same persistent
    kernel as the renderer (b200pt_trace_batch_device). The tree and the triangle records (~0.9 GB) do not fit L2 and
    every ray lands somewhere else, so node and triangle fetches are DRAM gathers."""
    import numpy as np
    import torch
    blobs, pc, cfg, info = make_workload(args)
    cp = new_pass(64, 64, blobs, 1)
    n = args.soup_rays
    g = torch.Generator(device=dev).manual_seed(0x5EED + rank)
    rays = torch.zeros(n, 8, dtype=torch.float32, device=dev)
    rays[:, 0:3] = torch.rand(n, 3, generator=g, device=dev) * 1.9 - 0.95
    d = torch.randn(n, 3, generator=g, device=dev)
    rays[:, 4:7] = d / d.norm(dim=1, keepdim=True)
    hits = torch.zeros(n, 12, dtype=torch.float32, device=dev)

    def step():
        cp.trace_batch_device(pc.buffer, rays.data_ptr(), n, hits.data_ptr())
    for _ in range(max(args.warmup, 1)):
        step()
    cp.set_option("profile_stages", 1)
    cp.stage_times(reset=True)
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0"))) if rank == 0 else None
    ms, t0, t1 = device_ms(lambda: [step() for _ in range(args.steps)])
    clocks = sampler.stop(t0, t1) if sampler else None
    stages = cp.stage_times(reset=True)
    cp.set_option("profile_stages", 0)
    cp.set_option("count_work", 1)
    cp.reset_counters()
    step()
    cp.sync()
    cw = cp.counters()
    cp.set_option("count_work", 0)
    hit_frac = float((hits.view(torch.int32)[:, 1] != 0).float().mean().item())
    if rank == 0:
        peaks = load_json("MEASURED_PEAKS.json")
        pm = load_json("profiles", "peaks_micro.json")
        tj = load_json("profiles", "traffic.json").get("soup", {})
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        ext_ms, launches = stages["extend"]
        nv, tt = cw["extend_node_visits"] / n, cw["extend_tri_tests"] / n
        algo = (nv * NODE_BYTES + tt * PRIM_BYTES) * n * args.steps
        achieved = algo / ext_ms / 1e6
        kk = tj.get("kernels", {}).get("k_extend") if tj.get("rays") == n and tj.get("triangles") == info["triangles"] else None
        traffic = kk["dram_bytes"] if kk else None
        gather = nearest_gather_peak(pm, "node80", 1025.0)
        roof = dict(bound="hbm", kernel="k_extend", achieved=achieved, peak=hbm, unit="GB/s", frac=achieved / hbm, traffic=traffic,
                    peak_source="MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                    launches=launches, avg_launch_ms=ext_ms / max(launches, 1), node_visits_per_ray=nv, tri_tests_per_ray=tt,
                    algorithmic_bytes_per_ray=nv * NODE_BYTES + tt * PRIM_BYTES,
                    note="algorithmic bytes = node visits x 80 B + triangle tests x 64 B counted by the kernel (count_work); records are 16-byte "
                         "aligned, DRAM moves 32-byte sectors, so traffic (ncu) is above this by the sector overhead of misaligned records")
        if traffic:
            roof["dram_frac"] = traffic / (ext_ms / max(launches, 1)) / 1e6 / hbm
        if gather:
            roof["random_gather_ceiling_gbs"] = gather[2]
            roof["frac_of_random_gather"] = achieved / gather[2]
        value = n * args.steps / ms / 1e3
        cfg.update(rays_per_step=n, ray_distribution="origins uniform in [-0.95, 0.95]^3, directions uniform on the sphere", hit_fraction=hit_frac,
                   l2="tree + triangle records %.0f MB >> 126 MB L2, rays incoherent by construction; no flush" % (sum(wide_sizes(info))))
        emit(dict(metric="Mrays_per_s_soup_closest_hit", value=value, unit="Mrays/s", n_gpus=N, steps=args.steps, warmup=args.warmup,
                              ms_per_step=ms / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                              config=cfg, clocks=clocks, gpu_launches=cw["launches"], roofline=roof))
    cp.close()


if __name__ == "__main__":
    main()
