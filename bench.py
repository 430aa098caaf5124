#!/usr/bin/env python
"""Headline benchmark: Mrays/s (and Mpaths/s) of the path-tracing hot path on B200.

Workload (BASELINE.json configs[1], "C2"): procedural 1,002,528-triangle terrain,
one diffuse material + ground plane, default sky, pinhole camera, 1920x1080.
One STEP = `--frames` steady-state frames, i.e. `--frames` x {trace, scatter} over
all 2,073,600 path slots (what the reference's Update() issues once per displayed
frame, src/application/application.cpp:114).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (contract in the task statement): `value` = whole-job Mrays/s
with everything resident in HBM; `e2e` = the same metric through the public API with
host-side inputs/outputs inside the timed region; `roofline` for the dominant kernel
(the trace kernel); `cpu_baseline` = oracle/_ref (the reference's own trace / scatter
GLSL compiled as C++, one forked process per host core; kind "reference") on a bounded
sample of the same workload, or the oracle port where that library is unavailable.
`--impl reference` times that CPU code alone (the reference's own Vulkan path is
unavailable on this box).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs.  C2 is the configuration the metric is quoted on and
# the default; the others are parity-test cases that can also be timed with --workload.
#   key: (description, scene name, scene params, width, height, tone map, path length limit, CPU tile stride)
WORKLOADS = {
    "C1": ("C1: Cornell box of planes/spheres/cubes, diffuse+mirror+glass, pinhole, 640x480", "cornell", (), 640, 480, "aces", 0, 1),
    "C2": ("C2: procedural 1,002,528-triangle terrain + ground plane, diffuse, pinhole, 1920x1080", "terrain", (708,), 1920, 1080, "aces", 0, 4),
    "C3": ("C3: textured 76,800-triangle room + 3 refractive spheres, thin lens, HDR sky texture, Hable, 1920x1080", "room", (1, 4096), 1920, 1080, "hable", 0, 4),
    "C4": ("C4: 9,994,240-triangle sphere field, mixed materials, 360 camera, ACES, 8-bounce, 4096x2048", "spherefield", (488, 5, 2), 4096, 2048, "aces", 8, 16),
    "C5": ("C5: 9,994,240-triangle sphere field, mixed materials, pinhole, 3840x2160", "spherefield", (488, 5, 0), 3840, 2160, "aces", 0, 16),
}
WORKLOAD, SCENE_NAME, SCENE_PARAMS, WIDTH, HEIGHT, TONE, PATH_LIMIT, CPU_STRIDE = WORKLOADS["C2"]
GRID = 708


def select_workload(args):
    global WORKLOAD, SCENE_NAME, SCENE_PARAMS, WIDTH, HEIGHT, TONE, PATH_LIMIT, CPU_STRIDE
    WORKLOAD, SCENE_NAME, SCENE_PARAMS, WIDTH, HEIGHT, TONE, PATH_LIMIT, CPU_STRIDE = WORKLOADS[args.workload]
    if args.workload == "C2" and args.grid != GRID:
        SCENE_PARAMS = (args.grid,)
        WORKLOAD += " (grid %d)" % args.grid
    if args.ref_tile_stride <= 0:
        args.ref_tile_stride = CPU_STRIDE


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(p.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.p = None
        self.lines = []
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        sm, mx, reasons = [], [], set()
        for ts, line in self.lines:
            if ts < t0 - 0.05 or ts > t1 + 0.05:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def oracle_sample(scene, frames, tile_stride, threads):
    """The oracle on every `tile_stride`-th 16x16 tile of the full-size film: restart + `frames` frames."""
    from oracle import binding as O
    r = O.Renderer(scene.view, WIDTH, HEIGHT, tie=O.TIE_FAITHFUL, threads=threads, tile_stride=tile_stride, path_length_limit=PATH_LIMIT)
    r.set_frame_index(0)
    r.reset()
    r.run(2)                                  # warm-up: restart rounds
    s0 = r.stats()
    t0 = time.perf_counter()
    r.run_frames(frames)
    dt = time.perf_counter() - t0
    s1 = r.stats()
    d = {k: s1[k] - s0[k] for k in s1}
    d["seconds"] = dt
    return d


def _ref_worker(view_addr, p, procs, stride, warm_frames, frames, barrier, out):
    """One host core's share of the reference-shader run: its own copy of libref.so's globals (fork), every
    (stride * procs)-th 16x16 work group starting at p * stride."""
    import ctypes as C
    from oracle.ref import binding as R
    from path_tracer_b200.types import SceneView
    view = SceneView.from_address(view_addr)
    r = R.Renderer(view, WIDTH, HEIGHT, path_length_limit=PATH_LIMIT)
    r.set_groups(stride * procs, p * stride)
    r.set_frame_index(0)
    r.reset()
    r.run(2)
    r.run_frames(warm_frames)
    rays0, paths0 = r.rays_traced(), float(r.film[..., 3].sum(dtype=np.float64))
    barrier.wait()
    t0 = time.perf_counter()
    r.run_frames(frames)
    t1 = time.perf_counter()
    out.put((p, r.rays_traced() - rays0, float(r.film[..., 3].sum(dtype=np.float64)) - paths0, t0, t1))


def reference_shader_sample(scene, stride, warm_frames, frames, procs):
    """oracle/_ref (the reference's own trace / scatter GLSL compiled as C++) on every `stride`-th 16x16 tile of the
    full-size film, the tiles dealt round-robin to `procs` forked processes (the library is single-threaded: the
    shaders' interface blocks are globals).  Returns None when the library is unavailable, the film needs more path
    slots than the reference's TRACE_COUNT (2 Mi), or the workload sets a path length limit."""
    import ctypes as C
    import multiprocessing as mp
    from oracle.ref import binding as R
    if not R.available():
        return None
    try:
        L = R.lib()
    except Exception:
        return None
    if R.Renderer.slots_needed(WIDTH, HEIGHT) > L.ref_slot_capacity() or WIDTH % 16 or PATH_LIMIT:
        return None                              # (the shader never reads PathLengthLimit: a limited workload needs the port)
    ctx = mp.get_context("fork")
    barrier, out = ctx.Barrier(procs), ctx.Queue()
    addr = C.addressof(scene.view)
    ws = [ctx.Process(target=_ref_worker, args=(addr, p, procs, stride, warm_frames, frames, barrier, out)) for p in range(procs)]
    for w in ws:
        w.start()
    res = [out.get() for _ in ws]
    for w in ws:
        w.join()
    dt = max(r[4] for r in res) - min(r[3] for r in res)
    return {"rays": sum(r[1] for r in res), "paths": sum(r[2] for r in res), "seconds": dt, "procs": procs}


def bytes_per_segment(counts):
    """32*N_node + 48*N_face + 144*N_shape + 24*[mesh hit] + 256 per ray segment,
    counts taken from the oracle's traversal of the reference BVH2 for the same workload."""
    rays = max(1, counts["rays"])
    trav = (32.0 * counts["node_fetches"] + 48.0 * counts["face_tests"] + 144.0 * counts["shape_visits"] + 24.0 * counts["mesh_hits"]) / rays
    return trav + 256.0, trav + 48.0          # whole segment; the trace kernel's share (ray 32 B in + hit 16 B out)


def run_reference(args, rank, world):
    if rank != 0:
        return
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    sc = S.Scene.make(SCENE_NAME, *SCENE_PARAMS)
    threads = os.cpu_count() or 1
    stride = args.ref_tile_stride
    kind, note = "reference", ("the reference's own basic_trace.glsl / basic_scatter.glsl compiled as C++ from its sources (oracle/_ref), "
                                "one process per host core; its Vulkan path is unavailable on this box")
    c = None if args.ref_port else reference_shader_sample(sc, stride, args.warmup * args.ref_frames, args.steps * args.ref_frames, threads)
    if c is not None:
        rays, paths, dt = c["rays"], c["paths"], c["seconds"]
    else:
        kind, note = "port", ("CPU restatement of the reference's trace/scatter shaders (oracle port: " + ("selected with --ref-port" if args.ref_port else
                              "oracle/_ref is not available for this workload") + "); the reference's own Vulkan/GLSL path is unavailable on this box")
        from oracle import binding as O
        r = O.Renderer(sc.view, WIDTH, HEIGHT, tie=O.TIE_FAITHFUL, threads=threads, tile_stride=stride, path_length_limit=PATH_LIMIT)
        r.set_frame_index(0)
        r.reset()
        r.run(2)
        for _ in range(args.warmup):
            r.run_frames(args.ref_frames)
        s0 = r.stats()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            r.run_frames(args.ref_frames)
        dt = time.perf_counter() - t0
        s1 = r.stats()
        rays, paths = s1["rays"] - s0["rays"], s1["paths"] - s0["paths"]
    mrays = rays / dt / 1e6
    sample = "every %d-th 16x16 tile of the %dx%d film (%d rays/frame), %d frames/step" % (stride, WIDTH, HEIGHT, rays // max(1, args.steps * args.ref_frames), args.ref_frames)
    _emit(({
        "impl": "reference", "metric": "Mrays/s (ray segments traced, whole job)", "value": mrays, "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": note},
        "mpaths_per_s": paths / dt / 1e6,
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


_RESULT_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version line on the first
    communicator), so everything else is sent to stderr and the result line goes to the original descriptor."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def _emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_RESULT_FD, line)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=16, help="steady-state frames (rounds) per step")
    ap.add_argument("--integrator", default="wavefront", choices=["wavefront", "megakernel"])
    ap.add_argument("--traversal", default="wide", choices=["wide", "bvh2"])
    ap.add_argument("--trace-variant", type=int, default=-1, help="PTB_TRACE_VARIANT_* (default: the library's)")
    ap.add_argument("--scatter-mode", default="auto", choices=["auto", "direct", "queues", "direct-staged"])
    ap.add_argument("--pipelines", type=int, default=0, choices=[0, 1, 2], help="0 = the library's choice")
    ap.add_argument("--builder", default="sah-host", choices=["sah-host", "lbvh-gpu", "ploc-gpu", "auto"], help="who flattens the scene into the BVH8 (ptb_scene_set_builder)")
    ap.add_argument("--grid", type=int, default=GRID)
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS), help="BASELINE.json config (default C2, the one the metric is quoted on)")
    ap.add_argument("--ref-frames", type=int, default=8, help="frames per step of the CPU reference arm")
    ap.add_argument("--ref-tile-stride", type=int, default=0, help="CPU arms process every n-th 16x16 tile (bounded sample)")
    ap.add_argument("--cpu-frames", type=int, default=96, help="frames of the cpu_baseline sample (about 10-30 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-port", action="store_true", help="time the oracle port instead of oracle/_ref (the reference's shaders compiled as C++)")
    ap.add_argument("--reduce-mode", default="p2p", choices=["p2p", "nccl", "sliced"], help="N > 1: how the end-to-end loop combines the films (ptb_film_reduce_resolve mode)")
    ap.add_argument("--spp-target", type=int, default=0, help="also measure the seconds from reset until the combined film holds this many samples per pixel")
    ap.add_argument("--spp-criterion", default="mean", choices=["mean", "min"], help="samples per pixel on average (finished paths / pixels: what 'N spp' means "
                    "for a progressive renderer, and what bin/ptb_headless --spp uses) or in the least-sampled pixel")
    args = ap.parse_args()
    select_workload(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    from path_tracer_b200 import dist as D
    from path_tracer_b200 import integrator as I
    from path_tracer_b200 import scene as S

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the integrator has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    rank, world = D.init_process_group()

    # ---- setup: the scene is packed and flattened ONCE (rank 0) and sent device-to-device to the other ranks ----
    S.ensure_spectrum_table()
    ctx = I.Context(local_rank)
    group = D.create_group(ctx, device)                  # NCCL communicator inside libptb.so (None for one GPU)
    builder = {"sah-host": I.BUILDER_SAH_HOST, "lbvh-gpu": I.BUILDER_LBVH_GPU, "ploc-gpu": I.BUILDER_PLOC_GPU, "auto": I.BUILDER_AUTO}[args.builder]
    setup = {}
    t0 = time.perf_counter()
    if rank == 0:
        scene = S.Scene.make(SCENE_NAME, *SCENE_PARAMS)
        setup["scene_make_s"] = time.perf_counter() - t0
        t1 = time.perf_counter()
        dscene = I.DeviceScene(ctx, scene, builder=builder)
        setup["upload_and_flatten_s"] = time.perf_counter() - t1
        small = (bytes(scene.cameras.tobytes()), bytes(np.asarray(scene.globals).tobytes()))
    else:
        scene, small = None, None
        dscene = I.DeviceScene(ctx)
    if group is not None:
        t1 = time.perf_counter()
        group.replicate_scene(dscene, 0)
        ctx.sync()
        setup["replicate_s"] = time.perf_counter() - t1
        small = D.broadcast_bytes(small, 0, device)
    setup["total_s"] = time.perf_counter() - t0
    # the per-step host inputs of the end-to-end loop (camera + globals records) as a view of their own, on every rank
    cam_host = np.frombuffer(small[0], dtype=np.uint8).copy()
    glob_host = np.frombuffer(small[1], dtype=np.uint8).copy()
    from path_tracer_b200.types import SceneView
    step_view = SceneView()
    step_view.Cameras, step_view.CameraCount, step_view.Globals = cam_host.ctypes.data, len(cam_host) // 160, glob_host.ctypes.data
    wide = dscene.wide_info()
    film = I.SampleBuffer(ctx, WIDTH, HEIGHT)             # library-owned: peer-addressable for the fused exchange
    total_film = I.SampleBuffer(ctx, WIDTH, HEIGHT) if (group is not None and rank == 0) else None
    r = I.BasicRenderer(ctx, dscene, film)
    r.Integrator = I.INTEGRATOR_MEGAKERNEL if args.integrator == "megakernel" else I.INTEGRATOR_WAVEFRONT
    r.Traversal = I.TRAVERSAL_WIDE if args.traversal == "wide" else I.TRAVERSAL_BVH2_REFERENCE
    r.TieMode = I.TIE_CANONICAL
    r.PathLengthLimit = PATH_LIMIT
    r.Pipelines = args.pipelines
    r.ScatterMode = {"auto": I.SCATTER_AUTO, "direct": I.SCATTER_DIRECT, "queues": I.SCATTER_QUEUES, "direct-staged": I.SCATTER_DIRECT_STAGED}[args.scatter_mode]
    if args.trace_variant >= 0:
        r.TraceVariant = args.trace_variant
    F = args.frames
    total_frames = F * (2 * max(3, args.warmup) + 2 * args.steps + 16) + max(3, args.warmup) + max(args.steps, 20) + 64   # all timed loops, their warm-ups, the profiled frames
    r.set_frame_index(D.frame_schedule(rank, world, total_frames))                      # disjoint seed block per rank
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)

    r.reset()
    r.run(2)
    for _ in range(max(3, args.warmup)):
        r.run_frames(F)
    r.sync()

    # ---- device-resident timing: K steps, CUDA events on the launching stream, max over ranks ----
    D.barrier()
    torch.cuda.synchronize()
    s0 = r.stats()
    sampler = ClockSampler(local_rank)
    time.sleep(0.25)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps):
            r.run_frames(F)
        if group is not None:
            group.reduce(film, 0, total_film)     # the one exchange step of the sample-split job: ptb_film_reduce (ncclReduce over NVLink)
        e1.record(stream)
    e1.synchronize()
    torch.cuda.synchronize()
    w1 = time.time()
    D.barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop(w0, w1)
    s1 = r.stats()
    rays = s1["rays_traced"] - s0["rays_traced"]
    paths = s1["paths_finished"] - s0["paths_finished"]
    launches = s1["kernel_launches"] - s0["kernel_launches"]
    ms_max = D.max_over_ranks(ms, device)
    rays_all = D.sum_over_ranks(float(rays), device)
    paths_all = D.sum_over_ranks(float(paths), device)
    value = rays_all / (ms_max * 1e-3) / 1e6

    # ---- end to end through the public API with host buffers.  Per step: the camera + globals records go host -> device
    #      (ptb_scene_update), the frames run, and ONE resolved sRGB8 frame of the whole job arrives in pinned host memory on
    #      rank 0 - at N > 1 that is ptb_film_reduce_resolve: every rank's film summed and tone-mapped by the fused
    #      peer-memory kernel (NCCL reduce + root resolve where peer access is unavailable) - plus the stats read-back. ----
    D.barrier()
    torch.cuda.synchronize()
    h2d = len(cam_host) + len(glob_host)
    d2h = WIDTH * HEIGHT * 4 + 4 * (r.slot_count // 256)
    frame_host = torch.empty((HEIGHT, WIDTH, 4), dtype=torch.uint8, pin_memory=True).numpy()     # pinned destination of the resolved frame
    tone = I.TONE_HABLE if TONE == "hable" else I.TONE_ACES
    exchange = {"mode": None, "ms": None}
    reduce_mode = {"p2p": I.REDUCE_P2P_FUSED, "nccl": I.REDUCE_NCCL, "sliced": I.REDUCE_NCCL_SLICED}[args.reduce_mode]

    def deliver_frame():
        nonlocal reduce_mode
        if group is None:
            film.resolve(1.0, tone, 1.0, want_linear=False, out_srgb8=frame_host)
            return
        try:
            group.reduce_resolve(film, (1.0, tone, 1.0), mode=reduce_mode, root=0, out_srgb8=frame_host if rank == 0 else None)
        except I.PtbError as e:              # the verdict is collective: every rank lands here together
            if reduce_mode == I.REDUCE_NCCL or e.code != -4:
                raise
            sys.stderr.write("bench.py: %s; falling back to ncclReduce + root resolve\n" % e)
            reduce_mode = I.REDUCE_NCCL
            group.reduce_resolve(film, (1.0, tone, 1.0), mode=reduce_mode, root=0, out_srgb8=frame_host if rank == 0 else None)

    def e2e_loop(frames, steps):
        for _ in range(max(3, args.warmup)):          # untimed warm-up of this path too (first resolve allocates and loads its kernel)
            dscene.update(step_view, 1 | 32)
            r.run_frames(frames)
            deliver_frame()
            r.stats()
        D.barrier()
        a = r.stats()
        t0 = time.perf_counter()
        for _ in range(steps):
            dscene.update(step_view, 1 | 32)          # SCENE_DIRTY_GLOBALS | SCENE_DIRTY_CAMERAS (synchronises, like the reference)
            r.run_frames(frames)
            deliver_frame()
            b = r.stats()
        dt = D.max_over_ranks(time.perf_counter() - t0, device)
        return D.sum_over_ranks(float(b["rays_traced"] - a["rays_traced"]), device) / dt / 1e6

    e2e_value = e2e_loop(F, args.steps)
    if group is not None:
        t = group.last_reduce()
        exchange = {"mode": {0: "ncclReduce + root resolve", 1: "reduce-scatter + sliced resolve + all-gather of sRGB8", 2: "fused peer-memory reduce + resolve kernel"}[t["mode"]],
                    "ms": D.max_over_ranks(t["exchange_ms"], device), "film_bytes": t["bytes_in"]}
    e2e_one = e2e_loop(1, max(args.steps, 20))        # what the reference does: present after EVERY frame (application.cpp:114)

    # ---- per-kernel device time (events around every launch) for the roofline of the trace kernel ----
    kt = r.profile_frames(F)
    trace_ms = (kt["trace_ms"] / max(1, kt["trace_launches"])) if kt["trace_launches"] else (kt["mega_ms"] / max(1, kt["mega_launches"]))

    # ---- instrumented launch of the default trace kernel over the current rays: what its own layout fetches ----
    tstats = None
    if args.integrator == "wavefront" and args.traversal == "wide":
        try:
            tstats = r.trace_stats()
        except I.PtbError:
            tstats = None

    # ---- optional: seconds from reset until every pixel of the COMBINED film holds >= target samples and the resolved frame
    #      is in rank 0's host memory (BASELINE config 5 at N > 1: strong scaling, the exchange inside) ----
    spp = None
    if args.spp_target > 0:
        class _Cai:                               # the library-owned film as a torch tensor (no copy), for min() on the device
            def __init__(self, ptr, shape):
                self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (ptr, False), "version": 2}
        film_t = torch.as_tensor(_Cai(film.device_ptr, (HEIGHT, WIDTH, 4)), device=device)
        per_rank = -(-args.spp_target // world)   # mean: the means add up; min: min over pixels of a sum >= sum of the per-rank minima
        reached = (lambda: float(film_t[..., 3].mean())) if args.spp_criterion == "mean" else (lambda: float(film_t[..., 3].min()))
        chunk = 64
        r.set_frame_index(rank << 22)             # disjoint seed blocks of 4 Mi frames
        D.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r.reset()
        frames_done = 0
        while True:
            r.run_frames(chunk)
            frames_done += chunk
            r.sync()
            if reached() >= per_rank or frames_done >= (1 << 22) - chunk:
                break
        t_render = time.perf_counter() - t0
        deliver_frame()
        seconds = D.max_over_ranks(time.perf_counter() - t0, device)
        spp = {"target": args.spp_target, "criterion": args.spp_criterion, "per_rank_target": per_rank, "seconds": seconds, "render_seconds_this_rank": t_render,
               "frames_this_rank": frames_done, "frames_max": int(D.max_over_ranks(float(frames_done), device)),
               "min_alpha_this_rank": float(film_t[..., 3].min()), "mean_alpha_this_rank": float(film_t[..., 3].mean()),
               "exchange_ms": (D.max_over_ranks(group.last_reduce()["exchange_ms"], device) if group is not None else 0.0),
               "what": "reset -> frames until the %s-over-pixels alpha >= ceil(target / n_gpus) on every rank -> exchange + resolve -> sRGB8 frame in rank 0's pinned host memory" % args.spp_criterion}
        del film_t

    if group is not None:
        group.close()
    D.shutdown()
    if rank != 0:
        return

    # ---- CPU baseline + algorithmic byte model from the oracle on a bounded sample (rank 0, N=1 only) ----
    hbm_peak, peak_src, sm_max = load_peaks()
    cpu = None
    seg_bytes, trace_bytes = None, None
    if not args.no_cpu_baseline and world == 1:
        threads = os.cpu_count() or 1
        # the reference's own shader code (oracle/_ref) gives the timed baseline; the oracle port, which counts what
        # the traversal touches, gives the algorithmic byte model (and the baseline where _ref is unavailable)
        shader = None if args.ref_port else reference_shader_sample(scene, args.ref_tile_stride, 2, args.cpu_frames, threads)
        c = oracle_sample(scene, args.cpu_frames if shader is None else max(8, args.cpu_frames // 4), args.ref_tile_stride, threads)
        seg_bytes, trace_bytes = bytes_per_segment(c)
        t = shader if shader is not None else c
        cpu = {"value": t["rays"] / t["seconds"] / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference" if shader is not None else "port",
               "sample": "every %d-th 16x16 tile of the same %dx%d film, %d frames, %d rays in %.1f s"
                         % (args.ref_tile_stride, WIDTH, HEIGHT, args.cpu_frames, t["rays"], t["seconds"]),
               "mpaths_per_s": t["paths"] / t["seconds"] / 1e6,
               "port_value": c["rays"] / c["seconds"] / 1e6,
               "per_ray": {"node_fetches": c["node_fetches"] / c["rays"], "face_tests": c["face_tests"] / c["rays"],
                           "shape_visits": c["shape_visits"] / c["rays"], "mesh_hit_fraction": c["mesh_hits"] / c["rays"]}}
    bytes_source = "oracle traversal counters of this run's CPU sample"
    if trace_bytes is None and args.workload == "C2" and args.grid == GRID:
        # no CPU sample in this run (N>1 or --no-cpu-baseline): the per-ray figure the oracle counted for this
        # workload in the committed full run (profiles/r01_v1_bench.json: 82.5 nodes, 7.05 faces, 1.92 shapes per ray)
        seg_bytes, trace_bytes = 3522.405, 3314.405
        bytes_source = "oracle traversal counters recorded in profiles/r01_v1_bench.json (same workload)"
    from path_tracer_b200 import _build
    build_hash = _build.source_hash()
    live = WIDTH * HEIGHT
    trace_s = trace_ms * 1e-3
    issue_peak = 148 * 4 * sm_max * 1e6 / 1e9            # G warp-instructions / s: 148 SMs x 4 schedulers x one issue per clock
    # ncu counters of the dominant kernels, per launch: accepted only from a capture of THIS build (source hash) of THIS workload
    ncu, ncu_note = None, "no ncu capture of this build for this workload in profiles/traffic.json"
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(args.workload)
        default_path = args.integrator == "wavefront" and args.traversal == "wide" and args.trace_variant < 0 and args.builder == "sah-host"
        if t and not default_path:
            ncu_note = "profiles/traffic.json describes the default kernels; this run selected others"
        elif t and t.get("build_hash") != build_hash:
            ncu_note = "profiles/traffic.json was captured from build %s, this is build %s: not quoted" % (t.get("build_hash"), build_hash)
        elif t:
            ncu, ncu_note = t, t.get("source")
    except Exception:
        pass
    roofline = None
    if kt["trace_launches"]:
        own = None
        if tstats and tstats["rays"]:
            own_bytes = tstats["node_bytes"] * tstats["node_steps"] + tstats["prim_bytes"] * tstats["prim_tests"] + tstats["state_bytes"] * tstats["rays"]
            own = {"bytes_per_launch": own_bytes, "bytes_per_ray": own_bytes / tstats["rays"], "node_steps_per_ray": tstats["node_steps"] / tstats["rays"],
                   "prim_tests_per_ray": tstats["prim_tests"] / tstats["rays"], "achieved_gbs": own_bytes / trace_s / 1e9, "frac_of_hbm_peak": own_bytes / trace_s / 1e9 / hbm_peak,
                   "lanes_per_node_step": tstats["node_steps"] / max(1, tstats["warp_iterations"]), "lanes_per_primitive_flush": tstats["prim_tests"] / max(1, tstats["flushes"]),
                   "what": "80 B x BVH8 node steps + 48 B x leaf primitive tests + 48 B wavefront state per ray, counted by one instrumented launch of the "
                           "same kernel over this run's rays (ptb_renderer_trace_stats); most of it is served by L1/L2, which is why it may exceed the HBM peak"}
        inst = ncu["trace"]["warp_instructions_per_launch"] if ncu else None
        achieved = inst / trace_s / 1e9 if inst else None
        roofline = {"bound": "sm_issue", "kernel": "k_trace_coop", "achieved": achieved, "peak": issue_peak, "unit": "Gwarp-inst/s",
                    "frac": achieved / issue_peak if achieved else None,
                    "traffic": ncu["trace"]["dram_bytes_per_launch"] if ncu else None,
                    "warp_instructions_per_launch": inst, "lanes_per_instruction": ncu["trace"]["lanes_per_instruction"] if ncu else None,
                    "frac_of_lane_issue_peak": (achieved / issue_peak * ncu["trace"]["lanes_per_instruction"] / 32.0) if achieved else None,
                    "kernel_ms": trace_ms, "rays_per_launch": live, "peak_source": "148 SMs x 4 schedulers x %.0f MHz (MEASURED_PEAKS.json sm_max_mhz)" % sm_max,
                    "counters_source": ncu_note, "build_hash": build_hash,
                    "own_layout": own,
                    "reference_layout_model": None if trace_bytes is None else {
                        "bytes_per_ray": trace_bytes, "achieved_gbs": live * trace_bytes / trace_s / 1e9, "bytes_source": bytes_source,
                        "what": "SURVEY 8(d): what the REFERENCE's uncached BVH2 walk would fetch for these rays (32 B nodes, 48 B faces, 144 B shapes); "
                                "a comparison figure, not this kernel's traffic"},
                    "note": "software BVH traversal on the SMs is bound by instruction issue, not by HBM (DRAM traffic per launch is a few % of what HBM could move in the "
                            "kernel's time); frac = warp instructions (ncu, this build) / (kernel time measured in this run x issue peak)"}
    roofline_scatter = None
    if kt["scatter_launches"]:
        scatter_s = kt["scatter_ms"] / kt["scatter_launches"] * 1e-3
        finished = paths / max(1, args.steps * F)            # paths finished per round in the timed loop (this rank)
        sb = 208.0 * live + 32.0 * finished                    # 7 x 16 B in, 5 x 16 B out... per slot: hit0, hit1, ray, path0-3 in; ray, path0-3 out (+ the pixel RMW of a finished path)
        roofline_scatter = {"bound": "hbm", "kernel": "k_scatter (direct) or k_classify + k_scatter_list per material queue", "achieved": sb / scatter_s / 1e9, "peak": hbm_peak,
                            "unit": "GB/s", "frac": sb / scatter_s / 1e9 / hbm_peak, "traffic": ncu["scatter"]["dram_bytes_per_launch"] if (ncu and ncu.get("scatter")) else None,
                            "kernel_ms": scatter_s * 1e3, "algorithmic_bytes_per_slot": 208, "algorithmic_bytes_per_finished_path": 32, "finished_paths_per_round": finished,
                            "peak_source": peak_src}
    out = {
        "metric": "Mrays/s (ray segments traced, whole job)", "value": value, "unit": "Mrays/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_max / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": F, "rays_per_step_per_gpu": WIDTH * HEIGHT * F,
                   "integrator": args.integrator, "traversal": args.traversal, "trace_variant": r.TraceVariant, "scatter_mode": args.scatter_mode, "pipelines": args.pipelines, "parallelism": "sample-split x%d" % world,
                   "l2": "per-round working set (232 MB wavefront state + ~170 MB scene) exceeds the 126 MB L2; no explicit flush",
                   "wide_bvh": dict({k: wide[k] for k in ("node_count", "prim_count", "depth", "build_seconds", "sah_cost", "device_build_ms")}, builder=args.builder)},
        "mpaths_per_s": paths_all / (ms_max * 1e-3) / 1e6,
        "grays_per_s_per_gpu": value / world / 1e3,
        "seconds_to_1024_paths_per_pixel_1080p": 1024.0 * WIDTH * HEIGHT / (paths_all / world / (ms_max * 1e-3)) if paths_all else None,
        "time_to_spp": spp,
        "gpu_launches": int(launches),
        "kernel_ms": {"trace": kt["trace_ms"] / max(1, kt["trace_launches"]), "scatter": kt["scatter_ms"] / max(1, kt["scatter_launches"]),
                      "mega": kt["mega_ms"] / max(1, kt["mega_launches"])},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "frames_per_step": F,
                "frames_per_step_1": {"value": e2e_one, "unit": "Mrays/s", "what": "the same loop with ONE frame per step: a resolved frame on the host after every frame, as the reference presents (application.cpp:114)"},
                "exchange": exchange,
                "what": "per step: ptb_scene_update(camera+globals from host) + run_frames + ONE resolved sRGB8 frame of the whole job in pinned host memory on rank 0 "
                        "(ptb_film_resolve; at N > 1 ptb_film_reduce_resolve over all ranks' films) + stats; d2h is rank 0's"},
        "setup_seconds": setup,
        "roofline": roofline,
        "roofline_scatter": roofline_scatter,
        "cpu_baseline": cpu,
    }
    _emit(out)


if __name__ == "__main__":
    main()
