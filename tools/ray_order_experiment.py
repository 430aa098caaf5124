#!/usr/bin/env python
"""Upper bound of what RAY REORDERING could buy the trace kernel.

Takes the rays of one steady-state round of a workload (the wavefront's own ray records, downloaded), and times the
same warp-cooperative trace kernel (ptb_raycast_device, device buffers, no copies) over them in different ORDERS:
  slot      the order the wavefront traces them in (tile-major slots; what ships)
  random    a random permutation (no coherence at all: the lower bound)
  octant    stable sort by direction octant (3 bits) - the binning the verdict proposes
  oct+mort  octant, then 30-bit Morton code of the origin
  morton    Morton code of the origin, then octant
  primary   (reference point) the camera rays right after a reset: the most coherent batch this renderer ever sees
A perfect reordering pass costs at least one extra read + write of the 16-byte ray records plus the sort itself and is
NOT included: the figures are what the kernel alone would gain.

    python tools/ray_order_experiment.py [--workload C2] [--reps 20]      (GPU box)"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def unpack_unit_vector(p):
    """common.glsl.inc:122-128 in numpy (float32)."""
    x = (p & 0xFFFF).astype(np.uint16).view(np.int16).astype(np.float32) / np.float32(32767.0)
    y = (p >> 16).astype(np.uint16).view(np.int16).astype(np.float32) / np.float32(32767.0)
    x, y = np.clip(x, -1, 1), np.clip(y, -1, 1)
    z = np.float32(1.0) - np.abs(x) - np.abs(y)
    neg = z < 0
    sx, sy = np.where(x >= 0, 1.0, -1.0).astype(np.float32), np.where(y >= 0, 1.0, -1.0).astype(np.float32)
    nx = np.where(neg, (1 - np.abs(y)) * sx, x)
    ny = np.where(neg, (1 - np.abs(x)) * sy, y)
    v = np.stack([nx, ny, z], -1).astype(np.float32)
    return v / np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)


def morton30(o, lo, hi):
    q = np.clip(((o - lo) / np.maximum(hi - lo, 1e-20) * 1023.0), 0, 1023).astype(np.uint32)

    def spread(v):
        v = v.astype(np.uint64)
        v = (v | (v << 16)) & 0x030000FF
        v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3
        v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--frames", type=int, default=24)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import bench
    from path_tracer_b200 import integrator as I
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    desc, name, params, w, h, tone, limit, _ = bench.WORKLOADS[args.workload]
    sc = S.Scene.make(name, *params)
    ctx = I.Context(0)
    dev = I.DeviceScene(ctx, sc)
    film = I.SampleBuffer(ctx, w, h)
    r = I.BasicRenderer(ctx, dev, film)
    r.PathLengthLimit = limit
    r.set_frame_index(0)
    r.reset()

    def rays_now():
        o = np.stack([r.state("OriginX"), r.state("OriginY"), r.state("OriginZ")], -1)
        d = unpack_unit_vector(r.state("PackedVelocity"))
        pix = r.state("ImagePosition")
        keep = np.isfinite(o).all(axis=1) & (pix >= 0)            # slots beyond the film hold no ray
        rays = np.zeros(int(keep.sum()), dtype=I.RAY_DTYPE)
        rays["o"], rays["d"], rays["duration"] = o[keep], d[keep], 1048576.0
        return rays

    primary = rays_now()
    r.run(2)
    r.run_frames(args.frames)
    ctx.sync()
    steady = rays_now()
    n = len(steady)
    lo, hi = steady["o"].min(axis=0), steady["o"].max(axis=0)
    octant = ((steady["d"][:, 0] < 0).astype(np.uint64) | ((steady["d"][:, 1] < 0).astype(np.uint64) << 1) | ((steady["d"][:, 2] < 0).astype(np.uint64) << 2))
    mort = morton30(steady["o"], lo, hi)
    rng = np.random.default_rng(0)
    orders = {
        "slot": np.arange(n),
        "random": rng.permutation(n),
        "octant": np.argsort(octant, kind="stable"),
        "oct+mort": np.argsort((octant << np.uint64(30)) | mort, kind="stable"),
        "morton": np.argsort((mort << np.uint64(3)) | octant, kind="stable"),
    }
    d_hits = torch.empty(max(n, len(primary)) * I.HIT_DTYPE.itemsize, dtype=torch.uint8, device="cuda:0")
    res = {"workload": desc, "rays": n, "reps": args.reps, "orders": {}}
    ref_hits = None

    def timed(rays_np, label):
        d_rays = torch.from_numpy(rays_np.view(np.uint8).reshape(-1).copy()).to("cuda:0")
        torch.cuda.synchronize()
        for _ in range(3):
            dev.raycast_device(d_rays.data_ptr(), len(rays_np), d_hits.data_ptr())
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            dev.raycast_device(d_rays.data_ptr(), len(rays_np), d_hits.data_ptr())
        ctx.sync()
        ms = (time.perf_counter() - t0) * 1e3 / args.reps
        hits = d_hits[: len(rays_np) * I.HIT_DTYPE.itemsize].cpu().numpy().view(I.HIT_DTYPE)
        res["orders"][label] = {"ms": ms, "mrays_per_s": len(rays_np) / ms / 1e3, "hit_fraction": float((hits["shape"] != 0xFFFFFFFF).mean())}
        print(label, res["orders"][label], flush=True)
        return hits

    for label, order in orders.items():
        hits = timed(steady[order], label)
        inv = np.empty(n, dtype=np.int64)
        inv[order] = np.arange(n)
        back = hits[inv]
        if ref_hits is None:
            ref_hits = back.copy()
        else:
            assert back.tobytes() == ref_hits.tobytes(), "hits depend on the ray order?"
    timed(primary, "primary")
    base = res["orders"]["slot"]["ms"]
    for k, v in res["orders"].items():
        v["vs_slot_order"] = v["ms"] / base if k != "primary" else None
    print(json.dumps(res))
    if args.out:
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
