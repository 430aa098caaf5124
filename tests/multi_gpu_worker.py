"""Worker of tests/test_multi_gpu.py: one rank of a sample-split render (own process, own GPU)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FRAMES = 6
W, H = 192, 112              # W*H divisible by 2, 4 and 8


def render_rank(I, ctx, dev, rank, world, frames=FRAMES, w=W, h=H):
    film = I.SampleBuffer(ctx, w, h)
    r = I.BasicRenderer(ctx, dev, film)
    r.set_frame_index(I.Group.first_frame(rank, world, frames))
    r.reset()
    r.run_frames(frames)
    r.sync()
    return film, r


def rank_main(rank, world, gid, scene_spec, out_q):
    try:
        from path_tracer_b200 import integrator as I
        from path_tracer_b200 import scene as S
        ctx = I.Context(rank)
        g = I.Group(ctx, gid, rank, world)
        name, params = scene_spec
        if rank == 0:
            S.ensure_spectrum_table()
            sc = S.Scene.make(name, *params)
            dev = I.DeviceScene(ctx, sc)
        else:
            dev = I.DeviceScene(ctx)                      # never packed, never built here
        g.replicate_scene(dev, 0)
        info = dev.wide_info()
        rng = np.random.default_rng(3)
        d = rng.normal(size=(20000, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        rays = np.zeros(20000, dtype=I.RAY_DTYPE)
        rays["o"] = rng.uniform(-3, 3, size=(20000, 3)).astype(np.float32)
        rays["d"] = d.astype(np.float32)
        rays["duration"] = 1048576.0
        hits = dev.raycast(rays)
        film, r = render_rank(I, ctx, dev, rank, world)
        own = film.download()
        st = r.stats()
        res = {"rank": rank, "own": own, "paths": st["paths_finished"], "rays": st["rays_traced"], "hits": hits.tobytes(),
               "nodes": info["node_count"], "prims": info["prim_count"], "first_frame": I.Group.first_frame(rank, world, FRAMES)}
        params_r = (1.0, I.TONE_ACES, 1.0)
        # 1. plain sum into a second film on the root
        total = I.SampleBuffer(ctx, W, H) if rank == 0 else None
        g.reduce(film, 0, total)
        if rank == 0:
            res["sum_nccl"] = total.download()
        # 2. the three fused exchange + resolve variants
        for mode, key in ((I.REDUCE_NCCL, "nccl"), (I.REDUCE_NCCL_SLICED, "sliced"), (I.REDUCE_P2P_FUSED, "p2p")):
            into = I.SampleBuffer(ctx, W, H) if (rank == 0 and mode != I.REDUCE_NCCL_SLICED) else None
            frame = g.reduce_resolve(film, params_r, mode=mode, root=0, into=into)
            if rank == 0:
                res["srgb_" + key] = frame.copy()
                if into is not None:
                    res["sum_" + key] = into.download()
            res["ms_" + key] = g.last_reduce()["exchange_ms"]
        # the source films must be untouched by all of that
        res["own_after"] = film.download()
        # 3. all-reduce in place: every rank ends with the total
        g.reduce(film, -1, film)
        res["allreduce"] = film.download()
        if rank == 0:
            lin, srgb = total.resolve(*params_r)
            res["srgb_root_resolve"] = srgb
        g.barrier()
        out_q.put(res)
        g.close()
    except Exception as e:      # noqa: BLE001 - the parent must hear about it instead of waiting for the timeout
        import traceback
        out_q.put({"rank": rank, "error": "%s\n%s" % (e, traceback.format_exc())})
