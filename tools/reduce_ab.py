#!/usr/bin/env python
"""A/B of the three exchange + resolve variants of ptb_film_reduce_resolve (include/ptb.h) at the two film sizes that matter:
1920x1080 (33 MB of float film per rank, C2) and 3840x2160 (133 MB, C5).  Run under torchrun on N GPUs of one box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29555 tools/reduce_ab.py

Per mode: device time of the exchange + resolve step (ptb_group_last_reduce, max over ranks, median of the repetitions) and the
wall time of the whole call on the root including the sRGB8 frame's copy to pinned host memory."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from path_tracer_b200 import dist as D
    from path_tracer_b200 import integrator as I
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    rank, world = D.init_process_group()
    ctx = I.Context(local)
    group = D.create_group(ctx, device)
    assert group is not None, "run under torchrun with at least 2 ranks"
    out = {"n_gpus": world, "nccl": I.nccl_version() if hasattr(I, "nccl_version") else None, "sizes": {}}
    reps = 15
    for w, h in ((1920, 1080), (3840, 2160)):
        rng = np.random.default_rng(rank)
        film = I.SampleBuffer(ctx, w, h)
        a = np.zeros((h, w, 4), np.float32)
        a[..., 3] = 64 + rank
        a[..., :3] = rng.uniform(0, 2, size=(h, w, 3)).astype(np.float32) * a[..., 3:]
        film.upload(a)
        res = {}
        for mode, name in ((I.REDUCE_NCCL, "nccl_reduce_then_resolve_on_root"), (I.REDUCE_NCCL_SLICED, "reduce_scatter_resolve_slices_allgather_srgb8"),
                           (I.REDUCE_P2P_FUSED, "fused_peer_memory_reduce_resolve")):
            dev_ms, wall_ms = [], []
            for k in range(reps + 3):
                group.barrier()
                t0 = time.perf_counter()
                frame = group.reduce_resolve(film, (1.0, I.TONE_ACES, 1.0), mode=mode, root=0)
                t1 = time.perf_counter()
                ms = D.max_over_ranks(group.last_reduce()["exchange_ms"], device)
                if k >= 3:
                    dev_ms.append(ms)
                    wall_ms.append((t1 - t0) * 1e3)
            res[name] = {"exchange_resolve_ms_device": float(np.median(dev_ms)), "call_ms_wall_on_this_rank": float(np.median(wall_ms)),
                         "checksum": int(frame.astype(np.uint64).sum()) if rank == 0 else None}
        out["sizes"]["%dx%d" % (w, h)] = {"film_bytes_per_rank": w * h * 16, "modes": res}
        film.close()
    group.close()
    D.shutdown()
    if rank == 0:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
