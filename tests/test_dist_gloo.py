"""Sample-split plumbing on CPU: world_size 2 over gloo.  Each rank renders its own
frame block (with the oracle standing in for the GPU integrator — this test is about
the schedule and the reduce, not the kernels) and the films are sum-reduced."""
import os
import sys

import numpy as np
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, FRAMES = 32, 32, 3


def render_block(first_frame):
    from oracle import binding as O
    from path_tracer_b200 import scene as S
    sc = S.Scene.make("cornell")
    r = O.Renderer(sc.view, W, H, tie=O.TIE_FAITHFUL, threads=1)
    r.set_frame_index(first_frame)
    r.reset()
    r.run_frames(FRAMES)
    return r.accum()


def worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from path_tracer_b200 import dist as D
    r, w = D.init_process_group("gloo")
    assert (r, w) == (rank, world)
    first = D.frame_schedule(rank, world, FRAMES)
    film = torch.from_numpy(render_block(first))
    D.barrier()
    D.reduce_film(film, dst=0)
    total = D.sum_over_ranks(float(rank + 1), "cpu")
    mx = D.max_over_ranks(float(rank + 1), "cpu")
    assert total == 3.0 and mx == 2.0
    if rank == 0:
        np.save(os.path.join(out_dir, "film.npy"), film.numpy())
    import torch.distributed as dist
    dist.destroy_process_group()


def test_frame_schedule_is_disjoint_and_gap_free():
    from path_tracer_b200 import dist as D
    for world in (1, 2, 4, 8):
        seeds = []
        for r in range(world):
            f0 = D.frame_schedule(r, world, 5, base=7)
            seeds += list(range(f0, f0 + 6))          # reset seed + 5 run seeds
        assert sorted(seeds) == list(range(7, 7 + 6 * world))


def test_two_rank_film_reduce(tmp_path):
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    port = 29600 + (os.getpid() % 300)
    mp.spawn(worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    film = np.load(os.path.join(str(tmp_path), "film.npy"))
    from path_tracer_b200 import dist as D
    expect = render_block(D.frame_schedule(0, 2, FRAMES)) + render_block(D.frame_schedule(1, 2, FRAMES))
    assert np.array_equal(film, expect)
    assert film[..., 3].sum() > 0
