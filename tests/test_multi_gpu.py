"""Multi-GPU path on real hardware (needs >= 2 B200s: run with `gpurun --gpus 2`; skipped on a 1-GPU box).

Checks what tests/test_dist_gloo.py cannot: that the NCCL communicators inside libptb.so (include/ptb.h "groups")
sum real per-rank films exactly, that the scene replicated device-to-device traces like the original, that rank seed
blocks are disjoint AS RENDERED, and that the three exchange + resolve variants deliver the same frame."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def I():
    from path_tracer_b200 import integrator
    return integrator


def _device_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _run_ranks(world, scene_spec):
    import multiprocessing as mp
    from path_tracer_b200 import integrator as I
    import multi_gpu_worker as Wk
    ctx = mp.get_context("spawn")
    gid = I.Group.unique_id()
    q = ctx.Queue()
    ps = [ctx.Process(target=Wk.rank_main, args=(r, world, gid, scene_spec, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = {}
    try:
        for _ in ps:
            r = q.get(timeout=600)
            assert "error" not in r, r["error"]
            res[r["rank"]] = r
    finally:
        for p in ps:
            p.join(timeout=60)
            if p.is_alive():
                p.kill()
    return res


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("scene_spec", [("cornell", ()), ("room", (1, 512))])
def test_sample_split_films_sum_exactly(I, gpu, scenes, world, scene_spec):
    if _device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    if world > 2 and scene_spec[0] != "cornell":
        pytest.skip("one scene is enough at this size")
    import multi_gpu_worker as Wk
    res = _run_ranks(world, scene_spec)
    own = [res[r]["own"] for r in range(world)]
    # Sigma alpha == Sigma paths_finished, per rank and in total
    for r in range(world):
        assert int(own[r][..., 3].sum(dtype=np.float64)) == res[r]["paths"] > 0
        assert res[r]["rays"] == Wk.FRAMES * Wk.W * Wk.H
        assert np.array_equal(res[r]["own_after"], own[r])
    seq = own[0].copy()
    for r in range(1, world):
        seq = seq + own[r]                                    # rank order, float32: what the peer-memory kernel computes
    total = np.sum(np.stack(own).astype(np.float64), axis=0)
    assert int(seq[..., 3].sum(dtype=np.float64)) == sum(res[r]["paths"] for r in range(world))
    root = res[0]
    assert np.array_equal(root["sum_p2p"], seq)               # bit for bit, any world size
    for key in ("sum_nccl", "sum_nccl"):
        if world == 2:
            assert np.array_equal(root[key], seq)             # two addends: addition is commutative, so NCCL's order cannot matter
        assert np.array_equal(root[key][..., 3], seq[..., 3])  # counts are small integers: exact in any order
        assert np.allclose(root[key], total, rtol=1e-6, atol=1e-6)
    for r in range(world):
        assert np.array_equal(res[r]["allreduce"][..., 3], seq[..., 3])
        assert np.allclose(res[r]["allreduce"], total, rtol=1e-6, atol=1e-6)
    # the three variants deliver the same frame (up to one sRGB8 step where the float sums differ in the last bit)
    ref = root["srgb_root_resolve"].astype(int)
    for key in ("srgb_nccl", "srgb_sliced", "srgb_p2p"):
        assert np.abs(root[key].astype(int) - ref).max() <= 1, key
        assert (root[key][..., 3] == 255).all()
    # replicated scene == original scene: same tree, same hits
    for r in range(1, world):
        assert res[r]["nodes"] == root["nodes"] and res[r]["prims"] == root["prims"]
        assert res[r]["hits"] == root["hits"]
    # seed blocks are disjoint as rendered: rank r's film is what ONE GPU renders from that rank's first frame index
    d = I.DeviceScene(gpu, scenes(*((scene_spec[0],) + scene_spec[1])))
    firsts = [res[r]["first_frame"] for r in range(world)]
    assert len(set(firsts)) == world and all(b - a == Wk.FRAMES + 1 for a, b in zip(firsts, firsts[1:]))
    for r in range(world):
        film = I.SampleBuffer(gpu, Wk.W, Wk.H)
        rr = I.BasicRenderer(gpu, d, film)
        rr.set_frame_index(firsts[r])
        rr.reset()
        rr.run_frames(Wk.FRAMES)
        assert np.array_equal(film.download(), own[r]), r
    for a in range(world):
        for b in range(a + 1, world):
            assert not np.array_equal(own[a], own[b])


def test_local_group_one_process_two_devices(I, gpu, scenes):
    """ptb_group_create_local: one process, one host thread per device (what bin/ptb_headless --devices does)."""
    if _device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import ctypes as C
    import threading
    import multi_gpu_worker as Wk
    L = I.lib()
    ctxs = [gpu, I.Context(1)]
    arr = (C.c_void_p * 2)(ctxs[0].h, ctxs[1].h)
    out = (C.c_void_p * 2)()
    assert L.ptb_group_create_local(arr, 2, out) == 0, L.ptb_last_error()
    groups = []
    for k in range(2):
        g = I.Group.__new__(I.Group)
        g.ctx, g.h, g.rank, g.world = ctxs[k], C.c_void_p(out[k]), k, 2
        groups.append(g)
    sc = scenes("cornell")
    res, errs = [None, None], []

    def work(k):
        try:
            dev = I.DeviceScene(ctxs[k], sc) if k == 0 else I.DeviceScene(ctxs[k])
            groups[k].replicate_scene(dev, 0)
            film, r = Wk.render_rank(I, ctxs[k], dev, k, 2)
            own = film.download()
            into = I.SampleBuffer(ctxs[k], Wk.W, Wk.H) if k == 0 else None
            frame = groups[k].reduce_resolve(film, (1.0, I.TONE_ACES, 1.0), mode=I.REDUCE_P2P_FUSED, root=0, into=into)
            res[k] = (own, into.download() if into is not None else None, frame)
        except Exception as e:      # noqa: BLE001
            errs.append(e)
    ts = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=300)
    assert not errs, errs
    assert np.array_equal(res[0][1], res[0][0] + res[1][0])
    assert res[0][2] is not None and res[0][2][..., :3].std() > 5
    for g in groups:
        g.close()
