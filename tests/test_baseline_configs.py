"""BASELINE.json configs C1-C5 AT THEIR STATED SIZES, rendered by the SHIPPED configuration (BVH8, the default
k_trace_coop variant, ScatterMode AUTO = material queues where materials mix, canonical ties) and compared with the
oracle on a bounded sample of the film: every k-th 16x16 tile (tiles are independent: a slot owns one pixel and carries
its own RNG stream, so the oracle's film on the sampled tiles IS the full-size film there).

Per config: hit IDs of >= 100k random rays bit-exact against the oracle's walk of the reference BVH2 (mismatches
adjudicated by the oracle's brute-force loop; the rate is recorded), alpha equal and radiance within 1e-4 after
reset + RunBasicRenderer(2) (primary + first bounce, application.cpp:104-110), and the same after further frames up to
the bounded fraction of paths that an ulp of libdevice-vs-glibc transcendentals sends another way.
Workload definitions are bench.py's (the ones the metric is quoted on)."""
import json
import os

import numpy as np
import pytest

from util import bits, ids_equal, random_rays

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

#       workload key: (tile stride - coprime with the tile columns so the sample spreads over the image, steady-state frames)
SAMPLE = {"C1": (1, 4), "C2": (7, 3), "C3": (7, 3), "C4": (29, 3), "C5": (29, 3)}


@pytest.fixture(scope="module")
def I():
    from path_tracer_b200 import integrator
    return integrator


def sampled_mask(w, h, stride):
    tiles_x = (w + 15) // 16
    ys, xs = np.mgrid[0:h, 0:w]
    return (((ys // 16) * tiles_x + xs // 16) % stride) == 0


def radiance_outliers(a, b, tol=1e-4):
    err = np.abs(a - b) / np.maximum(np.abs(b), 1.0)
    return err.max(axis=-1) > tol


@pytest.mark.parametrize("key", ["C1", "C2", "C3", "C4", "C5"])
def test_shipped_configuration_matches_oracle_at_full_size(I, gpu, oracle, key):
    import bench
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    desc, name, params, w, h, tone, limit, _ = bench.WORKLOADS[key]
    stride, frames = SAMPLE[key]
    sc = S.Scene.make(name, *params)
    d = I.DeviceScene(gpu, sc)                              # default builder (host SAH + BVH8 collapse)
    info = d.wide_info()
    assert info["status"] == 0 and info["builder"] == I.BUILDER_SAH_HOST

    # ---- ray casts: IDs bit-exact, t bit-exact (north_star asks 1e-5 relative) ----
    n = 120000
    rays = random_rays(sc, n, 1000 + ord(key[1]), oracle)
    out = d.raycast(rays)                                   # TRAVERSAL_WIDE, TIE_CANONICAL, default kernel variant
    ref = oracle.raycast(sc.view, rays, tie=oracle.TIE_CANONICAL)
    bad = np.nonzero(~(ids_equal(out, ref) & (bits(out["t"]) == bits(ref["t"]))))[0]
    assert len(bad) <= n * 1e-4, "wide vs reference traversal: %d of %d rays differ" % (len(bad), n)
    sample = np.concatenate([bad, np.arange(0, n, 1500)]).astype(np.int64)
    bf = oracle.raycast_bruteforce(sc.view, rays[sample])
    assert ids_equal(out[sample], bf).all() and (bits(out["t"][sample]) == bits(bf["t"])).all()
    ref_vs_bf = 0
    if len(bad):
        ref_vs_bf = int((~(ids_equal(ref[bad], bf[:len(bad)]) & (bits(ref["t"][bad]) == bits(bf["t"][:len(bad)])))).sum())

    # ---- film: shipped renderer vs oracle on every `stride`-th tile ----
    film = I.SampleBuffer(gpu, w, h)
    r = I.BasicRenderer(gpu, d, film)                       # defaults: WAVEFRONT, WIDE, CANONICAL, TRACE_VARIANT_DEFAULT, SCATTER_AUTO
    assert (r.Integrator, r.Traversal, r.TieMode, r.TraceVariant, r.ScatterMode) == \
        (I.INTEGRATOR_WAVEFRONT, I.TRAVERSAL_WIDE, I.TIE_CANONICAL, I.TRACE_VARIANT_DEFAULT, I.SCATTER_AUTO)
    r.PathLengthLimit = limit
    o = oracle.Renderer(sc.view, w, h, tie=oracle.TIE_CANONICAL, tile_stride=stride, path_length_limit=limit)
    m = sampled_mask(w, h, stride)
    assert m.sum() >= 250000
    for side in (o, r):
        side.set_frame_index(0)
        side.reset()
        side.run(2)
    a, b = film.download(), o.accum()
    assert (b[~m] == 0).all()                               # the oracle touched the sampled tiles only
    first_alpha = float(np.mean(a[m][:, 3] != b[m][:, 3]))
    first_out = radiance_outliers(a[m][:, :3], b[m][:, :3])
    assert first_alpha <= 1e-4, first_alpha
    assert first_out.mean() <= 2e-3, first_out.mean()
    r.run_frames(frames)
    o.run_frames(frames)
    a, b = film.download(), o.accum()
    later_alpha = float(np.mean(a[m][:, 3] != b[m][:, 3]))
    later_out = radiance_outliers(a[m][:, :3], b[m][:, :3])
    assert later_alpha <= 1e-2, later_alpha
    assert later_out.mean() <= 2e-2, later_out.mean()
    # bulk statistics of the sampled region agree far inside the 1 % converged-image bound
    sa, sb = a[m][:, :3].sum(axis=0, dtype=np.float64), b[m][:, :3].sum(axis=0, dtype=np.float64)
    assert (np.abs(sa - sb) <= 2e-3 * np.abs(sb)).all(), (sa, sb)
    st = r.stats()
    assert st["rays_traced"] == (2 + frames) * w * h and int(a[..., 3].sum(dtype=np.float64)) == st["paths_finished"]
    lin, srgb = film.resolve(1.0, I.TONE_HABLE if tone == "hable" else I.TONE_ACES, 1.0)
    assert np.isfinite(lin).all()

    rec = {"workload": desc, "rays": n, "wide_vs_reference_traversal_mismatches": int(len(bad)), "of_which_reference_traversal_wrong": ref_vs_bf,
           "mismatch_rate": len(bad) / n, "sampled_pixels": int(m.sum()), "tile_stride": stride,
           "first_bounce_alpha_mismatch": first_alpha, "first_bounce_radiance_outliers": float(first_out.mean()),
           "frames": frames, "later_alpha_mismatch": later_alpha, "later_radiance_outliers": float(later_out.mean()),
           "bvh8_nodes": info["node_count"], "bvh8_prims": info["prim_count"]}
    print("PARITY " + key + " " + json.dumps(rec))
    out_dir = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out_dir, exist_ok=True)
        with open(os.path.join(out_dir, "parity_%s.json" % key), "w") as f:
            json.dump(rec, f, indent=1)
    except OSError:
        pass


def test_c1_converged_image_as_stated(I, gpu, oracle):
    """BASELINE config 1 as written - Cornell box, 640x480, 64 spp - rendered by the shipped configuration until every pixel
    holds 64 samples on average, against the oracle at the same frame count (same seeds): north_star's third level, mean
    per-channel relative error <= 1 % (measured ~1e-4: the films differ only where a libdevice / glibc ulp sends a path
    another way)."""
    import bench
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    desc, name, params, w, h, tone, limit, _ = bench.WORKLOADS["C1"]
    sc = S.Scene.make(name, *params)
    d = I.DeviceScene(gpu, sc)
    film = I.SampleBuffer(gpu, w, h)
    r = I.BasicRenderer(gpu, d, film)
    r.set_frame_index(0)
    r.reset()
    r.run(2)
    frames = 0
    while r.stats()["paths_finished"] < 64 * w * h:
        r.run_frames(32)
        frames += 32
        assert frames < 5000
    stride = 2
    o = oracle.Renderer(sc.view, w, h, tie=oracle.TIE_CANONICAL, tile_stride=stride)
    o.set_frame_index(0)
    o.reset()
    o.run(2)
    o.run_frames(frames)
    m = sampled_mask(w, h, stride)
    a, b = film.download()[m], o.accum()[m]
    assert a[:, 3].mean() >= 60
    ia, ib = a[:, :3] / np.maximum(a[:, 3:], 1), b[:, :3] / np.maximum(b[:, 3:], 1)
    rel = np.abs(ia.mean(axis=0) - ib.mean(axis=0)) / ib.mean(axis=0)
    assert (rel <= 0.01).all(), rel
    assert np.abs(ia - ib).mean() / ib.mean() <= 0.01
    assert abs(a[:, 3].sum() - b[:, 3].sum()) / b[:, 3].sum() < 1e-3
    print("C1 64 spp: frames", frames, "mean-channel rel err", rel, "per-pixel", float(np.abs(ia - ib).mean() / ib.mean()))
