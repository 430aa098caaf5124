"""Generates tests/golden/ref_*.npz from the REFERENCE'S OWN SHADER SOURCES.

oracle/ref/ compiles the reference's basic_trace.glsl / basic_scatter.glsl (and the .glsl.inc files they include)
as C++ where they lie under /root/reference, into oracle/_ref/libref.so.  This script runs that library - Trace()
on seeded random rays, ResetBasicRenderer + RunBasicRenderer on small films, the resolve and preview fragment
shaders - and stores inputs and outputs.
The reference cannot travel to the GPU box and ships no vectors of its own, so these files are
what pins the oracle (tests/test_reference_shaders.py) and the CUDA path (tests/test_gpu_parity.py) to it there.

    python tests/golden/make_ref_golden.py          # needs /root/reference
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import binding as O                       # noqa: E402  (ray construction only)
from oracle.ref import binding as R                   # noqa: E402
from path_tracer_b200 import scene as S               # noqa: E402
from util import random_rays, make_test_scene, REF_CASES   # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
HIT_FIELDS = ("t", "shape", "prim", "material", "shape_type", "n", "tan", "uv", "bary", "scene_complexity", "mesh_complexity")
FILM_ARRAYS = ("ImagePosition", "NormalizedLambda0", "OriginX", "OriginY", "OriginZ", "PackedVelocity", "ShapeAndMaterialIndex")


def fixture(tag, name, params, n_rays, seed, w, h, camera, flags, term, frames):
    sc = make_test_scene(S, name, params)
    rays = random_rays(sc, n_rays, seed, O)
    hits = R.raycast(sc.view, rays)
    miss = hits["shape"] == 0xFFFFFFFF
    for f in ("prim", "material", "shape_type", "n", "tan", "uv", "bary"):     # the shader leaves these unwritten on a miss
        hits[f][miss] = 0
    out = {"rays": rays.view(np.float32).reshape(n_rays, 7), "size": np.array([w, h]), "frames": np.array([frames]),
           "render": np.array([camera, flags], dtype=np.int64), "termination": np.array([term], dtype=np.float32)}
    for f in HIT_FIELDS:
        out["hit_" + f] = hits[f]
    r = R.Renderer(sc.view, w, h, camera_index=camera, render_flags=flags, path_termination_probability=term)
    r.set_frame_index(0)
    r.reset()
    r.run(2)
    out["accum_first"] = r.accum()
    for a in FILM_ARRAYS:                                  # the wavefront state after the first two rounds
        out["state_" + a] = r.state(a)
    r.run_frames(frames)
    out["accum_frames"] = r.accum()
    np.savez_compressed(os.path.join(HERE, "ref_" + tag + ".npz"), **out)
    print("ref_" + tag, "hits", int((hits["shape"] != 0xFFFFFFFF).sum()), "of", n_rays, "paths", int(out["accum_frames"][..., 3].sum()),
          "mean", float(out["accum_frames"][..., :3].sum() / max(1.0, out["accum_frames"][..., 3].sum())))


def resolve_film():
    """A film with the edge cases of resolve.glsl:112-131: empty pixels, zero radiance, very bright, negative."""
    rng = np.random.default_rng(1)
    film = np.zeros((40, 56, 4), np.float32)
    film[..., 3] = rng.integers(0, 5, size=(40, 56))
    film[..., :3] = rng.gamma(1.0, 2.0, size=(40, 56, 3)) * film[..., 3:4]
    film[0, 0], film[0, 1], film[0, 2], film[0, 3] = [1, 2, 3, 0], [0, 0, 0, 4], [1e6, 1e6, 1e6, 1], [-0.5, 0.2, 0.1, 2]
    return film


RESOLVE_SETTINGS = [(1.0, 0, 1.0), (2.5, 0, 1.0), (1.0, 1, 1.0), (2.5, 1, 4.0), (1.0, 2, 1.0), (0.1, 2, 1.0), (1.0, 3, 1.0), (2.5, 3, 1.0)]
PREVIEW_CASES = ("cornell", "room", "mixed_lens")
PREVIEW_ARGS = dict(brightness=1.25, selected=2, mouse=(20, 17))


def resolve_and_preview_fixture():
    film = resolve_film()
    out = {"film": film, "settings": np.array(RESOLVE_SETTINGS, dtype=np.float64)}
    for k, (brightness, mode, white) in enumerate(RESOLVE_SETTINGS):
        out["linear_%d" % k] = R.resolve(film, brightness, mode, white)
    np.savez_compressed(os.path.join(HERE, "ref_resolve.npz"), **out)
    out = {"size": np.array([48, 32])}
    for tag in PREVIEW_CASES:
        name, params, *_ = REF_CASES[tag]
        sc = make_test_scene(S, name, params)
        cam = sc.cameras[0]
        for mode in range(7):
            img, pick = R.preview(sc.view, cam["To"].ravel(), cam["From"].ravel(), mode, 48, 32, **PREVIEW_ARGS)
            out["%s_mode%d" % (tag, mode)] = img
        out[tag + "_pick"] = np.array([pick], dtype=np.uint32)
    np.savez_compressed(os.path.join(HERE, "ref_preview.npz"), **out)
    print("ref_resolve, ref_preview written")


if __name__ == "__main__":
    S.ensure_spectrum_table()
    R.build(force=True)
    for tag, (name, params, camera, flags, term) in REF_CASES.items():
        fixture(tag, name, params, 2048, 21, 64, 48, camera, flags, term, 6)
    resolve_and_preview_fixture()
