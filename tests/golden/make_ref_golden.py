"""Generates tests/golden/ref_*.npz from the REFERENCE'S OWN SHADER SOURCES.

oracle/ref/ compiles the reference's basic_trace.glsl / basic_scatter.glsl (and the .glsl.inc files they include)
as C++ where they lie under /root/reference, into oracle/_ref/libref.so.  This script runs that library - Trace()
on seeded random rays, and ResetBasicRenderer + RunBasicRenderer on small films - and stores inputs and outputs.
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


if __name__ == "__main__":
    S.ensure_spectrum_table()
    R.build(force=True)
    for tag, (name, params, camera, flags, term) in REF_CASES.items():
        fixture(tag, name, params, 2048, 21, 64, 48, camera, flags, term, 6)
