"""Generates the golden fixtures under tests/golden/ from the oracle.

These files pin the ORACLE's behaviour at the time of writing (canonical tie mode and
brute force included, which the reference does not have): tests/test_oracle_kats.py
checks the oracle still reproduces them, and the -m gpu tests check the CUDA path
against the same numbers.  Vectors written by the reference's own shader sources are
make_ref_golden.py's ref_*.npz.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import binding as O          # noqa: E402
from path_tracer_b200 import scene as S  # noqa: E402
from util import random_rays             # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def raycast_fixture(tag, name, params, n, seed):
    sc = S.Scene.make(name, *params)
    rays = random_rays(sc, n, seed, O)
    out = {"scene": np.array([name]), "scene_params": np.array(params, dtype=np.int64), "rays": rays.view(np.float32).reshape(n, 7)}
    for tie, t in ((O.TIE_FAITHFUL, "faithful"), (O.TIE_CANONICAL, "canonical")):
        h = O.raycast(sc.view, rays, tie=tie)
        out[t + "_shape"], out[t + "_prim"], out[t + "_t"] = h["shape"], h["prim"], h["t"]
    bf = O.raycast_bruteforce(sc.view, rays)
    out["brute_shape"], out["brute_prim"], out["brute_t"] = bf["shape"], bf["prim"], bf["t"]
    np.savez_compressed(os.path.join(HERE, tag + ".npz"), **out)
    print(tag, "hits", int((out["canonical_shape"] != 0xFFFFFFFF).sum()), "of", n)


def first_bounce_fixture():
    w, h = 64, 48
    sc = S.Scene.make("cornell")      # keep alive: the view borrows its arrays
    r = O.Renderer(sc.view, w, h, tie=O.TIE_FAITHFUL)
    r.set_frame_index(0)
    r.reset()
    r.run(2)
    np.savez_compressed(os.path.join(HERE, "cornell_first_bounce.npz"), size=np.array([w, h]), accum=r.accum())
    print("cornell_first_bounce", r.stats())


if __name__ == "__main__":
    S.ensure_spectrum_table()
    raycast_fixture("cornell_raycast", "cornell", [], 4096, 11)
    raycast_fixture("terrain32_raycast", "terrain", [32], 4096, 12)
    first_bounce_fixture()
