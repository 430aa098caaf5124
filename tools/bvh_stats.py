#!/usr/bin/env python
"""Builder tuning aid (CPU only): node steps / primitive tests per ray of the wide BVH on the
in-flight rays of the C2 workload, via tests/host_emu (g++ build of the device functions)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import binding as O          # noqa: E402
from path_tracer_b200 import scene as S  # noqa: E402


def build_emu():
    d = os.path.join(ROOT, "tests", "host_emu")
    csrc = os.path.join(ROOT, "path_tracer_b200", "csrc")
    out = os.path.join(d, "_build", "libemu.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-I", csrc, "-I", os.path.join(ROOT, "include"),
                           "-I", "/usr/local/cuda/include", "-o", out, os.path.join(d, "emu.cpp"), os.path.join(csrc, "ptb_wide_build.cpp")] + sys.argv[2:])
    L = C.CDLL(out)
    L.emu_create.restype = C.c_void_p
    L.emu_create.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
    L.emu_wide_info.argtypes = [C.c_void_p, C.c_void_p]
    L.emu_wide_stats.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    return L


def inflight_rays(sc, w, h, rounds):
    r = O.Renderer(sc.view, w, h, tie=O.TIE_CANONICAL)
    r.set_frame_index(0)
    r.reset()
    r.run(2)
    r.run_frames(rounds)
    o = np.stack([r.state("OriginX"), r.state("OriginY"), r.state("OriginZ")], 1)
    pv = r.state("PackedVelocity")
    x = ((pv & 0xFFFF).astype(np.int16)).astype(np.float32) / 32767.0
    y = ((pv >> 16).astype(np.int16)).astype(np.float32) / 32767.0
    z = 1 - np.abs(x) - np.abs(y)
    neg = z < 0
    xx = np.where(neg, (1 - np.abs(y)) * np.where(x >= 0, 1, -1), x)
    yy = np.where(neg, (1 - np.abs(x)) * np.where(y >= 0, 1, -1), y)
    d = np.stack([xx, yy, z], 1)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return O.make_rays(o.astype(np.float32), d.astype(np.float32))


if __name__ == "__main__":
    grid = int(sys.argv[1]) if len(sys.argv) > 1 else 708
    L = build_emu()
    sc = S.Scene.make("terrain", grid)
    rays = inflight_rays(sc, 480, 270, 6)
    e = L.emu_create(C.addressof(sc.view), 16, 16)
    info = (C.c_uint32 * 5)()
    L.emu_wide_info(e, info)
    out = (C.c_uint64 * 3)()
    L.emu_wide_stats(e, rays.ctypes.data, len(rays), out)
    print("nodes %d prims %d depth %d | per ray: node steps %.2f prim tests %.2f max stack %d" % (info[0], info[1], info[2], out[0] / len(rays), out[1] / len(rays), out[2]))
