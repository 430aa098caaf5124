#!/usr/bin/env python
"""Debug aid for tools/fuzz_scatter.py --gpu: one seed, round by round, first slot whose state leaves the oracle's."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import fuzz_scatter as FS
from oracle import binding as oracle
from path_tracer_b200 import scene as S
from path_tracer_b200 import integrator as I
oracle.lib(); S.ensure_spectrum_table()
seed = int(sys.argv[1]); w, h = 48, 32
rng = np.random.default_rng(seed)
sc, n_cams = FS.random_scene(S, rng)
cam = int(rng.integers(0, n_cams)); flags = int(rng.integers(0, 4)); limit = int(rng.choice([0, 0, 1, 2, 5])); term = float(rng.choice([0.0, 0.0, 0.1, 0.5])); frame = int(rng.integers(0, 1000))
print("cam", cam, "flags", flags, "limit", limit, "term", term, "frame", frame, "shape types", sc.shapes["Type"], "materials of shapes", sc.shapes["MaterialIndex"])
md = np.asarray(sc.materials).view(np.uint32).reshape(-1)
print("material types by slot", {int(m): int(md[32 * int(m)]) for m in set(sc.shapes["MaterialIndex"])})
o = oracle.Renderer(sc.view, w, h, camera_index=cam, render_flags=flags, path_length_limit=limit, path_termination_probability=term, tie=oracle.TIE_CANONICAL)
ctx = I.Context(0); dev = I.DeviceScene(ctx, sc); film = I.SampleBuffer(ctx, w, h); r = I.BasicRenderer(ctx, dev, film)
r.CameraIndex, r.RenderFlags, r.PathLengthLimit, r.PathTerminationProbability = cam, flags, limit, term
r.TieMode, r.Traversal, r.Integrator = I.TIE_CANONICAL, I.TRAVERSAL_WIDE, I.INTEGRATOR_WAVEFRONT
for x in (o, r):
    x.set_frame_index(frame); x.reset()
names = ["OriginX", "OriginY", "OriginZ", "PackedVelocity", "Time", "ShapeAndMaterialIndex", "PackedNormal", "PackedTangentX", "TextureU", "TextureV", "ImagePosition", "NormalizedLambda0",
         "Throughput0", "Throughput1", "Throughput2", "Throughput3", "Probability0", "Probability1", "Probability2", "Probability3", "SampleR", "SampleG", "SampleB", "ActiveShapeIndex01", "ActiveShapeIndex23"]
prev = None
for step in range(8):
    for x in (o, r):
        x.run(1)
    r.sync()
    so = {n: o.state(n) for n in names}; sg = {n: r.state(n) for n in names}
    a, b = film.download(), o.accum()
    dalpha = np.flatnonzero((a[..., 3] != b[..., 3]).reshape(-1))
    diff = {n: int((so[n].view(np.uint32) != sg[n].view(np.uint32)).sum()) for n in names}
    print("round", step + 1, "alpha differs in", len(dalpha), "pixels; state words differing:", {k: v for k, v in diff.items() if v})
    bad = np.flatnonzero((so["ShapeAndMaterialIndex"] != sg["ShapeAndMaterialIndex"]) | (so["ImagePosition"] != sg["ImagePosition"]) | (so["ActiveShapeIndex01"] != sg["ActiveShapeIndex01"]))
    if len(bad):
        for s_ in bad[:3]:
            print(" slot", int(s_))
            if prev is not None:
                print("   before (oracle):", {n: prev[0][n][s_] for n in names})
                print("   before (kernels):", {n: prev[1][n][s_] for n in names})
            print("   after  (oracle):", {n: so[n][s_] for n in names})
            print("   after  (kernels):", {n: sg[n][s_] for n in names})
        break
    prev = (so, sg)
