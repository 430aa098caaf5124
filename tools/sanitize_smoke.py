#!/usr/bin/env python
"""Small end-to-end run for compute-sanitizer (memcheck / racecheck / synccheck): every kernel of the
default path on small films — cooperative trace, direct and queued scatter, restart, resolve, ray cast, preview."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from path_tracer_b200 import integrator as I          # noqa: E402
from path_tracer_b200 import scene as S               # noqa: E402

S.ensure_spectrum_table()
ctx = I.Context(0)
for name, params in (("cornell", ()), ("terrain", (24,)), ("spherefield", (6, 1, 2))):
    sc = S.Scene.make(name, *params)
    dev = I.DeviceScene(ctx, sc)
    for mode in (I.SCATTER_DIRECT, I.SCATTER_QUEUES, I.SCATTER_DIRECT_STAGED):
        film = I.SampleBuffer(ctx, 80, 48)
        r = I.BasicRenderer(ctx, dev, film)
        r.ScatterMode = mode
        r.reset()
        r.run(2)
        r.run_frames(3)
        a = film.download()
        assert np.isfinite(a).all()
        film.resolve(1.0, I.TONE_ACES, 1.0)
    rng = np.random.default_rng(1)
    rays = np.zeros(2048, dtype=I.RAY_DTYPE)
    lo, hi = sc.bounds()
    rays["o"] = rng.uniform(np.clip(lo, -6, 6), np.clip(hi, -6, 6), size=(2048, 3))
    d = rng.normal(size=(2048, 3))
    rays["d"] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays["duration"] = 1048576.0
    dev.raycast(rays)
    dev.raycast_noalloc(rays)
    cam = sc.cameras[0]
    dev.preview(cam["To"].ravel(), cam["From"].ravel(), 1, 64, 40)
print("sanitize_smoke ok")
