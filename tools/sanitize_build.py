"""Small run of the GPU builders and the spectrum kernels for compute-sanitizer --tool memcheck."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from path_tracer_b200 import integrator as I      # noqa: E402
from path_tracer_b200 import scene as S           # noqa: E402

ctx = I.Context(0)
S.ensure_spectrum_table()
rng = np.random.default_rng(0)
for name, params in (("cornell", ()), ("terrain", (48,)), ("spherefield", (9, 2, 2)), ("room", (1, 256))):
    sc = S.Scene.make(name, *params)
    ref = None
    for b in (I.BUILDER_SAH_HOST, I.BUILDER_LBVH_GPU, I.BUILDER_PLOC_GPU):
        d = I.DeviceScene(ctx, sc, builder=b)
        dirs = rng.normal(size=(20000, 3))
        dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
        rays = np.zeros(20000, dtype=I.RAY_DTYPE)
        rays["o"] = rng.uniform(-4, 4, size=(20000, 3))
        rays["d"] = dirs
        rays["duration"] = 1e30
        rng = np.random.default_rng(0)
        out = d.raycast(rays)
        if ref is None:
            ref = out
        assert out.tobytes() == ref.tobytes(), (name, b)
        film = I.SampleBuffer(ctx, 64, 32)
        r = I.BasicRenderer(ctx, d, film)
        r.reset()
        r.run(2)
        r.run_frames(3)
        r.sync()
    print(name, "ok", d.wide_info()["node_count"])
table, ms = I.build_spectrum_table(ctx)
out = I.upsample_texels(ctx, table, rng.uniform(0, 1, size=(5000, 4)).astype(np.float32), 2)
print("spectrum ok", ms, out.shape)
