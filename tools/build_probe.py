"""Scene flattening and spectral upsampling on the GPU, once each (profiled by tools/profile_round.sh)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from path_tracer_b200 import integrator as I      # noqa: E402
from path_tracer_b200 import scene as S           # noqa: E402

ctx = I.Context(0)
S.ensure_spectrum_table(ctx)
sc = S.Scene.make("terrain", 708)                 # C2 geometry: 1 002 528 triangles
for b in (I.BUILDER_LBVH_GPU, I.BUILDER_PLOC_GPU):
    d = I.DeviceScene(ctx, sc, builder=b)
    w = d.wide_info()
    print("builder", b, {k: w[k] for k in ("node_count", "depth", "bvh2_depth", "sah_cost", "build_seconds", "device_build_ms")})
table, ms = I.build_spectrum_table(ctx)
print("spectrum table: %.2f ms" % ms)
rgba = np.random.default_rng(0).uniform(0, 1, size=(1024 * 1024, 4)).astype(np.float32)
out = I.upsample_texels(ctx, table, rgba, 1)
print("upsampled", out.shape)
