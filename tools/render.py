#!/usr/bin/env python
"""Headless stand-in for the reference application's Update() loop (src/application/application.cpp:100-124):
build or import a scene, run the integrator for a number of samples, resolve, write a PNG (and optionally a PFM).

    python tools/render.py --scene cornell --spp 256 --out cornell.png
    python tools/render.py --obj model.obj --spp 512 --size 1280 640 --tone aces --out model.png
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="cornell", help="procedural scene name (cornell, terrain, room, spherefield)")
    ap.add_argument("--scene-params", type=float, nargs="*", default=[])
    ap.add_argument("--obj", default=None, help="Wavefront OBJ to import instead (placed on the default checker plane)")
    ap.add_argument("--size", type=int, nargs=2, default=[1024, 512], help="the reference's cameras assume a 2:1 film")
    ap.add_argument("--spp", type=int, default=256, help="frames to run (about one path per pixel and a few bounces each)")
    ap.add_argument("--camera", type=int, default=0)
    ap.add_argument("--tone", default="aces", choices=["clamp", "reinhard", "hable", "aces"])
    ap.add_argument("--brightness", type=float, default=1.0)
    ap.add_argument("--builder", default="sah-host", choices=["sah-host", "lbvh-gpu", "ploc-gpu", "auto"])
    ap.add_argument("--out", default="render.png")
    ap.add_argument("--pfm", default=None)
    args = ap.parse_args()

    from path_tracer_b200 import integrator as I
    from path_tracer_b200 import scene as S
    ctx = I.Context(0)
    S.ensure_spectrum_table(ctx)
    if args.obj:
        sc = S.Scene(with_defaults=True)
        prefab = sc.load_model(args.obj)
        sc.create_prefab_instance(prefab)
        sc.pack()
        lo, hi = sc.bounds()
        lo, hi = np.maximum(lo, -1e3), np.minimum(hi, 1e3)
        centre, radius = 0.5 * (lo + hi), 0.5 * float(np.linalg.norm(hi - lo))
        cam = sc.create_camera(tuple(centre + np.array([0.0, -2.2 * radius, 0.9 * radius])), tuple(centre), params=(45.0, 0.0))
        sc.pack()
        args.camera = sc.camera_index(cam)
    else:
        sc = S.Scene.make(args.scene, *args.scene_params)
    builder = {"sah-host": I.BUILDER_SAH_HOST, "lbvh-gpu": I.BUILDER_LBVH_GPU, "ploc-gpu": I.BUILDER_PLOC_GPU, "auto": I.BUILDER_AUTO}[args.builder]
    t0 = time.time()
    dev = I.DeviceScene(ctx, sc, builder=builder)
    info = dev.wide_info()
    w, h = args.size
    film = I.SampleBuffer(ctx, w, h)
    r = I.BasicRenderer(ctx, dev, film)
    r.CameraIndex = args.camera
    r.set_frame_index(0)
    r.reset()
    r.run(2)
    t1 = time.time()
    r.run_frames(args.spp)
    r.sync()
    t2 = time.time()
    tone = {"clamp": I.TONE_CLAMP, "reinhard": I.TONE_REINHARD, "hable": I.TONE_HABLE, "aces": I.TONE_ACES}[args.tone]
    lin, srgb = film.resolve(args.brightness, tone, 1.0)
    S.write_png(args.out, srgb)
    if args.pfm:
        S.write_pfm(args.pfm, lin)
    st = r.stats()
    print("scene upload + BVH8 (%d nodes, %d primitives): %.3f s; %d frames in %.3f s = %.0f Mrays/s; %.1f paths/pixel -> %s"
          % (info["node_count"], info["prim_count"], t1 - t0, args.spp, t2 - t1, st["rays_traced"] / (t2 - t1) / 1e6,
             st["paths_finished"] / float(w * h), args.out))


if __name__ == "__main__":
    main()
