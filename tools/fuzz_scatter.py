#!/usr/bin/env python
"""Differential fuzz of the product's SCATTER / path-state code against the oracle on random scenes.  CPU only.

tools/fuzz_traversal.py covers what a ray hits; this tool covers what happens next.  Random scenes with every material
type over its whole parameter range (roughness 0 and 1, anisotropy to both ends, IOR 1 to 3, Abbe numbers from strongly
dispersive to none, transmission depth and scattering colours that switch the interior medium on and off), textured and
untextured attributes (nearest and bilinear), dielectrics nested in dielectrics and overlapping each other (the
active-shape priority list, ACTIVE_SHAPE_LIMIT = 4), a participating medium on the scene root, textured and default sky
with and without sky sampling, all three camera models, every render-flag combination, path-length limits and
termination probabilities.  The product's per-thread device functions (csrc/*.cuh, compiled for the host by
tests/host_emu with the same libm) render a small film next to the oracle (oracle/sh_*.h, pinned against the
reference's own shaders) with the same seeds:

    the RGBA32F accumulators must be EQUAL, bit for bit, after reset + 2 rounds + 6 one-round frames.

    python tools/fuzz_scatter.py [--scenes 200] [--seed 1] [--width 40] [--height 24] [--reference]

--reference (where oracle/_ref/libref.so can be built or has travelled) renders the same scenes with the reference's OWN
shaders first and demands that the oracle equals them - film and every wavefront state array after every round - up to
the first round in which the reference reads a medium it never wrote (see the comment in main).  Exit code 1 and the
seed of the scene on the first difference.  Round 2: 800 scenes with --reference (seeds 5, 6), 6.6 M finished paths,
5445 rounds compared with the reference's shaders, no difference; 192 scenes reach the undefined medium."""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def random_texture(sc, S, rng, k, ttype):
    w, h = int(rng.integers(1, 9)), int(rng.integers(1, 9))
    px = rng.uniform(0, 1, size=(h, w, 4)).astype(np.float32)
    if ttype == S.TEXTURE_RADIANCE:
        px[..., :3] *= np.float32(rng.choice([0.5, 4.0, 50.0]))
        if rng.random() < 0.3:
            px[0, 0, :3] = 0.0
    if rng.random() < 0.2:
        px[..., 3] = (px[..., 3] > 0.5)                       # cut-out alpha
    return sc.create_texture("t%d" % k, ttype, px, nearest=bool(rng.integers(0, 2)))


def random_material(sc, S, rng, k, tex_r, tex_raw, unit_ior=True):
    t = rng.integers(0, 3)
    pick = lambda pool: pool[rng.integers(0, len(pool))] if pool and rng.random() < 0.4 else None
    col = lambda: tuple(rng.choice([0.0, 1.0], size=3)) if rng.random() < 0.15 else tuple(rng.uniform(0, 1, 3))
    rough = float(rng.choice([0.0, 0.0, 1e-3, 0.05, 0.3, 0.7, 1.0]))
    aniso = float(rng.choice([0.0, 0.0, -0.95, -0.4, 0.4, 0.95]))
    tx = {}
    if t == 0:
        if pick(tex_r) is not None:
            tx["BaseTexture"] = pick(tex_r) or tex_r[0]
        return sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "d%d" % k, BaseColor=col(), textures=tx or None)
    if t == 1:
        for name, pool in (("BaseTexture", tex_r), ("SpecularTexture", tex_r), ("RoughnessTexture", tex_raw), ("RoughnessAnisotropyTexture", tex_raw)):
            p = pick(pool)
            if p is not None:
                tx[name] = p
        return sc.create_material(S.MATERIAL_BASIC_METAL, "m%d" % k, BaseColor=col(), SpecularColor=col(), Roughness=rough, RoughnessAnisotropy=aniso,
                                  textures=tx or None)
    for name, pool in (("RoughnessTexture", tex_raw), ("RoughnessAnisotropyTexture", tex_raw)):
        p = pick(pool)
        if p is not None:
            tx[name] = p
    return sc.create_material(S.MATERIAL_BASIC_TRANSLUCENT, "t%d" % k, IOR=float(rng.choice([1.0 if unit_ior else 1.05, 1.05, 1.33, 1.5, 2.4, 3.0])),
                              AbbeNumber=float(rng.choice([5.0, 20.0, 60.0, 1e4])), Roughness=rough, RoughnessAnisotropy=aniso,
                              TransmissionColor=col(), TransmissionDepth=float(rng.choice([0.0, 0.0, 0.05, 0.5, 5.0])), ScatteringColor=col(),
                              ScatteringAnisotropy=float(rng.choice([0.0, -0.8, 0.3, 0.9])), textures=tx or None)


def random_scene(S, rng, unit_ior=True, make=None, with_meshes=True):
    import fuzz_traversal as FT
    sc = (make or S.Scene)()
    tex_r = [random_texture(sc, S, rng, k, S.TEXTURE_REFLECTANCE_WITH_ALPHA) for k in range(int(rng.integers(0, 3)))]
    tex_raw = [random_texture(sc, S, rng, 10 + k, S.TEXTURE_RAW) for k in range(int(rng.integers(0, 3)))]
    sky = random_texture(sc, S, rng, 20, S.TEXTURE_RADIANCE) if rng.random() < 0.5 else None
    sc.set_root(scatter_rate=float(rng.choice([0.0, 0.0, 0.02, 0.3])), sky_brightness=float(rng.choice([0.3, 1.0, 4.0])),
                sky_sampling_probability=float(rng.choice([0.0, 0.0, 0.3, 0.9])) if sky is not None else float(rng.choice([0.0, 0.5])), sky_texture=sky)
    mats = [random_material(sc, S, rng, k, tex_r, tex_raw, unit_ior) for k in range(int(rng.integers(1, 6)))]
    mat = lambda: mats[rng.integers(0, len(mats))]
    meshes = [sc.create_mesh("m%d" % k, *FT.random_mesh(rng, False)) for k in range(int(rng.integers(0, 3)) if with_meshes else 0)]
    if rng.random() < 0.7:
        sc.create_entity(S.ENTITY_PLANE, position=(0, 0, float(rng.uniform(-2, 0))), material=mat())
    for k in range(int(rng.integers(1, 9))):
        pos = tuple(rng.normal(size=3) * 1.5)
        rot = tuple(rng.uniform(-3, 3, size=3) * rng.choice([0.0, 1.0]))
        scale = tuple(rng.uniform(0.3, 1.6, size=3) * rng.choice([0.5, 1.0, 2.0]))
        t = rng.integers(0, 4)
        if t == 0:
            sc.create_entity(S.ENTITY_SPHERE, position=pos, rotation=rot, scale=scale, material=mat())
        elif t == 1:
            sc.create_entity(S.ENTITY_CUBE, position=pos, rotation=rot, scale=scale, material=mat())
        elif t == 2 and meshes:
            sc.create_entity(S.ENTITY_MESH_INSTANCE, position=pos, rotation=rot, scale=scale, material=mat(), mesh=meshes[rng.integers(0, len(meshes))])
        else:                                                  # nested and overlapping volumes around one centre
            for j in range(int(rng.integers(2, 6))):
                s = 2.0 * 0.7 ** j
                sc.create_entity((S.ENTITY_SPHERE, S.ENTITY_CUBE)[j & 1], position=tuple(np.array(pos) + rng.normal(size=3) * 0.1 * j), scale=(s, s, s), material=mat())
    cams = []
    for model, params in ((S.CAMERA_PINHOLE, (float(rng.uniform(20, 120)), float(rng.choice([0.0, 5.0])))),
                          (S.CAMERA_THIN_LENS, (36.0, 24.0, float(rng.uniform(20, 80)), float(rng.choice([1.0, 12.0])), float(rng.uniform(1, 8)))),
                          (S.CAMERA_360, ())):
        eye = rng.normal(size=3) * 4.0
        if rng.random() < 0.2:
            eye = rng.normal(size=3) * 0.3                     # inside the cluster: paths that start within volumes
        cams.append(sc.create_camera(tuple(eye), (0, 0, 0), model=model, params=params))
    sc.pack()
    return sc, len(cams)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", type=int, default=200)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--width", type=int, default=40)
    ap.add_argument("--height", type=int, default=24)
    ap.add_argument("--reference", action="store_true", help="also render with the reference's OWN shaders (oracle/_ref/libref.so, where /root/reference is "
                    "present or the library travelled): the oracle must equal them bit for bit on these scenes too (width must be a multiple of 16)")
    ap.add_argument("--gpu", action="store_true", help="on a GPU box: the shipped library through the C ABI (BVH8, cooperative trace kernel, scatter kernels "
                    "with material queues, then the megakernel) against the oracle, radiance within 1e-4 (CUDA's transcendentals are not glibc's)")
    args = ap.parse_args(argv)
    import fuzz_traversal as FT
    from oracle import binding as oracle
    from path_tracer_b200 import scene as S
    oracle.lib()
    S.ensure_spectrum_table()
    if args.gpu:
        return main_gpu(args, oracle, S)
    emu = FT.load_emu()
    emu.emu_set.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_uint32]
    emu.emu_reset.argtypes = [C.c_void_p]
    emu.emu_run.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int]
    emu.emu_accum.restype = C.c_void_p
    emu.emu_accum.argtypes = [C.c_void_p]
    w, h = args.width, args.height
    paths = 0
    ref_rounds = 0
    counts = {}
    R = None
    if args.reference:
        from oracle.ref import binding as R
        assert R.available(), "oracle/_ref/libref.so is not built and /root/reference is not here"
        R.lib()
        assert w % 16 == 0
    for k in range(args.scenes):
        seed = args.seed * 100003 + k
        rng = np.random.default_rng(seed)
        sc, n_cams = random_scene(S, rng)
        cam = int(rng.integers(0, n_cams))
        flags = int(rng.integers(0, 4))
        limit = int(rng.choice([0, 0, 1, 2, 5]))
        term = float(rng.choice([0.0, 0.0, 0.1, 0.5]))
        frame = int(rng.integers(0, 1000))
        if R is not None:                                      # the reference's shaders against the oracle (faithful ties are the reference's)
            # PathLengthLimit is a push constant the reference's shaders never read (basic.glsl.inc:76-85): no limit here
            o = oracle.Renderer(sc.view, w, h, camera_index=cam, render_flags=flags, path_length_limit=0, path_termination_probability=term, tie=oracle.TIE_FAITHFUL)
            r = R.Renderer(sc.view, w, h, camera_index=cam, render_flags=flags, path_length_limit=0, path_termination_probability=term)
            # The reference leaves `medium` UNWRITTEN for a diffuse or metal shape (BasicDiffuse_LoadMedium / BasicMetal_LoadMedium
            # return false without touching their `out` argument, basic_diffuse.glsl.inc:8-11, basic_metal.glsl.inc:33-36; ResolveMedium,
            # basic_scatter.glsl:44-64, uses it regardless): undefined in GLSL.  It is reached when a path is inside such a shape - it got
            # there through an alpha cut-out - while inside or leaving a dielectric.  The oracle and the product define it as vacuum
            # (IOR 1, no absorption, no scattering); this build of the shaders reads zeros.  The comparison stops for a scene at the first
            # round after which some path has such a shape in its active list.
            shapes = sc.shapes
            mat_words = np.asarray(sc.materials).view(np.uint32).reshape(-1)
            not_dielectric = np.array([mat_words[32 * int(m)] != 2 for m in shapes["MaterialIndex"]] + [False], bool)      # + none
            undefined_from = None
            for x in (o, r):
                x.set_frame_index(frame)
                x.reset()
            for step in range(7):
                act = np.concatenate([o.state("ActiveShapeIndex01") & 0xFFFF, o.state("ActiveShapeIndex01") >> 16,
                                      o.state("ActiveShapeIndex23") & 0xFFFF, o.state("ActiveShapeIndex23") >> 16])
                act = np.where(act == 0xFFFF, len(shapes), act)
                if not_dielectric[np.minimum(act, len(shapes))].any():
                    undefined_from = step
                    counts["scenes where the reference's medium becomes undefined"] = counts.get("scenes where the reference's medium becomes undefined", 0) + 1
                    break
                for x in (o, r):
                    x.run(2 if step == 0 else 1)
                a, b = r.accum(), o.accum()
                same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
                if not same.all():
                    y, xx, c = [int(v[0]) for v in np.nonzero(~same)]
                    print("seed", seed, "ORACLE != REFERENCE SHADERS after step", step, ": camera", cam, "flags", flags, "termination", term, "frame", frame,
                          ": %d of %d accumulator words differ, first at pixel (%d, %d) channel %d: reference %r, oracle %r" % (int((~same).sum()), same.size, xx, y, c, a[y, xx, c], b[y, xx, c]))
                    return 1
                for arr in R.STATE_ARRAYS:
                    xa, xb = r.state(arr).view(np.uint32), o.state(arr).view(np.uint32)
                    if arr in ("Time", "TextureU", "TextureV", "PackedNormal", "PackedTangentX"):     # unwritten on a miss
                        hit = r.state("ShapeAndMaterialIndex") != 0xFFFFFFFF
                        xa, xb = xa[hit], xb[hit]
                    if not np.array_equal(xa, xb):
                        print("seed", seed, "ORACLE != REFERENCE SHADERS after step", step, "in the wavefront state array", arr, ":", int((xa != xb).sum()), "slots")
                        return 1
                ref_rounds += 2 if step == 0 else 1
            o.close()
        for mode, tie in ((0, oracle.TIE_FAITHFUL), (1, oracle.TIE_CANONICAL)):
            e = emu.emu_create(C.addressof(sc.view), w, h)
            r = oracle.Renderer(sc.view, w, h, camera_index=cam, render_flags=flags, path_length_limit=limit, path_termination_probability=term, tie=tie)
            r.set_frame_index(frame)
            r.reset()
            r.run(2)
            r.run_frames(6)
            emu.emu_set(e, cam, flags, limit, term, frame)
            emu.emu_reset(e)
            emu.emu_run(e, 2, mode, 0)
            emu.emu_run(e, 6, mode, 1)
            a = r.accum()
            b = np.frombuffer((C.c_float * (w * h * 4)).from_address(emu.emu_accum(e)), dtype=np.float32).reshape(h, w, 4)
            same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
            if not same.all():
                y, x, c = [int(v[0]) for v in np.nonzero(~same)]
                print("seed", seed, "mode", mode, "camera", cam, "flags", flags, "limit", limit, "termination", term, "frame", frame,
                      ": %d of %d accumulator words differ, first at pixel (%d, %d) channel %d: oracle %r, product %r" % (int((~same).sum()), same.size, x, y, c, a[y, x, c], b[y, x, c]))
                return 1
            paths += int(np.nan_to_num(a[..., 3]).sum())
            emu.emu_destroy(e)
            r.close()
        if (k + 1) % 20 == 0:
            print("%d scenes (x 2 tie rules): accumulators equal, %d paths finished" % (k + 1, paths), flush=True)
    print("done: %d scenes, %d finished paths, every accumulator word equal%s" % (args.scenes, paths, " (reference shaders = oracle = product; %d rounds compared with the reference's shaders; %s)" % (ref_rounds, counts or "none undefined") if R is not None else ""))
    return 0


def main_gpu(args, oracle, S):
    """Same scenes, shipped kernels.  A random branch that flips on an ulp of a transcendental changes a whole sample and
    every later one of that slot, so a small fraction of pixels may deviate: counted per scene, bounded at 2 %."""
    from path_tracer_b200 import integrator as I
    ctx = I.Context(0)
    w, h = args.width, args.height
    worst = (0.0, None)
    pixels = outliers = alpha_diff = 0
    for k in range(args.scenes):
        seed = args.seed * 100003 + k
        rng = np.random.default_rng(seed)
        # No IOR of exactly 1 with a rough interface here: the refraction half vector -(eta_i in + eta_o out) is then a difference of
        # nearly equal vectors, and an ulp in sinf / cosf / logf (CUDA's are not glibc's) decides which way a path goes - for 0.5 % of
        # the pixels in the first round of such a scene (seed 600119 of round 2).  The CPU comparison, same libm, keeps it.
        sc, n_cams = random_scene(S, rng, unit_ior=False)
        cam = int(rng.integers(0, n_cams))
        flags = int(rng.integers(0, 4))
        limit = int(rng.choice([0, 0, 1, 2, 5]))
        term = float(rng.choice([0.0, 0.0, 0.1, 0.5]))
        frame = int(rng.integers(0, 1000))
        o = oracle.Renderer(sc.view, w, h, camera_index=cam, render_flags=flags, path_length_limit=limit, path_termination_probability=term, tie=oracle.TIE_CANONICAL)
        o.set_frame_index(frame)
        o.reset()
        o.run(2)
        o.run_frames(4)
        b = o.accum()
        dev = I.DeviceScene(ctx, sc)
        for integrator in (I.INTEGRATOR_WAVEFRONT, I.INTEGRATOR_MEGAKERNEL):
            film = I.SampleBuffer(ctx, w, h)
            r = I.BasicRenderer(ctx, dev, film)
            r.CameraIndex, r.RenderFlags, r.PathLengthLimit, r.PathTerminationProbability = cam, flags, limit, term
            r.TieMode, r.Traversal, r.Integrator = I.TIE_CANONICAL, I.TRAVERSAL_WIDE, integrator
            r.set_frame_index(frame)
            r.reset()
            r.run(2)
            r.run_frames(4)
            r.sync()
            a = film.download()
            both_nan = np.isnan(a) & np.isnan(b)
            err = np.abs(a - b) / np.maximum(np.abs(b), 1.0)
            err[both_nan] = 0.0
            err = np.nan_to_num(err, nan=np.inf)
            bad = err[..., :3].max(axis=-1) > 1e-4
            ad = a[..., 3] != b[..., 3]
            frac = float((bad | ad).mean())
            pixels += bad.size
            outliers += int(bad.sum())
            alpha_diff += int(ad.sum())
            if frac > worst[0]:
                worst = (frac, seed)
            if frac > 0.02:
                y, x = [int(v[0]) for v in np.nonzero(bad | ad)]
                print("seed", seed, "integrator", integrator, "camera", cam, "flags", flags, "limit", limit, "termination", term, "frame", frame,
                      ": %d of %d pixels beyond 1e-4 (%d with another sample count), first (%d, %d): oracle %r, kernels %r" % (int((bad | ad).sum()), bad.size, int(ad.sum()), x, y, b[y, x], a[y, x]))
                return 1
            r.close()
            film.close()
        dev.close()
        o.close()
        if (k + 1) % 20 == 0:
            print("%d scenes x 2 integrators: %d of %d pixels beyond 1e-4 (%d with another sample count), worst scene %.2f %% (seed %s)"
                  % (k + 1, outliers, pixels, alpha_diff, 100 * worst[0], worst[1]), flush=True)
    print("done (GPU, C ABI): %d scenes x 2 integrators, %d of %d pixels beyond 1e-4 of the oracle (%d with another sample count), worst scene %.2f %% (seed %s)"
          % (args.scenes, outliers, pixels, alpha_diff, 100 * worst[0], worst[1]))
    return 0


if __name__ == "__main__":
    sys.exit(main())
