"""Shared helpers for the parity tests."""
import numpy as np


def random_rays(scene, n, seed, oracle):
    """Uniform origins in the (clamped) scene bounds, uniform directions."""
    rng = np.random.default_rng(seed)
    lo, hi = scene.bounds()
    lo = np.maximum(lo, -20.0)
    hi = np.minimum(hi, 20.0)
    o = rng.uniform(lo - 1.0, hi + 1.0, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return oracle.make_rays(o, d.astype(np.float32))


def camera_rays(oracle, scene, w, h, seed=0):
    """Primary rays of camera 0 at pixel centres (through the oracle's camera model)."""
    import ctypes as C
    cams = scene.cameras
    rays = np.zeros(w * h, dtype=oracle.RAY_DTYPE)
    L = oracle.lib()
    for y in range(h):
        for x in range(w):
            L.orc_kat_camera_ray(cams.ctypes.data, (x + 0.5) / w, (y + 0.5) / h, seed + y * w + x,
                                 rays[y * w + x:].ctypes.data)
    return rays


def ids_equal(a, b):
    return (a["shape"] == b["shape"]) & (a["prim"] == b["prim"])


def bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def unpack_unit_vectors(oracle, packed):
    import ctypes as C
    out = np.zeros((len(packed), 3), np.float32)
    tmp = (C.c_float * 3)()
    L = oracle.lib()
    for i, p in enumerate(packed):
        L.orc_kat_unpack_unit_vector(int(p), tmp)
        out[i] = tmp[:]
    return out


def d65_observation_weights():
    """observer(lambda) * D65(lambda) / 10566.864005 at the 471 integer wavelengths 360..830 nm
    (spectrum.cpp:137-143,151-176,190-208), float64, for checking fitted spectra in colour space."""
    import os
    import re
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "path_tracer_b200", "host", "d65_table.inc")
    txt = "".join(l for l in open(path) if not l.lstrip().startswith("//"))
    d65 = np.array([float(x) for x in re.findall(r"([0-9.]+)f", txt)])
    assert len(d65) == 471
    lam = np.arange(360.0, 831.0)

    def g(mu, s1, s2):
        t = (lam - mu) * np.where(lam < mu, s1, s2)
        return np.exp(-0.5 * t * t)
    x = 0.362 * g(442.0, 0.0624, 0.0374) + 1.056 * g(599.8, 0.0264, 0.0323) - 0.065 * g(501.1, 0.0490, 0.0382)
    y = 0.821 * g(568.8, 0.0213, 0.0247) + 0.286 * g(530.9, 0.0613, 0.0322)
    z = 1.217 * g(437.0, 0.0845, 0.0278) + 0.681 * g(459.0, 0.0385, 0.0725)
    return lam, np.stack([x, y, z], axis=1) * (d65 / 10566.864005)[:, None]


def xyz_to_lab(c):
    """spectrum.cpp:213-233 (L driven by f(X), as the reference has it)."""
    d = 6 / 29.0

    def f(t):
        return np.where(t > d ** 3, np.cbrt(np.maximum(t, 0.0)), t / (3 * d * d) + 4 / 29.0)
    fx, fy, fz = f(c[..., 0] / 0.950489), f(c[..., 1]), f(c[..., 2] / 1.088840)
    return np.stack([116.0 * fx - 16.0, 500.0 * (fx - fy), 200.0 * (fy - fz)], axis=-1)


def check_wide_tree(nodes, prims, shapes):
    """Structural check of a flattened BVH8 (layout: csrc/ptb_wide.cuh): every node reachable once, every leaf
    primitive referenced once, and every child slot's dequantised box CONTAINS the world-space bounds of
    everything below it (so culling can never lose a hit).  Returns the depth.  numpy, level by level."""
    n_nodes, n_prims = len(nodes), len(prims)
    if n_nodes == 0:
        assert n_prims == 0
        return 0
    # world bounds of the leaf primitives (float64, exact transform of float32 inputs)
    tag = prims[:, 11].view(np.uint32)
    shape = prims[:, 7].view(np.uint32)
    kind = tag & 3
    to = shapes["To"].reshape(-1, 4, 4).astype(np.float64)            # column-major: to[s, col, row]
    corners = np.array([[(c & 1) * 2 - 1, ((c >> 1) & 1) * 2 - 1, ((c >> 2) & 1) * 2 - 1] for c in range(8)], np.float64)
    plo = np.full((n_prims, 3), np.inf)
    phi = np.full((n_prims, 3), -np.inf)
    m = to[shape]
    for pts in ([prims[:, 0:3], prims[:, 4:7], prims[:, 8:11]], [np.broadcast_to(c, (n_prims, 3)) for c in corners]):
        tri = pts is not None and len(pts) == 3
        sel = (kind == 0) if tri else (kind != 0)
        for p in pts:
            w = np.einsum("nk,nkr->nr", np.asarray(p, np.float64), m[:, :3, :3]) + m[:, 3, :3]
            plo[sel] = np.minimum(plo[sel], w[sel])
            phi[sel] = np.maximum(phi[sel], w[sel])
    w3 = nodes[:, 3]
    scale = np.stack([np.ldexp(1.0, ((w3 >> (8 * a)) & 0xFF).astype(np.int64) - 127) for a in range(3)], axis=1)
    origin = nodes[:, 0:3].view(np.float32).astype(np.float64)
    imask = (w3 >> 24) & 0xFF
    q = nodes[:, 8:20].copy().view(np.uint8).reshape(n_nodes, 6, 8).astype(np.float64)   # [node, qlo xyz | qhi xyz, slot]
    slot_lo = origin[:, :, None] + q[:, 0:3, :] * scale[:, :, None]
    slot_hi = origin[:, :, None] + q[:, 3:6, :] * scale[:, :, None]
    valid24 = nodes[:, 6]
    # top-down: levels and parent slots; bottom-up: true bounds
    level = [np.array([0])]
    seen_nodes = np.zeros(n_nodes, bool)
    seen_nodes[0] = True
    seen_prims = np.zeros(n_prims, np.int64)
    child_of = {}
    while True:
        cur = level[-1]
        nxt = []
        for s in range(8):
            inner = cur[(imask[cur] >> s) & 1 == 1]
            if len(inner):
                rel = np.array([bin(int(x) & ((1 << s) - 1)).count("1") for x in imask[inner]])
                ch = nodes[inner, 4] + rel
                assert ch.max() < n_nodes and not seen_nodes[ch].any() and len(np.unique(ch)) == len(ch)
                seen_nodes[ch] = True
                child_of[s] = child_of.get(s, []) + [(inner, ch)]
                nxt.append(ch)
        if not nxt:
            break
        level.append(np.concatenate(nxt))
        assert len(level) <= 64
    assert seen_nodes.all()
    blo = np.full((n_nodes, 3), np.inf)
    bhi = np.full((n_nodes, 3), -np.inf)
    for cur in reversed(level):
        for s in range(8):
            bits = (valid24[cur] >> (3 * s)) & 7
            leaf = cur[bits != 0]
            if len(leaf):
                assert ((imask[leaf] >> s) & 1 == 0).all()
                below = np.array([bin(int(v) & ((1 << (3 * s)) - 1)).count("1") for v in valid24[leaf]])
                for k in range(3):
                    has = leaf[((valid24[leaf] >> (3 * s + k)) & 1) == 1]
                    if not len(has):
                        continue
                    off = below[((valid24[leaf] >> (3 * s + k)) & 1) == 1] + np.array(
                        [bin(int(v >> (3 * s)) & ((1 << k) - 1)).count("1") for v in valid24[has]])
                    pi = nodes[has, 5] + off
                    assert pi.max() < n_prims
                    np.add.at(seen_prims, pi, 1)
                    assert (slot_lo[has, :, s] <= plo[pi]).all() and (slot_hi[has, :, s] >= phi[pi]).all(), "leaf slot box misses its primitive"
                    np.minimum.at(blo, has, plo[pi])
                    np.maximum.at(bhi, has, phi[pi])
            inner = cur[(imask[cur] >> s) & 1 == 1]
            if len(inner):
                rel = np.array([bin(int(x) & ((1 << s) - 1)).count("1") for x in imask[inner]])
                ch = nodes[inner, 4] + rel
                assert (slot_lo[inner, :, s] <= blo[ch]).all() and (slot_hi[inner, :, s] >= bhi[ch]).all(), "inner slot box misses its subtree"
                np.minimum.at(blo, inner, blo[ch])
                np.maximum.at(bhi, inner, bhi[ch])
    assert (seen_prims == 1).all()
    return len(level)


def mixed_material_scene(S, fog=0.0):
    """One small scene through every branch the basic integrator's scatter pass has: the three basic materials
    (textured base colour, textured + anisotropic roughness, dispersive glass with an interior scattering medium),
    an importance-sampled sky texture, a thin-lens camera and a 360 camera; fog > 0 fills the root with a scattering
    medium (no path then reaches the sky: paths end by Russian roulette only, radiance stays 0)."""
    sc = S.Scene()
    rng = np.random.default_rng(5)
    sky = np.zeros((16, 32, 4), np.float32)
    sky[..., :3] = rng.uniform(0.1, 0.6, size=(16, 32, 3))
    sky[3:6, 8:12, :3] = 20.0                                     # a bright patch for the von Mises-Fisher lobe
    sky[..., 3] = 1.0
    sky_t = sc.create_texture("sky", S.TEXTURE_TYPE_RADIANCE, sky)
    sc.set_root(scatter_rate=fog, sky_brightness=1.5, sky_sampling_probability=0.5, sky_texture=sky_t)
    check = sc.create_checker_texture("check", S.TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, (0.8, 0.8, 0.8, 1.0), (0.2, 0.3, 0.7, 1.0), nearest=True)
    rough = np.zeros((8, 8, 4), np.float32)
    rough[..., 0] = rng.uniform(0.05, 0.6, size=(8, 8))
    rough[..., 3] = 1.0
    rough_t = sc.create_texture("rough", S.TEXTURE_TYPE_RAW, rough)
    floor = sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "floor", textures={"BaseTexture": check}, BaseColor=(0.9, 0.9, 0.9))
    metal = sc.create_material(S.MATERIAL_BASIC_METAL, "metal", textures={"RoughnessTexture": rough_t}, BaseColor=(0.9, 0.6, 0.3),
                               SpecularColor=(1.0, 0.9, 0.8), Roughness=0.4, RoughnessAnisotropy=0.6)
    glass = sc.create_material(S.MATERIAL_BASIC_TRANSLUCENT, "glass", IOR=1.5, AbbeNumber=30.0, Roughness=0.1, TransmissionColor=(0.8, 0.9, 1.0),
                               TransmissionDepth=0.7, ScatteringColor=(0.9, 0.5, 0.4), ScatteringAnisotropy=0.3)
    mirror = sc.create_material(S.MATERIAL_BASIC_METAL, "mirror", BaseColor=(0.95, 0.95, 0.95), Roughness=0.0)
    sc.create_entity(S.ENTITY_PLANE, position=(0, 0, 0), scale=(6, 6, 1), material=floor)
    sc.create_entity(S.ENTITY_SPHERE, position=(-1.2, 0.3, 1.0), material=glass)
    sc.create_entity(S.ENTITY_CUBE, position=(1.3, 0.0, 0.8), rotation=(0.0, 0.0, 0.5), scale=(0.8, 0.8, 0.8), material=metal)
    sc.create_entity(S.ENTITY_SPHERE, position=(0.2, 2.0, 0.7), scale=(0.7, 0.7, 0.7), material=mirror)
    sc.create_entity(S.ENTITY_SPHERE, position=(0.0, -0.8, 0.4), scale=(0.4, 0.4, 0.4), material=glass)
    sc.create_camera((0.0, -5.0, 2.0), (0.0, 0.0, 0.8), model=S.CAMERA_THIN_LENS, params=(36.0, 24.0, 35.0, 8.0, 5.2))
    sc.create_camera((0.0, -1.5, 1.5), (0.0, 0.0, 0.8), model=S.CAMERA_360)
    sc.pack()
    return sc


# Cases of tests/golden/make_ref_golden.py: tag -> (scene, params, camera, render flags, termination probability).
# PathLengthLimit stays 0: the reference's scatter shader never reads it.
REF_CASES = {
    "cornell": ("cornell", (), 0, 3, 0.0),
    "terrain32": ("terrain", (32,), 0, 3, 0.0),
    "room": ("room", (1, 256), 0, 3, 0.0),
    "spherefield": ("spherefield", (6, 1, 2), 0, 3, 0.0),
    "mixed_lens": ("mixed", (0.0,), 0, 3, 0.25),
    "mixed_360": ("mixed", (0.0,), 1, 1, 0.0),
    "mixed_fog": ("mixed", (0.3,), 0, 3, 0.25),
}


def make_test_scene(S, name, params):
    return mixed_material_scene(S, *params) if name == "mixed" else S.Scene.make(name, *params)
