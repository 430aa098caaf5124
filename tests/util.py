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
