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
