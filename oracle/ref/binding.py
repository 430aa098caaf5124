"""ORACLE — TEST INFRASTRUCTURE ONLY.

oracle/_ref/libref.so: the reference's own trace / scatter shader sources compiled as C++ (see ref_api.cpp,
transpile.py, ref_glsl.h).  It can only be BUILT where /root/reference exists (this container); the built library
travels to the GPU box with gpurun like the other .so files, and the golden vectors generated from it
(tests/golden/ref_*.npz, tests/golden/make_ref_golden.py) are committed.  Same call shapes as oracle/binding.py.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(os.path.dirname(HERE), "_ref")
LIB = os.path.join(OUT, "libref.so")
REFERENCE_SRC = "/root/reference/src"

RAY_DTYPE = np.dtype([("o", "<f4", 3), ("d", "<f4", 3), ("duration", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("shape", "<u4"), ("prim", "<u4"), ("material", "<u4"), ("shape_type", "<u4"),
                      ("n", "<f4", 3), ("tan", "<f4", 3), ("uv", "<f4", 2), ("bary", "<f4", 3),
                      ("scene_complexity", "<u4"), ("mesh_complexity", "<u4")])
STATE_ARRAYS = {
    "OriginX": "f4", "OriginY": "f4", "OriginZ": "f4", "Duration": "f4", "PackedVelocity": "u4",
    "Time": "f4", "TextureU": "f4", "TextureV": "f4", "ShapeAndMaterialIndex": "u4", "PackedNormal": "u4", "PackedTangentX": "u4",
    "ImagePosition": "i4", "NormalizedLambda0": "f4", "Throughput0": "f4", "Throughput1": "f4", "Throughput2": "f4", "Throughput3": "f4",
    "Probability0": "f4", "Probability1": "f4", "Probability2": "f4", "Probability3": "f4", "SampleR": "f4", "SampleG": "f4", "SampleB": "f4",
    "ActiveShapeIndex01": "u4", "ActiveShapeIndex23": "u4",
}


class Params(C.Structure):
    _fields_ = [("CameraIndex", C.c_uint32), ("RenderFlags", C.c_uint32), ("PathLengthLimit", C.c_uint32),
                ("PathTerminationProbability", C.c_float)]


class PreviewParams(C.Structure):
    _fields_ = [("To", C.c_float * 16), ("From", C.c_float * 16), ("RenderMode", C.c_uint32), ("Brightness", C.c_float),
                ("SelectedShapeIndex", C.c_uint32), ("RenderSizeX", C.c_uint32), ("RenderSizeY", C.c_uint32),
                ("MouseX", C.c_uint32), ("MouseY", C.c_uint32)]


def available():
    """True when the library exists or can be built (the reference sources are present)."""
    return os.path.exists(LIB) or os.path.isdir(REFERENCE_SRC)


def build(force=False):
    """transpile.py + g++ on the reference's shader sources where they lie; outputs only under oracle/_ref/."""
    if not os.path.isdir(REFERENCE_SRC):
        if os.path.exists(LIB):
            return LIB
        raise RuntimeError("oracle/_ref: %s is not here and no prebuilt libref.so travelled with the repository" % REFERENCE_SRC)
    deps = [os.path.join(HERE, f) for f in ("ref_api.cpp", "ref_frag.cpp", "ref_bind_scene.inc", "ref_glsl.h", "transpile.py")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        os.makedirs(OUT, exist_ok=True)
        gen = os.path.join(OUT, "gen")
        subprocess.check_call([sys.executable, os.path.join(HERE, "transpile.py"), REFERENCE_SRC, gen], stdout=subprocess.DEVNULL)
        cmd = ["g++", "-std=c++20", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-w",
               "-I", gen, "-I", HERE, "-I", os.path.join(ROOT, "include"), "-o", LIB, os.path.join(HERE, "ref_api.cpp"), os.path.join(HERE, "ref_frag.cpp")]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("oracle/_ref build failed:\n" + r.stdout)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        vp, u32, u64, f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_float
        L.ref_bind_scene.argtypes = [vp]
        L.ref_raycast.argtypes = [vp, u64, vp]
        L.ref_bind_film.argtypes = [vp, u32, u32]
        L.ref_reset.argtypes = [C.POINTER(Params), u32]
        L.ref_run.argtypes = [C.POINTER(Params), u32, u32]
        L.ref_state_array.restype = vp
        L.ref_state_array.argtypes = [C.c_char_p]
        L.ref_pack_unit_vector.restype = u32
        L.ref_pack_unit_vector.argtypes = [f32, f32, f32]
        L.ref_unpack_unit_vector.argtypes = [u32, vp]
        L.ref_observer.argtypes = [f32, vp]
        L.ref_rng.argtypes = [u32, u32, vp]
        L.ref_set_groups.argtypes = [u32, u32]
        L.ref_rays_traced.restype = u64
        L.ref_slot_capacity.restype = C.c_int
        L.ref_resolve.argtypes = [vp, u32, u32, f32, u32, f32, vp]
        L.ref_preview.argtypes = [vp, C.POINTER(PreviewParams), vp, vp]
        _lib = L
    return _lib


def raycast(view, rays):
    """Trace() of the reference's scene.glsl.inc on caller rays (visiting order and tie rules are the reference's)."""
    L = lib()
    L.ref_bind_scene(C.addressof(view))
    rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
    out = np.zeros(len(rays), dtype=HIT_DTYPE)
    L.ref_raycast(rays.ctypes.data, len(rays), out.ctypes.data)
    return out


def resolve(accum, brightness=1.0, mode=0, white=1.0):
    """resolve.glsl's fragment stage over a film: (H, W, 3) linear RGB after tone mapping."""
    a = np.ascontiguousarray(accum, dtype=np.float32)
    h, w = a.shape[0], a.shape[1]
    lin = np.zeros((h, w, 3), dtype=np.float32)
    lib().ref_resolve(a.ctypes.data, w, h, brightness, mode, white, lin.ctypes.data)
    return lin


def preview(view, camera_to, camera_from, mode, width, height, brightness=1.0, selected=0xFFFFFFFF, mouse=(0, 0)):
    """preview_render.glsl's fragment stage: (H, W, 4) float image and the shape under the mouse pixel."""
    p = PreviewParams()
    p.To[:] = [float(x) for x in camera_to]
    p.From[:] = [float(x) for x in camera_from]
    p.RenderMode, p.Brightness, p.SelectedShapeIndex = mode, brightness, selected
    p.RenderSizeX, p.RenderSizeY, p.MouseX, p.MouseY = width, height, mouse[0], mouse[1]
    img = np.zeros((height, width, 4), dtype=np.float32)
    pick = C.c_uint32(0xFFFFFFFF)
    lib().ref_preview(C.addressof(view), C.byref(p), img.ctypes.data, C.addressof(pick))
    return img, pick.value


class Renderer:
    """ResetBasicRenderer / RunBasicRenderer (basic.cpp:285-332) over the reference's shaders.  One at a time: the
    shaders' buffers are globals of the library."""

    def __init__(self, view, width, height, camera_index=0, render_flags=3, path_length_limit=0, path_termination_probability=0.0):
        assert width % 16 == 0, "the reference's slot index formula (ImageSize.x / 16) assumes whole tiles per row"
        self.L = lib()
        assert self.slots_needed(width, height) <= self.L.ref_slot_capacity(), "more path slots than the reference's TRACE_COUNT"
        self.view = view
        self.w, self.h = width, height
        self.film = np.zeros((height, width, 4), dtype=np.float32)
        self.p = Params(camera_index, render_flags, path_length_limit, path_termination_probability)
        self.frame_index = 0

    def _bind(self):
        self.L.ref_bind_scene(C.addressof(self.view))
        self.L.ref_bind_film(self.film.ctypes.data, self.w, self.h)

    def set_frame_index(self, f):
        self.frame_index = f

    def reset(self):
        self._bind()
        self.L.ref_reset(C.byref(self.p), self.frame_index)

    def run(self, rounds=1):
        self._bind()
        self.frame_index += 1                       # basic.cpp:315
        self.L.ref_run(C.byref(self.p), self.frame_index, rounds)

    def run_frames(self, frames):
        for _ in range(frames):
            self.run(1)

    @staticmethod
    def slots_needed(width, height):
        return (width // 16) * ((height + 15) // 16) * 256

    @property
    def slot_count(self):
        return self.slots_needed(self.w, self.h)

    def set_groups(self, modulus, residue):
        """Only 16x16 work groups with index % modulus == residue are dispatched (a bounded / sharded sample)."""
        self.L.ref_set_groups(modulus, residue)

    def rays_traced(self):
        return int(self.L.ref_rays_traced())

    def accum(self):
        return self.film.copy()

    def state(self, name):
        ptr = self.L.ref_state_array(name.encode())
        n = self.slot_count
        buf = (C.c_char * (n * 4)).from_address(ptr)
        return np.frombuffer(buf, dtype=np.dtype("<" + STATE_ARRAYS[name]), count=n).copy()
