"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so (the CPU restatement of the reference
shaders).  Importable only from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; the product (path_tracer_b200/) never
imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")

TIE_FAITHFUL, TIE_CANONICAL = 0, 1

RAY_DTYPE = np.dtype([("o", "<f4", 3), ("d", "<f4", 3), ("duration", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("shape", "<u4"), ("prim", "<u4"), ("material", "<u4"), ("shape_type", "<u4"),
                      ("n", "<f4", 3), ("tan", "<f4", 3), ("uv", "<f4", 2), ("bary", "<f4", 3),
                      ("scene_complexity", "<u4"), ("mesh_complexity", "<u4"),
                      ("node_fetches", "<u4"), ("face_tests", "<u4"), ("shape_visits", "<u4")])
assert HIT_DTYPE.itemsize == 84


class RenderParams(C.Structure):
    _fields_ = [("camera_index", C.c_uint32), ("render_flags", C.c_uint32), ("path_length_limit", C.c_uint32),
                ("path_termination_probability", C.c_float), ("tie_mode", C.c_int), ("threads", C.c_int),
                ("tile_stride", C.c_uint32)]


class PreviewParams(C.Structure):
    """preview_render.glsl:48-60 push constants (+ tie rule); same bytes as ptb_preview_params up to `Tie`."""
    _fields_ = [("To", C.c_float * 16), ("From", C.c_float * 16), ("RenderMode", C.c_uint32), ("Brightness", C.c_float),
                ("SelectedShapeIndex", C.c_uint32), ("RenderSizeX", C.c_uint32), ("RenderSizeY", C.c_uint32),
                ("MouseX", C.c_uint32), ("MouseY", C.c_uint32), ("Tie", C.c_uint32)]


class OStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "paths", "node_fetches", "face_tests", "shape_visits", "mesh_hits")]


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".h", ".cpp", ".inc"))]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        r = subprocess.run(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("oracle build failed:\n" + r.stdout)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        vp, u32, u64, f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_float
        f32p = C.POINTER(C.c_float)
        L.orc_raycast.argtypes = [vp, vp, u64, C.c_int, vp, C.c_int]
        L.orc_raycast_bruteforce.argtypes = [vp, vp, u64, vp, C.c_int]
        L.orc_renderer_create.restype = vp
        L.orc_renderer_create.argtypes = [u32, u32]
        L.orc_renderer_destroy.argtypes = [vp]
        L.orc_renderer_set_frame_index.argtypes = [vp, u32]
        L.orc_renderer_frame_index.restype = u32
        L.orc_renderer_frame_index.argtypes = [vp]
        L.orc_renderer_reset.argtypes = [vp, vp, C.POINTER(RenderParams)]
        L.orc_renderer_run.argtypes = [vp, vp, C.POINTER(RenderParams), u32]
        L.orc_renderer_trace_only.argtypes = [vp, vp, C.POINTER(RenderParams)]
        L.orc_renderer_slot_count.restype = u32
        L.orc_renderer_slot_count.argtypes = [vp]
        L.orc_renderer_accum.restype = vp
        L.orc_renderer_accum.argtypes = [vp]
        L.orc_renderer_stats.argtypes = [vp, C.POINTER(OStats)]
        L.orc_renderer_state_array.restype = vp
        L.orc_renderer_state_array.argtypes = [vp, C.c_char_p]
        L.orc_resolve.argtypes = [vp, u32, u32, f32, u32, f32, vp, vp]
        L.orc_kat_rng.argtypes = [u32, u32, vp, vp]
        L.orc_kat_pack_unit_vector.restype = u32
        L.orc_kat_pack_unit_vector.argtypes = [f32, f32, f32]
        L.orc_kat_unpack_unit_vector.argtypes = [u32, f32p]
        L.orc_kat_pack_half2.restype = u32
        L.orc_kat_pack_half2.argtypes = [f32, f32]
        L.orc_kat_unpack_half2.argtypes = [u32, f32p]
        L.orc_kat_observer.argtypes = [f32, f32p]
        L.orc_kat_parametric_spectrum.restype = f32
        L.orc_kat_parametric_spectrum.argtypes = [f32, f32, f32, f32]
        L.orc_kat_observe_d65.argtypes = [f32, f32, f32, f32p]
        L.orc_kat_tonemap.argtypes = [f32, f32, f32, u32, f32, f32p]
        L.orc_kat_box.restype = f32
        L.orc_kat_box.argtypes = [vp, f32, f32p, f32p]
        L.orc_kat_camera_ray.argtypes = [vp, f32, f32, u32, vp]
        L.orc_kat_srgb8.restype = C.c_uint8
        L.orc_kat_srgb8.argtypes = [f32]
        L.orc_preview.argtypes = [vp, C.POINTER(PreviewParams), vp, vp, vp, C.c_int]
        _lib = L
    return _lib


def _view_ptr(view):
    """`view` is a ctypes ptb_scene_view (same bytes as orc::scene_view)."""
    assert C.sizeof(view) == lib().orc_sizeof_scene_view(), "scene_view layout mismatch"
    return C.addressof(view)


def make_rays(origins, directions, duration=1048576.0):
    r = np.zeros(len(origins), dtype=RAY_DTYPE)
    r["o"] = origins
    r["d"] = directions
    r["duration"] = duration
    return r


def raycast(view, rays, tie=TIE_FAITHFUL, threads=0):
    rays = np.ascontiguousarray(rays)
    out = np.zeros(len(rays), dtype=HIT_DTYPE)
    lib().orc_raycast(_view_ptr(view), rays.ctypes.data, len(rays), tie, out.ctypes.data, threads or os.cpu_count())
    return out


def raycast_bruteforce(view, rays, threads=0):
    rays = np.ascontiguousarray(rays)
    out = np.zeros(len(rays), dtype=HIT_DTYPE)
    lib().orc_raycast_bruteforce(_view_ptr(view), rays.ctypes.data, len(rays), out.ctypes.data, threads or os.cpu_count())
    return out


def preview(view, camera_to, camera_from, mode, width, height, brightness=1.0, selected=0xFFFFFFFF, mouse=(0, 0), tie=TIE_CANONICAL, threads=0):
    """preview_render.glsl on the CPU: (H, W, 4) float image, (H, W) shape indices, shape under the mouse pixel."""
    p = PreviewParams()
    p.To[:] = [float(x) for x in camera_to]
    p.From[:] = [float(x) for x in camera_from]
    p.RenderMode, p.Brightness, p.SelectedShapeIndex = mode, brightness, selected
    p.RenderSizeX, p.RenderSizeY, p.MouseX, p.MouseY, p.Tie = width, height, mouse[0], mouse[1], tie
    img = np.zeros((height, width, 4), dtype=np.float32)
    shapes = np.zeros((height, width), dtype=np.uint32)
    pick = C.c_uint32(0xFFFFFFFF)
    lib().orc_preview(_view_ptr(view), C.byref(p), img.ctypes.data, shapes.ctypes.data, C.addressof(pick), threads or os.cpu_count())
    return img, shapes, pick.value


STATE_ARRAYS = {
    "OriginX": "f4", "OriginY": "f4", "OriginZ": "f4", "Duration": "f4", "PackedVelocity": "u4",
    "Time": "f4", "TextureU": "f4", "TextureV": "f4", "ShapeAndMaterialIndex": "u4", "PackedNormal": "u4", "PackedTangentX": "u4",
    "ImagePosition": "i4", "NormalizedLambda0": "f4", "Throughput0": "f4", "Throughput1": "f4", "Throughput2": "f4", "Throughput3": "f4",
    "Probability0": "f4", "Probability1": "f4", "Probability2": "f4", "Probability3": "f4", "SampleR": "f4", "SampleG": "f4", "SampleB": "f4",
    "ActiveShapeIndex01": "u4", "ActiveShapeIndex23": "u4", "PathLength": "u4", "PrimitiveIndex": "u4",
}


class Renderer:
    """Reference-shaped wavefront renderer on the CPU: reset() / run(rounds)."""

    def __init__(self, view, width, height, camera_index=0, render_flags=3, path_length_limit=0,
                 path_termination_probability=0.0, tie=TIE_FAITHFUL, threads=0, tile_stride=1):
        self.L = lib()
        self.view = view
        self.w, self.h = width, height
        self.r = self.L.orc_renderer_create(width, height)
        self.p = RenderParams(camera_index, render_flags, path_length_limit, path_termination_probability, tie,
                              threads or os.cpu_count(), tile_stride)

    def close(self):
        if self.r:
            self.L.orc_renderer_destroy(self.r)
            self.r = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_frame_index(self, f):
        self.L.orc_renderer_set_frame_index(self.r, f)

    def reset(self):
        self.L.orc_renderer_reset(self.r, _view_ptr(self.view), C.byref(self.p))

    def run(self, rounds=1):
        self.L.orc_renderer_run(self.r, _view_ptr(self.view), C.byref(self.p), rounds)

    def run_frames(self, frames):
        for _ in range(frames):
            self.run(1)

    def trace_only(self):
        self.L.orc_renderer_trace_only(self.r, _view_ptr(self.view), C.byref(self.p))

    @property
    def slot_count(self):
        return self.L.orc_renderer_slot_count(self.r)

    def accum(self):
        ptr = self.L.orc_renderer_accum(self.r)
        buf = (C.c_float * (self.w * self.h * 4)).from_address(ptr)
        return np.frombuffer(buf, dtype=np.float32).reshape(self.h, self.w, 4).copy()

    def state(self, name):
        ptr = self.L.orc_renderer_state_array(self.r, name.encode())
        n = self.slot_count
        buf = (C.c_char * (n * 4)).from_address(ptr)
        return np.frombuffer(buf, dtype=np.dtype("<" + STATE_ARRAYS[name]), count=n).copy()

    def stats(self):
        s = OStats()
        self.L.orc_renderer_stats(self.r, C.byref(s))
        return {n: getattr(s, n) for n, _ in OStats._fields_}


def resolve(accum, brightness=1.0, mode=0, white=1.0):
    a = np.ascontiguousarray(accum, dtype=np.float32)
    h, w = a.shape[0], a.shape[1]
    lin = np.zeros((h, w, 3), dtype=np.float32)
    srgb = np.zeros((h, w, 4), dtype=np.uint8)
    lib().orc_resolve(a.ctypes.data, w, h, brightness, mode, white, lin.ctypes.data, srgb.ctypes.data)
    return lin, srgb
