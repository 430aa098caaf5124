"""Host-side mirror of the reference's integrator interface over the C ABI.

The reference drives its path tracer through seven free functions and a struct
of public tunables (src/integrator/integrator.hpp:51-60, src/integrator/basic.hpp
:6-32, called from src/application/application.cpp:100-124).  This module keeps
those names and argument meanings — CreateSampleBuffer, CreateBasicRenderer,
ResetBasicRenderer, RunBasicRenderer, RenderSampleBuffer, and a BasicRenderer
whose CameraIndex / RenderFlags / PathLengthLimit / PathTerminationProbability
fields the caller pokes before Reset/Run — and forwards them to libptb.so
(include/ptb.h), the hand-written sm_100a CUDA implementation.

There is no CPU fallback: if libptb.so cannot be built/loaded, or no B200 is
visible, the calls raise.
"""
import ctypes as C
import os

import numpy as np

from . import _build
from .types import (KernelTimes, PreviewParams, PREVIEW_MODES, RenderParams, ResolveParams, SceneView, Stats, WideInfo, HIT_DTYPE, RAY_DTYPE,
                    RENDER_FLAG_ACCUMULATE, RENDER_FLAG_SAMPLE_JITTER, INTEGRATOR_WAVEFRONT, INTEGRATOR_MEGAKERNEL,
                    TRAVERSAL_BVH2_REFERENCE, TRAVERSAL_WIDE, TIE_FAITHFUL, TIE_CANONICAL, SCENE_DIRTY_ALL, BUILDER_SAH_HOST, BUILDER_LBVH_GPU, BUILDER_PLOC_GPU, BUILDER_AUTO,
                    TRACE_VARIANT_SIMPLE, TRACE_VARIANT_DEFAULT, TRACE_VARIANT_MAX, SCATTER_AUTO, SCATTER_DIRECT, SCATTER_QUEUES, SCATTER_DIRECT_STAGED,
                    TONE_CLAMP, TONE_REINHARD, TONE_HABLE, TONE_ACES)

# every symbol include/ptb.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "ptb_last_error", "ptb_version",
    "ptb_context_create", "ptb_context_destroy", "ptb_context_sync", "ptb_context_stream",
    "ptb_scene_create", "ptb_scene_update", "ptb_scene_destroy", "ptb_scene_wide_info", "ptb_scene_set_builder", "ptb_scene_wide_download",
    "ptb_film_create", "ptb_film_destroy", "ptb_film_device_accum", "ptb_film_download_accum", "ptb_film_upload_accum",
    "ptb_film_resolve", "ptb_film_resolve_device",
    "ptb_renderer_create", "ptb_renderer_destroy", "ptb_renderer_set_params", "ptb_renderer_set_frame_index",
    "ptb_renderer_reset", "ptb_renderer_run", "ptb_renderer_run_frames", "ptb_renderer_profile_frames", "ptb_renderer_sync", "ptb_renderer_stats",
    "ptb_renderer_trace_stats",
    "ptb_renderer_download_state", "ptb_renderer_slot_count", "ptb_raycast", "ptb_raycast_device", "ptb_preview_render",
    "ptb_raycast_noalloc", "ptb_raycast_noalloc_max_rays",
    "ptb_spectrum_table_build", "ptb_spectrum_upsample",
    "ptb_group_unique_id", "ptb_group_create", "ptb_group_create_local", "ptb_group_destroy", "ptb_group_rank", "ptb_group_world",
    "ptb_group_barrier", "ptb_group_nccl_version", "ptb_group_first_frame", "ptb_scene_replicate", "ptb_film_reduce",
    "ptb_film_reduce_resolve", "ptb_group_last_reduce",
]

REDUCE_NCCL, REDUCE_NCCL_SLICED, REDUCE_P2P_FUSED = 0, 1, 2

STATE_ARRAYS = {
    "OriginX": "f4", "OriginY": "f4", "OriginZ": "f4", "Duration": "f4", "PackedVelocity": "u4",
    "Time": "f4", "TextureU": "f4", "TextureV": "f4", "ShapeAndMaterialIndex": "u4", "PackedNormal": "u4", "PackedTangentX": "u4",
    "ImagePosition": "i4", "NormalizedLambda0": "f4", "Throughput0": "f4", "Throughput1": "f4", "Throughput2": "f4", "Throughput3": "f4",
    "Probability0": "f4", "Probability1": "f4", "Probability2": "f4", "Probability3": "f4", "SampleR": "f4", "SampleG": "f4", "SampleB": "f4",
    "ActiveShapeIndex01": "u4", "ActiveShapeIndex23": "u4", "PathLength": "u4", "PrimitiveIndex": "u4", "ShapeType": "u4",
    "PrimitiveCoordinateX": "f4", "PrimitiveCoordinateY": "f4", "PrimitiveCoordinateZ": "f4",
}

_lib = None


class PtbError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("ptb error %d: %s" % (code, message))
        self.code = code


def lib():
    """Loads libptb.so (building it in-tree with nvcc when missing or stale)."""
    global _lib
    if _lib is None:
        # PTB_LIB: load another in-tree build of the same sources (A/B experiments with -D tuning macros)
        L = C.CDLL(os.environ.get("PTB_LIB") or _build.build_cuda())
        vp, u32, sz = C.c_void_p, C.c_uint32, C.c_size_t
        L.ptb_last_error.restype = C.c_char_p
        L.ptb_version.restype = C.c_char_p
        L.ptb_context_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.ptb_context_destroy.argtypes = [vp]
        L.ptb_context_destroy.restype = None
        L.ptb_context_sync.argtypes = [vp]
        L.ptb_context_stream.argtypes = [vp]
        L.ptb_context_stream.restype = vp
        L.ptb_scene_create.argtypes = [vp, C.POINTER(vp)]
        L.ptb_scene_update.argtypes = [vp, C.POINTER(SceneView), u32]
        L.ptb_scene_destroy.argtypes = [vp]
        L.ptb_scene_destroy.restype = None
        L.ptb_scene_wide_info.argtypes = [vp, C.POINTER(WideInfo)]
        L.ptb_scene_set_builder.argtypes = [vp, u32]
        L.ptb_scene_wide_download.argtypes = [vp, vp, sz, vp, sz]
        L.ptb_spectrum_table_build.argtypes = [vp, vp, sz, C.POINTER(C.c_float)]
        L.ptb_spectrum_upsample.argtypes = [vp, vp, sz, vp, vp, sz, u32]
        L.ptb_film_create.argtypes = [vp, u32, u32, vp, C.POINTER(vp)]
        L.ptb_film_destroy.argtypes = [vp]
        L.ptb_film_destroy.restype = None
        L.ptb_film_device_accum.argtypes = [vp]
        L.ptb_film_device_accum.restype = vp
        L.ptb_film_download_accum.argtypes = [vp, vp, sz]
        L.ptb_film_upload_accum.argtypes = [vp, vp, sz]
        L.ptb_film_resolve.argtypes = [vp, C.POINTER(ResolveParams), vp, vp]
        L.ptb_film_resolve_device.argtypes = [vp, C.POINTER(ResolveParams), vp]
        L.ptb_renderer_create.argtypes = [vp, vp, vp, C.POINTER(vp)]
        L.ptb_renderer_destroy.argtypes = [vp]
        L.ptb_renderer_destroy.restype = None
        L.ptb_renderer_set_params.argtypes = [vp, C.POINTER(RenderParams)]
        L.ptb_renderer_set_frame_index.argtypes = [vp, u32]
        L.ptb_renderer_reset.argtypes = [vp]
        L.ptb_renderer_run.argtypes = [vp, u32]
        L.ptb_renderer_run_frames.argtypes = [vp, u32]
        L.ptb_renderer_profile_frames.argtypes = [vp, u32, C.POINTER(KernelTimes)]
        L.ptb_renderer_sync.argtypes = [vp]
        L.ptb_renderer_trace_stats.argtypes = [vp, C.POINTER(TraceStats)]
        L.ptb_renderer_stats.argtypes = [vp, C.POINTER(Stats)]
        L.ptb_renderer_download_state.argtypes = [vp, C.c_char_p, vp, sz]
        L.ptb_renderer_slot_count.argtypes = [vp]
        L.ptb_renderer_slot_count.restype = u32
        L.ptb_raycast.argtypes = [vp, vp, sz, u32, u32, vp]
        L.ptb_raycast_device.argtypes = [vp, vp, sz, u32, u32, vp]
        L.ptb_raycast_noalloc.argtypes = [vp, vp, sz, u32, u32, vp]
        L.ptb_raycast_noalloc_max_rays.argtypes = []
        L.ptb_raycast_noalloc_max_rays.restype = sz
        L.ptb_preview_render.argtypes = [vp, C.POINTER(PreviewParams), vp, vp, vp]
        L.ptb_group_unique_id.argtypes = [vp]
        L.ptb_group_create.argtypes = [vp, vp, C.c_int, C.c_int, C.POINTER(vp)]
        L.ptb_group_create_local.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(vp)]
        L.ptb_group_destroy.argtypes = [vp]
        L.ptb_group_destroy.restype = None
        L.ptb_group_rank.argtypes = [vp]
        L.ptb_group_world.argtypes = [vp]
        L.ptb_group_barrier.argtypes = [vp]
        L.ptb_group_nccl_version.argtypes = [C.POINTER(C.c_int)]
        L.ptb_group_first_frame.argtypes = [C.c_int, C.c_int, u32, u32]
        L.ptb_group_first_frame.restype = u32
        L.ptb_scene_replicate.argtypes = [vp, vp, C.c_int]
        L.ptb_film_reduce.argtypes = [vp, vp, C.c_int, vp]
        L.ptb_film_reduce_resolve.argtypes = [vp, vp, C.c_int, u32, C.POINTER(ResolveParams), vp, vp]
        L.ptb_group_last_reduce.argtypes = [vp, C.POINTER(ReduceTimes)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise PtbError(rc, lib().ptb_last_error().decode(errors="replace"))


class Context:
    """One CUDA device + stream (stands in for the `vulkan*` every reference call takes)."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        _check(lib().ptb_context_create(device, C.byref(self.h)))
        self.device = device

    def sync(self):
        _check(lib().ptb_context_sync(self.h))

    @property
    def stream(self):
        return lib().ptb_context_stream(self.h)

    def close(self):
        if self.h:
            lib().ptb_context_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


SPECTRUM_TABLE_SHAPE = (3, 64, 64, 64, 3)      # float[3][SCALE][COLOR][COLOR][3], spectrum.cpp:413-437


def build_spectrum_table(ctx, save_path=None):
    """BuildParametricSpectrumTableForSRGB on the GPU (spectrum.cpp:365-411).  Returns (table, device_ms);
    with save_path the table is written in the reference's raw file format."""
    table = np.zeros(SPECTRUM_TABLE_SHAPE, dtype=np.float32)
    ms = C.c_float(0.0)
    _check(lib().ptb_spectrum_table_build(ctx.h, table.ctypes.data, table.size, C.byref(ms)))
    if save_path:
        table.tofile(save_path)
    return table, float(ms.value)


def upsample_texels(ctx, table, rgba, texture_type):
    """The atlas packer's texel conversion (scene.cpp:1187-1212) on the GPU; rgba is (..., 4) float32."""
    table = np.ascontiguousarray(table, dtype=np.float32)
    src = np.ascontiguousarray(rgba, dtype=np.float32)
    out = np.empty_like(src)
    _check(lib().ptb_spectrum_upsample(ctx.h, table.ctypes.data, table.size, src.ctypes.data, out.ctypes.data, src.size // 4, texture_type))
    return out


class TraceStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "node_steps", "prim_tests", "warp_iterations", "flushes")] + \
               [(n, C.c_uint32) for n in ("node_bytes", "prim_bytes", "state_bytes", "_pad")]


class ReduceTimes(C.Structure):
    _fields_ = [("exchange_ms", C.c_float), ("mode", C.c_uint32), ("bytes_in", C.c_uint64)]


class Group:
    """Communicator of the sample-split multi-GPU job (include/ptb.h "groups"; NCCL inside libptb.so).

    One process per GPU: rank 0 calls Group.unique_id(), the launcher hands the 128 bytes to the other ranks
    (torch.distributed, MPI, a pipe ...), every rank constructs Group(ctx, id, rank, world).  Every method is
    collective: all ranks call it in the same order."""

    def __init__(self, ctx, group_id, rank, world):
        self.ctx = ctx
        self.h = C.c_void_p()
        buf = C.create_string_buffer(bytes(group_id), 128)
        _check(lib().ptb_group_create(ctx.h, buf, rank, world, C.byref(self.h)))
        self.rank, self.world = rank, world

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        _check(lib().ptb_group_unique_id(buf))
        return bytes(buf.raw)

    @staticmethod
    def first_frame(rank, world, frames_per_rank, base=0):
        """First FrameIndex (= RNG seed, basic.cpp:315,323) of `rank`: disjoint, gap-free seed blocks."""
        return lib().ptb_group_first_frame(rank, world, frames_per_rank, base)

    def barrier(self):
        _check(lib().ptb_group_barrier(self.h))

    def replicate_scene(self, device_scene, root=0):
        """The root's flattened device scene, device-to-device to every other rank (which passes an empty DeviceScene)."""
        _check(lib().ptb_scene_replicate(device_scene.h, self.h, root))

    def reduce(self, film, root=0, into=None):
        """Sum of all ranks' films into `into` on root (root < 0: everywhere); in place when `into` is the film itself."""
        _check(lib().ptb_film_reduce(film.h, self.h, root, into.h if into is not None else None))

    def reduce_resolve(self, film, params, mode=REDUCE_P2P_FUSED, root=0, into=None, out_srgb8=None):
        """Exchange + RenderSampleBuffer of the combined film; the sRGB8 frame arrives in out_srgb8 on the root."""
        p = ResolveParams(*params)
        if self.rank == root and out_srgb8 is None:
            out_srgb8 = np.empty((film.Height, film.Width, 4), dtype=np.uint8)
        _check(lib().ptb_film_reduce_resolve(film.h, self.h, root, mode, C.byref(p), into.h if into is not None else None,
                                             out_srgb8.ctypes.data if (self.rank == root and out_srgb8 is not None) else None))
        return out_srgb8 if self.rank == root else None

    def last_reduce(self):
        t = ReduceTimes()
        _check(lib().ptb_group_last_reduce(self.h, C.byref(t)))
        return {"exchange_ms": float(t.exchange_ms), "mode": int(t.mode), "bytes_in": int(t.bytes_in)}

    def close(self):
        if self.h:
            lib().ptb_group_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_version():
    v = C.c_int(0)
    _check(lib().ptb_group_nccl_version(C.byref(v)))
    return v.value


class DeviceScene:
    """CreateVulkanScene / UpdateVulkanScene / DestroyVulkanScene (scene.hpp:440-442)."""

    def __init__(self, ctx, host_scene=None, builder=None):
        self.ctx = ctx
        self.h = C.c_void_p()
        _check(lib().ptb_scene_create(ctx.h, C.byref(self.h)))
        if builder is not None:
            self.set_builder(builder)
        if host_scene is not None:
            self.update(host_scene, SCENE_DIRTY_ALL)

    def update(self, host_scene, dirty_flags=SCENE_DIRTY_ALL):
        """host_scene: path_tracer_b200.scene.Scene (packed) or a raw SceneView."""
        view = host_scene if isinstance(host_scene, SceneView) else host_scene.view
        _check(lib().ptb_scene_update(self.h, C.byref(view), dirty_flags))

    def set_builder(self, builder):
        """BUILDER_SAH_HOST (default), BUILDER_LBVH_GPU or BUILDER_PLOC_GPU: who flattens the scene at the next update."""
        _check(lib().ptb_scene_set_builder(self.h, builder))

    def wide_download(self):
        """(nodes[N, 20] uint32, prims[P, 12] float32) of the flattened tree, for structural checks."""
        info = self.wide_info()
        nodes = np.zeros((info["node_count"], 20), dtype=np.uint32)
        prims = np.zeros((info["prim_count"], 12), dtype=np.float32)
        _check(lib().ptb_scene_wide_download(self.h, nodes.ctypes.data, nodes.size, prims.ctypes.data, prims.size))
        return nodes, prims

    def wide_info(self):
        w = WideInfo()
        _check(lib().ptb_scene_wide_info(self.h, C.byref(w)))
        return {n: getattr(w, n) for n, _ in WideInfo._fields_}

    def raycast(self, rays, traversal=TRAVERSAL_WIDE, tie=TIE_CANONICAL):
        """Trace() on caller rays (host arrays in, host hits out)."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        out = np.zeros(len(rays), dtype=HIT_DTYPE)
        _check(lib().ptb_raycast(self.h, rays.ctypes.data, len(rays), traversal, tie, out.ctypes.data))
        return out

    def raycast_noalloc(self, rays, traversal=TRAVERSAL_WIDE, tie=TIE_CANONICAL):
        """ptb_raycast_noalloc: the same query through the library's static device scratch (no cudaMalloc anywhere on the
        path; at most ptb_raycast_noalloc_max_rays() rays per call; calls on one device must not overlap)."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        out = np.zeros(len(rays), dtype=HIT_DTYPE)
        _check(lib().ptb_raycast_noalloc(self.h, rays.ctypes.data, len(rays), traversal, tie, out.ctypes.data))
        return out

    def preview(self, camera_to, camera_from, mode, width, height, brightness=1.0, selected=0xFFFFFFFF, mouse=(0, 0),
                traversal=TRAVERSAL_WIDE):
        """The reference's preview visualiser (preview_render.glsl:96-178) as a CUDA kernel: returns the (H, W, 4)
        float image, the (H, W) shape indices and the shape under the mouse pixel (the pick query)."""
        p = PreviewParams()
        p.To[:] = [float(x) for x in camera_to]
        p.From[:] = [float(x) for x in camera_from]
        p.RenderMode, p.Brightness, p.SelectedShapeIndex = mode, brightness, selected
        p.RenderSizeX, p.RenderSizeY, p.MouseX, p.MouseY, p.Traversal = width, height, mouse[0], mouse[1], traversal
        img = np.empty((height, width, 4), dtype=np.float32)
        shapes = np.empty((height, width), dtype=np.uint32)
        pick = C.c_uint32(0xFFFFFFFF)
        _check(lib().ptb_preview_render(self.h, C.byref(p), img.ctypes.data, shapes.ctypes.data, C.addressof(pick)))
        return img, shapes, pick.value

    def raycast_device(self, d_rays, n, d_hits, traversal=TRAVERSAL_WIDE, tie=TIE_CANONICAL):
        """Same on device-resident buffers (raw pointers); asynchronous on the context stream."""
        _check(lib().ptb_raycast_device(self.h, d_rays, n, traversal, tie, d_hits))

    def close(self):
        if self.h:
            lib().ptb_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SampleBuffer:
    """vulkan_sample_buffer (integrator.hpp:21-41): the RGBA32F accumulator, headless."""

    def __init__(self, ctx, width, height, external_device_ptr=None):
        self.ctx, self.Width, self.Height = ctx, width, height
        self.h = C.c_void_p()
        _check(lib().ptb_film_create(ctx.h, width, height, external_device_ptr, C.byref(self.h)))

    @property
    def device_ptr(self):
        return lib().ptb_film_device_accum(self.h)

    def download(self):
        a = np.empty((self.Height, self.Width, 4), dtype=np.float32)
        _check(lib().ptb_film_download_accum(self.h, a.ctypes.data, a.size))
        return a

    def upload(self, accum):
        a = np.ascontiguousarray(accum, dtype=np.float32)
        _check(lib().ptb_film_upload_accum(self.h, a.ctypes.data, a.size))

    def resolve(self, Brightness=1.0, ToneMappingMode=TONE_CLAMP, ToneMappingWhiteLevel=1.0, want_linear=True, want_srgb8=True,
                out_srgb8=None, out_linear=None):
        """RenderSampleBuffer (integrator.cpp:110-159) into host arrays.  `out_*`: caller-owned destinations
        (e.g. pinned memory, so the device-to-host copy is one DMA); allocated here when omitted."""
        p = ResolveParams(Brightness, ToneMappingMode, ToneMappingWhiteLevel)
        lin = srgb = None
        if want_linear:
            lin = out_linear if out_linear is not None else np.empty((self.Height, self.Width, 3), dtype=np.float32)
            assert lin.dtype == np.float32 and lin.size == self.Height * self.Width * 3 and lin.flags["C_CONTIGUOUS"]
        if want_srgb8:
            srgb = out_srgb8 if out_srgb8 is not None else np.empty((self.Height, self.Width, 4), dtype=np.uint8)
            assert srgb.dtype == np.uint8 and srgb.size == self.Height * self.Width * 4 and srgb.flags["C_CONTIGUOUS"]
        _check(lib().ptb_film_resolve(self.h, C.byref(p), srgb.ctypes.data if want_srgb8 else None, lin.ctypes.data if want_linear else None))
        return lin, srgb

    def close(self):
        if self.h:
            lib().ptb_film_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BasicRenderer:
    """basic_renderer (basic.hpp:6-26).  The five public fields are poked by the
    caller before Reset/Run, as the reference application does (application.cpp:104-107)."""

    def __init__(self, ctx, scene, sample_buffer):
        self.ctx, self.Scene, self.SampleBuffer = ctx, scene, sample_buffer
        self.CameraIndex = 0
        self.RenderFlags = RENDER_FLAG_ACCUMULATE | RENDER_FLAG_SAMPLE_JITTER
        self.PathLengthLimit = 0
        self.PathTerminationProbability = 0.0
        # the build's own knobs
        self.Integrator = INTEGRATOR_WAVEFRONT
        self.Traversal = TRAVERSAL_WIDE
        self.TieMode = TIE_CANONICAL
        self.TraceVariant = TRACE_VARIANT_DEFAULT
        self.ScatterMode = SCATTER_AUTO
        self.Pipelines = 0
        self.h = C.c_void_p()
        _check(lib().ptb_renderer_create(ctx.h, scene.h, sample_buffer.h, C.byref(self.h)))

    def _push(self):
        p = RenderParams(self.CameraIndex, self.RenderFlags, self.PathLengthLimit, self.PathTerminationProbability,
                         self.Integrator, self.Traversal, self.TieMode, self.TraceVariant, self.ScatterMode, self.Pipelines)
        _check(lib().ptb_renderer_set_params(self.h, C.byref(p)))

    def set_frame_index(self, f):
        _check(lib().ptb_renderer_set_frame_index(self.h, f))

    def reset(self):
        self._push()
        _check(lib().ptb_renderer_reset(self.h))

    def run(self, rounds=1):
        self._push()
        _check(lib().ptb_renderer_run(self.h, rounds))

    def run_frames(self, frames):
        self._push()
        _check(lib().ptb_renderer_run_frames(self.h, frames))

    def profile_frames(self, frames):
        """run_frames with CUDA events around every kernel: summed device ms per kernel."""
        self._push()
        t = KernelTimes()
        _check(lib().ptb_renderer_profile_frames(self.h, frames, C.byref(t)))
        return {n: getattr(t, n) for n, _ in KernelTimes._fields_}

    def sync(self):
        _check(lib().ptb_renderer_sync(self.h))

    def trace_stats(self):
        """One instrumented launch of the default trace kernel over the current rays: what its own layout fetches."""
        self._push()
        t = TraceStats()
        _check(lib().ptb_renderer_trace_stats(self.h, C.byref(t)))
        return {n: int(getattr(t, n)) for n, _ in TraceStats._fields_ if n != "_pad"}

    def stats(self):
        s = Stats()
        _check(lib().ptb_renderer_stats(self.h, C.byref(s)))
        return {n: getattr(s, n) for n, _ in Stats._fields_}

    @property
    def slot_count(self):
        return lib().ptb_renderer_slot_count(self.h)

    def state(self, name):
        out = np.empty(self.slot_count, dtype=np.dtype("<" + STATE_ARRAYS[name]))
        _check(lib().ptb_renderer_download_state(self.h, name.encode(), out.ctypes.data, out.nbytes))
        return out

    def close(self):
        if self.h:
            lib().ptb_renderer_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- the reference's free functions (same names, `vulkan*` -> Context) ----
def CreateSampleBuffer(ctx, Width, Height):              # integrator.hpp:51
    return SampleBuffer(ctx, Width, Height)


def DestroySampleBuffer(ctx, SampleBuffer_):             # integrator.hpp:53
    SampleBuffer_.close()


def RenderSampleBuffer(ctx, SampleBuffer_, Parameters):  # integrator.hpp:55-60; Parameters = (Brightness, ToneMappingMode, ToneMappingWhiteLevel)
    return SampleBuffer_.resolve(*Parameters)


def CreateBasicRenderer(ctx, Scene, SampleBuffer_):      # basic.hpp:28
    return BasicRenderer(ctx, Scene, SampleBuffer_)


def DestroyBasicRenderer(ctx, Renderer):                 # basic.hpp:29
    Renderer.close()


def ResetBasicRenderer(ctx, Renderer):                   # basic.hpp:31
    Renderer.reset()


def RunBasicRenderer(ctx, Renderer, Rounds):             # basic.hpp:32
    Renderer.run(Rounds)
