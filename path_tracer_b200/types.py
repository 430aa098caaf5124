"""ctypes mirrors of the PODs declared in include/ptb.h."""
import ctypes as C

import numpy as np

u32 = C.c_uint32


class SceneView(C.Structure):
    """ptb_scene_view: {pointer, count} pairs over the packed scene arrays."""
    _fields_ = [
        ("Textures", C.c_void_p), ("TextureCount", u32), ("_r0", u32),
        ("MaterialData", C.c_void_p), ("MaterialWordCount", u32), ("_r1", u32),
        ("Shapes", C.c_void_p), ("ShapeCount", u32), ("_r2", u32),
        ("ShapeNodes", C.c_void_p), ("ShapeNodeCount", u32), ("_r3", u32),
        ("MeshFaces", C.c_void_p), ("MeshFaceCount", u32), ("_r4", u32),
        ("MeshVertices", C.c_void_p), ("MeshVertexCount", u32), ("_r5", u32),
        ("MeshNodes", C.c_void_p), ("MeshNodeCount", u32), ("_r6", u32),
        ("Cameras", C.c_void_p), ("CameraCount", u32), ("_r7", u32),
        ("AtlasLayers", C.c_void_p), ("AtlasLayerCount", u32),
        ("AtlasWidth", u32),
        ("AtlasHeight", u32), ("_r8", u32),
        ("Globals", C.c_void_p),
    ]


class ResolveParams(C.Structure):
    _fields_ = [("Brightness", C.c_float), ("ToneMappingMode", u32), ("ToneMappingWhiteLevel", C.c_float)]


class PreviewParams(C.Structure):
    """ptb_preview_params = the push constants of preview_render.glsl:48-60 (+ traversal)."""
    _fields_ = [("To", C.c_float * 16), ("From", C.c_float * 16), ("RenderMode", u32), ("Brightness", C.c_float),
                ("SelectedShapeIndex", u32), ("RenderSizeX", u32), ("RenderSizeY", u32), ("MouseX", u32), ("MouseY", u32),
                ("Traversal", u32)]


PREVIEW_MODES = ("BASE_COLOR", "BASE_COLOR_SHADED", "NORMAL", "MATERIAL_INDEX", "PRIMITIVE_INDEX", "MESH_COMPLEXITY", "SCENE_COMPLEXITY")


class RenderParams(C.Structure):
    _fields_ = [
        ("CameraIndex", u32), ("RenderFlags", u32), ("PathLengthLimit", u32),
        ("PathTerminationProbability", C.c_float),
        ("Integrator", u32), ("Traversal", u32), ("TieMode", u32), ("TraceVariant", u32),
        ("ScatterMode", u32), ("Pipelines", u32), ("Reserved", u32 * 2),
    ]


class Stats(C.Structure):
    _fields_ = [("rays_traced", C.c_uint64), ("paths_finished", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("frame_index", C.c_uint64)]


class KernelTimes(C.Structure):
    _fields_ = [("trace_ms", C.c_float), ("scatter_ms", C.c_float), ("mega_ms", C.c_float),
                ("trace_launches", u32), ("scatter_launches", u32), ("mega_launches", u32)]


class WideInfo(C.Structure):
    _fields_ = [("status", C.c_int32), ("node_count", u32), ("prim_count", u32), ("plane_count", u32), ("depth", u32),
                ("bvh2_depth", u32), ("sah_cost", C.c_double), ("build_seconds", C.c_double), ("builder", u32), ("device_build_ms", C.c_float)]


BUILDER_SAH_HOST, BUILDER_LBVH_GPU, BUILDER_PLOC_GPU, BUILDER_AUTO = 0, 1, 2, 3


# numpy record layouts of the packed arrays (src/scene/scene.hpp:82-173)
RAY_DTYPE = np.dtype([("o", "<f4", 3), ("d", "<f4", 3), ("duration", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("shape", "<u4"), ("prim", "<u4"), ("material", "<u4"), ("shape_type", "<u4"),
                      ("n", "<f4", 3), ("tan", "<f4", 3), ("uv", "<f4", 2), ("bary", "<f4", 3)])
SHAPE_DTYPE = np.dtype([("Type", "<i4"), ("MaterialIndex", "<u4"), ("MeshRootNodeIndex", "<u4"), ("Pad0", "<u4"),
                        ("To", "<f4", 16), ("From", "<f4", 16)])
SHAPE_NODE_DTYPE = np.dtype([("Minimum", "<f4", 3), ("ChildNodeIndices", "<u4"), ("Maximum", "<f4", 3), ("ShapeIndex", "<u4")])
MESH_FACE_DTYPE = np.dtype([("Position0", "<f4", 3), ("VertexIndex0", "<u4"), ("Position1", "<f4", 3), ("VertexIndex1", "<u4"),
                            ("Position2", "<f4", 3), ("VertexIndex2", "<u4")])
MESH_VERTEX_DTYPE = np.dtype([("PackedNormal", "<u4"), ("PackedUV", "<u4")])
MESH_NODE_DTYPE = np.dtype([("Minimum", "<f4", 3), ("FaceBeginOrNodeIndex", "<u4"), ("Maximum", "<f4", 3), ("FaceEndIndex", "<u4")])
TEXTURE_DTYPE = np.dtype([("AtlasPlacementMinimum", "<f4", 2), ("AtlasPlacementMaximum", "<f4", 2),
                          ("AtlasImageIndex", "<u4"), ("Type", "<u4"), ("Flags", "<u4"), ("Unused0", "<u4")])
CAMERA_DTYPE = np.dtype([("Model", "<u4"), ("FocalLength", "<f4"), ("ApertureRadius", "<f4"), ("SensorDistance", "<f4"),
                         ("SensorSize", "<f4", 2), ("Pad", "<u4", 2), ("To", "<f4", 16), ("From", "<f4", 16)])
GLOBALS_DTYPE = np.dtype([("SkyboxMeanDirection", "<f4", 3), ("SkyboxConcentration", "<f4"),
                          ("SkyboxSamplingProbability", "<f4"), ("SkyboxBrightness", "<f4"),
                          ("SkyboxTextureIndex", "<u4"), ("ShapeCount", "<u4"), ("SceneScatterRate", "<f4"), ("Pad", "<u4", 3)])

assert SHAPE_DTYPE.itemsize == 144 and SHAPE_NODE_DTYPE.itemsize == 32 and MESH_FACE_DTYPE.itemsize == 48
assert MESH_VERTEX_DTYPE.itemsize == 8 and MESH_NODE_DTYPE.itemsize == 32 and TEXTURE_DTYPE.itemsize == 32
assert CAMERA_DTYPE.itemsize == 160 and GLOBALS_DTYPE.itemsize == 48
assert RAY_DTYPE.itemsize == 28 and HIT_DTYPE.itemsize == 64

# enums (include/ptb.h)
RENDER_FLAG_ACCUMULATE = 1
RENDER_FLAG_SAMPLE_JITTER = 2
TONE_CLAMP, TONE_REINHARD, TONE_HABLE, TONE_ACES = 0, 1, 2, 3
INTEGRATOR_WAVEFRONT, INTEGRATOR_MEGAKERNEL = 0, 1
TRAVERSAL_BVH2_REFERENCE, TRAVERSAL_WIDE = 0, 1
TIE_FAITHFUL, TIE_CANONICAL = 0, 1
TRACE_VARIANT_SIMPLE, TRACE_VARIANT_DEFAULT, TRACE_VARIANT_MAX = 0, 14, 19
SCATTER_AUTO, SCATTER_DIRECT, SCATTER_QUEUES, SCATTER_DIRECT_STAGED = 0, 1, 2, 3
SCENE_DIRTY_ALL = 0xFFFFFFFF
