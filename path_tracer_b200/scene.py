"""Python face of the host scene layer (libptb_scene.so).

Mirrors the reference's scene API (src/scene/scene.hpp:410-442): CreateScene,
CreateEntity, CreateMaterial, CreateCheckerTexture, PackSceneData — plus the
procedural scenes of the BASELINE.json configs.  CPU only.
"""
import ctypes as C
import os

import numpy as np

from . import _build
from .types import (SceneView, SHAPE_DTYPE, SHAPE_NODE_DTYPE, MESH_FACE_DTYPE, MESH_VERTEX_DTYPE,
                    MESH_NODE_DTYPE, TEXTURE_DTYPE, CAMERA_DTYPE, GLOBALS_DTYPE)

ENTITY_ROOT, ENTITY_CONTAINER, ENTITY_CAMERA, ENTITY_MESH_INSTANCE, ENTITY_PLANE, ENTITY_SPHERE, ENTITY_CUBE = range(7)
MATERIAL_BASIC_DIFFUSE, MATERIAL_BASIC_METAL, MATERIAL_BASIC_TRANSLUCENT, MATERIAL_OPENPBR = range(4)
TEXTURE_RAW, TEXTURE_REFLECTANCE_WITH_ALPHA, TEXTURE_RADIANCE = range(3)
CAMERA_PINHOLE, CAMERA_THIN_LENS, CAMERA_360 = range(3)
TEXTURE_TYPE_RAW, TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, TEXTURE_TYPE_RADIANCE = range(3)     # scene.hpp:10-16

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = _build.build_scene()
        L = C.CDLL(path)
        vp, u32, f32p = C.c_void_p, C.c_uint32, C.POINTER(C.c_float)
        L.ptbh_make_scene.restype = vp
        L.ptbh_make_scene.argtypes = [C.c_char_p, C.POINTER(C.c_double), C.c_int]
        L.ptbh_scene_create.restype = vp
        L.ptbh_scene_create.argtypes = [C.c_int]
        L.ptbh_scene_destroy.argtypes = [vp]
        L.ptbh_scene_set_atlas_size.argtypes = [vp, u32, u32]
        L.ptbh_scene_pack.restype = u32
        L.ptbh_scene_pack.argtypes = [vp]
        L.ptbh_scene_mark_dirty.argtypes = [vp, u32]
        L.ptbh_scene_view.argtypes = [vp, C.POINTER(SceneView)]
        L.ptbh_root_set.argtypes = [vp, C.c_float, C.c_float, C.c_float, vp]
        L.ptbh_create_entity.restype = vp
        L.ptbh_create_entity.argtypes = [vp, C.c_int, vp]
        L.ptbh_destroy_entity.argtypes = [vp, vp]
        L.ptbh_entity_set_transform.argtypes = [vp, vp, f32p, f32p, f32p]
        L.ptbh_entity_look_at.argtypes = [vp, vp, f32p, f32p]
        L.ptbh_entity_set_active.argtypes = [vp, vp, C.c_int]
        L.ptbh_entity_set_material.argtypes = [vp, vp, vp]
        L.ptbh_entity_set_mesh.argtypes = [vp, vp, vp]
        L.ptbh_entity_shape_index.restype = u32
        L.ptbh_entity_shape_index.argtypes = [vp]
        L.ptbh_entity_camera_index.restype = u32
        L.ptbh_entity_camera_index.argtypes = [vp]
        L.ptbh_camera_set.argtypes = [vp, vp, C.c_int, f32p]
        L.ptbh_create_material.restype = vp
        L.ptbh_create_material.argtypes = [vp, C.c_int, C.c_char_p]
        L.ptbh_material_set.argtypes = [vp, vp, C.c_char_p, f32p]
        L.ptbh_material_set_texture.argtypes = [vp, vp, C.c_char_p, vp]
        L.ptbh_material_slot.restype = u32
        L.ptbh_material_slot.argtypes = [vp]
        L.ptbh_create_texture.restype = vp
        L.ptbh_create_texture.argtypes = [vp, C.c_char_p, C.c_int, u32, u32, f32p, C.c_int]
        L.ptbh_create_checker_texture.restype = vp
        L.ptbh_create_checker_texture.argtypes = [vp, C.c_char_p, C.c_int, f32p, f32p, C.c_int]
        L.ptbh_create_mesh.restype = vp
        L.ptbh_create_mesh.argtypes = [vp, C.c_char_p, f32p, u32, C.POINTER(u32), u32]
        for n in ("ptbh_mesh_depth", "ptbh_mesh_node_count", "ptbh_mesh_face_count"):
            getattr(L, n).restype = u32
            getattr(L, n).argtypes = [vp]
        L.ptbh_scene_mesh_count.restype = u32
        L.ptbh_scene_mesh_count.argtypes = [vp]
        L.ptbh_scene_mesh.restype = vp
        L.ptbh_scene_mesh.argtypes = [vp, u32]
        L.ptbh_mesh_faces.argtypes = [vp, C.POINTER(u32)]
        L.ptbh_load_model.restype = vp
        L.ptbh_load_model.argtypes = [vp, C.c_char_p, C.c_char_p, vp, f32p]
        L.ptbh_create_prefab_instance.restype = vp
        L.ptbh_create_prefab_instance.argtypes = [vp, vp, vp]
        L.ptbh_destroy_prefab.argtypes = [vp, vp]
        L.ptbh_prefab_mesh_count.restype = u32
        L.ptbh_prefab_mesh_count.argtypes = [vp]
        L.ptbh_load_texture.restype = vp
        L.ptbh_load_texture.argtypes = [vp, C.c_char_p, C.c_int, C.c_char_p]
        L.ptbh_decode_image.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), vp, C.c_size_t]
        L.ptbh_write_png.argtypes = [C.c_char_p, u32, u32, vp]
        L.ptbh_write_pfm.argtypes = [C.c_char_p, u32, u32, vp]
        L.ptbh_mesh_vertex_count.restype = u32
        L.ptbh_mesh_vertex_count.argtypes = [vp]
        L.ptbh_mesh_vertices.argtypes = [vp, f32p]
        L.ptbh_last_error.restype = C.c_char_p
        L.ptbh_save_scene.argtypes = [vp, C.c_char_p]
        L.ptbh_load_scene.restype = vp
        L.ptbh_load_scene.argtypes = [C.c_char_p]
        L.ptbh_set_spectrum_table_path.argtypes = [C.c_char_p]
        L.ptbh_rgb_to_spectrum.argtypes = [f32p, f32p]
        L.ptbh_fit_spectrum.argtypes = [f32p, f32p]
        L.ptbh_sample_spectrum.restype = C.c_float
        L.ptbh_sample_spectrum.argtypes = [f32p, C.c_float]
        L.ptbh_make_transform.argtypes = [f32p, f32p, f32p, f32p]
        L.ptbh_inverse.argtypes = [f32p, f32p]
        L.ptbh_pack_unit_vector.restype = u32
        L.ptbh_pack_unit_vector.argtypes = [C.c_float] * 3
        L.ptbh_pack_half2.restype = u32
        L.ptbh_pack_half2.argtypes = [C.c_float] * 2
        os.makedirs(os.path.dirname(_build.SPECTRUM_TABLE), exist_ok=True)
        L.ptbh_set_spectrum_table_path(_build.SPECTRUM_TABLE.encode())
        _lib = L
    return _lib


def _f3(v):
    return (C.c_float * len(v))(*[float(x) for x in v])


def ensure_spectrum_table(ctx=None):
    """Loads the sRGB->spectrum table.  When the file is missing it is fitted first: on the GPU through
    ptb_spectrum_table_build when a Context is given (about 55 ms on a B200), else on all host threads (seconds)."""
    L = lib()
    if ctx is not None and not os.path.exists(_build.SPECTRUM_TABLE):
        from . import integrator
        integrator.build_spectrum_table(ctx, _build.SPECTRUM_TABLE)
    return L.ptbh_ensure_spectrum_table()


def rgb_to_spectrum(rgb):
    out = (C.c_float * 3)()
    lib().ptbh_rgb_to_spectrum(_f3(rgb), out)
    return np.array(out[:], dtype=np.float32)


def fit_spectrum(rgb):
    out = (C.c_float * 3)()
    lib().ptbh_fit_spectrum(_f3(rgb), out)
    return np.array(out[:], dtype=np.float32)


def decode_image(path):
    """What LoadTexture reads from an image file (the reference: stbi_loadf(path, ..., 4), scene.cpp:294-297):
    linear RGBA32F (H, W, 4) from PNG, JPEG, Radiance HDR (or PPM / PFM); None when the file cannot be decoded."""
    L = lib()
    w, h = C.c_int(0), C.c_int(0)
    if L.ptbh_decode_image(os.fsencode(path), C.byref(w), C.byref(h), None, 0) != 0:
        return None
    out = np.zeros((h.value, w.value, 4), dtype=np.float32)
    L.ptbh_decode_image(os.fsencode(path), C.byref(w), C.byref(h), out.ctypes.data, out.size)
    return out


def write_png(path, srgb8):
    """The resolved 8-bit frame (H, W, 4) as a PNG file (the reference presents to a swapchain instead)."""
    a = np.ascontiguousarray(srgb8, dtype=np.uint8)
    if lib().ptbh_write_png(os.fsencode(path), a.shape[1], a.shape[0], a.ctypes.data) != 0:
        raise OSError("cannot write " + str(path))


def write_pfm(path, linear_rgb):
    """Linear float RGB (H, W, 3) as a PFM file."""
    a = np.ascontiguousarray(linear_rgb, dtype=np.float32)
    if lib().ptbh_write_pfm(os.fsencode(path), a.shape[1], a.shape[0], a.ctypes.data) != 0:
        raise OSError("cannot write " + str(path))


def _view_array(ptr, count, dtype):
    if not ptr or count == 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count)


class Scene:
    """A host scene (source description + packed arrays)."""

    def __init__(self, handle=None, with_defaults=False):
        self._L = lib()
        self.h = handle if handle is not None else self._L.ptbh_scene_create(1 if with_defaults else 0)
        if not self.h:
            raise RuntimeError("scene creation failed")
        self._view = None

    # -- named procedural scenes (BASELINE.json configs) --
    @classmethod
    def make(cls, name, *params):
        L = lib()
        arr = (C.c_double * max(1, len(params)))(*[float(p) for p in params])
        h = L.ptbh_make_scene(name.encode(), arr, len(params))
        if not h:
            raise ValueError("unknown scene " + name)
        return cls(handle=h)

    # -- LoadScene / SaveScene (scene.hpp:432-433): the reference's JSON + side-file format --
    @classmethod
    def load(cls, path):
        h = lib().ptbh_load_scene(os.fsencode(path))
        if not h:
            raise OSError(lib().ptbh_last_error().decode(errors="replace"))
        return cls(handle=h)

    def save(self, path):
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        if self._L.ptbh_save_scene(self.h, os.fsencode(path)) != 0:
            raise OSError(self._L.ptbh_last_error().decode(errors="replace"))

    def close(self):
        if self.h:
            self._L.ptbh_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- source description (scene.hpp:410-424) --
    def set_atlas_size(self, w, h):
        self._L.ptbh_scene_set_atlas_size(self.h, w, h)

    def set_root(self, scatter_rate=0.0, sky_brightness=1.0, sky_sampling_probability=0.0, sky_texture=None):
        self._L.ptbh_root_set(self.h, scatter_rate, sky_brightness, sky_sampling_probability, sky_texture)

    def create_entity(self, etype, parent=None, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1), material=None, mesh=None):
        e = self._L.ptbh_create_entity(self.h, etype, parent)
        self._L.ptbh_entity_set_transform(self.h, e, _f3(position), _f3(rotation), _f3(scale))
        if material is not None:
            self._L.ptbh_entity_set_material(self.h, e, material)
        if mesh is not None:
            self._L.ptbh_entity_set_mesh(self.h, e, mesh)
        return e

    def create_camera(self, position, look_at, model=CAMERA_PINHOLE, params=(90.0, 0.0)):
        e = self._L.ptbh_create_entity(self.h, ENTITY_CAMERA, None)
        self._L.ptbh_entity_look_at(self.h, e, _f3(position), _f3(look_at))
        p = list(params) + [0.0] * 5
        self._L.ptbh_camera_set(self.h, e, model, _f3(p))
        return e

    def set_transform(self, e, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1)):
        self._L.ptbh_entity_set_transform(self.h, e, _f3(position), _f3(rotation), _f3(scale))

    def set_active(self, e, active):
        self._L.ptbh_entity_set_active(self.h, e, 1 if active else 0)

    def destroy_entity(self, e):
        self._L.ptbh_destroy_entity(self.h, e)

    def create_material(self, mtype, name="Material", textures=None, **fields):
        m = self._L.ptbh_create_material(self.h, mtype, name.encode())
        for k, v in fields.items():
            vv = v if hasattr(v, "__len__") else [v]
            if self._L.ptbh_material_set(self.h, m, k.encode(), _f3(vv)) != 0:
                raise KeyError(k)
        for k, t in (textures or {}).items():
            if self._L.ptbh_material_set_texture(self.h, m, k.encode(), t) != 0:
                raise KeyError(k)
        return m

    def create_texture(self, name, ttype, pixels, nearest=False):
        px = np.ascontiguousarray(pixels, dtype=np.float32)
        h, w = px.shape[0], px.shape[1]
        return self._L.ptbh_create_texture(self.h, name.encode(), ttype, w, h, px.ctypes.data_as(C.POINTER(C.c_float)), int(nearest))

    def create_checker_texture(self, name, ttype, a, b, nearest=False):
        return self._L.ptbh_create_checker_texture(self.h, name.encode(), ttype, _f3(a), _f3(b), int(nearest))

    def create_mesh(self, name, positions, normals, uvs, faces):
        n = len(positions)
        v = np.zeros((n, 8), dtype=np.float32)
        v[:, 0:3] = positions
        v[:, 3:6] = normals
        v[:, 6:8] = uvs
        f = np.ascontiguousarray(faces, dtype=np.uint32)
        return self._L.ptbh_create_mesh(self.h, name.encode(), v.ctypes.data_as(C.POINTER(C.c_float)), n,
                                        f.ctypes.data_as(C.POINTER(C.c_uint32)), len(f))

    # -- model import and prefabs (scene.hpp:415,428-429) --
    def load_model(self, path, name=None, default_material=None, vertex_transform=None):
        """LoadModelAsPrefab: Wavefront OBJ (+ MTL) -> prefab handle.  vertex_transform: 4x4 column-major or None."""
        vt = None
        if vertex_transform is not None:
            vt = np.ascontiguousarray(vertex_transform, dtype=np.float32).reshape(16).ctypes.data_as(C.POINTER(C.c_float))
        p = self._L.ptbh_load_model(self.h, os.fsencode(path), (name or "").encode(), default_material, vt)
        if not p:
            raise OSError("cannot load model " + str(path))
        return p

    def create_prefab_instance(self, prefab, parent=None, position=None, rotation=(0, 0, 0), scale=(1, 1, 1)):
        e = self._L.ptbh_create_prefab_instance(self.h, prefab, parent)
        if position is not None:
            self._L.ptbh_entity_set_transform(self.h, e, _f3(position), _f3(rotation), _f3(scale))
        return e

    def destroy_prefab(self, prefab):
        self._L.ptbh_destroy_prefab(self.h, prefab)

    def prefab_mesh_count(self, prefab):
        return self._L.ptbh_prefab_mesh_count(prefab)

    def load_texture(self, path, ttype, name=None):
        """LoadTexture (scene.cpp:294-318): PNG, JPEG, Radiance HDR as stbi_loadf delivers them; also PPM / PFM."""
        t = self._L.ptbh_load_texture(self.h, os.fsencode(path), ttype, (name or "").encode())
        if not t:
            raise OSError("cannot load texture " + str(path))
        return t

    def source_mesh_vertices(self, i=0):
        m = self._L.ptbh_scene_mesh(self.h, i)
        out = np.zeros((self._L.ptbh_mesh_vertex_count(m), 8), dtype=np.float32)
        if len(out):
            self._L.ptbh_mesh_vertices(m, out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def mesh_count(self):
        return self._L.ptbh_scene_mesh_count(self.h)

    def atlas_layer(self, i=0):
        v = self.view
        n = v.AtlasWidth * v.AtlasHeight * 4
        ptrs = (C.c_void_p * v.AtlasLayerCount).from_address(v.AtlasLayers)
        buf = (C.c_float * n).from_address(ptrs[i])
        return np.frombuffer(buf, dtype=np.float32).reshape(v.AtlasHeight, v.AtlasWidth, 4)

    def shape_index(self, e):
        return self._L.ptbh_entity_shape_index(e)

    def camera_index(self, e):
        return self._L.ptbh_entity_camera_index(e)

    def material_slot(self, m):
        return self._L.ptbh_material_slot(m)

    def mesh_info(self, i=0):
        m = self._L.ptbh_scene_mesh(self.h, i)
        return dict(depth=self._L.ptbh_mesh_depth(m), nodes=self._L.ptbh_mesh_node_count(m), faces=self._L.ptbh_mesh_face_count(m))

    # -- PackSceneData (scene.hpp:436) --
    def pack(self):
        self._view = None
        return self._L.ptbh_scene_pack(self.h)

    def mark_dirty(self, flags):
        self._L.ptbh_scene_mark_dirty(self.h, flags)

    @property
    def view(self):
        """ptb_scene_view over the packed host arrays (valid until the next edit/pack)."""
        if self._view is None:
            v = SceneView()
            self._L.ptbh_scene_view(self.h, C.byref(v))
            self._view = v
        return self._view

    # numpy views of the packs (zero-copy, read-only use)
    @property
    def shapes(self):
        return _view_array(self.view.Shapes, self.view.ShapeCount, SHAPE_DTYPE)

    @property
    def shape_nodes(self):
        return _view_array(self.view.ShapeNodes, self.view.ShapeNodeCount, SHAPE_NODE_DTYPE)

    @property
    def mesh_faces(self):
        return _view_array(self.view.MeshFaces, self.view.MeshFaceCount, MESH_FACE_DTYPE)

    @property
    def mesh_vertices(self):
        return _view_array(self.view.MeshVertices, self.view.MeshVertexCount, MESH_VERTEX_DTYPE)

    @property
    def mesh_nodes(self):
        return _view_array(self.view.MeshNodes, self.view.MeshNodeCount, MESH_NODE_DTYPE)

    @property
    def textures(self):
        return _view_array(self.view.Textures, self.view.TextureCount, TEXTURE_DTYPE)

    @property
    def cameras(self):
        return _view_array(self.view.Cameras, self.view.CameraCount, CAMERA_DTYPE)

    @property
    def materials(self):
        return _view_array(self.view.MaterialData, self.view.MaterialWordCount, np.dtype("<u4"))

    @property
    def globals(self):
        return _view_array(self.view.Globals, 1, GLOBALS_DTYPE)[0]

    def bounds(self):
        n = self.shape_nodes
        return n[0]["Minimum"].copy(), n[0]["Maximum"].copy()
