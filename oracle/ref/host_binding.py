"""ORACLE - TEST INFRASTRUCTURE ONLY.

oracle/_ref/libref_host.so: the reference's own host scene code (src/scene/scene.cpp's non-Vulkan half,
src/core/spectrum.cpp, the material packers, the vendored stb_image / stb_rect_pack / tiny_obj_loader) compiled where it
lies against oracle/ref/hostshim (stand-in for the un-vendored glm 1.0.1; stub core/vulkan.hpp).  See ref_host.cpp.
It exports path_tracer_b200/host/capi.cpp's ptbh_* names, so `RefScene` below is path_tracer_b200.scene.Scene driving
the REFERENCE's CreateEntity / CreateMaterial / LoadModelAsPrefab / PackSceneData.  Built only where /root/reference
exists; the built library travels with gpurun but nothing on the GPU box needs it (CPU suite only).
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(os.path.dirname(HERE), "_ref")
LIB = os.path.join(OUT, "libref_host.so")
REFERENCE_SRC = "/root/reference/src"
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def available():
    return os.path.exists(LIB) or os.path.isdir(REFERENCE_SRC)


def build(force=False):
    if not os.path.isdir(REFERENCE_SRC):
        if os.path.exists(LIB):
            return LIB
        raise RuntimeError("oracle/_ref: %s is not here and no prebuilt libref_host.so travelled with the repository" % REFERENCE_SRC)
    deps = [os.path.join(HERE, "ref_host.cpp"), os.path.join(HERE, "ref_serializer.cpp"), os.path.join(HERE, "hostgen.py"), os.path.join(HERE, "hostshim", "glm", "glm.hpp"),
            os.path.join(HERE, "hostshim", "core", "vulkan.hpp"), os.path.join(ROOT, "include", "ptb.h")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        gen = os.path.join(OUT, "gen")
        os.makedirs(gen, exist_ok=True)
        subprocess.check_call([sys.executable, os.path.join(HERE, "hostgen.py"), REFERENCE_SRC, gen])
        flags = ["-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-fpermissive", "-w", "-I", os.path.join(HERE, "hostshim"), "-I", gen, "-I", REFERENCE_SRC,
                 "-I", os.path.join(ROOT, "include")]
        miniz, ser = os.path.join(gen, "miniz.o"), os.path.join(gen, "ref_serializer.o")
        for cmd in (["gcc", "-O2", "-fPIC", "-w", "-c", "-o", miniz, os.path.join(REFERENCE_SRC, "core", "miniz.c")],
                    ["g++", "-std=c++20"] + flags + ["-c", "-o", ser, os.path.join(HERE, "ref_serializer.cpp")],
                    ["g++", "-std=c++20"] + flags + ["-shared", "-o", LIB, os.path.join(HERE, "ref_host.cpp"), ser, miniz]):
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                raise RuntimeError("oracle/_ref host build failed:\n" + r.stdout)
    return LIB


_lib = None


def lib():
    """The library with the product binding's prototypes copied onto every ptbh_* function both libraries export."""
    global _lib
    if _lib is None:
        from path_tracer_b200 import scene as S
        from path_tracer_b200 import _build
        P = S.lib()
        S.ensure_spectrum_table()
        R = C.CDLL(build())
        for name, fn in list(vars(P).items()):
            if name.startswith("ptbh_") and hasattr(R, name):
                g = getattr(R, name)
                g.restype, g.argtypes = fn.restype, fn.argtypes
        R.refh_stbi_loadf.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_size_t]
        R.refh_spectrum_column.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
        R.refh_spectrum_lab_error.restype = C.c_double
        R.refh_spectrum_lab_error.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
        R.ptbh_set_spectrum_table_path(_build.SPECTRUM_TABLE.encode())      # same raw file format (spectrum.cpp:426-448)
        assert R.ptbh_ensure_spectrum_table() == 1
        _lib = R
    return _lib


def ref_scene_class():
    from path_tracer_b200 import scene as S

    class RefScene(S.Scene):
        """path_tracer_b200.scene.Scene over the reference's own scene.cpp."""

        def __init__(self, with_defaults=False):
            self._L = lib()
            self.h = self._L.ptbh_scene_create(1 if with_defaults else 0)
            if not self.h:
                raise RuntimeError("reference scene creation failed")
            self._view = None

        @classmethod
        def load(cls, path):
            """The reference's LoadScene (serializer.cpp:501-516)."""
            h = lib().ptbh_load_scene(os.fsencode(path))
            if not h:
                raise OSError("the reference's LoadScene failed on " + str(path))
            sc = cls.__new__(cls)
            sc._L, sc.h, sc._view = lib(), h, None
            return sc

        def save(self, path):
            """The reference's SaveScene (serializer.cpp:518-529)."""
            if self._L.ptbh_save_scene(self.h, os.fsencode(path)) != 0:
                raise OSError("the reference's SaveScene failed on " + str(path))

        def set_atlas_size(self, w, h):
            assert (w, h) == (4096, 4096), "the reference's atlas size is a constant (scene.cpp:1122-1123)"

        def create_camera(self, *a, **k):
            raise NotImplementedError("use create_entity(ENTITY_CAMERA, rotation=...) + set_camera: the reference has no look-at helper")

    return RefScene


def stbi_loadf(path):
    """stbi_loadf(path, &w, &h, &n, 4) of the reference's vendored stb_image.h: (H, W, 4) float32 or None."""
    L = lib()
    w, h = C.c_int(0), C.c_int(0)
    if L.refh_stbi_loadf(os.fsencode(path), C.byref(w), C.byref(h), None, 0) != 0:
        return None
    out = np.zeros((h.value, w.value, 4), dtype=np.float32)
    assert L.refh_stbi_loadf(os.fsencode(path), C.byref(w), C.byref(h), out.ctypes.data, out.size) == 0
    return out


def spectrum_column(i, j, l):
    out = np.zeros((64, 3), dtype=np.float32)
    lib().refh_spectrum_column(i, j, l, out.ctypes.data)
    return out


def spectrum_lab_error(rgb, beta):
    a = (C.c_float * 3)(*[float(x) for x in rgb])
    b = (C.c_float * 3)(*[float(x) for x in beta])
    return lib().refh_spectrum_lab_error(a, b)
