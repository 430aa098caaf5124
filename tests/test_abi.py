"""The C-ABI library loads and exports every symbol include/ptb.h declares; without a
GPU every compute entry fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "ptb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(ptb_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_symbols_are_exported():
    from path_tracer_b200 import integrator as I
    L = I.lib()
    declared = header_functions()
    assert len(declared) >= 28
    for name in declared:
        assert hasattr(L, name), "libptb.so does not export " + name
    assert sorted(I.ABI_SYMBOLS) == declared


def test_pod_sizes_match_header():
    from path_tracer_b200 import types as T
    assert C.sizeof(T.RenderParams) == 48 and C.sizeof(T.ResolveParams) == 12 and C.sizeof(T.Stats) == 32
    assert C.sizeof(T.WideInfo) == 48
    assert T.RAY_DTYPE.itemsize == 28 and T.HIT_DTYPE.itemsize == 64
    assert C.sizeof(T.SceneView) == 8 * 16 + 8 + 4 + 4 + 4 + 4 + 8


def test_product_does_not_touch_the_oracle():
    """Nothing under path_tracer_b200/ may import, include or link oracle/."""
    pkg = os.path.join(ROOT, "path_tracer_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(d, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f
                assert not re.search(r'#include\s+"[^"]*oracle', text), f
                assert "liboracle" not in text, f


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from path_tracer_b200 import integrator as I
    with pytest.raises(I.PtbError) as e:
        I.Context(0)
    assert e.value.code == -2
