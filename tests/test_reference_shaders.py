"""The oracle pinned against the reference's own shader sources.

oracle/ref/ compiles basic_trace.glsl / basic_scatter.glsl and their includes as C++ straight from /root/reference
(oracle/_ref/libref.so, git-ignored).  Where that library is available - this container, or a box the prebuilt
file travelled to - the oracle restatement must reproduce it BIT FOR BIT: hit records of random rays, the film and
the wavefront state arrays after several frames.  Everywhere, the oracle must reproduce the vectors that library
wrote into tests/golden/ref_*.npz (tests/golden/make_ref_golden.py).
"""
import ctypes as C
import os

import numpy as np
import pytest

from util import REF_CASES, make_test_scene, random_rays

GOLD = os.path.join(os.path.dirname(__file__), "golden")
MISS = 0xFFFFFFFF
HIT_DETAIL = ("Time", "TextureU", "TextureV", "PackedNormal", "PackedTangentX")     # unwritten garbage on a miss


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def ref():
    from oracle.ref import binding as R
    if not R.available():
        pytest.skip("oracle/_ref/libref.so is not built and /root/reference is not here")
    R.lib()
    return R


@pytest.fixture(scope="module")
def case_scene(scenes):
    from path_tracer_b200 import scene as S
    cache = {}

    def get(name, params):
        key = (name,) + tuple(params)
        if key not in cache:
            scenes("cornell")                            # makes sure the spectrum table is there
            cache[key] = make_test_scene(S, name, params)
        return cache[key]
    return get


def test_small_functions_agree(ref, oracle):
    L, K = ref.lib(), oracle.lib()
    n = 4096
    a, b, u = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint32)
    for seed in (0, 1, 277803737, 0xFFFFFFFF):
        L.ref_rng(seed, n, a.ctypes.data)
        K.orc_kat_rng(C.c_uint(seed), C.c_uint(n), u.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p))
        assert np.array_equal(bits(a), bits(b))
    rng = np.random.default_rng(2)
    v = rng.normal(size=(2000, 3)).astype(np.float32)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    K.orc_kat_pack_unit_vector.restype = C.c_uint32
    K.orc_kat_pack_unit_vector.argtypes = [C.c_float] * 3
    K.orc_kat_unpack_unit_vector.argtypes = [C.c_uint32, C.c_void_p]
    K.orc_kat_observer.argtypes = [C.c_float, C.c_void_p]
    x, y = np.zeros(3, np.float32), np.zeros(3, np.float32)
    for p in v:
        pa, pb = L.ref_pack_unit_vector(*map(float, p)), K.orc_kat_pack_unit_vector(*map(float, p))
        assert pa == pb
        L.ref_unpack_unit_vector(pa, x.ctypes.data)
        K.orc_kat_unpack_unit_vector(pb, y.ctypes.data)
        assert np.array_equal(bits(x), bits(y))
    for lam in np.linspace(340.0, 850.0, 300, dtype=np.float32):
        L.ref_observer(float(lam), x.ctypes.data)
        K.orc_kat_observer(float(lam), y.ctypes.data)
        assert np.array_equal(bits(x), bits(y))


@pytest.mark.parametrize("tag", sorted(REF_CASES))
def test_oracle_raycast_is_bit_identical_to_the_reference_shader(ref, oracle, case_scene, tag):
    name, params, *_ = REF_CASES[tag]
    sc = case_scene(name, params)
    rays = random_rays(sc, 20000, 3, oracle)
    a, b = ref.raycast(sc.view, rays), oracle.raycast(sc.view, rays, tie=oracle.TIE_FAITHFUL)
    assert np.array_equal(a["shape"], b["shape"]) and np.array_equal(bits(a["t"]), bits(b["t"]))
    assert np.array_equal(a["scene_complexity"], b["scene_complexity"]) and np.array_equal(a["mesh_complexity"], b["mesh_complexity"])
    hit = a["shape"] != MISS
    assert 0.2 < hit.mean() < 0.9
    for f in ("prim", "material", "shape_type", "n", "tan", "uv", "bary"):
        assert np.array_equal(bits(a[f][hit]), bits(b[f][hit])), f


@pytest.mark.parametrize("tag", sorted(REF_CASES))
def test_oracle_render_is_bit_identical_to_the_reference_shaders(ref, oracle, case_scene, tag):
    name, params, camera, flags, term = REF_CASES[tag]
    sc = case_scene(name, params)
    w, h = 64, 48
    o = oracle.Renderer(sc.view, w, h, camera_index=camera, render_flags=flags, path_termination_probability=term, tie=oracle.TIE_FAITHFUL)
    r = ref.Renderer(sc.view, w, h, camera_index=camera, render_flags=flags, path_termination_probability=term)
    o.set_frame_index(0)
    r.set_frame_index(0)
    o.reset()
    r.reset()
    o.run(2)
    r.run(2)
    for step in range(3):
        a, b = r.accum(), o.accum()
        assert np.array_equal(bits(a), bits(b)), step
        hit = r.state("ShapeAndMaterialIndex") != MISS
        for arr in ref.STATE_ARRAYS:
            x, y = bits(r.state(arr)), bits(o.state(arr))
            if arr in HIT_DETAIL:
                x, y = x[hit], y[hit]
            assert np.array_equal(x, y), (step, arr)
        o.run_frames(10)
        r.run_frames(10)
    assert a[..., 3].sum() > 500


@pytest.mark.parametrize("tag", sorted(REF_CASES))
def test_oracle_reproduces_reference_golden_vectors(oracle, case_scene, tag):
    """Needs no reference at run time: the vectors were written by oracle/_ref in the container that has it."""
    name, params, camera, flags, term = REF_CASES[tag]
    g = np.load(os.path.join(GOLD, "ref_" + tag + ".npz"))
    sc = case_scene(name, params)
    rays = np.ascontiguousarray(g["rays"]).view(oracle.RAY_DTYPE).reshape(-1)
    hits = oracle.raycast(sc.view, rays, tie=oracle.TIE_FAITHFUL)
    assert np.array_equal(hits["shape"], g["hit_shape"]) and np.array_equal(bits(hits["t"]), bits(g["hit_t"]))
    hit = hits["shape"] != MISS
    for f in ("prim", "material", "shape_type", "n", "tan", "uv", "bary"):
        assert np.array_equal(bits(hits[f][hit]), bits(g["hit_" + f][hit])), f
    w, h = int(g["size"][0]), int(g["size"][1])
    o = oracle.Renderer(sc.view, w, h, camera_index=camera, render_flags=flags, path_termination_probability=term, tie=oracle.TIE_FAITHFUL)
    o.set_frame_index(0)
    o.reset()
    o.run(2)
    a = o.accum()
    assert np.array_equal(a[..., 3], g["accum_first"][..., 3])
    assert np.allclose(a, g["accum_first"], rtol=1e-5, atol=1e-6)            # libm may differ across boxes by an ulp
    for arr in ("ImagePosition", "ShapeAndMaterialIndex"):
        assert np.array_equal(o.state(arr), g["state_" + arr]), arr
    for arr in ("NormalizedLambda0", "OriginX", "OriginY", "OriginZ"):
        assert np.allclose(o.state(arr), g["state_" + arr], rtol=1e-5, atol=1e-6), arr
    o.run_frames(int(g["frames"][0]))
    b = o.accum()
    same = bits(b) == bits(g["accum_frames"])
    assert same.mean() > 0.99                                                # bit-exact on this box; an ulp of libm can fork a path elsewhere
    assert abs(b[..., 3].sum() - g["accum_frames"][..., 3].sum()) <= 0.01 * g["accum_frames"][..., 3].sum()


# ---------------------------------------------------------------- the two fragment shaders: resolve.glsl, preview_render.glsl
def nan_or_bit_equal(a, b):
    return ((bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))).all()


def test_oracle_resolve_is_bit_identical_to_the_reference_shader(ref, oracle):
    from golden.make_ref_golden import resolve_film
    film = resolve_film()
    for mode in range(4):
        for brightness, white in ((1.0, 1.0), (2.5, 4.0), (0.1, 0.5)):
            assert nan_or_bit_equal(ref.resolve(film, brightness, mode, white), oracle.resolve(film, brightness, mode, white)[0]), (mode, brightness)


@pytest.mark.parametrize("tag", ["cornell", "room", "mixed_lens", "spherefield", "terrain32"])
def test_oracle_preview_is_bit_identical_to_the_reference_shader(ref, oracle, case_scene, tag):
    name, params, *_ = REF_CASES[tag]
    sc = case_scene(name, params)
    cam = sc.cameras[0]
    to, fr = cam["To"].ravel(), cam["From"].ravel()
    for mode in range(7):
        a, pick_a = ref.preview(sc.view, to, fr, mode, 48, 32, brightness=1.25, selected=2, mouse=(20, 17))
        b, _, pick_b = oracle.preview(sc.view, to, fr, mode, 48, 32, brightness=1.25, selected=2, mouse=(20, 17), tie=oracle.TIE_FAITHFUL)
        assert np.array_equal(bits(a), bits(b)) and pick_a == pick_b, mode


def test_oracle_reproduces_reference_resolve_and_preview_vectors(oracle, case_scene):
    g = np.load(os.path.join(GOLD, "ref_resolve.npz"))
    for k, (brightness, mode, white) in enumerate(g["settings"]):
        lin, _ = oracle.resolve(g["film"], float(brightness), int(mode), float(white))
        assert np.allclose(lin, g["linear_%d" % k], rtol=1e-6, atol=1e-7, equal_nan=True)
    g = np.load(os.path.join(GOLD, "ref_preview.npz"))
    w, h = int(g["size"][0]), int(g["size"][1])
    for tag in ("cornell", "room", "mixed_lens"):
        name, params, *_ = REF_CASES[tag]
        sc = case_scene(name, params)
        cam = sc.cameras[0]
        for mode in range(7):
            img, _, pick = oracle.preview(sc.view, cam["To"].ravel(), cam["From"].ravel(), mode, w, h, brightness=1.25, selected=2, mouse=(20, 17),
                                          tie=oracle.TIE_FAITHFUL)
            assert np.allclose(img, g["%s_mode%d" % (tag, mode)], rtol=1e-6, atol=1e-7) and pick == int(g[tag + "_pick"][0]), (tag, mode)


def test_oracle_equals_the_reference_shaders_on_random_scenes(ref):
    """A slice of tools/fuzz_scatter.py --reference: scenes nobody designed (see its docstring), film and all wavefront
    state arrays after every round, up to where the reference's medium of a non-dielectric shape becomes undefined."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_scatter
    assert fuzz_scatter.main(["--scenes", "40", "--seed", "12", "--width", "48", "--height", "32", "--reference"]) == 0


def test_oracle_preview_equals_the_reference_shader_on_random_scenes(ref, oracle):
    """preview_render.glsl on the random scenes of the fuzzers (nested transforms, degenerate meshes, every material, textured sky,
    all three camera entities as viewpoints), all seven modes, pick query and selection tint included."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_scatter
    import fuzz_traversal
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    done = 0
    for seed in range(24):
        rng = np.random.default_rng(7000 + seed)
        sc = fuzz_scatter.random_scene(S, rng)[0] if seed % 2 else fuzz_traversal.random_scene(S, rng, False)
        if not len(sc.cameras):
            continue
        cam = sc.cameras[seed % len(sc.cameras)]
        to, fr = cam["To"].ravel(), cam["From"].ravel()
        for mode in range(7):
            sel = int(rng.integers(0, max(1, len(sc.shapes))))
            mouse = (int(rng.integers(0, 40)), int(rng.integers(0, 24)))
            a, pick_a = ref.preview(sc.view, to, fr, mode, 40, 24, brightness=0.8, selected=sel, mouse=mouse)
            b, _, pick_b = oracle.preview(sc.view, to, fr, mode, 40, 24, brightness=0.8, selected=sel, mouse=mouse, tie=oracle.TIE_FAITHFUL)
            same = (bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))
            assert same.all() and pick_a == pick_b, (seed, mode, int((~same).sum()))
        done += 1
    assert done >= 12
