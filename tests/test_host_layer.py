"""Host logic on CPU: packed layouts, the reference-order mesh BVH2, and the wide-BVH
flattening (checked through tests/host_emu, a g++ build of the product's per-thread
device functions — test infrastructure, see tests/host_emu/emu.cpp)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from util import ids_equal, random_rays

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU_DIR = os.path.join(ROOT, "tests", "host_emu")


@pytest.fixture(scope="module")
def emu():
    out = os.path.join(EMU_DIR, "_build", "libemu.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    csrc = os.path.join(ROOT, "path_tracer_b200", "csrc")
    srcs = [os.path.join(EMU_DIR, "emu.cpp"), os.path.join(csrc, "ptb_wide_build.cpp")]
    deps = srcs + [os.path.join(csrc, f) for f in os.listdir(csrc)]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-I", csrc,
                               "-I", os.path.join(ROOT, "include"), "-I", "/usr/local/cuda/include", "-o", out] + srcs)
    L = C.CDLL(out)
    L.emu_create.restype = C.c_void_p
    L.emu_create.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
    L.emu_destroy.argtypes = [C.c_void_p]
    L.emu_raycast.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
    L.emu_wide_info.argtypes = [C.c_void_p, C.c_void_p]
    L.emu_wide_download.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.emu_set.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_uint32]
    L.emu_reset.argtypes = [C.c_void_p]
    L.emu_run.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int]
    L.emu_accum.restype = C.c_void_p
    L.emu_accum.argtypes = [C.c_void_p]
    return L


def test_snorm_unpack_equals_the_division_for_every_input(emu):
    """ptb_vec.cuh unpacks snorm16 with a multiply and two fused multiply-adds; GLSL defines it as clamp(x / 32767, -1, 1)."""
    out = np.zeros(2, np.float32)
    x = np.arange(-32768, 32768, dtype=np.int32)
    want = np.clip(x.astype(np.float32) / np.float32(32767.0), -1.0, 1.0).astype(np.float32)
    got = np.zeros((65536, 2), np.float32)
    for k, v in enumerate(x):
        emu.emu_unpack_snorm2x16(C.c_uint32((int(v) & 0xFFFF) | ((int(-v - 1) & 0xFFFF) << 16)), out.ctypes.data_as(C.c_void_p))
        got[k] = out
    assert np.array_equal(got[:, 0].view(np.uint32), want.view(np.uint32))
    assert np.array_equal(got[:, 1].view(np.uint32), want[::-1].view(np.uint32))


def test_magic_division_is_exact(emu):
    """udivmod (ptb_scatter.cuh): multiply-high by floor(2^32 / d) plus one correction equals n // d, n % d."""
    rng = np.random.default_rng(4)
    out = (C.c_uint32 * 2)()
    ds = [1, 2, 3, 7, 16, 17, 120, 240, 256, 1920, 3840, 4096, 65535, 65536, 2**31, 2**32 - 1] + [int(x) for x in rng.integers(1, 2**32, 40)]
    for d in ds:
        ns = [0, 1, d - 1, d, d + 1, 2 * d - 1, 2**32 - 1, 2**32 - d] + [int(x) for x in rng.integers(0, 2**32, 200)]
        for n in ns:
            n &= 0xFFFFFFFF
            emu.emu_udivmod(C.c_uint32(n), C.c_uint32(d), out)
            assert (out[0], out[1]) == (n // d, n % d), (n, d)


def test_packed_layouts(scenes):
    sc = scenes("terrain", 32)
    v = sc.view
    assert v.MeshFaceCount == 2 * 32 * 32 and v.ShapeCount == 2
    assert sc.shapes.dtype.itemsize == 144 and sc.mesh_nodes.dtype.itemsize == 32 and sc.mesh_faces.dtype.itemsize == 48
    assert v.MaterialWordCount % 32 == 0
    # material slots 0-1 are the fallback OpenPBR record (scene.cpp:1241-1250): first user slot is 2
    assert int(sc.shapes["MaterialIndex"].min()) >= 2
    g = sc.globals
    assert g["ShapeCount"] == v.ShapeCount and g["SkyboxTextureIndex"] == 0xFFFFFFFF


def test_mesh_bvh2_invariants(scenes):
    """Children adjacent, leaves partition the face range, boxes contain their faces (scene.cpp:435-599)."""
    sc = scenes("terrain", 32)
    nodes, faces = sc.mesh_nodes, sc.mesh_faces
    root = int(sc.shapes[sc.shapes["Type"] == 0]["MeshRootNodeIndex"][0])
    seen = np.zeros(len(faces), dtype=np.int32)
    stack = [root]
    while stack:
        n = nodes[stack.pop()]
        if n["FaceEndIndex"] > 0:
            b, e = int(n["FaceBeginOrNodeIndex"]), int(n["FaceEndIndex"])
            seen[b:e] += 1
            for k in ("Position0", "Position1", "Position2"):
                p = faces[k][b:e]
                assert (p >= n["Minimum"] - 1e-6).all() and (p <= n["Maximum"] + 1e-6).all()
        else:
            c = int(n["FaceBeginOrNodeIndex"])
            stack += [c, c + 1]
    assert (seen == 1).all()


def test_wide_flattening_of_a_large_mesh_on_several_threads(emu, oracle, scenes):
    """Above 262 144 primitives the host builder cuts the passes of a node into pieces, hands subtrees to other threads
    and emits the wide tree level by level on all threads (csrc/ptb_wide_build.cpp).  The tree must not depend on the
    thread schedule, must contain everything below every slot, and its closest hits must be the brute-force ones."""
    sc = scenes("terrain", 370)                               # 273 800 triangles + the ground plane
    trees = []
    for _ in range(2):
        e = emu.emu_create(C.addressof(sc.view), 16, 16)
        assert e
        info = (C.c_uint32 * 5)()
        emu.emu_wide_info(e, info)
        assert info[1] == 2 * 370 * 370 and info[2] <= 30
        nodes, prims = np.zeros((info[0], 20), np.uint32), np.zeros((info[1], 12), np.float32)
        emu.emu_wide_download(e, nodes.ctypes.data, prims.ctypes.data)
        trees.append((e, nodes, prims, info[2]))
    assert trees[0][1].tobytes() == trees[1][1].tobytes() and trees[0][2].tobytes() == trees[1][2].tobytes()
    from util import check_wide_tree
    e, nodes, prims, depth = trees[0]
    assert check_wide_tree(nodes, prims, sc.shapes) == depth
    rank = prims[:, 11].view(np.uint32) >> 3                  # canonical (shape, face) order, the plane ranked among them
    key = (prims[:, 7].view(np.uint32).astype(np.uint64) << np.uint64(32)) | prims[:, 3].view(np.uint32).astype(np.uint64)
    assert len(np.unique(rank)) == len(rank) and np.array_equal(np.argsort(rank), np.argsort(key))
    n = 400
    rays = random_rays(sc, n, 33, oracle)
    from path_tracer_b200.types import HIT_DTYPE
    out = np.zeros(n, dtype=HIT_DTYPE)
    emu.emu_raycast(e, rays.ctypes.data, n, 2, out.ctypes.data)
    bf = oracle.raycast_bruteforce(sc.view, rays)
    assert ids_equal(out, bf).all() and (out["t"].view(np.uint32) == bf["t"].view(np.uint32)).all()
    for t in trees:
        emu.emu_destroy(t[0])


def test_resolve_on_random_accumulators(emu, oracle, scenes):
    """resolve_pixel (ptb_resolve.cuh, host build) against the oracle's restatement of resolve.glsl and - where it is here -
    against the reference's own fragment stage, on accumulators nobody would render: huge, tiny, negative, zero sample counts,
    infinities and NaNs, every tone-mapping mode, random brightness and white level.  Bit for bit (same libm)."""
    sc = scenes("cornell")
    w, h = 64, 48
    e = emu.emu_create(C.addressof(sc.view), w, h)
    emu.emu_resolve.argtypes = [C.c_void_p, C.c_float, C.c_uint32, C.c_float, C.c_void_p, C.c_void_p]
    acc = np.frombuffer((C.c_float * (w * h * 4)).from_address(emu.emu_accum(e)), dtype=np.float32).reshape(h, w, 4)
    try:
        from oracle.ref import binding as R
        ref = R if R.available() else None
        if ref:
            ref.lib()
    except Exception:
        ref = None
    rng = np.random.default_rng(8)
    for trial in range(24):
        a = (rng.uniform(0, 1, size=(h, w, 4)) * 10.0 ** rng.integers(-6, 7, size=(h, w, 1))).astype(np.float32)
        a[..., 3] = rng.integers(0, 5, size=(h, w)) * rng.choice([1.0, 1000.0])         # sample counts, zeros included
        a[rng.random(size=(h, w)) < 0.05, :3] *= -1.0
        for v in (np.inf, -np.inf, np.nan, 0.0, -0.0):
            a[rng.integers(0, h), rng.integers(0, w), rng.integers(0, 4)] = v
        acc[...] = a
        mode = trial % 4
        brightness, white = float(rng.choice([0.0, 0.01, 1.0, 3.7, 1e4])), float(rng.choice([0.1, 1.0, 11.2, 1e3]))
        lin = np.zeros((h, w, 3), np.float32)
        srgb = np.zeros((h, w, 4), np.uint8)
        emu.emu_resolve(e, brightness, mode, white, lin.ctypes.data, srgb.ctypes.data)
        o_lin, o_srgb = oracle.resolve(a, brightness, mode, white)
        same = (lin.view(np.uint32) == o_lin.view(np.uint32)) | (np.isnan(lin) & np.isnan(o_lin))
        assert same.all(), (trial, mode, np.argwhere(~same)[0], lin[~same][:3], o_lin[~same][:3])
        assert np.array_equal(srgb, o_srgb), (trial, mode)
        if ref:
            r_lin = ref.resolve(a, brightness, mode, white)
            same = (r_lin.view(np.uint32) == o_lin.view(np.uint32)) | (np.isnan(r_lin) & np.isnan(o_lin))
            assert same.all(), ("reference", trial, mode, np.argwhere(~same)[0], r_lin[~same][:3], o_lin[~same][:3])
    emu.emu_destroy(e)


def test_destroying_a_container_removes_everything_in_it():
    """DestroyEntity on a container with children (the reference's walks the child list while the children erase themselves from it,
    scene.cpp:256-268: undefined behaviour; host/scene.cpp copies the list first).  The packed scene must equal one that never had
    the container."""
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()

    def build(with_box):
        sc = S.Scene()
        m = sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "d", BaseColor=(0.5, 0.6, 0.7))
        sc.create_entity(S.ENTITY_PLANE, material=m)
        box = None
        if with_box:
            box = sc.create_entity(S.ENTITY_CONTAINER, position=(1, 2, 3), rotation=(0.1, 0.2, 0.3), scale=(2, 2, 2))
            sc.create_entity(S.ENTITY_SPHERE, parent=box, position=(1, 0, 0), material=m)
            inner = sc.create_entity(S.ENTITY_CONTAINER, parent=box, position=(0, 1, 0))
            sc.create_entity(S.ENTITY_CUBE, parent=inner, position=(0, 0, 1), material=m)
            sc.create_entity(S.ENTITY_SPHERE, parent=inner, position=(0, 0, 2), material=m)
            sc.create_camera((5, 5, 5), (0, 0, 0))           # a camera outside the container stays
            cam_in = sc._L.ptbh_create_entity(sc.h, S.ENTITY_CAMERA, inner)
            assert cam_in
        else:
            sc.create_camera((5, 5, 5), (0, 0, 0))
        sc.create_entity(S.ENTITY_CUBE, position=(-3, 0, 1), material=m)
        return sc, box

    a, box = build(True)
    a.pack()
    assert len(a.shapes) == 5 and len(a.cameras) == 2
    a.destroy_entity(box)
    a.pack()
    b, _ = build(False)
    b.pack()
    for name in ("shapes", "shape_nodes", "cameras", "materials"):
        assert getattr(a, name).tobytes() == getattr(b, name).tobytes(), name
    assert len(a.shapes) == 2 and len(a.cameras) == 1


def test_random_scenes_scatter_like_the_oracle(emu):
    """A slice of tools/fuzz_scatter.py: random materials over their whole parameter range, nested and overlapping
    dielectrics, media, textured sky, three cameras, every flag combination, path-length limits and termination
    probabilities - the product's scatter code (host build, same libm) must leave the accumulator the oracle leaves."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import fuzz_scatter
    assert fuzz_scatter.main(["--scenes", "40", "--seed", "11"]) == 0


def test_random_scenes_traverse_like_brute_force(emu):
    """A slice of tools/fuzz_traversal.py: random shapes, nested non-uniform transforms, triangle soups, slivers, duplicate
    and zero-area faces; uniform, axis-parallel, signed-zero and vertex-aimed rays.  The BVH8 hit must be the brute-force
    hit unless the reference's own traversal loses it too; the reference-layout traversals must equal the oracle's."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import fuzz_traversal
    assert fuzz_traversal.main(["--scenes", "150", "--rays", "1000", "--seed", "7"]) == 0


@pytest.mark.parametrize("name,params", [("cornell", ()), ("terrain", (48,)), ("room", (1, 256)), ("spherefield", (10, 2, 2))])
def test_wide_flattening_equals_brute_force(emu, oracle, scenes, name, params):
    """The wide BVH is only a culling structure: its closest hit must be the brute-force
    closest hit of the reference's primitive tests (canonical ties), bit for bit."""
    sc = scenes(name, *params)
    e = emu.emu_create(C.addressof(sc.view), 16, 16)
    assert e
    info = (C.c_uint32 * 5)()
    emu.emu_wide_info(e, info)
    assert info[2] <= 30                                    # depth fits the traversal stack
    # structure: every primitive once, every child slot box contains the world bounds of everything below it
    from util import check_wide_tree
    nodes, prims = np.zeros((info[0], 20), np.uint32), np.zeros((info[1], 12), np.float32)
    emu.emu_wide_download(e, nodes.ctypes.data, prims.ctypes.data)
    assert check_wide_tree(nodes, prims, sc.shapes) == info[2]
    n = 6000
    rays = random_rays(sc, n, 21, oracle)
    # axis-parallel and signed-zero directions are the classic slab-test traps
    rays["d"][:64] = np.eye(3, dtype=np.float32)[np.arange(64) % 3] * np.where(np.arange(64) % 2, -1, 1)[:, None]
    rays["d"][64:96, 0] = -0.0
    rays["d"][64:96] /= np.linalg.norm(rays["d"][64:96], axis=1, keepdims=True)
    from path_tracer_b200.types import HIT_DTYPE
    out = np.zeros(n, dtype=HIT_DTYPE)
    emu.emu_raycast(e, rays.ctypes.data, n, 2, out.ctypes.data)
    bf = oracle.raycast_bruteforce(sc.view, rays)
    assert ids_equal(out, bf).all()
    assert (out["t"].view(np.uint32) == bf["t"].view(np.uint32)).all()
    # the reference-layout traversal compiled from the same headers is bit-identical to the oracle's
    for mode, tie in ((0, oracle.TIE_FAITHFUL), (1, oracle.TIE_CANONICAL)):
        ref = oracle.raycast(sc.view, rays, tie=tie)
        emu.emu_raycast(e, rays.ctypes.data, n, mode, out.ctypes.data)
        assert ids_equal(out, ref).all() and (out["t"].view(np.uint32) == ref["t"].view(np.uint32)).all()
        hit = ref["shape"] != 0xFFFFFFFF
        assert np.array_equal(out["n"][hit], ref["n"][hit]) and np.array_equal(out["uv"][hit], ref["uv"][hit])
    emu.emu_destroy(e)


def test_device_functions_render_like_the_oracle(emu, oracle, scenes):
    """Scatter/regenerate logic of the product headers (g++ build) against the oracle, same libm: exact."""
    sc = scenes("cornell")
    w, h = 48, 32
    e = emu.emu_create(C.addressof(sc.view), w, h)
    r = oracle.Renderer(sc.view, w, h, tie=oracle.TIE_FAITHFUL)
    r.set_frame_index(0)
    r.reset()
    r.run(2)
    r.run_frames(5)
    emu.emu_set(e, 0, 3, 0, 0.0, 0)
    emu.emu_reset(e)
    emu.emu_run(e, 2, 0, 0)
    emu.emu_run(e, 5, 0, 1)
    b = np.frombuffer((C.c_float * (w * h * 4)).from_address(emu.emu_accum(e)), dtype=np.float32).reshape(h, w, 4)
    assert np.array_equal(r.accum(), b)
    emu.emu_destroy(e)
