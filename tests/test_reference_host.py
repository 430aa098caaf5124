"""Scene packing pinned against the REFERENCE'S OWN HOST CODE.

oracle/_ref/libref_host.so is the reference's src/scene/scene.cpp (everything before its Vulkan half), src/core/spectrum.cpp,
the material packers and the vendored stb_rect_pack / stb_image / tiny_obj_loader, compiled where they lie (oracle/ref/
ref_host.cpp; glm 1.0.1 is un-vendored and stood in for by oracle/ref/hostshim).  The same scene is built twice with the same
calls - through path_tracer_b200.scene.Scene over this repository's host layer and over that library - and every packed
array PackSceneData produces is compared byte for byte: shapes (transforms and their inverses), the shape BVH, mesh faces /
vertices / nodes (BuildMeshNode's partition order), material attribute slots, texture placements in the atlas
(stb_rect_pack) and the texels written there (spectral upsampling), cameras and globals.

CPU only.  Runs where the library can be built (/root/reference present) or has travelled."""
import os

import numpy as np
import pytest

from oracle.ref import host_binding as H
from path_tracer_b200 import scene as S

pytestmark = pytest.mark.skipif(not H.available(), reason="oracle/_ref/libref_host.so cannot be built here")


@pytest.fixture(scope="module")
def Ref():
    S.ensure_spectrum_table()
    H.lib()
    return H.ref_scene_class()


def set_camera(sc, e, model, params):
    p = list(params) + [0.0] * 5
    sc._L.ptbh_camera_set(sc.h, e, model, S._f3(p))


def torus(nu, nv, R=1.0, r=0.35):
    u, v = np.meshgrid(np.arange(nu) * 2 * np.pi / nu, np.arange(nv) * 2 * np.pi / nv, indexing="ij")
    p = np.stack([(R + r * np.cos(v)) * np.cos(u), (R + r * np.cos(v)) * np.sin(u), r * np.sin(v)], -1).reshape(-1, 3)
    n = np.stack([np.cos(v) * np.cos(u), np.cos(v) * np.sin(u), np.sin(v)], -1).reshape(-1, 3)
    uv = np.stack([u / (2 * np.pi), v / (2 * np.pi)], -1).reshape(-1, 2)
    f = []
    for i in range(nu):
        for j in range(nv):
            a, b, c, d = i * nv + j, ((i + 1) % nu) * nv + j, ((i + 1) % nu) * nv + (j + 1) % nv, i * nv + (j + 1) % nv
            f += [(a, b, c), (a, c, d)]
    return p.astype(np.float32), n.astype(np.float32), uv.astype(np.float32), np.array(f, dtype=np.uint32)


def heightfield(n, seed):
    rng = np.random.default_rng(seed)
    x, y = np.meshgrid(np.linspace(-2, 2, n), np.linspace(-2, 2, n), indexing="ij")
    z = 0.3 * np.sin(2.1 * x) * np.cos(1.7 * y) + 0.05 * rng.normal(size=x.shape)
    p = np.stack([x, y, z], -1).reshape(-1, 3).astype(np.float32)
    nrm = np.tile(np.array([0, 0, 1], np.float32), (len(p), 1))
    uv = np.stack([(x + 2) / 4, (y + 2) / 4], -1).reshape(-1, 2).astype(np.float32)
    f = []
    for i in range(n - 1):
        for j in range(n - 1):
            a, b, c, d = i * n + j, (i + 1) * n + j, (i + 1) * n + j + 1, i * n + j + 1
            f += [(a, b, c), (a, c, d)]
    return p, nrm, uv, np.array(f, dtype=np.uint32)


def build_mixed(sc, texture_sizes, seed):
    """Every entity / material / texture kind, nested containers, two meshes instanced several times."""
    rng = np.random.default_rng(seed)
    tex = []
    for k, (w, h) in enumerate(texture_sizes):
        px = rng.uniform(0, 1, size=(h, w, 4)).astype(np.float32)
        ttype = (S.TEXTURE_REFLECTANCE_WITH_ALPHA, S.TEXTURE_RAW, S.TEXTURE_RADIANCE)[k % 3]
        if ttype == S.TEXTURE_RADIANCE:
            px[..., :3] *= 6.0
            px[0, 0, :3] = 0.0                                   # the "too dark" branch (scene.cpp:1200-1210)
        tex.append(sc.create_texture("tex%d" % k, ttype, px, nearest=bool(k & 1)))
    chk = sc.create_checker_texture("checker", S.TEXTURE_REFLECTANCE_WITH_ALPHA, (1, 1, 1, 1), (0.5, 0.25, 0.125, 1), nearest=True)
    sky = [t for k, t in enumerate(tex) if k % 3 == 2]
    sc.set_root(scatter_rate=0.02, sky_brightness=1.7, sky_sampling_probability=0.4, sky_texture=sky[0] if sky else None)
    m_d = sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "d", BaseColor=(0.8, 0.3, 0.1), textures={"BaseTexture": tex[0]})
    m_c = sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "c", BaseColor=(1, 1, 1), textures={"BaseTexture": chk})
    m_m = sc.create_material(S.MATERIAL_BASIC_METAL, "m", BaseColor=(0.9, 0.8, 0.2), SpecularColor=(1.0, 0.9, 0.7), Roughness=0.2,
                             RoughnessAnisotropy=0.3, textures={"RoughnessTexture": tex[1]} if len(tex) > 1 else None)
    m_t = sc.create_material(S.MATERIAL_BASIC_TRANSLUCENT, "t", IOR=1.45, AbbeNumber=35.0, Roughness=0.05, TransmissionColor=(0.9, 0.95, 1.0),
                             TransmissionDepth=0.4, ScatteringColor=(0.7, 0.2, 0.2), ScatteringAnisotropy=0.3)
    m_p = sc.create_material(S.MATERIAL_OPENPBR, "pbr")          # two slots, numbering only (scene.hpp:468-483)
    m_e = sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "e", BaseColor=(0.2, 0.9, 0.3))
    mesh_a = sc.create_mesh("torus", *torus(24, 12))
    mesh_b = sc.create_mesh("field", *heightfield(33, seed))
    sc.create_entity(S.ENTITY_PLANE, material=m_c)
    box = sc.create_entity(S.ENTITY_CONTAINER, position=(0.5, -0.25, 0.1), rotation=(0.1, 0.2, 0.3), scale=(1.5, 1.5, 1.5))
    sc.create_entity(S.ENTITY_SPHERE, parent=box, position=(1, 2, 1), scale=(0.7, 0.7, 0.7), material=m_t)
    sc.create_entity(S.ENTITY_CUBE, parent=box, position=(-2, 0.5, 0.6), rotation=(0.3, -0.4, 1.1), scale=(0.5, 0.8, 0.6), material=m_m)
    inner = sc.create_entity(S.ENTITY_CONTAINER, parent=box, position=(0, 0, 2), rotation=(0, 0, 0.7))
    sc.create_entity(S.ENTITY_MESH_INSTANCE, parent=inner, position=(0, 3, 0), rotation=(1.0, 0.2, -0.5), scale=(1.2, 1.2, 1.2), material=m_d, mesh=mesh_a)
    sc.create_entity(S.ENTITY_MESH_INSTANCE, position=(-3, -3, 0.5), scale=(2, 2, 1), material=m_e, mesh=mesh_b)
    sc.create_entity(S.ENTITY_MESH_INSTANCE, position=(4, -1, 1.5), rotation=(0.0, 1.3, 0.0), material=m_p, mesh=mesh_a)
    hidden = sc.create_entity(S.ENTITY_SPHERE, position=(9, 9, 9), material=m_d)
    sc.set_active(hidden, False)
    for k in range(7):
        pos = rng.uniform(-6, 6, 3)
        sc.create_entity((S.ENTITY_SPHERE, S.ENTITY_CUBE)[k & 1], position=pos, rotation=rng.uniform(-3, 3, 3), scale=rng.uniform(0.2, 1.2, 3),
                         material=(m_d, m_m, m_t, m_e)[k % 4])
    cams = []
    for k, (model, params) in enumerate(((S.CAMERA_PINHOLE, (70.0, 0.0)), (S.CAMERA_THIN_LENS, (36.0, 24.0, 50.0, 12.0, 4.5)), (S.CAMERA_360, ()),
                                         (S.CAMERA_PINHOLE, (45.0, 3.0)))):
        e = sc.create_entity(S.ENTITY_CAMERA, position=(0.5 * k, -7 + k, 2), rotation=(1.2, 0.1 * k, 0.3))
        set_camera(sc, e, model, params)
        cams.append(e)
    return cams


def rows_equal(name, a, b):
    """Byte equality of every named field except padding (the reference leaves std430 padding words unwritten)."""
    assert a.dtype == b.dtype and len(a) == len(b), (name, len(a), len(b))
    if a.dtype.names:
        keep = [f for f in a.dtype.names if not f.lower().startswith(("pad", "_", "unused"))]
        if len(keep) != len(a.dtype.names):
            a, b = np.ascontiguousarray(a[keep]), np.ascontiguousarray(b[keep])
            from numpy.lib import recfunctions as rf
            a, b = rf.repack_fields(a), rf.repack_fields(b)
    if a.tobytes() != b.tobytes():
        bad = [i for i in range(len(a)) if a[i].tobytes() != b[i].tobytes()]
        raise AssertionError("%s: %d of %d records differ, first %d:\n%s\n%s" % (name, len(bad), len(a), bad[0], a[bad[0]], b[bad[0]]))


def compare_packs(ours, ref):
    a, b = ours.shapes.copy(), ref.shapes.copy()
    for x in (a, b):                                         # only mesh instances have a root node; the reference leaves the word
        x["MeshRootNodeIndex"][x["Type"] != 0] = 0           # unwritten for the analytic shapes (scene.cpp:1341-1362)
    rows_equal("ShapePack", a, b)
    if len(ours.shapes):                                     # with no shape left the reference keeps the STALE root of its previous pack
        rows_equal("ShapeNodePack", ours.shape_nodes, ref.shape_nodes)      # (scene.cpp:1349 resize(1) keeps element 0); here it is an empty box
    else:
        assert len(ours.shape_nodes) == len(ref.shape_nodes) == 1 and not np.asarray(ours.shape_nodes[0].tolist()[1::2]).any()
    rows_equal("MeshFacePack", ours.mesh_faces, ref.mesh_faces)
    rows_equal("MeshVertexPack", ours.mesh_vertices, ref.mesh_vertices)
    rows_equal("MeshNodePack", ours.mesh_nodes, ref.mesh_nodes)
    a, b = ours.cameras.copy(), ref.cameras.copy()
    for x in (a, b):                                         # fields the reference leaves unwritten per model (scene.cpp:1512-1533)
        x["FocalLength"][x["Model"] != S.CAMERA_THIN_LENS] = 0
        for f in ("ApertureRadius", "SensorDistance", "SensorSize"):
            x[f][x["Model"] == S.CAMERA_360] = 0
    rows_equal("CameraPack", a, b)
    a, b = ours.materials.copy(), ref.materials.copy()
    assert len(a) == len(b) and len(a) % 32 == 0
    k = 0
    while k < len(a):                                        # slot numbering and types must agree; OpenPBR attribute words (its shader is
        assert a[k] == b[k]                                  # disabled in the reference build, scene.glsl.inc:685) are not restated: DESIGN 9
        words = 64 if a[k] == S.MATERIAL_OPENPBR else 32
        if a[k] == S.MATERIAL_OPENPBR:
            a[k + 1:k + 64] = 0
            b[k + 1:k + 64] = 0
        k += words
    rows_equal("MaterialAttributePack", a, b)
    g0, g1 = ours.globals, ref.globals
    for f in ("SkyboxMeanDirection", "SkyboxConcentration", "SkyboxSamplingProbability", "SkyboxBrightness", "SkyboxTextureIndex", "ShapeCount", "SceneScatterRate"):
        assert np.array_equal(np.asarray(g0[f]).view(np.uint32), np.asarray(g1[f]).view(np.uint32)), (f, g0[f], g1[f])
    # textures: placement, layer and flags (the reference leaves `Type` unwritten, scene.cpp:1163-1177)
    t0, t1 = ours.textures, ref.textures
    assert len(t0) == len(t1)
    for f in ("AtlasPlacementMinimum", "AtlasPlacementMaximum", "AtlasImageIndex", "Flags"):
        assert np.array_equal(t0[f].view(np.uint32), t1[f].view(np.uint32)), (f, t0[f], t1[f])
    v0, v1 = ours.view, ref.view
    assert (v0.AtlasLayerCount, v0.AtlasWidth, v0.AtlasHeight) == (v1.AtlasLayerCount, v1.AtlasWidth, v1.AtlasHeight)
    for i in range(len(t0)):
        layer = int(t0["AtlasImageIndex"][i])
        x0 = int(round(float(t0["AtlasPlacementMinimum"][i][0]) * v0.AtlasWidth - 0.5))
        x1 = int(round(float(t0["AtlasPlacementMaximum"][i][0]) * v0.AtlasWidth + 0.5))
        y0 = int(round(float(t0["AtlasPlacementMaximum"][i][1]) * v0.AtlasHeight - 0.5))
        y1 = int(round(float(t0["AtlasPlacementMinimum"][i][1]) * v0.AtlasHeight + 0.5))
        a, b = ours.atlas_layer(layer)[y0:y1, x0:x1], ref.atlas_layer(layer)[y0:y1, x0:x1]
        assert a.size > 0 and a.tobytes() == b.tobytes(), "texels of texture %d differ" % i


@pytest.mark.parametrize("texture_sizes", [
    [(64, 64), (32, 16), (128, 64)],
    [(256, 256)] * 5 + [(300, 200), (17, 900), (1024, 8), (40, 40)],          # equal sizes: ties in stb_rect_pack's height sort
    [(3000, 2500), (2048, 2048), (1500, 3000), (900, 700)],                    # more than one atlas layer
])
def test_pack_scene_data_equals_the_reference(Ref, texture_sizes):
    ours, ref = S.Scene(), Ref()
    cams0 = build_mixed(ours, texture_sizes, 5)
    cams1 = build_mixed(ref, texture_sizes, 5)
    assert ours.pack() == ref.pack()                         # the dirty flags PackSceneData reports
    compare_packs(ours, ref)
    assert [ours.camera_index(c) for c in cams0] == [ref.camera_index(c) for c in cams1]
    assert len(ours.shapes) == 13 and len(ours.mesh_faces) == 2 * 24 * 12 + 2 * 32 * 32
    # an edit and a re-pack: only what the dirty flags select is rebuilt, and it still agrees
    for sc, cams in ((ours, cams0), (ref, cams1)):
        sc.set_transform(cams[0], position=(1, -9, 3), rotation=(1.1, 0.0, 0.2))
        sc.create_entity(S.ENTITY_CUBE, position=(0, 0, 5), scale=(0.3, 0.3, 0.3))
    assert ours.pack() == ref.pack()
    compare_packs(ours, ref)


def test_default_scene_equals_the_reference(Ref):
    """CreateScene's default checker plane + camera (scene.cpp:912-943)."""
    ours, ref = S.Scene(with_defaults=True), Ref(with_defaults=True)
    ours.pack(), ref.pack()
    compare_packs(ours, ref)
    assert len(ours.shapes) == 1 and len(ours.cameras) == 1 and len(ours.textures) == 1


@pytest.mark.parametrize("faces", ["degenerate-run", "coincident", "tiny"])
def test_build_mesh_node_edge_cases(Ref, faces):
    """BuildMeshNode (scene.cpp:435-599) where its split search has nothing to choose from."""
    if faces == "tiny":
        p = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0.5]], np.float32)
        f = np.array([[0, 1, 2], [1, 3, 2]], np.uint32)
    elif faces == "coincident":                               # many identical triangles: every split has the same cost
        p = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
        f = np.tile(np.array([[0, 1, 2]], np.uint32), (40, 1))
    else:                                                     # a strip along one axis with zero extent on the others
        n = 64
        p = np.stack([np.arange(n + 2, dtype=np.float32), np.zeros(n + 2, np.float32), np.zeros(n + 2, np.float32)], -1)
        f = np.array([[i, i + 1, i + 2] for i in range(n)], np.uint32)
    nrm = np.tile(np.array([0, 0, 1], np.float32), (len(p), 1))
    uv = np.zeros((len(p), 2), np.float32)
    packs = []
    for sc in (S.Scene(), Ref()):
        m = sc.create_mesh("m", p, nrm, uv, f)
        sc.create_entity(S.ENTITY_MESH_INSTANCE, mesh=m, position=(0, 0, 1))
        e = sc.create_entity(S.ENTITY_CAMERA, position=(0, -3, 1), rotation=(1.5, 0, 0))
        sc.pack()
        packs.append(sc)
    compare_packs(*packs)
    assert packs[0].mesh_info() == packs[1].mesh_info()
    assert mesh_faces_of(packs[0], 0).tobytes() == mesh_faces_of(packs[1], 0).tobytes()


@pytest.mark.parametrize("kind", ["field", "torus"])
def test_threaded_mesh_build_equals_the_references_sequential_one(Ref, kind):
    """Large meshes are built on several threads here (host/scene.cpp: subtrees as tasks from 2048 faces per side, the
    passes of a node cut into pieces from 131 072 faces); nodes, their numbering and the face permutation must still
    be the reference's, whose BuildMeshNode is one sequential recursion."""
    p, nrm, uv, f = heightfield(300, 4) if kind == "field" else torus(400, 170)
    assert len(f) >= 131072
    packs = []
    for sc in (S.Scene(), Ref()):
        m = sc.create_mesh("m", p, nrm, uv, f)
        sc.create_entity(S.ENTITY_MESH_INSTANCE, mesh=m, position=(0, 0, 1))
        sc.create_entity(S.ENTITY_CAMERA, position=(0, -3, 1), rotation=(1.5, 0, 0))
        sc.pack()
        packs.append(sc)
    rows_equal("MeshNodePack", packs[0].mesh_nodes, packs[1].mesh_nodes)
    rows_equal("MeshFacePack", packs[0].mesh_faces, packs[1].mesh_faces)
    assert packs[0].mesh_info() == packs[1].mesh_info()
    assert mesh_faces_of(packs[0], 0).tobytes() == mesh_faces_of(packs[1], 0).tobytes()


def write_obj(path, two_materials_in_one_shape):
    p, n, uv, f = torus(16, 8)
    q, _, quv, g = heightfield(9, 1)
    with open(os.path.join(os.path.dirname(path), "m.mtl"), "w") as o:
        o.write("newmtl red\nKd 0.8 0.1 0.1\nnewmtl blue\nKd 0.1 0.1 0.8\n")
    with open(path, "w") as o:
        o.write("mtllib m.mtl\no ring\n")
        for v in p:
            o.write("v %.9g %.9g %.9g\n" % tuple(v))
        for t in uv:
            o.write("vt %.9g %.9g\n" % tuple(t))
        for v in n:
            o.write("vn %.9g %.9g %.9g\n" % tuple(v))
        o.write("usemtl red\n")
        for k, t in enumerate(f):
            if two_materials_in_one_shape and k == len(f) // 2:
                o.write("usemtl blue\n")
            o.write("f " + " ".join("%d/%d/%d" % (i + 1, i + 1, i + 1) for i in t) + "\n")
        o.write("o field\n")                                  # no normals: accumulated from face normals (scene.cpp:640-690)
        for v in q:
            o.write("v %.9g %.9g %.9g\n" % (v[0] + 5, v[1], v[2]))
        for t in quv:
            o.write("vt %.9g %.9g\n" % tuple(t))
        o.write("usemtl blue\n")
        for t in g:
            o.write("f " + " ".join("%d/%d" % (i + 1 + len(p), i + 1 + len(uv)) for i in t) + "\n")


def mesh_faces_of(sc, i):
    import ctypes as C
    m = sc._L.ptbh_scene_mesh(sc.h, i)
    out = np.zeros((sc._L.ptbh_mesh_face_count(m), 3), dtype=np.uint32)
    sc._L.ptbh_mesh_faces(m, out.ctypes.data_as(C.POINTER(C.c_uint32)))
    return out


def test_obj_import_equals_the_reference(Ref, tmp_path):
    """LoadModelAsPrefab (scene.cpp:601-903) through tinyobjloader on the reference side: welded vertices, generated
    normals, one mesh per (shape, material), recentred shapes, prefab instancing; then PackSceneData on the result.
    Materials of the file are OpenPBR in the reference and BasicDiffuse here, so a default material is
    given and the material words are not part of this comparison."""
    write_obj(str(tmp_path / "m.obj"), False)
    packs = []
    for sc in (S.Scene(), Ref()):
        mat = sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "all", BaseColor=(0.5, 0.6, 0.7))
        prefab = sc.load_model(str(tmp_path / "m.obj"), default_material=mat)
        assert sc.prefab_mesh_count(prefab) == 2
        sc.create_prefab_instance(prefab, position=(0.5, 1.0, 0.25), rotation=(0.2, 0.1, 0.9), scale=(1, 1, 1))
        sc.create_prefab_instance(prefab, position=(-4, 0, 1), rotation=(0, 0, 0), scale=(2, 2, 2))
        sc.create_entity(S.ENTITY_CAMERA, position=(0, -6, 2), rotation=(1.4, 0, 0))
        sc.pack()
        packs.append(sc)
    ours, ref = packs
    assert ours.mesh_count() == ref.mesh_count() == 2
    for i in range(2):
        assert ours.source_mesh_vertices(i).tobytes() == ref.source_mesh_vertices(i).tobytes(), i
        assert mesh_faces_of(ours, i).tobytes() == mesh_faces_of(ref, i).tobytes(), i
    for name in ("mesh_faces", "mesh_vertices", "mesh_nodes", "shape_nodes"):
        rows_equal(name, getattr(ours, name), getattr(ref, name))
    a, b = ours.shapes, ref.shapes
    assert len(a) == 4
    for fld in ("Type", "MeshRootNodeIndex", "To", "From"):
        assert a[fld].tobytes() == b[fld].tobytes(), fld


def test_obj_import_two_materials_in_one_shape(Ref, tmp_path):
    """A shape that uses two materials becomes two meshes.  The reference takes the material groups out of a
    std::unordered_set (scene.cpp:773-778), so their ORDER is the standard library's; here it is order of first use
    (host/model.cpp header).  Mesh by mesh, whichever way round, vertices, faces and BVH are the reference's."""
    write_obj(str(tmp_path / "m.obj"), True)
    got = []
    for sc in (S.Scene(), Ref()):
        prefab = sc.load_model(str(tmp_path / "m.obj"))
        assert sc.prefab_mesh_count(prefab) == 3 and sc.mesh_count() == 3
        got.append([(sc.source_mesh_vertices(i).tobytes(), mesh_faces_of(sc, i).tobytes(), sc.mesh_info(i)["nodes"]) for i in range(3)])
    assert sorted(got[0]) == sorted(got[1])
    assert got[0][2] == got[1][2]                            # the single-material shape comes last either way


def test_spectrum_lookup_and_table_columns(Ref):
    """GetParametricSpectrumCoefficients bit for bit on the same table; and the table this repository fits against the
    reference's own OptimizeSpectrum on sampled columns: both fits reproduce their target colour (CIELAB distance to the
    colour the entry stands for, measured by the reference's ObserveSpectrumUnderD65) equally well - including the
    saturated corners where neither can (the sigmoid family has no such spectrum; CIELAB distance 7-12 in both)."""
    rng = np.random.default_rng(2)
    L = H.lib()
    colours = np.concatenate([rng.uniform(0, 1, (3000, 3)), rng.uniform(0, 0.02, (300, 3)), np.eye(3), np.zeros((1, 3)), np.ones((1, 3)),
                              rng.uniform(-0.5, 1.5, (300, 3))]).astype(np.float32)
    import ctypes as C
    for c in colours:
        a, b = (C.c_float * 3)(), (C.c_float * 3)()
        S.lib().ptbh_rgb_to_spectrum(S._f3(c), a)
        L.ptbh_rgb_to_spectrum(S._f3(c), b)
        assert bytes(a) == bytes(b), c
    from path_tracer_b200 import _build
    table = np.fromfile(_build.SPECTRUM_TABLE, dtype=np.float32).reshape(3, 64, 64, 64, 3)     # [L][K][J][I]
    worst = 0.0
    for (i, j, l) in ((0, 0, 0), (63, 63, 2), (20, 41, 1), (63, 0, 0), (7, 55, 2)):
        col = H.spectrum_column(i, j, l)
        for k in range(0, 64, 3):
            r = k / 63.0
            s = r * r * (3 - 2 * r)
            t = s * s * (3 - 2 * s)                                                             # IndexToScale (spectrum.cpp:306-313)
            rgb = np.zeros(3)
            rgb[l] = 1.0
            rgb[(l + 1) % 3] = i / 63.0
            rgb[(l + 2) % 3] = j / 63.0
            rgb = (rgb * t).astype(np.float32)
            e_ref = H.spectrum_lab_error(rgb, col[k])
            e_ours = H.spectrum_lab_error(rgb, table[l, k, j, i])
            assert e_ours <= e_ref + 1e-3, (i, j, l, k, e_ours, e_ref)      # measured: within 1.6e-4 of the reference's own fit
            worst = max(worst, e_ours - e_ref)
    assert worst <= 1e-3


def test_random_scenes_pack_like_the_reference(Ref):
    """A slice of tools/fuzz_pack.py: the random scenes of the traversal and scatter fuzzers (nested non-uniform transforms,
    degenerate meshes, every material, odd-sized textures, three cameras) through both host layers, every packed array
    byte for byte.  It found glm's half conversion (ties away from zero) and the zero signs of T * R * S."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_pack
    assert fuzz_pack.main(["--scenes", "25", "--seed", "3"]) == 0


def test_half_conversion_is_glms(Ref):
    """glm::detail::toFloat16 rounds ties away from zero; IEEE (numpy) rounds them to even."""
    import ctypes as C
    L = S.Scene()._L
    L.ptbh_pack_half2.restype = C.c_uint32
    L.ptbh_pack_half2.argtypes = [C.c_float, C.c_float]
    tie_down = np.uint32(0x3F001000).view(np.float32)        # 0.5 + half an ulp of binary16, even mantissa below: IEEE keeps 0x3800
    tie_up = np.uint32(0x3F003000).view(np.float32)          # odd mantissa below: both go up
    assert L.ptbh_pack_half2(float(tie_down), float(tie_up)) == (0x3801 | (0x3802 << 16))
    assert int(np.float16(tie_down).view(np.uint16)) == 0x3800
    rng = np.random.default_rng(1)
    for v in rng.uniform(-3, 3, 2000).astype(np.float32):    # away from ties the two agree
        if (int(v.view(np.uint32)) & 0x1FFF) != 0x1000:
            assert (L.ptbh_pack_half2(float(v), 0.0) & 0xFFFF) == int(np.float16(v).view(np.uint16))


def test_random_obj_files_import_like_the_reference(Ref):
    """A slice of tools/fuzz_obj.py: random Wavefront files (quads, larger polygons, relative indices, several shapes and
    materials, with and without normals) through both importers.  It found tinyobjloader's triangulation (quads along the
    shorter diagonal, ear clipping beyond) and glm's normalize (v * inversesqrt) in the generated normals."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_obj
    assert fuzz_obj.main(["--files", "60", "--seed", "2"]) == 0


def test_quad_is_cut_along_its_shorter_diagonal(tmp_path):
    """tiny_obj_loader.h:1510-1602: (0,1,2)(0,2,3) if |p2-p0|^2 < |p3-p1|^2, else (0,1,3)(1,2,3)."""
    for k, (p3, want) in enumerate((((0.0, 1.0, 0.0), [(0, 1, 3), (1, 2, 3)]),          # square: equal diagonals -> second form
                                    ((-2.0, 1.0, 0.0), [(0, 1, 2), (0, 2, 3)]))):         # p3 far away: diagonal 1-3 is the long one
        path = str(tmp_path / ("q%d.obj" % k))
        with open(path, "w") as o:
            o.write("v 0 0 0\nv 1 0 0\nv 1 1 0\nv %g %g %g\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\n" % p3)
        sc = S.Scene()
        sc.load_model(path)
        v = sc.source_mesh_vertices(0)[:, :3]
        src = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], p3], np.float32)
        centre = (src.min(0) + src.max(0)) / 2                 # shapes are recentred (scene.cpp:742-768)
        index_of = {tuple(np.round(p - centre, 5)): i for i, p in enumerate(src)}
        faces = sorted(tuple(sorted(index_of[tuple(np.round(v[i], 5))] for i in f)) for f in mesh_faces_of(sc, 0))
        assert faces == sorted(tuple(sorted(f)) for f in want), (k, faces)


def test_scene_files_are_interchangeable_with_the_reference(Ref, tmp_path, monkeypatch):
    """LoadScene / SaveScene (serializer.cpp:501-529) of the reference itself - its serializer.cpp, json.hpp and miniz compiled
    into oracle/_ref/libref_host.so - against host/serializer.cpp: a scene saved by either is loaded by both, and all four
    packed results equal the original's, byte for byte.  Three limits, all the reference's: it does not save mesh vertices
    (serializer.cpp:294-295), so scenes with meshes cannot go through it at all; it does not save SkyboxSamplingProbability
    (:342-349); and its size prefix is `unsigned long`, 8 bytes in this LP64 build and 4 where the reference is built (MSVC) -
    the product writes 4 by default and 8 on request, and reads both."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_pack
    import fuzz_scatter
    monkeypatch.setenv("PTB_SCENE_FILE_SIZE_PREFIX", "8")
    Ours, Theirs = fuzz_pack.plain_cameras(S.Scene, S), fuzz_pack.plain_cameras(Ref, S)
    done = 0
    for seed in range(40):
        pair = []
        for make in (Ours, Theirs):
            sc, _ = fuzz_scatter.random_scene(S, np.random.default_rng(900 + seed), make=make, with_meshes=False)
            pair.append(sc)
        if float(pair[0].globals["SkyboxSamplingProbability"]) != 0.0:
            continue
        compare_packs(*pair)
        d = tmp_path / ("s%d" % seed)
        (d / "ours").mkdir(parents=True)
        (d / "theirs").mkdir(parents=True)
        pair[0].save(str(d / "ours" / "scene.json"))
        pair[1].save(str(d / "theirs" / "scene.json"))
        for written_by in ("ours", "theirs"):
            a = S.Scene.load(str(d / written_by / "scene.json"))
            b = Ref.load(str(d / written_by / "scene.json"))
            a.pack(), b.pack()
            compare_packs(a, pair[1])
            compare_packs(pair[0], b)
            for sc in (a, b):
                sc.close()
        for sc in pair:
            sc.close()
        done += 1
        if done == 8:
            break
    assert done == 8
    # the default prefix is the 4-byte one of the reference's own platform; the product reads it back
    monkeypatch.delenv("PTB_SCENE_FILE_SIZE_PREFIX")
    sc, _ = fuzz_scatter.random_scene(S, np.random.default_rng(5), make=Ours, with_meshes=True)
    sc.save(str(tmp_path / "four" / "scene.json"))
    back = S.Scene.load(str(tmp_path / "four" / "scene.json"))
    back.pack()
    for name in ("shapes", "shape_nodes", "mesh_faces", "mesh_vertices", "mesh_nodes", "cameras", "materials"):
        assert getattr(back, name).tobytes() == getattr(sc, name).tobytes(), name
