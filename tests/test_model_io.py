"""Model import (LoadModelAsPrefab, src/scene/scene.cpp:601-903), prefab instancing, texture and image
files — host logic, CPU only.  The imported scene is then traced by the oracle to check that the packed
arrays are a usable input for the hot path."""
import struct
import zlib

import numpy as np
import pytest

OBJ = """# two shapes: a unit cube given as quads (two materials), a textured quad with vt; no normals in the file
mtllib model.mtl
o Box
v -1 -1 -1
v  1 -1 -1
v  1  1 -1
v -1  1 -1
v -1 -1  1
v  1 -1  1
v  1  1  1
v -1  1  1
usemtl red
f 1 4 3 2
f 5 6 7 8
f 1 2 6 5
usemtl green
f 2 3 7 6
f 3 4 8 7
f 4 1 5 8
o Card
v 4 -1 0
v 6 -1 0
v 6  1 0
v 4  1 0
vt 0 0
vt 1 0
vt 1 1
vt 0 1
usemtl card
f -4/1 -3/2 -2/3 -1/4
"""
MTL = """newmtl red
Kd 0.8 0.1 0.1
newmtl green
Kd 0.1 0.8 0.1
newmtl card
Kd 1 1 1
map_Kd card.ppm
"""


@pytest.fixture()
def model_dir(tmp_path):
    (tmp_path / "model.obj").write_text(OBJ)
    (tmp_path / "model.mtl").write_text(MTL)
    px = np.array([[[255, 0, 0], [0, 255, 0]], [[0, 0, 255], [188, 188, 188]]], dtype=np.uint8)
    with open(tmp_path / "card.ppm", "wb") as f:
        f.write(b"P6\n# comment\n2 2\n255\n" + px.tobytes())
    return tmp_path


def test_load_model_as_prefab(model_dir, oracle):
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    sc = S.Scene()
    prefab = sc.load_model(str(model_dir / "model.obj"))
    assert sc.prefab_mesh_count(prefab) == 3 and sc.mesh_count() == 3          # Box/red, Box/green, Card
    assert [sc.mesh_info(i)["faces"] for i in range(3)] == [6, 6, 2]
    # welding: each material group of the cube (bottom, top, one side / the other three sides) touches all
    # 8 corners once although its 6 triangles have 18 corners; the card has 4 vertices
    assert [len(sc.source_mesh_vertices(i)) for i in range(3)] == [8, 8, 4]
    v = sc.source_mesh_vertices(0)
    # shapes are recentred on their bounding-box centre; the cube's is the origin already
    assert np.abs(v[:, :3]).max() == 1.0
    # no normals in the file for the cube: sum of the unit normals of the faces around each corner, normalised
    assert np.allclose(np.linalg.norm(v[:, 3:6], axis=1), 1.0, atol=1e-6)
    # (a quad is two triangles, so a face counts once or twice at a corner: components are in ratio 1:2 at most)
    w = np.abs(v[:, 3:6]) / np.abs(v[:, 3:6]).min(axis=1, keepdims=True)
    assert np.allclose(w, np.round(w), atol=1e-5) and w.max() <= 2.0
    assert (np.sign(v[:, 3:6]) == np.sign(v[:, :3])).all()                     # outward
    card = sc.source_mesh_vertices(2)
    assert np.allclose(card[:, :3].mean(axis=0), 0.0, atol=1e-6)               # recentred (file centre is x = 5)
    assert np.array_equal(card[:, 3:6], np.tile([0, 0, 1], (4, 1))) and set(map(tuple, card[:, 6:8])) == {(0, 0), (1, 0), (1, 1), (0, 1)}

    # nothing is in the scene until the prefab is instantiated (scene.cpp:253-258); instances share meshes
    sc.create_camera((0, -8, 2), (0, 0, 0))
    sc.pack()
    assert len(sc.shapes) == 0
    sc.create_prefab_instance(prefab)
    sc.create_prefab_instance(prefab, position=(0, 0, 5))
    sc.pack()
    assert len(sc.shapes) == 6 and len(sc.mesh_faces) == 14 and sc.mesh_count() == 3
    # the card instance sits where the file put it (x = 5), the second copy 5 units higher
    rays = oracle.make_rays(np.array([[5, 0, 3], [5, 0, 8], [0, 0, 3], [20, 0, 3]], np.float32), np.array([[0, 0, -1]] * 4, np.float32))
    hits = oracle.raycast(sc.view, rays, tie=oracle.TIE_CANONICAL)
    assert hits["shape"][3] == 0xFFFFFFFF and (hits["shape"][:3] != 0xFFFFFFFF).all()
    assert np.allclose(hits["t"][:3], [3.0, 3.0, 2.0], atol=1e-6)
    bf = oracle.raycast_bruteforce(sc.view, rays)
    assert (hits["shape"] == bf["shape"]).all() and (hits["prim"] == bf["prim"]).all()
    # the card's material carries the texture from map_Kd: 2x2 PPM, sRGB-decoded, converted to spectra in the atlas
    assert len(sc.textures) == 1
    # materials: Kd became BaseColor of BasicDiffuse slots
    assert len(sc.materials) >= 3 * 32
    sc.destroy_prefab(prefab)
    sc.close()


def test_file_normals_are_used_and_missing_ones_stay_zero(model_dir):
    """With any `vn` in the file nothing is recomputed (scene.cpp:615): corners without a normal index keep (0,0,0)."""
    from path_tracer_b200 import scene as S
    (model_dir / "n.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\nvn 0 0 -1\nf 1//1 3//1 2//1\nf 1 2 4\n")
    sc = S.Scene()
    sc.load_model(str(model_dir / "n.obj"))
    v = sc.source_mesh_vertices(0)
    assert sc.mesh_info(0)["faces"] == 2 and len(v) == 6          # same positions, different normals: not welded
    assert sorted(map(tuple, v[:, 3:6])) == [(0, 0, -1)] * 3 + [(0, 0, 0)] * 3
    sc.close()


def test_load_model_errors_and_options(model_dir):
    from path_tracer_b200 import scene as S
    sc = S.Scene()
    with pytest.raises(OSError):
        sc.load_model(str(model_dir / "missing.obj"))
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = m[1, 1] = m[2, 2] = 2.0                                            # uniform scale 2 (column-major = symmetric here)
    sc.load_model(str(model_dir / "model.obj"), name="Scaled", vertex_transform=m)
    assert np.abs(sc.source_mesh_vertices(0)[:, :3]).max() == 2.0
    sc.close()


def test_ppm_and_pfm_textures(model_dir, tmp_path):
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    sc = S.Scene()
    sc.set_atlas_size(64, 64)
    sc.load_texture(str(model_dir / "card.ppm"), S.TEXTURE_RAW)
    lin = np.arange(2 * 3 * 3, dtype=np.float32).reshape(2, 3, 3) / 7.0
    S.write_pfm(str(tmp_path / "t.pfm"), lin)
    sc.load_texture(str(tmp_path / "t.pfm"), S.TEXTURE_RAW)
    with pytest.raises(OSError):
        sc.load_texture(str(model_dir / "model.obj"), S.TEXTURE_RAW)
    sc.pack()
    layer = sc.atlas_layer(0)
    tex = sc.textures
    assert len(tex) == 2

    def texels(t, w, h):
        x0 = int(round(t["AtlasPlacementMinimum"][0] * 64 - 0.5))
        y0 = int(round(t["AtlasPlacementMaximum"][1] * 64 - 0.5))
        return layer[y0:y0 + h, x0:x0 + w]
    a = texels(tex[0], 2, 2)
    assert np.allclose(a[0, 0], [1, 0, 0, 1]) and np.allclose(a[1, 0], [0, 0, 1, 1])
    assert np.allclose(a[1, 1, :3], (188 / 255.0) ** 2.2, atol=1e-6)              # stbi_loadf's ldr_to_hdr, as for every 8-bit format
    b = texels(tex[1], 3, 2)
    assert np.array_equal(b[..., :3], lin) and (b[..., 3] == 1).all()             # PFM round trip, row 0 at the top
    sc.close()


def test_png_writer_is_a_valid_png(tmp_path):
    from path_tracer_b200 import scene as S
    rng = np.random.default_rng(1)
    for w, h in ((7, 5), (300, 120)):                                            # the second needs several stored blocks
        img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
        path = tmp_path / ("i%d.png" % w)
        S.write_png(str(path), img)
        data = path.read_bytes()
        assert data[:8] == b"\x89PNG\r\n\x1a\n"
        pos, chunks = 8, []
        while pos < len(data):
            n, = struct.unpack(">I", data[pos:pos + 4])
            kind, body = data[pos + 4:pos + 8], data[pos + 8:pos + 8 + n]
            crc, = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
            assert crc == zlib.crc32(kind + body)
            chunks.append((kind, body))
            pos += 12 + n
        assert [k for k, _ in chunks] == [b"IHDR", b"sRGB", b"IDAT", b"IEND"]
        assert struct.unpack(">IIBBBBB", chunks[0][1]) == (w, h, 8, 6, 0, 0, 0)
        raw = np.frombuffer(zlib.decompress(chunks[2][1]), dtype=np.uint8).reshape(h, 1 + 4 * w)
        assert (raw[:, 0] == 0).all() and np.array_equal(raw[:, 1:].reshape(h, w, 4), img)


def _packs_equal(a, b):
    for name in ("shapes", "shape_nodes", "mesh_faces", "mesh_vertices", "mesh_nodes", "textures", "cameras", "materials"):
        x, y = getattr(a, name), getattr(b, name)
        assert x.tobytes() == y.tobytes(), name
    assert a.globals.tobytes() == b.globals.tobytes()
    assert a.view.AtlasLayerCount == b.view.AtlasLayerCount
    for i in range(a.view.AtlasLayerCount):
        assert np.array_equal(a.atlas_layer(i), b.atlas_layer(i))


@pytest.mark.parametrize("name,params", [("cornell", ()), ("room", (1, 256)), ("spherefield", (6, 1, 2))])
def test_scene_file_round_trip_is_byte_identical(tmp_path, name, params):
    """SaveScene -> LoadScene -> PackSceneData gives the same packed arrays, byte for byte (every input of the hot path)."""
    import json
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    a = S.Scene.make(name, *params)
    path = tmp_path / "saved" / "scene.json"
    a.save(str(path))
    doc = json.loads(path.read_text())                       # a real JSON document with the reference's top-level keys
    assert set(doc) == {"Textures", "Materials", "Meshes", "Prefabs", "Root"}
    assert doc["Root"]["Type"] == 0 and isinstance(doc["Root"]["Children"], list)
    assert (tmp_path / "saved" / "spectrum.dat").exists()
    b = S.Scene.load(str(path))
    b.set_atlas_size(a.view.AtlasWidth, a.view.AtlasHeight)
    b.pack()
    _packs_equal(a, b)
    b.save(str(tmp_path / "again" / "scene.json"))           # saving the loaded scene reproduces the document
    assert (tmp_path / "again" / "scene.json").read_text() == path.read_text()
    a.close()
    b.close()


def test_scene_file_with_prefabs_and_side_files(model_dir, tmp_path):
    import struct as st
    import zlib as z
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    a = S.Scene(with_defaults=True)
    a.set_atlas_size(128, 128)
    prefab = a.load_model(str(model_dir / "model.obj"))
    a.create_prefab_instance(prefab, position=(0, 0, 1))
    a.pack()
    a.save(str(tmp_path / "s.json"))
    # side-file layout (serializer.cpp:271-313): 16-byte header, then size-prefixed zlib streams
    assert (tmp_path / "Box_2.mesh").exists()                # the shape's second material group: renamed, not overwritten
    raw = (tmp_path / "Box.mesh").read_bytes()
    magic, version, faces, nodes = st.unpack("<IIII", raw[:16])
    assert magic == 0x4D455348 and version == 0 and faces == 6
    n, = st.unpack("<I", raw[16:20])
    assert len(z.decompress(raw[20:20 + n])) == 12 * faces
    b = S.Scene.load(str(tmp_path / "s.json"))
    b.set_atlas_size(128, 128)
    b.pack()
    _packs_equal(a, b)
    # a mesh file as the reference writes it (no vertex block) is refused with a message, not rendered wrongly
    m, = st.unpack("<I", raw[20 + n:24 + n])
    (tmp_path / "Box.mesh").write_bytes(raw[:24 + n + m])
    with pytest.raises(OSError, match="vertex block"):
        S.Scene.load(str(tmp_path / "s.json"))
    with pytest.raises(OSError):
        S.Scene.load(str(tmp_path / "missing.json"))
    (tmp_path / "bad.json").write_text('{"Textures": [}')
    with pytest.raises(OSError):
        S.Scene.load(str(tmp_path / "bad.json"))
    a.close()
    b.close()


def test_mtl_is_read_the_way_tinyobjloader_reads_it(tmp_path):
    """Defaults and quirks of the reference's loader that decide what a model looks like: a material without Kd is black, unless it
    has a map_Kd and no Kd statement has been seen IN THE FILE so far (then 0.6); options in front of a texture name are skipped by
    argument count; a texture name runs to the end of the line; material names are one word (newmtl and usemtl alike)."""
    from path_tracer_b200 import scene as S
    from PIL import Image
    S.ensure_spectrum_table()
    Image.fromarray(np.full((2, 2, 3), 200, np.uint8)).save(tmp_path / "a b.png")
    (tmp_path / "m.mtl").write_text("newmtl first extra words\nmap_Kd -s 2 2 1 -clamp on a b.png\n"       # no Kd seen yet: 0.6, texture found
                                    "newmtl second\n"                                                     # nothing: black
                                    "newmtl third\nKd 0.25 0.5 0.75\n"
                                    "newmtl fourth\nmap_Kd a b.png\n")                                   # a Kd was seen before in the file: stays black
    (tmp_path / "m.obj").write_text("mtllib m.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nvt 0 0\nvt 1 0\nvt 0 1\n"
                                    "usemtl first ignored\nf 1/1 2/2 3/3\nusemtl second\nf 2/1 4/2 3/3\nusemtl third\nf 1/1 2/2 4/3\nusemtl fourth\nf 1/1 4/2 3/3\n")
    sc = S.Scene()
    sc.set_atlas_size(64, 64)
    prefab = sc.load_model(str(tmp_path / "m.obj"))
    assert sc.prefab_mesh_count(prefab) == 4
    sc.create_prefab_instance(prefab, position=(0, 0, 0))
    sc.pack()
    want = S.Scene()
    want.set_atlas_size(64, 64)
    t = want.load_texture(str(tmp_path / "a b.png"), S.TEXTURE_REFLECTANCE_WITH_ALPHA)
    want.create_material(S.MATERIAL_BASIC_DIFFUSE, "first", BaseColor=(0.6, 0.6, 0.6), textures={"BaseTexture": t})
    want.create_material(S.MATERIAL_BASIC_DIFFUSE, "second", BaseColor=(0, 0, 0))
    want.create_material(S.MATERIAL_BASIC_DIFFUSE, "third", BaseColor=(0.25, 0.5, 0.75))
    want.create_material(S.MATERIAL_BASIC_DIFFUSE, "fourth", BaseColor=(0, 0, 0), textures={"BaseTexture": t})
    want.pack()
    assert len(sc.textures) == 1
    assert np.asarray(sc.materials).tobytes() == np.asarray(want.materials).tobytes()
