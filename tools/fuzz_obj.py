#!/usr/bin/env python
"""Differential fuzz of the OBJ importer (host/model.cpp) against the REFERENCE'S OWN LoadModelAsPrefab
(scene.cpp:601-903) over its vendored tiny_obj_loader, on random Wavefront files.  CPU only.

Random files: vertices with and without normals and texture coordinates (per face or not at all), triangles, quads and
larger polygons (tinyobjloader triangulates them), absolute and negative (relative) indices, several `o` / `g` blocks,
`usemtl` switches inside a shape, MTL files with texture maps (options, blanks in names, shared and missing files, emission maps), smoothing-group lines, comments, blank lines, trailing spaces, exponents and signs in
numbers, faces that repeat a vertex (zero-area), shapes without faces.  The same file is loaded through both host layers
(oracle/_ref/libref_host.so is the reference's scene.cpp and tiny_obj_loader.h compiled where they lie), the prefab is
instanced, the scene packed, and per mesh - the reference takes a shape's material groups out of a std::unordered_set, so
meshes are matched as a multiset - welded vertices, face order after BuildMeshNode and node counts must be equal byte for
byte.

    python tools/fuzz_obj.py [--files 300] [--seed 1]

Exit code 1 and the path of the file on the first difference."""
import argparse
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def number(rng, v):
    style = rng.integers(0, 5)
    if style == 0:
        return "%.6f" % v
    if style == 1:
        return "%.9g" % v
    if style == 2:
        return "%e" % v
    if style == 3:
        return ("+" if v >= 0 else "") + "%.4f" % v
    if rng.random() < 0.3:                                     # spellings exporters produce now and then: .5  5.  1E+2  -.25e-1  007.50
        t = "%.4f" % v
        neg, t = t.startswith("-"), t.lstrip("-")
        form = rng.integers(0, 5)
        if form == 0 and t.startswith("0."):
            t = t[1:]
        elif form == 1:
            t = "%d." % int(abs(v))
        elif form == 2:
            t = ("%.3E" % abs(v))
        elif form == 3:
            t = ("%.2e" % abs(v)).replace("e-0", "e-").replace("e+0", "e+")
        else:
            t = "00" + t
        return ("-" if neg else "") + t
    return "%d" % int(round(v * 4)) if rng.random() < 0.5 else "%.3f" % v


def write_random_obj(path, rng):
    n_mats = int(rng.integers(0, 4))
    lines = ["# fuzz"]
    if n_mats:
        # texture statements as exporters write them: plain, with options in front (tinyobjloader skips them by argument count),
        # a name with a blank in it, the same file under two materials, a file that is not there, emission maps
        names = ["tex_a.png", "tex b.png", "tex_c.png", "missing.png"]
        with open(os.path.join(os.path.dirname(path), "f.mtl"), "w") as m:
            for k in range(n_mats):
                m.write("newmtl m%d\n" % k)
                if rng.random() < 0.8:
                    m.write("Kd %.3f %.3f %.3f\n" % tuple(rng.uniform(0, 1, 3)))
                for key in ("map_Kd", "map_Ke"):
                    if rng.random() < (0.6 if key == "map_Kd" else 0.25):
                        opts = rng.choice(["", "", "-s 2 2 1 ", "-o 0.5 0 0 -s 1 1 1 ", "-clamp on ", "-blendu off -mm 0 1 ", "-bm 0.5 ", "-colorspace sRGB "])
                        m.write("%s %s%s%s\n" % (key, opts, names[rng.integers(0, len(names))], "  " if rng.random() < 0.2 else ""))
        lines.append("mtllib f.mtl" if rng.random() < 0.8 else "mtllib nothere.mtl f.mtl")
    nv = nt = nn = 0
    for shape in range(int(rng.integers(1, 5))):
        if rng.random() < 0.8:
            lines.append(("o " if rng.random() < 0.6 else "g ") + "shape%d" % shape)
        has_vt, has_vn = rng.random() < 0.6, rng.random() < 0.5
        count = int(rng.integers(3, 40))
        centre = rng.normal(size=3) * 3
        shape_colours = int(rng.choice([0, 0, 0, 1, 2]))
        for _ in range(count):
            v = centre + rng.normal(size=3)
            extra = ""
            if shape_colours:                                  # "v x y z r g b": vertex colours, or a weight as a fourth number
                extra = " " + " ".join("%.3f" % c for c in rng.uniform(0, 1, 3 if shape_colours == 1 else 1))
            lines.append("v " + " ".join(number(rng, x) for x in v) + extra + (" " if rng.random() < 0.1 else ""))
        if rng.random() < 0.15:                                # exact duplicates: welded or not, like the reference
            lines.append(lines[-1])
            count += 1
        ct = int(rng.integers(1, count + 1)) if has_vt else 0
        for _ in range(ct):
            lines.append("vt " + " ".join(number(rng, x) for x in rng.uniform(-0.5, 1.5, 3 if rng.random() < 0.2 else 2)))
        cn = int(rng.integers(1, count + 1)) if has_vn else 0
        for _ in range(cn):
            n = rng.normal(size=3)
            lines.append("vn " + " ".join(number(rng, x) for x in n / max(np.linalg.norm(n), 1e-6)))
        if rng.random() < 0.3:
            lines.append("s %s" % rng.choice(["1", "off", "2"]))
        if rng.random() < 0.1:
            lines.append("")
        for f in range(int(rng.integers(0, 30))):
            if n_mats and rng.random() < 0.25:
                lines.append("usemtl m%d" % rng.integers(0, n_mats + (1 if rng.random() < 0.2 else 0)))      # now and then a name the MTL does not define
            if rng.random() < 0.03:
                lines.append("f %d %d" % (nv + 1, nv + 2))         # a "face" of two corners: dropped
            if rng.random() < 0.08:                            # statements an importer has to step over
                lines.append(rng.choice(["l 1 2 3", "p 1", "usemap off", "mg 1 0.5", "# a comment", "bevel on", "vp 0.5 0.5", "g", "   "]))
            k = int(rng.choice([3, 3, 3, 4, 4, 5, 6]))
            vs = rng.integers(0, count, size=k) if rng.random() < 0.1 else rng.permutation(count)[:k]
            relative = rng.random() < 0.3
            toks = []
            for vi in vs:
                a = int(vi) - count if relative else nv + int(vi) + 1
                t = (int(rng.integers(0, ct)) - ct if relative else nt + int(rng.integers(0, ct)) + 1) if ct else None
                n = (int(rng.integers(0, cn)) - cn if relative else nn + int(rng.integers(0, cn)) + 1) if cn else None
                if t is not None and n is not None:
                    toks.append("%d/%d/%d" % (a, t, n))
                elif t is not None:
                    toks.append("%d/%d" % (a, t))
                elif n is not None:
                    toks.append("%d//%d" % (a, n))
                else:
                    toks.append("%d" % a)
            lines.append("f " + " ".join(toks))
        nv, nt, nn = nv + count, nt + ct, nn + cn
    eol = "\r\n" if rng.random() < 0.25 else "\n"           # files from Windows exporters
    if rng.random() < 0.2:
        lines = [l.replace(" ", "\t") if l[:2] in ("v ", "f ", "vt", "vn") and rng.random() < 0.5 else l for l in lines]
    with open(path, "w", newline="") as o:
        o.write(eol.join(lines) + (eol if rng.random() < 0.9 else ""))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--files", type=int, default=300)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args(argv)
    import test_reference_host as T
    from oracle.ref import host_binding as H
    from path_tracer_b200 import scene as S
    assert H.available(), "oracle/_ref/libref_host.so cannot be built here"
    S.ensure_spectrum_table()
    H.lib()
    Ref = H.ref_scene_class()
    d = tempfile.mkdtemp(prefix="fuzz_obj_")
    from PIL import Image
    for k, name in enumerate(("tex_a.png", "tex b.png", "tex_c.png")):
        Image.fromarray(np.random.default_rng(70 + k).integers(0, 256, size=(5 + 3 * k, 9 - 2 * k, 4), dtype=np.uint8)).save(os.path.join(d, name))
    meshes = faces = refused = 0
    for k in range(args.files):
        rng = np.random.default_rng(args.seed * 100003 + k)
        path = os.path.join(d, "f%05d.obj" % k)
        write_random_obj(path, rng)
        got = []
        for sc in (S.Scene(), Ref()):
            mat = sc.create_material(S.MATERIAL_BASIC_DIFFUSE, "all", BaseColor=(0.5, 0.6, 0.7))
            prefab = sc.load_model(path, default_material=mat)
            if not prefab:
                got.append(None)
                continue
            n = sc.prefab_mesh_count(prefab)
            sc.create_prefab_instance(prefab, position=(0.5, 1.0, 0.25), rotation=(0.2, 0.1, 0.9), scale=(1, 1.5, 1))
            sc.create_entity(S.ENTITY_CAMERA, position=(0, -6, 2), rotation=(1.4, 0, 0))
            sc.pack()
            tex = sc.textures
            texels = []
            for i in range(len(tex)):
                layer = sc.atlas_layer(int(tex["AtlasImageIndex"][i]))
                aw, ah = sc.view.AtlasWidth, sc.view.AtlasHeight
                x0, x1 = int(round(float(tex["AtlasPlacementMinimum"][i][0]) * aw - 0.5)), int(round(float(tex["AtlasPlacementMaximum"][i][0]) * aw + 0.5))
                y0, y1 = int(round(float(tex["AtlasPlacementMaximum"][i][1]) * ah - 0.5)), int(round(float(tex["AtlasPlacementMinimum"][i][1]) * ah + 0.5))
                texels.append((x0, x1, y0, y1, int(tex["AtlasImageIndex"][i]), int(tex["Flags"][i]), layer[y0:y1, x0:x1].tobytes()))
            per_mesh = sorted((sc.source_mesh_vertices(i).tobytes(), T.mesh_faces_of(sc, i).tobytes(), sc.mesh_info(i)["nodes"]) for i in range(sc.mesh_count()))   # mesh::Depth: the reference never writes it
            got.append((n, sc.mesh_count(), per_mesh, len(sc.shapes), len(sc.mesh_faces), len(sc.mesh_vertices), len(sc.mesh_nodes), texels))
            sc.close()
        if got[0] is None or got[1] is None:
            if (got[0] is None) != (got[1] is None):
                print(path, ": one importer refuses the file, the other does not (ours: %s, reference: %s)" % (got[0] is not None, got[1] is not None))
                return 1
            refused += 1
            continue
        if got[0] != got[1]:
            what = [name for name, a, b in zip(("prefab meshes", "scene meshes", "per-mesh vertices / face order / nodes", "shapes", "packed faces", "packed vertices", "packed nodes", "textures: order, atlas placement, texels"), got[0], got[1]) if a != b]
            print(path, ": DIFFERS in", what, "ours", [x for x in got[0] if not isinstance(x, list)], "reference", [x for x in got[1] if not isinstance(x, list)])
            return 1
        meshes += got[0][1]
        faces += got[0][4]
        if (k + 1) % 50 == 0:
            print("%d files: equal (%d meshes, %d faces, %d files refused by both)" % (k + 1, meshes, faces, refused), flush=True)
    print("done: %d files, %d meshes, %d faces, %d refused by both: welded vertices, face order and BVH equal to the reference's" % (args.files, meshes, faces, refused))
    return 0


if __name__ == "__main__":
    sys.exit(main())
