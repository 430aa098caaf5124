#!/usr/bin/env python
"""Differential fuzz of the host scene layer (PackSceneData, BuildMeshNode, ShapeBounds, the shape BVH, material packers, atlas
packing and spectral upsampling of texels) against the REFERENCE'S OWN HOST CODE on random scenes.  CPU only.

oracle/_ref/libref_host.so is the reference's src/scene/scene.cpp (non-Vulkan half), spectrum.cpp, the material packers
and its vendored stb_rect_pack compiled where they lie (oracle/ref/ref_host.cpp).  The random scenes of
tools/fuzz_traversal.py (nested non-uniform transforms, flattened and tiny shapes, triangle soups, slivers, duplicate and
zero-area faces) and of tools/fuzz_scatter.py (every material over its whole parameter range, textures of odd sizes,
textured sky, root medium, three camera models) are built twice with the same calls - through this repository's host
layer and through that library - and every packed array must be equal BYTE FOR BYTE (tests/test_reference_host.py
compare_packs: the fields the reference leaves unwritten are masked there).

    python tools/fuzz_pack.py [--scenes 300] [--seed 1]

Exit code 1 and the seed on the first difference."""
import argparse
import os
import shutil
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def plain_cameras(cls, S):
    """The reference has no look-at helper: cameras get a position and Euler angles, the same through both layers."""
    class Scene(cls):
        def create_camera(self, position, look_at, model=None, params=(90.0, 0.0)):
            model = S.CAMERA_PINHOLE if model is None else model
            rot = tuple(float(x) for x in (np.asarray(look_at, np.float64) - np.asarray(position, np.float64)) * 0.37)
            e = self.create_entity(S.ENTITY_CAMERA, position=tuple(float(x) for x in position), rotation=rot)
            p = list(params) + [0.0] * 5
            self._L.ptbh_camera_set(self.h, e, model, S._f3(p))
            return e

        # handles of what the generators create, for the edit rounds
        def create_entity(self, etype, parent=None, **kw):
            e = super().create_entity(etype, parent=parent, **kw)
            made = self.__dict__.setdefault("made_entities", [])
            owner = next((r for r in reversed(made) if r["handle"] == parent and not r["gone"]), None) if parent is not None else None
            made.append({"handle": e, "type": etype, "owner": owner, "gone": False})      # handles are addresses: a freed one can come back
            return e

        def create_material(self, mtype, name="Material", textures=None, **fields):
            m = super().create_material(mtype, name, textures=textures, **fields)
            self.__dict__.setdefault("made_materials", []).append((m, mtype))
            return m
    return Scene


def edit_round(sc, S, rng):
    """Random edits through the scene API (the dirty flags are the ones the reference's editor sets for the same edits)."""
    ents = sc.__dict__.setdefault("made_entities", [])
    mats = sc.__dict__.setdefault("made_materials", [])

    def is_gone(r):                                            # destroyed itself, or inside a destroyed container (any depth)
        while r is not None:
            if r["gone"]:
                return True
            r = r["owner"]
        return False

    def living():
        return [r for r in ents if not is_gone(r)]
    alive = living()
    for _ in range(int(rng.integers(1, 5))):
        what = rng.integers(0, 6)
        if what == 0 and alive:
            e = alive[rng.integers(0, len(alive))]["handle"]
            sc.set_transform(e, position=tuple(rng.normal(size=3) * 3), rotation=tuple(rng.uniform(-3, 3, 3)), scale=tuple(rng.uniform(0.3, 2.0, 3)))
        elif what == 1 and alive:
            sc.set_active(alive[rng.integers(0, len(alive))]["handle"], bool(rng.integers(0, 2)))
        elif what == 2 and alive:
            # Containers are left alone: the reference's DestroyEntity walks Entity->Children while every child erases itself from that
            # very vector (scene.cpp:256-268) - undefined behaviour, a crash in this build - so a container with children cannot be
            # destroyed through it.  host/scene.cpp copies the list first (tests/test_host_layer.py, test_destroying_a_container_removes_everything_in_it).
            leaves = [r for r in alive if r["type"] != S.ENTITY_CONTAINER]
            if leaves:
                r = leaves[rng.integers(0, len(leaves))]
                sc.destroy_entity(r["handle"])
                r["gone"] = True
                alive = living()
        elif what == 3:
            parents = [None] + [r["handle"] for r in alive if r["type"] == S.ENTITY_CONTAINER]
            sc.create_entity((S.ENTITY_SPHERE, S.ENTITY_CUBE, S.ENTITY_PLANE)[rng.integers(0, 3)], parent=parents[rng.integers(0, len(parents))],
                             position=tuple(rng.normal(size=3) * 2), rotation=tuple(rng.uniform(-3, 3, 3)), scale=tuple(rng.uniform(0.3, 2.0, 3)),
                             material=mats[rng.integers(0, len(mats))][0] if mats else None)
            alive = living()
        elif what == 4 and mats:
            m, mtype = mats[rng.integers(0, len(mats))]
            if mtype in (S.MATERIAL_BASIC_DIFFUSE, S.MATERIAL_BASIC_METAL):
                sc._L.ptbh_material_set(sc.h, m, b"BaseColor", S._f3(rng.uniform(0, 1, 3)))
            else:
                sc._L.ptbh_material_set(sc.h, m, b"Roughness", S._f3([float(rng.uniform(0, 1)), 0, 0]))
        else:
            sc.set_root(scatter_rate=float(rng.choice([0.0, 0.1])), sky_brightness=float(rng.uniform(0.5, 2)), sky_sampling_probability=0.0, sky_texture=None)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", type=int, default=300)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args(argv)
    import fuzz_scatter as FS
    import fuzz_traversal as FT
    import test_reference_host as T
    from oracle.ref import host_binding as H
    from path_tracer_b200 import scene as S
    assert H.available(), "oracle/_ref/libref_host.so cannot be built here"
    S.ensure_spectrum_table()
    H.lib()
    Ours, Ref = plain_cameras(S.Scene, S), plain_cameras(H.ref_scene_class(), S)
    faces = shapes = 0
    tmp = tempfile.mkdtemp(prefix="fuzz_pack_")
    for k in range(args.scenes):
        seed = args.seed * 100003 + k
        for which in ("geometry", "geometry-extreme", "materials"):
            pair = []
            for make in (Ours, Ref):
                rng = np.random.default_rng(seed)
                if which == "materials":
                    sc, _ = FS.random_scene(S, rng, make=make)
                else:
                    sc = FT.random_scene(S, rng, which.endswith("extreme"), make=make)
                pair.append(sc)
            try:
                T.compare_packs(*pair)
                n_meshes = [sc._L.ptbh_scene_mesh_count(sc.h) for sc in pair]
                assert n_meshes[0] == n_meshes[1]
                for i in range(n_meshes[0]):
                    assert pair[0].mesh_info(i) == pair[1].mesh_info(i), ("mesh", i)
                    assert T.mesh_faces_of(pair[0], i).tobytes() == T.mesh_faces_of(pair[1], i).tobytes(), ("face order of mesh", i)
            except AssertionError as ex:
                print("seed", seed, which, "PACKED ARRAYS DIFFER:", str(ex)[:1500])
                return 1
            # edit rounds: only what the dirty flags select is repacked, and it must still agree
            for r in range(3):
                flags = []
                for sc in pair:
                    edit_round(sc, S, np.random.default_rng(seed * 7 + r))
                    flags.append(sc.pack())
                try:
                    assert flags[0] == flags[1], ("dirty flags reported by PackSceneData", flags)
                    T.compare_packs(*pair)
                except AssertionError as ex:
                    print("seed", seed, which, "after edit round", r, "PACKED ARRAYS DIFFER:", str(ex)[:1500])
                    return 1
            # the product's own scene file (with meshes, which the reference cannot carry): save -> load -> pack is the identity
            if k % 4 == 0:
                file = os.path.join(tmp, "s%d_%s" % (seed, which), "scene.json")
                pair[0].save(file)
                back = S.Scene.load(file)
                back.pack()
                for name in ("shapes", "shape_nodes", "mesh_faces", "mesh_vertices", "mesh_nodes", "cameras", "materials", "textures"):
                    if getattr(back, name).tobytes() != getattr(pair[0], name).tobytes():
                        print("seed", seed, which, "SAVE / LOAD ROUND TRIP changes", name)
                        return 1
                back.close()
                shutil.rmtree(os.path.dirname(file))
            faces += len(pair[0].mesh_faces)
            shapes += len(pair[0].shapes)
            for sc in pair:
                sc.close()
        if (k + 1) % 20 == 0:
            print("%d seeds x 3 generators: every packed array equal (%d shapes, %d mesh faces so far)" % (k + 1, shapes, faces), flush=True)
    print("done: %d seeds x 3 generators, %d shapes, %d mesh faces: every packed array equal to the reference's, byte for byte" % (args.scenes, shapes, faces))
    return 0


if __name__ == "__main__":
    sys.exit(main())
