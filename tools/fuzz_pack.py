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
import sys

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
    return Scene


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
