// Deterministic procedural scenes for the BASELINE.json configs (SURVEY.md
// §8(d), Appendix D).  Every generator emits a *source description* through the
// scene API, so PackSceneData/BuildMeshNode produce the packed arrays and the
// primitive numbering the reference would produce for the same source data.
#pragma once
#include "scene.hpp"

namespace ptb { namespace host {

// C1: open Cornell-style box of planes / spheres / cubes, diffuse + mirror + glass.
scene* MakeCornellScene();
// C2: height-field terrain of 2*grid*grid triangles (grid=708 -> 1,002,528), one
// diffuse material, ground plane, default white sky, pinhole camera.
scene* MakeTerrainScene(uint32_t Grid = 708);
// C3: textured room (~100k triangles at Detail=1) + refractive spheres, thin
// lens, procedural HDR sky, sky-lobe sampling on.
scene* MakeRoomScene(uint32_t Detail = 1, uint32_t AtlasSize = 4096);
// C4/C5: `Copies` displaced icospheres of subdivision `Level` (20*4^Level
// triangles each; 488 x level 5 = 9,994,240), materials cycled
// diffuse/metal/translucent.  CameraModel: 2 = 360 (C4), 0 = pinhole (C5).
scene* MakeSphereFieldScene(uint32_t Copies = 488, uint32_t Level = 5, int CameraModel = 2);

// Euler ZYX rotation that points a camera (looks down -z, +y up) from `From` to `To` with world z up.
vec3 LookAtRotation(vec3 From, vec3 To);

}}  // namespace ptb::host
