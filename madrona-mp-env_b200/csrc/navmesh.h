// navmesh.h -- host-side navmesh loading and the A* next-triangle lookup table
// used by the scripted bot (policy id -1; reference planAStarAISystem,
// src/sim.cpp:5041-5172).
//
// Reference pipeline (src/mgr.cpp:1300-1330): importNavmesh
// (map_importer.cpp:421-466) -> meshoptimizer vertex de-duplication ->
// madrona::Navmesh::initFromPolygons -> buildAStarLookup (mgr.cpp:946-1211).
// meshoptimizer and madrona::Navmesh are NOT vendored in the reference tree
// (parity unpinned for their conventions); this file states the
// conventions it uses:
//   * de-duplication: bitwise-equal positions collapse, new indices in order of
//     first appearance (meshopt_generateVertexRemap's documented behaviour);
//   * polygons are fan-triangulated from their first vertex (simple_map's 155
//     faces are all triangles, so face i == triangle i);
//   * triAdjacency[3 t + e] = the first triangle owning the opposite half-edge
//     (v_{e+1} -> v_e), -1 = none.
#pragma once

#include <stdint.h>
#include <string>
#include <vector>

#include "mp_defs.h"

namespace mpenv {

struct NavmeshAssets {
    std::vector<Vec3> verts;          // de-duplicated
    std::vector<uint32_t> tris;       // 3 per triangle
    std::vector<int32_t> adjacency;   // 3 per triangle, -1 = border
    std::vector<uint32_t> lookup;     // numTris x numTris: next triangle from [start][goal]
    uint32_t numTris() const { return (uint32_t)(tris.size() / 3); }
};

// Reads <path> (navmesh.bin: u32 nV, Vec3[nV], u32 nF, u32 counts[nF], u32 nI,
// u32 idx[nI]; map_importer.cpp:421-466), de-duplicates, triangulates, builds
// adjacency and the A* table.  A missing file is not an error for the step
// (bots are optional): returns false with err set, caller decides.
bool loadNavmesh(const std::string &path, NavmeshAssets *out, std::string *err);

// buildAStarLookup (mgr.cpp:1139-1211) without the on-disk cache.
void buildAStarLookup(NavmeshAssets *nav);

}  // namespace mpenv
