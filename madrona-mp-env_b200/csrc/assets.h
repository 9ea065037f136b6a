// assets.h -- host-side map loading and static-mesh BVH construction.
//
// File formats follow the reference's readers (reference src/map_importer.cpp:
// collisions.bin :35-124 + :223-256, spawns.bin :508-541, zones.bin :543-567).
// The reference builds its BVH with Embree 4 (mesh_bvh_builder.cpp:338-368),
// which is not available here; closest-hit results do not depend on tree
// topology, so this project has its own SAH builder that emits the reference's
// 4-wide quantised node format (mesh_bvh.hpp:61-86) with <= 2 triangles per
// leaf (MADRONA_BLAS_LEAF_WIDTH, mesh_bvh.hpp:16-18).
#pragma once

#include <string>
#include <vector>

#include "mp_defs.h"
#include "navmesh.h"

namespace mpenv {

struct MapAssets {
    AABB worldBounds;
    // static mesh BVH
    std::vector<BVHNode> nodes;
    std::vector<Vec3> verts;      // 3 per triangle, leaf order, padded by one triangle
    uint32_t numTris = 0;
    uint32_t numLeaves = 0;
    AABB rootAABB;
    // spawns / zones
    std::vector<Spawn> aSpawns, bSpawns, respawns;
    std::vector<AABB> zoneBoxes;
    std::vector<float> zoneRots;
    // navmesh + A* table for the scripted bot (optional: navmesh.bin may be absent)
    NavmeshAssets nav;
    bool hasNav = false;
};

// Loads <scene_dir>/{collisions,spawns,zones}.bin and builds the BVH.
// Returns false and fills err on failure.
bool loadMapAssets(const std::string &scene_dir, MapAssets *out, std::string *err);

// Builds the BVH from a de-indexed triangle soup (3 vertices per triangle).
void buildMeshBVH(const std::vector<Vec3> &tri_verts, MapAssets *out);

// Fills the device-constant block from loaded assets (goal regions
// mgr.cpp:913-944, weapon mgr.cpp:1383-1395, frustum sim.cpp:5869-5882,
// lidar direction tables sim.cpp:3347-3350).
bool fillMapConsts(const MapAssets &assets, MapConsts *out, std::string *err);

}  // namespace mpenv
