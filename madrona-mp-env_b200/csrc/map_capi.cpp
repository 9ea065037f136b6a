// map_capi.cpp -- C-ABI accessors for the host-side map assets (include/mpenv.h,
// "map" section).  No GPU needed: the CPU test-suite uses these to hand the
// SAME BVH the kernels traverse to the oracle (the reference's own BVH comes
// from Embree, which is unavailable; see assets.h).
#include <string>
#include <cstring>

#include "assets.h"
#include "../../include/mpenv.h"

using namespace mpenv;

struct mpenv_map {
    MapAssets assets;
    MapConsts consts;
};

static thread_local std::string g_map_error;

extern "C" {

int mpenv_map_load(const char *scene_dir, mpenv_map **out)
{
    if (!scene_dir || !out) { g_map_error = "null argument"; return MPENV_ERR_INVALID; }
    mpenv_map *m = new mpenv_map();
    std::string err;
    if (!loadMapAssets(scene_dir, &m->assets, &err) || !fillMapConsts(m->assets, &m->consts, &err)) {
        g_map_error = err;
        delete m;
        return MPENV_ERR_IO;
    }
    *out = m;
    return MPENV_OK;
}

void mpenv_map_free(mpenv_map *m) { delete m; }

const char *mpenv_map_last_error(void) { return g_map_error.c_str(); }

int mpenv_map_info(const mpenv_map *m, mpenv_map_info_t *info)
{
    if (!m || !info) return MPENV_ERR_INVALID;
    info->num_nodes = (uint32_t)m->assets.nodes.size();
    info->num_tris = m->assets.numTris;
    info->num_leaves = m->assets.numLeaves;
    memcpy(info->root_aabb, &m->assets.rootAABB, 24);
    memcpy(info->world_bounds, &m->assets.worldBounds, 24);
    info->num_a_spawns = (uint32_t)m->assets.aSpawns.size();
    info->num_b_spawns = (uint32_t)m->assets.bSpawns.size();
    info->num_respawns = (uint32_t)m->assets.respawns.size();
    info->num_zones = (uint32_t)m->assets.zoneBoxes.size();
    return MPENV_OK;
}

const void *mpenv_map_nodes(const mpenv_map *m) { return m ? m->assets.nodes.data() : nullptr; }
const float *mpenv_map_verts(const mpenv_map *m) { return m ? (const float *)m->assets.verts.data() : nullptr; }

int mpenv_map_navmesh(const mpenv_map *m, mpenv_navmesh_t *out)
{
    if (!m || !out) return MPENV_ERR_INVALID;
    if (!m->assets.hasNav) { g_map_error = "navmesh.bin not found in the scene directory"; return MPENV_ERR_IO; }
    const NavmeshAssets &n = m->assets.nav;
    out->num_verts = (uint32_t)n.verts.size();
    out->num_tris = n.numTris();
    out->verts = (const float *)n.verts.data();
    out->tri_indices = n.tris.data();
    out->tri_adjacency = n.adjacency.data();
    out->astar_lookup = n.lookup.data();
    return MPENV_OK;
}

}  // extern "C"

namespace mpenv {
const MapAssets &mapAssetsOf(const mpenv_map *m) { return m->assets; }
const MapConsts &mapConstsOf(const mpenv_map *m) { return m->consts; }
}
