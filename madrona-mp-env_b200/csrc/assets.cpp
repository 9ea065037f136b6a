#include <cfloat>
// assets.cpp -- see assets.h.
#include "assets.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <numeric>

#include "mp_math.h"

namespace mpenv {

namespace {

bool readFile(const std::string &path, std::vector<char> *out)
{
    std::ifstream f(path, std::ios::binary);
    if (!f.is_open()) return false;
    f.seekg(0, f.end);
    size_t n = (size_t)f.tellg();
    f.seekg(0, f.beg);
    out->resize(n);
    f.read(out->data(), (std::streamsize)n);
    return (size_t)f.gcount() == n;
}

struct Cursor {
    const std::vector<char> &buf;
    size_t off = 0;
    bool ok = true;
    template <typename T> void take(T *dst, size_t count = 1)
    {
        size_t n = sizeof(T) * count;
        if (!ok || off + n > buf.size()) { ok = false; return; }
        memcpy(dst, buf.data() + off, n);
        off += n;
    }
    void skip(size_t n)
    {
        if (!ok || off + n > buf.size()) { ok = false; return; }
        off += n;
    }
};

// collisions.bin: AABB, materials (u64 n, u64 nameBytes, names, u32 flags[n]),
// meshes (u64 nMesh, nVert, nTri, Vector3[nVert], u32[3*nTri], u32 mats[nTri],
// MeshInfo{vOff,nV,tOff,nT}[nMesh]).  Triangles whose material is BulletsOnly
// (flag value 1) are dropped and the rest de-indexed, mesh by mesh, with
// mesh-local indices (map_importer.cpp:126-221).
bool loadCollisions(const std::string &path, AABB *world_bounds, std::vector<Vec3> *tris,
                    std::string *err)
{
    std::vector<char> buf;
    if (!readFile(path, &buf)) { *err = "failed to open " + path; return false; }
    Cursor c { buf };
    c.take(world_bounds);
    uint64_t num_mats = 0, num_name_bytes = 0;
    c.take(&num_mats);
    c.take(&num_name_bytes);
    c.skip(num_name_bytes);
    if (!c.ok || num_mats > (1u << 20)) { *err = "corrupt materials in " + path; return false; }
    std::vector<uint32_t> mat_flags(num_mats);
    c.take(mat_flags.data(), num_mats);
    uint64_t num_meshes = 0, num_verts = 0, num_tris = 0;
    c.take(&num_meshes);
    c.take(&num_verts);
    c.take(&num_tris);
    if (!c.ok || num_verts > (1ull << 28) || num_tris > (1ull << 28) || num_meshes > (1ull << 24)) {
        *err = "corrupt mesh header in " + path; return false;
    }
    std::vector<Vec3> verts(num_verts);
    std::vector<uint32_t> indices(num_tris * 3), tri_mats(num_tris);
    struct MeshInfo { uint32_t vOff, nV, tOff, nT; };
    std::vector<MeshInfo> meshes(num_meshes);
    c.take(verts.data(), num_verts);
    c.take(indices.data(), num_tris * 3);
    c.take(tri_mats.data(), num_tris);
    c.take(meshes.data(), num_meshes);
    if (!c.ok) { *err = "truncated " + path; return false; }

    tris->clear();
    for (const MeshInfo &m : meshes) {
        for (uint32_t i = 0; i < m.nT; i++) {
            uint64_t t = (uint64_t)m.tOff + i;
            if (t >= num_tris) { *err = "triangle index out of range in " + path; return false; }
            uint32_t mat = tri_mats[t];
            if (mat < num_mats && mat_flags[mat] == 1u) continue;   // BulletsOnly
            for (int k = 0; k < 3; k++) {
                uint64_t vi = (uint64_t)m.vOff + indices[3 * t + k];
                if (vi >= num_verts) { *err = "vertex index out of range in " + path; return false; }
                tris->push_back(verts[vi]);
            }
        }
    }
    if (tris->empty()) { *err = "no collision triangles in " + path; return false; }
    return true;
}

bool loadSpawns(const std::string &path, MapAssets *out, std::string *err)
{
    std::vector<char> buf;
    if (!readFile(path, &buf)) { *err = "failed to open " + path; return false; }
    Cursor c { buf };
    std::vector<Spawn> *lists[3] = { &out->aSpawns, &out->bSpawns, &out->respawns };
    for (auto *l : lists) {
        uint32_t n = 0;
        c.take(&n);
        if (!c.ok || n > kMaxSpawns) { *err = "bad spawn count in " + path; return false; }
        l->resize(n);
        c.take(l->data(), n);
    }
    if (!c.ok) { *err = "truncated " + path; return false; }
    if (out->aSpawns.empty() || out->bSpawns.empty()) {
        *err = "map needs at least one spawn per team: " + path; return false;
    }
    return true;
}

bool loadZones(const std::string &path, MapAssets *out, std::string *err)
{
    std::vector<char> buf;
    if (!readFile(path, &buf)) { *err = "failed to open " + path; return false; }
    Cursor c { buf };
    uint32_t n = 0;
    c.take(&n);
    if (!c.ok || n == 0 || n > kMaxZones) { *err = "bad zone count in " + path; return false; }
    out->zoneBoxes.resize(n);
    out->zoneRots.resize(n);
    c.take(out->zoneBoxes.data(), n);
    c.take(out->zoneRots.data(), n);
    if (!c.ok) { *err = "truncated " + path; return false; }
    return true;
}

// ------------------------------------------------------------- BVH builder
struct Box {
    float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
    float hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    void grow(const Vec3 &p)
    {
        const float v[3] = { p.x, p.y, p.z };
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], v[a]); hi[a] = std::max(hi[a], v[a]); }
    }
    void grow(const Box &b)
    {
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); }
    }
    float area() const
    {
        float d[3] = { hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2] };
        if (d[0] < 0.f) return 0.f;
        return 2.f * (d[0] * d[1] + d[1] * d[2] + d[2] * d[0]);
    }
};

struct BinNode {
    Box box;
    int left = -1, right = -1;     // children (binary)
    int first = 0, count = 0;      // leaf range in the ordered triangle list
    bool leaf() const { return left < 0; }
};

struct Builder {
    const std::vector<Vec3> &tv;
    std::vector<Box> triBox;
    std::vector<float> centroid[3];
    std::vector<int> order;
    std::vector<BinNode> bin;

    explicit Builder(const std::vector<Vec3> &t) : tv(t)
    {
        int n = (int)t.size() / 3;
        triBox.resize(n);
        for (int a = 0; a < 3; a++) centroid[a].resize(n);
        for (int i = 0; i < n; i++) {
            for (int k = 0; k < 3; k++) triBox[i].grow(t[3 * i + k]);
            for (int a = 0; a < 3; a++) centroid[a][i] = 0.5f * (triBox[i].lo[a] + triBox[i].hi[a]);
        }
        order.resize(n);
        std::iota(order.begin(), order.end(), 0);
    }

    int build(int first, int count)
    {
        int id = (int)bin.size();
        bin.emplace_back();
        Box box;
        for (int i = 0; i < count; i++) box.grow(triBox[order[first + i]]);
        bin[id].box = box;
        bin[id].first = first;
        bin[id].count = count;
        if (count <= 2) return id;

        // Full-sweep SAH over the three axes.
        float best_cost = FLT_MAX;
        int best_axis = -1, best_split = -1;
        std::vector<int> idx(count);
        std::vector<float> right_area(count);
        for (int axis = 0; axis < 3; axis++) {
            for (int i = 0; i < count; i++) idx[i] = order[first + i];
            std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) {
                return centroid[axis][a] < centroid[axis][b];
            });
            Box acc;
            for (int i = count - 1; i > 0; i--) { acc.grow(triBox[idx[i]]); right_area[i] = acc.area(); }
            Box left;
            for (int i = 1; i < count; i++) {
                left.grow(triBox[idx[i - 1]]);
                float cost = left.area() * (float)i + right_area[i] * (float)(count - i);
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = i; }
            }
        }
        for (int i = 0; i < count; i++) idx[i] = order[first + i];
        std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) {
            return centroid[best_axis][a] < centroid[best_axis][b];
        });
        for (int i = 0; i < count; i++) order[first + i] = idx[i];

        int l = build(first, best_split);
        int r = build(first + best_split, count - best_split);
        bin[id].left = l;
        bin[id].right = r;
        return id;
    }
};

// Smallest exponent e with 255 * 2^e >= extent (clamped so that the IEEE
// exponent field (e + 127) stays valid and e fits an int8).
int8_t chooseExp(float extent)
{
    if (!(extent > 0.f)) return -100;
    int e = (int)std::ceil(std::log2((double)extent / 255.0));
    while (255.0 * std::ldexp(1.0, e) < (double)extent) e++;
    e = std::max(-100, std::min(100, e));
    return (int8_t)e;
}

float expScale(int8_t e) { return u2f((uint32_t)((int32_t)e + 127) << 23); }

}  // namespace

void buildMeshBVH(const std::vector<Vec3> &tri_verts, MapAssets *out)
{
    const int num_tris = (int)tri_verts.size() / 3;
    Builder b(tri_verts);
    b.bin.reserve((size_t)num_tris * 2 + 2);
    int root = b.build(0, num_tris);

    // De-index in leaf order.
    out->verts.resize((size_t)num_tris * 3 + 3);
    for (int i = 0; i < num_tris; i++) {
        for (int k = 0; k < 3; k++) out->verts[(size_t)3 * i + k] = tri_verts[(size_t)3 * b.order[i] + k];
    }
    // sphereCastLeaf reads two triangles from every leaf (mesh_bvh.inl:867):
    // pad with a copy of the last triangle so the last leaf's read is defined.
    for (int k = 0; k < 3; k++) out->verts[(size_t)3 * num_tris + k] = out->verts[(size_t)3 * (num_tris - 1) + k];
    out->numTris = (uint32_t)num_tris;

    Box rb = b.bin[root].box;
    out->rootAABB.pMin = { rb.lo[0], rb.lo[1], rb.lo[2] };
    out->rootAABB.pMax = { rb.hi[0], rb.hi[1], rb.hi[2] };

    // Collapse the binary tree into 4-wide nodes: start from a node's two
    // children and keep opening the internal child with the largest surface
    // area until four slots are used.
    out->nodes.clear();
    out->numLeaves = 0;
    struct Pending { int bin; int wide; };
    std::vector<Pending> queue;
    out->nodes.emplace_back();
    queue.push_back({ root, 0 });
    std::vector<int> parent_of(1, -1);
    for (size_t qi = 0; qi < queue.size(); qi++) {
        Pending p = queue[qi];
        std::vector<int> kids;
        if (b.bin[p.bin].leaf()) {
            kids.push_back(p.bin);      // degenerate: the whole mesh is one leaf
        } else {
            kids.push_back(b.bin[p.bin].left);
            kids.push_back(b.bin[p.bin].right);
            while (kids.size() < 4) {
                int best = -1;
                float best_area = -1.f;
                for (size_t k = 0; k < kids.size(); k++) {
                    if (b.bin[kids[k]].leaf()) continue;
                    float a = b.bin[kids[k]].box.area();
                    if (a > best_area) { best_area = a; best = (int)k; }
                }
                if (best < 0) break;
                int open = kids[best];
                kids[best] = b.bin[open].left;
                kids.insert(kids.begin() + best + 1, b.bin[open].right);
            }
        }

        BVHNode node;
        memset(&node, 0, sizeof(node));
        Box nb;
        for (int k : kids) nb.grow(b.bin[k].box);
        node.minX = nb.lo[0]; node.minY = nb.lo[1]; node.minZ = nb.lo[2];
        node.expX = chooseExp(nb.hi[0] - nb.lo[0]);
        node.expY = chooseExp(nb.hi[1] - nb.lo[1]);
        node.expZ = chooseExp(nb.hi[2] - nb.lo[2]);
        const float node_min[3] = { node.minX, node.minY, node.minZ };
        const float scale[3] = { expScale(node.expX), expScale(node.expY), expScale(node.expZ) };
        node.parentID = parent_of[p.wide];
        for (int s = 0; s < 4; s++) {
            node.children[s] = kBVHSentinel;
            node.triSize[s] = 0;
        }
        uint8_t *qmin[3] = { node.qMinX, node.qMinY, node.qMinZ };
        uint8_t *qmax[3] = { node.qMaxX, node.qMaxY, node.qMaxZ };
        for (size_t s = 0; s < kids.size(); s++) {
            const BinNode &kid = b.bin[kids[s]];
            for (int a = 0; a < 3; a++) {
                int lo = (int)std::floor(((double)kid.box.lo[a] - (double)node_min[a]) / (double)scale[a]);
                int hi = (int)std::ceil(((double)kid.box.hi[a] - (double)node_min[a]) / (double)scale[a]);
                lo = std::max(0, std::min(255, lo));
                hi = std::max(0, std::min(255, hi));
                // Conservative under the float expression the traversal uses.
                while (lo > 0 && node_min[a] + scale[a] * (float)lo > kid.box.lo[a]) lo--;
                while (hi < 255 && node_min[a] + scale[a] * (float)hi < kid.box.hi[a]) hi++;
                qmin[a][s] = (uint8_t)lo;
                qmax[a][s] = (uint8_t)hi;
            }
            if (kid.leaf()) {
                node.children[s] = (int32_t)(0x80000000u | (uint32_t)kid.first);
                node.triSize[s] = (uint8_t)kid.count;
                out->numLeaves++;
            } else {
                int wide = (int)out->nodes.size();
                out->nodes.emplace_back();
                parent_of.push_back(p.wide);
                node.children[s] = wide;
                node.internalNodes |= (uint8_t)(1u << s);
                queue.push_back({ kids[s], wide });
            }
        }
        out->nodes[p.wide] = node;
    }
}

bool loadMapAssets(const std::string &scene_dir, MapAssets *out, std::string *err)
{
    std::vector<Vec3> tris;
    if (!loadCollisions(scene_dir + "/collisions.bin", &out->worldBounds, &tris, err)) return false;
    if (!loadSpawns(scene_dir + "/spawns.bin", out, err)) return false;
    if (!loadZones(scene_dir + "/zones.bin", out, err)) return false;
    buildMeshBVH(tris, out);
    // The reference requires navmesh.bin (map_importer.cpp:426 asserts); here it is
    // needed only when an agent's policy id is -1 (the A* bot).
    std::string nav_err;
    out->hasNav = loadNavmesh(scene_dir + "/navmesh.bin", &out->nav, &nav_err);
    return true;
}

bool fillMapConsts(const MapAssets &assets, MapConsts *mc, std::string *err)
{
    memset(mc, 0, sizeof(*mc));
    mc->worldBounds = assets.worldBounds;
    mc->maxDist = length(assets.worldBounds.pMax - assets.worldBounds.pMin);
    {
        // sim.cpp:5869-5882: 90 degree vertical fov, 16:9.
        float aspect = 16.f / 9.f;
        constexpr float mult = kPi / 180.f;
        float f = 1.f / tanf(mult * (90.f / 2.f));
        float wx = f / aspect, wy = 1.f;
        float hx = f, hy = 1.f;
        float wi = 1.f / sqrtf(wx * wx + wy * wy);
        float hi = 1.f / sqrtf(hx * hx + hy * hy);
        mc->frustum[0] = wx * wi; mc->frustum[1] = wy * wi;
        mc->frustum[2] = hx * hi; mc->frustum[3] = hy * hi;
    }
    {   // vertex occupancy grid (mp_geom.h: sphereSweepMayTouchMirroredVertex); every vertex the casts can
        // read, the pad triangle included
        float lo[2] = { FLT_MAX, FLT_MAX }, hi[2] = { -FLT_MAX, -FLT_MAX };
        for (const Vec3 &v : assets.verts) {
            lo[0] = fminf(lo[0], v.x); hi[0] = fmaxf(hi[0], v.x);
            lo[1] = fminf(lo[1], v.y); hi[1] = fmaxf(hi[1], v.y);
        }
        if (assets.verts.empty()) { lo[0] = lo[1] = 0.f; hi[0] = hi[1] = 1.f; }
        mc->vertGridMinX = lo[0];
        mc->vertGridMinY = lo[1];
        mc->vertGridInvCellX = (float)kVertGrid / fmaxf(hi[0] - lo[0], 1.f);
        mc->vertGridInvCellY = (float)kVertGrid / fmaxf(hi[1] - lo[1], 1.f);
        for (const Vec3 &v : assets.verts) {
            int i = (int)((v.x - lo[0]) * mc->vertGridInvCellX), j = (int)((v.y - lo[1]) * mc->vertGridInvCellY);
            i = i < 0 ? 0 : (i >= kVertGrid ? kVertGrid - 1 : i);
            j = j < 0 ? 0 : (j >= kVertGrid ? kVertGrid - 1 : j);
            mc->vertGrid[(j * kVertGrid + i) >> 5] |= 1u << ((j * kVertGrid + i) & 31);
        }
    }
    if (assets.aSpawns.size() > kMaxSpawns || assets.bSpawns.size() > kMaxSpawns ||
        assets.respawns.size() >= kMaxSpawns || assets.zoneBoxes.size() > kMaxZones) {
        *err = "too many spawns / zones";
        return false;
    }
    mc->numASpawns = (uint32_t)assets.aSpawns.size();
    mc->numBSpawns = (uint32_t)assets.bSpawns.size();
    mc->numRespawns = (uint32_t)assets.respawns.size();
    mc->numZones = (uint32_t)assets.zoneBoxes.size();
    for (size_t i = 0; i < assets.aSpawns.size(); i++) mc->aSpawns[i] = assets.aSpawns[i];
    for (size_t i = 0; i < assets.bSpawns.size(); i++) mc->bSpawns[i] = assets.bSpawns[i];
    for (size_t i = 0; i < assets.respawns.size(); i++) mc->respawns[i] = assets.respawns[i];
    for (size_t i = 0; i < assets.zoneBoxes.size(); i++) {
        mc->zoneBoxes[i] = assets.zoneBoxes[i];
        mc->zoneRots[i] = assets.zoneRots[i];
    }
    // mgr.cpp:913-944
    const float top = -56.f + kStandHeight * 1.5f;
    {   // createPersistentEntities, level_gen.cpp:279-322 (SimFlags::SubZones); a map with fewer than
        // three zones cannot run that flag (mpenv_create refuses it)
        const size_t nz = assets.zoneBoxes.size();
        for (int k = 0; k < 2; k++) {
            const size_t z = (size_t)k + 1 < nz ? (size_t)k + 1 : 0;
            mc->subZones[k] = { assets.zoneBoxes[z].pMin, assets.zoneBoxes[z].pMax, assets.zoneRots[z] };
        }
        mc->subZones[2] = { { -950, -500, 0 }, { -50, 500, 1000 }, 0.f };
        mc->subZones[3] = { { 50, -500, 0 }, { 950, 500, 1000 }, 0.f };
        mc->subZones[4] = { { -1000, -1650, 0 }, { -50, -600, 1000 }, 0.f };
        mc->subZones[5] = { { 50, -1650, 0 }, { 1000, -600, 1000 }, 0.f };
        mc->subZones[6] = { { -1000, 600, 0 }, { -50, 1650, 1000 }, 0.f };
        mc->subZones[7] = { { 1000, 600, 0 }, { 50, 1650, 1000 }, 0.f };      // pMin.x > pMax.x in the reference too
    }
    mc->numGoalRegions = 2;
    mc->goalRegions[0].sub[0] = { { 625, 510, -64 }, { 900, 540, top }, 0.f };
    mc->goalRegions[0].numSub = 1;
    mc->goalRegions[0].attackerTeam = 1;
    mc->goalRegions[0].rewardStrength = 1.f;
    mc->goalRegions[1].sub[0] = { { 938, 440, -56 }, { 1030, 539, top }, 0.f };
    mc->goalRegions[1].sub[1] = { { 545, 102, -64 }, { 630, 134, top }, 0.f };
    mc->goalRegions[1].numSub = 2;
    mc->goalRegions[1].attackerTeam = 1;
    mc->goalRegions[1].rewardStrength = 1.f;
    // mgr.cpp:1383-1395
    mc->numWeaponTypes = 1;
    mc->weapons[0] = { 30, 30, 10.f, 0.005f };
    // sim.cpp:3347-3350 with the arguments of :3485-3487 and :3500-3502.
    for (int w = 0; w < kFwdLidarW; w++) {
        float theta = (0.75f * kPi) * ((float)w / (float)(kFwdLidarW - 1)) +
                      0.5f * (1.f - 0.75f) * kPi;
        mc->fwdLidarX[w] = -mp_cosf(theta);
        mc->fwdLidarY[w] = mp_sinf(theta);
    }
    for (int w = 0; w < kRearLidarW; w++) {
        float theta = (-kPi) * ((float)w / (float)(kRearLidarW - 1)) + 0.f;
        mc->rearLidarX[w] = -mp_cosf(theta);
        mc->rearLidarY[w] = mp_sinf(theta);
    }
    return true;
}

}  // namespace mpenv
