// navmesh.cpp -- see navmesh.h.
#include "navmesh.h"

#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <set>

#include "mp_math.h"

namespace mpenv {

namespace {

struct VertKey {
    uint32_t w[3];
    bool operator<(const VertKey &o) const { return memcmp(w, o.w, sizeof(w)) < 0; }
};

Vec3 triCenter(const NavmeshAssets &nav, int tri)
{
    // NavUtils::CenterOfTri (mgr.cpp:969-977): sum of v / 3
    Vec3 c = v3(0.f, 0.f, 0.f);
    for (int i = 0; i < 3; i++) c = c + nav.verts[nav.tris[tri * 3 + i]] / 3.0f;
    return c;
}

// Search state shared by the comparator, as in the reference (a function-local
// static vector referenced from the ordering lambda, mgr.cpp:1066-1074).
struct SearchNode { int cameFrom; float startDist; float score; };
std::vector<SearchNode> g_state;
struct ByScore {
    bool operator()(int l, int r) const { return g_state[l].score < g_state[r].score; }
};

// EarlyOutData (mgr.cpp:1000-1048, 1109-1137): for navmeshes above 400 triangles,
// far goals are replaced by one precomputed goal per compass direction.
struct EarlyOut {
    int forDir[9];
    float totalDistance = 0.f;
    float reasonableDistance2 = 0.f;
    bool enabled = false;

    void init(const NavmeshAssets &nav)
    {
        const int reasonable = 400;
        for (int &d : forDir) d = -1;
        if ((int)nav.numTris() <= reasonable) { enabled = false; return; }
        enabled = true;
        Vec3 mins = v3(FLT_MAX, FLT_MAX, FLT_MAX), maxs = v3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
        for (const Vec3 &p : nav.verts) {
            mins = v3(fminf(mins.x, p.x), fminf(mins.y, p.y), fminf(mins.z, p.z));
            maxs = v3(fmaxf(maxs.x, p.x), fmaxf(maxs.y, p.y), fmaxf(maxs.z, p.z));
        }
        totalDistance = length(maxs - mins);
        reasonableDistance2 = totalDistance * (float)reasonable / (float)nav.numTris();
        reasonableDistance2 *= reasonableDistance2;
    }

    int get(Vec3 start, Vec3 target) const
    {
        if (!enabled) return -1;
        const Vec3 d = start - target;
        if (dot(d, d) < reasonableDistance2) return -1;
        const Vec3 dir = normalize(target - start);
        const int dx = (int)(dir.x * 1.9f), dy = (int)(dir.y * 1.9f);
        return forDir[(dx + 1) + (dy + 1) * 3];
    }
};

// NavUtils::AStarPathfindToTri (mgr.cpp:1051-1107): the FIRST triangle of the
// path start -> goal.  The open list is a std::set ordered by the live scores,
// which drops a candidate whose score ties an element already in the set and can
// hold a triangle twice after its score improved; the same container and the
// same operation sequence are used here so that ties resolve identically.
int firstStepTowards(const NavmeshAssets &nav, const EarlyOut &early, int start_tri, int goal_tri)
{
    const Vec3 start = triCenter(nav, start_tri);
    const Vec3 goal = triCenter(nav, goal_tri);
    const int shortcut = early.get(start, goal);
    if (shortcut != -1) return shortcut;

    const int n = (int)nav.numTris();
    g_state.resize(n);
    for (int t = 0; t < n; t++) g_state[t] = SearchNode { -1, FLT_MAX, FLT_MAX };
    g_state[start_tri].startDist = 0.f;

    std::set<int, ByScore> open;
    open.insert(start_tri);
    while (!open.empty()) {
        const int cur = *open.begin();
        if (cur == goal_tri) {
            int back = cur;
            while (back != start_tri && back != -1) {
                if (g_state[back].cameFrom == start_tri) return back;
                back = g_state[back].cameFrom;
            }
            return goal_tri;
        }
        open.erase(open.begin());
        const Vec3 center = cur == start_tri ? start : triCenter(nav, cur);
        for (int e = 0; e < 3; e++) {
            const int nb = nav.adjacency[cur * 3 + e];
            if (nb == -1) continue;
            const Vec3 nb_pos = triCenter(nav, nb);
            const float through = g_state[cur].startDist + distance(center, nb_pos);
            if (through < g_state[nb].startDist) {
                g_state[nb].cameFrom = cur;
                g_state[nb].startDist = through;
                g_state[nb].score = through + distance(nb_pos, goal);
                open.insert(nb);
            }
        }
    }
    return -1;
}

void prepareEarlyOut(const NavmeshAssets &nav, EarlyOut &early, int start_tri)
{
    if (!early.enabled) return;
    early.enabled = false;          // the searches below must not shortcut
    for (int dx = -1; dx <= 1; dx++) {
        for (int dy = -1; dy <= 1; dy++) {
            if (dx == 0 && dy == 0) continue;
            const Vec3 dir = normalize(v3((float)dx, (float)dy, 0.f));
            const Vec3 start = triCenter(nav, start_tri);
            const Vec3 end = start + dir * early.totalDistance * 2.0f;
            float best = FLT_MAX;
            int best_tri = -1;
            for (int t = 0; t < (int)nav.numTris(); t++) {
                const float d = length(triCenter(nav, t) - end);
                if (d < best) { best = d; best_tri = t; }
            }
            early.forDir[(dx + 1) + (dy + 1) * 3] = firstStepTowards(nav, early, start_tri, best_tri);
        }
    }
    early.enabled = true;
}

}  // namespace

void buildAStarLookup(NavmeshAssets *nav)
{
    const uint32_t n = nav->numTris();
    nav->lookup.assign((size_t)n * n, 0xFFFFFFFFu);
    EarlyOut early;
    early.init(*nav);
    for (uint32_t s = 0; s < n; s++) {
        prepareEarlyOut(*nav, early, (int)s);
        for (uint32_t g = 0; g < n; g++)
            nav->lookup[(size_t)s * n + g] = (uint32_t)firstStepTowards(*nav, early, (int)s, (int)g);
    }
}

bool loadNavmesh(const std::string &path, NavmeshAssets *out, std::string *err)
{
    std::ifstream f(path, std::ios::binary);
    if (!f.is_open()) { *err = "failed to open " + path; return false; }
    auto rd = [&](void *dst, size_t n) { f.read((char *)dst, (std::streamsize)n); return (size_t)f.gcount() == n; };
    uint32_t nv = 0, nf = 0, ni = 0;
    if (!rd(&nv, 4) || nv > (1u << 24)) { *err = "corrupt navmesh " + path; return false; }
    std::vector<Vec3> verts(nv);
    if (!rd(verts.data(), sizeof(Vec3) * nv) || !rd(&nf, 4) || nf > (1u << 24)) { *err = "corrupt navmesh " + path; return false; }
    std::vector<uint32_t> counts(nf);
    if (!rd(counts.data(), 4 * (size_t)nf) || !rd(&ni, 4) || ni > (1u << 26)) { *err = "corrupt navmesh " + path; return false; }
    std::vector<uint32_t> idx(ni);
    if (!rd(idx.data(), 4 * (size_t)ni)) { *err = "corrupt navmesh " + path; return false; }

    // de-duplicate (mgr.cpp:1300-1318)
    std::vector<uint32_t> remap(nv);
    std::map<VertKey, uint32_t> seen;
    out->verts.clear();
    for (uint32_t i = 0; i < nv; i++) {
        VertKey k;
        memcpy(k.w, &verts[i], 12);
        auto it = seen.find(k);
        if (it == seen.end()) {
            it = seen.emplace(k, (uint32_t)out->verts.size()).first;
            out->verts.push_back(verts[i]);
        }
        remap[i] = it->second;
    }

    // fan triangulation of each face
    out->tris.clear();
    size_t off = 0;
    for (uint32_t fc = 0; fc < nf; fc++) {
        const uint32_t c = counts[fc];
        if (off + c > ni || c < 3) { *err = "corrupt navmesh faces in " + path; return false; }
        for (uint32_t k = 1; k + 1 < c; k++) {
            const uint32_t tri[3] = { idx[off], idx[off + k], idx[off + k + 1] };
            for (uint32_t v : tri) {
                if (v >= nv) { *err = "navmesh index out of range in " + path; return false; }
                out->tris.push_back(remap[v]);
            }
        }
        off += c;
    }

    // adjacency through directed half-edges: the neighbour across edge (a -> b) is the
    // first triangle that owns the opposite half-edge (b -> a).  Coincident duplicate
    // triangles (simple_map has a few) therefore both link to the real neighbour
    // instead of pairing up with each other.
    const uint32_t nt = out->numTris();
    out->adjacency.assign((size_t)nt * 3, -1);
    std::map<std::pair<uint32_t, uint32_t>, int32_t> half_edges;
    for (uint32_t t = 0; t < nt; t++)
        for (int e = 0; e < 3; e++)
            half_edges.emplace(std::make_pair(out->tris[t * 3 + e], out->tris[t * 3 + (e + 1) % 3]), (int32_t)t);
    for (uint32_t t = 0; t < nt; t++) {
        for (int e = 0; e < 3; e++) {
            auto it = half_edges.find(std::make_pair(out->tris[t * 3 + (e + 1) % 3], out->tris[t * 3 + e]));
            if (it != half_edges.end() && it->second != (int32_t)t) out->adjacency[t * 3 + e] = it->second;
        }
    }
    buildAStarLookup(out);
    return true;
}

}  // namespace mpenv
