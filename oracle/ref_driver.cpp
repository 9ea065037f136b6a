// TEST INFRASTRUCTURE (oracle) -- never linked into or called by the product.
//
// Drives the reference's OWN simulation code (sim.cpp / utils.cpp /
// level_gen.cpp / mesh_bvh.inl compiled from /root/reference/src against
// oracle/shim) behind a small C API so that tests and bench.py's CPU-baseline
// leg can step it from Python.
//
// What this file restates (because mgr.cpp cannot be compiled here -- it
// needs the Madrona executors, Embree and meshoptimizer):
//   * asset readers              map_importer.cpp:35-124, 223-256, 421-567
//   * hard-coded weapon / flags  mgr.cpp:1383-1413
//   * goal regions               mgr.cpp:913-944
//   * RNG key derivation         mgr.cpp:1736-1738
//   * TaskConfig                 mgr.cpp:1752-1785
//   * Manager::init / step       mgr.cpp:1923-1951, CPUImpl::run mgr.cpp:242-273
// The MeshBVH (reference node / vertex format, mesh_bvh.hpp:61-99) is handed
// in by the caller: the reference builds it with Embree, which is absent.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>
#include <thread>
#include <atomic>
#include <fstream>
#include <new>

#include "sim.hpp"
#include "utils.hpp"
#include "level_gen.hpp"

#include "../include/mpenv_state.h"

using namespace madrona;
using namespace madronaMPEnv;

namespace madronaMPEnv {
// dnn.hpp symbols referenced by sim.cpp (policyWeights is always null here).
struct PolicyWeights {};
PolicyWeights * loadPolicyWeights(const char *) { return nullptr; }
void addPolicyEvalTasks(TaskGraphBuilder &) {}
namespace VizSystem {
void registerTypes(ECSRegistry &) {}
void initWorld(Context &, VizState *) {}
void setupGameTasks(VizState *, madrona::TaskGraphBuilder &) {}
}
}

namespace {

struct RefConfig {
    uint32_t numWorlds;
    uint32_t randSeed;
    int32_t autoReset;
    uint32_t simFlags;
    uint32_t taskType;
    uint32_t teamSize;
    uint32_t numPBTPolicies;
    uint32_t policyHistorySize;
    uint32_t worldIDOffset;   // global id of world 0 (multi-GPU shards)
    uint32_t numThreads;
    uint32_t enableEventLog;  // TaskConfig::eventGlobalState (mgr.cpp:1227, 1777)
    uint32_t enableRecordLog; // TaskConfig::recordLog (mgr.cpp:1774)
    uint32_t enableReplayLog; // TaskConfig::replayLog (mgr.cpp:1775)
    uint32_t numCurriculumSnapshots;      // TrajectoryCurriculum (mgr.cpp:1419-1441)
    const void *curriculumSnapshots;      // numCurriculumSnapshots x sizeof(CurriculumSnapshot), copied
    // Navmesh + A* table (LevelData, types.hpp:136-143), handed in by the caller like the BVH:
    // madrona::Navmesh::initFromPolygons and meshoptimizer are not vendored, and
    // buildAStarLookup lives in mgr.cpp, which cannot be compiled here.  All copied.
    uint32_t numNavVerts, numNavTris;
    const float *navVerts;                // numNavVerts x 3
    const uint32_t *navTriIndices;        // numNavTris x 3
    const int32_t *navTriAdjacency;       // numNavTris x 3
    const uint32_t *navLookup;            // numNavTris x numNavTris
    uint32_t trainFlank;                  // Config::trainFlank -> RewardMode::Flank (mgr.cpp:1745-1750)
};

struct RefSim {
    RefConfig cfg;
    TaskConfig taskCfg;
    TaskGraphManager graphs;
    std::vector<shim::WorldState> worlds;
    std::vector<Sim *> sims;

    MeshBVH bvh;
    std::vector<MeshBVH::Node> bvhNodes;
    std::vector<MeshBVH::BVHVertex> bvhVerts;

    std::vector<Spawn> aSpawns, bSpawns, commonRespawns;
    std::vector<AABB> zoneAABBs;
    std::vector<float> zoneRots;
    std::vector<GoalRegion> goalRegions;
    WeaponStats weapon;
    TrainControl trainCtrl;
    SpawnCurriculum spawnCurriculum;
    EventLogGlobalState eventGlobal { 0, 0 };
    std::vector<StepLog> recordLog, replayLog;
    std::vector<CurriculumSnapshot> curriculum;
    std::vector<math::Vector3> navVerts;
    std::vector<uint32_t> navTris, navAdj, navLookup;
    std::string error;
};

bool readFile(const std::string &path, std::vector<char> &out)
{
    std::ifstream f(path, std::ios::binary);
    if (!f.is_open()) return false;
    f.seekg(0, f.end);
    size_t n = (size_t)f.tellg();
    f.seekg(0, f.beg);
    out.resize(n);
    f.read(out.data(), (std::streamsize)n);
    return true;
}

template <typename T>
bool take(const std::vector<char> &buf, size_t &off, T *dst, size_t count = 1)
{
    size_t n = sizeof(T) * count;
    if (off + n > buf.size()) return false;
    memcpy(dst, buf.data() + off, n);
    off += n;
    return true;
}

// map_importer.cpp:508-541
bool loadSpawns(const std::string &path, RefSim &s)
{
    std::vector<char> buf;
    if (!readFile(path, buf)) return false;
    size_t off = 0;
    std::vector<Spawn> *lists[3] = { &s.aSpawns, &s.bSpawns, &s.commonRespawns };
    for (auto *l : lists) {
        uint32_t n;
        if (!take(buf, off, &n)) return false;
        l->resize(n);
        if (!take(buf, off, l->data(), n)) return false;
    }
    return true;
}

// map_importer.cpp:543-567
bool loadZones(const std::string &path, RefSim &s)
{
    std::vector<char> buf;
    if (!readFile(path, buf)) return false;
    size_t off = 0;
    uint32_t n;
    if (!take(buf, off, &n)) return false;
    s.zoneAABBs.resize(n);
    s.zoneRots.resize(n);
    if (!take(buf, off, s.zoneAABBs.data(), n)) return false;
    if (!take(buf, off, s.zoneRots.data(), n)) return false;
    return true;
}

// map_importer.cpp:223-233 (only the world AABB is needed here; the geometry
// arrives already built into the BVH).
bool loadWorldBounds(const std::string &path, AABB *out)
{
    std::vector<char> buf;
    if (!readFile(path, buf)) return false;
    size_t off = 0;
    return take(buf, off, out);
}

void runGraph(RefSim &s, TaskGraphID id)
{
    TaskGraphBuilder &g = s.graphs.graph((size_t)id);
    uint32_t nw = s.cfg.numWorlds;
    uint32_t nt = s.cfg.numThreads;
    if (nt <= 1 || nw < 2) {
        for (uint32_t w = 0; w < nw; w++) g.run(s.worlds[w]);
        return;
    }
    // One world per task, like ThreadPoolExecutor.
    std::atomic<uint32_t> next { 0 };
    std::vector<std::thread> threads;
    threads.reserve(nt);
    for (uint32_t t = 0; t < nt; t++) {
        threads.emplace_back([&]() {
            for (;;) {
                uint32_t w = next.fetch_add(1, std::memory_order_relaxed);
                if (w >= nw) break;
                g.run(s.worlds[w]);
            }
        });
    }
    for (auto &t : threads) t.join();
}

int32_t agentIndexOf(Sim &sim, Entity e)
{
    if (e == Entity::none()) return -1;
    for (uint32_t i = 0; i < sim.numAgents; i++) {
        if (sim.agents[i] == e) return (int32_t)i;
    }
    return -2;
}

} // namespace

extern "C" {

// bvh_nodes: num_nodes x 64 B (MeshBVH::Node), bvh_verts: num_verts x 20 B
// (MeshBVH::BVHVertex), root_aabb: 6 floats.
void * ref_create(const RefConfig *cfg, const char *scene_dir,
                  const void *bvh_nodes, uint32_t num_nodes,
                  const void *bvh_verts, uint32_t num_verts,
                  uint32_t num_leaves, const float *root_aabb)
{
    static_assert(sizeof(MeshBVH::Node) == 64);
    static_assert(sizeof(MeshBVH::BVHVertex) == 20);
    static_assert(sizeof(Spawn) == 32);

    RefSim *s = new RefSim {};
    s->cfg = *cfg;
    if (s->cfg.numThreads == 0) {
        s->cfg.numThreads = std::thread::hardware_concurrency();
    }

    std::string dir(scene_dir);
    AABB world_bounds;
    if (!loadWorldBounds(dir + "/collisions.bin", &world_bounds) ||
        !loadSpawns(dir + "/spawns.bin", *s) ||
        !loadZones(dir + "/zones.bin", *s)) {
        fprintf(stderr, "ref_create: failed to read assets under %s\n", scene_dir);
        delete s;
        return nullptr;
    }

    s->bvhNodes.resize(num_nodes);
    memcpy(s->bvhNodes.data(), bvh_nodes, sizeof(MeshBVH::Node) * num_nodes);
    // sphereCastLeaf always reads numTrisPerLeaf triangles (mesh_bvh.inl:867):
    // pad by one zero-area copy of the last triangle so the over-read is defined.
    s->bvhVerts.resize(num_verts + 3);
    memcpy(s->bvhVerts.data(), bvh_verts, sizeof(MeshBVH::BVHVertex) * num_verts);
    for (int i = 0; i < 3; i++) {
        s->bvhVerts[num_verts + i] = s->bvhVerts[num_verts - 3 + i];
    }

    s->bvh.nodes = s->bvhNodes.data();
    s->bvh.leafMats = nullptr;
    s->bvh.vertices = s->bvhVerts.data();
    s->bvh.rootAABB = AABB {
        { root_aabb[0], root_aabb[1], root_aabb[2] },
        { root_aabb[3], root_aabb[4], root_aabb[5] },
    };
    s->bvh.numNodes = num_nodes;
    s->bvh.numLeaves = num_leaves;
    s->bvh.numVerts = num_verts;
    s->bvh.materialIDX = -1;
    s->bvh.magic = 0;

    // mgr.cpp:913-944
    {
        GoalRegion r0 {};
        r0.subRegions[0] = ZOBB {
            { 625, 510, -64 },
            { 900, 540, -56 + consts::standHeight * 1.5f },
            0.f,
        };
        GoalRegion r1 {};
        r1.subRegions[0] = ZOBB {
            { 938, 440, -56 },
            { 1030, 539, -56 + consts::standHeight * 1.5f },
            0.f,
        };
        r1.subRegions[1] = ZOBB {
            { 545, 102, -64 },
            { 630, 134, -56 + consts::standHeight * 1.5f },
            0.f,
        };
        r1.numSubRegions = 2;
        s->goalRegions.push_back(r0);
        s->goalRegions.push_back(r1);
    }

    // mgr.cpp:1383-1395
    s->weapon = WeaponStats {
        .magSize = 30,
        .reloadTime = 30,
        .dmgPerBullet = 10.f,
        .accuracyScale = 0.005f,
    };

    // mgr.cpp:1397-1413
    SimFlags flags = SimFlags(cfg->simFlags);
    s->trainCtrl = TrainControl {
        .evalMode = (flags & SimFlags::SimEvalMode) == SimFlags::SimEvalMode ? 1 : 0,
        .randomizeEpisodeLengthAfterReset =
            (flags & SimFlags::StaggerStarts) == SimFlags::StaggerStarts ? 1 : 0,
        .randomizeTeamSides =
            (flags & SimFlags::RandomFlipTeams) == SimFlags::RandomFlipTeams ? 1 : 0,
    };

    // The spawn curriculum (mgr.cpp:685-911) needs madrona::Navmesh BFS; its
    // only live consumer is curCurriculumSpawnIdx (unused: utils.cpp:811-818 is
    // #if 0).  numSpawns = 0 keeps the RNG advance count identical.
    memset(&s->spawnCurriculum, 0, sizeof(SpawnCurriculum));

    // mgr.cpp:1736-1738
    RandKey init_key = rand::initKey(cfg->randSeed);
    RandKey sim_init_key = rand::split_i(init_key, 0);

    StandardSpawns spawns {};
    spawns.aSpawns = s->aSpawns.data();
    spawns.bSpawns = s->bSpawns.data();
    spawns.commonRespawns = s->commonRespawns.data();
    spawns.numDefaultASpawns = (uint32_t)s->aSpawns.size();
    spawns.numDefaultBSpawns = (uint32_t)s->bSpawns.size();
    spawns.numExtraASpawns = 0;
    spawns.numExtraBSpawns = 0;
    spawns.numCommonRespawns = (uint32_t)s->commonRespawns.size();
    spawns.respawnRegions = nullptr;
    spawns.numRespawnRegions = 0;

    Zones zones {};
    zones.bboxes = s->zoneAABBs.data();
    zones.rotations = s->zoneRots.data();
    zones.numZones = (uint32_t)s->zoneAABBs.size();

    Navmesh navmesh {};
    AStarLookup astar { nullptr };
    if (cfg->numNavTris > 0 && cfg->navVerts && cfg->navTriIndices && cfg->navLookup) {
        s->navVerts.resize(cfg->numNavVerts);
        memcpy(s->navVerts.data(), cfg->navVerts, sizeof(float) * 3 * cfg->numNavVerts);
        s->navTris.assign(cfg->navTriIndices, cfg->navTriIndices + 3 * (size_t)cfg->numNavTris);
        s->navAdj.resize(3 * (size_t)cfg->numNavTris);
        if (cfg->navTriAdjacency) memcpy(s->navAdj.data(), cfg->navTriAdjacency, 4 * s->navAdj.size());
        s->navLookup.assign(cfg->navLookup, cfg->navLookup + (size_t)cfg->numNavTris * cfg->numNavTris);
        navmesh.vertices = s->navVerts.data();
        navmesh.triIndices = s->navTris.data();
        navmesh.triAdjacency = s->navAdj.data();
        navmesh.triSampleCDF = nullptr;
        navmesh.numVerts = cfg->numNavVerts;
        navmesh.numTris = cfg->numNavTris;
        astar.data = s->navLookup.data();
    }

    // mgr.cpp:1752-1785
    TaskConfig &tc = s->taskCfg;
    tc = TaskConfig {};
    tc.autoReset = cfg->autoReset != 0;
    tc.showSpawns = true;
    tc.simFlags = flags;
    tc.initRandKey = sim_init_key;
    tc.numPBTPolicies = cfg->numPBTPolicies;
    tc.policyHistorySize = cfg->policyHistorySize;
    tc.worldBounds = world_bounds;
    tc.pTeamSize = cfg->teamSize;
    tc.eTeamSize = cfg->teamSize;
    tc.staticMeshes = &s->bvh;
    tc.numStaticMeshes = 1;
    tc.navmesh = navmesh;
    tc.aStarLookup = astar;
    tc.standardSpawns = spawns;
    tc.spawnCurriculum = s->spawnCurriculum;
    tc.zones = zones;
    tc.task = Task(cfg->taskType);
    tc.rewardMode = cfg->trainFlank ? RewardMode::Flank : RewardMode::Default;
    tc.highlevelMove = false;
    tc.viz = nullptr;
    if (cfg->enableRecordLog) s->recordLog.resize(cfg->numWorlds);
    if (cfg->enableReplayLog) s->replayLog.resize(cfg->numWorlds);
    tc.recordLog = cfg->enableRecordLog ? s->recordLog.data() : nullptr;
    tc.replayLog = cfg->enableReplayLog ? s->replayLog.data() : nullptr;
    tc.eventGlobalState = cfg->enableEventLog ? &s->eventGlobal : nullptr;
    tc.goalRegions = s->goalRegions.data();
    tc.numGoalRegions = (int32_t)s->goalRegions.size();
    tc.numEpisodes = 0;
    tc.episodes = nullptr;
    tc.numWeaponTypes = 1;
    tc.weaponTypeStats = &s->weapon;
    tc.trainControl = &s->trainCtrl;
    static_assert(sizeof(CurriculumSnapshot) == 176);
    if (cfg->numCurriculumSnapshots > 0 && cfg->curriculumSnapshots) {
        s->curriculum.resize(cfg->numCurriculumSnapshots);
        memcpy(s->curriculum.data(), cfg->curriculumSnapshots, sizeof(CurriculumSnapshot) * cfg->numCurriculumSnapshots);
    }
    tc.trajectoryCurriculum = TrajectoryCurriculum { (i32)s->curriculum.size(),
                                                     s->curriculum.empty() ? nullptr : s->curriculum.data() };
    tc.policyWeights = nullptr;

    // The registry is process-global (type ids are static); registering the
    // same types again for a second RefSim is idempotent.
    ECSRegistry registry;
    Sim::registerTypes(registry, tc);
    Sim::setupTasks(s->graphs, tc);

    s->worlds.resize(cfg->numWorlds);
    s->sims.resize(cfg->numWorlds);
    for (uint32_t w = 0; w < cfg->numWorlds; w++) {
        shim::WorldState &ws = s->worlds[w];
        ws.init((int32_t)(cfg->worldIDOffset + w));
        void *mem = aligned_alloc(alignof(Sim), (sizeof(Sim) + 63) / 64 * 64);
        memset(mem, 0, sizeof(Sim));
        ws.userData = mem;
        Engine ctx(&ws);
        Sim::WorldInit init {};
        s->sims[w] = new (mem) Sim(ctx, tc, init);
    }

    return s;
}

void ref_destroy(void *h)
{
    RefSim *s = (RefSim *)h;
    if (!s) return;
    for (Sim *sim : s->sims) free(sim);
    delete s;
}

// Manager::init (mgr.cpp:1923-1946): force a reset on every world and run the
// Init graph.
void ref_init(void *h)
{
    RefSim *s = (RefSim *)h;
    for (auto &w : s->worlds) {
        Engine ctx(&w);
        ctx.singleton<WorldReset>().reset = 1;
    }
    runGraph(*s, TaskGraphID::Init);
}

// Manager::step (mgr.cpp:1948-1951).
void ref_step(void *h)
{
    RefSim *s = (RefSim *)h;
    runGraph(*s, TaskGraphID::Step);
}

uint32_t ref_num_threads(void *h) { return ((RefSim *)h)->cfg.numThreads; }

// The step's GameEvent / PackedStepSnapshot rows (the temporaries cleared at the
// start of every Step graph, sim.cpp:5344-5345), gathered world by world in
// table order -- what CPUImpl::run hands to writeGameEvents (mgr.cpp:259-268)
// for a single-threaded executor.  which: 0 = events, 1 = step snapshots.
// Returns the number of rows, or -1 if `cap_rows` is too small.
int64_t ref_read_event_log(void *h, uint32_t which, void *dst, uint64_t cap_rows)
{
    RefSim *s = (RefSim *)h;
    const uint32_t arch = which == 0 ? shim::TypeID<GameEventEntity>::get()
                                     : shim::TypeID<PackedStepSnapshotEntity>::get();
    const uint32_t comp = which == 0 ? shim::TypeID<GameEvent>::get()
                                     : shim::TypeID<PackedStepSnapshot>::get();
    const uint64_t row_bytes = which == 0 ? sizeof(GameEvent) : sizeof(PackedStepSnapshot);
    uint64_t n = 0;
    for (auto &w : s->worlds) {
        shim::Table &t = w.tables[arch];
        int col = w.colIndex(arch, comp);
        if (n + (uint64_t)t.numRows > cap_rows) return -1;
        memcpy((char *)dst + n * row_bytes, t.cols[col], (uint64_t)t.numRows * row_bytes);
        n += (uint64_t)t.numRows;
    }
    return (int64_t)n;
}

// The global counters logEvent / writePackedStepSnapshot bump (sim.cpp:31-32, 58-59);
// reset to zero as CPUImpl::run does after writing (mgr.cpp:267-268).
void ref_event_log_counts(void *h, uint32_t *num_events, uint32_t *num_step_states)
{
    RefSim *s = (RefSim *)h;
    *num_events = s->eventGlobal.numEvents;
    *num_step_states = s->eventGlobal.numStepStates;
    s->eventGlobal.numEvents = 0;
    s->eventGlobal.numStepStates = 0;
}

uint32_t ref_event_row_bytes(uint32_t which)
{
    return which == 0 ? (uint32_t)sizeof(GameEvent) : (uint32_t)sizeof(PackedStepSnapshot);
}

// StepLog buffers (numWorlds x sizeof(StepLog)): the record log is filled by
// pvpRecordSystem every step; the replay log must be filled by the caller
// before each step (CPUImpl::run, mgr.cpp:244-257).
void * ref_record_log(void *h) { RefSim *s = (RefSim *)h; return s->recordLog.empty() ? nullptr : s->recordLog.data(); }
void * ref_replay_log(void *h) { RefSim *s = (RefSim *)h; return s->replayLog.empty() ? nullptr : s->replayLog.data(); }
uint32_t ref_step_log_bytes() { return (uint32_t)sizeof(StepLog); }

// Bytes per row of exported slot `export_id` (ExportID, sim.hpp:15-66) and
// the number of rows per world; 0 if the slot is not exported.
uint32_t ref_export_row_bytes(void *h, uint32_t export_id)
{
    (void)h;
    auto &g = shim::globalRegistry();
    if (export_id >= g.exports.size() || !g.exports[export_id].valid) return 0;
    return g.exports[export_id].numBytes;
}

uint32_t ref_export_rows_per_world(void *h, uint32_t export_id)
{
    RefSim *s = (RefSim *)h;
    auto &g = shim::globalRegistry();
    if (export_id >= g.exports.size() || !g.exports[export_id].valid) return 0;
    const shim::ExportDesc &e = g.exports[export_id];
    if (e.singleton) return 1;
    return (uint32_t)s->worlds[0].tables[e.archetype].numRows;
}

static char * exportRowPtr(shim::WorldState &w, const shim::ExportDesc &e)
{
    if (e.singleton) return w.singletons[e.typeID];
    int col = w.colIndex(e.archetype, e.typeID);
    return w.tables[e.archetype].cols[col];
}

// Gather slot `export_id` of every world into one [numWorlds * rows, bytes]
// world-major buffer (the layout Manager::exportTensor hands out).
int ref_read_export(void *h, uint32_t export_id, void *dst, uint64_t dst_bytes)
{
    RefSim *s = (RefSim *)h;
    auto &g = shim::globalRegistry();
    if (export_id >= g.exports.size() || !g.exports[export_id].valid) return -1;
    const shim::ExportDesc &e = g.exports[export_id];
    uint64_t off = 0;
    for (auto &w : s->worlds) {
        uint32_t rows = e.singleton ? 1 : (uint32_t)w.tables[e.archetype].numRows;
        uint64_t n = (uint64_t)rows * e.numBytes;
        if (off + n > dst_bytes) return -2;
        memcpy((char *)dst + off, exportRowPtr(w, e), n);
        off += n;
    }
    return 0;
}

int ref_write_export(void *h, uint32_t export_id, const void *src, uint64_t src_bytes)
{
    RefSim *s = (RefSim *)h;
    auto &g = shim::globalRegistry();
    if (export_id >= g.exports.size() || !g.exports[export_id].valid) return -1;
    const shim::ExportDesc &e = g.exports[export_id];
    uint64_t off = 0;
    for (auto &w : s->worlds) {
        uint32_t rows = e.singleton ? 1 : (uint32_t)w.tables[e.archetype].numRows;
        uint64_t n = (uint64_t)rows * e.numBytes;
        if (off + n > src_bytes) return -2;
        memcpy(exportRowPtr(w, e), (const char *)src + off, n);
        off += n;
    }
    return 0;
}

// simControlTensor (mgr.cpp:287-293): live TrainControl, 3 x i32.
void ref_set_sim_control(void *h, const int32_t *v)
{
    RefSim *s = (RefSim *)h;
    s->trainCtrl.evalMode = v[0];
    s->trainCtrl.randomizeEpisodeLengthAfterReset = v[1];
    s->trainCtrl.randomizeTeamSides = v[2];
}

uint32_t ref_state_bytes() { return (uint32_t)sizeof(mpenv_world_state); }

int ref_get_world_state(void *h, uint32_t world, mpenv_world_state *out)
{
    RefSim *s = (RefSim *)h;
    if (world >= s->cfg.numWorlds) return -1;
    shim::WorldState &w = s->worlds[world];
    Engine ctx(&w);
    Sim &sim = ctx.data();
    memset(out, 0, sizeof(*out));

    const MatchInfo &mi = ctx.singleton<MatchInfo>();
    out->team_a = mi.teamA;
    out->cur_step = mi.curStep;
    out->is_finished = mi.isFinished;
    out->enable_spawn_curriculum = mi.enableSpawnCurriculum;
    out->cur_curriculum_tier = mi.curCurriculumTier;
    out->cur_curriculum_spawn_idx = mi.curCurriculumSpawnIdx;

    const ZoneState &zs = ctx.singleton<ZoneState>();
    out->zone_cur = zs.curZone;
    out->zone_controlling_team = zs.curControllingTeam;
    out->zone_contested = zs.isContested;
    out->zone_captured = zs.isCaptured;
    out->zone_earned_point = zs.earnedPoint;
    out->zone_steps_remaining = zs.zoneStepsRemaining;
    out->zone_steps_until_point = zs.stepsUntilPoint;

    static_assert(sizeof(MatchResult) == 30 * sizeof(int32_t));
    memcpy(out->match_result, &ctx.singleton<MatchResult>(), sizeof(MatchResult));
    static_assert(sizeof(sim.zoneStats) == MPENV_MAX_ZONES * 5 * sizeof(int32_t));
    memcpy(out->zone_stats, sim.zoneStats.data(), sizeof(sim.zoneStats));

    static_assert(sizeof(RNG) == 12);
    uint32_t rng_raw[3];
    memcpy(rng_raw, &sim.baseRNG, 12);
    out->base_rng_key[0] = rng_raw[0];
    out->base_rng_key[1] = rng_raw[1];
    out->base_rng_count = rng_raw[2];
    out->cur_episode_idx = sim.curEpisodeIdx;
    out->world_episode_counter = sim.worldEpisodeCounter;
    out->world_curriculum = (int32_t)ctx.singleton<WorldCurriculum>();
    out->reset_flag = ctx.singleton<WorldReset>().reset;

    for (int t = 0; t < 2; t++) {
        out->filters_active[t] = sim.filtersState[t].active;
        for (int f = 0; f < 3; f++) {
            out->filters_last_matches[t][f] = sim.filtersState[t].lastMatches[f];
        }
        out->filters_last_matched_step[t] = sim.filtersLastMatchedStep[t];
    }

    const GoalRegionsState &gr = ctx.singleton<GoalRegionsState>();
    out->goal_min_dist[0] = gr.minDistToRegions[0];
    out->goal_min_dist[1] = gr.minDistToRegions[1];
    out->goal_team_step_rewards[0] = gr.teamStepRewards[0];
    out->goal_team_step_rewards[1] = gr.teamStepRewards[1];
    const TeamRewardState &tr = ctx.singleton<TeamRewardState>();
    out->team_rewards[0] = tr.teamRewards[0];
    out->team_rewards[1] = tr.teamRewards[1];
    for (int z = 0; z < 8; z++) {
        const SubZone &sz = ctx.get<SubZone>(sim.subZones[z]);
        out->sub_zones[z][0] = sz.curControllingTeam;
        out->sub_zones[z][1] = sz.isContested;
        out->sub_zones[z][2] = sz.isCaptured;
    }

    const SpawnUsageCounter &su = ctx.singleton<SpawnUsageCounter>();
    memcpy(out->spawn_a_last_used, su.initASpawnsLastUsedTick, sizeof(uint32_t) * 128);
    memcpy(out->spawn_b_last_used, su.initBSpawnsLastUsedTick, sizeof(uint32_t) * 128);
    memcpy(out->respawn_last_used, su.respawnLastUsedTick, sizeof(uint32_t) * 128);

    uint32_t crumb_arch = shim::TypeID<BreadcrumbEntity>::get();
    shim::Table &ct = w.tables[crumb_arch];
    int crumb_col = w.colIndex(crumb_arch, shim::TypeID<Breadcrumb>::get());
    out->num_breadcrumbs = ct.numRows;
    for (int32_t i = 0; i < ct.numRows && i < MPENV_MAX_BREADCRUMBS; i++) {
        const Breadcrumb &b = ((Breadcrumb *)ct.cols[crumb_col])[i];
        out->breadcrumbs[i].pos[0] = b.pos.x;
        out->breadcrumbs[i].pos[1] = b.pos.y;
        out->breadcrumbs[i].pos[2] = b.pos.z;
        out->breadcrumbs[i].penalty = b.penalty;
        out->breadcrumbs[i].team = b.teamInfo.team;
        out->breadcrumbs[i].team_offset = b.teamInfo.offset;
    }

    for (uint32_t i = 0; i < sim.numAgents && i < MPENV_AGENTS_PER_WORLD; i++) {
        Entity e = sim.agents[i];
        mpenv_agent_state &a = out->agents[i];
        Vector3 p = ctx.get<Position>(e);
        a.pos[0] = p.x; a.pos[1] = p.y; a.pos[2] = p.z;
        Quat r = ctx.get<Rotation>(e);
        a.rot[0] = r.w; a.rot[1] = r.x; a.rot[2] = r.y; a.rot[3] = r.z;
        Vector3 v = ctx.get<AgentVelocity>(e);
        a.vel[0] = v.x; a.vel[1] = v.y; a.vel[2] = v.z;
        const IntermediateMoveState &ms = ctx.get<IntermediateMoveState>(e);
        a.move_new_pos[0] = ms.newPosition.x; a.move_new_pos[1] = ms.newPosition.y;
        a.move_new_pos[2] = ms.newPosition.z;
        a.move_new_vel[0] = ms.newVelocity.x; a.move_new_vel[1] = ms.newVelocity.y;
        a.move_new_vel[2] = ms.newVelocity.z;
        a.move_max_vel = ms.maxVelocity;
        const Aim &aim = ctx.get<Aim>(e);
        a.aim_yaw = aim.yaw; a.aim_pitch = aim.pitch;
        a.aim_rot[0] = aim.rot.w; a.aim_rot[1] = aim.rot.x;
        a.aim_rot[2] = aim.rot.y; a.aim_rot[3] = aim.rot.z;
        a.hp = ctx.get<HP>(e).hp;
        a.alive = ctx.get<Alive>(e).mask;
        a.mag_bullets = ctx.get<Magazine>(e).numBullets;
        a.mag_reloading = ctx.get<Magazine>(e).isReloading;
        for (int j = 0; j < 6; j++) a.dmg[j] = ctx.get<DamageDealt>(e).dmg[j];
        const StandState &ss = ctx.get<StandState>(e);
        a.pose_cur = (int32_t)ss.curPose; a.pose_tgt = (int32_t)ss.tgtPose;
        a.pose_transition = ss.transitionRemaining;
        const CombatState &cs = ctx.get<CombatState>(e);
        memcpy(rng_raw, &cs.rng, 12);
        a.rng_key[0] = rng_raw[0]; a.rng_key[1] = rng_raw[1]; a.rng_count = rng_raw[2];
        a.landed_shot_on = agentIndexOf(sim, cs.landedShotOn);
        a.remaining_respawn_steps = cs.remainingRespawnSteps;
        a.remaining_autoheal_steps = cs.remainingStepsBeforeAutoheal;
        a.successful_kill = cs.successfulKill;
        a.was_shot_count = cs.wasShotCount;
        a.was_killed = cs.wasKilled;
        a.fired_shot_t = cs.firedShotT;
        a.in_zone = cs.inZone;
        a.min_dist_to_zone = cs.minDistToZone;
        a.has_died = cs.hasDiedDuringEpisode;
        a.reloaded_full_mag = cs.reloadedFullMag;
        a.weapon_type = cs.weaponType;
        a.in_sub_zone = cs.inSubZone;
        a.min_dist_to_sub_zone = cs.minDistToSubZone;
        const BreadcrumbAgentState &bs = ctx.get<BreadcrumbAgentState>(e);
        a.crumb_total_penalty = bs.totalPenalty;
        a.crumb_steps_since_new = bs.stepsSinceLastNewBreadcrumb;
        a.crumb_last = -1;
        if (bs.lastBreadcrumb != Entity::none() && w.entityValid(bs.lastBreadcrumb)) {
            a.crumb_last = w.slots[bs.lastBreadcrumb.id].row;
        }
        Vector3 sp = ctx.get<StartPos>(e);
        a.start_pos[0] = sp.x; a.start_pos[1] = sp.y; a.start_pos[2] = sp.z;
        a.explore_new_cells = ctx.get<ExploreTracker>(e).numNewCellsVisited;
        for (int j = 0; j < 6; j++) {
            a.can_see[j] = ctx.get<OpponentsVisibility>(e).canSee[j];
        }
        a.team = ctx.get<TeamInfo>(e).team;
        a.team_offset = ctx.get<TeamInfo>(e).offset;
        a.policy = ctx.get<AgentPolicy>(e).idx;
        memcpy(a.reward_coefs, &ctx.get<RewardHyperParams>(e), sizeof(float) * 9);
    }
    return 0;
}

// ---- geometry probes used to pin the BVH / traversal restatements ---------

// MeshBVH::traceRay on the static mesh only (mesh_bvh.inl:110-208).
int ref_trace_ray(void *h, const float *o, const float *d, float t_max, float *t_out)
{
    RefSim *s = (RefSim *)h;
    TraversalStack stack;
    stack.size = 0;
    MeshBVH::HitInfo hit;
    bool ok = s->bvh.traceRay({ o[0], o[1], o[2] }, { d[0], d[1], d[2] }, &hit,
                              &stack, t_max);
    *t_out = ok ? hit.tHit : FLT_MAX;
    return ok ? 1 : 0;
}

// MeshBVH::sphereCast (mesh_bvh.inl:743-815); normal_out untouched on a miss.
float ref_sphere_cast(void *h, const float *o, const float *d, float r, float *normal_out)
{
    RefSim *s = (RefSim *)h;
    Vector3 n { normal_out[0], normal_out[1], normal_out[2] };
    float t = s->bvh.sphereCast({ o[0], o[1], o[2] }, { d[0], d[1], d[2] }, r, &n);
    normal_out[0] = n.x; normal_out[1] = n.y; normal_out[2] = n.z;
    return t;
}

// Brute force over every triangle with the reference's own triangle routines
// (rayTriangleIntersection mesh_bvh.inl:433-554): pins traversal against
// topology.
int ref_trace_ray_brute(void *h, const float *o, const float *d, float *t_out)
{
    RefSim *s = (RefSim *)h;
    Vector3 ro { o[0], o[1], o[2] }, rd { d[0], d[1], d[2] };
    math::Diag3x3 inv_d = math::Diag3x3::fromVec(rd).inv();
    MeshBVH::RayIsectTxfm tx = MeshBVH::computeRayIsectTxfm(ro, rd, inv_d, s->bvh.rootAABB);
    float t_max = FLT_MAX;
    bool any = false;
    uint32_t num_tris = s->bvh.numVerts / 3;
    for (uint32_t i = 0; i < num_tris; i++) {
        float t;
        Vector3 bary, n;
        bool hit = s->bvh.rayTriangleIntersection(
            s->bvhVerts[3 * i].pos, s->bvhVerts[3 * i + 1].pos, s->bvhVerts[3 * i + 2].pos,
            tx.kx, tx.ky, tx.kz, tx.Sx, tx.Sy, tx.Sz, ro, t_max, &t, &bary, &n);
        if (hit) { any = true; t_max = t; }
    }
    *t_out = any ? t_max : FLT_MAX;
    return any ? 1 : 0;
}

float ref_sphere_cast_brute(void *h, const float *o, const float *d, float r)
{
    RefSim *s = (RefSim *)h;
    Vector3 ro { o[0], o[1], o[2] }, rd { d[0], d[1], d[2] };
    float t = FLT_MAX;
    uint32_t num_tris = s->bvh.numVerts / 3;
    Vector3 n;
    for (uint32_t i = 0; i < num_tris; i++) {
        t = s->bvh.sphereCastTriangle(
            s->bvhVerts[3 * i].pos, s->bvhVerts[3 * i + 1].pos, s->bvhVerts[3 * i + 2].pos,
            ro, rd, t, r, &n);
    }
    return t;
}

// rand::split_i / sampleI32 / sampleUniform probes for the Threefry KATs.
// Batched probes (one C loop instead of one ctypes call per ray): the >= 10^6-ray
// brute-force pins of tests/test_oracle_pins.py.  mode 0 = traversal, 1 = brute force.
void ref_trace_rays(void *h, int64_t n, const float *o, const float *d, int32_t mode,
                    float *t_out, int32_t *hit_out)
{
    for (int64_t i = 0; i < n; i++) {
        hit_out[i] = mode == 0 ? ref_trace_ray(h, o + 3 * i, d + 3 * i, FLT_MAX, t_out + i)
                               : ref_trace_ray_brute(h, o + 3 * i, d + 3 * i, t_out + i);
    }
}

// normal_out [n][3]: in = the caller's initial normal, out = contact normal (untouched on a miss)
void ref_sphere_casts(void *h, int64_t n, const float *o, const float *d, float r, int32_t mode,
                      float *t_out, float *normal_out)
{
    for (int64_t i = 0; i < n; i++) {
        t_out[i] = mode == 0 ? ref_sphere_cast(h, o + 3 * i, d + 3 * i, r, normal_out + 3 * i)
                             : ref_sphere_cast_brute(h, o + 3 * i, d + 3 * i, r);
    }
}

void ref_rand_split(uint32_t a, uint32_t b, uint32_t idx, uint32_t idx_upper, uint32_t *out)
{
    RandKey k = rand::split_i(RandKey { a, b }, idx, idx_upper);
    out[0] = k.a; out[1] = k.b;
}

int32_t ref_rand_sample_i32(uint32_t a, uint32_t b, int32_t lo, int32_t hi)
{
    return rand::sampleI32(RandKey { a, b }, lo, hi);
}

float ref_rand_uniform(uint32_t a, uint32_t b)
{
    return rand::sampleUniform(RandKey { a, b });
}

} // extern "C"
