// TEST INFRASTRUCTURE -- never built into, loaded by or linked with the product
// library (libmpenv_b200.so has no CPU execution path and fails without CUDA).
//
// hostsim compiles the SAME stage functions the CUDA kernels wrap
// (madrona-mp-env_b200/csrc/mp_stages*.h, mp_geom.h, mp_math.h) for the host,
// so the CPU test-suite (`pytest -m "not gpu"`) can diff the kernels' logic
// against the oracle in this GPU-less container.  The GPU parity tests diff
// the real kernels against the same oracle.
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../madrona-mp-env_b200/csrc/assets.h"
#include "../../madrona-mp-env_b200/csrc/mp_stages.h"
#include "../../madrona-mp-env_b200/csrc/sim_layout.h"

using namespace mpenv;

namespace {

struct HostSim {
    MapAssets assets;
    MapConsts consts;
    Ctx ctx;
    std::vector<void *> allocs;
    std::vector<CurriculumSnapshot> curriculum;
    uint32_t numWorlds;
    bool fullTeam;
};

void obsPhase(HostSim &s)
{
    const Ctx &c = s.ctx;
    const int W = (int)s.numWorlds;
    for (int w = 0; w < W; w++)
        for (int a = 0; a < kAgents; a++)
            for (int j = 0; j < kTeamSize; j++)
                c.p.canSee[(size_t)(w * kAgents + a) * 8 + j] =
                    visibilityPair(c, w, a, j, c.p.pos + (size_t)w * kAgents) ? 1 : 0;
    for (int w = 0; w < W; w++)
        for (int a = 0; a < kAgents; a++) stageObs(c, w, a);
    if (s.fullTeam)
        for (int w = 0; w < W; w++)
            for (int t = 0; t < 2; t++) stageFullTeamObs(c, w, t);
    for (int w = 0; w < W; w++)
        for (int a = 0; a < kAgents; a++) {
            const int idx = w * kAgents + a;
            for (int r = 0; r < 80; r++) {
                LidarSample smp = lidarRay(c, w, a, r, c.p.pos + (size_t)w * kAgents);
                if (r < 64) c.p.fwdLidar[(size_t)idx * 64 + r] = smp;
                else c.p.rearLidar[(size_t)idx * 16 + (r - 64)] = smp;
            }
        }
}

}  // namespace

extern "C" {

struct HostSimConfig {
    uint32_t numWorlds, randSeed;
    int32_t autoReset;
    uint32_t simFlags, worldIDOffset, enableFullTeam;
    uint32_t enableEventLog, enableRecordLog, enableReplayLog;
    uint32_t numCurriculumSnapshots;
    const void *curriculumSnapshots;
    uint32_t trainFlank;
    uint32_t taskType;          // 2 = Zone, 4 = ZoneCaptureDefend
};

void *hostsim_create(const HostSimConfig *cfg, const char *scene_dir)
{
    HostSim *s = new HostSim();
    std::string err;
    if (!loadMapAssets(scene_dir, &s->assets, &err) || !fillMapConsts(s->assets, &s->consts, &err)) {
        fprintf(stderr, "hostsim_create: %s\n", err.c_str());
        delete s;
        return nullptr;
    }
    s->numWorlds = cfg->numWorlds;
    s->fullTeam = cfg->enableFullTeam != 0;
    Ctx &c = s->ctx;
    memset(&c, 0, sizeof(c));
    c.cfg.numWorlds = cfg->numWorlds;
    c.cfg.worldIDOffset = cfg->worldIDOffset;
    c.cfg.simFlags = cfg->simFlags;
    c.cfg.autoReset = cfg->autoReset;
    c.cfg.initRandKey = randSplit(randInitKey(cfg->randSeed), 0);    // mgr.cpp:1736-1738
    c.cfg.eventLog = cfg->enableEventLog ? 1 : 0;
    c.cfg.recordLog = cfg->enableRecordLog ? 1 : 0;
    c.cfg.replayLog = cfg->enableReplayLog ? 1 : 0;
    c.cfg.rewardFlank = cfg->trainFlank ? 1 : 0;
    c.cfg.taskCaptureDefend = cfg->taskType == 4 ? 1 : 0;
    c.mc = &s->consts;
    c.bvh.nodes = s->assets.nodes.data();
    c.bvh.verts = s->assets.verts.data();
    c.bvh.rootAABB = s->assets.rootAABB;
    c.bvh.numNodes = (uint32_t)s->assets.nodes.size();
    c.bvh.numTris = s->assets.numTris;
    if (s->assets.hasNav) {
        c.nav.verts = s->assets.nav.verts.data();
        c.nav.tris = s->assets.nav.tris.data();
        c.nav.lookup = s->assets.nav.lookup.data();
        c.nav.numTris = s->assets.nav.numTris();
    }
    int n;
    const BufferDesc *table = bufferTable(&n);
    for (int i = 0; i < n; i++) {
        void *mem = calloc(bufferRows(table[i], cfg->numWorlds), table[i].rowBytes);
        *bufferSlot(&c.p, table[i]) = mem;
        s->allocs.push_back(mem);
    }
    if (cfg->numCurriculumSnapshots > 0 && cfg->curriculumSnapshots) {
        s->curriculum.resize(cfg->numCurriculumSnapshots);
        memcpy(s->curriculum.data(), cfg->curriculumSnapshots, sizeof(CurriculumSnapshot) * cfg->numCurriculumSnapshots);
        c.p.curriculum = s->curriculum.data();
        c.cfg.numCurriculumSnapshots = (int32_t)cfg->numCurriculumSnapshots;
    }
    // mgr.cpp:1397-1413
    c.p.trainCtrl->evalMode = (cfg->simFlags & kFlagSimEvalMode) ? 1 : 0;
    c.p.trainCtrl->randomizeEpisodeLengthAfterReset = (cfg->simFlags & kFlagStaggerStarts) ? 1 : 0;
    c.p.trainCtrl->randomizeTeamSides = (cfg->simFlags & kFlagRandomFlipTeams) ? 1 : 0;
    for (uint32_t w = 0; w < cfg->numWorlds; w++) ctorWorld(c, (int)w);
    return s;
}

void hostsim_destroy(void *h)
{
    HostSim *s = (HostSim *)h;
    if (!s) return;
    for (void *m : s->allocs) free(m);
    delete s;
}

void hostsim_init(void *h)
{
    HostSim *s = (HostSim *)h;
    for (uint32_t w = 0; w < s->numWorlds; w++) s->ctx.p.resetFlag[w] = 1;
    for (uint32_t w = 0; w < s->numWorlds; w++) resetSystem(s->ctx, (int)w);
    obsPhase(*s);
}

void hostsim_step(void *h)
{
    HostSim *s = (HostSim *)h;
    const Ctx &c = s->ctx;
    const int W = (int)s->numWorlds;
    if (!c.cfg.replayLog) {
        for (int i = 0; i < W * kAgents; i++) stageMove(c, i);
        for (int w = 0; w < W; w++)
            for (int a = 0; a < kAgents; a++) stageFire(c, w, a, c.p.pos + (size_t)w * kAgents);
    } else if (c.cfg.eventLog) {
        // no fireSystem in the replay graph: no fire events this step
        for (size_t i = 0; i < (size_t)W * kAgents * kEventSlotsPerAgent; i++) c.p.evAgent[i].type = kEvNone;
    }
    for (int w = 0; w < W; w++) stageWorld(c, w);
    obsPhase(*s);
}

// The step's events / step snapshots compacted in world-major, agent-major slot
// order (what k_event_scan / k_event_scatter produce on the device).
int64_t hostsim_read_event_log(void *h, uint32_t which, void *dst, uint64_t cap_rows)
{
    HostSim *s = (HostSim *)h;
    const SimPtrs &p = s->ctx.p;
    uint64_t n = 0;
    for (uint32_t w = 0; w < s->numWorlds; w++) {
        if (which == 0) {
            for (int k = 0; k < kEventSlotsPerWorld; k++) {
                GameEvent e = k < kAgents * kEventSlotsPerAgent
                    ? p.evAgent[(size_t)w * kAgents * kEventSlotsPerAgent + k] : p.evWorld[w];
                if (e.type == kEvNone) continue;
                if (n >= cap_rows) return -1;
                e.pad0 = 0;
                ((GameEvent *)dst)[n++] = e;
            }
        } else if (p.evWorld[w].pad0) {
            if (n >= cap_rows) return -1;
            ((PackedStepSnapshot *)dst)[n++] = p.stepSnap[w];
        }
    }
    return (int64_t)n;
}

void *hostsim_record_log(void *h) { return ((HostSim *)h)->ctx.p.recordLog; }
void *hostsim_replay_log(void *h) { return ((HostSim *)h)->ctx.p.replayLog; }

int hostsim_export_bytes(void *h, uint32_t id, uint64_t *bytes)
{
    HostSim *s = (HostSim *)h;
    ExportDesc d;
    if (!exportDesc(s->ctx.p, s->numWorlds, id, &d)) return -1;
    uint64_t n = 4;
    for (int i = 0; i < d.ndim; i++) n *= (uint64_t)d.dims[i];
    *bytes = n;
    return 0;
}

int hostsim_read_export(void *h, uint32_t id, void *dst, uint64_t dst_bytes)
{
    HostSim *s = (HostSim *)h;
    ExportDesc d;
    uint64_t n;
    if (!exportDesc(s->ctx.p, s->numWorlds, id, &d) || hostsim_export_bytes(h, id, &n)) return -1;
    if (n > dst_bytes) return -2;
    memcpy(dst, d.ptr, n);
    return 0;
}

int hostsim_write_export(void *h, uint32_t id, const void *src, uint64_t src_bytes)
{
    HostSim *s = (HostSim *)h;
    ExportDesc d;
    uint64_t n;
    if (!exportDesc(s->ctx.p, s->numWorlds, id, &d) || hostsim_export_bytes(h, id, &n)) return -1;
    if (n != src_bytes) return -2;
    memcpy(d.ptr, src, n);
    return 0;
}

int hostsim_get_world_state(void *h, uint32_t w, mpenv_world_state *out)
{
    HostSim *s = (HostSim *)h;
    if (w >= s->numWorlds) return -1;
    SimPtrs one = s->ctx.p;
    int n;
    const BufferDesc *table = bufferTable(&n);
    for (int i = 0; i < n; i++) {
        size_t rows_per_world = table[i].kind == kPerAgent ? kAgents : (table[i].kind == kPerTeam ? 2 : 1);
        if (table[i].kind == kGlobal) continue;
        char *basep = (char *)*bufferSlot(&s->ctx.p, table[i]);
        *bufferSlot(&one, table[i]) = basep + (size_t)w * rows_per_world * table[i].rowBytes;
    }
    fillWorldState(one, out);
    return 0;
}

// ---- geometry probes ------------------------------------------------------
int hostsim_trace_ray(void *h, const float *o, const float *d, float *t_out)
{
    HostSim *s = (HostSim *)h;
    return bvhTraceRay(s->ctx.bvh, v3(o[0], o[1], o[2]), v3(d[0], d[1], d[2]), t_out) ? 1 : 0;
}

float hostsim_sphere_cast(void *h, const float *o, const float *d, float r, float *n_io)
{
    HostSim *s = (HostSim *)h;
    Vec3 n = v3(n_io[0], n_io[1], n_io[2]);
    float t = bvhSphereCast(s->ctx.bvh, v3(o[0], o[1], o[2]), v3(d[0], d[1], d[2]), r, &n);
    n_io[0] = n.x; n_io[1] = n.y; n_io[2] = n.z;
    return t;
}

void hostsim_trace_rays(void *h, int64_t n, const float *o, const float *d, float *t_out, int32_t *hit_out)
{
    for (int64_t i = 0; i < n; i++) hit_out[i] = hostsim_trace_ray(h, o + 3 * i, d + 3 * i, t_out + i);
}

void hostsim_sphere_casts(void *h, int64_t n, const float *o, const float *d, float r, float *t_out, float *n_io)
{
    for (int64_t i = 0; i < n; i++) t_out[i] = hostsim_sphere_cast(h, o + 3 * i, d + 3 * i, r, n_io + 3 * i);
}

void hostsim_rand_split(uint32_t a, uint32_t b, uint32_t idx, uint32_t idx_upper, uint32_t *out)
{
    RandKey k; k.a = a; k.b = b;
    RandKey r = randSplit(k, idx, idx_upper);
    out[0] = r.a; out[1] = r.b;
}
int32_t hostsim_rand_sample_i32(uint32_t a, uint32_t b, int32_t lo, int32_t hi)
{
    RandKey k; k.a = a; k.b = b;
    return randI32FromKey(k, lo, hi);
}
float hostsim_rand_uniform(uint32_t a, uint32_t b)
{
    RandKey k; k.a = a; k.b = b;
    return randUniformFromKey(k);
}

}  // extern "C"
