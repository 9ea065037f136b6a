// mpenv_capi.cu -- the C ABI of libmpenv_b200.so (include/mpenv.h): device
// state ownership, kernel launch sequence, CUDA-graph capture of the step, and
// the tensor table.  Replaces Manager::Impl / CUDAImpl of the reference
// (src/mgr.cpp:305-662, 1213-1912) for the per-step simulation path.
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include <cstdlib>

#include <nvtx3/nvToolsExt.h>

#include "step_kernels.cuh"
#include "lidar_packet.cuh"
#include "assets.h"
#include "sim_layout.h"
#include "../../include/mpenv.h"

using namespace mpenv;

namespace mpenv {
const MapAssets &mapAssetsOf(const mpenv_map *m);
}

static thread_local std::string g_create_error;

struct mpenv_handle {
    mpenv_config cfg;
    std::string sceneDir;
    std::string error;
    int device = 0;
    int numSMs = 0;
    cudaStream_t stream = nullptr;
    MapAssets assets;
    MapConsts hostConsts;
    MapConsts *devConsts = nullptr;
    BVHNode *devNodes = nullptr;
    Vec3 *devVerts = nullptr;
    KernelArgs ka;
    std::vector<void *> allocs;
    struct Guarded { char *raw; size_t bytes; };
    std::vector<Guarded> guarded;            // buffers between guard bands (mpenv_check_guards)
    float *agentMap = nullptr;               // [N,16,16,4] zeros (never written, mgr.cpp:2422)
    unsigned long long *devStats = nullptr;
    size_t bvhSmemBytes = 0;
    cudaGraphExec_t stepGraph = nullptr;
    cudaGraphExec_t initGraph = nullptr;
    bool fullTeam = false;
    // pipelined host read-back (mpenv_read_tensors_pipelined)
    cudaStream_t copyStream = nullptr;
    cudaStream_t lidarStream = nullptr;      // captured into the graphs as the lidar branch (launchObsPhase)
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    cudaEvent_t evSnap[2] = { nullptr, nullptr }, evRead[2] = { nullptr, nullptr }, evCopied = nullptr;
    char *snapshot[2] = { nullptr, nullptr };    // double-buffered: snapshot t+1 is taken while t goes to the host
    size_t snapshotBytes = 0;
    int snapNext = 0;                            // buffer the next snapshot goes into
    int snapReady = -1;                          // buffer holding a snapshot that has not been read yet
    uint32_t snapIds[64];
    uint32_t snapCount = 0;
    bool snapRead[2] = { false, false };         // a D2H from this buffer has been enqueued (evRead valid)
    bool copyInFlight = false;
    int packetLidar = 0;                     // packet traversal (lidar_packet.cuh): CTAs per SM (2, 1), 0 = per-thread k_lidar
    size_t lidarPacketSmem = 0;
    int gridLidarPacket = 0;
    bool poolLos = false;
    size_t losPoolSmem = 0;
    bool fastFire = false;                   // k_fire_fast: the shot on the FastBVH in shared memory (ray_kernels.cuh)
    size_t fireSmem = 0;
    bool fastMove = false;                   // k_move<true>: SphereNodes in shared memory (sphere_fast.cuh)
    size_t moveSmem = 0;
    int losWarps = 0;                        // warps per k_los_pool CTA (what fits beside the BVH)
    int gridLosPool = 0;
    // launch geometry
    int gridMove = 0, gridFire = 0, gridWorld = 0, gridLos = 0, gridObs = 0, gridTeam = 0, gridLidar = 0,
        gridSmall = 0;
    uint32_t stepLaunches = 0, initLaunches = 0;
    // logs (event log + step snapshots, record log, replay log)
    bool eventLog = false, recordLog = false, replayLog = false;
    std::string eventDir, recordPath, replayPath;
    FILE *evFile = nullptr, *stepsFile = nullptr, *recFile = nullptr, *replayFile = nullptr;
    uint32_t *devEvOffset = nullptr, *devSnapOffset = nullptr, *devEvTotals = nullptr;
    GameEvent *devEvOut = nullptr;
    PackedStepSnapshot *devSnapOut = nullptr;
    uint32_t *hostEvTotals = nullptr;        // pinned: [0] events, [1] step snapshots
    GameEvent *hostEvents = nullptr;         // pinned
    PackedStepSnapshot *hostSnaps = nullptr; // pinned
    StepLog *hostRecord = nullptr;           // pinned
    StepLog *hostReplay = nullptr;           // pinned
    std::string curriculumPath;
    uint32_t lastNumEvents = 0, lastNumSnaps = 0;
    bool cold = false;                       // any log / replay / curriculum feature on: the kCold kernel variants
    bool replayPending = false;              // mpenv_set_replay_log() data waiting for the next step
};

constexpr size_t kGuardBytes = 256;

// ---- episode statistics without any allocation (mpenv_episode_stats_static) ------------------
// k_stats accumulates into a cudaMalloc'ed buffer with atomics.  This variant keeps its scratch --
// one row of partial sums per CTA plus a ticket -- in a file-scope device array that every GPU
// loading the library holds until exit (kStaticStatsBytes = 33 KB), and writes the totals to the
// CALLER's buffer: each CTA stores its partials, the last CTA to take a ticket adds the rows up in
// CTA order.  In host code `g_static_stats` is only a stand-in for the symbol: the device address
// comes from cudaGetSymbolAddress, once per device (staticStatsScratch below).
constexpr int kStaticStatsMaxCTAs = 256;
constexpr int kStaticStatsRow = MPENV_NUM_STATS;                       // u64 per CTA
constexpr size_t kStaticStatsWords = (size_t)kStaticStatsMaxCTAs * kStaticStatsRow + 32;   // + the ticket's own 256 B
constexpr size_t kStaticStatsBytes = kStaticStatsWords * sizeof(unsigned long long);
static __device__ __align__(256) unsigned long long g_static_stats[kStaticStatsWords];

namespace mpenv {
__global__ void __launch_bounds__(256) k_stats_static(KernelArgs ka, unsigned long long *scratch, unsigned long long *out)
{
    __shared__ long long warp_sums[8][13];
    __shared__ bool last_cta;
    const Ctx &c = ka.ctx;
    const int W = (int)c.cfg.numWorlds;
    unsigned long long *ticket = scratch + (size_t)kStaticStatsMaxCTAs * kStaticStatsRow;
    long long acc[13];
    for (int k = 0; k < 13; k++) acc[k] = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < W; w += gridDim.x * blockDim.x) {
        acc[0] += 1;
        const MatchResult &mr = c.p.matchResult[w];
        if (c.p.done[w * kAgents]) {
            acc[1] += 1;
            if (mr.winResult >= 0 && mr.winResult <= 2) acc[2 + mr.winResult] += 1;
            acc[5] += mr.teamTotalKills[0]; acc[6] += mr.teamTotalKills[1];
            acc[7] += mr.teamObjectivePoints[0]; acc[8] += mr.teamObjectivePoints[1];
        }
        for (int a = 0; a < kAgents; a++) {
            const int idx = w * kAgents + a;
            acc[9] += c.p.alive[idx] != 0.f ? 1 : 0;
            acc[10] += c.p.done[idx];
            acc[11] += (long long)llrintf(c.p.reward[idx] * 1.0e6f);
        }
        acc[12] += c.p.world[w].match.curStep;
    }
    for (int k = 0; k < 13; k++) {
        long long v = acc[k];
        for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, off);
        if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < kStaticStatsRow) {
        long long v = 0;
        if (threadIdx.x < 13)
            for (int wp = 0; wp < 8; wp++) v += warp_sums[wp][threadIdx.x];
        scratch[(size_t)blockIdx.x * kStaticStatsRow + threadIdx.x] = (unsigned long long)v;   // every word of the row
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) last_cta = atomicAdd(ticket, 1ull) == (unsigned long long)gridDim.x - 1ull;
    __syncthreads();
    if (!last_cta) return;
    __threadfence();
    if (threadIdx.x < kStaticStatsRow) {
        unsigned long long v = 0;
        for (unsigned b = 0; b < gridDim.x; b++) v += __ldcg(scratch + (size_t)b * kStaticStatsRow + threadIdx.x);
        out[threadIdx.x] = v;
    }
    if (threadIdx.x == 0) *ticket = 0;
}
}  // namespace mpenv

namespace {

// NVTX range over a host-side entry point (nsys / ncu timelines; a no-op without a profiler attached).
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

bool cudaOk(mpenv_handle *h, cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return true;
    char buf[512];
    snprintf(buf, sizeof(buf), "%s: %s", what, cudaGetErrorString(e));
    if (h) h->error = buf;
    g_create_error = buf;
    return false;
}

#define MP_CUDA(h, call) do { if (!cudaOk((h), (call), #call)) return MPENV_ERR_CUDA; } while (0)

int fail(mpenv_handle *h, int code, const std::string &msg)
{
    if (h) h->error = msg;
    g_create_error = msg;
    return code;
}

// Depth of the 4-wide tree (root = 1): bounds the packet kernel's shared stack.
int bvhDepth(const std::vector<BVHNode> &nodes)
{
    int best = 0;
    std::vector<std::pair<int, int>> todo(1, { 0, 1 });
    while (!todo.empty()) {
        const auto [n, d] = todo.back();
        todo.pop_back();
        if (d > best) best = d;
        for (int k = 0; k < 4; k++) {
            const int32_t ch = nodes[(size_t)n].children[k];
            if (ch != kBVHSentinel && !(ch & 0x80000000)) todo.push_back({ (int)ch, d + 1 });
        }
    }
    return best;
}

template <typename K>
int gridFor(mpenv_handle *h, K kernel, int block, size_t smem, long long items)
{
    int per_sm = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, smem) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    long long want = (items + block - 1) / block;
    long long cap = (long long)h->numSMs * per_sm;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

void launchMove(mpenv_handle *h, cudaStream_t s)
{
    if (h->fastMove) k_move<true><<<h->gridMove, 128, h->moveSmem, s>>>(h->ka);
    else k_move<false><<<h->gridMove, 128, h->bvhSmemBytes, s>>>(h->ka);
}

void launchFire(mpenv_handle *h, cudaStream_t s)
{
    if (h->fastFire) {
        if (h->cold) k_fire_fast<true><<<h->gridFire, 128, h->fireSmem, s>>>(h->ka);
        else k_fire_fast<false><<<h->gridFire, 128, h->fireSmem, s>>>(h->ka);
    } else {
        if (h->cold) k_fire<true><<<h->gridFire, 128, h->bvhSmemBytes, s>>>(h->ka);
        else k_fire<false><<<h->gridFire, 128, h->bvhSmemBytes, s>>>(h->ka);
    }
}

void launchLidar(mpenv_handle *h, cudaStream_t s)
{
    if (h->packetLidar == 2)
        k_lidar_packet<kPacketWarpsPerCTA, MPENV_PACKET_MINB><<<h->gridLidarPacket, kPacketWarpsPerCTA * 32, h->lidarPacketSmem, s>>>(h->ka);
    else if (h->packetLidar == 1)
        k_lidar_packet<32, 1><<<h->gridLidarPacket, 1024, h->lidarPacketSmem, s>>>(h->ka);
    else k_lidar<<<h->gridLidar, 160, h->bvhSmemBytes, s>>>(h->ka);
}

void launchLos(mpenv_handle *h, cudaStream_t s)
{
    if (h->poolLos) k_los_pool<10, 8><<<h->gridLosPool, h->losWarps * 32, h->losPoolSmem, s>>>(h->ka);
    else k_los<<<h->gridLos, 128, h->bvhSmemBytes, s>>>(h->ka);
}

// The obs phase shared by the Init and Step graphs (resetAndObsTasks,
// sim.cpp:5174-5320): visibility -> masks + observations -> full-team -> lidar.
uint32_t launchObsPhase(mpenv_handle *h, cudaStream_t s)
{
    uint32_t n = 0;
    if (h->lidarStream) {
        // Lidar needs nothing from visibility / observations (it reads what k_world left), so the
        // graph forks: {visibility -> observation rows} beside {lidar}.  The visibility CTAs and the
        // lidar CTAs each fill an SM, so those two still run one after the other, but the hand-out
        // tail of visibility and the whole latency-bound observation kernel (24 % issue) now run
        // under lidar's ramp-up instead of in front of it: 1.233 -> 1.207 ms per step.
        cudaEventRecord(h->evFork, s);
        cudaStreamWaitEvent(h->lidarStream, h->evFork, 0);
        launchLos(h, s); n++;
        launchLidar(h, h->lidarStream); n++;
        k_obs<<<h->gridObs, 128, 0, s>>>(h->ka); n++;
        cudaEventRecord(h->evJoin, h->lidarStream);
        cudaStreamWaitEvent(s, h->evJoin, 0);
        return n;
    }
    // with the full-team tensors: k_fullteam gathers the PREVIOUS step's lidar rows (the reference's
    // task order, sim.cpp:5174-5320), so lidar stays behind it
    launchLos(h, s); n++;
    k_obs<<<h->gridObs, 128, 0, s>>>(h->ka); n++;
    if (h->fullTeam) { k_fullteam<<<h->gridTeam, 128, 0, s>>>(h->ka); n++; }
    launchLidar(h, s); n++;
    return n;
}

uint32_t launchStep(mpenv_handle *h, cudaStream_t s)
{
    uint32_t n = 0;
    if (!h->replayLog) {
        launchMove(h, s); n++;
        launchFire(h, s);
        n++;
    } else if (h->eventLog) {
        // pvpReplayLogic has no fireSystem: no fire events this step
        cudaMemsetAsync(h->ka.ctx.p.evAgent, 0,
                        (size_t)h->cfg.num_worlds * kAgents * kEventSlotsPerAgent * sizeof(GameEvent), s);
    }
    if (h->cold) k_world<true><<<h->gridWorld, 128, 0, s>>>(h->ka);
    else k_world<false><<<h->gridWorld, 128, 0, s>>>(h->ka);
    n++;
    if (h->eventLog) {
        k_event_scan<<<1, 1024, 0, s>>>(h->ka, h->devEvOffset, h->devSnapOffset, h->devEvTotals); n++;
        k_event_scatter<<<h->gridSmall, 128, 0, s>>>(h->ka, h->devEvOffset, h->devSnapOffset, h->devEvOut,
                                                      h->devSnapOut); n++;
    }
    n += launchObsPhase(h, s);
    return n;
}

// Host side of the logs around one step (CPUImpl::run mgr.cpp:242-271, CUDAImpl::saveEvents
// mgr.cpp:417-470).  Before: the replay log's next StepLog[W] goes to the device.
int logsBeforeStep(mpenv_handle *h, cudaStream_t s)
{
    if (!h->replayLog) return MPENV_OK;
    const size_t bytes = (size_t)h->cfg.num_worlds * sizeof(StepLog);
    if (h->replayFile) {
        if (fread(h->hostReplay, 1, bytes, h->replayFile) != bytes)
            return fail(h, MPENV_ERR_IO, "replay log exhausted: " + h->replayPath);
    } else if (!h->replayPending) {
        return fail(h, MPENV_ERR_INVALID, "replay from memory: call mpenv_set_replay_log() before every step");
    }
    h->replayPending = false;
    MP_CUDA(h, cudaMemcpyAsync(h->ka.ctx.p.replayLog, h->hostReplay, bytes, cudaMemcpyHostToDevice, s));
    return MPENV_OK;
}

// After: counts, then the compacted events / snapshots and the record log come
// back to pinned host memory and are appended to their files.  Synchronises `s`.
int logsAfterStep(mpenv_handle *h, cudaStream_t s)
{
    if (!h->eventLog && !h->recordLog) {
        // replay: hostReplay is reused by the next step's read
        if (h->replayLog) MP_CUDA(h, cudaStreamSynchronize(s));
        return MPENV_OK;
    }
    const size_t W = h->cfg.num_worlds;
    if (h->eventLog)
        MP_CUDA(h, cudaMemcpyAsync(h->hostEvTotals, h->devEvTotals, 8, cudaMemcpyDeviceToHost, s));
    if (h->recordLog)
        MP_CUDA(h, cudaMemcpyAsync(h->hostRecord, h->ka.ctx.p.recordLog, W * sizeof(StepLog), cudaMemcpyDeviceToHost, s));
    MP_CUDA(h, cudaStreamSynchronize(s));
    if (h->eventLog) {
        h->lastNumEvents = h->hostEvTotals[0];
        h->lastNumSnaps = h->hostEvTotals[1];
        if (h->lastNumEvents)
            MP_CUDA(h, cudaMemcpyAsync(h->hostEvents, h->devEvOut, (size_t)h->lastNumEvents * sizeof(GameEvent),
                                       cudaMemcpyDeviceToHost, s));
        if (h->lastNumSnaps)
            MP_CUDA(h, cudaMemcpyAsync(h->hostSnaps, h->devSnapOut, (size_t)h->lastNumSnaps * sizeof(PackedStepSnapshot),
                                       cudaMemcpyDeviceToHost, s));
        MP_CUDA(h, cudaStreamSynchronize(s));
        // writeGameEvents (mgr.cpp:104-116)
        if (h->stepsFile && fwrite(h->hostSnaps, sizeof(PackedStepSnapshot), h->lastNumSnaps, h->stepsFile) != h->lastNumSnaps)
            return fail(h, MPENV_ERR_IO, "short write to " + h->eventDir + "/steps.bin");
        if (h->evFile && fwrite(h->hostEvents, sizeof(GameEvent), h->lastNumEvents, h->evFile) != h->lastNumEvents)
            return fail(h, MPENV_ERR_IO, "short write to " + h->eventDir + "/events.bin");
    }
    if (h->recFile && fwrite(h->hostRecord, sizeof(StepLog), W, h->recFile) != W)
        return fail(h, MPENV_ERR_IO, "short write to " + h->recordPath);
    return MPENV_OK;
}

uint32_t launchInit(mpenv_handle *h, cudaStream_t s)
{
    uint32_t n = 0;
    k_force_reset<<<h->gridSmall, 128, 0, s>>>(h->ka); n++;
    k_reset<<<h->gridSmall, 128, 0, s>>>(h->ka); n++;
    n += launchObsPhase(h, s);
    return n;
}

int captureGraph(mpenv_handle *h, bool step, cudaGraphExec_t *out, uint32_t *launches)
{
    cudaGraph_t graph = nullptr;
    MP_CUDA(h, cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    *launches = step ? launchStep(h, h->stream) : launchInit(h, h->stream);
    MP_CUDA(h, cudaStreamEndCapture(h->stream, &graph));
    MP_CUDA(h, cudaGraphInstantiate(out, graph, 0));
    MP_CUDA(h, cudaGraphDestroy(graph));
    return MPENV_OK;
}

const mpenv_named_tensor kTrainInterface[] = {
    // inputs (mgr.cpp:2386-2396)
    { "discrete", MPENV_EXPORT_PVP_DISCRETE_ACTION, 1 },
    { "aim", MPENV_EXPORT_PVP_DISCRETE_AIM_ACTION, 1 },
    { "resets", MPENV_EXPORT_RESET, 1 },
    { "sim_ctrl", MPENV_EXPORT_SIM_CONTROL, 1 },
    { "policy_assignments", MPENV_EXPORT_AGENT_POLICY, 1 },
    // outputs (mgr.cpp:2398-2430)
    { "fwd_lidar", MPENV_EXPORT_FWD_LIDAR, 0 },
    { "rear_lidar", MPENV_EXPORT_REAR_LIDAR, 0 },
    { "hp", MPENV_EXPORT_HP, 0 },
    { "magazine", MPENV_EXPORT_MAGAZINE, 0 },
    { "alive", MPENV_EXPORT_ALIVE, 0 },
    { "self", MPENV_EXPORT_SELF_OBSERVATION, 0 },
    { "filters_state", MPENV_EXPORT_FILTERS_STATE_OBSERVATION, 0 },
    { "teammates", MPENV_EXPORT_TEAMMATE_OBSERVATIONS, 0 },
    { "opponents", MPENV_EXPORT_OPPONENT_OBSERVATIONS, 0 },
    { "opponents_last_known", MPENV_EXPORT_OPPONENT_LAST_KNOWN_OBSERVATIONS, 0 },
    { "self_pos", MPENV_EXPORT_SELF_POSITION_OBSERVATION, 0 },
    { "teammate_positions", MPENV_EXPORT_TEAMMATE_POSITION_OBSERVATIONS, 0 },
    { "opponent_positions", MPENV_EXPORT_OPPONENT_POSITION_OBSERVATIONS, 0 },
    { "opponent_last_known_positions", MPENV_EXPORT_OPPONENT_LAST_KNOWN_POSITION_OBSERVATIONS, 0 },
    { "opponent_masks", MPENV_EXPORT_OPPONENT_MASKS, 0 },
    { "agent_map", MPENV_EXPORT_AGENT_MAP, 0 },
    { "unmasked_agent_map", MPENV_EXPORT_AGENT_MAP, 0 },     // bound to agentMapTensor (mgr.cpp:2422)
    { "reward_coefs", MPENV_EXPORT_REWARD_HYPER_PARAMS, 0 },
    { "rewards", MPENV_EXPORT_REWARD, 0 },
    { "dones", MPENV_EXPORT_DONE, 0 },
    { "episode_results", MPENV_EXPORT_MATCH_RESULT, 0 },
};
constexpr uint32_t kNumTrainTensors = sizeof(kTrainInterface) / sizeof(kTrainInterface[0]);

size_t tensorBytes(const mpenv_tensor_desc &d)
{
    size_t n = 4;
    for (int i = 0; i < d.ndim; i++) n *= (size_t)d.dims[i];
    return n;
}

int copyInterface(mpenv_handle *h, cudaStream_t s, void **buffers, bool inputs)
{
    for (uint32_t i = 0; i < kNumTrainTensors; i++) {
        if ((kTrainInterface[i].is_input != 0) != inputs) continue;
        if (!buffers[i]) continue;          // NULL = caller aliases the sim tensor directly
        mpenv_tensor_desc d;
        int rc = mpenv_tensor(h, kTrainInterface[i].export_id, &d);
        if (rc != MPENV_OK) return rc;
        if (buffers[i] == d.dev_ptr) continue;
        if (inputs) MP_CUDA(h, cudaMemcpyAsync(d.dev_ptr, buffers[i], tensorBytes(d), cudaMemcpyDeviceToDevice, s));
        else MP_CUDA(h, cudaMemcpyAsync(buffers[i], d.dev_ptr, tensorBytes(d), cudaMemcpyDeviceToDevice, s));
    }
    return MPENV_OK;
}

}  // namespace

namespace {

// Device address of g_static_stats on `device` (which must be current): resolved once per device.
constexpr int kStaticStatsMaxDevices = 64;
unsigned long long *g_static_stats_ptr[kStaticStatsMaxDevices] = {};
std::mutex g_static_stats_mutex;

int staticStatsScratch(mpenv_handle *h, unsigned long long **out)
{
    if (h->device < 0 || h->device >= kStaticStatsMaxDevices)
        return fail(h, MPENV_ERR_INVALID, "mpenv_episode_stats_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_static_stats_mutex);
    if (!g_static_stats_ptr[h->device]) {
        unsigned long long *ptr = nullptr;
        MP_CUDA(h, cudaGetSymbolAddress((void **)&ptr, g_static_stats));
        if (!ptr) return fail(h, MPENV_ERR_CUDA, "mpenv_episode_stats_static: cudaGetSymbolAddress returned a null address");
        g_static_stats_ptr[h->device] = ptr;
    }
    *out = g_static_stats_ptr[h->device];
    return MPENV_OK;
}

}  // namespace

extern "C" {

const char *mpenv_last_error(mpenv_handle *h) { return h ? h->error.c_str() : g_create_error.c_str(); }

void mpenv_destroy(mpenv_handle *h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->stepGraph) cudaGraphExecDestroy(h->stepGraph);
    if (h->initGraph) cudaGraphExecDestroy(h->initGraph);
    for (void *p : h->allocs) cudaFree(p);
    for (FILE *f : { h->evFile, h->stepsFile, h->recFile, h->replayFile }) if (f) fclose(f);
    for (void *p : { (void *)h->hostEvTotals, (void *)h->hostEvents, (void *)h->hostSnaps, (void *)h->hostRecord,
                     (void *)h->hostReplay }) if (p) cudaFreeHost(p);
    if (h->copyStream) { cudaStreamSynchronize(h->copyStream); cudaStreamDestroy(h->copyStream); }
    for (int k = 0; k < 2; k++) {
        if (h->evSnap[k]) cudaEventDestroy(h->evSnap[k]);
        if (h->evRead[k]) cudaEventDestroy(h->evRead[k]);
        if (h->snapshot[k]) cudaFree(h->snapshot[k]);
    }
    if (h->evCopied) cudaEventDestroy(h->evCopied);
    if (h->evFork) cudaEventDestroy(h->evFork);
    if (h->evJoin) cudaEventDestroy(h->evJoin);
    if (h->lidarStream) cudaStreamDestroy(h->lidarStream);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int mpenv_create(const mpenv_config *cfg, mpenv_handle **out)
{
    NvtxRange nvtx("mpenv_create");
    if (!cfg || !out || !cfg->scene_dir) return fail(nullptr, MPENV_ERR_INVALID, "null argument");
    if (cfg->task_type != MPENV_TASK_ZONE && cfg->task_type != MPENV_TASK_ZONE_CAPTURE_DEFEND)
        return fail(nullptr, MPENV_ERR_INVALID, "only Task::Zone and Task::ZoneCaptureDefend are implemented");
    if (cfg->team_size != (uint32_t)kTeamSize) return fail(nullptr, MPENV_ERR_INVALID, "team_size must be 6");
    if (cfg->num_worlds == 0 || cfg->num_worlds > (1u << 22)) return fail(nullptr, MPENV_ERR_INVALID, "num_worlds out of range");
    const uint32_t unsupported = MPENV_FLAG_SPAWN_IN_MIDDLE | MPENV_FLAG_NAVMESH_SPAWN |
        MPENV_FLAG_HARDCODED_SPAWNS | MPENV_FLAG_FULL_TEAM_POLICY;
    if (cfg->sim_flags & unsupported) return fail(nullptr, MPENV_ERR_INVALID, "unsupported sim_flags bit set");

    int dev_count = 0;
    cudaError_t e = cudaGetDeviceCount(&dev_count);
    if (e != cudaSuccess || dev_count == 0) {
        return fail(nullptr, MPENV_ERR_CUDA, std::string("no usable CUDA device: ") + cudaGetErrorString(e) +
                    " (libmpenv_b200 has no CPU execution path)");
    }
    if (cfg->gpu_id < 0 || cfg->gpu_id >= dev_count) return fail(nullptr, MPENV_ERR_INVALID, "gpu_id out of range");

    mpenv_handle *h = new mpenv_handle();
    h->cfg = *cfg;
    h->sceneDir = cfg->scene_dir;
    h->cfg.scene_dir = h->sceneDir.c_str();
    h->device = cfg->gpu_id;
    h->fullTeam = cfg->enable_full_team != 0;
    h->eventLog = cfg->event_log_dir != nullptr;
    h->recordLog = cfg->record_log_path != nullptr;
    h->replayLog = cfg->replay_log_path != nullptr;
    h->cold = h->eventLog || h->recordLog || h->replayLog || cfg->curriculum_data_path != nullptr ||
              cfg->train_flank != 0 || cfg->task_type == MPENV_TASK_ZONE_CAPTURE_DEFEND ||
              (cfg->sim_flags & MPENV_FLAG_SUBZONES) != 0;
    if (h->eventLog) h->eventDir = cfg->event_log_dir;
    if (h->recordLog) h->recordPath = cfg->record_log_path;
    if (h->replayLog) h->replayPath = cfg->replay_log_path;
    if (cfg->curriculum_data_path) h->curriculumPath = cfg->curriculum_data_path;
    h->cfg.curriculum_data_path = cfg->curriculum_data_path ? h->curriculumPath.c_str() : nullptr;
    h->cfg.event_log_dir = h->eventLog ? h->eventDir.c_str() : nullptr;
    h->cfg.record_log_path = h->recordLog ? h->recordPath.c_str() : nullptr;
    h->cfg.replay_log_path = h->replayLog ? h->replayPath.c_str() : nullptr;
    auto bail = [&](int code) { std::string m = h->error; mpenv_destroy(h); g_create_error = m; return code; };

    std::string err;
    if (!loadMapAssets(h->sceneDir, &h->assets, &err) || !fillMapConsts(h->assets, &h->hostConsts, &err)) {
        h->error = err;
        return bail(MPENV_ERR_IO);
    }
    if ((cfg->sim_flags & MPENV_FLAG_SUBZONES) && h->hostConsts.numZones < 3) {
        // sub-zones 0 and 1 are zones 1 and 2 of the map (level_gen.cpp:283-293)
        h->error = "SimFlags::SubZones needs a map with at least 3 zones";
        g_create_error = h->error;
        return bail(MPENV_ERR_INVALID);
    }
    if (cfg->task_type == MPENV_TASK_ZONE_CAPTURE_DEFEND && h->hostConsts.numZones < 4) {
        // initWorld pins the objective to zone 3 (sim.cpp:822-825)
        h->error = "Task::ZoneCaptureDefend needs a map with at least 4 zones (zones.bin has " +
                   std::to_string(h->hostConsts.numZones) + ")";
        g_create_error = h->error;
        return bail(MPENV_ERR_INVALID);
    }

#define MP_CREATE(call) do { if (!cudaOk(h, (call), #call)) return bail(MPENV_ERR_CUDA); } while (0)
    MP_CREATE(cudaSetDevice(h->device));
    cudaDeviceProp prop;
    MP_CREATE(cudaGetDeviceProperties(&prop, h->device));
    h->numSMs = prop.multiProcessorCount;
    MP_CREATE(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));

    // Every device buffer sits between two 256-byte guard bands filled with 0xA5
    // (mpenv_check_guards): compute-sanitizer is not available on every pool, the bands catch an
    // out-of-bounds write of any kernel at the cost of 512 bytes per buffer.
    auto devAlloc = [&](void **ptr, size_t bytes) -> bool {
        char *raw = nullptr;
        const size_t padded = (bytes + 255) & ~(size_t)255;
        if (!cudaOk(h, cudaMalloc((void **)&raw, padded + 2 * kGuardBytes), "cudaMalloc")) return false;
        h->allocs.push_back(raw);
        h->guarded.push_back({raw, bytes});
        *ptr = raw + kGuardBytes;
        return cudaOk(h, cudaMemsetAsync(raw, 0xA5, padded + 2 * kGuardBytes, h->stream), "cudaMemsetAsync") &&
               cudaOk(h, cudaMemsetAsync(*ptr, 0, bytes, h->stream), "cudaMemsetAsync");
    };

    // static assets
    const size_t node_bytes = h->assets.nodes.size() * sizeof(BVHNode);
    const size_t vert_bytes = h->assets.verts.size() * sizeof(Vec3);
    if (!devAlloc((void **)&h->devNodes, node_bytes) || !devAlloc((void **)&h->devVerts, vert_bytes + 16) ||
        !devAlloc((void **)&h->devConsts, sizeof(MapConsts)))
        return bail(MPENV_ERR_CUDA);
    MP_CREATE(cudaMemcpyAsync(h->devNodes, h->assets.nodes.data(), node_bytes, cudaMemcpyHostToDevice, h->stream));
    MP_CREATE(cudaMemcpyAsync(h->devVerts, h->assets.verts.data(), vert_bytes, cudaMemcpyHostToDevice, h->stream));
    MP_CREATE(cudaMemcpyAsync(h->devConsts, &h->hostConsts, sizeof(MapConsts), cudaMemcpyHostToDevice, h->stream));

    // state + export buffers
    KernelArgs &ka = h->ka;
    memset(&ka, 0, sizeof(ka));
    int nbuf;
    const BufferDesc *table = bufferTable(&nbuf);
    for (int i = 0; i < nbuf; i++) {
        size_t bytes = bufferRows(table[i], cfg->num_worlds) * table[i].rowBytes;
        if (!devAlloc(bufferSlot(&ka.ctx.p, table[i]), bytes)) return bail(MPENV_ERR_CUDA);
    }
    if (!devAlloc((void **)&h->devStats, sizeof(unsigned long long) * MPENV_NUM_STATS)) return bail(MPENV_ERR_CUDA);

    ka.ctx.cfg.numWorlds = cfg->num_worlds;
    ka.ctx.cfg.worldIDOffset = cfg->world_id_offset;
    ka.ctx.cfg.simFlags = cfg->sim_flags;
    ka.ctx.cfg.autoReset = cfg->auto_reset;
    ka.ctx.cfg.initRandKey = randSplit(randInitKey(cfg->rand_seed), 0);     // mgr.cpp:1736-1738
    ka.ctx.cfg.eventLog = h->eventLog ? 1 : 0;
    ka.ctx.cfg.recordLog = h->recordLog ? 1 : 0;
    ka.ctx.cfg.replayLog = h->replayLog ? 1 : 0;
    ka.ctx.cfg.rewardFlank = cfg->train_flank ? 1 : 0;
    ka.ctx.cfg.taskCaptureDefend = cfg->task_type == MPENV_TASK_ZONE_CAPTURE_DEFEND ? 1 : 0;

    // navmesh + A* table for the scripted bot (LevelData, sim.cpp:5895-5898)
    if (h->assets.hasNav) {
        const NavmeshAssets &nav = h->assets.nav;
        Vec3 *dv = nullptr; uint32_t *dt = nullptr, *dl = nullptr;
        if (!devAlloc((void **)&dv, nav.verts.size() * sizeof(Vec3)) || !devAlloc((void **)&dt, nav.tris.size() * 4) ||
            !devAlloc((void **)&dl, nav.lookup.size() * 4))
            return bail(MPENV_ERR_CUDA);
        MP_CREATE(cudaMemcpyAsync(dv, nav.verts.data(), nav.verts.size() * sizeof(Vec3), cudaMemcpyHostToDevice, h->stream));
        MP_CREATE(cudaMemcpyAsync(dt, nav.tris.data(), nav.tris.size() * 4, cudaMemcpyHostToDevice, h->stream));
        MP_CREATE(cudaMemcpyAsync(dl, nav.lookup.data(), nav.lookup.size() * 4, cudaMemcpyHostToDevice, h->stream));
        ka.ctx.nav.verts = dv; ka.ctx.nav.tris = dt; ka.ctx.nav.lookup = dl; ka.ctx.nav.numTris = nav.numTris();
    }

    // trajectory curriculum (mgr.cpp:1424-1441, 1643-1650)
    if (cfg->curriculum_data_path) {
        FILE *f = fopen(h->curriculumPath.c_str(), "rb");
        if (!f) { h->error = "failed to open curriculum data " + h->curriculumPath; return bail(MPENV_ERR_IO); }
        std::vector<char> data;
        char chunk[65536];
        size_t got;
        while ((got = fread(chunk, 1, sizeof(chunk), f)) > 0) data.insert(data.end(), chunk, chunk + got);
        fclose(f);
        const size_t n = data.size() / sizeof(CurriculumSnapshot);
        if (n > 0) {
            CurriculumSnapshot *dev = nullptr;
            if (!devAlloc((void **)&dev, n * sizeof(CurriculumSnapshot))) return bail(MPENV_ERR_CUDA);
            MP_CREATE(cudaMemcpyAsync(dev, data.data(), n * sizeof(CurriculumSnapshot), cudaMemcpyHostToDevice, h->stream));
            MP_CREATE(cudaStreamSynchronize(h->stream));     // `data` is pageable and about to go away
            ka.ctx.p.curriculum = dev;
        }
        ka.ctx.cfg.numCurriculumSnapshots = (int32_t)n;
    }

    // logs: files as the reference opens them (mgr.cpp:200-231; FATAL there, an error here)
    {
        auto openFile = [&](const std::string &path, const char *mode, FILE **out) -> bool {
            *out = fopen(path.c_str(), mode);
            if (!*out) h->error = "failed to open log file " + path;
            return *out != nullptr;
        };
        const size_t W = cfg->num_worlds;
        if (h->eventLog) {
            if (!h->eventDir.empty() && (!openFile(h->eventDir + "/events.bin", "wb", &h->evFile) ||
                                         !openFile(h->eventDir + "/steps.bin", "wb", &h->stepsFile)))
                return bail(MPENV_ERR_IO);
            if (!devAlloc((void **)&h->devEvOffset, W * 4) || !devAlloc((void **)&h->devSnapOffset, W * 4) ||
                !devAlloc((void **)&h->devEvTotals, 8) ||
                !devAlloc((void **)&h->devEvOut, W * kEventSlotsPerWorld * sizeof(GameEvent)) ||
                !devAlloc((void **)&h->devSnapOut, W * sizeof(PackedStepSnapshot)))
                return bail(MPENV_ERR_CUDA);
            MP_CREATE(cudaMallocHost((void **)&h->hostEvTotals, 8));
            MP_CREATE(cudaMallocHost((void **)&h->hostEvents, W * kEventSlotsPerWorld * sizeof(GameEvent)));
            MP_CREATE(cudaMallocHost((void **)&h->hostSnaps, W * sizeof(PackedStepSnapshot)));
        }
        if (h->recordLog) {
            if (!h->recordPath.empty() && !openFile(h->recordPath, "wb", &h->recFile)) return bail(MPENV_ERR_IO);
            MP_CREATE(cudaMallocHost((void **)&h->hostRecord, W * sizeof(StepLog)));
        }
        if (h->replayLog) {
            if (!h->replayPath.empty() && !openFile(h->replayPath, "rb", &h->replayFile)) return bail(MPENV_ERR_IO);
            MP_CREATE(cudaMallocHost((void **)&h->hostReplay, W * sizeof(StepLog)));
        }
    }
    ka.ctx.mc = h->devConsts;
    ka.ctx.bvh.nodes = h->devNodes;
    ka.ctx.bvh.verts = h->devVerts;
    ka.ctx.bvh.rootAABB = h->assets.rootAABB;
    ka.ctx.bvh.numNodes = (uint32_t)h->assets.nodes.size();
    ka.ctx.bvh.numTris = h->assets.numTris;
    ka.one = 1.0f;
    ka.bvhNodeWords = (uint32_t)(node_bytes / 4);
    ka.bvhVertWords = (uint32_t)(vert_bytes / 4);
    h->bvhSmemBytes = node_bytes + ((vert_bytes + 15) / 16) * 16;
    ka.bvhInSmem = h->bvhSmemBytes <= 160 * 1024 ? 1 : 0;
    if (!devAlloc((void **)&ka.workCounters, 64)) return bail(MPENV_ERR_CUDA);
    if (!ka.bvhInSmem) h->bvhSmemBytes = 0;

    // TrainControl from the flags (mgr.cpp:1397-1413)
    TrainControl tc;
    tc.evalMode = (cfg->sim_flags & MPENV_FLAG_SIM_EVAL_MODE) ? 1 : 0;
    tc.randomizeEpisodeLengthAfterReset = (cfg->sim_flags & MPENV_FLAG_STAGGER_STARTS) ? 1 : 0;
    tc.randomizeTeamSides = (cfg->sim_flags & MPENV_FLAG_RANDOM_FLIP_TEAMS) ? 1 : 0;
    MP_CREATE(cudaMemcpyAsync(ka.ctx.p.trainCtrl, &tc, sizeof(tc), cudaMemcpyHostToDevice, h->stream));

    if (h->bvhSmemBytes > 48 * 1024) {
        MP_CREATE(cudaFuncSetAttribute(k_move<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->bvhSmemBytes));
        MP_CREATE(cudaFuncSetAttribute(k_fire<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->bvhSmemBytes));
        MP_CREATE(cudaFuncSetAttribute(k_fire<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->bvhSmemBytes));
        MP_CREATE(cudaFuncSetAttribute(k_los, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->bvhSmemBytes));
        MP_CREATE(cudaFuncSetAttribute(k_lidar, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->bvhSmemBytes));
    }

    const long long W = cfg->num_worlds, N = W * kAgents;
    h->gridMove = gridFor(h, k_move<false>, 128, h->bvhSmemBytes, N);
    h->gridFire = h->cold ? gridFor(h, k_fire<true>, 128, h->bvhSmemBytes, N) : gridFor(h, k_fire<false>, 128, h->bvhSmemBytes, N);
    h->gridWorld = h->cold ? gridFor(h, k_world<true>, 128, 0, W * 16) : gridFor(h, k_world<false>, 128, 0, W * 16);
    h->gridLos = gridFor(h, k_los, 128, h->bvhSmemBytes, N * kTeamSize);
    h->gridObs = gridFor(h, k_obs, 128, 0, N * kObsRowsPerAgent);
    h->gridTeam = gridFor(h, k_fullteam, 128, 0, W * 2);
    h->gridLidar = gridFor(h, k_lidar, 160, h->bvhSmemBytes, N * 80);
    h->gridSmall = gridFor(h, k_reset, 128, 0, W);

    // Warp-pooled ray kernels: need the re-encoded BVH + per-warp pools in shared memory.  A map
    // whose BVH does not fit runs the per-thread k_los / k_lidar of step_kernels.cuh on the
    // reference-format BVH (shared memory if that fits, else L2); MPENV_KERNELS=legacy forces
    // that path so that the tests can cover it on simple_map.
    const char *kmode = getenv("MPENV_KERNELS");
    const bool legacy = kmode && strcmp(kmode, "legacy") == 0;
    // sphere casts on radius-expanded float bounds (sphere_fast.cuh) while five CTAs' copies fit in an SM
    h->moveSmem = sphereBVHBytes(ka.ctx.bvh.numNodes, ka.bvhVertWords);
    h->fastMove = !legacy && h->moveSmem <= 44 * 1024;
    if (h->fastMove) {
        MP_CREATE(cudaFuncSetAttribute(k_move<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->moveSmem));
        h->gridMove = gridFor(h, k_move<true>, 128, h->moveSmem, N);
    }
    // hitscan on the re-encoded BVH while five CTAs' copies fit in an SM
    h->fireSmem = fastBVHBytes(ka.ctx.bvh.numNodes, ka.ctx.bvh.numTris);
    h->fastFire = !legacy && h->fireSmem <= 44 * 1024;
    if (h->fastFire) {
        MP_CREATE(cudaFuncSetAttribute(k_fire_fast<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fireSmem));
        MP_CREATE(cudaFuncSetAttribute(k_fire_fast<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fireSmem));
        h->gridFire = h->cold ? gridFor(h, k_fire_fast<true>, 128, h->fireSmem, N) : gridFor(h, k_fire_fast<false>, 128, h->fireSmem, N);
    }
    // one CTA per SM: as many warps as have room for their pool beside the BVH (24 on simple_map)
    constexpr size_t kSmemPerCTA = 226 * 1024, kSmemHalfSM = 113 * 1024;
    h->losWarps = losPoolWarps(ka.ctx.bvh.numNodes, ka.ctx.bvh.numTris, kSmemPerCTA);
    h->poolLos = !legacy && h->losWarps >= 8;
    if (h->poolLos) {
        h->losPoolSmem = losPoolSmemBytes(ka.ctx.bvh.numNodes, ka.ctx.bvh.numTris, h->losWarps);
        MP_CREATE(cudaFuncSetAttribute(k_los_pool<10, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->losPoolSmem));
        h->gridLosPool = gridFor(h, k_los_pool<10, 8>, h->losWarps * 32, h->losPoolSmem, W * 2 * 32);
    }
    {
        // two 16-warp CTAs per SM while two copies of the BVH fit, else one 32-warp CTA; the shared
        // stack holds <= 3 pending children per tree level
        const size_t two = lidarPacketSmemBytes(ka.ctx.bvh.numNodes, ka.ctx.bvh.numTris, kPacketWarpsPerCTA);
        const size_t one = lidarPacketSmemBytes(ka.ctx.bvh.numNodes, ka.ctx.bvh.numTris, 32);
        if (!legacy && 3 * bvhDepth(h->assets.nodes) + 1 <= kPacketStack)
            h->packetLidar = two <= kSmemHalfSM ? 2 : one <= kSmemPerCTA ? 1 : 0;
        if (h->packetLidar == 2) {
            h->lidarPacketSmem = two;
            auto k = k_lidar_packet<kPacketWarpsPerCTA, MPENV_PACKET_MINB>;
            MP_CREATE(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)two));
            h->gridLidarPacket = gridFor(h, k, kPacketWarpsPerCTA * 32, two, (N / 2) * 32);
        } else if (h->packetLidar == 1) {
            h->lidarPacketSmem = one;
            auto k = k_lidar_packet<32, 1>;
            MP_CREATE(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)one));
            h->gridLidarPacket = gridFor(h, k, 1024, one, (N / 2) * 32);
        }
    }
    // The re-encoded BVH images are built once here; the kernels stage them with plain 16-byte copies.
    ka.fastImage = nullptr;
    ka.sphereImage = nullptr;
    if (h->fastFire || h->poolLos || h->packetLidar) {
        void *img = nullptr;
        if (!devAlloc(&img, fastBVHBytes(ka.ctx.bvh.numNodes, ka.ctx.bvh.numTris))) return bail(MPENV_ERR_CUDA);
        k_encode_fast_bvh<<<1, 256, 0, h->stream>>>(ka.ctx.bvh, img);
        MP_CREATE(cudaGetLastError());
        ka.fastImage = reinterpret_cast<const uint4 *>(img);
    }
    if (h->fastMove) {
        void *img = nullptr;
        if (!devAlloc(&img, h->moveSmem)) return bail(MPENV_ERR_CUDA);
        k_encode_sphere_bvh<<<1, 256, 0, h->stream>>>(ka.ctx.bvh, ka.bvhVertWords, kAgentRadius, img);
        MP_CREATE(cudaGetLastError());
        ka.sphereImage = reinterpret_cast<const uint4 *>(img);
    }
    if (!h->fullTeam) {
        MP_CREATE(cudaStreamCreateWithFlags(&h->lidarStream, cudaStreamNonBlocking));
        MP_CREATE(cudaEventCreateWithFlags(&h->evFork, cudaEventDisableTiming));
        MP_CREATE(cudaEventCreateWithFlags(&h->evJoin, cudaEventDisableTiming));
    }
    // Sim::Sim for every world (sim.cpp:5850-5980)
    k_ctor<<<h->gridSmall, 128, 0, h->stream>>>(ka);
    MP_CREATE(cudaGetLastError());
    MP_CREATE(cudaStreamSynchronize(h->stream));

    if (captureGraph(h, true, &h->stepGraph, &h->stepLaunches) != MPENV_OK ||
        captureGraph(h, false, &h->initGraph, &h->initLaunches) != MPENV_OK)
        return bail(MPENV_ERR_CUDA);
#undef MP_CREATE
    *out = h;
    return MPENV_OK;
}

int mpenv_init_async(mpenv_handle *h, void *cuda_stream)
{
    if (!h) return MPENV_ERR_INVALID;
    NvtxRange nvtx("mpenv_init (Init graph: reset -> visibility -> observations -> lidar)");
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaGraphLaunch(h->initGraph, (cudaStream_t)cuda_stream));
    return MPENV_OK;
}

int mpenv_step_async(mpenv_handle *h, void *cuda_stream)
{
    if (!h) return MPENV_ERR_INVALID;
    NvtxRange nvtx("mpenv_step (Step graph)");
    MP_CUDA(h, cudaSetDevice(h->device));
    int rc = logsBeforeStep(h, (cudaStream_t)cuda_stream);
    if (rc != MPENV_OK) return rc;
    MP_CUDA(h, cudaGraphLaunch(h->stepGraph, (cudaStream_t)cuda_stream));
    return logsAfterStep(h, (cudaStream_t)cuda_stream);     // no-op (and no host sync) without logs
}

int mpenv_event_log(mpenv_handle *h, const void **events, uint32_t *num_events, const void **step_states,
                    uint32_t *num_step_states)
{
    if (!h) return MPENV_ERR_INVALID;
    if (!h->eventLog) return fail(h, MPENV_ERR_INVALID, "event log not enabled (mpenv_config.event_log_dir)");
    if (events) *events = h->hostEvents;
    if (num_events) *num_events = h->lastNumEvents;
    if (step_states) *step_states = h->hostSnaps;
    if (num_step_states) *num_step_states = h->lastNumSnaps;
    return MPENV_OK;
}

int mpenv_record_log(mpenv_handle *h, const void **step_logs, uint32_t *num_worlds)
{
    if (!h) return MPENV_ERR_INVALID;
    if (!h->recordLog) return fail(h, MPENV_ERR_INVALID, "record log not enabled (mpenv_config.record_log_path)");
    if (step_logs) *step_logs = h->hostRecord;
    if (num_worlds) *num_worlds = h->cfg.num_worlds;
    return MPENV_OK;
}

int mpenv_set_replay_log(mpenv_handle *h, const void *step_logs, uint64_t bytes)
{
    if (!h || !step_logs) return MPENV_ERR_INVALID;
    if (!h->replayLog || h->replayFile)
        return fail(h, MPENV_ERR_INVALID, "replay from memory not enabled (mpenv_config.replay_log_path == \"\")");
    if (bytes != (uint64_t)h->cfg.num_worlds * sizeof(StepLog))
        return fail(h, MPENV_ERR_INVALID, "mpenv_set_replay_log: size mismatch");
    memcpy(h->hostReplay, step_logs, bytes);                // every replay step ends with a stream sync
    h->replayPending = true;
    return MPENV_OK;
}

void *mpenv_stream(mpenv_handle *h) { return h ? (void *)h->stream : nullptr; }

int mpenv_sync(mpenv_handle *h)
{
    if (!h) return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaStreamSynchronize(h->stream));
    return MPENV_OK;
}

int mpenv_init(mpenv_handle *h)
{
    int rc = mpenv_init_async(h, h ? h->stream : nullptr);
    return rc != MPENV_OK ? rc : mpenv_sync(h);
}

int mpenv_step(mpenv_handle *h)
{
    int rc = mpenv_step_async(h, h ? h->stream : nullptr);
    return rc != MPENV_OK ? rc : mpenv_sync(h);
}

int mpenv_stream_init(mpenv_handle *h, void *cuda_stream, void **buffers)
{
    if (!h || !buffers) return MPENV_ERR_INVALID;
    int rc = mpenv_init_async(h, cuda_stream);
    if (rc != MPENV_OK) return rc;
    return copyInterface(h, (cudaStream_t)cuda_stream, buffers, false);
}

int mpenv_stream_step(mpenv_handle *h, void *cuda_stream, void **buffers)
{
    if (!h || !buffers) return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    int rc = copyInterface(h, (cudaStream_t)cuda_stream, buffers, true);
    if (rc != MPENV_OK) return rc;
    rc = mpenv_step_async(h, cuda_stream);
    if (rc != MPENV_OK) return rc;
    return copyInterface(h, (cudaStream_t)cuda_stream, buffers, false);
}

int mpenv_check_guards(mpenv_handle *h, uint32_t *num_damaged)
{
    if (!h || !num_damaged) return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaStreamSynchronize(h->stream));
    std::vector<unsigned char> host(2 * kGuardBytes);
    uint32_t bad = 0;
    for (const mpenv_handle::Guarded &g : h->guarded) {
        const size_t padded = (g.bytes + 255) & ~(size_t)255;
        MP_CUDA(h, cudaMemcpy(host.data(), g.raw, kGuardBytes, cudaMemcpyDeviceToHost));
        // the tail band starts right after the buffer's last byte (the padding belongs to it)
        const size_t tail = kGuardBytes + padded - g.bytes;
        std::vector<unsigned char> t(tail);
        MP_CUDA(h, cudaMemcpy(t.data(), g.raw + kGuardBytes + g.bytes, tail, cudaMemcpyDeviceToHost));
        bool ok = true;
        for (size_t i = 0; i < kGuardBytes; i++) ok = ok && host[i] == 0xA5;
        for (size_t i = 0; i < tail; i++) ok = ok && t[i] == 0xA5;
        if (!ok) bad++;
    }
    *num_damaged = bad;
    return MPENV_OK;
}

// XLA GPU custom-call entry points (the ABI XLA binds: stream, buffers, opaque bytes).
static mpenv_handle *xlaHandle(const char *opaque, size_t opaque_len)
{
    mpenv_handle *h = nullptr;
    if (opaque && opaque_len == sizeof(h)) memcpy(&h, opaque, sizeof(h));
    return h;
}

void mpenv_xla_init(void *cuda_stream, void **buffers, const char *opaque, size_t opaque_len)
{
    mpenv_handle *h = xlaHandle(opaque, opaque_len);
    if (!h) { g_create_error = "mpenv_xla_init: opaque must hold the 8-byte mpenv_handle pointer"; return; }
    if (mpenv_stream_init(h, cuda_stream, buffers) != MPENV_OK)
        fprintf(stderr, "mpenv_xla_init: %s\n", h->error.c_str());
}

void mpenv_xla_step(void *cuda_stream, void **buffers, const char *opaque, size_t opaque_len)
{
    mpenv_handle *h = xlaHandle(opaque, opaque_len);
    if (!h) { g_create_error = "mpenv_xla_step: opaque must hold the 8-byte mpenv_handle pointer"; return; }
    if (mpenv_stream_step(h, cuda_stream, buffers) != MPENV_OK)
        fprintf(stderr, "mpenv_xla_step: %s\n", h->error.c_str());
}

int mpenv_tensor(mpenv_handle *h, uint32_t export_id, mpenv_tensor_desc *out)
{
    if (!h || !out) return MPENV_ERR_INVALID;
    ExportDesc d;
    if (export_id == MPENV_EXPORT_AGENT_MAP || export_id == MPENV_EXPORT_UNMASKED_AGENT_MAP) {
        if (!h->agentMap) {
            MP_CUDA(h, cudaSetDevice(h->device));
            size_t bytes = (size_t)h->cfg.num_worlds * kAgents * 16 * 16 * 4 * sizeof(float);
            MP_CUDA(h, cudaMalloc((void **)&h->agentMap, bytes));
            h->allocs.push_back(h->agentMap);
            MP_CUDA(h, cudaMemset(h->agentMap, 0, bytes));
        }
        out->dev_ptr = h->agentMap; out->dtype = MPENV_DTYPE_F32; out->ndim = 4;
        out->dims[0] = (int64_t)h->cfg.num_worlds * kAgents; out->dims[1] = 16; out->dims[2] = 16; out->dims[3] = 4;
        out->gpu_id = h->device;
        return MPENV_OK;
    }
    if (!exportDesc(h->ka.ctx.p, h->cfg.num_worlds, export_id, &d)) {
        return fail(h, MPENV_ERR_INVALID, "tensor not exported by this build (export id " + std::to_string(export_id) + ")");
    }
    out->dev_ptr = d.ptr; out->dtype = d.dtype; out->ndim = d.ndim;
    for (int i = 0; i < 4; i++) out->dims[i] = d.dims[i];
    if (export_id == MPENV_EXPORT_SIM_CONTROL) { out->ndim = 1; out->dims[0] = 3; out->dims[1] = 0; }
    out->gpu_id = h->device;
    return MPENV_OK;
}

int mpenv_write_tensor_async(mpenv_handle *h, uint32_t export_id, const void *host_src, uint64_t bytes)
{
    if (!h || !host_src) return MPENV_ERR_INVALID;
    mpenv_tensor_desc d;
    int rc = mpenv_tensor(h, export_id, &d);
    if (rc != MPENV_OK) return rc;
    if (bytes != tensorBytes(d)) return fail(h, MPENV_ERR_INVALID, "mpenv_write_tensor_async: size mismatch");
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaMemcpyAsync(d.dev_ptr, host_src, bytes, cudaMemcpyHostToDevice, h->stream));
    return MPENV_OK;
}

int mpenv_read_tensor_async(mpenv_handle *h, uint32_t export_id, void *host_dst, uint64_t bytes)
{
    if (!h || !host_dst) return MPENV_ERR_INVALID;
    mpenv_tensor_desc d;
    int rc = mpenv_tensor(h, export_id, &d);
    if (rc != MPENV_OK) return rc;
    if (bytes != tensorBytes(d)) return fail(h, MPENV_ERR_INVALID, "mpenv_read_tensor_async: size mismatch");
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaMemcpyAsync(host_dst, d.dev_ptr, bytes, cudaMemcpyDeviceToHost, h->stream));
    return MPENV_OK;
}

int mpenv_snapshot_tensors_async(mpenv_handle *h, const uint32_t *export_ids, uint32_t n)
{
    if (!h || !export_ids || n == 0 || n > 64) return MPENV_ERR_INVALID;
    NvtxRange nvtx("mpenv_snapshot_tensors_async");
    MP_CUDA(h, cudaSetDevice(h->device));
    mpenv_tensor_desc d[64];
    size_t total = 0;
    for (uint32_t i = 0; i < n; i++) {
        int rc = mpenv_tensor(h, export_ids[i], &d[i]);
        if (rc != MPENV_OK) return rc;
        total += (tensorBytes(d[i]) + 255) & ~(size_t)255;
    }
    if (!h->copyStream) {
        MP_CUDA(h, cudaStreamCreateWithFlags(&h->copyStream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; k++) {
            MP_CUDA(h, cudaEventCreateWithFlags(&h->evSnap[k], cudaEventDisableTiming));
            MP_CUDA(h, cudaEventCreateWithFlags(&h->evRead[k], cudaEventDisableTiming));
        }
        MP_CUDA(h, cudaEventCreateWithFlags(&h->evCopied, cudaEventDisableTiming));
    }
    if (total > h->snapshotBytes) {
        MP_CUDA(h, cudaStreamSynchronize(h->copyStream));
        for (int k = 0; k < 2; k++) {
            if (h->snapshot[k]) cudaFree(h->snapshot[k]);
            h->snapshot[k] = nullptr;
            h->snapRead[k] = false;
        }
        h->snapshotBytes = 0;
        for (int k = 0; k < 2; k++) MP_CUDA(h, cudaMalloc((void **)&h->snapshot[k], total));
        h->snapshotBytes = total;
    }
    const int b = h->snapNext;
    // this buffer may not be overwritten before the transfer that last read it has left it
    if (h->snapRead[b]) MP_CUDA(h, cudaStreamWaitEvent(h->stream, h->evRead[b], 0));
    size_t off = 0;
    for (uint32_t i = 0; i < n; i++) {
        MP_CUDA(h, cudaMemcpyAsync(h->snapshot[b] + off, d[i].dev_ptr, tensorBytes(d[i]), cudaMemcpyDeviceToDevice, h->stream));
        off += (tensorBytes(d[i]) + 255) & ~(size_t)255;
        h->snapIds[i] = export_ids[i];
    }
    h->snapCount = n;
    MP_CUDA(h, cudaEventRecord(h->evSnap[b], h->stream));
    h->snapReady = b;
    h->snapNext = b ^ 1;
    return MPENV_OK;
}

int mpenv_read_snapshot_async(mpenv_handle *h, void *const *host_dst, uint32_t n)
{
    if (!h || !host_dst || n == 0 || n > 64) return MPENV_ERR_INVALID;
    if (h->snapReady < 0 || n != h->snapCount)
        return fail(h, MPENV_ERR_INVALID, "mpenv_read_snapshot_async: no snapshot of that many tensors is pending");
    NvtxRange nvtx("mpenv_read_snapshot_async (D2H)");
    MP_CUDA(h, cudaSetDevice(h->device));
    const int b = h->snapReady;
    MP_CUDA(h, cudaStreamWaitEvent(h->copyStream, h->evSnap[b], 0));
    size_t off = 0;
    for (uint32_t i = 0; i < n; i++) {
        if (!host_dst[i]) return fail(h, MPENV_ERR_INVALID, "mpenv_read_snapshot_async: null host buffer");
        mpenv_tensor_desc d;
        int rc = mpenv_tensor(h, h->snapIds[i], &d);
        if (rc != MPENV_OK) return rc;
        MP_CUDA(h, cudaMemcpyAsync(host_dst[i], h->snapshot[b] + off, tensorBytes(d), cudaMemcpyDeviceToHost, h->copyStream));
        off += (tensorBytes(d) + 255) & ~(size_t)255;
    }
    MP_CUDA(h, cudaEventRecord(h->evRead[b], h->copyStream));
    MP_CUDA(h, cudaEventRecord(h->evCopied, h->copyStream));
    h->snapRead[b] = true;
    h->snapReady = -1;
    h->copyInFlight = true;
    return MPENV_OK;
}

int mpenv_read_tensors_pipelined(mpenv_handle *h, const uint32_t *export_ids, void *const *host_dst, uint32_t n)
{
    if (!h || !export_ids || !host_dst || n == 0 || n > 64) return MPENV_ERR_INVALID;
    NvtxRange nvtx("mpenv_read_tensors_pipelined (snapshot + D2H)");
    int rc = mpenv_snapshot_tensors_async(h, export_ids, n);
    return rc != MPENV_OK ? rc : mpenv_read_snapshot_async(h, host_dst, n);
}

int mpenv_read_wait(mpenv_handle *h)
{
    if (!h) return MPENV_ERR_INVALID;
    if (!h->copyInFlight) return MPENV_OK;
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaEventSynchronize(h->evCopied));
    return MPENV_OK;
}

int mpenv_host_alloc(uint64_t bytes, void **out)
{
    if (!out) return MPENV_ERR_INVALID;
    if (!cudaOk(nullptr, cudaMallocHost(out, bytes), "cudaMallocHost")) return MPENV_ERR_CUDA;
    return MPENV_OK;
}

void mpenv_host_free(void *p) { if (p) cudaFreeHost(p); }

int mpenv_train_interface(mpenv_handle *h, const mpenv_named_tensor **out, uint32_t *count)
{
    if (!h || !out || !count) return MPENV_ERR_INVALID;
    *out = kTrainInterface;
    *count = kNumTrainTensors;
    return MPENV_OK;
}

int mpenv_trigger_reset(mpenv_handle *h, int32_t world_idx)
{
    if (!h || world_idx < 0 || (uint32_t)world_idx >= h->cfg.num_worlds) return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    int32_t one = 1;
    MP_CUDA(h, cudaMemcpy(h->ka.ctx.p.resetFlag + world_idx, &one, 4, cudaMemcpyHostToDevice));
    return MPENV_OK;
}

int mpenv_set_pvp_action(mpenv_handle *h, int32_t world_idx, int32_t agent_idx, const int32_t discrete[4],
                         const float aim[2], const int32_t discrete_aim[2])
{
    if (!h || world_idx < 0 || (uint32_t)world_idx >= h->cfg.num_worlds || agent_idx < 0 || agent_idx >= kAgents)
        return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    const size_t idx = (size_t)world_idx * kAgents + agent_idx;
    MP_CUDA(h, cudaMemcpy(h->ka.ctx.p.actDiscrete + idx, discrete, 16, cudaMemcpyHostToDevice));
    MP_CUDA(h, cudaMemcpy(h->ka.ctx.p.actAim + idx, aim, 8, cudaMemcpyHostToDevice));
    MP_CUDA(h, cudaMemcpy(h->ka.ctx.p.actDiscreteAim + idx, discrete_aim, 8, cudaMemcpyHostToDevice));
    return MPENV_OK;
}

int mpenv_set_hp(mpenv_handle *h, int32_t world_idx, int32_t agent_idx, int32_t hp)
{
    if (!h || world_idx < 0 || (uint32_t)world_idx >= h->cfg.num_worlds || agent_idx < 0 || agent_idx >= kAgents)
        return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    float v = (float)hp;
    MP_CUDA(h, cudaMemcpy(h->ka.ctx.p.hp + (size_t)world_idx * kAgents + agent_idx, &v, 4, cudaMemcpyHostToDevice));
    return MPENV_OK;
}

int mpenv_set_uniform_agent_policy(mpenv_handle *h, int32_t policy_idx)
{
    if (!h) return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    std::vector<int32_t> v((size_t)h->cfg.num_worlds * kAgents, policy_idx);
    MP_CUDA(h, cudaMemcpy(h->ka.ctx.p.policy, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    return MPENV_OK;
}

int mpenv_episode_stats_async(mpenv_handle *h, void *cuda_stream, int64_t **dev_out)
{
    if (!h || !dev_out) return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)cuda_stream;
    MP_CUDA(h, cudaMemsetAsync(h->devStats, 0, sizeof(unsigned long long) * MPENV_NUM_STATS, s));
    int grid = (int)((h->cfg.num_worlds + 255) / 256);
    if (grid > h->numSMs * 4) grid = h->numSMs * 4;
    k_stats<<<grid, 256, 0, s>>>(h->ka, h->devStats);
    MP_CUDA(h, cudaGetLastError());
    *dev_out = (int64_t *)h->devStats;
    return MPENV_OK;
}

int mpenv_episode_stats(mpenv_handle *h, int64_t *host_out)
{
    if (!h || !host_out) return MPENV_ERR_INVALID;
    int64_t *dev = nullptr;
    int rc = mpenv_episode_stats_async(h, h->stream, &dev);
    if (rc != MPENV_OK) return rc;
    MP_CUDA(h, cudaMemcpyAsync(host_out, dev, sizeof(int64_t) * MPENV_NUM_STATS, cudaMemcpyDeviceToHost, h->stream));
    MP_CUDA(h, cudaStreamSynchronize(h->stream));
    return MPENV_OK;
}

int mpenv_episode_stats_static(mpenv_handle *h, void *cuda_stream, int64_t *dev_out)
{
    if (!h || !dev_out) return MPENV_ERR_INVALID;
    if (h->cfg.num_worlds > MPENV_STATIC_STATS_MAX_WORLDS) {
        char buf[160];
        snprintf(buf, sizeof(buf), "mpenv_episode_stats_static: %u worlds exceed the static scratch's maximum of %u "
                 "(use mpenv_episode_stats_async)", (unsigned)h->cfg.num_worlds, (unsigned)MPENV_STATIC_STATS_MAX_WORLDS);
        return fail(h, MPENV_ERR_INVALID, buf);
    }
    MP_CUDA(h, cudaSetDevice(h->device));
    unsigned long long *scratch = nullptr;
    const int rc = staticStatsScratch(h, &scratch);
    if (rc != MPENV_OK) return rc;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    int grid = (int)((h->cfg.num_worlds + 255) / 256);
    if (grid > kStaticStatsMaxCTAs) grid = kStaticStatsMaxCTAs;
    if (grid < 1) grid = 1;
    // the ticket must start at zero whatever an earlier (possibly aborted) call left; the partial rows
    // are overwritten in full by the CTAs that the last one reads
    MP_CUDA(h, cudaMemsetAsync(scratch + (size_t)kStaticStatsMaxCTAs * kStaticStatsRow, 0, sizeof(unsigned long long), s));
    k_stats_static<<<grid, 256, 0, s>>>(h->ka, scratch, reinterpret_cast<unsigned long long *>(dev_out));
    MP_CUDA(h, cudaGetLastError());
    return MPENV_OK;
}

int mpenv_get_world_state(mpenv_handle *h, uint32_t world_idx, mpenv_world_state *out)
{
    if (!h || !out || world_idx >= h->cfg.num_worlds) return MPENV_ERR_INVALID;
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaStreamSynchronize(h->stream));
    int nbuf;
    const BufferDesc *table = bufferTable(&nbuf);
    SimPtrs one;
    memset(&one, 0, sizeof(one));
    std::vector<std::vector<char>> staging(nbuf);
    for (int i = 0; i < nbuf; i++) {
        if (table[i].kind == kGlobal) continue;
        if (strcmp(table[i].name, "exploreGrid") == 0) continue;     // 26 KB / agent, not inspected
        size_t rows = table[i].kind == kPerAgent ? kAgents : (table[i].kind == kPerTeam ? 2 : 1);
        size_t bytes = rows * table[i].rowBytes;
        staging[i].resize(bytes);
        const char *src = (const char *)*bufferSlot(&h->ka.ctx.p, table[i]) + (size_t)world_idx * bytes;
        MP_CUDA(h, cudaMemcpy(staging[i].data(), src, bytes, cudaMemcpyDeviceToHost));
        *bufferSlot(&one, table[i]) = staging[i].data();
    }
    fillWorldState(one, out);
    return MPENV_OK;
}

int mpenv_step_timed(mpenv_handle *h, float *kernel_ms, const char **kernel_names, uint32_t cap, uint32_t *count)
{
    if (!h || !kernel_ms || !count) return MPENV_ERR_INVALID;
    if (h->cold)
        return fail(h, MPENV_ERR_INVALID, "mpenv_step_timed: not available with logs / replay / curriculum enabled");
    MP_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t s = h->stream;
    struct Item { const char *name; };
    const char *names[8];
    cudaEvent_t ev[9];
    for (int i = 0; i < 9; i++) MP_CUDA(h, cudaEventCreate(&ev[i]));
    uint32_t n = 0;
    MP_CUDA(h, cudaEventRecord(ev[0], s));
    launchMove(h, s); names[n++] = "k_move"; cudaEventRecord(ev[n], s);
    launchFire(h, s); names[n++] = "k_fire"; cudaEventRecord(ev[n], s);
    k_world<false><<<h->gridWorld, 128, 0, s>>>(h->ka); names[n++] = "k_world"; cudaEventRecord(ev[n], s);
    launchLos(h, s); names[n++] = "k_los"; cudaEventRecord(ev[n], s);
    k_obs<<<h->gridObs, 128, 0, s>>>(h->ka); names[n++] = "k_obs"; cudaEventRecord(ev[n], s);
    if (h->fullTeam) {
        k_fullteam<<<h->gridTeam, 128, 0, s>>>(h->ka); names[n++] = "k_fullteam"; cudaEventRecord(ev[n], s);
    }
    launchLidar(h, s); names[n++] = "k_lidar"; cudaEventRecord(ev[n], s);
    MP_CUDA(h, cudaGetLastError());
    MP_CUDA(h, cudaStreamSynchronize(s));
    for (uint32_t i = 0; i < n && i < cap; i++) {
        MP_CUDA(h, cudaEventElapsedTime(&kernel_ms[i], ev[i], ev[i + 1]));
        if (kernel_names) kernel_names[i] = names[i];
    }
    for (int i = 0; i < 9; i++) cudaEventDestroy(ev[i]);
    *count = n < cap ? n : cap;
    return MPENV_OK;
}

int mpenv_ray_counts(mpenv_handle *h, uint64_t *out, int32_t reset)
{
    if (!h || !out) return MPENV_ERR_INVALID;
#if defined(MPENV_COUNT_RAYS)
    MP_CUDA(h, cudaSetDevice(h->device));
    MP_CUDA(h, cudaStreamSynchronize(h->stream));
    unsigned long long host[4];
    MP_CUDA(h, cudaMemcpyFromSymbol(host, g_rayCounts, sizeof(host)));
    for (int i = 0; i < 4; i++) out[i] = host[i];
    if (reset) {
        unsigned long long zero[4] = { 0, 0, 0, 0 };
        MP_CUDA(h, cudaMemcpyToSymbol(g_rayCounts, zero, sizeof(zero)));
    }
    return MPENV_OK;
#else
    (void)reset;
    return fail(h, MPENV_ERR_INVALID, "mpenv_ray_counts: not an instrumented build (use libmpenv_b200_count.so)");
#endif
}

int mpenv_launches_per_step(mpenv_handle *h, uint32_t *step_launches, uint32_t *init_launches)
{
    if (!h) return MPENV_ERR_INVALID;
    if (step_launches) *step_launches = h->stepLaunches;
    if (init_launches) *init_launches = h->initLaunches;
    return MPENV_OK;
}

int mpenv_kernel_variants(mpenv_handle *h, int32_t *los_pool_warps, int32_t *lidar_packet_ctas)
{
    if (!h) return MPENV_ERR_INVALID;
    if (los_pool_warps) *los_pool_warps = h->poolLos ? h->losWarps : 0;
    if (lidar_packet_ctas) *lidar_packet_ctas = h->packetLidar;
    return MPENV_OK;
}

}  // extern "C"
