// step_kernels.cuh -- the sm_100a kernels of one MadronaMPEnv step, one per
// stage group of the reference Step task graph (sim.cpp:5342-5842):
//
//   k_move      movement + aim + swept-sphere collision + fall   (thread / agent)
//   k_fire      hitscan fire                                     (thread / agent)
//   k_world     damage, respawn, zone, breadcrumbs, match info,
//               rewards, done, reset                             (thread / world)
//   k_los       line-of-sight visibility                         (thread / (agent, opponent))
//   k_obs       opponent masks + observation packing             (thread / agent)
//   k_fullteam  full-team observation tensors (optional)         (warp / team)
//   k_lidar     80 lidar rays per agent                          (thread / ray)
//
// All kernels are persistent grid-stride loops sized from the SM count; the
// static-mesh BVH (nodes + de-indexed triangle positions, ~13 KB on
// simple_map) is staged into shared memory once per CTA, so traversal never
// touches global memory; world state is read / written with plain coalesced
// loads and stores.  Compiled with -fmad=false: the arithmetic is the stage
// functions' (mp_stages*.h), operation for operation.
#pragma once

#include <cuda_runtime.h>

#include "mp_stages.h"
#include "sphere_fast.cuh"
#include "world_coop.cuh"

namespace mpenv {

struct KernelArgs {
    Ctx ctx;                 // ctx.bvh points at the GLOBAL copies
    uint32_t bvhNodeWords;   // 32-bit words of node data
    uint32_t bvhVertWords;   // 32-bit words of vertex data (incl. the pad triangle)
    int32_t bvhInSmem;       // 1: stage into shared memory, 0: traverse from global / L2
    uint32_t *workCounters;  // device-wide task counters of the dynamically scheduled kernels (zero between launches)
    const uint4 *fastImage;  // FastBVH image (ray_fast.cuh), null when no kernel uses it
    const uint4 *sphereImage;// SphereBVH image (sphere_fast.cuh), null when k_move runs on the reference format
    float one;               // 1.0f, opaque to ptxas: keeps the two roundings of q * dq + oq apart (ray_fast.cuh)
};

// ---- dynamic task hand-out --------------------------------------------------
// A kernel whose tasks vary in cost draws them from a device-wide counter pair
// {next task, warps finished} (KernelArgs::workCounters, zero between launches)
// instead of a static stride: warps that get cheap tasks take more of them.
// Returns the next task id (>= the task count once they have run out).
__device__ __forceinline__ int fetchWarpTask(uint32_t *pair)
{
    int t = 0;
    if ((threadIdx.x & 31) == 0) t = (int)atomicAdd(&pair[0], 1u);
    return __shfl_sync(0xFFFFFFFFu, t, 0);
}

// Every warp of the grid calls this once, after its last fetch; the last one re-arms the pair.
__device__ __forceinline__ void warpTasksDone(uint32_t *pair, uint32_t total_warps)
{
    if ((threadIdx.x & 31) == 0) {
        __threadfence();
        if (atomicAdd(&pair[1], 1u) == total_warps - 1u) {
            pair[0] = 0;
            pair[1] = 0;
        }
    }
}

// Copies the BVH into dynamic shared memory and returns a view onto it.
__device__ __forceinline__ BVHView stageBVH(const KernelArgs &ka, uint32_t *smem)
{
    BVHView v = ka.ctx.bvh;
    if (!ka.bvhInSmem) return v;
    const uint32_t *gn = reinterpret_cast<const uint32_t *>(ka.ctx.bvh.nodes);
    const uint32_t *gv = reinterpret_cast<const uint32_t *>(ka.ctx.bvh.verts);
    uint32_t *sn = smem;
    uint32_t *sv = smem + ka.bvhNodeWords;
    // 128-bit copies: node data is a multiple of 64 bytes; the vertex tail is
    // finished word by word.
    const uint32_t n4 = ka.bvhNodeWords / 4;
    for (uint32_t i = threadIdx.x; i < n4; i += blockDim.x)
        reinterpret_cast<uint4 *>(sn)[i] = __ldg(reinterpret_cast<const uint4 *>(gn) + i);
    const uint32_t v4 = ka.bvhVertWords / 4;
    for (uint32_t i = threadIdx.x; i < v4; i += blockDim.x)
        reinterpret_cast<uint4 *>(sv)[i] = __ldg(reinterpret_cast<const uint4 *>(gv) + i);
    for (uint32_t i = v4 * 4 + threadIdx.x; i < ka.bvhVertWords; i += blockDim.x) sv[i] = __ldg(gv + i);
    __syncthreads();
    v.nodes = reinterpret_cast<const BVHNode *>(sn);
    v.verts = reinterpret_cast<const Vec3 *>(sv);
    return v;
}

__global__ void __launch_bounds__(128) k_ctor(KernelArgs ka)
{
    const int W = (int)ka.ctx.cfg.numWorlds;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < W; w += gridDim.x * blockDim.x)
        ctorWorld(ka.ctx, w);
}

__global__ void __launch_bounds__(128) k_force_reset(KernelArgs ka)
{
    const int W = (int)ka.ctx.cfg.numWorlds;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < W; w += gridDim.x * blockDim.x)
        ka.ctx.p.resetFlag[w] = 1;
}

__global__ void __launch_bounds__(128) k_reset(KernelArgs ka)
{
    const int W = (int)ka.ctx.cfg.numWorlds;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < W; w += gridDim.x * blockDim.x)
        resetSystem(ka.ctx, w);
}

#ifndef MPENV_MOVE_MINB
#define MPENV_MOVE_MINB 6       // 80 registers, 24 warps / SM (measured 5: 0.101 ms, 6: 0.096, 7: 0.102)
#endif
// kFast: the sphere casts run on SphereNodes staged in shared memory (sphere_fast.cuh); else on the
// reference-format BVH (shared memory or, for a map that does not fit, L2).
template <bool kFast>
__global__ void __launch_bounds__(128, MPENV_MOVE_MINB) k_move(KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    Ctx c = ka.ctx;
    const int N = (int)c.cfg.numWorlds * kAgents;
    // 1 .. 11 sphere casts per agent: 32-agent chunks are drawn dynamically (a static stride
    // gave 1.3 agents per thread, i.e. a second, 30 %-full round)
    const int chunks = (N + 31) / 32;
    if (kFast) {
        const SphereBVH sb = stageSphereBVH(ka.sphereImage, c.bvh.numNodes, ka.bvhVertWords, smem_raw);
        for (;;) {
            const int chunk = fetchWarpTask(ka.workCounters + 4);
            if (chunk >= chunks) break;
            const int i = chunk * 32 + (int)(threadIdx.x & 31);
            if (i < N) stageMove(c, sb, i);
        }
    } else {
        c.bvh = stageBVH(ka, reinterpret_cast<uint32_t *>(smem_raw));
        for (;;) {
            const int chunk = fetchWarpTask(ka.workCounters + 4);
            if (chunk >= chunks) break;
            const int i = chunk * 32 + (int)(threadIdx.x & 31);
            if (i < N) stageMove(c, i);
        }
    }
    warpTasksDone(ka.workCounters + 4, gridDim.x * (blockDim.x >> 5));
}

template <bool kCold>
__global__ void __launch_bounds__(128) k_fire(KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    Ctx c = ka.ctx;
    c.bvh = stageBVH(ka, reinterpret_cast<uint32_t *>(smem_raw));
    const int N = (int)c.cfg.numWorlds * kAgents;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const int w = i / kAgents, a = i - w * kAgents;
        stageFire<kCold>(c, w, a, c.p.pos + (size_t)w * kAgents);
    }
}

// stageWorld (mp_stages_world.h) executed by a 16-lane group per world: the
// per-agent systems (damage, autoheal, explore grid, zone reward) run on lanes
// 0..11 in parallel -- each touches only its own agent's rows -- and the systems
// with sequential-in-world semantics (respawn draw order, zone state machine,
// breadcrumb pool order, match info, team reward, reset) run on lane 0, in task
// graph order (sim.cpp:5508-5842).  __syncwarp() orders the phases' global-memory
// traffic.  16x the threads of a thread-per-world launch: latency is what bounds
// this stage.
template <bool kCold>
__global__ void __launch_bounds__(128) k_world(KernelArgs ka)
{
    const Ctx &c = ka.ctx;
    const SimPtrs &p = c.p;
    const int W = (int)c.cfg.numWorlds;
    const int lane = threadIdx.x & 15;
    const int half = (threadIdx.x >> 4) & 1;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int num_warps = (gridDim.x * blockDim.x) >> 5;
    for (int wp = warp; wp * 2 < W; wp += num_warps) {
        const int w = wp * 2 + half;
        const bool valid = w < W;
        const int base = w * kAgents;
        const bool agent = valid && lane < kAgents;
        const bool lead = valid && lane == 0;
        if (kCold && c.cfg.replayLog) {
            // pvpReplayLogic (sim.cpp:5587-5604): logged state, then the zone state machine
            if (lead) p.world[w].match.curStep = p.replayLog[w].curStep;
            if (agent) replayAgent(c, w, lane);
            __syncwarp();
            zoneSystemCoop(c, w, lane, half, valid);
        } else {
            if (agent) applyDmgSystem(p, base + lane);
            __syncwarp();
            if (lead && !(c.cfg.simFlags & kFlagNoRespawn)) spawnAgents(c, w, true);
            __syncwarp();
            if (agent) autoHealSystem(p, base + lane);
            __syncwarp();
            zoneSystemCoop(c, w, lane, half, valid);
            if (kCold && (c.cfg.simFlags & kFlagSubZones)) {           // sim.cpp:5549-5555
                __syncwarp();
                if (lead) subzoneSystems(c, w);
                __syncwarp();
            }
            if (kCold && c.cfg.recordLog && agent) recordAgent(c, w, lane);
            breadcrumbSystemsCoop(c, w, lane, half, valid);
        }
        if (lead) zoneMatchInfoSystem<kCold>(c, w);
        __syncwarp();
        evaluateGoalRegionsCoop(c, w, lane, half, valid);
        if (agent) {
            exploreVisitedSystem(c, w, base + lane);
            if (kCold && c.cfg.taskCaptureDefend) zoneCaptureDefendRewardSystem(c, w, lane);
            else if (kCold && c.cfg.rewardFlank) flankRewardSystem(c, w, lane);
            else if (kCold && (c.cfg.simFlags & kFlagSubZones)) subzoneRewardSystem(c, w, lane);
            else zoneRewardSystem(c, w, lane);
        }
        __syncwarp();
        if (lead) {
            teamAndFinalReward(c, w);
            const int32_t finished = p.world[w].match.isFinished;
            for (int i = 0; i < kAgents; i++) p.done[base + i] = finished;
            // fullTeamDoneRewardSystem (sim.cpp:4720-4748)
            for (int t = 0; t < 2; t++) {
                float r = 0.f;
                for (int i = 0; i < kTeamSize; i++) r += p.reward[base + t * kTeamSize + i];
                p.ftReward[w * 2 + t] = r;
                p.ftDone[w * 2 + t] = finished ? 1 : 0;
            }
            resetSystem<kCold>(c, w);
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(128) k_los(KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    Ctx c = ka.ctx;
    c.bvh = stageBVH(ka, reinterpret_cast<uint32_t *>(smem_raw));
    const int P = (int)c.cfg.numWorlds * kAgents * kTeamSize;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
        const int idx = i / kTeamSize, j = i - idx * kTeamSize;
        const int w = idx / kAgents, a = idx - w * kAgents;
        c.p.canSee[(size_t)idx * 8 + j] = visibilityPair(c, w, a, j, c.p.pos + (size_t)w * kAgents) ? 1 : 0;
    }
}

// One thread per output ROW (12 per agent: self, 5 teammates, 6 opponents), rows
// enumerated TYPE-major -- all self rows, then all teammate rows, then all opponent
// rows -- so a warp runs one row function instead of three (9.8 -> ~30 active lanes),
// and consecutive lanes still write consecutive 128-byte rows of one tensor.
// (128, 8) -> 64 registers and (128, 10) -> 48 were measured: no gain; an explicit (128, 1) costs 12 %
__global__ void __launch_bounds__(128) k_obs(KernelArgs ka)
{
    const Ctx &c = ka.ctx;
    const long long N = (long long)c.cfg.numWorlds * kAgents;
    const long long R = N * kObsRowsPerAgent;
    const long long first_opponent = N * kTeamSize;       // N self + 5 N teammate rows come first
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < R;
         g += (long long)gridDim.x * blockDim.x) {
        int i, r;
        if (g < N) { i = (int)g; r = 0; }
        else if (g < first_opponent) {
            const long long t = g - N;
            i = (int)(t / (kTeamSize - 1)); r = 1 + (int)(t - (long long)i * (kTeamSize - 1));
        } else {
            const long long t = g - first_opponent;
            i = (int)(t / kTeamSize); r = kTeamSize + (int)(t - (long long)i * kTeamSize);
        }
        const int w = i / kAgents, a = i - w * kAgents;
        stageObsRow(c, w, a, r);
    }
}

__global__ void __launch_bounds__(128) k_fullteam(KernelArgs ka)
{
    const Ctx &c = ka.ctx;
    const int T = (int)c.cfg.numWorlds * 2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < T; i += gridDim.x * blockDim.x)
        stageFullTeamObs(c, i / 2, i & 1);
}

__global__ void __launch_bounds__(160) k_lidar(KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    Ctx c = ka.ctx;
    c.bvh = stageBVH(ka, reinterpret_cast<uint32_t *>(smem_raw));
    const long long R = (long long)c.cfg.numWorlds * kAgents * 80;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < R;
         g += (long long)gridDim.x * blockDim.x) {
        const int idx = (int)(g / 80), ray = (int)(g - (long long)idx * 80);
        const int w = idx / kAgents, a = idx - w * kAgents;
        LidarSample s = lidarRay(c, w, a, ray, c.p.pos + (size_t)w * kAgents);
        float4 v = make_float4(s.depth, s.isWall, s.isTeammate, s.isOpponent);
        if (ray < 64) reinterpret_cast<float4 *>(c.p.fwdLidar)[(size_t)idx * 64 + ray] = v;
        else reinterpret_cast<float4 *>(c.p.rearLidar)[(size_t)idx * 16 + (ray - 64)] = v;
    }
}

// ---- event log compaction (replaces the reference's atomic append, sim.cpp:23-39,
// with a deterministic order) -------------------------------------------------
// k_event_scan: ONE CTA; per world counts the used event slots and whether a step
// snapshot was written, exclusive-scans both over worlds (chunks of 1024 with a
// carry) and leaves the totals in totals[0..1].
__global__ void __launch_bounds__(1024) k_event_scan(KernelArgs ka, uint32_t *ev_offset, uint32_t *snap_offset,
                                                     uint32_t *totals)
{
    __shared__ uint32_t warp_sums[2][32];
    __shared__ uint32_t carry[2];
    const SimPtrs &p = ka.ctx.p;
    const int W = (int)ka.ctx.cfg.numWorlds;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { carry[0] = 0; carry[1] = 0; }
    __syncthreads();
    for (int w0 = 0; w0 < W; w0 += 1024) {
        const int w = w0 + threadIdx.x;
        uint32_t ne = 0, ns = 0;
        if (w < W) {
            const GameEvent *ev = p.evAgent + (size_t)w * kAgents * kEventSlotsPerAgent;
            for (int k = 0; k < kAgents * kEventSlotsPerAgent; k++) ne += ev[k].type != kEvNone ? 1u : 0u;
            const GameEvent cap = p.evWorld[w];
            ne += cap.type != kEvNone ? 1u : 0u;
            ns = cap.pad0 ? 1u : 0u;
        }
        uint32_t se = ne, ss = ns;                 // inclusive warp scans
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t te = __shfl_up_sync(0xFFFFFFFFu, se, off), ts = __shfl_up_sync(0xFFFFFFFFu, ss, off);
            if (lane >= off) { se += te; ss += ts; }
        }
        if (lane == 31) { warp_sums[0][warp] = se; warp_sums[1][warp] = ss; }
        __syncthreads();
        if (warp == 0) {
            uint32_t ve = warp_sums[0][lane], vs = warp_sums[1][lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t te = __shfl_up_sync(0xFFFFFFFFu, ve, off), ts = __shfl_up_sync(0xFFFFFFFFu, vs, off);
                if (lane >= off) { ve += te; vs += ts; }
            }
            warp_sums[0][lane] = ve; warp_sums[1][lane] = vs;
        }
        __syncthreads();
        const uint32_t base_e = carry[0] + (warp ? warp_sums[0][warp - 1] : 0u);
        const uint32_t base_s = carry[1] + (warp ? warp_sums[1][warp - 1] : 0u);
        if (w < W) { ev_offset[w] = base_e + se - ne; snap_offset[w] = base_s + ss - ns; }
        __syncthreads();
        if (threadIdx.x == 0) { carry[0] += warp_sums[0][31]; carry[1] += warp_sums[1][31]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { totals[0] = carry[0]; totals[1] = carry[1]; }
}

// k_event_scatter: thread / world copies its used slots (slot order = agent order,
// Capture last) and its step snapshot to the scanned offsets.
__global__ void __launch_bounds__(128) k_event_scatter(KernelArgs ka, const uint32_t *ev_offset,
                                                       const uint32_t *snap_offset, GameEvent *ev_out,
                                                       PackedStepSnapshot *snap_out)
{
    const SimPtrs &p = ka.ctx.p;
    const int W = (int)ka.ctx.cfg.numWorlds;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < W; w += gridDim.x * blockDim.x) {
        uint32_t o = ev_offset[w];
        const GameEvent *ev = p.evAgent + (size_t)w * kAgents * kEventSlotsPerAgent;
        for (int k = 0; k < kAgents * kEventSlotsPerAgent; k++) {
            const GameEvent e = ev[k];
            if (e.type != kEvNone) ev_out[o++] = e;
        }
        GameEvent cap = p.evWorld[w];
        if (cap.pad0) {
            const uint4 *src = reinterpret_cast<const uint4 *>(p.stepSnap + w);
            uint4 *dst = reinterpret_cast<uint4 *>(snap_out + snap_offset[w]);
#pragma unroll
            for (int k = 0; k < (int)(sizeof(PackedStepSnapshot) / 16); k++) dst[k] = src[k];
        }
        if (cap.type != kEvNone) { cap.pad0 = 0; ev_out[o] = cap; }
    }
}

// Episode statistics for the multi-GPU allreduce: per-device sums
// over worlds.  out[0] worlds, [1] episodes finished this step, [2..4] wins
// team0 / team1 / draw among them, [5..6] kills, [7..8] objective points of the
// finished episodes, [9] agents alive, [10] sum(done), [11] sum(reward) in 1e-6
// units, [12] sum(cur_step); the rest reserved.
__global__ void __launch_bounds__(256) k_stats(KernelArgs ka, unsigned long long *out)
{
    const Ctx &c = ka.ctx;
    const int W = (int)c.cfg.numWorlds;
    long long acc[13];
    for (int k = 0; k < 13; k++) acc[k] = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < W; w += gridDim.x * blockDim.x) {
        acc[0] += 1;
        const MatchResult &mr = c.p.matchResult[w];
        // done is exported per agent; row 0 of the world carries the episode flag
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
        if ((threadIdx.x & 31) == 0 && v != 0) atomicAdd(out + k, (unsigned long long)v);
    }
}

}  // namespace mpenv
