// ray_kernels.cuh -- the ray-heavy step kernels on the warp-pooled engine of
// ray_fast.cuh.  Results are bit-identical to the per-thread formulations in
// step_kernels.cuh (k_lidar / k_los), which stay as the path for maps whose BVH
// does not fit in shared memory.
#pragma once

#include "ray_fast.cuh"
#include "step_kernels.cuh"

namespace mpenv {

// ------------------------------------------------------------------- lidar
// pvpLidarSystem (sim.cpp:3324-3506): the kernel is k_lidar_packet (lidar_packet.cuh).
// A task is TWO agents of one world = 160 rays: 2 x (2 heights x 32 forward columns +
// 2 heights x 8 rear columns).  The two heights share their directions, so a task has
// 80 unique directions and 4 origins.
constexpr int kLidarRaysPerAgent = kFwdLidarH * kFwdLidarW + kRearLidarH * kRearLidarW;   // 80
constexpr int kLidarTaskRays = 2 * kLidarRaysPerAgent;                                    // 160

// ----------------------------------------------------------------- hitscan
// k_fire (step_kernels.cuh) with the shot traced on the FastBVH (FastMeshTrace).
template <bool kCold>
__global__ void __launch_bounds__(128) k_fire_fast(KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    const Ctx &c = ka.ctx;
    FastMeshTrace mesh;
    mesh.bvh = stageFastBVH(ka.fastImage, c.bvh.numNodes, c.bvh.numTris, smem_raw);
    mesh.smem = reinterpret_cast<const char *>(smem_raw);
    mesh.one = ka.one;
    const int N = (int)c.cfg.numWorlds * kAgents;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const int w = i / kAgents, a = i - w * kAgents;
        stageFire<kCold>(c, w, a, c.p.pos + (size_t)w * kAgents, mesh);
    }
}

// --------------------------------------------------------------------- LOS
// opponentsWriteVisibilitySystem (sim.cpp:2526-2560) + isAgentVisible
// (utils.cpp:186-271).  One warp owns one TEAM of a world as observers: 6 agents x 6
// opponents x 4 sample points = 144 candidates (a world per task left each warp with
// 2.3 tasks at 8192 worlds: the hand-out's tail was a third of a task long).  Most die in the alive / view-space cull; survivors
// are compacted into the warp's ray pool (ballot rank), traversed with dynamic
// refill.  A sample sees its target iff the closest hit of traceRayAgainstWorld
// IS the target; that is decided exactly without finishing every traversal:
// the capsule tests run first (target must win among the capsules), and a ray
// stops at the first ACCEPTED mesh hit with t <= t_target -- accepted hits only
// decrease, so such a hit exists iff the final closest mesh hit is <= t_target.
// canSee[pair] = OR of its four samples (pure per sample, order irrelevant).
#ifndef MPENV_LOS_WARPS
#define MPENV_LOS_WARPS 28      // one CTA per SM: the BVH image (27 KB on simple_map) staged once + 28 x 6.4 KB of pools; 24: 0.168 ms, 28: 0.163
#endif
constexpr int kLosWarpsPerCTA = MPENV_LOS_WARPS;
constexpr int kLosPoolCap = 96;
constexpr int kLosCandidates = kTeamSize * kTeamSize * 4;    // 144

struct __align__(16) LosWarpPool {
    RayDirRec dirs[kLosPoolCap];
    float4 origin[kAgents];          // eye position of each agent
    float4 invRot[kAgents];          // inverse aim rotation (x, y, z, w)
    float4 dRight[kAgents];          // 0.9 r along the observer's right axis
    float4 target[kAgents];          // position + view height of each agent as a target
    float tmesh[kLosPoolCap];
    float tstop[kLosPoolCap];        // distance to the target capsule: a mesh hit at t <= tstop blocks
    int32_t meta[kLosPoolCap];       // observer | target << 4 | pair << 8
    uint32_t seen[4];                // bit per (observer, opponent slot) pair of the task's team: 36 bits
    int32_t alive;                   // bit per agent
    int32_t pad[3];
};

struct LosPoolAccess {
    LosWarpPool *P;
    int32_t n;
    __device__ __forceinline__ int count() const { return n; }
    uint32_t dirsOff;                // shared-memory offset of P->dirs
    __device__ __forceinline__ float load(int ray, RayRegs &r)
    {
        const float4 o = P->origin[P->meta[ray] & 15];
        const RayDirRec &d = P->dirs[ray];
        setRay(r, o.x, o.y, o.z, d.invX, d.invY, d.invZ,
               dirsOff + (uint32_t)ray * (uint32_t)sizeof(RayDirRec) + (uint32_t)offsetof(RayDirRec, tri));
        r.tStop = P->tstop[ray];
        return r.tStop * 1.0003f;        // nothing beyond the target matters (see traversePool)
    }
    __device__ __forceinline__ void finish(int ray, float t_max) { P->tmesh[ray] = t_max; }
};

template <int kBatch, int kRefill>
#ifndef MPENV_LOS_MINB
#define MPENV_LOS_MINB 1        // 28 warps at <= 73 registers (72 used)
#endif
__global__ void __launch_bounds__(kLosWarpsPerCTA * 32, MPENV_LOS_MINB) k_los_pool(KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    const Ctx &c = ka.ctx;
    const FastBVH bvh = stageFastBVH(ka.fastImage, c.bvh.numNodes, c.bvh.numTris, smem_raw);
    const size_t bvh_bytes = fastBVHBytes(c.bvh.numNodes, c.bvh.numTris);
    LosWarpPool *pools = reinterpret_cast<LosWarpPool *>(reinterpret_cast<char *>(smem_raw) + bvh_bytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lane_lt = (1u << lane) - 1u;
    LosWarpPool &P = pools[warp];
    const SimPtrs &p = c.p;
    const float *fr = c.mc->frustum;
    const int num_tasks = (int)c.cfg.numWorlds * 2;

    for (;;) {
        const int task = fetchWarpTask(ka.workCounters + 2);
        if (task >= num_tasks) break;
        const int w = task >> 1, team = task & 1;
        const int base = w * kAgents;
        // ---- per-agent terms (utils.cpp:240-251)
        if (lane < 4) P.seen[lane] = 0;
        const bool alive_l = lane < kAgents && p.alive[base + lane] != 0.f;
        const unsigned alive_m = __ballot_sync(0xFFFFFFFFu, alive_l);
        if (lane < kAgents) {
            const int idx = base + lane;
            const Vec3 pos = p.pos[idx];
            const float vh = viewHeight(p.stand[idx].cur);
            const Quat rot = p.aim[idx].rot;
            const Quat inv = quatInv(rot);
            const Vec3 aim_right = rotateVec(rot, kRight());
            const Vec3 dr = aim_right * 0.9f * kAgentRadius;
            P.origin[lane] = make_float4(pos.x, pos.y, pos.z + vh, 0.f);
            P.invRot[lane] = make_float4(inv.x, inv.y, inv.z, inv.w);
            P.dRight[lane] = make_float4(dr.x, dr.y, dr.z, 0.f);
            P.target[lane] = make_float4(pos.x, pos.y, pos.z, vh);
        }
        __syncwarp();

        int32_t cand = 0;
        while (cand < kLosCandidates) {
            // ---- phase 1: cull candidates, append survivors to the pool
            int32_t n = 0;
            while (cand < kLosCandidates && n + 32 <= kLosPoolCap) {
                const int32_t cd = cand + lane;
                cand += 32;
                const int32_t pair = cd >> 2, s = cd & 3;
                const int32_t a_off = pair / kTeamSize, j = pair - a_off * kTeamSize;
                const int32_t a = team * kTeamSize + a_off;
                const int32_t opp = (team ^ 1) * kTeamSize + j;
                bool ok = cd < kLosCandidates && ((alive_m >> a) & 1u) && ((alive_m >> opp) & 1u);   // 144 = 4.5 warps
                Vec3 dir = v3(0.f, 1.f, 0.f);
                if (ok) {
                    const float4 o4 = P.origin[a], t4 = P.target[opp], q4 = P.invRot[a], r4 = P.dRight[a];
                    const Vec3 o = v3(o4.x, o4.y, o4.z);
                    const Vec3 pt = losSamplePoint(v3(t4.x, t4.y, t4.z), t4.w, v3(r4.x, r4.y, r4.z), s);
                    Quat inv; inv.x = q4.x; inv.y = q4.y; inv.z = q4.z; inv.w = q4.w;
                    // losSampleVisible up to the ray (utils.cpp:191-219)
                    Vec3 to_test = pt - o;
                    const Vec3 vp = rotateVec(inv, to_test);
                    ok = vp.y > 0.f;
                    ok = ok && (vp.y * fr[1] - fabsf(vp.x) * fr[0] > -kAgentRadius);
                    ok = ok && (vp.y * fr[3] - fabsf(vp.z) * fr[2] > -kAgentRadius);
                    const float len = length(to_test);
                    ok = ok && !(len < kAgentRadius);
                    dir = to_test / len;
                }
#if defined(MPENV_COUNT_RAYS)
                {   // the reference traces every candidate that survives the view-space cull
                    const unsigned cm = __ballot_sync(0xFFFFFFFFu, ok);
                    if (lane == 0 && cm) MP_COUNT_RAYS(1, __popc(cm));
                }
#endif
                // The agent capsules first (utils.cpp:39-61): the sample sees its target only
                // if, among the capsules, the target is the one traceRayAgainstWorld would
                // pick -- earlier agents must be strictly farther, later ones not nearer --
                // and then only if no mesh hit lies at t <= t_target.
                float t_tgt = 0.f;
                if (ok) {
                    const float4 o4 = P.origin[a];
                    const float a2 = dir.x * dir.x + dir.y * dir.y;
                    const float dd = a2 + dir.z * dir.z;
                    {
                        const float4 t4 = P.target[opp];
                        t_tgt = rayCapsuleRec(makeCapsuleRec(v3(o4.x, o4.y, o4.z), v3(t4.x, t4.y, t4.z)),
                                              dir.x, dir.y, dir.z, a2, dd);
                    }
                    ok = t_tgt != 0.f;
                    // the other capsules whose axis the ray passes within r + 2 of (for the rest t would be 0):
                    // a branch-free 12-bit mask, then a loop over the few set bits
                    uint32_t near = 0;
#pragma unroll
                    for (int i = 0; i < kAgents; i++) {
                        const float2 t2 = *reinterpret_cast<const float2 *>(&P.target[i]);
                        near |= capsuleMayHit(o4.x - t2.x, o4.y - t2.y, dir.x, dir.y, a2) ? 1u << i : 0u;
                    }
                    near &= ~(1u << opp);
                    while (near && ok) {
                        const int i = __ffs(near) - 1;
                        near &= near - 1u;
                        const float4 t4 = P.target[i];
                        const float t = rayCapsuleRec(makeCapsuleRec(v3(o4.x, o4.y, o4.z), v3(t4.x, t4.y, t4.z)),
                                                      dir.x, dir.y, dir.z, a2, dd);
                        if (t != 0.f && (i < opp ? !(t_tgt < t) : (t < t_tgt))) ok = false;
                    }
                }
                const unsigned m = __ballot_sync(0xFFFFFFFFu, ok);
                if (ok) {
                    const int32_t slot = n + __popc(m & lane_lt);
                    P.dirs[slot] = makeRayDirRec(dir);
                    P.tstop[slot] = t_tgt;
                    P.meta[slot] = a | (opp << 4) | (pair << 8);
                }
                n += __popc(m);
            }
            __syncwarp();
            if (n == 0) continue;

            // ---- phase 2: closest mesh hit
            LosPoolAccess acc{&P, n,
                              (uint32_t)(reinterpret_cast<const char *>(P.dirs) - reinterpret_cast<const char *>(smem_raw))};
            traversePool<kBatch, kRefill>(reinterpret_cast<const char *>(smem_raw), bvh, acc, ka.one);
            __syncwarp();

            // ---- phase 3: visible unless a mesh hit at t <= t_target was accepted
            for (int32_t ray = lane; ray < n; ray += 32) {
                const int32_t pair = P.meta[ray] >> 8;
                if (!(P.tmesh[ray] <= P.tstop[ray])) atomicOr(&P.seen[pair >> 5], 1u << (pair & 31));
            }
            __syncwarp();
        }
        // ---- export: canSee[agent][8] bytes, the team's 6 agents x 8 B contiguous
        if (lane < kTeamSize * 2) {
            const int32_t a = lane >> 1, half = lane & 1;     // a: offset within the team, as in `pair`
            uint32_t word = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int32_t j = half * 4 + k;
                if (j < kTeamSize) {
                    const int32_t pair = a * kTeamSize + j;
                    if ((P.seen[pair >> 5] >> (pair & 31)) & 1u) word |= 1u << (8 * k);
                }
            }
            reinterpret_cast<uint32_t *>(p.canSee)[(size_t)(base + team * kTeamSize) * 2 + lane] = word;
        }
        __syncwarp();
    }
    warpTasksDone(ka.workCounters + 2, gridDim.x * (blockDim.x >> 5));
}

// The CTA is launched with as many warps (<= kLosWarpsPerCTA) as have room for their pool beside the BVH.
inline size_t losPoolSmemBytes(uint32_t num_nodes, uint32_t num_tris, int warps)
{
    return fastBVHBytes(num_nodes, num_tris) + (size_t)warps * sizeof(LosWarpPool);
}
inline int losPoolWarps(uint32_t num_nodes, uint32_t num_tris, size_t smem_limit)
{
    const size_t bvh = fastBVHBytes(num_nodes, num_tris);
    if (bvh >= smem_limit) return 0;
    const size_t w = (smem_limit - bvh) / sizeof(LosWarpPool);
    return (int)(w < (size_t)kLosWarpsPerCTA ? w : (size_t)kLosWarpsPerCTA);
}

}  // namespace mpenv
