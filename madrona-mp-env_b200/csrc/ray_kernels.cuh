// ray_kernels.cuh -- the ray-heavy step kernels on the warp-pooled engine of
// ray_fast.cuh.  Results are bit-identical to the per-thread formulations in
// step_kernels.cuh (k_lidar / k_los), which stay as the path for maps whose BVH
// does not fit in shared memory.
#pragma once

#include "ray_fast.cuh"
#include "step_kernels.cuh"

namespace mpenv {

// ------------------------------------------------------------------- lidar
// pvpLidarSystem (sim.cpp:3324-3506).  One warp owns a task of TWO agents (same
// world) = 160 rays: 2 x (2 heights x 32 forward columns + 2 heights x 8 rear
// columns).  The two heights share their directions, so a task has 80 unique
// directions and 4 origins.
constexpr int kLidarWarpsPerCTA = 8;
#ifndef MPENV_LEAF_BATCH
#define MPENV_LEAF_BATCH 10
#endif
constexpr int kLeafBatch = MPENV_LEAF_BATCH;
constexpr int kLidarRaysPerAgent = kFwdLidarH * kFwdLidarW + kRearLidarH * kRearLidarW;   // 80
constexpr int kLidarDirsPerAgent = kFwdLidarW + kRearLidarW;                              // 40
constexpr int kLidarTaskRays = 2 * kLidarRaysPerAgent;                                    // 160

struct __align__(16) LidarWarpPool {
    RayDirRec dirs[2 * kLidarDirsPerAgent];
    CapsuleRec caps[4][kAgents];
    float4 origin[4];                    // [agent * 2 + height]
    float tmesh[kLidarTaskRays];         // closest mesh hit, -1 = miss
    int32_t counter;
    int32_t pad[3];
};

// ray index within a task -> (agent 0/1, ray 0..79, height, direction slot)
__device__ __forceinline__ void lidarDecode(int ray, int *agent, int *rr, int *h, int *dir)
{
    const int ag = ray >= kLidarRaysPerAgent ? 1 : 0;
    const int r = ray - ag * kLidarRaysPerAgent;
    int hh, d;
    if (r < kFwdLidarH * kFwdLidarW) { hh = r >> 5; d = r & 31; }
    else { const int r2 = r - kFwdLidarH * kFwdLidarW; hh = r2 >> 3; d = kFwdLidarW + (r2 & 7); }
    *agent = ag; *rr = r; *h = hh; *dir = ag * kLidarDirsPerAgent + d;
}

struct LidarPoolAccess {
    LidarWarpPool *P;
    __device__ __forceinline__ int next()
    {
        const int i = atomicAdd(&P->counter, 1);
        return i < kLidarTaskRays ? i : -1;
    }
    __device__ __forceinline__ void load(int ray, RayRegs &r)
    {
        int ag, rr, h, d;
        lidarDecode(ray, &ag, &rr, &h, &d);
        const float4 o = P->origin[ag * 2 + h];
        setRay(r, o.x, o.y, o.z, P->dirs[d]);
    }
    __device__ __forceinline__ void finish(int ray, bool hit, float t) { P->tmesh[ray] = hit ? t : -1.f; }
};

template <int kBatch>
__global__ void __launch_bounds__(kLidarWarpsPerCTA * 32) k_lidar_pool(KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    const Ctx &c = ka.ctx;
    const FastBVH bvh = stageFastBVH(c.bvh, smem_raw);
    const size_t bvh_bytes = fastBVHBytes(c.bvh.numNodes, c.bvh.numTris);
    LidarWarpPool *pools = reinterpret_cast<LidarWarpPool *>(reinterpret_cast<char *>(smem_raw) + bvh_bytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    LidarWarpPool &P = pools[warp];
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int num_tasks = (int)c.cfg.numWorlds * (kAgents / 2);

    for (int task = blockIdx.x * kLidarWarpsPerCTA + warp; task < num_tasks;
         task += gridDim.x * kLidarWarpsPerCTA) {
        const int idx0 = task * 2;                       // first agent's global index
        const int base = (idx0 / kAgents) * kAgents;     // the world's first agent

        // ---- phase 1: origins, directions, capsule records (converged lanes)
        if (lane < 4) {
            const int idx = idx0 + (lane >> 1), h = lane & 1;
            // sim.cpp:3402-3437: two sample heights between foot + r and eye level
            const float top_height = viewHeight(p.stand[idx].cur) + kAgentRadius;
            const float range = top_height - 2.f * kAgentRadius;
            const float h_off = kAgentRadius + range * (float(h) / float(kFwdLidarH - 1));
            Vec3 o = p.pos[idx];
            o.z += h_off;
            P.origin[lane] = make_float4(o.x, o.y, o.z, 0.f);
        }
        if (lane == 0) P.counter = 0;
        for (int d = lane; d < 2 * kLidarDirsPerAgent; d += 32) {
            const int ag = d >= kLidarDirsPerAgent ? 1 : 0;
            const int col = d - ag * kLidarDirsPerAgent;
            const int idx = idx0 + ag;
            const bool is_fwd = col < kFwdLidarW;
            const Quat q = is_fwd ? p.aim[idx].rot : p.rot[idx];
            const Vec3 fwd = rotateVec(q, kFwd());
            const Vec3 right = rotateVec(q, kRight());
            const float x = is_fwd ? mc.fwdLidarX[col] : mc.rearLidarX[col - kFwdLidarW];
            const float y = is_fwd ? mc.fwdLidarY[col] : mc.rearLidarY[col - kFwdLidarW];
            P.dirs[d] = makeRayDirRec(normalize(x * right + y * fwd));
        }
        __syncwarp();
        for (int e = lane; e < 4 * kAgents; e += 32) {
            const int oi = e / kAgents, i = e - oi * kAgents;
            const float4 o = P.origin[oi];
            P.caps[oi][i] = makeCapsuleRec(v3(o.x, o.y, o.z), p.pos[base + i]);
        }
        __syncwarp();

        // ---- phase 2: closest mesh hit of the 160 rays
        LidarPoolAccess acc{&P};
        traversePool<kBatch>(bvh, acc);
        __syncwarp();

        // ---- phase 3: agent capsules, classification, coalesced export
        for (int ray = lane; ray < kLidarTaskRays; ray += 32) {
            int ag, rr, h, d;
            lidarDecode(ray, &ag, &rr, &h, &d);
            const RayDirRec &dr = P.dirs[d];
            const float tm = P.tmesh[ray];
            float hit_t;
            int32_t hit_agent;
            const bool hit = closestWithCapsules(P.caps[ag * 2 + h], tm >= 0.f, tm, dr.dx, dr.dy, dr.dz,
                                                 dr.a2, dr.dd, &hit_t, &hit_agent);
            const int idx = idx0 + ag, a = idx - base;
            float4 v;
            if (hit) {
                v.x = fminf(hit_t, mc.maxDist);
                v.y = hit_agent < 0 ? 1.f : 0.f;
                v.z = (hit_agent >= 0 && teamOf(hit_agent) == teamOf(a)) ? 1.f : 0.f;
                v.w = (hit_agent >= 0 && teamOf(hit_agent) != teamOf(a)) ? 1.f : 0.f;
            } else {
                v = make_float4(-1.f, 0.f, 0.f, 0.f);
            }
            if (rr < kFwdLidarH * kFwdLidarW)
                reinterpret_cast<float4 *>(p.fwdLidar)[(size_t)idx * (kFwdLidarH * kFwdLidarW) + rr] = v;
            else
                reinterpret_cast<float4 *>(p.rearLidar)[(size_t)idx * (kRearLidarH * kRearLidarW) +
                                                        (rr - kFwdLidarH * kFwdLidarW)] = v;
        }
        __syncwarp();
    }
}

inline size_t lidarPoolSmemBytes(uint32_t num_nodes, uint32_t num_tris)
{
    return fastBVHBytes(num_nodes, num_tris) + kLidarWarpsPerCTA * sizeof(LidarWarpPool);
}

}  // namespace mpenv
