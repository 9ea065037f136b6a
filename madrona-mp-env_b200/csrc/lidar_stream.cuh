// lidar_stream.cuh -- pvpLidarSystem (sim.cpp:3324-3506) as a STREAM of tasks per
// warp.  Same rays, same per-ray arithmetic and visiting order as k_lidar_pool
// (ray_kernels.cuh); what changes is the schedule.  With one 160-ray pool at a
// time, a third of the traversal iterations ran after the pool's last ray had
// been handed out, at 6.7 of 32 lanes (measured with loop counters in the
// instrumented build).  Here two tasks are resident per warp: lanes that run
// out of rays of task j start on task j+1 at once, and the warp leaves the
// traversal loop only for the converged work of a finished task (agent capsules,
// export) and to set up the task after next in the buffer that just fell free.
#pragma once

#include "ray_kernels.cuh"

namespace mpenv {

#ifndef MPENV_STREAM_WARPS
// one CTA of 28 warps per SM: 72 registers x 896 threads fills the register file, 206 KB of
// shared memory (BVH staged once per SM); measured against 3 CTAs x 8 warps: 1.243 -> 1.165 ms
#define MPENV_STREAM_WARPS 28
#define MPENV_STREAM_MINB 1
#endif
constexpr int kStreamWarpsPerCTA = MPENV_STREAM_WARPS;
#ifndef MPENV_STREAM_LEAF
#define MPENV_STREAM_LEAF 10      // leaf-test / refill batch thresholds (see traversePool)
#define MPENV_STREAM_REFILL 8
#endif

// The traversal half of RayDirRec; the direction itself is only needed by the
// capsule phase and is recomputed there (same expressions, see lidarDir).
struct __align__(16) RayTravRec {
    float invX, invY, invZ;
    int32_t pad;
    RayTriRec tri;
};
static_assert(sizeof(RayTravRec) == 32, "RayTravRec is 2 x 16 bytes");

struct __align__(16) LidarStreamPool {
    RayTravRec trav[2][2 * kLidarDirsPerAgent];      // [buffer][direction]; a finished task's half is
                                                     // re-used for its capsule records + directions
    float4 origin[2][4];                             // [buffer][agent * 2 + height]
    float tmesh[2][kLidarTaskRays];                  // closest mesh hit, -1 = miss
    int32_t st[12];                                  // fin, prep, iss, next_ray, remaining[2], task[0], 0, task[1]
};
static_assert(sizeof(RayTravRec) * 2 * kLidarDirsPerAgent >=
              sizeof(CapsuleRec) * 4 * kAgents + sizeof(float) * 3 * 2 * kLidarDirsPerAgent,
              "capsule records + directions must fit in one traversal buffer");

// Direction `d` (0..79: agent 0 forward 32 + rear 8, then agent 1) of the task
// whose first agent is idx0 (sim.cpp:3439-3470).
__device__ __forceinline__ Vec3 lidarDir(const Ctx &c, int idx0, int d)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int ag = d >= kLidarDirsPerAgent ? 1 : 0;
    const int col = d - ag * kLidarDirsPerAgent;
    const int idx = idx0 + ag;
    const bool is_fwd = col < kFwdLidarW;
    const Quat q = is_fwd ? p.aim[idx].rot : p.rot[idx];
    const Vec3 fwd = rotateVec(q, kFwd());
    const Vec3 right = rotateVec(q, kRight());
    const float x = is_fwd ? mc.fwdLidarX[col] : mc.rearLidarX[col - kFwdLidarW];
    const float y = is_fwd ? mc.fwdLidarY[col] : mc.rearLidarY[col - kFwdLidarW];
    return normalize(x * right + y * fwd);
}

// Origins and per-direction traversal constants of `task` into buffer b.
__device__ __noinline__ void lidarStreamSetup(const Ctx &c, LidarStreamPool &P, int b, int task)
{
    const SimPtrs &p = c.p;
    const int lane = threadIdx.x & 31;
    const int idx0 = task * 2;
    if (lane < 4) {
        const int idx = idx0 + (lane >> 1), h = lane & 1;
        // sim.cpp:3402-3437: two sample heights between foot + r and eye level
        const float top_height = viewHeight(p.stand[idx].cur) + kAgentRadius;
        const float range = top_height - 2.f * kAgentRadius;
        const float h_off = kAgentRadius + range * (float(h) / float(kFwdLidarH - 1));
        Vec3 o = p.pos[idx];
        o.z += h_off;
        P.origin[b][lane] = make_float4(o.x, o.y, o.z, 0.f);
    }
    for (int d = lane; d < 2 * kLidarDirsPerAgent; d += 32) {
        const RayDirRec r = makeRayDirRec(lidarDir(c, idx0, d));
        RayTravRec t;
        t.invX = r.invX; t.invY = r.invY; t.invZ = r.invZ;
        t.pad = 0;
        t.tri = r.tri;
        P.trav[b][d] = t;
    }
    if (lane == 0) MP_COUNT_RAYS(0, kLidarTaskRays);
    __syncwarp();
}

// All 160 rays of the task in buffer b have their mesh result: agent capsules,
// classification and the coalesced float4 export (the second half of
// traceRayAgainstWorld, utils.cpp:39-61, and sim.cpp:3472-3500).
__device__ __noinline__ void lidarStreamExport(const Ctx &c, LidarStreamPool &P, int b, int task)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int lane = threadIdx.x & 31;
    const int idx0 = task * 2;
    const int base = (idx0 / kAgents) * kAgents;
    // no ray of this task reads trav[b] any more
    CapsuleRec *caps = reinterpret_cast<CapsuleRec *>(P.trav[b]);
    float *dvec = reinterpret_cast<float *>(caps + 4 * kAgents);
    for (int e = lane; e < 4 * kAgents; e += 32) {
        const int oi = e / kAgents, i = e - oi * kAgents;
        const float4 o = P.origin[b][oi];
        caps[e] = makeCapsuleRec(v3(o.x, o.y, o.z), p.pos[base + i]);
    }
    for (int d = lane; d < 2 * kLidarDirsPerAgent; d += 32) {
        const Vec3 dir = lidarDir(c, idx0, d);
        dvec[d * 3 + 0] = dir.x; dvec[d * 3 + 1] = dir.y; dvec[d * 3 + 2] = dir.z;
    }
    __syncwarp();
    for (int ray = lane; ray < kLidarTaskRays; ray += 32) {
        int ag, rr, h, d;
        lidarDecode(ray, &ag, &rr, &h, &d);
        const float dx = dvec[d * 3 + 0], dy = dvec[d * 3 + 1], dz = dvec[d * 3 + 2];
        const float a2 = dx * dx + dy * dy;          // as makeRayDirRec
        const float dd = a2 + dz * dz;
        const float tm = P.tmesh[b][ray];
        float hit_t;
        int32_t hit_agent;
        const bool hit = closestWithCapsules(caps + (ag * 2 + h) * kAgents, tm >= 0.f, tm, dx, dy, dz,
                                             a2, dd, &hit_t, &hit_agent);
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

// Per-lane traversal state (everything that lives across loop iterations).
#define MP_LANE_STATE(F, I) \
    F(0, r.ox) F(1, r.oy) F(2, r.oz) F(3, r.invX) F(4, r.invY) F(5, r.invZ) I(6, r.offX) I(7, r.offY) I(8, r.offZ) \
    F(9, t_max) F(10, tn[0]) F(11, tn[1]) F(12, tn[2]) F(13, tn[3]) I(14, nd) I(15, rem) I(16, leaf) \
    I(17, sp) I(18, ray)
#define MP_STASH_F(i, x) sv[i] = __float_as_int(x);
#define MP_STASH_I(i, x) sv[i] = (int32_t)(x);
#define MP_UNSTASH_F(i, x) x = __int_as_float(sv[i]);
#define MP_UNSTASH_I(i, x) x = sv[i];
constexpr int kLaneStateWords = 19;

// Tasks are handed out by a device-wide counter (warps that draw cheap tasks take more of
// them); a warp numbers the tasks it draws 0, 1, ... and task seq lives in buffer seq & 1.  Warp-uniform stream state:
//   fin   oldest task whose export has not run;     prep  tasks set up so far (<= fin + 2)
//   iss   task whose rays are being handed out;     next_ray  its next ray
//   remaining0/1  rays of the task in buffer b that have not been retired
// (kept in LidarStreamPool::st).  The node visit / child walk / leaf test are traversePool's, statement for
// statement (ray_fast.cuh); only the refill step knows about tasks.
template <int kLeafBatch, int kRefillBatch>
__global__ void __launch_bounds__(kStreamWarpsPerCTA * 32, MPENV_STREAM_MINB) k_lidar_stream(const __grid_constant__ KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    const Ctx &c = ka.ctx;
    const FastBVH bvh = stageFastBVH(c.bvh, smem_raw);
    const size_t bvh_bytes = fastBVHBytes(c.bvh.numNodes, c.bvh.numTris);
    LidarStreamPool *pools = reinterpret_cast<LidarStreamPool *>(reinterpret_cast<char *>(smem_raw) + bvh_bytes);
    const int warp = threadIdx.x >> 5;
    LidarStreamPool &P = pools[warp];
    const int num_tasks = (int)c.cfg.numWorlds * (kAgents / 2);

    // stream state lives in shared memory: it is touched only by the refill step, and the
    // traversal loop has no registers to spare (80 at 3 CTAs / SM)
    {
        int32_t prep = 0;
        for (; prep < 2; prep++) {
            const int task = fetchWarpTask(ka.workCounters);
            if (task >= num_tasks) break;
            lidarStreamSetup(c, P, prep, task);
            if ((threadIdx.x & 31) == 0) P.st[6 + 2 * prep] = task;
        }
        if ((threadIdx.x & 31) == 0) {
            P.st[0] = 0; P.st[1] = prep; P.st[2] = 0; P.st[3] = 0;
            P.st[4] = kLidarTaskRays; P.st[5] = kLidarTaskRays;
            P.st[7] = 0;
        }
        __syncwarp();
        if (prep == 0) {
            warpTasksDone(ka.workCounters, gridDim.x * kStreamWarpsPerCTA);
            return;                                      // no CTA-wide barrier below
        }
    }

    int32_t stack[32 + kLaneStateWords];     // traversal stack + the stash used around a task change
    int32_t sp = 0;
    int32_t ray = -1;                // the lane's ray | buffer of its task << 8
    bool can_issue = true;
    RayRegs r;
    r.tStop = -1.f;                  // closest hit wanted: never stop early
    float t_max = FLT_MAX;           // FLT_MAX = no accepted hit yet (no hit lies that far in a bounded map)
    uint32_t nd = 0;                 // the current node (shared-memory offset)
    float tn[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t rem = 0;
    int32_t leaf = -1;               // leaf waiting for its triangle tests: first triangle | count << 24
    const char *smem = reinterpret_cast<const char *>(smem_raw);
    const uint32_t trav_off = (uint32_t)(reinterpret_cast<const char *>(P.trav) - smem);

    for (;;) {
        const bool idle = leaf < 0 && rem == 0 && sp == 0;
        const unsigned idle_m = __ballot_sync(0xFFFFFFFFu, idle);
        const unsigned adv_m = __ballot_sync(0xFFFFFFFFu, !idle && leaf < 0);
        const unsigned fresh_m = __ballot_sync(0xFFFFFFFFu, idle && ray >= 0);
        if ((can_issue ? idle_m : fresh_m) != 0 && (__popc(idle_m) >= kRefillBatch || adv_m == 0)) {
            int32_t fin = P.st[0], prep = P.st[1], iss = P.st[2], next_ray = P.st[3];
            int32_t remaining0 = P.st[4], remaining1 = P.st[5];
            // retire finished rays
            const unsigned fresh1_m = __ballot_sync(0xFFFFFFFFu, idle && ray >= 256);
            if (idle && ray >= 0) P.tmesh[ray >> 8][ray & 255] = t_max != FLT_MAX ? t_max : -1.f;
            remaining1 -= __popc(fresh1_m);
            remaining0 -= __popc(fresh_m) - __popc(fresh1_m);
            if (idle) ray = -1;
            __syncwarp();
            // converged work of finished tasks; the freed buffer takes the task after next.  The
            // lanes' traversal state is parked in local memory across it (through an index the
            // compiler cannot resolve), so that the export / setup code does not compete for
            // registers with the traversal loop -- left to ptxas, tn[] / ch[] lived on the stack.
            if (fin < prep && ((fin & 1) ? remaining1 : remaining0) == 0) {
                {
                    int32_t *sv = stack + 32 + *reinterpret_cast<volatile int32_t *>(&P.st[7]);
                    MP_LANE_STATE(MP_STASH_F, MP_STASH_I)
                }
                do {
                    lidarStreamExport(c, P, fin & 1, P.st[6 + 2 * (fin & 1)]);
                    fin++;
                    // fin + 1 == prep, or the tasks have run out (then no further fetch: prep stays)
                    const int task = fin + 1 == prep ? fetchWarpTask(ka.workCounters) : num_tasks;
                    if (task < num_tasks) {
                        lidarStreamSetup(c, P, prep & 1, task);
                        if ((threadIdx.x & 31) == 0) P.st[6 + 2 * (prep & 1)] = task;
                        if (prep & 1) remaining1 = kLidarTaskRays; else remaining0 = kLidarTaskRays;
                        prep++;
                    }
                } while (fin < prep && ((fin & 1) ? remaining1 : remaining0) == 0);
                if (fin >= prep) break;                  // nothing resident and nothing left to fetch
                {
                    const int32_t *sv = stack + 32 + *reinterpret_cast<volatile int32_t *>(&P.st[7]);
                    MP_LANE_STATE(MP_UNSTASH_F, MP_UNSTASH_I)
                }
            }
            // hand out rays of the issuing task (ballot rank, no atomics); lanes ranked past its
            // end stay idle and are served from the next task on the next round
            if (iss < prep) {
                if (idle) {
                    const int32_t cand = next_ray + __popc(idle_m & ((1u << (threadIdx.x & 31)) - 1u));
                    if (cand < kLidarTaskRays) {
                        const int32_t b = iss & 1;
                        ray = cand | (b << 8);
                        int ag, rr, h, d;
                        lidarDecode(cand, &ag, &rr, &h, &d);
                        {
                            const float4 o = P.origin[b][ag * 2 + h];
                            const float4 tv = *reinterpret_cast<const float4 *>(&P.trav[b][d]);
                            setRay(r, o.x, o.y, o.z, tv.x, tv.y, tv.z, 0);
                        }
                        t_max = FLT_MAX;
                        stack[0] = 0;
                        sp = 1;
                    }
                }
                next_ray += __popc(idle_m);
                if (next_ray >= kLidarTaskRays) { next_ray = 0; iss++; }
            }
            can_issue = iss < prep;
            __syncwarp();
            if ((threadIdx.x & 31) == 0) {
                P.st[0] = fin; P.st[1] = prep; P.st[2] = iss; P.st[3] = next_ray;
                P.st[4] = remaining0; P.st[5] = remaining1;
            }
            __syncwarp();
        }
        if (leaf < 0 && rem == 0 && sp > 0) {
            nd = (uint32_t)stack[--sp] * (uint32_t)sizeof(FastNode);
            if (r.offX >= 0) {
                visitNode(smem + nd, r, ka.one, tn);
            } else {
                const float4 v = visitNodeRef(smem + nd, r.ox, r.oy, r.oz, r.invX, r.invY, r.invZ);
                tn[0] = v.x; tn[1] = v.y; tn[2] = v.z; tn[3] = v.w;
            }
            rem = 0xFu;
        }
        if (leaf < 0 && rem != 0) leaf = walkChildren(smem + nd, tn, t_max, rem, stack, sp);
        const unsigned pend = __ballot_sync(0xFFFFFFFFu, leaf >= 0);
        const unsigned can_advance = __ballot_sync(0xFFFFFFFFu, leaf < 0 && (rem != 0 || sp > 0));
        if (leaf >= 0 && (__popc(pend) >= kLeafBatch || can_advance == 0)) {
            int ag, rr, h, d;
            lidarDecode(ray & 255, &ag, &rr, &h, &d);
            r.tri = trav_off + (uint32_t)((ray >> 8) * 2 * kLidarDirsPerAgent + d) * (uint32_t)sizeof(RayTravRec) +
                    (uint32_t)offsetof(RayTravRec, tri);
            const TriConsts tc = loadTriConsts(smem, r);
            const FastTri *tri = bvh.tris + (leaf & 0xFFFFFF);
            for (int32_t k = 0; k < (leaf >> 24); k++) {
                float t;
                if (rayTriangleFast(tri + k, tc, t_max, &t)) t_max = t;
            }
            leaf = -1;
        }
    }
    warpTasksDone(ka.workCounters, gridDim.x * kStreamWarpsPerCTA);
}

inline size_t lidarStreamSmemBytes(uint32_t num_nodes, uint32_t num_tris)
{
    return fastBVHBytes(num_nodes, num_tris) + kStreamWarpsPerCTA * sizeof(LidarStreamPool);
}

}  // namespace mpenv
