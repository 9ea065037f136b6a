// lidar_packet.cuh -- pvpLidarSystem (sim.cpp:3324-3506) as warp-wide ray PACKETS
// walking the BVH together in the reference's visiting order.
//
// The reference's traversal order does not depend on the ray: children are taken
// in slot order, leaves are tested inline, internal children are pushed and popped
// LIFO (mesh_bvh.inl:142-201).  So 32 rays can share ONE stack of (node, lane mask)
// entries: a node is visited once for all the lanes whose own traversal would visit
// it, each lane applies its own t_near <= t_max test at each child's turn, a leaf's
// triangles are tested by the lanes that hit the leaf, and an internal child is
// pushed with the mask of the lanes that hit it.  Every lane sees exactly the nodes
// and triangles, in exactly the order, of its own reference traversal -- results are
// bit-identical -- while control flow is warp-uniform: no per-lane stack in local
// memory, no parking / refill ballots, no walk state machine, node headers and
// children read at one address by all lanes.  A lane keeps ONE ray for the whole
// packet, so its triangle constants live in registers.
//
// The price is lane utilisation on nodes that only some lanes visit.  It pays for
// coherent packets: the forward fan is split into 16 columns x 2 heights per packet
// (measured on rollout states, tools/exp/exp_packet.py: 9.3 - 10.2 node visits and
// 10 - 11 triangle tests per 32 rays -- the same warp-level counts as the pooled
// loop, at less than half its instructions per visit).  The rear fans (8 columns
// over 180 degrees x 2 heights) are incoherent; two agents' rear fans share a packet.
#pragma once

#include "ray_kernels.cuh"

namespace mpenv {

#ifndef MPENV_PACKET_WARPS
#define MPENV_PACKET_WARPS 16
#define MPENV_PACKET_MINB 2
#endif
constexpr int kPacketWarpsPerCTA = MPENV_PACKET_WARPS;
constexpr int kPacketStack = 48;         // > 3 pending children per level of any tree the engine accepts

struct __align__(16) LidarPacketPool {
    CapsuleRec caps[4][kAgents];         // [agent * 2 + height][capsule]
    uint2 stack[kPacketStack];           // (node, mask of the lanes that visit it)
};

// The packet traversal: every lane enters with its ray in `r` / `tc`; returns the
// lane's closest mesh hit in t_max (FLT_MAX = none).
__device__ __forceinline__ float tracePacket(const char *smem, const FastBVH &bvh, uint2 *stack, const RayRegs &r,
                                             const TriConsts &tc, float one)
{
    const int lane = threadIdx.x & 31;
    float t_max = FLT_MAX;
    int sp = 1;                          // warp-uniform
    if (lane == 0) stack[0] = make_uint2(0u, 0xFFFFFFFFu);
    __syncwarp();
    while (sp > 0) {
        const uint2 e = stack[--sp];     // one address for the warp: a broadcast
        const char *nd = smem + e.x * (uint32_t)sizeof(FastNode);
        const bool active = (e.y >> lane) & 1u;
        float tn[4];
        if (r.offX >= 0) {
            visitNode(nd, r, one, tn);
        } else {
            const float4 v = visitNodeRef(nd, r.ox, r.oy, r.oz, r.invX, r.invY, r.invZ);
            tn[0] = v.x; tn[1] = v.y; tn[2] = v.z; tn[3] = v.w;
        }
        const uint32_t meta = reinterpret_cast<const uint32_t *>(nd)[3];
        const int4 ch = reinterpret_cast<const int4 *>(nd)[2];
        const int32_t child[4] = { ch.x, ch.y, ch.z, ch.w };
        // a lane that does not visit this node compares against -1: no child is ever hit
        float t_eff = active ? t_max : -1.f;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            // the lane's own test at this child's turn (t_max may have shrunk at an earlier leaf)
            const bool hit = tn[i] <= t_eff;
            const unsigned m = __ballot_sync(0xFFFFFFFFu, hit);
            if (m == 0) continue;
            if ((meta >> i) & 1u) {
                // traceRayLeaf (mesh_bvh.inl:360-431): a later triangle replaces an earlier one
                // whenever T <= t_cur * det
                const FastTri *tri = bvh.tris + (child[i] & 0x7FFFFFFF);
                const int ntri = (int)((meta >> (4 + 7 * i)) & 0x7Fu);
                if (hit) {
                    for (int k = 0; k < ntri; k++) {
                        float t;
                        if (rayTriangleFast(tri + k, tc, t_max, &t)) t_max = t;
                    }
                    t_eff = t_max;
                }
            } else {
                if (lane == 0) stack[sp] = make_uint2((uint32_t)child[i], m);
                sp++;
            }
        }
        __syncwarp();
    }
    return t_max;
}

// <16 warps, 2 CTAs per SM> while two copies of the BVH fit in an SM's shared memory, else <32 warps, 1>.
template <int kWarps, int kMinB>
__global__ void __launch_bounds__(kWarps * 32, kMinB) k_lidar_packet(const __grid_constant__ KernelArgs ka)
{
    extern __shared__ uint4 smem_raw[];
    const Ctx &c = ka.ctx;
    const FastBVH bvh = stageFastBVH(ka.fastImage, c.bvh.numNodes, c.bvh.numTris, smem_raw);
    const size_t bvh_bytes = fastBVHBytes(c.bvh.numNodes, c.bvh.numTris);
    const char *smem = reinterpret_cast<const char *>(smem_raw);
    LidarPacketPool &P = reinterpret_cast<LidarPacketPool *>(reinterpret_cast<char *>(smem_raw) + bvh_bytes)[threadIdx.x >> 5];
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int lane = threadIdx.x & 31;
    const int num_tasks = (int)c.cfg.numWorlds * (kAgents / 2);

    for (;;) {
        const int task = fetchWarpTask(ka.workCounters);
        if (task >= num_tasks) break;
        const int idx0 = task * 2;                       // first agent's global index
        const int base = (idx0 / kAgents) * kAgents;     // the world's first agent
        if (lane == 0) MP_COUNT_RAYS(0, kLidarTaskRays);

        // capsule records of the task's four origins (2 agents x 2 heights) against the world's 12 agents
        for (int e = lane; e < 4 * kAgents; e += 32) {
            const int oi = e / kAgents, i = e - oi * kAgents;
            const int idx = idx0 + (oi >> 1);
            const float top_height = viewHeight(p.stand[idx].cur) + kAgentRadius;
            const float range = top_height - 2.f * kAgentRadius;
            Vec3 o = p.pos[idx];
            o.z += kAgentRadius + range * (float(oi & 1) / float(kFwdLidarH - 1));
            P.caps[oi][i] = makeCapsuleRec(o, p.pos[base + i]);
        }
        __syncwarp();

        // packets 0..3: one agent's forward fan, 16 columns x 2 heights; packet 4: both agents' rear fans
#pragma unroll 1
        for (int pk = 0; pk < 5; pk++) {
            int ag, h, rr;
            bool is_fwd;
            int col;
            if (pk < 4) {
                ag = pk >> 1; h = lane >> 4; col = (pk & 1) * 16 + (lane & 15);
                rr = h * kFwdLidarW + col; is_fwd = true;
            } else {
                ag = lane >> 4; h = (lane >> 3) & 1; col = lane & 7;
                rr = kFwdLidarH * kFwdLidarW + h * kRearLidarW + col; is_fwd = false;
            }
            const int idx = idx0 + ag, a = idx - base;
            // sim.cpp:3402-3470: origin (two sample heights between foot + r and eye level) and direction
            const float top_height = viewHeight(p.stand[idx].cur) + kAgentRadius;
            const float range = top_height - 2.f * kAgentRadius;
            Vec3 o = p.pos[idx];
            o.z += kAgentRadius + range * (float(h) / float(kFwdLidarH - 1));
            const Quat q = is_fwd ? p.aim[idx].rot : p.rot[idx];
            const Vec3 fwd = rotateVec(q, kFwd());
            const Vec3 right = rotateVec(q, kRight());
            const float x = is_fwd ? mc.fwdLidarX[col] : mc.rearLidarX[col];
            const float y = is_fwd ? mc.fwdLidarY[col] : mc.rearLidarY[col];
            const Vec3 d = normalize(x * right + y * fwd);
            const RayDirRec rec = makeRayDirRec(d);
            RayRegs r;
            setRay(r, o.x, o.y, o.z, rec.invX, rec.invY, rec.invZ, 0);
            r.tStop = -1.f;
            TriConsts tc;
            tc.Sx = rec.tri.Sx; tc.Sy = rec.tri.Sy; tc.Sz = rec.tri.Sz;
            tc.kx = rec.tri.kperm & 3; tc.ky = (rec.tri.kperm >> 2) & 3; tc.kz = (rec.tri.kperm >> 4) & 3;
            tc.okx = sel3(tc.kx, o.x, o.y, o.z);
            tc.oky = sel3(tc.ky, o.x, o.y, o.z);
            tc.okz = sel3(tc.kz, o.x, o.y, o.z);

            const float tm = tracePacket(smem, bvh, P.stack, r, tc, ka.one);

            // agent capsules, classification, export (utils.cpp:39-61, sim.cpp:3472-3500)
            float hit_t;
            int32_t hit_agent;
            const bool hit = closestWithCapsules(P.caps[ag * 2 + h], tm != FLT_MAX, tm, d.x, d.y, d.z, rec.a2, rec.dd,
                                                 &hit_t, &hit_agent);
            float4 v;
            if (hit) {
                v.x = fminf(hit_t, mc.maxDist);
                v.y = hit_agent < 0 ? 1.f : 0.f;
                v.z = (hit_agent >= 0 && teamOf(hit_agent) == teamOf(a)) ? 1.f : 0.f;
                v.w = (hit_agent >= 0 && teamOf(hit_agent) != teamOf(a)) ? 1.f : 0.f;
            } else {
                v = make_float4(-1.f, 0.f, 0.f, 0.f);
            }
            if (is_fwd)
                reinterpret_cast<float4 *>(p.fwdLidar)[(size_t)idx * (kFwdLidarH * kFwdLidarW) + rr] = v;
            else
                reinterpret_cast<float4 *>(p.rearLidar)[(size_t)idx * (kRearLidarH * kRearLidarW) +
                                                        (rr - kFwdLidarH * kFwdLidarW)] = v;
        }
        __syncwarp();
    }
    warpTasksDone(ka.workCounters, gridDim.x * kWarps);
}

inline size_t lidarPacketSmemBytes(uint32_t num_nodes, uint32_t num_tris, int warps)
{
    return fastBVHBytes(num_nodes, num_tris) + (size_t)warps * sizeof(LidarPacketPool);
}

}  // namespace mpenv
