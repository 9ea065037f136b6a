// world_coop.cuh -- 16-lane cooperative forms of the per-world systems whose
// inner loops run over the 12 agents or the breadcrumb pool (device only).
// Same results as the sequential forms in mp_stages_world.h, which stay the
// statement of the semantics (and what tests/hostsim runs):
//   * per-agent terms are computed by lane = agent with the same expressions;
//   * order-sensitive float sums keep their order (each agent's breadcrumb
//     penalty is summed over crumbs in pool order by that agent's lane);
//   * min / count reductions are order independent;
//   * sequential allocation (new crumbs, uids, stable compaction) is done by
//     ballot rank, which reproduces agent / pool order.
// Every function must be called by all 32 lanes of the warp (two worlds per
// warp, `half` selects the 16-lane group); `valid` masks a missing world.
#pragma once

#include "mp_stages.h"

namespace mpenv {

__device__ __forceinline__ unsigned groupBallot(bool pred, int half)
{
    return (__ballot_sync(0xFFFFFFFFu, pred) >> (half * 16)) & 0xFFFFu;
}

__device__ __forceinline__ float groupMin(float v)
{
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) v = fminf(v, __shfl_xor_sync(0xFFFFFFFFu, v, off, 16));
    return v;
}

// zoneSystem (sim.cpp:1892-1976): the OBB containment of the 12 agents in parallel.
__device__ __forceinline__ void zoneSystemCoop(const Ctx &c, int w, int lane, int half, bool valid)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    const bool agent = valid && lane < kAgents;
    if (valid && lane == 0) {
        ZoneState &zs = p.world[w].zone;
        if (zs.curControllingTeam != -1) zs.zoneStepsRemaining -= 1;
        if (zs.zoneStepsRemaining == 0) {
            zs.curZone += 1;
            if (zs.curZone == (int32_t)mc.numZones) zs.curZone = 0;
            zs.isCaptured = 0;
            zs.zoneStepsRemaining = kNumStepsPerZone;
            zs.stepsUntilPoint = kZonePointInterval;
            AABB zb = mc.zoneBoxes[zs.curZone];
            Vec3 zone_center = (zb.pMax + zb.pMin) / 2.f;
            for (int i = 0; i < kAgents; i++)
                p.combat[base + i].minDistToZone = distance(p.pos[base + i], zone_center);
        }
    }
    __syncwarp();
    bool in_zone = false;
    if (agent) {
        Quat to_zone; AABB zbox;
        zoneFrame(mc, p.world[w].zone.curZone, &to_zone, &zbox);
        Vec3 pos = p.pos[base + lane];
        pos.z += kStandHeight / 2.f;
        in_zone = aabbContains(zbox, rotateVec(to_zone, pos));
        p.combat[base + lane].inZone = in_zone ? 1 : 0;
    }
    const unsigned m = groupBallot(in_zone, half);
    if (valid && lane == 0) {
        ZoneState &zs = p.world[w].zone;
        const int num_a = __popc(m & 0x3Fu), num_b = __popc(m & 0xFC0u);
        zs.stepsUntilPoint -= 1;
        zs.isContested = (num_a > 0 && num_b > 0) ? 1 : 0;
        if (zs.isContested || (num_a == 0 && num_b == 0)) {
            zs.curControllingTeam = -1;
            zs.isCaptured = 0;
            zs.stepsUntilPoint = kZonePointInterval;
        } else if (num_a > 0 && num_b == 0) {
            if (zs.curControllingTeam != 0) {
                zs.curControllingTeam = 0; zs.isCaptured = 0; zs.stepsUntilPoint = kZonePointInterval;
            }
        } else if (num_a == 0 && num_b > 0) {
            if (zs.curControllingTeam != 1) {
                zs.curControllingTeam = 1; zs.isCaptured = 0; zs.stepsUntilPoint = kZonePointInterval;
            }
        }
    }
    __syncwarp();
}

// leaveBreadcrumbsSystem + accumulateBreadcrumbPenaltiesSystem (sim.cpp:4845-4926).
__device__ __forceinline__ void breadcrumbSystemsCoop(const Ctx &c, int w, int lane, int half, bool valid)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const bool agent = valid && lane < kAgents;
    Crumb *pool = p.crumbs + (size_t)(valid ? w : 0) * kMaxCrumbs;
    const int n0 = valid ? p.world[w].numCrumbs : 0;
    const uint32_t uid0 = valid ? p.world[w].nextCrumbUID : 0u;

    // -- leave: refresh the own last crumb, or ask for a new one
    bool want_new = false;
    Vec3 pos = v3(0.f, 0.f, 0.f);
    if (agent) {
        CrumbAgentState &as = p.crumbAgent[base + lane];
        pos = p.pos[base + lane];
        as.totalPenalty = 0.f;
        bool updated_last = false;
        if (as.lastCrumbUID != 0) {
            for (int k = 0; k < n0; k++) {
                if (pool[k].uid != as.lastCrumbUID) continue;
                if (distance(pos, pool[k].pos) < kAgentRadius * 4) {
                    pool[k].penalty = 1.f;
                    updated_last = true;
                    as.stepsSinceNew = 0;
                }
                break;
            }
        }
        if (!updated_last) {
            as.stepsSinceNew += 1;
            want_new = as.stepsSinceNew > 10;
        }
    }
    // sequential allocation in agent order: slot n0 + rank while the pool has room
    const unsigned want_m = groupBallot(want_new, half);
    const int rank = __popc(want_m & ((1u << lane) - 1u));
    const bool create = want_new && (n0 + rank < kMaxCrumbs);
    if (create) {
        Crumb cr; cr.pos = pos; cr.penalty = 1.f; cr.agent = lane; cr.uid = uid0 + (uint32_t)rank + 1u;
        pool[n0 + rank] = cr;
        CrumbAgentState &as = p.crumbAgent[base + lane];
        as.stepsSinceNew = 0;
        as.lastCrumbUID = cr.uid;
    }
    const int created = __popc(groupBallot(create, half));
    const int n = n0 + created;
    __syncwarp();

    // -- accumulate: each agent sums the penalties of its teammates' crumbs in pool order
    if (agent) {
        float total = 0.f;          // as.totalPenalty was zeroed above
        const int team = teamOf(lane);
        for (int k = 0; k < n; k++) {
            const Crumb cr = pool[k];
            if (teamOf(cr.agent) != team || cr.agent == lane) continue;
            if (distance(pos, cr.pos) <= kAgentRadius * 4.f) total += cr.penalty;
        }
        p.crumbAgent[base + lane].totalPenalty = total;
    }
    __syncwarp();

    // -- decay + stable compaction, 16 crumbs per pass
    int out = 0;
    // n is uniform within a group, not within the warp: both groups run the longer loop
    const int n_warp = max(n, __shfl_xor_sync(0xFFFFFFFFu, n, 16));
    for (int k0 = 0; k0 < n_warp; k0 += 16) {
        const int k = k0 + lane;
        Crumb cr;
        bool keep = false;
        if (k < n) {
            cr = pool[k];
            cr.penalty -= 0.025f;
            keep = !(cr.penalty <= 0.f);
        }
        const unsigned keep_m = groupBallot(keep, half);
        __syncwarp();
        if (keep) pool[out + __popc(keep_m & ((1u << lane) - 1u))] = cr;
        out += __popc(keep_m);
        __syncwarp();
    }
    if (valid && lane == 0) {
        p.world[w].numCrumbs = out;
        p.world[w].nextCrumbUID = uid0 + (uint32_t)created;
    }
    __syncwarp();
}

// evaluateGoalRegionsSystem (sim.cpp:4023-4087): the per-agent OBB distances in parallel.
__device__ __forceinline__ void evaluateGoalRegionsCoop(const Ctx &c, int w, int lane, int half, bool valid)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    const bool agent = valid && lane < kAgents;
    const int32_t attacker_team = valid ? p.world[w].match.teamA : 0;
    const Vec3 pos = agent ? p.pos[base + lane] : v3(0.f, 0.f, 0.f);
    float step_rewards[2] = { 0.f, 0.f };
    for (int r = 0; r < mc.numGoalRegions; r++) {
        const GoalRegion &gr = mc.goalRegions[r];
        const int32_t region_team = gr.attackerTeam ? attacker_team : (attacker_team ^ 1);
        float max_min = -FLT_MAX;
        for (int s = 0; s < gr.numSub; s++) {
            float d = FLT_MAX;
            if (agent && teamOf(lane) == region_team) d = distToZOBB(gr.sub[s], pos);
            const float min_d = groupMin(d);
            if (min_d > max_min) max_min = min_d;
        }
        if (valid && lane == 0) {
            WorldState &ws = p.world[w];
            const float prev = ws.goalMinDist[r];
            if (prev == FLT_MAX) {
                ws.goalMinDist[r] = max_min;
            } else {
                const float diff = prev - max_min;
                if (diff > 0.f) {
                    ws.goalMinDist[r] = max_min;
                    step_rewards[region_team] += diff * gr.rewardStrength;
                }
            }
        }
    }
    if (valid && lane == 0) {
        p.world[w].goalTeamStepRewards[0] = step_rewards[0];
        p.world[w].goalTeamStepRewards[1] = step_rewards[1];
    }
    __syncwarp();
}

}  // namespace mpenv
