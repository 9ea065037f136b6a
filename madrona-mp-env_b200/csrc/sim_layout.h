// sim_layout.h -- the table of state / export buffers of one simulator
// instance (row size and row count of every SimPtrs member), the export-id ->
// tensor mapping (reference ExportID, src/sim.hpp:15-66; shapes and dtypes
// src/mgr.cpp:1965-2381) and the world-state inspection record filler.
//
// Host-only helpers; the allocator is supplied by the caller (cudaMalloc in
// the product library, calloc in tests/hostsim).
#pragma once

#include <stddef.h>
#include <string.h>

#include "mp_defs.h"
#include "../../include/mpenv_state.h"

namespace mpenv {

enum RowKind { kPerAgent, kPerWorld, kPerTeam, kGlobal };

struct BufferDesc {
    const char *name;
    size_t ptrOffset;       // offset of the pointer inside SimPtrs
    size_t rowBytes;
    RowKind kind;
};

#define MP_BUF(member, bytes, kind) { #member, offsetof(SimPtrs, member), (size_t)(bytes), kind }

inline const BufferDesc *bufferTable(int *count)
{
    static const BufferDesc table[] = {
        MP_BUF(pos, sizeof(Vec3), kPerAgent),
        MP_BUF(rot, sizeof(Quat), kPerAgent),
        MP_BUF(vel, sizeof(Vec3), kPerAgent),
        MP_BUF(move, sizeof(MoveState), kPerAgent),
        MP_BUF(aim, sizeof(Aim), kPerAgent),
        MP_BUF(hp, 4, kPerAgent),
        MP_BUF(alive, 4, kPerAgent),
        MP_BUF(mag, sizeof(Magazine), kPerAgent),
        MP_BUF(dmg, 4 * kTeamSize, kPerAgent),
        MP_BUF(stand, sizeof(StandState), kPerAgent),
        MP_BUF(combat, sizeof(CombatState), kPerAgent),
        MP_BUF(crumbAgent, sizeof(CrumbAgentState), kPerAgent),
        MP_BUF(subZoneAgent, sizeof(SubZoneAgentState), kPerAgent),
        MP_BUF(startPos, sizeof(Vec3), kPerAgent),
        MP_BUF(exploreGrid, 4 * kExploreGridW * kExploreGridW, kPerAgent),
        MP_BUF(exploreNew, 4, kPerAgent),
        MP_BUF(canSee, 8, kPerAgent),
        MP_BUF(policy, 4, kPerAgent),
        MP_BUF(rewardCoefs, sizeof(RewardHyperParams), kPerAgent),
        MP_BUF(actDiscrete, sizeof(DiscreteAction), kPerAgent),
        MP_BUF(actAim, sizeof(AimAction), kPerAgent),
        MP_BUF(actDiscreteAim, sizeof(DiscreteAimAction), kPerAgent),
        MP_BUF(reward, 4, kPerAgent),
        MP_BUF(done, 4, kPerAgent),
        MP_BUF(selfObs, sizeof(SelfObs), kPerAgent),
        MP_BUF(teammateObs, sizeof(TeammateObs) * (kTeamSize - 1), kPerAgent),
        MP_BUF(opponentObs, sizeof(OpponentObs) * kTeamSize, kPerAgent),
        MP_BUF(lastKnownObs, sizeof(OpponentObs) * kTeamSize, kPerAgent),
        MP_BUF(selfPos, sizeof(NormPosObs), kPerAgent),
        MP_BUF(teammatePos, sizeof(NormPosObs) * (kTeamSize - 1), kPerAgent),
        MP_BUF(opponentPos, sizeof(NormPosObs) * kTeamSize, kPerAgent),
        MP_BUF(lastKnownPos, sizeof(NormPosObs) * kTeamSize, kPerAgent),
        MP_BUF(masks, 4 * kTeamSize, kPerAgent),
        MP_BUF(filtersObs, 4, kPerAgent),
        MP_BUF(fwdLidar, sizeof(LidarSample) * kFwdLidarH * kFwdLidarW, kPerAgent),
        MP_BUF(rearLidar, sizeof(LidarSample) * kRearLidarH * kRearLidarW, kPerAgent),
        MP_BUF(world, sizeof(WorldState), kPerWorld),
        MP_BUF(spawnUsage, sizeof(SpawnUsage), kPerWorld),
        MP_BUF(crumbs, sizeof(Crumb) * kMaxCrumbs, kPerWorld),
        MP_BUF(matchResult, sizeof(MatchResult), kPerWorld),
        MP_BUF(resetFlag, 4, kPerWorld),
        MP_BUF(worldCurriculum, 4, kPerWorld),
        MP_BUF(subZones, sizeof(SubZoneState) * kNumSubZones, kPerWorld),
        MP_BUF(ftGlobal, sizeof(FullTeamGlobalObs), kPerTeam),
        MP_BUF(ftPlayers, sizeof(FullTeamPlayerObs) * kTeamSize, kPerTeam),
        MP_BUF(ftEnemies, sizeof(FullTeamEnemyObs) * kTeamSize, kPerTeam),
        MP_BUF(ftLastKnown, sizeof(FullTeamCommonObs) * kTeamSize, kPerTeam),
        MP_BUF(ftFwdLidar, sizeof(LidarSample) * kFwdLidarH * kFwdLidarW * kTeamSize, kPerTeam),
        MP_BUF(ftRearLidar, sizeof(LidarSample) * kRearLidarH * kRearLidarW * kTeamSize, kPerTeam),
        MP_BUF(ftReward, 4, kPerTeam),
        MP_BUF(ftDone, 4, kPerTeam),
        MP_BUF(evAgent, sizeof(GameEvent) * kEventSlotsPerAgent, kPerAgent),
        MP_BUF(evWorld, sizeof(GameEvent), kPerWorld),
        MP_BUF(stepSnap, sizeof(PackedStepSnapshot), kPerWorld),
        MP_BUF(recordLog, sizeof(StepLog), kPerWorld),
        MP_BUF(replayLog, sizeof(StepLog), kPerWorld),
        MP_BUF(trainCtrl, sizeof(TrainControl), kGlobal),
    };
    *count = (int)(sizeof(table) / sizeof(table[0]));
    return table;
}
#undef MP_BUF

inline size_t bufferRows(const BufferDesc &d, uint32_t num_worlds)
{
    switch (d.kind) {
    case kPerAgent: return (size_t)num_worlds * kAgents;
    case kPerWorld: return num_worlds;
    case kPerTeam: return (size_t)num_worlds * 2;
    default: return 1;
    }
}

inline void **bufferSlot(SimPtrs *p, const BufferDesc &d) { return (void **)((char *)p + d.ptrOffset); }

// ---- exported tensors -------------------------------------------------------
// ids = reference ExportID (sim.hpp:15-66); kExportSimControl is the extra
// simControlTensor (mgr.cpp:287-293).
enum ExportId : uint32_t {
    kExReset = 0, kExWorldCurriculum, kExExploreAction, kExPvPDiscreteAction, kExPvPAimAction,
    kExPvPDiscreteAimAction, kExReward, kExDone, kExMatchResult, kExAgentPolicy,
    kExSelfObservation, kExTeammateObservations, kExOpponentObservations,
    kExOpponentLastKnownObservations, kExSelfPositionObservation,
    kExTeammatePositionObservations, kExOpponentPositionObservations,
    kExOpponentLastKnownPositionObservations, kExOpponentMasks, kExFwdLidar, kExRearLidar,
    kExAgentMap, kExUnmaskedAgentMap, kExHP, kExAlive, kExMagazine, kExFullTeamActions,
    kExFullTeamGlobal, kExFullTeamPlayers, kExFullTeamEnemies, kExFullTeamLastKnownEnemies,
    kExFullTeamFwdLidar, kExFullTeamRearLidar, kExFullTeamReward, kExFullTeamDone,
    kExFullTeamPolicyAssignments, kExEventLog, kExPackedStepSnapshot,
    kExFiltersStateObservation, kExRewardHyperParams, kExNumExports,
    kExportSimControl = 1000,
};

enum DType : int32_t { kDTypeI32 = 0, kDTypeF32 = 1 };

struct ExportDesc {
    void *ptr;
    int32_t dtype;
    int32_t ndim;
    int64_t dims[4];
};

// Shapes as in mgr.cpp:1965-2381 (N = numWorlds * 12, W = numWorlds).
inline bool exportDesc(const SimPtrs &p, uint32_t num_worlds, uint32_t id, ExportDesc *out)
{
    const int64_t N = (int64_t)num_worlds * kAgents, W = num_worlds, T = (int64_t)num_worlds * 2;
    auto set = [&](void *ptr, int32_t dt, int64_t d0, int64_t d1, int64_t d2 = -1, int64_t d3 = -1) {
        out->ptr = ptr; out->dtype = dt; out->dims[0] = d0; out->dims[1] = d1;
        out->dims[2] = d2 < 0 ? 0 : d2; out->dims[3] = d3 < 0 ? 0 : d3;
        out->ndim = d3 >= 0 ? 4 : (d2 >= 0 ? 3 : 2);
        return true;
    };
    switch (id) {
    case kExReset: return set(p.resetFlag, kDTypeI32, W, 1);
    case kExWorldCurriculum: return set(p.worldCurriculum, kDTypeI32, W, 1);
    case kExPvPDiscreteAction: return set(p.actDiscrete, kDTypeI32, N, 4);
    case kExPvPAimAction: return set(p.actAim, kDTypeF32, N, 1, 2);
    case kExPvPDiscreteAimAction: return set(p.actDiscreteAim, kDTypeI32, N, 2);
    case kExReward: return set(p.reward, kDTypeF32, N, 1);
    case kExDone: return set(p.done, kDTypeI32, N, 1);
    case kExMatchResult: return set(p.matchResult, kDTypeI32, W, 30);
    case kExAgentPolicy: return set(p.policy, kDTypeI32, N, 1);
    case kExSelfObservation: return set(p.selfObs, kDTypeF32, N, 43);
    case kExTeammateObservations: return set(p.teammateObs, kDTypeF32, N, kTeamSize - 1, 32);
    case kExOpponentObservations: return set(p.opponentObs, kDTypeF32, N, kTeamSize, 32);
    case kExOpponentLastKnownObservations: return set(p.lastKnownObs, kDTypeF32, N, kTeamSize, 32);
    case kExSelfPositionObservation: return set(p.selfPos, kDTypeF32, N, 3);
    case kExTeammatePositionObservations: return set(p.teammatePos, kDTypeF32, N, kTeamSize - 1, 3);
    case kExOpponentPositionObservations: return set(p.opponentPos, kDTypeF32, N, kTeamSize, 3);
    case kExOpponentLastKnownPositionObservations: return set(p.lastKnownPos, kDTypeF32, N, kTeamSize, 3);
    case kExOpponentMasks: return set(p.masks, kDTypeF32, N, kTeamSize, 1);
    case kExFwdLidar: return set(p.fwdLidar, kDTypeF32, N, kFwdLidarH, kFwdLidarW, 4);
    case kExRearLidar: return set(p.rearLidar, kDTypeF32, N, kRearLidarH, kRearLidarW, 4);
    case kExHP: return set(p.hp, kDTypeF32, N, 1);
    case kExAlive: return set(p.alive, kDTypeF32, N, 1);
    case kExMagazine: return set(p.mag, kDTypeI32, N, 2);
    case kExFullTeamGlobal: return set(p.ftGlobal, kDTypeF32, T, 16);
    case kExFullTeamPlayers: return set(p.ftPlayers, kDTypeF32, T, kTeamSize, 28);
    case kExFullTeamEnemies: return set(p.ftEnemies, kDTypeF32, T, kTeamSize, 33);
    case kExFullTeamLastKnownEnemies: return set(p.ftLastKnown, kDTypeF32, T, kTeamSize, 24);
    case kExFullTeamFwdLidar: return set(p.ftFwdLidar, kDTypeF32, T, kTeamSize * kFwdLidarH, kFwdLidarW, 4);
    case kExFullTeamRearLidar: return set(p.ftRearLidar, kDTypeF32, T, kTeamSize * kRearLidarH, kRearLidarW, 4);
    case kExFullTeamReward: return set(p.ftReward, kDTypeF32, T, 1);
    case kExFullTeamDone: return set(p.ftDone, kDTypeI32, T, 1);
    case kExFiltersStateObservation: return set(p.filtersObs, kDTypeF32, N, 1);
    case kExRewardHyperParams: return set(p.rewardCoefs, kDTypeF32, N, 9);
    case kExportSimControl: return set(p.trainCtrl, kDTypeI32, 3, 1);
    default: return false;
    }
}

// ---- inspection -------------------------------------------------------------
// Fills the plain-C inspection record from HOST copies of one world's rows:
// `p` points at arrays holding just that world (world index 0).
inline void fillWorldState(const SimPtrs &p, mpenv_world_state *out)
{
    memset(out, 0, sizeof(*out));
    const WorldState &ws = p.world[0];
    out->team_a = ws.match.teamA;
    out->cur_step = ws.match.curStep;
    out->is_finished = ws.match.isFinished;
    out->enable_spawn_curriculum = ws.match.enableSpawnCurriculum;
    out->cur_curriculum_tier = ws.match.curCurriculumTier;
    out->cur_curriculum_spawn_idx = ws.match.curCurriculumSpawnIdx;
    out->zone_cur = ws.zone.curZone;
    out->zone_controlling_team = ws.zone.curControllingTeam;
    out->zone_contested = ws.zone.isContested;
    out->zone_captured = ws.zone.isCaptured;
    out->zone_earned_point = ws.zone.earnedPoint;
    out->zone_steps_remaining = ws.zone.zoneStepsRemaining;
    out->zone_steps_until_point = ws.zone.stepsUntilPoint;
    memcpy(out->match_result, &p.matchResult[0], sizeof(MatchResult));
    memcpy(out->zone_stats, ws.zoneStats, sizeof(ws.zoneStats));
    out->base_rng_key[0] = ws.baseRNG.k.a;
    out->base_rng_key[1] = ws.baseRNG.k.b;
    out->base_rng_count = ws.baseRNG.count;
    out->cur_episode_idx = ws.curEpisodeIdx;
    out->world_episode_counter = ws.worldEpisodeCounter;
    out->world_curriculum = p.worldCurriculum[0];
    out->reset_flag = p.resetFlag[0];
    for (int t = 0; t < 2; t++) {
        out->filters_active[t] = ws.filtersActive[t];
        for (int f = 0; f < 3; f++) out->filters_last_matches[t][f] = ws.filtersLastMatches[t][f];
        out->filters_last_matched_step[t] = ws.filtersLastMatchedStep[t];
    }
    out->goal_min_dist[0] = ws.goalMinDist[0];
    out->goal_min_dist[1] = ws.goalMinDist[1];
    out->goal_team_step_rewards[0] = ws.goalTeamStepRewards[0];
    out->goal_team_step_rewards[1] = ws.goalTeamStepRewards[1];
    out->team_rewards[0] = ws.teamRewards[0];
    out->team_rewards[1] = ws.teamRewards[1];
    for (int z = 0; z < kNumSubZones; z++) {
        out->sub_zones[z][0] = p.subZones[z].curControllingTeam;
        out->sub_zones[z][1] = p.subZones[z].isContested;
        out->sub_zones[z][2] = p.subZones[z].isCaptured;
    }
    memcpy(out->spawn_a_last_used, p.spawnUsage[0].a, sizeof(uint32_t) * kMaxSpawns);
    memcpy(out->spawn_b_last_used, p.spawnUsage[0].b, sizeof(uint32_t) * kMaxSpawns);
    memcpy(out->respawn_last_used, p.spawnUsage[0].respawn, sizeof(uint32_t) * kMaxSpawns);
    out->num_breadcrumbs = ws.numCrumbs;
    for (int i = 0; i < ws.numCrumbs && i < MPENV_MAX_BREADCRUMBS; i++) {
        const Crumb &c = p.crumbs[i];
        out->breadcrumbs[i].pos[0] = c.pos.x; out->breadcrumbs[i].pos[1] = c.pos.y;
        out->breadcrumbs[i].pos[2] = c.pos.z;
        out->breadcrumbs[i].penalty = c.penalty;
        out->breadcrumbs[i].team = c.agent / kTeamSize;
        out->breadcrumbs[i].team_offset = c.agent % kTeamSize;
    }
    for (int i = 0; i < kAgents; i++) {
        mpenv_agent_state &a = out->agents[i];
        memcpy(a.pos, &p.pos[i], 12);
        memcpy(a.rot, &p.rot[i], 16);
        memcpy(a.vel, &p.vel[i], 12);
        memcpy(a.move_new_pos, &p.move[i].newPos, 12);
        memcpy(a.move_new_vel, &p.move[i].newVel, 12);
        a.move_max_vel = p.move[i].maxVel;
        a.aim_yaw = p.aim[i].yaw; a.aim_pitch = p.aim[i].pitch;
        memcpy(a.aim_rot, &p.aim[i].rot, 16);
        a.hp = p.hp[i]; a.alive = p.alive[i];
        a.mag_bullets = p.mag[i].numBullets; a.mag_reloading = p.mag[i].isReloading;
        for (int j = 0; j < kTeamSize; j++) a.dmg[j] = p.dmg[i * kTeamSize + j];
        a.pose_cur = p.stand[i].cur; a.pose_tgt = p.stand[i].tgt; a.pose_transition = p.stand[i].transition;
        const CombatState &cs = p.combat[i];
        a.rng_key[0] = cs.rng.k.a; a.rng_key[1] = cs.rng.k.b; a.rng_count = cs.rng.count;
        a.landed_shot_on = cs.landedShotOn;
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
        a.in_sub_zone = p.subZoneAgent[i].inSubZone;
        a.min_dist_to_sub_zone = p.subZoneAgent[i].minDistToSubZone;
        a.crumb_total_penalty = p.crumbAgent[i].totalPenalty;
        a.crumb_steps_since_new = p.crumbAgent[i].stepsSinceNew;
        a.crumb_last = -1;
        if (p.crumbAgent[i].lastCrumbUID != 0) {
            for (int k = 0; k < ws.numCrumbs; k++) {
                if (p.crumbs[k].uid == p.crumbAgent[i].lastCrumbUID) { a.crumb_last = k; break; }
            }
        }
        memcpy(a.start_pos, &p.startPos[i], 12);
        a.explore_new_cells = p.exploreNew[i];
        for (int j = 0; j < kTeamSize; j++) a.can_see[j] = p.canSee[i * 8 + j];
        a.team = i / kTeamSize;
        a.team_offset = i % kTeamSize;
        a.policy = p.policy[i];
        memcpy(a.reward_coefs, &p.rewardCoefs[i], 36);
    }
}

}  // namespace mpenv
