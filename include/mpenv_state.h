/*
 * mpenv_state.h -- plain-C inspection record for one simulated world.
 *
 * Test/debug surface only: it is how tests diff the CUDA step against the
 * oracle (the reference's own sim.cpp run under oracle/shim).  Both sides
 * fill the same record so a memcmp-style field diff is meaningful.
 *
 * Field provenance (reference components, /root/reference/src/types.hpp):
 *   Position/Rotation (madrona base), AgentVelocity :465, IntermediateMoveState :471,
 *   Aim :455, HP :477, Alive :490, Magazine :481, DamageDealt :486, StandState :221,
 *   CombatState :510, BreadcrumbAgentState :772, StartPos :494, ExploreTracker :500,
 *   OpponentsVisibility :451, TeamInfo :270, AgentPolicy :266, RewardHyperParams :254,
 *   MatchInfo :127, ZoneState :563, MatchResult :246, ZoneStats :239,
 *   SpawnUsageCounter :95, GoalRegionsState :808, TeamRewardState :227,
 *   Breadcrumb :766; Sim members sim.hpp:102-193.
 */
#ifndef MPENV_STATE_H
#define MPENV_STATE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPENV_AGENTS_PER_WORLD 12
#define MPENV_TEAM_SIZE 6
#define MPENV_MAX_ZONES 5
#define MPENV_MAX_SPAWNS 128
#define MPENV_MAX_BREADCRUMBS 192

typedef struct mpenv_agent_state {
    float pos[3];
    float rot[4];            /* w x y z */
    float vel[3];
    float move_new_pos[3];
    float move_new_vel[3];
    float move_max_vel;
    float aim_yaw;
    float aim_pitch;
    float aim_rot[4];        /* w x y z */
    float hp;
    float alive;
    int32_t mag_bullets;
    int32_t mag_reloading;
    float dmg[MPENV_TEAM_SIZE];
    int32_t pose_cur;
    int32_t pose_tgt;
    int32_t pose_transition;
    uint32_t rng_key[2];
    uint32_t rng_count;
    int32_t landed_shot_on;          /* agent index in world, -1 = none */
    int32_t remaining_respawn_steps;
    int32_t remaining_autoheal_steps;
    int32_t successful_kill;
    int32_t was_shot_count;
    int32_t was_killed;
    float fired_shot_t;
    int32_t in_zone;
    float min_dist_to_zone;
    int32_t has_died;
    int32_t reloaded_full_mag;
    int32_t weapon_type;
    float crumb_total_penalty;
    int32_t crumb_steps_since_new;
    int32_t crumb_last;              /* index into breadcrumbs[], -1 = none/stale */
    float start_pos[3];
    uint32_t explore_new_cells;
    int32_t can_see[MPENV_TEAM_SIZE];
    int32_t team;
    int32_t team_offset;
    int32_t policy;
    float reward_coefs[9];
    int32_t in_sub_zone;             /* CombatState::inSubZone / minDistToSubZone (SimFlags::SubZones) */
    float min_dist_to_sub_zone;
} mpenv_agent_state;

typedef struct mpenv_breadcrumb_state {
    float pos[3];
    float penalty;
    int32_t team;
    int32_t team_offset;
} mpenv_breadcrumb_state;

typedef struct mpenv_world_state {
    /* MatchInfo */
    int32_t team_a;
    int32_t cur_step;
    int32_t is_finished;
    int32_t enable_spawn_curriculum;
    uint32_t cur_curriculum_tier;
    uint32_t cur_curriculum_spawn_idx;
    /* ZoneState */
    int32_t zone_cur;
    int32_t zone_controlling_team;
    int32_t zone_contested;
    int32_t zone_captured;
    int32_t zone_earned_point;
    int32_t zone_steps_remaining;
    int32_t zone_steps_until_point;
    /* MatchResult (exported, 30 x i32) */
    int32_t match_result[30];
    /* Sim::zoneStats (5 x {swaps, capA, capB, contested, active}) */
    int32_t zone_stats[MPENV_MAX_ZONES * 5];
    /* episode bookkeeping */
    uint32_t base_rng_key[2];
    uint32_t base_rng_count;
    uint32_t cur_episode_idx;
    uint32_t world_episode_counter;
    int32_t world_curriculum;
    int32_t reset_flag;
    /* filters */
    uint64_t filters_active[2];
    int32_t filters_last_matches[2][3];
    int32_t filters_last_matched_step[2];
    /* goal regions / team reward */
    float goal_min_dist[2];
    float goal_team_step_rewards[2];
    float team_rewards[2];
    /* SubZone x 8: curControllingTeam, isContested, isCaptured (types.hpp:791-796) */
    int32_t sub_zones[8][3];
    /* SpawnUsageCounter */
    uint32_t spawn_a_last_used[MPENV_MAX_SPAWNS];
    uint32_t spawn_b_last_used[MPENV_MAX_SPAWNS];
    uint32_t respawn_last_used[MPENV_MAX_SPAWNS];
    /* breadcrumbs, table order */
    int32_t num_breadcrumbs;
    mpenv_breadcrumb_state breadcrumbs[MPENV_MAX_BREADCRUMBS];
    mpenv_agent_state agents[MPENV_AGENTS_PER_WORLD];
} mpenv_world_state;

#ifdef __cplusplus
}
#endif

#endif
