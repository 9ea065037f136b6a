// mp_stages.h -- the per-step simulation stages of MadronaMPEnv (Task::Zone,
// 6v6) as plain functions over structure-of-component-arrays world state.
// The CUDA kernels in step_kernels.cu are thin thread-mapping wrappers around
// these; tests/hostsim compiles the same functions for the CPU test-suite.
//
// Stage order and semantics follow the reference Step task graph
// (reference src/sim.cpp:5342-5842); each function cites the system it
// restates.  Agents of world w are rows w*12 .. w*12+11; rows 0-5 are team 0,
// 6-11 team 1 (level_gen.cpp:195-207), so TeamInfo is implicit.
#pragma once

#include "mp_geom.h"

namespace mpenv {

struct Ctx {
    SimConfig cfg;
    const MapConsts *mc;
    BVHView bvh;
    SimPtrs p;
    NavView nav;             // numTris == 0: no navmesh loaded (bots stand still)
};

MP_HD int teamOf(int a) { return a / kTeamSize; }
MP_HD int offsetOf(int a) { return a % kTeamSize; }

// utils.hpp:36-54
MP_HD float viewHeight(int32_t cur_pose)
{
    float top = cur_pose == 0 ? kStandHeight : (cur_pose == 1 ? kCrouchHeight : kProneHeight);
    return top - kAgentRadius;
}

// utils.cpp:140-167
MP_HD Aim computeAim(float yaw, float pitch)
{
    if (yaw < -kPi) yaw += 2.f * kPi;
    else if (yaw > kPi) yaw -= 2.f * kPi;
    if (pitch < -0.25f * kPi) pitch = -0.25f * kPi;
    if (pitch > 0.25f * kPi) pitch = 0.25f * kPi;
    Quat r = quatMul(angleAxis(yaw, kUp()), angleAxis(pitch, kRight()));
    r = quatNormalize(r);
    Aim a; a.yaw = yaw; a.pitch = pitch; a.rot = r;
    return a;
}

MP_HD int32_t clampi(int32_t v, int32_t lo, int32_t hi) { return v < lo ? lo : (v > hi ? hi : v); }

// ============================================================ stage: move
// pvpMovementSystem (sim.cpp:2093-2199)
MP_HD void movementSystem(const DiscreteAction &action, Quat rot, Vec3 &vel, float alive,
                          CombatState &combat, StandState &stand, MoveState &move)
{
    if (alive == 0.f) return;
    {
        Vec3 v = vel;
        float v_len = length(v);
        if (v_len > 0.f) {
            Vec3 norm_v = v / v_len;
            v_len -= kDeaccelerateRate * kDeltaT;
            v_len = fmaxf(0.f, v_len);
            vel = norm_v * v_len;
        }
    }
    if (stand.transition > 0) {
        stand.transition -= 1;
        if (stand.transition == 0) stand.cur = stand.tgt;
    }
    int32_t action_pose = clampi(action.stand, 0, 2);
    if (action_pose != stand.tgt) {
        stand.tgt = action_pose;
        int dst = stand.tgt - stand.cur;
        if (dst < 0) dst = -dst;
        stand.transition = dst * (kPoseTransitionSpeed / 2);
    }
    int32_t discrete_move_amount = action.moveAmount;
    int32_t discrete_move_angle = action.moveAngle;
    float move_accel_max = 3000;
    if (stand.cur == 1) move_accel_max = 100;
    else if (stand.cur == 2) move_accel_max = 50;
    float move_amount = (float)discrete_move_amount *
        (move_accel_max / (float)(kNumMoveAmountBuckets - 1));
    constexpr float move_angle_per_bucket = 2.f * kPi / float(kNumMoveAngleBuckets);
    float move_angle = float(discrete_move_angle) * move_angle_per_bucket;
    float f_x = move_amount * mp_sinf(move_angle);
    float f_y = move_amount * mp_cosf(move_angle);
    vel = vel + rotateVec(rot, v3(f_x, f_y, 0)) * kDeltaT;
    if (move_amount != 0) combat.remainingRespawnSteps = 0;
    float v_len = length(vel);
    if (v_len == 0.f) return;
    {
        constexpr float max_vel_change_rate = 510.f;
        float tgt_max_vel;
        if (stand.cur == 0) tgt_max_vel = discrete_move_amount == 2 ? kMaxRunVelocity : kMaxWalkVelocity;
        else if (stand.cur == 1) tgt_max_vel = kMaxCrouchVelocity;
        else tgt_max_vel = kMaxProneVelocity;
        float max_vel_diff = tgt_max_vel - move.maxVel;
        float max_vel_adjust = fmaxf(fminf(max_vel_diff, max_vel_change_rate), -max_vel_change_rate);
        move.maxVel += max_vel_adjust;
    }
    Vec3 v_norm = vel / v_len;
    v_len = fminf(v_len, move.maxVel);
    vel = v_norm * v_len;
}

// pvpContinuousAimSystem (sim.cpp:2266-2282) then pvpDiscreteAimSystem
// (sim.cpp:2284-2334): both run every step on the same Aim.
MP_HD void aimSystems(AimAction cont, DiscreteAimAction disc, Quat &rot, Aim &aim, float alive)
{
    if (alive == 0.f) return;
    aim.yaw += cont.yaw * kDeltaT;
    aim.pitch += cont.pitch * kDeltaT;
    aim = computeAim(aim.yaw, aim.pitch);
    rot = quatNormalize(angleAxis(aim.yaw, kUp()));

    const float yaw_turn[7] = { 0, 0.00390625f * kPi, 0.0078125f * kPi, 0.015625f * kPi,
                                0.03125f * kPi, 0.0625f * kPi, 0.125f * kPi };
    const float pitch_turn[4] = { 0, 0.0078125f * kPi, 0.015625f * kPi, 0.03125f * kPi };
    int32_t yaw_bucket = clampi(disc.yaw, 0, kAimYawBuckets - 1) - kAimYawBuckets / 2;
    int32_t pitch_bucket = clampi(disc.pitch, 0, kAimPitchBuckets - 1) - kAimPitchBuckets / 2;
    if (yaw_bucket < 0) aim.yaw -= yaw_turn[-yaw_bucket];
    else aim.yaw += yaw_turn[yaw_bucket];
    if (pitch_bucket < 0) aim.pitch -= pitch_turn[-pitch_bucket];
    else aim.pitch += pitch_turn[pitch_bucket];
    aim = computeAim(aim.yaw, aim.pitch);
    rot = quatNormalize(angleAxis(aim.yaw, kUp()));
}

// Rotate2D (sim.cpp:874-879)
MP_HD Vec3 rotate2D(Vec3 dir, float radians)
{
    float c = mp_cosf(radians);
    float s = mp_sinf(radians);
    return v3(c * dir.x - s * dir.y, s * dir.x + c * dir.y, 0);
}

// applyVelocitySystem (sim.cpp:889-1028): swept-sphere character controller
// against the static mesh only (utils.cpp:75-128 has the agent capsules
// compiled out).
// BVH: BVHView (reference format) or SphereBVH (sphere_fast.cuh) -- bvhSphereCast is overloaded.
template <typename BVH>
MP_HD void applyVelocitySystem(const BVH &bvh, const MapConsts &mc, Vec3 pos, Vec3 vel, int32_t cur_pose,
                               MoveState &move)
{
    Vec3 x = pos;
    Vec3 v = vel;
    v.z = 0;
    move.newPos = x;
    move.newVel = v3(0, 0, 0);
    float v_len = length(v);
    if (v_len == 0.f) return;
    Vec3 v_norm = v / v_len;
    float move_dist = v_len * kDeltaT;

    const float buffer = 0.05f * kAgentRadius;
    const float r = kAgentRadius;
    float top_of_capsule = kStandHeight - r;
    float low_check_height = kProneHeight;
    if (cur_pose == 1) {
        top_of_capsule = kCrouchHeight - r;
    } else if (cur_pose == 2) {
        top_of_capsule = low_check_height;
        low_check_height = kProneHeight - r + buffer;
    }

    Vec3 ray_o = x;
    ray_o.z += top_of_capsule;
    // The reference leaves `normal` uninitialised when this cast misses
    // (sim.cpp:927-930); defined here as zero, which makes the gate false.
    Vec3 normal = v3(0, 0, 0);
    bvhSphereCast(bvh, ray_o, -kUp(), r, &normal);
    if (normal.z > 0.0f && (double)normal.z < 0.7 && dot(normal, v_norm) < 0.0f) return;

    // Everything below uses the forward casts only through fmaxf(0, t - buffer) compared
    // with / clamped to move_dist (and t == 0), so hits past move_dist + buffer are misses.
    // (Bounded only where that is exact: sphereSweepBelow, mp_geom.h.)
    const float fwd_bound = (move_dist + buffer) * 1.001f + 0.01f;
    ray_o = x + v_norm * buffer * 0.5f;
    ray_o.z += low_check_height;
    float low_dist = sphereSweepBelow(bvh, mc, ray_o, v_norm, r, &normal, fwd_bound);
    float high_dist = low_dist;
    bool high_hit = false;
    if (cur_pose != 2) {
        ray_o.z = x.z + top_of_capsule;
        Vec3 high_normal = v3(0, 0, 0);
        high_dist = sphereSweepBelow(bvh, mc, ray_o, v_norm, r, &high_normal, fwd_bound);
        if (high_dist < low_dist) {
            low_dist = high_dist;
            normal = high_normal;
            high_hit = true;
        }
    }
    bool stuck = low_dist == 0.0f || high_dist == 0.0f;
    low_dist = fmaxf(0.0f, low_dist - buffer);
    high_dist = fmaxf(0.0f, high_dist - buffer);
    Vec3 hit_pos = x + v_norm * fminf(low_dist, move_dist);

    if (move_dist > low_dist) {
        Vec3 slide_dir = normalize(cross(kUp(), normal));
        if (dot(slide_dir, v_norm) < 0) slide_dir = -slide_dir;
        ray_o = x + v_norm * low_dist;
        ray_o.z += high_hit ? top_of_capsule : low_check_height;
        Vec3 tmp_n = v3(0, 0, 0);
        float max_move = move_dist - low_dist;
        // clamped to max_move below: same argument
        float slide_dist = sphereSweepBelow(bvh, mc, ray_o, slide_dir, r, &tmp_n, (max_move + buffer) * 1.001f + 0.01f);
        slide_dist = fmaxf(0.0f, slide_dist - buffer);
        slide_dist = fminf(slide_dist, max_move);
        if (slide_dist > 0.0f) hit_pos = hit_pos + slide_dir * slide_dist;
    }

    Vec3 ground_check_pos = hit_pos;
    ground_check_pos.z += top_of_capsule;
    Vec3 tmp_n;
    float ground_dist = bvhSphereCast(bvh, ground_check_pos, -kUp(), r, &tmp_n);
    if (ground_dist == FLT_MAX) return;

    if (ground_dist <= 0.0f || stuck) {
        float furthest_hit = 0.0f;
        int best_dir = -1;
        for (int dir = 0; dir < 4; dir++) {
            Vec3 dir_vec = rotate2D(v_norm, (float)dir * 3.14159f * 0.5f);
            ray_o = x - dir_vec * r * 2.0f;
            ray_o.z += low_check_height;
            float hit_dist = bvhSphereCast(bvh, ray_o, dir_vec, r, &tmp_n);
            if (hit_dist > furthest_hit) { furthest_hit = hit_dist; best_dir = dir; }
        }
        if (best_dir != -1) {
            Vec3 dir_vec = rotate2D(v_norm, (float)best_dir * 3.14159f * 0.5f);
            hit_pos = x + dir_vec * (fminf(furthest_hit - r * 2.0f, -buffer));
            ground_check_pos = hit_pos;
            ground_check_pos.z += top_of_capsule;
            ground_dist = bvhSphereCast(bvh, ground_check_pos, -kUp(), r, &tmp_n);
            if (ground_dist == FLT_MAX) return;
        }
    }

    float fall_dist = fminf(ground_dist, top_of_capsule) + r;
    Vec3 new_pos = ground_check_pos;
    new_pos.z -= fall_dist;
    Vec3 to_new_pos = new_pos - x;
    float to_new_dist = length(to_new_pos);
    if (to_new_dist == 0.f) return;
    move.newPos = new_pos;
    move.newVel = to_new_pos / kDeltaT;
}

// fallSystem (sim.cpp:1041-1095): constant fall rate, one downward cast.
template <typename BVH>
MP_HD void fallSystem(const BVH &bvh, Vec3 pos, MoveState &move, float alive)
{
    if (alive == 0.f) { move.newPos = pos; return; }
    const float FALL_RATE = 386.08858267717f;
    const float cast_offset = kAgentRadius;
    Vec3 ray_o = pos;
    ray_o.z += kAgentRadius + cast_offset;
    Vec3 tmp_n;
    float ground_dist = bvhSphereCast(bvh, ray_o, -kUp(), kAgentRadius, &tmp_n);
    if (ground_dist == FLT_MAX || ground_dist < cast_offset) { move.newPos = pos; return; }
    float fall_dist = fminf(ground_dist - cast_offset, FALL_RATE * kDeltaT);
    Vec3 new_pos = pos;
    new_pos.z -= fall_dist;
    move.newPos = new_pos;
}

}  // namespace mpenv
#include "mp_stages_bot.h"
namespace mpenv {

// Per-agent fusion of Step-graph nodes 3-8 (sim.cpp:5403-5460): movement,
// both aim systems, applyVelocity, updateMoveState, fall, post-fall update.
// No cross-agent reads.
template <typename BVH>
MP_HD void stageMove(const Ctx &c, const BVH &bvh, int idx)
{
    const SimPtrs &p = c.p;
    float alive = p.alive[idx];
    Vec3 vel = p.vel[idx];
    Quat rot = p.rot[idx];
    CombatState combat = p.combat[idx];
    StandState stand = p.stand[idx];
    MoveState move = p.move[idx];
    Aim aim = p.aim[idx];
    Vec3 pos = p.pos[idx];

    stageBot(c, idx);            // policy id -1: plan + apply the scripted bot's action rows
    movementSystem(p.actDiscrete[idx], rot, vel, alive, combat, stand, move);
    aimSystems(p.actAim[idx], p.actDiscreteAim[idx], rot, aim, alive);
    applyVelocitySystem(bvh, *c.mc, pos, vel, stand.cur, move);
    pos = move.newPos;          // updateMoveStateSystem (sim.cpp:1030-1039)
    vel = move.newVel;
    fallSystem(bvh, pos, move, alive);
    pos = move.newPos;          // updateMoveStatePostFallSystem (sim.cpp:1097-1104)

    p.pos[idx] = pos;
    p.vel[idx] = vel;
    p.rot[idx] = rot;
    p.aim[idx] = aim;
    p.stand[idx] = stand;
    p.move[idx] = move;
    p.combat[idx].remainingRespawnSteps = combat.remainingRespawnSteps;
}

MP_HD void stageMove(const Ctx &c, int idx) { stageMove(c, c.bvh, idx); }

}  // namespace mpenv

#include "mp_stages_obs.h"
#include "mp_stages_world.h"
