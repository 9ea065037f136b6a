// mp_stages_bot.h -- the scripted A* bot (agents whose policy id is -1;
// included by mp_stages.h).  Reference: planAStarAISystem sim.cpp:5041-5172 with
// NavUtils::{CenterOfTri, NearestNavTri, PathfindToPoint} sim.cpp:4958-5037, and
// applyBotActionsSystem sim.cpp:2057-2091.  In the reference the plan is the
// first node of the Step graph (it reads the PREVIOUS step's lidar / visibility)
// and the apply node opens pvpGameplayLogic; both touch only the agent's own
// rows, so they are fused here and run at the top of the agent's move stage.
#pragma once

namespace mpenv {

constexpr int32_t kAStarPolicyID = -1;     // consts.hpp:68
constexpr int kNumTurnBuckets = 5;         // consts.hpp:49

MP_HD Vec3 navTriCenter(const NavView &nav, int tri)
{
    Vec3 c = v3(0.f, 0.f, 0.f);
    for (int i = 0; i < 3; i++) c = c + nav.verts[nav.tris[tri * 3 + i]] / 3.0f;
    return c;
}

// NearestNavTri (sim.cpp:4975-5013): first triangle whose three edge cross
// products agree in sign, else the running "closest" (the reference compares
// squared distances against a value it overwrites with |c.z|; kept as is).
MP_HD int nearestNavTri(const NavView &nav, Vec3 pos)
{
    float closest = FLT_MAX;
    int closest_idx = -1;
    for (int tri = 0; tri < (int)nav.numTris; tri++) {
        bool contained = true;
        bool gtz = false;
        for (int i = 0; i < 3; i++) {
            const Vec3 v1 = nav.verts[nav.tris[tri * 3 + i]];
            const Vec3 v2 = nav.verts[nav.tris[tri * 3 + ((i + 1) % 3)]];
            const Vec3 e = v2 - v1;
            const Vec3 vp = pos - v1;
            const Vec3 c = cross(e, vp);
            if ((c.z > 0.0f) != gtz && i > 0) contained = false;
            gtz = c.z > 0.0f;
            float distsq = length2(v1 - pos);
            if (distsq < closest) {
                const float dir = dot(e, vp);
                const Vec3 perp = vp * (-dir / dot(e, e)) + e;
                distsq = dot(perp, perp);
                if (distsq < closest) {
                    closest = fabsf(c.z);
                    closest_idx = tri;
                }
            }
        }
        if (contained) return tri;
    }
    return closest_idx;
}

// PathfindToPoint (sim.cpp:5015-5035)
MP_HD Vec3 pathfindToPoint(const NavView &nav, Vec3 start, Vec3 pos)
{
    const int start_tri = nearestNavTri(nav, start);
    const int goal_tri = nearestNavTri(nav, pos);
    const int next_tri = (int)nav.lookup[(size_t)start_tri * nav.numTris + goal_tri];
    if (next_tri == -1) return v3(0.f, 0.f, 0.f);
    if (next_tri == goal_tri) return pos;
    return navTriCenter(nav, next_tri);
}

// Plans and applies one bot agent's action; returns false (and touches nothing)
// for agents driven by a policy.
MP_HD bool stageBot(const Ctx &c, int idx)
{
    const SimPtrs &p = c.p;
    if (p.policy[idx] != kAStarPolicyID) return false;
    if (c.nav.numTris == 0) return false;           // no navmesh.bin: the reference cannot run at all
    const int w = idx / kAgents;

    RNG &rng = p.combat[idx].rng;
    int move_amount = rngI32(rng, 0, 2);
    int move_angle = rngI32(rng, 0, 2);
    int r_yaw = rngI32(rng, 0, 5);
    const int r_pitch = rngI32(rng, 0, 2);
    const int r = p.mag[idx].numBullets == 0 ? 1 : 0;
    int stand = rngI32(rng, 0, 2);

    int f = 0;
    for (int i = 0; i < kAgents / 2; i++)
        if (p.canSee[(size_t)idx * 8 + i]) f = 1;

    const AABB zb = c.mc->zoneBoxes[p.world[w].zone.curZone];
    Vec3 center = 0.5f * (zb.pMin + zb.pMax);
    const Vec3 agent_pos = p.pos[idx];
    const Vec3 pos = v3(agent_pos.x, agent_pos.y, 0.0f);
    center = pathfindToPoint(c.nav, pos, center);

    center.z = 0.0f;
    const float yaw = p.aim[idx].yaw;
    const Vec3 fwd = v3(-mp_sinf(yaw), mp_cosf(yaw), 0.f);
    const Vec3 tgt_dir = normalize(center - pos);
    move_amount = dot(fwd, tgt_dir) > 0.6f ? 1 : 0;
    r_yaw = cross(fwd, tgt_dir).z < 0.0f ? 0 + move_amount : 4 - move_amount;
    move_amount *= 2;
    move_angle = 0;
    stand = 0;

    // anything (incl. a miss, depth -1) closer than 16 in the forward lidar: back off
    float collision_ang = 0.0f, collision_norm = 0.0f;
    const LidarSample *lidar = p.fwdLidar + (size_t)idx * (kFwdLidarH * kFwdLidarW);
    for (int y = 0; y < kFwdLidarH; y++) {
        for (int x = 0; x < kFwdLidarW; x++) {
            if (lidar[y * kFwdLidarW + x].depth < 16.0f) {
                collision_norm++;
                collision_ang += x;
            }
        }
    }
    if (collision_norm > 0.0f) {
        collision_ang /= collision_norm;
        move_amount = 1;
        switch ((int)(collision_ang / kFwdLidarW * 8.0f)) {
        case 0: move_angle = 2; break;
        case 1: case 2: move_angle = 3; break;
        case 3: case 4: move_angle = 4; move_amount = 2; break;
        case 5: case 6: move_angle = 5; break;
        case 7: move_angle = 6; break;
        }
    }
    if (r) f = 0;
    if (f) r_yaw = 2;

    // applyBotActionsSystem: the `reload` member of the plan is never applied
    DiscreteAction da;
    da.moveAmount = move_amount; da.moveAngle = move_angle; da.fire = f; da.stand = stand;
    p.actDiscrete[idx] = da;
    const float turn_delta = 10.f / float(kNumTurnBuckets / 2);       // discreteTurnDelta, sim.cpp:293-298
    AimAction aa;
    aa.yaw = turn_delta * (r_yaw - kNumTurnBuckets / 2);
    aa.pitch = turn_delta * (r_pitch - kNumTurnBuckets / 2);
    p.actAim[idx] = aa;
    return true;
}

}  // namespace mpenv
