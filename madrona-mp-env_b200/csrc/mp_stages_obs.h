// mp_stages_obs.h -- line-of-sight, opponent masks, observation packing and
// lidar (included by mp_stages.h).  Reference: utils.cpp isAgentVisible
// :186-271 / inFrustum :169-184; sim.cpp opponentsWriteVisibilitySystem
// :2526-2560, pvpOpponentMasksSystem :2562-2614, pvpObservationsSystem
// :2645-3052, fullTeamObservationsSystem :3054-3301, pvpLidarSystem :3324-3506.
#pragma once

namespace mpenv {

// One of the four line-of-sight sample rays of isAgentVisible
// (utils.cpp:191-219): view-space cull, then the ray must hit `target`.
MP_HD bool losSampleVisible(const Ctx &c, const Vec3 *world_pos, Vec3 o, Quat aim_rot_inv,
                            Vec3 test_pos, int target)
{
    const float *fr = c.mc->frustum;
    Vec3 to_test = test_pos - o;
    Vec3 vp = rotateVec(aim_rot_inv, to_test);
    if (vp.y <= 0.f) return false;
    bool in_frustum = vp.y * fr[1] - fabsf(vp.x) * fr[0] > -kAgentRadius;
    in_frustum = in_frustum && (vp.y * fr[3] - fabsf(vp.z) * fr[2] > -kAgentRadius);
    if (!in_frustum) return false;
    float len = length(to_test);
    if (len < kAgentRadius) return false;
    to_test = to_test / len;
    float hit_t;
    int32_t hit_agent;
    bool hit = traceRayWorld(c.bvh, world_pos, o, to_test, &hit_t, &hit_agent);
    if (!hit) return false;
    return hit_agent == target;
}

// The sample point s (0 bottom, 1 top, 2 left, 3 right; utils.cpp:221-238).
MP_HD Vec3 losSamplePoint(Vec3 base_pos, float view_h, Vec3 delta_right, int s)
{
    Vec3 pt = base_pos;
    if (s == 0) { pt.z += kAgentRadius; return pt; }
    pt.z += view_h;
    if (s == 2) pt = pt - delta_right;
    else if (s == 3) pt = pt + delta_right;
    return pt;
}

// isAgentVisible (utils.cpp:186-271): any of the four sample points of agent `target`
// (index in world) seen from `o` looking along `aim`.
MP_HD bool agentVisible(const Ctx &c, int w, const Vec3 *world_pos, Vec3 o, const Aim &aim, int target)
{
    const Quat inv = quatInv(aim.rot);
    const float view_h = viewHeight(c.p.stand[w * kAgents + target].cur);
    const Vec3 delta_right = rotateVec(aim.rot, kRight()) * 0.9f * kAgentRadius;
    bool any = false;
    for (int s = 0; s < 4; s++) {      // bottom, top, left, right (utils.cpp:253-267)
        const Vec3 pt = losSamplePoint(world_pos[target], view_h, delta_right, s);
        if (losSampleVisible(c, world_pos, o, inv, pt, target)) any = true;
    }
    return any;
}

// opponentsWriteVisibilitySystem (sim.cpp:2526-2560) for one (agent, opponent
// slot) pair.  Returns canSee.
MP_HD bool visibilityPair(const Ctx &c, int w, int a, int j, const Vec3 *world_pos)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const int idx = base + a;
    if (p.alive[idx] == 0.f) return false;
    const int opp = (teamOf(a) ^ 1) * kTeamSize + j;
    if (p.alive[base + opp] == 0.f) return false;
    Vec3 ray_o = world_pos[a];
    ray_o.z += viewHeight(p.stand[idx].cur);
    const Aim aim = p.aim[idx];
    const Quat inv = quatInv(aim.rot);
    const float view_h = viewHeight(p.stand[base + opp].cur);
    Vec3 aim_right = rotateVec(aim.rot, kRight());
    Vec3 delta_right = aim_right * 0.9f * kAgentRadius;
    bool any = false;
    // order of utils.cpp:253-267: bottom, top, left, right
    for (int s = 0; s < 4; s++) {
        Vec3 pt = losSamplePoint(world_pos[opp], view_h, delta_right, s);
        if (losSampleVisible(c, world_pos, ray_o, inv, pt, opp)) any = true;
    }
    return any;
}

// ------------------------------------------------------------- observations
MP_HD float clamp01(float v) { return v < 0.f ? 0.f : (v > 1.f ? 1.f : v); }

MP_HD Vec3 normalizedPos(const MapConsts &mc, Vec3 q, bool clamp)
{
    const AABB &wb = mc.worldBounds;
    float x = (q.x - wb.pMin.x) / (wb.pMax.x - wb.pMin.x);
    float y = (q.y - wb.pMin.y) / (wb.pMax.y - wb.pMin.y);
    float z = (q.z - wb.pMin.z) / (wb.pMax.z - wb.pMin.z);
    if (clamp) { x = clamp01(x); y = clamp01(y); z = clamp01(z); }
    return v3(x, y, z);
}

// computeStandObs (sim.cpp:2442-2489)
MP_HD StandObs standObs(StandState s)
{
    StandObs o;
    o.curS = s.cur == 0 ? 1.f : 0.f; o.curC = s.cur == 1 ? 1.f : 0.f; o.curP = s.cur == 2 ? 1.f : 0.f;
    o.tgtS = s.tgt == 0 ? 1.f : 0.f; o.tgtC = s.tgt == 1 ? 1.f : 0.f; o.tgtP = s.tgt == 2 ? 1.f : 0.f;
    o.transition = (float)s.transition / (float)kPoseTransitionSpeed;
    return o;
}

// Stores one 128-byte observation row with 16-byte stores (the rows are 16-byte
// aligned: cudaMalloc base + a multiple of 128).  Per-field 4-byte stores made
// k_obs LSU-bound: 45 M sector writes per step for 4.3 M sectors of payload.
template <typename T> MP_HD void storeRow(T *dst, const T &src)
{
#if defined(__CUDA_ARCH__)
    static_assert(sizeof(T) % 16 == 0, "row is a multiple of 16 bytes");
    const float *s = reinterpret_cast<const float *>(&src);
    float4 *d = reinterpret_cast<float4 *>(dst);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 16); k++)
        d[k] = make_float4(s[4 * k], s[4 * k + 1], s[4 * k + 2], s[4 * k + 3]);
#else
    *dst = src;
#endif
}

template <typename T> MP_HD void zeroFloats(T *o)
{
    float *f = (float *)o;
    for (int k = 0; k < (int)(sizeof(T) / 4); k++) f[k] = 0.f;
}

MP_HD void wrapYaw(float &d)
{
    if (d > kPi) d -= 2.f * kPi;
    else if (d < -kPi) d += 2.f * kPi;
}

// direction-to-point block shared by the zone and other-player observations
// (sim.cpp:2821-2848, 2955-2982)
MP_HD void toTargetObs(Vec3 to, Aim self_aim, float *dist_out, float *yaw_out, float *pitch_out)
{
    float dist = length(to);
    if (dist < 1e-2f) { *dist_out = 0.f; *yaw_out = 0.f; *pitch_out = 0.f; return; }
    to = to / dist;
    float new_yaw = -mp_atan2f(to.x, to.y);
    float cz = to.z < -1.f ? -1.f : (to.z > 1.f ? 1.f : to.z);
    float new_pitch = mp_asinf(cz);
    float yaw_delta = new_yaw - self_aim.yaw;
    float pitch_delta = new_pitch - self_aim.pitch;
    wrapYaw(yaw_delta);
    *dist_out = dist; *yaw_out = yaw_delta; *pitch_out = pitch_delta;
}

// fillCommonOb (sim.cpp:2720-2774).  Returns false for a dead agent.
MP_HD bool fillCommonOb(const Ctx &c, PlayerCommonObs &ob, NormPosObs &pos_ob, int agent_idx,
                        Quat self_rot_inv)
{
    const SimPtrs &p = c.p;
    ob.isValid = 1.f;
    if (p.alive[agent_idx] == 0.f) return false;
    Vec3 pos = p.pos[agent_idx];
    Aim aim = p.aim[agent_idx];
    ob.isAlive = 1.f;
    Vec3 np = normalizedPos(*c.mc, pos, true);
    ob.globalX = np.x; ob.globalY = np.y; ob.globalZ = np.z;
    pos_ob.x = np.x; pos_ob.y = np.y; pos_ob.z = np.z;
    ob.facingYaw = 0.5f * ((aim.yaw / kPi) + 1.f);
    ob.facingPitch = 0.5f * (aim.pitch / (0.25f * kPi) + 1.f);
    Vec3 rel_vel = rotateVec(self_rot_inv, p.vel[agent_idx]);
    ob.velocityX = rel_vel.x; ob.velocityY = rel_vel.y; ob.velocityZ = rel_vel.z;
    ob.yawVelocity = 0.f;       // PvPDiscreteAimState is never non-zero (sim.cpp:2284-2334)
    ob.pitchVelocity = 0.f;
    ob.stand = standObs(p.stand[agent_idx]);
    const CombatState &cs = p.combat[agent_idx];
    ob.inZone = cs.inZone ? 1.f : 0.f;
    ob.weaponTypeObs[cs.weaponType] = 1.f;
    return true;
}

MP_HD void fillCombatOb(const SimPtrs &p, CombatObs &ob, int agent_idx)
{
    ob.hp = p.hp[agent_idx] / 100.f;
    ob.magazine = (float)p.mag[agent_idx].numBullets;
    ob.isReloading = (float)p.mag[agent_idx].isReloading;
    ob.timeBeforeAutoheal = float(p.combat[agent_idx].remainingStepsBeforeAutoheal) /
                            float(kAutohealDelaySteps);
}

MP_HD void fillOtherOb(const SimPtrs &p, OtherPlayerCommonObs &ob, int agent_idx, Vec3 self_pos,
                       Aim self_aim)
{
    Aim other_aim = p.aim[agent_idx];
    toTargetObs(p.pos[agent_idx] - self_pos, self_aim, &ob.toPlayerDist, &ob.toPlayerYaw,
                &ob.toPlayerPitch);
    float rfy = other_aim.yaw - self_aim.yaw;
    float rfp = other_aim.pitch - self_aim.pitch;
    wrapYaw(rfy);
    ob.relativeFacingYaw = rfy;
    ob.relativeFacingPitch = rfp;
}

// pvpOpponentMasksSystem + pvpObservationsSystem (sim.cpp:2645-3052) for agent a,
// split by output row so that the CUDA kernel can give every row its own thread
// (12 rows per agent: self, 5 teammates, 6 opponents).  The three parts write
// disjoint memory and read only state that this stage does not write.

// One opponent-mask entry (sim.cpp:2562-2614).
MP_HD float opponentMask(const Ctx &c, int w, int a, int i)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const int idx = base + a;
    const int team = teamOf(a);
    if (p.alive[idx] == 0.f) return 0.f;
    const int opp = base + (team ^ 1) * kTeamSize + i;
    if (p.alive[opp] == 0.f) return 0.f;
    float m = 0.f;
    for (int k = 0; k < kTeamSize; k++) {
        if (p.canSee[(size_t)(base + team * kTeamSize + k) * 8 + i]) { m = 1.f; break; }
    }
    if (p.combat[opp].firedShotT >= 0) m = 1.f;
    return m;
}

// SelfObservation, SelfPositionObservation, FiltersStateObservation.
MP_HD void stageObsSelf(const Ctx &c, int w, int a)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int idx = w * kAgents + a;
    const int team = teamOf(a);

    const WorldState &ws = p.world[w];
    p.filtersObs[idx] = (ws.match.curStep - ws.filtersLastMatchedStep[team] < 5) ? 1.f : 0.f;

    const Vec3 self_pos = p.pos[idx];
    const Aim self_aim = p.aim[idx];
    const Quat self_rot_inv = quatInv(p.rot[idx]);
    SelfObs self_ob;
    zeroFloats(&self_ob);
    NormPosObs self_pos_ob; self_pos_ob.x = -1000.f; self_pos_ob.y = -1000.f; self_pos_ob.z = -1000.f;
    const bool self_alive = fillCommonOb(c, self_ob, self_pos_ob, idx, self_rot_inv);
    if (self_alive) {
        fillCombatOb(p, self_ob.combat, idx);
        ZoneObs &zo = self_ob.zone;
        const ZoneState &zs = ws.zone;
        AABB zb = mc.zoneBoxes[zs.curZone];
        Vec3 zone_center = (zb.pMax + zb.pMin) / 2.f;
        Vec3 nc = normalizedPos(mc, zone_center, true);
        zo.centerX = nc.x; zo.centerY = nc.y; zo.centerZ = nc.z;
        toTargetObs(zone_center - self_pos, self_aim, &zo.toCenterDist, &zo.toCenterYaw, &zo.toCenterPitch);
        zo.myTeamControlling = (zs.curControllingTeam == team) ? 1.f : 0.f;
        zo.enemyTeamControlling = (zs.curControllingTeam != -1 && zs.curControllingTeam != team) ? 1.f : 0.f;
        zo.isContested = zs.isContested ? 1.f : 0.f;
        zo.isCaptured = zs.isCaptured ? 1.f : 0.f;
        zo.stepsUntilPoint = float(zs.stepsUntilPoint) / float(kZonePointInterval);
        zo.stepsRemaining = float(zs.zoneStepsRemaining) / float(kNumStepsPerZone);
        if (zs.curZone >= 0 && zs.curZone < 4) zo.id[zs.curZone] = 1.f;
    }
    p.selfObs[idx] = self_ob;
    p.selfPos[idx] = self_pos_ob;
}

// TeammateObservations / TeammatePositionObservations row `slot` (0..4).
MP_HD void stageObsTeammate(const Ctx &c, int w, int a, int slot)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const int idx = base + a;
    const int team = teamOf(a);
    const int off = slot < offsetOf(a) ? slot : slot + 1;       // teammates in offset order, self skipped
    const int mate = base + team * kTeamSize + off;
    TeammateObs ob; zeroFloats(&ob);
    NormPosObs pos_ob; pos_ob.x = -1000.f; pos_ob.y = -1000.f; pos_ob.z = -1000.f;
    if (p.alive[idx] != 0.f) {
        const Quat self_rot_inv = quatInv(p.rot[idx]);
        if (fillCommonOb(c, ob, pos_ob, mate, self_rot_inv)) {
            fillOtherOb(p, ob, mate, p.pos[idx], p.aim[idx]);
            fillCombatOb(p, ob.combat, mate);
        }
    }
    storeRow(&p.teammateObs[(size_t)idx * (kTeamSize - 1) + slot], ob);
    p.teammatePos[(size_t)idx * (kTeamSize - 1) + slot] = pos_ob;
}

// OpponentObservations / OpponentLastKnownObservations (+ position) row i and
// OpponentMasks entry i.
MP_HD void stageObsOpponent(const Ctx &c, int w, int a, int i)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const int idx = base + a;
    const int team = teamOf(a);
    const float mask = opponentMask(c, w, a, i);
    p.masks[(size_t)idx * kTeamSize + i] = mask;

    const int opp = base + (team ^ 1) * kTeamSize + i;
    const size_t row = (size_t)idx * kTeamSize + i;
    NormPosObs def_pos; def_pos.x = -1000.f; def_pos.y = -1000.f; def_pos.z = -1000.f;
    OpponentObs ob; zeroFloats(&ob);
    NormPosObs pos_ob = def_pos;
    if (p.alive[idx] == 0.f) {          // dead observer: zero rows, last-known rows untouched
        storeRow(&p.opponentObs[row], ob);
        p.opponentPos[row] = def_pos;
        return;
    }
    const Quat self_rot_inv = quatInv(p.rot[idx]);
    OpponentObs zero_ob; zeroFloats(&zero_ob);
    if (!fillCommonOb(c, ob, pos_ob, opp, self_rot_inv)) {
        storeRow(&p.lastKnownObs[row], zero_ob);
        p.lastKnownPos[row] = def_pos;
        storeRow(&p.opponentObs[row], ob);
        p.opponentPos[row] = pos_ob;
        return;
    }
    fillOtherOb(p, ob, opp, p.pos[idx], p.aim[idx]);
    const CombatState &ocs = p.combat[opp];
    if (ocs.wasKilled) {
        storeRow(&p.lastKnownObs[row], zero_ob);
        p.lastKnownPos[row] = def_pos;
    }
    ob.wasHit = (float)ocs.wasShotCount;
    ob.firedShot = ocs.firedShotT >= 0.f ? 1.f : 0.f;
    ob.hasLOS = p.canSee[(size_t)idx * 8 + i] ? 1.f : 0.f;
    const bool team_knows = mask == 1.f;
    ob.teamKnowsLocation = team_knows ? 1.f : 0.f;
    if (team_knows) {
        storeRow(&p.lastKnownObs[row], ob);
        p.lastKnownPos[row] = pos_ob;
    }
    storeRow(&p.opponentObs[row], ob);
    p.opponentPos[row] = pos_ob;
}

// Row r of agent a: 0 = self, 1..5 = teammates, 6..11 = opponents.
constexpr int kObsRowsPerAgent = 2 * kTeamSize;
MP_HD void stageObsRow(const Ctx &c, int w, int a, int r)
{
    if (r == 0) stageObsSelf(c, w, a);
    else if (r < kTeamSize) stageObsTeammate(c, w, a, r - 1);
    else stageObsOpponent(c, w, a, r - kTeamSize);
}

MP_HD void stageObs(const Ctx &c, int w, int a)
{
    for (int r = 0; r < kObsRowsPerAgent; r++) stageObsRow(c, w, a, r);
}

// fullTeamObservationsSystem (sim.cpp:3054-3301) for team t of world w.  Runs
// BEFORE the lidar stage, so the lidar copies are one step old (sim.cpp:5283-5309).
MP_HD void stageFullTeamObs(const Ctx &c, int w, int t)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    const size_t trow = (size_t)w * 2 + t;
    const WorldState &ws = p.world[w];

    FullTeamGlobalObs g;
    g.teamID[0] = t == 0 ? 0.f : 1.f;
    g.teamID[1] = t == 0 ? 1.f : 0.f;
    g.fractionMatchRemaining = float(kEpisodeLen - ws.match.curStep) / (float)kEpisodeLen;
    {
        const ZoneState &zs = ws.zone;
        AABB zb = mc.zoneBoxes[zs.curZone];
        Vec3 nc = normalizedPos(mc, (zb.pMax + zb.pMin) / 2.f, false);
        FullTeamZoneObs &zo = g.zone;
        zo.centerX = nc.x; zo.centerY = nc.y; zo.centerZ = nc.z;
        zo.myTeamControlling = (zs.curControllingTeam == t) ? 1.f : 0.f;
        zo.enemyTeamControlling = (zs.curControllingTeam != -1 && zs.curControllingTeam != t) ? 1.f : 0.f;
        zo.isContested = zs.isContested ? 1.f : 0.f;
        zo.isCaptured = zs.isCaptured ? 1.f : 0.f;
        zo.stepsUntilPoint = float(zs.stepsUntilPoint) / float(kZonePointInterval);
        zo.stepsRemaining = float(zs.zoneStepsRemaining) / float(kNumStepsPerZone);
        for (int k = 0; k < 4; k++) zo.id[k] = 0.f;
        if (zs.curZone >= 0 && zs.curZone < 4) zo.id[zs.curZone] = 1.f;
    }
    p.ftGlobal[trow] = g;

    auto fillCommon = [&](FullTeamCommonObs &ob, int agent_idx, int i) -> bool {
        ob.isValid = 1.f;
        ob.id[i] = 1.f;
        if (p.alive[agent_idx] == 0.f) return false;
        Aim aim = p.aim[agent_idx];
        ob.isAlive = 1.f;
        Vec3 np = normalizedPos(mc, p.pos[agent_idx], false);
        ob.globalX = np.x; ob.globalY = np.y; ob.globalZ = np.z;
        ob.facingYaw = 0.5f * ((aim.yaw / kPi) + 1.f);
        ob.facingPitch = 0.5f * (aim.pitch / (0.25f * kPi) + 1.f);
        Vec3 vel = p.vel[agent_idx];
        ob.velocityX = vel.x; ob.velocityY = vel.y; ob.velocityZ = vel.z;
        ob.stand = standObs(p.stand[agent_idx]);
        ob.inZone = p.combat[agent_idx].inZone ? 1.f : 0.f;
        return true;
    };

    for (int i = 0; i < kTeamSize; i++) {
        const int ag = base + t * kTeamSize + i;
        FullTeamPlayerObs ob; zeroFloats(&ob);
        if (fillCommon(ob, ag, i)) {
            ob.hp = p.hp[ag] / 100.f;
            ob.magazine = (float)p.mag[ag].numBullets / 30;
            ob.isReloading = (float)p.mag[ag].isReloading;
            ob.timeBeforeAutoheal = float(p.combat[ag].remainingStepsBeforeAutoheal) / float(kAutohealDelaySteps);
        }
        p.ftPlayers[trow * kTeamSize + i] = ob;
    }
    for (int i = 0; i < kTeamSize; i++) {
        const int ag = base + (t ^ 1) * kTeamSize + i;
        FullTeamEnemyObs ob; zeroFloats(&ob);
        FullTeamCommonObs last; zeroFloats(&last);     // cleared every step (sim.cpp:3072-3074)
        if (fillCommon(ob, ag, i)) {
            const CombatState &cs = p.combat[ag];
            ob.wasHit = (float)cs.wasShotCount;
            ob.firedShot = cs.firedShotT >= 0.f ? 1.f : 0.f;
            bool knows = ob.firedShot != 0.f;
            for (int j = 0; j < kTeamSize; j++) {
                if (p.canSee[(size_t)(base + t * kTeamSize + j) * 8 + i]) { ob.hasLOS[j] = 1.f; knows = true; }
            }
            ob.teamKnowsLocation = knows ? 1.f : 0.f;
            if (knows) last = ob;       // slices to the common prefix
        }
        p.ftEnemies[trow * kTeamSize + i] = ob;
        p.ftLastKnown[trow * kTeamSize + i] = last;
    }
    for (int i = 0; i < kTeamSize; i++) {
        const int ag = base + t * kTeamSize + i;
        const LidarSample *sf = p.fwdLidar + (size_t)ag * (kFwdLidarH * kFwdLidarW);
        LidarSample *df = p.ftFwdLidar + (trow * kTeamSize + i) * (kFwdLidarH * kFwdLidarW);
        for (int k = 0; k < kFwdLidarH * kFwdLidarW; k++) df[k] = sf[k];
        const LidarSample *sr = p.rearLidar + (size_t)ag * (kRearLidarH * kRearLidarW);
        LidarSample *dr = p.ftRearLidar + (trow * kTeamSize + i) * (kRearLidarH * kRearLidarW);
        for (int k = 0; k < kRearLidarH * kRearLidarW; k++) dr[k] = sr[k];
    }
}

// -------------------------------------------------------------------- lidar
// One lidar sample of pvpLidarSystem (sim.cpp:3324-3506).  ray in [0, 80):
// 0..63 = forward fan [h][w] about the aim rotation (135 degrees), 64..79 =
// rear fan [h][w] about the body rotation (180 degrees).
MP_HD LidarSample lidarRay(const Ctx &c, int w, int a, int ray, const Vec3 *world_pos)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int idx = w * kAgents + a;
    const bool is_fwd = ray < kFwdLidarH * kFwdLidarW;
    const int r = is_fwd ? ray : ray - kFwdLidarH * kFwdLidarW;
    const int width = is_fwd ? kFwdLidarW : kRearLidarW;
    const int h = r / width, col = r % width;
    const Quat q = is_fwd ? p.aim[idx].rot : p.rot[idx];
    Vec3 fwd = rotateVec(q, kFwd());
    Vec3 right = rotateVec(q, kRight());
    float top_height = viewHeight(p.stand[idx].cur) + kAgentRadius;
    float range = top_height - 2.f * kAgentRadius;
    // fwdLidarHeight == rearLidarHeight == 2 (sim.cpp:3423-3437)
    float h_off = kAgentRadius + range * (float(h) / float(kFwdLidarH - 1));
    Vec3 ray_o = world_pos[a];
    ray_o.z += h_off;
    float x = is_fwd ? mc.fwdLidarX[col] : mc.rearLidarX[col];
    float y = is_fwd ? mc.fwdLidarY[col] : mc.rearLidarY[col];
    Vec3 ray_dir = normalize(x * right + y * fwd);
    float hit_t;
    int32_t hit_agent;
    bool hit = traceRayWorld(c.bvh, world_pos, ray_o, ray_dir, &hit_t, &hit_agent);
    LidarSample s;
    if (hit) {
        s.depth = fminf(hit_t, mc.maxDist);
        s.isWall = hit_agent < 0 ? 1.f : 0.f;
        s.isTeammate = (hit_agent >= 0 && teamOf(hit_agent) == teamOf(a)) ? 1.f : 0.f;
        s.isOpponent = (hit_agent >= 0 && teamOf(hit_agent) != teamOf(a)) ? 1.f : 0.f;
    } else {
        s.depth = -1.f; s.isWall = 0.f; s.isTeammate = 0.f; s.isOpponent = 0.f;
    }
    return s;
}

}  // namespace mpenv
