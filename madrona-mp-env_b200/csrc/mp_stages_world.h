// mp_stages_world.h -- combat and per-world bookkeeping stages (included by
// mp_stages.h).  Reference: sim.cpp fireSystem :1443-1615, applyDmgSystem
// :1794-1836, respawnSystem :1838-1848, autoHealSystem :1875-1890, zoneSystem
// :1892-1976, breadcrumbs :4845-4926, zoneMatchInfoSystem :4470-4673,
// updateFiltersState :128-291, evaluateGoalRegionsSystem :4023-4087,
// exploreVisitedSystem :3508-3536, zoneRewardSystem :3849-3996, team/final
// reward :4292-4339, doneSystem :4712-4717, resetSystem/initWorld :732-872;
// utils.cpp spawnAgents :734-948, standardSpawnPoint :273-479; level_gen.cpp
// resetPersistentEntities :330-582.
#pragma once

namespace mpenv {

// ------------------------------------------------------------- event log
// logEvent (sim.cpp:23-39).  The reference appends to one global table with an
// atomic counter, so its order across worlds is whatever the executor produced;
// here every potential event has a fixed slot ([agent][2] for fireSystem, one
// Capture per world) and k_event_* compacts the used slots in world-major,
// agent-major order -- the order the reference's CPU path produces per world.
MP_HD uint64_t matchIDOf(const Ctx &c, int w)
{
    // initWorld (sim.cpp:736-738); ~0 until the first reset after construction (sim.cpp:5971)
    if (!c.p.world[w].matchIDValid) return ~0ull;
    return ((uint64_t)(c.cfg.worldIDOffset + (uint32_t)w) << 32) | (uint64_t)c.p.world[w].curEpisodeIdx;
}

MP_HD GameEvent makeEvent(uint32_t type, uint64_t match_id, int32_t step, uint8_t b0, uint8_t b1, uint16_t h1)
{
    GameEvent e;
    e.type = type; e.pad0 = 0; e.matchID = match_id; e.step = (uint32_t)step; e.b0 = b0; e.b1 = b1; e.h1 = h1;
    return e;
}

// Cold paths (event log, record / replay, trajectory curriculum) are compiled in only
// when the template flag kCold is set: the kernels are instantiated both ways and the
// host launches the lean variant when none of those features is configured, so the
// plain step keeps its register budget (k_world: 56 vs 157 registers).
MP_HD void logAgentEvent(const Ctx &c, int w, int a, int slot, uint32_t type, uint8_t b0, uint8_t b1)
{
    c.p.evAgent[(size_t)(w * kAgents + a) * kEventSlotsPerAgent + slot] =
        makeEvent(type, matchIDOf(c, w), c.p.world[w].match.curStep, b0, b1, 0);
}

// ============================================================ stage: fire
// fireSystem (sim.cpp:1443-1615).  Reads the world's 12 post-move positions
// and the target's HP / respawn timer; writes only this agent's rows plus
// DamageDealt[target][own team offset] (conflict-free: one slot per shooter).
// mesh: the static-mesh ray cast (RefMeshTrace, or a kernel's own traversal of the same BVH).
template <bool kCold, class MeshTrace>
MP_HD void stageFire(const Ctx &c, int w, int a, const Vec3 *world_pos, const MeshTrace &mesh)
{
    const SimPtrs &p = c.p;
    const int idx = w * kAgents + a;
    CombatState cs = p.combat[idx];
    cs.landedShotOn = -1;
    cs.successfulKill = 0;
    cs.firedShotT = -FLT_MAX;
    cs.reloadedFullMag = 0;
    const WeaponStats ws = c.mc->weapons[cs.weaponType];
    const bool log_events = kCold && c.cfg.eventLog != 0;
    if (log_events) {
        GameEvent *ev = p.evAgent + (size_t)idx * kEventSlotsPerAgent;
        ev[0].type = kEvNone;
        ev[1].type = kEvNone;
    }

    if (p.alive[idx] == 0.f) { p.combat[idx] = cs; return; }

    Magazine mag = p.mag[idx];
    const int32_t fire = p.actDiscrete[idx].fire;
    if (fire == 2) {
        // sim.cpp:1467-1476 (a reloading agent cannot also shoot this step, so slot 0 is free)
        if (log_events) logAgentEvent(c, w, a, 0, kEvReload, (uint8_t)a, (uint8_t)mag.numBullets);
        if (mag.numBullets == ws.magSize) cs.reloadedFullMag = 1;
        mag.numBullets = ws.magSize;
        mag.isReloading = ws.reloadTime;
    }
    bool reload_in_progress = mag.isReloading > 0;
    if (reload_in_progress) mag.isReloading -= 1;
    bool should_fire = false;
    if (!reload_in_progress && mag.numBullets > 0) should_fire = fire == 1;
    if (!should_fire) { p.mag[idx] = mag; p.combat[idx] = cs; return; }

    mag.numBullets -= 1;
    p.mag[idx] = mag;

    Vec3 fire_from = world_pos[a];
    fire_from.z += viewHeight(p.stand[idx].cur);

    float u1 = rngUniform(cs.rng);
    float u2 = rngUniform(cs.rng);
    float z1 = sqrtf(-2.f * mp_logf(u1)) * mp_cosf(2.f * kPi * u2);
    float z2 = sqrtf(-2.f * mp_logf(u1)) * mp_sinf(2.f * kPi * u2);
    float accuracy_scale = ws.accuracyScale;
    float upwardBias = 1.5f;
    float up_delta = fminf(fmaxf((z1 + upwardBias) * accuracy_scale, 0.f), 4.f * accuracy_scale);
    float right_delta = fminf(fmaxf(z2 * accuracy_scale, -4.f * accuracy_scale), 4.f * accuracy_scale);

    Aim aim = p.aim[idx];
    aim.yaw += right_delta;
    aim.pitch += up_delta;
    aim = computeAim(aim.yaw, aim.pitch);
    p.aim[idx] = aim;

    Vec3 fire_dir = rotateVec(aim.rot, kFwd());
    float hit_t;
    int32_t hit_agent;
    MP_COUNT_RAYS(2, 1);
    bool hit = traceRayWorldWith(mesh, world_pos, fire_from, fire_dir, &hit_t, &hit_agent);
    cs.firedShotT = hit ? hit_t : FLT_MAX;

    bool hit_success = hit;
    if (hit_agent < 0) hit_success = false;
    else {
        if (teamOf(hit_agent) == teamOf(a)) hit_success = false;
        if (hit_success && p.combat[w * kAgents + hit_agent].remainingRespawnSteps > 0) hit_success = false;
    }
    if (hit_success) {
        cs.landedShotOn = hit_agent;
        float tgt_hp = p.hp[w * kAgents + hit_agent];
        // sim.cpp:1580-1609
        if (log_events) logAgentEvent(c, w, a, 0, kEvPlayerShot, (uint8_t)a, (uint8_t)hit_agent);
        if (tgt_hp <= ws.dmgPerBullet) {
            cs.successfulKill = 1;
            if (log_events) logAgentEvent(c, w, a, 1, kEvKill, (uint8_t)a, (uint8_t)hit_agent);
        }
        p.dmg[(size_t)(w * kAgents + hit_agent) * kTeamSize + offsetOf(a)] = ws.dmgPerBullet;
    }
    p.combat[idx] = cs;
}

template <bool kCold = true>
MP_HD void stageFire(const Ctx &c, int w, int a, const Vec3 *world_pos)
{
    stageFire<kCold>(c, w, a, world_pos, RefMeshTrace{c.bvh});
}

// ============================================================ world stages
// applyDmgSystem (sim.cpp:1794-1836)
MP_HD void applyDmgSystem(const SimPtrs &p, int idx)
{
    CombatState &cs = p.combat[idx];
    cs.wasShotCount = 0;
    cs.wasKilled = 0;
    if (cs.remainingRespawnSteps > 0) cs.remainingRespawnSteps -= 1;
    float hp = p.hp[idx];
    for (int i = 0; i < kTeamSize; i++) {
        float d = p.dmg[(size_t)idx * kTeamSize + i];
        if (d > 0.f) {
            cs.wasShotCount += 1;
            cs.remainingStepsBeforeAutoheal = kAutohealDelaySteps;
        }
        hp -= d;
        p.dmg[(size_t)idx * kTeamSize + i] = 0.f;
    }
    if (p.alive[idx] == 1.f && hp <= 0.f) {
        cs.wasKilled = 1;
        cs.hasDiedDuringEpisode = 1;
    }
    if (hp <= 0.f) {
        hp = 0.f;
        p.alive[idx] = 0.f;
        p.pos[idx] = v3(0, 0, 10000.f);
        p.vel[idx] = v3(0, 0, 0);
    } else {
        p.alive[idx] = 1.f;
    }
    p.hp[idx] = hp;
}

// standardSpawnPoint (utils.cpp:273-479), Zone task.
MP_HD void standardSpawnPoint(const Ctx &c, int w, int a, bool is_respawn, Vec3 *out_pt,
                              float *out_yaw)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    RNG &agent_rng = p.combat[base + a].rng;
    SpawnUsage &usage = p.spawnUsage[w];
    const MatchInfo &mi = p.world[w].match;
    const Spawn *options;
    int32_t chosen;

    if (!is_respawn || mc.numRespawns == 0) {
        uint32_t *tracker;
        int32_t num_spawns;
        if (teamOf(a) == mi.teamA) { options = mc.aSpawns; num_spawns = (int32_t)mc.numASpawns; tracker = usage.a; }
        else { options = mc.bSpawns; num_spawns = (int32_t)mc.numBSpawns; tracker = usage.b; }
        int32_t init_idx = -1;
        for (int i = 0; i < 5; i++) {
            int32_t s = rngI32(agent_rng, 0, num_spawns);
            if (tracker[s] == (uint32_t)mi.curStep) continue;
            init_idx = s;
            break;
        }
        if (init_idx == -1) init_idx = rngI32(agent_rng, 0, num_spawns);
        chosen = init_idx;
        tracker[chosen] = (uint32_t)mi.curStep;
    } else {
        options = mc.respawns;
        const int32_t num_spawns = (int32_t)mc.numRespawns;
        AABB zb = mc.zoneBoxes[p.world[w].zone.curZone];
        Vec3 zone_center = 0.5f * (zb.pMin + zb.pMax);
        float best_score = FLT_MAX;
        int32_t best_idx = -1;
        for (int32_t s = 0; s < num_spawns; s++) {
            uint32_t last_used = usage.respawn[s];
            if (last_used == (uint32_t)mi.curStep) continue;
            float score = 0.f;
            // utils.cpp:418-419: the product is truncated into a uint32_t.
            uint32_t elapsed = (uint32_t)(kDeltaT * float((uint32_t)mi.curStep - last_used));
            constexpr float elapsed_weight = 0.1f;
            constexpr float dist_weight = 0.01f;
            if ((float)elapsed < 3.f) score += elapsed_weight * (3.f - (float)elapsed);
            Spawn sp = options[s];
            Vec3 spawn_pt = 0.5f * (sp.region.pMin + sp.region.pMax);
            for (int j = 0; j < kAgents; j++) {
                if (j == a) continue;
                if (p.alive[base + j] == 0.f) continue;
                float dist = distance(spawn_pt, p.pos[base + j]);
                if (dist < 4.f * kAgentRadius) score += 100000.f;
                else {
                    if (teamOf(j) == teamOf(a)) continue;
                    score += dist_weight * (1.f / dist);
                }
            }
            if (distance(spawn_pt, zone_center) < 100.f) score += 1000000.f;
            if (score < best_score) { best_idx = s; best_score = score; }
        }
        // The reference asserts best_idx != -1 (utils.cpp:475); with fewer
        // unused respawn points than simultaneous deaths fall back to 0.
        if (best_idx < 0) best_idx = 0;
        chosen = best_idx;
        usage.respawn[chosen] = (uint32_t)mi.curStep;
    }

    Spawn sp = options[chosen];
    float x_rnd = rngUniform(agent_rng);
    float y_rnd = rngUniform(agent_rng);
    float z_rnd = rngUniform(agent_rng);
    float yaw_rnd = rngUniform(agent_rng);
    float x_min = sp.region.pMin.x, x_diff = sp.region.pMax.x - x_min;
    float y_min = sp.region.pMin.y, y_diff = sp.region.pMax.y - y_min;
    float z_min = sp.region.pMin.z, z_diff = sp.region.pMax.z - z_min;
    *out_pt = v3(x_min + x_rnd * x_diff, y_min + y_rnd * y_diff, z_min + z_rnd * z_diff);
    *out_yaw = sp.yawMin + yaw_rnd * (sp.yawMax - sp.yawMin);
}

// The sub-zone an agent is assigned to: its policy id clamped to the sub-zone range (sim.cpp:1996-1997).
MP_HD int subZoneOf(const SimPtrs &p, int idx)
{
    const int32_t k = p.policy[idx];
    return k < 0 ? 0 : (k > kNumSubZones - 1 ? kNumSubZones - 1 : k);
}

// Zone-frame AABB of the current zone (shared by zoneSystem / spawnAgents).
MP_HD void zoneFrame(const MapConsts &mc, int32_t zone, Quat *to_zone, AABB *box)
{
    AABB zb = mc.zoneBoxes[zone];
    Quat q = quatInv(angleAxis(mc.zoneRots[zone], kUp()));
    zb.pMin = rotateVec(q, zb.pMin);
    zb.pMax = rotateVec(q, zb.pMax);
    *to_zone = q;
    *box = zb;
}

// spawnAgents (utils.cpp:734-948) for the supported flag set (standard spawns;
// RandomizeHPMagazine and the LearnShooting curriculum spawn honoured; NavmeshSpawn / HardcodedSpawns / SpawnInMiddle
// / curriculum spawns are rejected at create time).
MP_HD void spawnAgents(const Ctx &c, int w, bool is_respawn)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    int dead[kAgents];
    int num_dead = 0;
    for (int i = 0; i < kAgents; i++) if (p.alive[base + i] == 0.f) dead[num_dead++] = i;
    if (num_dead == 0) return;

    RNG &base_rng = p.world[w].baseRNG;
    (void)rngI32(base_rng, 0, 0);           // episodes[sampleI32(0, numEpisodes == 0)]
    const bool randomize_hp_mag = (c.cfg.simFlags & kFlagRandomizeHPMagazine) != 0;

    for (int k = 0; k < num_dead; k++) {
        const int a = dead[k];
        const int idx = base + a;
        Vec3 spawn_pt;
        float spawn_yaw;
        const float spawn_pitch = 0.f;
        standardSpawnPoint(c, w, a, is_respawn, &spawn_pt, &spawn_yaw);
        if ((c.cfg.simFlags & kFlagEnableCurriculum) && p.world[w].episodeCurriculum == 0) {
            // LearnShooting curriculum (utils.cpp:819-837): re-draw the point in
            // the half of the arena the standard spawn fell in.
            float rx = rngUniform(base_rng);
            float ry = rngUniform(base_rng);
            if (spawn_pt.y > 0.f) spawn_pt = v3(-700.f + rx * 1400.f, ry * 350.f, 0.f);
            else spawn_pt = v3(-700.f + rx * 1400.f, -ry * 350.f, 0.f);
        }

        p.pos[idx] = spawn_pt;
        p.rot[idx] = quatNormalize(angleAxis(spawn_yaw, kUp()));
        p.aim[idx] = computeAim(spawn_yaw, spawn_pitch);
        p.vel[idx] = v3(0, 0, 0);

        CombatState &cs = p.combat[idx];
        cs.weaponType = rngI32(base_rng, 0, mc.numWeaponTypes);
        const WeaponStats ws = mc.weapons[cs.weaponType];
        if (randomize_hp_mag) {
            int32_t tenth_hp = rngI32(base_rng, 1, 11);
            p.hp[idx] = float(tenth_hp * 10);
            p.mag[idx].numBullets = rngI32(base_rng, 0, ws.magSize);
            p.mag[idx].isReloading = 0;
        } else {
            p.hp[idx] = 100.f;
            p.mag[idx].numBullets = ws.magSize;
            p.mag[idx].isReloading = 0;
        }
        cs.remainingRespawnSteps = is_respawn ? 0 : kRespawnInvincibleSteps;
        cs.remainingStepsBeforeAutoheal = 0;
        {
            const int32_t zone = p.world[w].zone.curZone;
            AABB zb = mc.zoneBoxes[zone];
            Vec3 zone_center = (zb.pMax + zb.pMin) / 2.f;
            Quat to_zone; AABB zbox;
            zoneFrame(mc, zone, &to_zone, &zbox);
            // utils.cpp:899-903: containment uses the un-raised point, the
            // distance the point raised by half the stand height.
            Vec3 pos_in_zone = rotateVec(to_zone, spawn_pt);
            spawn_pt.z += kStandHeight / 2.f;
            cs.inZone = aabbContains(zbox, pos_in_zone) ? 1 : 0;
            cs.minDistToZone = distance(spawn_pt, zone_center);
        }
        if (c.cfg.simFlags & kFlagSubZones) {
            // utils.cpp:906-926: the agent's sub-zone is its policy id clamped to 0..7; spawn_pt has
            // already been raised by half the stand height above and is raised once more here
            const ZOBB &sz = mc.subZones[subZoneOf(p, idx)];
            Vec3 zone_center = (sz.pMax + sz.pMin) / 2.f;
            Quat to_zone = quatInv(angleAxis(sz.rotation, kUp()));
            AABB zbox; zbox.pMin = rotateVec(to_zone, sz.pMin); zbox.pMax = rotateVec(to_zone, sz.pMax);
            Vec3 pos_in_zone = rotateVec(to_zone, spawn_pt);
            spawn_pt.z += kStandHeight / 2.f;
            p.subZoneAgent[idx].inSubZone = aabbContains(zbox, pos_in_zone) ? 1 : 0;
            p.subZoneAgent[idx].minDistToSubZone = distance(spawn_pt, zone_center);
        }
        StandState st; st.cur = 0; st.tgt = 0; st.transition = 0;
        p.stand[idx] = st;
        MoveState ms; ms.newPos = p.pos[idx]; ms.newVel = v3(0, 0, 0); ms.maxVel = kMaxWalkVelocity;
        p.move[idx] = ms;
        p.alive[idx] = 1.f;
    }
}

// autoHealSystem (sim.cpp:1875-1890)
MP_HD void autoHealSystem(const SimPtrs &p, int idx)
{
    if (p.alive[idx] == 0.f) return;
    CombatState &cs = p.combat[idx];
    if (cs.remainingStepsBeforeAutoheal == 0 && p.hp[idx] < 100.f) {
        p.hp[idx] = fminf(100.f, p.hp[idx] + kAutohealPerStep);
    } else if (cs.remainingStepsBeforeAutoheal > 0) {
        cs.remainingStepsBeforeAutoheal -= 1;
    }
}

// zoneSystem (sim.cpp:1892-1976)
MP_HD void zoneSystem(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
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
        for (int i = 0; i < kAgents; i++) {
            p.combat[base + i].minDistToZone = distance(p.pos[base + i], zone_center);
        }
    }
    Quat to_zone; AABB zbox;
    zoneFrame(mc, zs.curZone, &to_zone, &zbox);
    int num_a = 0, num_b = 0;
    for (int i = 0; i < kAgents; i++) {
        Vec3 pos = p.pos[base + i];
        pos.z += kStandHeight / 2.f;
        Vec3 pz = rotateVec(to_zone, pos);
        if (!aabbContains(zbox, pz)) { p.combat[base + i].inZone = 0; continue; }
        p.combat[base + i].inZone = 1;
        if (teamOf(i) == 0) num_a += 1; else num_b += 1;
    }
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

// subzoneSystem (sim.cpp:1978-2043), all eight sub-zones of world w (SimFlags::SubZones).
MP_HD void subzoneSystems(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    for (int z = 0; z < kNumSubZones; z++) {
        const ZOBB &sz = mc.subZones[z];
        Quat to_zone = quatInv(angleAxis(sz.rotation, kUp()));
        AABB zbox; zbox.pMin = rotateVec(to_zone, sz.pMin); zbox.pMax = rotateVec(to_zone, sz.pMax);
        int num_a = 0, num_b = 0;
        for (int i = 0; i < kAgents; i++) {
            if (subZoneOf(p, base + i) != z) continue;
            Vec3 pos = p.pos[base + i];
            pos.z += kStandHeight / 2.f;
            if (!aabbContains(zbox, rotateVec(to_zone, pos))) { p.subZoneAgent[base + i].inSubZone = 0; continue; }
            p.subZoneAgent[base + i].inSubZone = 1;
            p.subZoneAgent[base + i].minDistToSubZone = 0.f;
            if (teamOf(i) == 0) num_a++; else num_b++;
        }
        SubZoneState &st = p.subZones[(size_t)w * kNumSubZones + z];
        st.isContested = (num_a > 0 && num_b > 0) ? 1 : 0;
        if (st.isContested || (num_a == 0 && num_b == 0)) {
            st.curControllingTeam = -1;
            st.isCaptured = 0;
        } else if (num_a > 0 && num_b == 0) {
            if (st.curControllingTeam != 0) { st.curControllingTeam = 0; st.isCaptured = 0; }
        } else if (num_a == 0 && num_b > 0) {
            if (st.curControllingTeam != 1) { st.curControllingTeam = 1; st.isCaptured = 0; }
        }
    }
}

// leaveBreadcrumbsSystem + accumulateBreadcrumbPenaltiesSystem
// (sim.cpp:4845-4926).  The reference's breadcrumb entities live in an ECS
// table in creation order; here they are a per-world pool kept in that same
// order (stable compaction), each crumb carrying a unique id that stands in
// for the entity handle + generation check (ctx.getCheck, sim.cpp:4857).
MP_HD void breadcrumbSystems(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    WorldState &ws = p.world[w];
    Crumb *pool = p.crumbs + (size_t)w * kMaxCrumbs;
    int n = ws.numCrumbs;

    for (int a = 0; a < kAgents; a++) {
        CrumbAgentState &as = p.crumbAgent[base + a];
        Vec3 pos = p.pos[base + a];
        as.totalPenalty = 0.f;
        bool updated_last = false;
        if (as.lastCrumbUID != 0) {
            for (int k = 0; k < n; k++) {
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
            if (as.stepsSinceNew > 10 && n < kMaxCrumbs) {
                Crumb cr; cr.pos = pos; cr.penalty = 1.f; cr.agent = a; cr.uid = ++ws.nextCrumbUID;
                pool[n++] = cr;
                as.stepsSinceNew = 0;
                as.lastCrumbUID = cr.uid;
            }
        }
    }

    int out = 0;
    for (int k = 0; k < n; k++) {
        Crumb cr = pool[k];
        const int team = teamOf(cr.agent);
        for (int off = 0; off < kTeamSize; off++) {
            if (off == offsetOf(cr.agent)) continue;
            const int j = base + team * kTeamSize + off;
            if (distance(p.pos[j], cr.pos) <= kAgentRadius * 4.f) {
                p.crumbAgent[j].totalPenalty += cr.penalty;
            }
        }
        cr.penalty -= 0.025f;
        if (cr.penalty <= 0.f) continue;        // destroyEntity + compaction
        pool[out++] = cr;
    }
    ws.numCrumbs = out;
}

// updateFiltersState (sim.cpp:128-291): three hard-coded analytics filters.
MP_HD void updateFiltersState(const Ctx &c, int w, int32_t cur_step)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    WorldState &ws = p.world[w];
    for (int f = 0; f < 3; f++) {
        for (int t = 0; t < 2; t++) {
            if ((ws.filtersActive[t] & (1u << f)) != 0) {
                if ((uint32_t)cur_step - (uint32_t)ws.filtersLastMatches[t][f] > 0u) {
                    ws.filtersActive[t] &= ~(1u << f);
                }
            }
        }
        if (f < 2) {
            // PlayerInRegion, i16 region bounds compared as floats.
            const float min_x = f == 0 ? -1272.f : 852.f, min_y = f == 0 ? -866.f : -851.f;
            const float max_x = f == 0 ? -825.f : 1280.f, max_y = f == 0 ? 696.f : 593.f;
            const int min_in_region = f == 0 ? 5 : 1;
            int num_in[2] = { 0, 0 };
            for (int i = 0; i < kAgents; i++) {
                Vec3 pos = p.pos[base + i];
                if (pos.x < min_x || pos.y < min_y || pos.x > max_x || pos.y > max_y) continue;
                num_in[teamOf(i)] += 1;
            }
            for (int t = 0; t < 2; t++) {
                if (num_in[t] >= min_in_region) {
                    ws.filtersActive[t] |= 1u << f;
                    ws.filtersLastMatches[t][f] = cur_step;
                }
            }
        } else {
            // PlayerShotEvent with the default (whole-map) regions.
            for (int i = 0; i < kAgents; i++) {
                int32_t tgt = p.combat[base + i].landedShotOn;
                if (tgt < 0) continue;
                Vec3 ap = p.pos[base + i], tp = p.pos[base + tgt];
                if (ap.x < -32768.f || ap.y < -32768.f || ap.x > 32767.f || ap.y > 32767.f ||
                    tp.x < -32768.f || tp.y < -32768.f || tp.x > 32767.f || tp.y > 32767.f) continue;
                ws.filtersActive[teamOf(i)] |= 1u << f;
                ws.filtersLastMatches[teamOf(i)][f] = cur_step;
            }
        }
    }
    for (int t = 0; t < 2; t++) {
        uint32_t m = ws.filtersActive[t];
        int pc = (int)((m & 1u) + ((m >> 1) & 1u) + ((m >> 2) & 1u));
        if (pc == 3) ws.filtersLastMatchedStep[t] = cur_step;
    }
}

// float -> i16 the way the reference's x86 build does it (cvttss2si to 32 bits,
// then truncation), not CUDA's saturating cvt.s16.f32.
MP_HD int16_t f2i16(float v) { return (int16_t)(int32_t)v; }

// writePackedStepSnapshot (sim.cpp:41-126).  eventLoggedInStep is a 0/1 flag
// there (atomic store of 1), eventMask the OR of the step's event types; both are
// cleared by every snapshot, and a snapshot is written every step, so they are
// exactly "this step's" events.
MP_HD void writePackedStepSnapshot(const Ctx &c, int w, bool captured)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const WorldState &ws = p.world[w];
    uint32_t mask = captured ? (uint32_t)kEvCapture : 0u;
    for (int i = 0; i < kAgents * kEventSlotsPerAgent; i++) mask |= p.evAgent[(size_t)base * kEventSlotsPerAgent + i].type;
    PackedStepSnapshot s;
    s.numEvents = mask != 0 ? 1u : 0u;
    s.eventMask = mask;
    s.matchID = matchIDOf(c, w);
    s.matchState.step = (uint16_t)ws.match.curStep;
    s.matchState.curZone = (uint8_t)ws.zone.curZone;
    s.matchState.curZoneController = (int8_t)(ws.zone.isCaptured ? ws.zone.curControllingTeam : -1);
    s.matchState.zoneStepsRemaining = (uint16_t)ws.zone.zoneStepsRemaining;
    s.matchState.stepsUntilPoint = (uint16_t)ws.zone.stepsUntilPoint;
    for (int i = 0; i < kAgents; i++) {
        PackedPlayerSnapshot &ps = s.players[i];
        const int idx = base + i;
        const Vec3 pos = p.pos[idx];
        ps.pos[0] = f2i16(pos.x); ps.pos[1] = f2i16(pos.y); ps.pos[2] = f2i16(pos.z);
        const Aim aim = p.aim[idx];
        ps.yaw = f2i16(aim.yaw * 32768 / kPi);
        ps.pitch = f2i16(aim.pitch * 32768 / kPi);
        const Magazine mag = p.mag[idx];
        ps.magNumBullets = (uint8_t)mag.numBullets;
        ps.isReloading = (uint8_t)mag.isReloading;
        ps.hp = (uint8_t)(int32_t)p.hp[idx];
        uint8_t flags = 0;
        if (p.combat[idx].landedShotOn >= 0) flags |= 1u << 1;      // FiredShot
        const int32_t pose = p.stand[idx].cur;
        if (pose == 1) flags |= 1u << 2;                             // Crouch
        else if (pose == 2) flags |= 1u << 3;                        // Prone
        ps.flags = flags;
    }
    p.stepSnap[w] = s;
}

// The event-log tail of zoneMatchInfoSystem (sim.cpp:4592-4634): the Capture event and
// the packed step snapshot.  evWorld[w]: the Capture event (type 0 = none); pad0 = 1
// when a step snapshot was written.
MP_HD void logWorldEvents(const Ctx &c, int w, bool new_captured)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const WorldState &ws = p.world[w];
    const ZoneState &zs = ws.zone;
    GameEvent cap = makeEvent(kEvNone, 0, 0, 0, 0, 0);
    if (ws.matchIDValid && new_captured) {
        Quat to_zone; AABB zbox;
        zoneFrame(*c.mc, zs.curZone, &to_zone, &zbox);
        uint32_t in_zone_mask = 0;
        for (int i = 0; i < kAgents; i++) {
            if (teamOf(i) != zs.curControllingTeam) continue;
            Vec3 pos = p.pos[base + i];
            pos.z += kStandHeight / 2.f;
            if (!aabbContains(zbox, rotateVec(to_zone, pos))) continue;
            in_zone_mask |= 1u << (uint32_t)i;
        }
        cap = makeEvent(kEvCapture, matchIDOf(c, w), ws.match.curStep, (uint8_t)zs.curZone,
                        (uint8_t)zs.curControllingTeam, (uint16_t)in_zone_mask);
    }
    if (ws.matchIDValid) {
        cap.pad0 = 1;
        writePackedStepSnapshot(c, w, new_captured);
    }
    p.evWorld[w] = cap;
}

// zoneMatchInfoSystem (sim.cpp:4470-4673), Task::Zone.
template <bool kCold = true>
MP_HD void zoneMatchInfoSystem(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    WorldState &ws = p.world[w];
    MatchInfo &mi = ws.match;
    int32_t cur_step = mi.curStep + 1;
    bool match_finished = false;
    if (cur_step >= kEpisodeLen || p.resetFlag[w] == 1) match_finished = true;
    MatchResult &mr = p.matchResult[w];
    if (cur_step == 1) {
        mr.winResult = -1;
        mr.teamTotalKills[0] = 0; mr.teamTotalKills[1] = 0;
        mr.teamObjectivePoints[0] = 0; mr.teamObjectivePoints[1] = 0;
    }
    for (int i = 0; i < kAgents; i++) {
        if (p.combat[base + i].wasKilled) mr.teamTotalKills[teamOf(i) ^ 1] += 1;
    }
    ZoneState &zs = ws.zone;
    zs.earnedPoint = 0;
    bool new_captured = false;
    if (zs.stepsUntilPoint == 0) {
        zs.stepsUntilPoint = kZonePointInterval;
        if (!zs.isCaptured) { zs.isCaptured = 1; new_captured = true; }
        // The reference asserts curControllingTeam != -1 here (sim.cpp:4521).
        if (zs.curControllingTeam >= 0) mr.teamObjectivePoints[zs.curControllingTeam] += 1;
        zs.earnedPoint = 1;
    }
    if (mr.teamObjectivePoints[0] >= kZoneWinPoints || mr.teamObjectivePoints[1] >= kZoneWinPoints) {
        match_finished = true;
    }
    // Task::ZoneCaptureDefend (sim.cpp:4534-4573): the attackers (team A) win on their first
    // objective point, the defenders at 8 or once every attacker has died during the episode.
    constexpr int32_t capdefend_attacker_win = 1, capdefend_defender_win = 8;
    const int32_t attacker_team = mi.teamA == 1 ? 1 : 0, defender_team = attacker_team ^ 1;
    bool team_has_all_died[2] = { true, true };
    const bool capdefend = kCold && c.cfg.taskCaptureDefend;
    if (capdefend) {
        if (mr.teamObjectivePoints[attacker_team] == capdefend_attacker_win) match_finished = true;
        if (mr.teamObjectivePoints[defender_team] == capdefend_defender_win) match_finished = true;
        for (int i = 0; i < kAgents; i++) {
            if (!p.combat[base + i].hasDiedDuringEpisode) team_has_all_died[teamOf(i)] = false;
        }
        if (team_has_all_died[attacker_team]) match_finished = true;
    }
    {
        ZoneStats &st = ws.zoneStats[zs.curZone];
        st.activeSteps += 1;
        if (zs.isCaptured && zs.curControllingTeam >= 0) st.capturedSteps[zs.curControllingTeam] += 1;
        if (zs.isContested) st.contestedSteps += 1;
        if (new_captured) st.numSwaps += 1;
        updateFiltersState(c, w, cur_step);
        if (kCold && c.cfg.eventLog) logWorldEvents(c, w, new_captured);      // sim.cpp:4592-4634
    }
    if (match_finished) {
        if (capdefend) {                                   // sim.cpp:4637-4646
            if (mr.teamObjectivePoints[attacker_team] == capdefend_attacker_win) mr.winResult = attacker_team;
            else if (mr.teamObjectivePoints[defender_team] == capdefend_defender_win ||
                     team_has_all_died[attacker_team]) mr.winResult = defender_team;
            else mr.winResult = 2;
        } else if (mr.teamObjectivePoints[0] > mr.teamObjectivePoints[1]) mr.winResult = 0;
        else if (mr.teamObjectivePoints[1] > mr.teamObjectivePoints[0]) mr.winResult = 1;
        else mr.winResult = 2;
        for (int z = 0; z < kMaxZones; z++) {
            mr.zoneStats[z] = ws.zoneStats[z];
            ZoneStats zero; zero.numSwaps = 0; zero.capturedSteps[0] = 0; zero.capturedSteps[1] = 0;
            zero.contestedSteps = 0; zero.activeSteps = 0;
            ws.zoneStats[z] = zero;
        }
    }
    mi.curStep = cur_step;
    mi.isFinished = match_finished ? 1 : 0;
}

// distToZOBB (sim.cpp:3998-4021)
MP_HD float distToZOBB(ZOBB z, Vec3 pos)
{
    Quat q = quatInv(angleAxis(z.rotation, kUp()));
    Vec3 p_min = rotateVec(q, z.pMin), p_max = rotateVec(q, z.pMax);
    Vec3 pf = rotateVec(q, pos);
    float sq = 0.f;
    for (int i = 0; i < 3; i++) {
        float v = comp(pf, i), lo = comp(p_min, i), hi = comp(p_max, i);
        if (v < lo) sq += (lo - v) * (lo - v);
        if (v > hi) sq += (v - hi) * (v - hi);
    }
    return sqrtf(sq);
}

// evaluateGoalRegionsSystem (sim.cpp:4023-4087)
MP_HD void evaluateGoalRegionsSystem(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    WorldState &ws = p.world[w];
    ws.goalTeamStepRewards[0] = 0.f;
    ws.goalTeamStepRewards[1] = 0.f;
    const int32_t attacker_team = ws.match.teamA;
    for (int r = 0; r < mc.numGoalRegions; r++) {
        const GoalRegion &gr = mc.goalRegions[r];
        int32_t region_team = gr.attackerTeam ? attacker_team : (attacker_team ^ 1);
        float max_min = -FLT_MAX;
        for (int s = 0; s < gr.numSub; s++) {
            float min_d = FLT_MAX;
            for (int i = 0; i < kAgents; i++) {
                if (teamOf(i) != region_team) continue;
                float d = distToZOBB(gr.sub[s], p.pos[base + i]);
                if (d < min_d) min_d = d;
            }
            if (min_d > max_min) max_min = min_d;
        }
        float prev = ws.goalMinDist[r];
        if (prev == FLT_MAX) {
            ws.goalMinDist[r] = max_min;
        } else {
            float diff = prev - max_min;
            if (diff > 0.f) {
                ws.goalMinDist[r] = max_min;
                ws.goalTeamStepRewards[region_team] += diff * gr.rewardStrength;
            }
        }
    }
}

// exploreVisitedSystem (sim.cpp:3508-3536)
MP_HD void exploreVisitedSystem(const Ctx &c, int w, int idx)
{
    const SimPtrs &p = c.p;
    Vec3 delta = p.pos[idx] - p.startPos[idx];
    int32_t x = int32_t((delta.x + 0.5f) / (kAgentRadius * 2.f));
    int32_t y = int32_t((delta.y + 0.5f) / (kAgentRadius * 2.f));
    int32_t cell_x = x + kExploreGridMax, cell_y = y + kExploreGridMax;
    if (cell_x < 0 || cell_x >= kExploreGridW || cell_y < 0 || cell_y >= kExploreGridW) return;
    uint32_t *cell = p.exploreGrid + (size_t)idx * (kExploreGridW * kExploreGridW) +
                     cell_y * kExploreGridW + cell_x;
    uint32_t cur = p.world[w].curEpisodeIdx;
    if (*cell != cur) {
        *cell = cur;
        if (length2(delta) > 2.f) p.exploreNew[idx] += 1;
    }
}

// flankRewardSystem (sim.cpp:4211-4290): RewardMode::Flank (Manager::Config::trainFlank)
// replaces zoneRewardSystem in the Step graph (sim.cpp:5733-5748).  Note the reference
// tests "can the opponent see me" along the AGENT's own aim (sim.cpp:4261).
MP_HD void flankRewardSystem(const Ctx &c, int w, int a)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    const int idx = base + a;
    const Vec3 *world_pos = p.pos + base;
    const Vec3 pos = p.pos[idx];
    const Aim aim = p.aim[idx];
    float reward = 0.f;
    Vec3 vis_check_pos = pos;
    vis_check_pos.z += viewHeight(p.stand[idx].cur);
    constexpr float flank_dist = 100.f;
    const int team = teamOf(a);
    float teammate_reward = 0.f;
    for (int i = 0; i < kTeamSize; i++) {
        const int mate = team * kTeamSize + i;
        if (mate == a) continue;
        const Vec3 dir = world_pos[mate] - pos;
        const bool visible = agentVisible(c, w, world_pos, vis_check_pos, aim, mate);
        if (length2(dir) >= flank_dist * flank_dist || !visible) teammate_reward += 0.001f;
    }
    reward += teammate_reward;
    float opponent_reward = 0.f;
    for (int i = 0; i < kTeamSize; i++) {
        const int opp = (team ^ 1) * kTeamSize + i;
        Vec3 opp_pos = world_pos[opp];
        opp_pos.z += viewHeight(p.stand[base + opp].cur);
        if (!agentVisible(c, w, world_pos, opp_pos, aim, a)) opponent_reward += 0.001f;
    }
    reward += opponent_reward;
    const CombatState &cs = p.combat[idx];
    if (cs.landedShotOn >= 0) {
        const float yaw_diff = fabsf(p.aim[base + cs.landedShotOn].yaw - aim.yaw);
        if (yaw_diff > kPi) reward += cs.successfulKill ? 1.f : 0.2f;
    }
    const uint32_t num_new_cells = p.exploreNew[idx];
    p.exploreNew[idx] = 0;
    if (num_new_cells > 0) reward += float(num_new_cells) * p.rewardCoefs[idx].exploreScale;
    p.reward[idx] = reward;
}

// zoneRewardSystem (sim.cpp:3849-3996) incl. learnShootingRewardSystem
// (:3707-3732) under the LearnShooting world curriculum.
MP_HD void zoneRewardSystem(const Ctx &c, int w, int a)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    const int idx = base + a;
    CombatState &cs = p.combat[idx];
    float reward = 0.f;
    if (p.worldCurriculum[w] == 0) {
        if (cs.landedShotOn != -1) reward += 0.5f;
        else if (cs.firedShotT >= 0.f) reward -= 0.05f;
        if (cs.reloadedFullMag) reward -= 0.5f;
        p.reward[idx] = reward;
        return;
    }
    const RewardHyperParams hp = p.rewardCoefs[idx];
    reward -= hp.breadcrumbScale * p.crumbAgent[idx].totalPenalty;
    if (cs.reloadedFullMag) reward -= 0.5f;
    if (cs.successfulKill) reward += 1.f;
    if (cs.landedShotOn != -1) reward += hp.shotScale * 1.f;
    if (cs.wasKilled) reward -= 1.5f;
    if (cs.wasShotCount > 0) reward -= hp.shotScale * 1.f;
    uint32_t num_new_cells = p.exploreNew[idx];
    p.exploreNew[idx] = 0;
    if (num_new_cells > 0) reward += float(num_new_cells) * hp.exploreScale;

    const ZoneState &zs = p.world[w].zone;
    const Vec3 pos = p.pos[idx];
    if (cs.inZone) {
        reward += hp.inZoneScale;
    } else {
        AABB zb = mc.zoneBoxes[zs.curZone];
        Vec3 center = (zb.pMax + zb.pMin) / 2.f;
        float dist = distance(center, pos);
        if (dist < cs.minDistToZone) {
            float zone_dist_scale = hp.zoneDistScale;
            if (!cs.hasDiedDuringEpisode) zone_dist_scale *= 10.f;
            reward += zone_dist_scale * (cs.minDistToZone - dist);
            cs.minDistToZone = dist;
        }
    }
    if (zs.curControllingTeam != -1) {
        if (zs.curControllingTeam == teamOf(a)) {
            reward += hp.zoneTeamCtrlScale;
            if (zs.earnedPoint) reward += hp.zoneEarnedPointScale;
        } else {
            reward -= hp.zoneTeamCtrlScale;
            if (zs.earnedPoint) reward -= hp.zoneEarnedPointScale;
        }
    }
    if (p.alive[idx] == 0.f) {
        cs.successfulKill = 0;
        cs.landedShotOn = -1;
        cs.wasKilled = 0;
        cs.wasShotCount = 0;
        cs.firedShotT = -FLT_MAX;
        p.reward[idx] = reward;
        return;
    }
    {
        // Team-spread bonus: fan of triangles (self, teammate i, teammate i+1)
        // over the teammate list in team order without self (level_gen.cpp:209-223).
        float poly_area2x = 0.f;
        int mates[kTeamSize - 1];
        int n = 0;
        for (int off = 0; off < kTeamSize; off++) if (off != offsetOf(a)) mates[n++] = base + teamOf(a) * kTeamSize + off;
        for (int i = 0; i < kTeamSize - 2; i++) {
            Vec3 t1 = p.pos[mates[i]], t2 = p.pos[mates[i + 1]];
            float e1x = t1.x - pos.x, e1y = t1.y - pos.y;
            float e2x = t2.x - pos.x, e2y = t2.y - pos.y;
            float triarea2x = e1x * e2y - e1y * e2x;
            poly_area2x += fabsf(triarea2x);
        }
        float dx = mc.worldBounds.pMax.x - mc.worldBounds.pMin.x;
        float dy = mc.worldBounds.pMax.y - mc.worldBounds.pMin.y;
        float bounds_area = dx * dy;
        float fraction_controlled = poly_area2x / (2.f * bounds_area);
        reward += fraction_controlled * 1e-2f;
    }
    p.reward[idx] = reward;
}

// subzoneRewardSystem (sim.cpp:3734-3848): SimFlags::SubZones replaces zoneRewardSystem in the
// Step graph (sim.cpp:5749-5766) -- the zone terms are taken against the agent's own sub-zone.
MP_HD void subzoneRewardSystem(const Ctx &c, int w, int a)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int idx = w * kAgents + a;
    CombatState &cs = p.combat[idx];
    float reward = 0.f;
    if (p.worldCurriculum[w] == 0) {
        if (cs.landedShotOn != -1) reward += 0.5f;
        else if (cs.firedShotT >= 0.f) reward -= 0.05f;
        if (cs.reloadedFullMag) reward -= 0.5f;
        p.reward[idx] = reward;
        return;
    }
    const RewardHyperParams hp = p.rewardCoefs[idx];
    reward -= hp.breadcrumbScale * p.crumbAgent[idx].totalPenalty;
    if (cs.reloadedFullMag) reward -= 0.5f;
    if (cs.successfulKill) reward += 3.f;
    if (cs.landedShotOn != -1) reward += hp.shotScale * 1.f;
    if (cs.wasKilled) reward -= 1.5f;
    if (cs.wasShotCount > 0) reward -= hp.shotScale * 1.f;
    uint32_t num_new_cells = p.exploreNew[idx];
    p.exploreNew[idx] = 0;
    if (num_new_cells > 0) reward += float(num_new_cells) * hp.exploreScale;
    const int z = subZoneOf(p, idx);
    const SubZoneState &st = p.subZones[(size_t)w * kNumSubZones + z];
    SubZoneAgentState &sa = p.subZoneAgent[idx];
    if (sa.inSubZone) {
        reward += hp.inZoneScale;
    } else {
        const ZOBB &sz = mc.subZones[z];
        Vec3 center = (sz.pMax + sz.pMin) / 2.f;
        float dist = distance(center, p.pos[idx]);
        if (dist < sa.minDistToSubZone) {
            float zone_dist_scale = hp.zoneDistScale;
            if (!cs.hasDiedDuringEpisode) zone_dist_scale *= 10.f;
            reward += zone_dist_scale * (sa.minDistToSubZone - dist);
            sa.minDistToSubZone = dist;
        }
    }
    if (st.curControllingTeam != -1) {
        if (st.curControllingTeam == teamOf(a)) reward += hp.zoneTeamCtrlScale;
        else reward -= hp.zoneTeamCtrlScale;
    }
    if (p.alive[idx] == 0.f) {
        cs.successfulKill = 0;
        cs.landedShotOn = -1;
        cs.wasKilled = 0;
        cs.wasShotCount = 0;
        cs.firedShotT = -FLT_MAX;
    }
    p.reward[idx] = reward;
}

// zoneCaptureDefendRewardSystem (sim.cpp:4089-4209): Task::ZoneCaptureDefend's per-agent reward
// (goal-region progress of the agent's team, kills / hits, exploration, zone control for the
// controlling team only, +-20 / -5 at the end of the match).
MP_HD void zoneCaptureDefendRewardSystem(const Ctx &c, int w, int a)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int idx = w * kAgents + a;
    CombatState &cs = p.combat[idx];
    const WorldState &ws = p.world[w];
    const RewardHyperParams hp = p.rewardCoefs[idx];
    float reward = 0.f;
    reward += 0.02f * ws.goalTeamStepRewards[teamOf(a)];
    if (cs.reloadedFullMag) reward -= 0.01f;
    if (cs.successfulKill) reward += 1.f;
    if (cs.landedShotOn != -1) reward += hp.shotScale * 1.f;
    if (cs.wasKilled) reward -= 1.f;
    if (cs.wasShotCount > 0) reward -= hp.shotScale * 1.f;
    uint32_t num_new_cells = p.exploreNew[idx];
    p.exploreNew[idx] = 0;
    if (num_new_cells > 0) reward += float(num_new_cells) * hp.exploreScale;
    const ZoneState &zs = ws.zone;
    if (!cs.inZone) {
        AABB zb = mc.zoneBoxes[zs.curZone];
        Vec3 center = (zb.pMax + zb.pMin) / 2.f;
        float dist = distance(center, p.pos[idx]);
        if (dist < cs.minDistToZone) cs.minDistToZone = dist;
    }
    if (zs.curControllingTeam != -1 && zs.curControllingTeam == teamOf(a)) {
        reward += hp.zoneTeamCtrlScale;
        if (zs.earnedPoint) reward += hp.zoneEarnedPointScale;
    }
    if (ws.match.isFinished) {
        const int32_t win = p.matchResult[w].winResult;
        if (win == 2) reward -= 5.f;
        else if (win == teamOf(a)) reward += 20.f;
        else reward -= 20.f;
    }
    if (p.alive[idx] == 0.f) {
        cs.successfulKill = 0;
        cs.landedShotOn = -1;
        cs.wasKilled = 0;
        cs.wasShotCount = 0;
        cs.firedShotT = -FLT_MAX;
    }
    p.reward[idx] = reward;
}

// pvpTeamRewardSystem + pvpFinalRewardSystem (sim.cpp:4292-4339)
MP_HD void teamAndFinalReward(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    float team_rewards[2] = { 0, 0 };
    for (int i = 0; i < kAgents; i++) team_rewards[teamOf(i)] += p.reward[base + i];
    team_rewards[0] /= float(kTeamSize);
    team_rewards[1] /= float(kTeamSize);
    p.world[w].teamRewards[0] = team_rewards[0];
    p.world[w].teamRewards[1] = team_rewards[1];
    for (int i = 0; i < kAgents; i++) {
        float my_reward = p.reward[base + i];
        float team_spirit = p.rewardCoefs[base + i].teamSpirit;
        p.reward[base + i] = my_reward * (1.f - team_spirit) + team_rewards[teamOf(i)] * team_spirit;
    }
}

// level_gen.cpp:502-581: overwrite zone state, step counter and the agents with a
// random recorded snapshot (teams mapped through teamA).
MP_HD void restoreCurriculumSnapshot(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    WorldState &ws = p.world[w];
    ZoneState &zs = ws.zone;
    MatchInfo &mi = ws.match;
    const int32_t snapshot_idx = rngI32(ws.baseRNG, 0, c.cfg.numCurriculumSnapshots);
    const CurriculumSnapshot snap = p.curriculum[snapshot_idx];
    zs.curZone = snap.matchState.curZone;
    if (snap.matchState.curZoneController == -1) {
        zs.curControllingTeam = -1;
        zs.isCaptured = 0;
    } else {
        zs.isCaptured = 1;
        zs.curControllingTeam = snap.matchState.curZoneController;
        zs.stepsUntilPoint = snap.matchState.stepsUntilPoint;
        zs.zoneStepsRemaining = snap.matchState.zoneStepsRemaining;
    }
    mi.curStep = snap.matchState.step;
    for (int i = 0; i < kAgents; i++) {
        const int a = mi.teamA == 0 ? i : (i < kAgents / 2 ? kAgents / 2 + i : i - kAgents / 2);
        const int idx = base + a;
        const PackedPlayerSnapshot ps = snap.players[i];
        p.pos[idx] = v3((float)ps.pos[0], (float)ps.pos[1], (float)ps.pos[2]);
        const Aim aim = computeAim((float)ps.yaw * kPi / 32768.f, (float)ps.pitch * kPi / 32768.f);
        p.aim[idx] = aim;
        p.rot[idx] = quatNormalize(angleAxis(aim.yaw, kUp()));
        p.hp[idx] = (float)ps.hp;
        Magazine mag; mag.numBullets = ps.magNumBullets; mag.isReloading = ps.isReloading;
        p.mag[idx] = mag;
        if (ps.flags & (1u << 2)) { StandState st; st.cur = 1; st.tgt = 1; st.transition = 0; p.stand[idx] = st; }
        if (ps.flags & (1u << 3)) { StandState st; st.cur = 2; st.tgt = 2; st.transition = 0; p.stand[idx] = st; }
    }
}

// initWorld (sim.cpp:732-833) + resetPersistentEntities (level_gen.cpp:330-582)
template <bool kCold = true>
MP_HD void initWorld(const Ctx &c, int w, bool triggered_reset)
{
    const SimPtrs &p = c.p;
    const MapConsts &mc = *c.mc;
    const int base = w * kAgents;
    WorldState &ws = p.world[w];
    ws.episodeCurriculum = p.worldCurriculum[w];
    const TrainControl tc = *p.trainCtrl;
    MatchInfo &mi = ws.match;

    RandKey episode_key = randSplit(c.cfg.initRandKey, ws.curEpisodeIdx, c.cfg.worldIDOffset + (uint32_t)w);
    ws.baseRNG = rngInit(randSplit(episode_key, 0));
    RNG &base_rng = ws.baseRNG;

    bool flip_teams = false;
    if (tc.randomizeTeamSides) flip_teams = rngUniform(base_rng) < 0.5f;
    mi.teamA = flip_teams ? 1 : 0;
    if (triggered_reset && tc.randomizeEpisodeLengthAfterReset) mi.curStep = rngI32(base_rng, 0, kEpisodeLen - 1);
    else mi.curStep = 0;
    mi.isFinished = 0;

    // CurriculumState (sim.cpp:5915-5924): prob 1.0, tiers {0,0,.3,.3,.4}.
    mi.enableSpawnCurriculum = rngUniform(base_rng) < 1.0f ? 1 : 0;
    const float tier_probs[5] = { 0.f, 0.f, 0.3f, 0.3f, 0.4f };
    float tier_cdf[5];
    float running = 0.f;
    for (int i = 0; i < 5; i++) { running += tier_probs[i]; tier_cdf[i] = running; }
    float tier_sel = running * rngUniform(base_rng);
    for (uint32_t i = 0; i < 5; i++) {
        if (tier_sel < tier_cdf[i]) { mi.curCurriculumTier = i; break; }
    }
    // Spawn curriculum tiers are empty in this build (the navmesh BFS that
    // fills them feeds only #if 0 code, utils.cpp:811-818): numSpawns == 0.
    mi.curCurriculumSpawnIdx = (uint32_t)rngI32(base_rng, 0, 0);

    ZoneState &zs = ws.zone;
    zs.curZone = rngI32(base_rng, 0, (int32_t)mc.numZones);
    zs.curControllingTeam = -1;
    zs.isContested = 0;
    zs.isCaptured = 0;
    zs.earnedPoint = 0;
    zs.zoneStepsRemaining = kNumStepsPerZone;
    zs.stepsUntilPoint = kZonePointInterval;
    for (int z = 0; z < kNumSubZones; z++) {            // sim.cpp:815-820 (whatever the flags)
        SubZoneState &st = p.subZones[(size_t)w * kNumSubZones + z];
        st.curControllingTeam = -1; st.isContested = 0; st.isCaptured = 0;
    }
    if (c.cfg.taskCaptureDefend) zs.curZone = 3;       // sim.cpp:822-825 (after the draw: the RNG count is unchanged)

    // ---- resetPersistentEntities
    for (int i = 0; i < kAgents; i++) {
        const int idx = base + i;
        p.pos[idx] = v3(FLT_MAX, FLT_MAX, FLT_MAX);
        CombatState &cs = p.combat[idx];
        cs.rng = rngInit(randSplit(episode_key, (uint32_t)(i + 1)));
        cs.landedShotOn = -1;
        cs.remainingRespawnSteps = 0;
        cs.remainingStepsBeforeAutoheal = 0;
        cs.successfulKill = 0;
        cs.wasShotCount = 0;
        cs.wasKilled = 0;
        cs.firedShotT = -FLT_MAX;
        cs.hasDiedDuringEpisode = 0;
        // The reference builds this CombatState in an UNINITIALISED local (level_gen.cpp:343-355): reloadedFullMag
        // (like inZone / minDistToZone, which zoneSystem rewrites before any use) is whatever the stack held --
        // zero in every build of the oracle observed.  Nothing reads it before fireSystem clears it next step.
        cs.reloadedFullMag = 0;
        p.alive[idx] = 0.f;
        OpponentObs zero_ob;
        {
            float *z = (float *)&zero_ob;
            for (int k = 0; k < 32; k++) z[k] = 0.f;
        }
        NormPosObs def_pos; def_pos.x = -1000.f; def_pos.y = -1000.f; def_pos.z = -1000.f;
        for (int j = 0; j < kTeamSize; j++) {
            p.lastKnownObs[(size_t)idx * kTeamSize + j] = zero_ob;
            p.lastKnownPos[(size_t)idx * kTeamSize + j] = def_pos;
        }
        CrumbAgentState as; as.totalPenalty = 0.f; as.lastCrumbUID = 0; as.stepsSinceNew = 0;
        p.crumbAgent[idx] = as;
    }
    {
        SpawnUsage &su = p.spawnUsage[w];
        for (int i = 0; i < kMaxSpawns; i++) { su.a[i] = 0xFFFFFFFFu; su.b[i] = 0xFFFFFFFFu; su.respawn[i] = 0xFFFFFFFFu; }
    }
    spawnAgents(c, w, false);
    for (int i = 0; i < kAgents; i++) {
        const int idx = base + i;
        p.startPos[idx] = p.pos[idx];
        DiscreteAction da; da.moveAmount = 0; da.moveAngle = 0; da.fire = 0; da.stand = 0;
        p.actDiscrete[idx] = da;
        AimAction aa; aa.yaw = 0; aa.pitch = 0;
        p.actAim[idx] = aa;
        p.exploreNew[idx] = 0;
        // Nine per-agent coefficient draws are consumed and then overwritten
        // by the defaults (level_gen.cpp:427-446).
        for (int k = 0; k < 9; k++) (void)rngUniform(base_rng);
        RewardHyperParams d;
        d.teamSpirit = 0.f; d.shotScale = 0.5f; d.exploreScale = 0.005f; d.inZoneScale = 0.05f;
        d.zoneTeamContestScale = 0.01f; d.zoneTeamCtrlScale = 0.1f; d.zoneDistScale = 0.0005f;
        d.zoneEarnedPointScale = 1.f; d.breadcrumbScale = 0.1f;
        p.rewardCoefs[idx] = d;
    }
    for (int r = 0; r < mc.numGoalRegions; r++) ws.goalMinDist[r] = FLT_MAX;
    ws.goalTeamStepRewards[0] = 0.f;
    ws.goalTeamStepRewards[1] = 0.f;
    for (int t = 0; t < 2; t++) {
        FullTeamCommonObs zero;
        float *z = (float *)&zero;
        for (int k = 0; k < 24; k++) z[k] = 0.f;
        for (int j = 0; j < kTeamSize; j++) p.ftLastKnown[((size_t)w * 2 + t) * kTeamSize + j] = zero;
    }
    // Trajectory curriculum (level_gen.cpp:498-581): half of the episodes (not in
    // eval mode) start from a recorded mid-match snapshot instead of the spawn state.
    if (kCold && c.cfg.numCurriculumSnapshots > 0 && rngUniform(base_rng) < 0.5f && !tc.evalMode)
        restoreCurriculumSnapshot(c, w);
    // back in initWorld (sim.cpp:829-832)
    for (int t = 0; t < 2; t++) { ws.filtersActive[t] = 0; ws.filtersLastMatchedStep[t] = 0; }
    ws.matchIDValid = 1;
}

// resetSystem (sim.cpp:835-872)
template <bool kCold = true>
MP_HD void resetSystem(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    WorldState &ws = p.world[w];
    int32_t force_reset = p.resetFlag[w];
    int32_t should_reset = force_reset;
    if (c.cfg.autoReset && ws.match.isFinished) should_reset = 1;
    if (should_reset != 0) {
        p.resetFlag[w] = 0;
        ws.curEpisodeIdx = ws.worldEpisodeCounter++;
        if (c.cfg.simFlags & kFlagEnableCurriculum) {
            if (ws.curEpisodeIdx < 50) {
                if (rngUniform(ws.baseRNG) < (float)(ws.curEpisodeIdx + 1) / (float)50) p.worldCurriculum[w] = 1;
                else p.worldCurriculum[w] = 0;
            } else {
                p.worldCurriculum[w] = 1;
            }
        }
        initWorld<kCold>(c, w, force_reset == 1);
    }
}

// World construction: Sim::Sim (sim.cpp:5850-5980) + createPersistentEntities
// (level_gen.cpp:148-189).  Buffers arrive zero-filled (ECS tables are).
MP_HD void ctorWorld(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    WorldState &ws = p.world[w];
    ws.curEpisodeIdx = 0;
    ws.worldEpisodeCounter = 0;
    p.resetFlag[w] = 0;
    for (int i = 0; i < kAgents; i++) {
        const int idx = base + i;
        // level_gen.cpp:166-171: only the [0,40)^2 quadrant is pre-filled.
        uint32_t *grid = p.exploreGrid + (size_t)idx * (kExploreGridW * kExploreGridW);
        for (int y = 0; y < kExploreGridMax; y++)
            for (int x = 0; x < kExploreGridMax; x++) grid[y * kExploreGridW + x] = 0xFFFFFFFFu;
        p.policy[idx] = 0;
        DiscreteAimAction da; da.yaw = kAimYawBuckets / 2; da.pitch = kAimPitchBuckets / 2;
        p.actDiscreteAim[idx] = da;
    }
    p.worldCurriculum[w] = 1;           // WorldCurriculum::FullMatch (sim.cpp:5959)
    initWorld(c, w, true);
    for (int t = 0; t < 2; t++) { ws.filtersActive[t] = 0; ws.filtersLastMatchedStep[t] = -1; }
    ws.matchIDValid = 0;                // Sim::Sim leaves matchID = ~0 (sim.cpp:5971)
}

// pvpRecordSystem (sim.cpp:4750-4792), one agent's row; runs after zoneSystem
// (sim.cpp:5558-5563).
MP_HD void recordAgent(const Ctx &c, int w, int a)
{
    const SimPtrs &p = c.p;
    const int idx = w * kAgents + a;
    AgentLogData d;
    d.position = p.pos[idx];
    d.aim = p.aim[idx];
    d.hp = p.hp[idx];
    d.mag = p.mag[idx];
    d.standState = p.stand[idx];
    const CombatState &cs = p.combat[idx];
    d.shotAgentIdx = cs.landedShotOn;
    d.firedShotT = cs.firedShotT;
    d.wasKilled = cs.wasKilled ? 1 : 0;
    d.successfullKill = cs.successfulKill ? 1 : 0;
    d.pad[0] = 0; d.pad[1] = 0;
    p.recordLog[w].agentData[a] = d;
    if (a == 0) p.recordLog[w].curStep = p.world[w].match.curStep;
}

// pvpReplaySystem (sim.cpp:4794-4843), one agent's row: the gameplay systems
// are replaced by the logged state (pvpReplayLogic, sim.cpp:5587-5604).
MP_HD void replayAgent(const Ctx &c, int w, int a)
{
    const SimPtrs &p = c.p;
    const int idx = w * kAgents + a;
    const AgentLogData d = p.replayLog[w].agentData[a];
    p.pos[idx] = d.position;
    p.aim[idx] = d.aim;
    p.rot[idx] = quatNormalize(angleAxis(d.aim.yaw, kUp()));
    p.hp[idx] = d.hp;
    p.mag[idx] = d.mag;
    p.stand[idx] = d.standState;
    CombatState &cs = p.combat[idx];
    cs.landedShotOn = d.shotAgentIdx;
    cs.firedShotT = d.firedShotT;
    cs.wasKilled = d.wasKilled ? 1 : 0;
    cs.successfulKill = d.successfullKill ? 1 : 0;
    if (cs.wasKilled) cs.hasDiedDuringEpisode = 1;
}

// All per-world sequential work of one step after the fire stage, in task
// graph order (sim.cpp:5508-5842): damage, respawn, autoheal, zone,
// breadcrumbs, match info, goal regions, explore, rewards, done, reset.
MP_HD void stageWorld(const Ctx &c, int w)
{
    const SimPtrs &p = c.p;
    const int base = w * kAgents;
    if (c.cfg.replayLog) {
        p.world[w].match.curStep = p.replayLog[w].curStep;
        for (int i = 0; i < kAgents; i++) replayAgent(c, w, i);
        zoneSystem(c, w);
    } else {
        for (int i = 0; i < kAgents; i++) applyDmgSystem(p, base + i);
        if (!(c.cfg.simFlags & kFlagNoRespawn)) spawnAgents(c, w, true);
        for (int i = 0; i < kAgents; i++) autoHealSystem(p, base + i);
        zoneSystem(c, w);
        if (c.cfg.simFlags & kFlagSubZones) subzoneSystems(c, w);       // sim.cpp:5549-5555
        if (c.cfg.recordLog)
            for (int i = 0; i < kAgents; i++) recordAgent(c, w, i);
        breadcrumbSystems(c, w);
    }
    zoneMatchInfoSystem(c, w);
    evaluateGoalRegionsSystem(c, w);
    for (int i = 0; i < kAgents; i++) exploreVisitedSystem(c, w, base + i);
    for (int i = 0; i < kAgents; i++) {
        if (c.cfg.taskCaptureDefend) zoneCaptureDefendRewardSystem(c, w, i);
        else if (c.cfg.rewardFlank) flankRewardSystem(c, w, i);
        else if (c.cfg.simFlags & kFlagSubZones) subzoneRewardSystem(c, w, i);
        else zoneRewardSystem(c, w, i);
    }
    teamAndFinalReward(c, w);
    const int32_t finished = p.world[w].match.isFinished;
    for (int i = 0; i < kAgents; i++) p.done[base + i] = finished;
    // fullTeamDoneRewardSystem (sim.cpp:4720-4748)
    for (int t = 0; t < 2; t++) {
        float r = 0.f;
        for (int i = 0; i < kTeamSize; i++) r += p.reward[base + t * kTeamSize + i];
        p.ftReward[w * 2 + t] = r;
        p.ftDone[w * 2 + t] = finished ? 1 : 0;
    }
    resetSystem(c, w);
}

}  // namespace mpenv
