"""CPU parity of the step LOGIC: the stage functions the CUDA kernels wrap
(madrona-mp-env_b200/csrc/mp_stages*.h, compiled for the host by tests/hostsim,
test infrastructure only) against the oracle (the reference's own simulation code
under oracle/shim, deterministic-transcendental build).  Bar: every exported
tensor and every field of the world state bit-identical -- integer and float.

The same comparisons run against the real kernels in test_gpu_parity.py; this
file lets the GPU-less container catch logic regressions.
"""
import json
import os

import numpy as np
import pytest

import simlib as S

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def assert_same(a, b, W, tag, full_team=True, worlds=None):
    names = S.STEP_OUTPUTS + (S.FULL_TEAM_OUTPUTS if full_team else [])
    for name in names:
        x, y = a.export(name), b.export(name)
        if not np.array_equal(x.view(np.int32), y.view(np.int32)):
            bad = np.argwhere(x.view(np.int32) != y.view(np.int32))
            raise AssertionError(f"{tag}: tensor {name}: {len(bad)} words differ, first at {bad[0].tolist()}: "
                                 f"ours={x[tuple(bad[0])]!r} oracle={y[tuple(bad[0])]!r}")
    for w in (range(W) if worlds is None else worlds):
        d = S.diff_world_state(a.world_state(w), b.world_state(w))
        assert not d, f"{tag}: world {w}: {d[:8]}"


def rollout(W, steps, policy, flags=0, auto_reset=False, check_every=1, seed=10):
    host = S.HostSim(W, rand_seed=seed, sim_flags=flags, auto_reset=auto_reset)
    orc = S.Oracle(W, rand_seed=seed, sim_flags=flags, auto_reset=auto_reset, det=True, num_threads=0)
    try:
        assert_same(host, orc, W, "ctor")
        host.init(); orc.init()
        assert_same(host, orc, W, "init")
        for s in range(steps):
            d, a = policy(orc, s, W)
            for sim in (host, orc):
                sim.set_actions(d, a)
                sim.step()
            if (s + 1) % check_every == 0 or s == steps - 1:
                assert_same(host, orc, W, f"step {s}")
        return orc.export("match_result")
    finally:
        host.close(); orc.close()


def test_larger_map(built, tmp_path):
    """simple_map plus 400 pillars (5052 triangles, a 1015-node BVH): asset reading, the BVH builder and
    the stage functions on a map that is not the reference's own -- against the oracle on that map."""
    scene = S.make_scene(str(tmp_path / "pillars"), extra_boxes=S.pillar_boxes(400))
    W = 4
    host = S.HostSim(W, auto_reset=True, scene=scene)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0, scene=scene)
    try:
        host.init(); orc.init()
        assert_same(host, orc, W, "init")
        for s in range(120):
            d, a = S.seek_policy(orc, s, W)
            for sim in (host, orc):
                sim.set_actions(d, a); sim.step()
            if s % 20 == 19:
                assert_same(host, orc, W, f"step {s}")
    finally:
        host.close(); orc.close()


def test_grazing_wall_bounded_sweep_regression(built):
    """World 165 of the 256-world soak (tools/soak.py), step 462: an agent slides along a 1300-unit wall at
    15.008 units and the reference's edge test reports the wall's diagonal at t = 6.88 although the sphere
    reaches the wall's plane at t = 7.16 -- inside the step's reach, where a sweep that was simply bounded at the
    distance the step can use returned a miss (mp_geom.h: kSweepSlack, sphereSweepBelow)."""
    W, off = 1, 165
    host = S.HostSim(W, auto_reset=True, full_team=False, world_id_offset=off)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=1, world_id_offset=off)
    try:
        host.init(); orc.init()
        for s in range(470):
            d, a = S.seek_policy(orc, s, W, world_id_offset=off) if (s // 100) % 2 == 0 else S.random_actions(s, W, off)
            for sim in (host, orc):
                sim.set_actions(d, a); sim.step()
            if s >= 455:
                assert_same(host, orc, W, f"step {s}", full_team=False)
    finally:
        host.close(); orc.close()


def test_episode_end_with_random_actions(built):
    """A full 3000-step episode under random actions, then the auto-reset: a third of the agents reload with a full
    magazine on any step, so CombatState::reloadedFullMag is set when initWorld rebuilds the agents (the reference
    leaves that field uninitialised there, level_gen.cpp:343-355; the oracle builds read it as 0)."""
    W = 1
    host = S.HostSim(W, auto_reset=True, full_team=False)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=1)
    try:
        host.init(); orc.init()
        for s in range(3004):
            d, a = S.random_actions(s, W)
            for sim in (host, orc):
                sim.set_actions(d, a); sim.step()
            if s % 500 == 499 or s >= 2996:
                assert_same(host, orc, W, f"step {s}", full_team=False)
    finally:
        host.close(); orc.close()


def random_policy(sim, s, W):
    return S.random_actions(s, W)


def test_random_actions_parity_run(built):
    """BASELINE.json configs[0] at its full length (random actions, 1000 steps) on 8 worlds: the stage
    functions run single-threaded on the host; the 32-world run is the GPU test's
    (test_gpu_parity.py::test_random_actions_32_worlds_bit_exact)."""
    rollout(8, 1000, random_policy, check_every=100)


def test_combat_policy(built):
    mr = rollout(8, 500, S.seek_policy, auto_reset=True, check_every=25)
    assert mr[:, 1:5].sum() > 0, "scenario produced no kills / points: combat path untested"


@pytest.mark.parametrize("flags", [2, 8, 16, 32, 128, 256, 1024, 2 | 16 | 128])
def test_sim_flags(built, flags):
    """RandomizeHPMagazine, NoRespawn, StaggerStarts, EnableCurriculum, RandomFlipTeams,
    StaticFlipTeams, SimEvalMode (sim_flags.hpp:7-21; mgr.cpp:1397-1413)."""
    rollout(4, 150, S.seek_policy, flags=flags, auto_reset=True, check_every=15)


def test_forced_reset_and_episode_end(built):
    """triggerReset mid-episode (mgr.cpp:2484-2500) and the 3000-step episode limit with autoReset."""
    W = 2
    host = S.HostSim(W, auto_reset=True)
    orc = S.Oracle(W, auto_reset=True, det=True)
    try:
        host.init(); orc.init()
        for s in range(60):
            d, a = S.random_actions(s, W)
            if s == 30:
                for sim in (host, orc):
                    sim.set_export("reset", np.array([[0], [1]], np.int32))
            for sim in (host, orc):
                sim.set_actions(d, a); sim.step()
            assert_same(host, orc, W, f"step {s}")
        st = orc.world_state(1)
        assert st["cur_step"] < 40, "world 1 was reset at step 30"
    finally:
        host.close(); orc.close()


def test_world_id_offset_shards_equal_big_batch(built):
    """a shard created with world_id_offset=k reproduces worlds [k, k+n)."""
    big, shard = S.HostSim(6), S.HostSim(3, world_id_offset=3)
    orc = S.Oracle(3, world_id_offset=3, det=True)
    try:
        for sim in (big, shard, orc):
            sim.init()
        for s in range(40):
            d, a = S.random_actions(s, 6)
            big.set_actions(d, a); big.step()
            for sim in (shard, orc):
                sim.set_actions(d[36:], a[36:]); sim.step()
        for name in S.STEP_OUTPUTS:
            x, y, z = big.export(name), shard.export(name), orc.export(name)
            assert np.array_equal(x[len(x) // 2:].view(np.int32), y.view(np.int32)), name
            assert np.array_equal(y.view(np.int32), z.view(np.int32)), name
    finally:
        big.close(); shard.close(); orc.close()


def test_golden_fixture(built):
    """tests/golden/state_combat_w8_first2.npz (raw tensors from the oracle, committed)."""
    spec = json.load(open(os.path.join(GOLDEN, "rollouts.json")))["combat_w8"]
    gold = np.load(os.path.join(GOLDEN, "state_combat_w8_first2.npz"))
    W = spec["worlds"]
    host = S.HostSim(W, rand_seed=spec["seed"], auto_reset=spec["auto_reset"])
    try:
        host.init()
        for s in range(spec["steps"]):
            d, a = S.seek_policy(host, s, W)
            host.set_actions(d, a); host.step()
        for name in S.STEP_OUTPUTS:
            x = host.export(name)[: len(gold[name])]
            assert np.array_equal(x.view(np.int32), gold[name].view(np.int32)), name
    finally:
        host.close()


def test_libm_oracle_within_tolerance(built):
    """The det oracle (the checker of every bit-exact test) against the libm oracle (the
    reference's own arithmetic), BASELINE.json configs[0]: 32 worlds x 1000 steps of the
    random-action stream, compared EVERY step (tests/bridge.py).  The two differ only
    through <= 2 ulp transcendentals.  north_star bar: integer exports identical;
    floats within 1e-4 * max(1, |row|_inf) except a counted set of discontinuity flips
    (a lidar ray grazing a silhouette edge lands on another surface), budgets in
    bridge.FLIP_BUDGET.  Measured in this container: integer exports identical for all
    1000 steps; fwd_lidar 7.8e-4 of rays outside the tolerance (first at step 38),
    rear_lidar 1.6e-4, every other tensor <= 5.2e-6 of its elements (first at step 205).
    The same comparison runs against the CUDA path in test_gpu_parity.py."""
    import bridge
    stats = bridge.run_bridge(lambda: S.Oracle(32, det=True, num_threads=0), W=32, steps=1000)
    stats.check_budgets()
    assert stats.flips["fwd_lidar"] > 0, "no discontinuity flip in 1000 steps: the comparison is not seeing the lidar"


def zone_capture_defend_rollout(make_ours, scene_dir, variant, flags, assert_same_fn):
    """Task::ZoneCaptureDefend (sim.cpp:822-825, 4089-4209, 4534-4573, 4637-4646) on a
    synthetic 4-zone copy of simple_map; the objective is pinned to zone 3.  variant:
      north   zone 3 next to the defenders' side: they win at 8 objective points;
      south   zone 3 next to the attackers' side: they win on their first point;
      sky     zone 3 out of reach (no points): at step 40 every attacker of worlds 0 and 2 gets
              hp = 0 through the exported hp tensor -> "all attackers have died" ends the match."""
    zones = list(S.FOUR_ZONES)
    if variant == "south":
        zones[3] = (-600.0, -1500.0, -10.0, 700.0, -500.0, 380.0, 0.3)
    elif variant == "sky":
        zones[3] = (-100.0, -100.0, 300.0, 100.0, 100.0, 380.0, 0.0)
    scene = S.make_scene(scene_dir, zones)
    W = 4
    ours = make_ours(W, flags, scene)
    orc = S.Oracle(W, sim_flags=flags, auto_reset=True, det=True, scene=scene, task=4)
    outcomes = set()
    try:
        ours.init(); orc.init()
        assert_same_fn(ours, orc, W, "init")
        for s in range(260 if variant != "sky" else 60):
            d, a = S.seek_policy(orc, s, W)
            if variant == "sky" and s == 40:
                hp = orc.export("hp")
                team_a = [orc.world_state(w)["team_a"] for w in (0, 2)]
                for w, ta in zip((0, 2), team_a):
                    hp[w * 12 + ta * 6: w * 12 + ta * 6 + 6] = 0.0
                for sim in (ours, orc):
                    sim.set_export("hp", hp)
            for sim in (ours, orc):
                sim.set_actions(d, a); sim.step()
            if s % 20 == 19 or (variant == "sky" and s in (40, 41)):
                assert_same_fn(ours, orc, W, f"step {s}")
            done = orc.export("done").reshape(W, 12)[:, 0]
            mr = orc.export("match_result")
            for w in np.flatnonzero(done):
                outcomes.add((int(mr[w, 0]) == orc.world_state(int(w))["team_a"], int(mr[w, 3] + mr[w, 4])))
        assert orc.world_state(1)["zone_cur"] == 3
    finally:
        ours.close(); orc.close()
    return outcomes        # {(attacker won?, objective points scored)}


@pytest.mark.parametrize("variant,flags", [("north", 0), ("south", 128), ("sky", 0), ("sky", 8)])
def test_zone_capture_defend(built, tmp_path, variant, flags):
    outcomes = zone_capture_defend_rollout(
        lambda W, f, scene: S.HostSim(W, sim_flags=f, auto_reset=True, scene=scene, task=4),
        str(tmp_path / "four_zones"), variant, flags, assert_same)
    expect = {"north": (False, 8), "south": (True, 1), "sky": (False, 0)}[variant]
    assert expect in outcomes, f"{variant}: outcomes {outcomes}"


def subzones_rollout(make_ours, assert_same_fn, steps=240):
    """SimFlags::SubZones (sim.cpp:1978-2043 subzoneSystem, 3734-3848 subzoneRewardSystem,
    utils.cpp:906-926, level_gen.cpp:279-322): every agent's sub-zone is its policy id (clamped to
    0..7), so the policy-assignment tensor spreads the agents over all eight sub-zones."""
    W = 4
    flags = 1 << 11
    ours = make_ours(W, flags)
    orc = S.Oracle(W, sim_flags=flags, auto_reset=True, det=True)
    try:
        pol = (np.arange(W * 12, dtype=np.int32) % 10 - 1).reshape(-1, 1)     # -1 .. 8: both clamps
        pol[pol == -1] = 0        # policy -1 is the scripted bot (another row of 8f): keep learned agents
        for sim in (ours, orc):
            sim.set_export("agent_policy", pol)
        ours.init(); orc.init()
        assert_same_fn(ours, orc, W, "init")
        # sub-zone boxes as level_gen.cpp:279-322 lays them out on simple_map (zones 1, 2 + six constants)
        boxes = [S.FOUR_ZONES[1][:6], S.FOUR_ZONES[2][:6], (-950, -500, 0, -50, 500, 1000), (50, -500, 0, 950, 500, 1000),
                 (-1000, -1650, 0, -50, -600, 1000), (50, -1650, 0, 1000, -600, 1000),
                 (-1000, 600, 0, -50, 1650, 1000), (1000, 600, 0, 50, 1650, 1000)]
        inside = controlled = 0
        for s in range(steps):
            d, a = S.seek_policy(orc, s, W)
            if s in (80, 160):
                # re-assign every agent to a sub-zone it is standing in (if any): the policy tensor is a
                # live input, so containment, control and contest all get exercised from here on
                pol = orc.export("agent_policy").copy()
                for w in range(W):
                    for i, ag in enumerate(orc.world_state(w)["agents"]):
                        x, y, z = ag["pos"][0], ag["pos"][1], ag["pos"][2] + 25.0
                        for k, b in enumerate(boxes):
                            if b[0] <= x <= b[3] and b[1] <= y <= b[4] and b[2] <= z <= b[5]:
                                pol[w * 12 + i, 0] = k
                                break
                for sim in (ours, orc):
                    sim.set_export("agent_policy", pol)
            for sim in (ours, orc):
                sim.set_actions(d, a); sim.step()
            if s % 20 == 19 or s in (80, 81, 160, 161):
                assert_same_fn(ours, orc, W, f"step {s}")
            for w in range(W):
                st = orc.world_state(w)
                inside += sum(ag["in_sub_zone"] for ag in st["agents"])
                controlled += int((np.asarray(st["sub_zones"])[:, 0] >= 0).sum())
        assert inside > 0 and controlled > 0, "no agent ever stood in its sub-zone: the system was not exercised"
    finally:
        ours.close(); orc.close()


def test_subzones(built):
    subzones_rollout(lambda W, f: S.HostSim(W, sim_flags=f, auto_reset=True), assert_same)
