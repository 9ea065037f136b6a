"""SURVEY 8(f-1, f-2) on the CPU: event log + packed step snapshots
(sim.cpp:23-126, 1467-1609, 4592-4634), record log (sim.cpp:4750-4792) and replay
(sim.cpp:4794-4843, 5587-5604) of the stage functions (tests/hostsim) against the
oracle (the reference's own sim code).  The GPU versions are in test_logs_gpu.py.

Order: the oracle's tables are gathered world by world in append order, i.e. the
order of the reference's CPU executor for ONE world; ours is world-major by
construction, so the arrays are compared element for element.  Bytes the
reference leaves unspecified (struct padding, the unused half of the payload
union) are masked by canonical_events().
"""
import numpy as np

import simlib as S


def assert_logs_equal(ours, orc, tag):
    ev_a, sn_a = ours.event_log()
    ev_b, sn_b = orc.event_log()
    assert len(ev_a) == len(ev_b), f"{tag}: {len(ev_a)} events vs oracle {len(ev_b)}"
    a, b = S.canonical_events(ev_a), S.canonical_events(ev_b)
    assert a.tobytes() == b.tobytes(), f"{tag}: events differ:\n{a[a != b][:4]}\n{b[a != b][:4]}"
    assert len(sn_a) == len(sn_b), f"{tag}: {len(sn_a)} snapshots vs oracle {len(sn_b)}"
    assert sn_a.tobytes() == sn_b.tobytes(), f"{tag}: step snapshots differ at {np.argwhere(sn_a != sn_b)[:4].tolist()}"
    return ev_a, sn_a


def test_event_log_and_snapshots_match_oracle(built):
    W = 8
    host = S.HostSim(W, auto_reset=True, event_log=True)
    orc = S.Oracle(W, auto_reset=True, det=True, event_log=True)
    seen = 0
    try:
        host.init(); orc.init()
        total = 0
        for s in range(400):
            d, a = S.seek_policy(orc, s, W)
            for sim in (host, orc):
                sim.set_actions(d, a)
                sim.step()
            ev, sn = assert_logs_equal(host, orc, f"step {s}")
            assert len(sn) == W
            total += len(ev)
            for t in ev["type"]:
                seen |= int(t)
            # the snapshot's numEvents is a 0/1 flag and its mask the OR of the world's event types
            for w in range(W):
                mine = ev[(ev["matchID"] >> np.uint64(32)) == w]
                mask = int(np.bitwise_or.reduce(mine["type"])) if len(mine) else 0
                assert sn[w]["eventMask"] == mask and sn[w]["numEvents"] == (1 if mask else 0)
        assert total > 50
        assert seen == (S.EV_CAPTURE | S.EV_RELOAD | S.EV_KILL | S.EV_PLAYER_SHOT), f"event types seen: {seen:#x}"
    finally:
        host.close(); orc.close()


def test_event_log_random_actions_with_offset(built):
    """matchID carries the GLOBAL world id (sim.cpp:736-738) and the episode index."""
    W, off = 4, 1000
    host = S.HostSim(W, auto_reset=True, event_log=True, world_id_offset=off)
    orc = S.Oracle(W, auto_reset=True, det=True, event_log=True, world_id_offset=off)
    try:
        host.init(); orc.init()
        for s in range(120):
            d, a = S.random_actions(s, W, off)
            for sim in (host, orc):
                sim.set_actions(d, a)
                sim.step()
            _, sn = assert_logs_equal(host, orc, f"step {s}")
            assert ((sn["matchID"] >> np.uint64(32)) == np.arange(off, off + W)).all()
            if s == 60:
                for sim in (host, orc):
                    sim.set_export("reset", np.array([[0], [1], [0], [0]], np.int32))
        assert sn["matchID"][1] & np.uint64(0xFFFFFFFF) == 1, "episode index after the forced reset"
    finally:
        host.close(); orc.close()


def test_record_log_matches_oracle_and_replays(built):
    """Record on both sides, byte-compare; then feed the oracle's record to a replaying
    hostsim AND a replaying oracle: same tensors, and the replayed agent state equals
    the recorded one."""
    W, steps = 4, 150
    host = S.HostSim(W, auto_reset=True, record_log=True)
    orc = S.Oracle(W, auto_reset=True, det=True, record_log=True)
    logs = []
    try:
        host.init(); orc.init()
        for s in range(steps):
            d, a = S.seek_policy(orc, s, W)
            for sim in (host, orc):
                sim.set_actions(d, a)
                sim.step()
            ra, rb = S.canonical_step_logs(host.record_log()), S.canonical_step_logs(orc.record_log())
            assert ra.tobytes() == rb.tobytes(), f"step {s}: record logs differ"
            logs.append(rb)
    finally:
        host.close(); orc.close()
    assert any((l["agentData"]["shotAgentIdx"] >= 0).any() for l in logs), "no landed shot recorded"

    rhost = S.HostSim(W, auto_reset=True, replay_log=True, event_log=True)
    rorc = S.Oracle(W, auto_reset=True, det=True, replay_log=True, event_log=True)
    try:
        rhost.init(); rorc.init()
        for s, log in enumerate(logs):
            for sim in (rhost, rorc):
                sim.set_replay_log(log)
                sim.step()
            for name in S.STEP_OUTPUTS:
                x, y = rhost.export(name), rorc.export(name)
                assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"replay step {s}: {name}"
            for w in range(W):
                dd = S.diff_world_state(rhost.world_state(w), rorc.world_state(w))
                assert not dd, f"replay step {s} world {w}: {dd[:6]}"
            assert_logs_equal(rhost, rorc, f"replay step {s}")
            # the replayed state IS the log (positions of agents that were not re-spawned by a reset)
            st = rhost.world_state(0)
            if st["cur_step"] == log[0]["curStep"] + 1:
                for i in range(12):
                    assert st["agents"][i]["hp"] == log[0]["agentData"][i]["hp"]
    finally:
        rhost.close(); rorc.close()


def _make_curriculum(built, W=4, steps=120):
    """CurriculumSnapshot records cut from the oracle's own step snapshots of a combat rollout."""
    orc = S.Oracle(W, auto_reset=True, det=True, event_log=True)
    snaps = []
    try:
        orc.init()
        for s in range(steps):
            d, a = S.seek_policy(orc, s, W)
            orc.set_actions(d, a)
            orc.step()
            sn = orc.event_log()[1]          # read every step: it also resets the global counters
            if s % 10 == 9:
                snaps.append(sn)
    finally:
        orc.close()
    cur = S.curriculum_from_step_snapshots(np.concatenate(snaps))
    # make sure the pose branches and a captured zone are exercised
    cur["players"]["flags"][0, 3] |= 1 << 2
    cur["players"]["flags"][1, 7] |= 1 << 3
    cur["curZoneController"][2] = 1
    return cur


def test_trajectory_curriculum_matches_oracle(built):
    """level_gen.cpp:498-581: half of the resets restore a recorded snapshot (zone state,
    step counter, per-agent position / aim / hp / magazine / pose), teams mapped through teamA."""
    cur = _make_curriculum(built)
    W = 8
    restored = 0
    for flags in (0, 128):          # RandomFlipTeams exercises the team remap
        host = S.HostSim(W, auto_reset=True, sim_flags=flags, curriculum=cur)
        orc = S.Oracle(W, auto_reset=True, det=True, sim_flags=flags, curriculum=cur)
        try:
            host.init(); orc.init()
            for s in range(60):
                restored += sum(orc.world_state(w)["cur_step"] > 0 for w in range(W)) if s == 0 else 0
                d, a = S.random_actions(s, W)
                for sim in (host, orc):
                    sim.set_actions(d, a)
                    if s % 20 == 10:
                        sim.set_export("reset", np.ones((W, 1), np.int32))
                    sim.step()
                for name in S.STEP_OUTPUTS:
                    x, y = host.export(name), orc.export(name)
                    assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"flags {flags} step {s}: {name}"
                # reloadedFullMag is left UNINITIALISED by the reference's reset (a local CombatState
                # with that field never assigned, level_gen.cpp:344-355) and re-assigned by the next
                # fireSystem before anything reads it: not comparable on a reset step.
                for w in range(W):
                    dd = S.diff_world_state(host.world_state(w), orc.world_state(w), skip=("reloaded_full_mag",))
                    assert not dd, f"flags {flags} step {s} world {w}: {dd[:6]}"
        finally:
            host.close(); orc.close()
    assert restored > 0, "no world started from a snapshot"


def test_flank_reward_mode_matches_oracle(built):
    """Manager::Config::trainFlank -> RewardMode::Flank: flankRewardSystem (sim.cpp:4211-4290)
    replaces zoneRewardSystem; 11 more isAgentVisible checks per agent."""
    W = 4
    host = S.HostSim(W, auto_reset=True, train_flank=True)
    orc = S.Oracle(W, auto_reset=True, det=True, train_flank=True)
    plain = S.Oracle(W, auto_reset=True, det=True)
    differs = False
    try:
        for sim in (host, orc, plain):
            sim.init()
        for s in range(200):
            d, a = S.seek_policy(orc, s, W)
            for sim in (host, orc, plain):
                sim.set_actions(d, a)
                sim.step()
            for name in S.STEP_OUTPUTS:
                x, y = host.export(name), orc.export(name)
                assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"step {s}: {name}"
            if s % 20 == 19:
                for w in range(W):
                    dd = S.diff_world_state(host.world_state(w), orc.world_state(w))
                    assert not dd, f"step {s} world {w}: {dd[:6]}"
            differs |= not np.array_equal(orc.export("reward"), plain.export("reward"))
        assert differs, "flank rewards never differed from the default reward mode"
    finally:
        host.close(); orc.close(); plain.close()
