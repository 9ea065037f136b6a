"""SURVEY 8(f-1, f-2) on the B200 through the C ABI: event log + packed step
snapshots, record log and replay of the CUDA path against the oracle (the
reference's own sim code).  See test_logs_host.py for the ordering argument."""
import os

import numpy as np
import pytest

import simlib as S
from test_logs_host import assert_logs_equal

pytestmark = pytest.mark.gpu


def test_event_log_gpu_matches_oracle(tmp_path):
    """In-memory log vs the oracle every step, and the files the reference writes
    (events.bin / steps.bin, mgr.cpp:104-116, 218-231) hold the concatenation."""
    W = 8
    gpu = S.GpuSim(W, auto_reset=True, event_log=str(tmp_path))
    orc = S.Oracle(W, auto_reset=True, det=True, event_log=True)
    all_ev, all_sn, seen = [], [], 0
    try:
        gpu.init(); orc.init()
        for s in range(400):
            d, a = S.seek_policy(orc, s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a)
                sim.step()
            ev, sn = assert_logs_equal(gpu, orc, f"step {s}")
            assert len(sn) == W
            all_ev.append(ev); all_sn.append(sn)
            for t in ev["type"]:
                seen |= int(t)
        assert seen == (S.EV_CAPTURE | S.EV_RELOAD | S.EV_KILL | S.EV_PLAYER_SHOT), f"event types seen: {seen:#x}"
    finally:
        gpu.close(); orc.close()
    ev_file = np.fromfile(os.path.join(tmp_path, "events.bin"), S.EVENT_DT)
    sn_file = np.fromfile(os.path.join(tmp_path, "steps.bin"), S.SNAP_DT)
    assert ev_file.tobytes() == np.concatenate(all_ev).tobytes()
    assert sn_file.tobytes() == np.concatenate(all_sn).tobytes()


def test_event_log_large_batch_scan():
    """The single-CTA scan over more than one 1024-world chunk: offsets are exact
    (events stay grouped by world, ascending) and the count equals the used slots."""
    W = 3000
    gpu = S.GpuSim(W, auto_reset=True, event_log=True, full_team=False)
    try:
        gpu.init()
        total = 0
        for s in range(40):
            d, a = S.random_actions(s, W)
            gpu.set_actions(d, a)
            gpu.step()
            ev, sn = gpu.event_log()
            assert len(sn) == W and (sn["matchID"] >> np.uint64(32) == np.arange(W)).all()
            wid = (ev["matchID"] >> np.uint64(32)).astype(np.int64)
            assert (np.diff(wid) >= 0).all(), "events not grouped by ascending world"
            # a reload event for exactly the alive agents whose action was fire == 2
            n_reload = int((ev["type"] == S.EV_RELOAD).sum())
            alive = gpu.export("alive").reshape(-1) != 0
            assert n_reload <= int((d[:, 2] == 2).sum())
            mask_from_events = np.zeros(W, np.uint32)
            np.bitwise_or.at(mask_from_events, wid, ev["type"])
            assert (sn["eventMask"] == mask_from_events).all()
            total += len(ev)
        assert total > W, "random actions reload ~5% of the time"
    finally:
        gpu.close()


def test_record_and_replay_gpu(tmp_path):
    W, steps = 4, 150
    rec_path = os.path.join(tmp_path, "record.bin")
    gpu = S.GpuSim(W, auto_reset=True, record_log=rec_path)
    orc = S.Oracle(W, auto_reset=True, det=True, record_log=True)
    logs = []
    try:
        gpu.init(); orc.init()
        for s in range(steps):
            d, a = S.seek_policy(orc, s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a)
                sim.step()
            ra, rb = S.canonical_step_logs(gpu.record_log()), S.canonical_step_logs(orc.record_log())
            assert ra.tobytes() == rb.tobytes(), f"step {s}: record logs differ"
            logs.append(rb)
    finally:
        gpu.close(); orc.close()
    on_disk = np.fromfile(rec_path, S.STEP_LOG_DT)
    assert S.canonical_step_logs(on_disk).tobytes() == np.concatenate(logs).tobytes()

    # replay from the FILE on the GPU, from memory in the oracle
    rgpu = S.GpuSim(W, auto_reset=True, replay_log=rec_path, event_log=True)
    rorc = S.Oracle(W, auto_reset=True, det=True, replay_log=True, event_log=True)
    try:
        rgpu.init(); rorc.init()
        for s, log in enumerate(logs):
            rorc.set_replay_log(log)
            rgpu.step(); rorc.step()
            for name in S.STEP_OUTPUTS:
                x, y = rgpu.export(name), rorc.export(name)
                assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"replay step {s}: {name}"
            for w in range(W):
                dd = S.diff_world_state(rgpu.world_state(w), rorc.world_state(w))
                assert not dd, f"replay step {s} world {w}: {dd[:6]}"
            assert_logs_equal(rgpu, rorc, f"replay step {s}")
        with pytest.raises(RuntimeError, match="replay log exhausted"):
            rgpu.step()
    finally:
        rgpu.close(); rorc.close()


def test_python_manager_logs(tmp_path):
    """SimManager(event_log_path=..., record_log_path=...) (bindings.cpp:93-108) writes the files."""
    import madrona_mp_env as M
    W = 16
    rec = os.path.join(tmp_path, "rec.bin")
    mgr = M.SimManager(M.ExecMode.CUDA, 0, W, 10, True, M.SimFlags.Default, M.Task.Zone, 6, 0, 1,
                       S.SCENE, event_log_path=str(tmp_path), record_log_path=rec)
    mgr.init()
    for _ in range(5):
        mgr.step()
    del mgr
    assert os.path.getsize(os.path.join(tmp_path, "steps.bin")) == 5 * W * S.SNAP_DT.itemsize
    assert os.path.getsize(rec) == 5 * W * S.STEP_LOG_DT.itemsize
    assert os.path.getsize(os.path.join(tmp_path, "events.bin")) % S.EVENT_DT.itemsize == 0


def test_trajectory_curriculum_gpu(tmp_path, built):
    """curriculum_data_path (mgr.cpp:1424-1441) -> level_gen.cpp:498-581 on the device."""
    from test_logs_host import _make_curriculum
    cur = _make_curriculum(built)
    path = os.path.join(tmp_path, "curriculum.bin")
    cur.tofile(path)
    W = 8
    for flags in (0, 128):
        gpu = S.GpuSim(W, auto_reset=True, sim_flags=flags, curriculum_path=path)
        orc = S.Oracle(W, auto_reset=True, det=True, sim_flags=flags, curriculum=cur)
        try:
            gpu.init(); orc.init()
            assert sum(orc.world_state(w)["cur_step"] > 0 for w in range(W)) > 0 or flags
            for s in range(60):
                d, a = S.random_actions(s, W)
                for sim in (gpu, orc):
                    sim.set_actions(d, a)
                    if s % 20 == 10:
                        sim.set_export("reset", np.ones((W, 1), np.int32))
                    sim.step()
                for name in S.STEP_OUTPUTS:
                    x, y = gpu.export(name), orc.export(name)
                    assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"flags {flags} step {s}: {name}"
                for w in range(W):
                    dd = S.diff_world_state(gpu.world_state(w), orc.world_state(w), skip=("reloaded_full_mag",))
                    assert not dd, f"flags {flags} step {s} world {w}: {dd[:6]}"
        finally:
            gpu.close(); orc.close()


def test_flank_reward_mode_gpu():
    """train_flank (bindings.cpp:104, mgr.cpp:1745-1750): flankRewardSystem on the device."""
    W = 8
    gpu = S.GpuSim(W, auto_reset=True, train_flank=True)
    orc = S.Oracle(W, auto_reset=True, det=True, train_flank=True)
    try:
        gpu.init(); orc.init()
        for s in range(200):
            d, a = S.seek_policy(orc, s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a)
                sim.step()
            for name in S.STEP_OUTPUTS:
                x, y = gpu.export(name), orc.export(name)
                assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"step {s}: {name}"
        for w in range(W):
            dd = S.diff_world_state(gpu.world_state(w), orc.world_state(w))
            assert not dd, f"world {w}: {dd[:6]}"
        assert np.abs(orc.export("reward")).sum() > 0
    finally:
        gpu.close(); orc.close()


def test_ray_census_instrumented_build():
    """libmpenv_b200_count.so (same sources, -DMPENV_COUNT_RAYS): exact counts where they are
    known a priori, same simulation results as the product library, and the product refuses."""
    W, steps = 16, 30
    cnt = S.GpuSim(W, auto_reset=True, instrumented=True)
    gpu = S.GpuSim(W, auto_reset=True)
    try:
        with pytest.raises(RuntimeError, match="not an instrumented build"):
            gpu.ray_counts()
        cnt.init(); gpu.init()
        cnt.ray_counts(reset=True)
        fired = 0
        for s in range(steps):
            d, a = S.seek_policy(gpu, s, W)
            for sim in (cnt, gpu):
                sim.set_actions(d, a)
                sim.step()
            ws = [gpu.world_state(w) for w in range(W)]
            fired += sum(ag["fired_shot_t"] != -np.finfo(np.float32).max for st in ws for ag in st["agents"])
        for name in S.STEP_OUTPUTS:
            assert np.array_equal(cnt.export(name).view(np.int32), gpu.export(name).view(np.int32)), name
        c = cnt.ray_counts()
        assert c[0] == W * 12 * 80 * steps, "80 lidar rays per agent per step"
        assert c[2] == fired, "one hitscan ray per shot fired"
        assert 0 < c[1] <= W * 12 * 24 * steps, "at most 24 line-of-sight rays per agent-step"
        assert W * 12 * steps <= c[3] <= W * 12 * 11 * steps, "1..11 swept-sphere casts per agent-step"
    finally:
        cnt.close(); gpu.close()
