"""GPU parity tests proper: the CUDA kernels, called through the C ABI of
libmpenv_b200.so with host buffers, against the oracle (the reference's own
sim.cpp / utils.cpp / level_gen.cpp / mesh_bvh.inl under oracle/shim, built with
the shared deterministic transcendentals).  Bar: bit-exact for every exported
tensor and every field of the world state, integer and float alike."""
import numpy as np
import pytest

import simlib as S

pytestmark = pytest.mark.gpu


def assert_same(gpu, orc, W, tag, full_team=True):
    names = S.STEP_OUTPUTS + (S.FULL_TEAM_OUTPUTS if full_team else [])
    for name in names:
        x, y = gpu.export(name), orc.export(name)
        if not np.array_equal(x.view(np.int32), y.view(np.int32)):
            bad = np.argwhere(x.view(np.int32) != y.view(np.int32))
            raise AssertionError(f"{tag}: tensor {name}: {len(bad)} words differ, first at {bad[0].tolist()}: "
                                 f"gpu={x[tuple(bad[0])]!r} oracle={y[tuple(bad[0])]!r}")
    for w in range(W):
        d = S.diff_world_state(gpu.world_state(w), orc.world_state(w))
        assert not d, f"{tag}: world {w}: {d[:8]}"


def rollout(W, steps, policy, flags=0, auto_reset=False, check_every=1, seed=10):
    gpu = S.GpuSim(W, rand_seed=seed, sim_flags=flags, auto_reset=auto_reset)
    orc = S.Oracle(W, rand_seed=seed, sim_flags=flags, auto_reset=auto_reset, det=True, num_threads=0)
    try:
        assert_same(gpu, orc, W, "ctor")
        gpu.init(); orc.init()
        assert_same(gpu, orc, W, "init")
        for s in range(steps):
            d, a = policy(orc, s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a)
                sim.step()
            if (s + 1) % check_every == 0 or s == steps - 1:
                assert_same(gpu, orc, W, f"step {s}")
        assert gpu.damaged_guards() == 0, "a kernel wrote outside its buffers (guard bands damaged)"
        return orc.export("match_result")
    finally:
        gpu.close(); orc.close()


def random_policy(sim, s, W):
    return S.random_actions(s, W)


def test_random_actions_32_worlds_bit_exact():
    # BASELINE.json configs[0] at its full length: 32 worlds, random actions, 1000 steps
    rollout(32, 1000, random_policy, check_every=50)


def test_gpu_vs_libm_reference_every_step():
    """The parity bridge (tests/bridge.py): the CUDA path against the UNMODIFIED-arithmetic
    reference (libm oracle), configs[0] at full length, every step: integer exports
    bit-identical, floats within 1e-4 * max(1, |row|_inf) up to the committed budget of
    lidar discontinuity flips (reference sim.cpp:3439-3500, utils.cpp:10-72)."""
    import bridge
    stats = bridge.run_bridge(lambda: S.GpuSim(32), W=32, steps=1000)
    stats.check_budgets()


def test_forced_reset_and_episode_end_gpu():
    """triggerReset mid-episode (mgr.cpp:2484-2500) on some worlds, then 40 more steps."""
    W = 6
    gpu = S.GpuSim(W, auto_reset=True)
    orc = S.Oracle(W, auto_reset=True, det=True)
    try:
        gpu.init(); orc.init()
        for s in range(70):
            d, a = S.random_actions(s, W)
            if s == 30:
                for sim in (gpu, orc):
                    sim.set_export("reset", np.array([[0], [1], [0], [1], [1], [0]], np.int32))
            for sim in (gpu, orc):
                sim.set_actions(d, a); sim.step()
            assert_same(gpu, orc, W, f"step {s}")
        assert orc.world_state(1)["cur_step"] < 45 and orc.world_state(0)["cur_step"] >= 70
    finally:
        gpu.close(); orc.close()


def test_combat_policy_bit_exact():
    mr = rollout(8, 400, S.seek_policy, auto_reset=True, check_every=10)
    assert mr[:, 1:5].sum() > 0, "scenario produced no kills / points: combat path untested"


@pytest.mark.parametrize("flags", [16, 128, 2, 8, 32, 2 | 16 | 128])
def test_sim_flags_bit_exact(flags):
    rollout(4, 120, S.seek_policy, flags=flags, auto_reset=True, check_every=10)


def test_world_id_offset_shards_match_one_big_batch():
    """Multi-GPU sharding rule: a shard created with world_id_offset=k
    reproduces worlds [k, k+n) of the big batch world for world."""
    big = S.GpuSim(8)
    shard = S.GpuSim(4, world_id_offset=4)
    try:
        big.init(); shard.init()
        for s in range(30):
            d, a = S.random_actions(s, 8)
            big.set_actions(d, a); shard.set_actions(d[48:], a[48:])
            big.step(); shard.step()
        for name in S.STEP_OUTPUTS:
            x, y = big.export(name), shard.export(name)
            assert np.array_equal(x[len(x) // 2:].view(np.int32), y.view(np.int32)), name
    finally:
        big.close(); shard.close()


def test_large_batch_properties():
    """BASELINE.json configs[1] size (8192 worlds): size-independent properties."""
    W = 8192
    gpu = S.GpuSim(W, auto_reset=True, full_team=False)
    try:
        gpu.init()
        for s in range(20):
            d, a = S.random_actions(s, W)
            gpu.set_actions(d, a)
            gpu.step()
        hp = gpu.export("hp"); alive = gpu.export("alive"); lidar = gpu.export("fwd_lidar")
        assert np.all((hp >= 0) & (hp <= 100)) and np.all((alive == 0) | (alive == 1))
        depth = lidar.reshape(-1, 4)[:, 0]
        assert np.all((depth == -1) | (depth > 0)) and np.isfinite(lidar).all()
        flags = lidar.reshape(-1, 4)[:, 1:]
        assert np.all(flags.sum(axis=1) == (depth != -1)), "exactly one class flag per hit"
        sp = gpu.export("self_pos")
        assert np.all((sp >= 0) & (sp <= 1))
        # oracle shards (world_id_offset) against the big batch: the first CTAs, the middle and
        # the very end of the batch (the dynamic task counters' tail), every exported tensor
        acts = [S.random_actions(s, W) for s in range(20)]
        for off, n in ((0, 32), (4095, 16), (8160, 32)):
            orc = S.Oracle(n, auto_reset=True, det=True, num_threads=0, world_id_offset=off)
            orc.init()
            for s in range(20):
                d, a = acts[s]
                orc.set_actions(d[off * 12:(off + n) * 12], a[off * 12:(off + n) * 12]); orc.step()
            for name in S.STEP_OUTPUTS:
                x, y = gpu.export(name), orc.export(name)
                per = len(y) // n
                assert np.array_equal(x[off * per:(off + n) * per].view(np.int32), y.view(np.int32)), (off, name)
            for w in (0, n - 1):
                dd = S.diff_world_state(gpu.world_state(off + w), orc.world_state(w))
                assert not dd, f"world {off + w}: {dd[:8]}"
            orc.close()
        assert gpu.damaged_guards() == 0, "a kernel wrote outside its buffers (guard bands damaged)"
    finally:
        gpu.close()


def test_golden_fixture_gpu():
    """The CUDA path against the committed oracle fixture (tests/golden/, made by
    make_golden.py in the container that has the reference)."""
    import json, os
    gdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    spec = json.load(open(os.path.join(gdir, "rollouts.json")))["combat_w8"]
    gold = np.load(os.path.join(gdir, "state_combat_w8_first2.npz"))
    W = spec["worlds"]
    gpu = S.GpuSim(W, rand_seed=spec["seed"], auto_reset=spec["auto_reset"])
    try:
        gpu.init()
        for s in range(spec["steps"]):
            d, a = S.seek_policy(gpu, s, W)
            gpu.set_actions(d, a); gpu.step()
        for name in S.STEP_OUTPUTS:
            x = gpu.export(name)[: len(gold[name])]
            assert np.array_equal(x.view(np.int32), gold[name].view(np.int32)), name
        assert gpu.export("match_result")[:, 1:5].sum() == spec["match_result_sum"]
    finally:
        gpu.close()


def test_raycast_stress_bit_exact():
    """BASELINE.json configs[3] at oracle-checkable size: every agent fires every step
    (reload when the magazine reads 0), LOS + lidar on."""
    W = 16
    gpu = S.GpuSim(W, auto_reset=True)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0)
    try:
        gpu.init(); orc.init()
        for s in range(150):
            d, a = S.random_actions(s, W, stress=True, magazine=orc.export("magazine"))
            for sim in (gpu, orc):
                sim.set_actions(d, a); sim.step()
            if s % 10 == 9:
                assert_same(gpu, orc, W, f"stress step {s}")
    finally:
        gpu.close(); orc.close()


def test_episode_stats_kernel():
    """k_stats (input of the multi-GPU allreduce) against its numpy mirror."""
    W = 16
    gpu = S.GpuSim(W, auto_reset=True)
    try:
        gpu.init()
        for s in range(40):
            d, a = S.seek_policy(gpu, s, W)
            if s == 20:
                gpu.set_export("reset", (np.arange(W) % 3 == 0).astype(np.int32).reshape(W, 1))
            gpu.set_actions(d, a); gpu.step()
            assert np.array_equal(gpu.episode_stats(), S.episode_stats_host(gpu)), f"step {s}"
    finally:
        gpu.close()


def test_python_manager_zero_copy_and_stream_step():
    """BASELINE.json configs[4] coupling: the Python SimManager mirror (bindings.cpp) --
    action tensors aliased zero-copy by torch (DLPack / __cuda_array_interface__),
    stream-ordered step, and gpuStreamStep-style buffer copies (mgr.cpp:614-645)."""
    import ctypes as C
    import torch
    import madrona_mp_env as mp
    W = 8
    n = W * 12
    mgr = mp.SimManager(mp.ExecMode.CUDA, 0, W, 10, False, 0, mp.Task.Zone, 6, 0, 1, S.SCENE)
    orc = S.Oracle(W, det=True, num_threads=0)
    try:
        act = mgr.pvp_discrete_action_tensor().to_torch()
        aim = torch.from_dlpack(mgr.pvp_discrete_aim_action_tensor())
        assert act.shape == (n, 4) and act.dtype == torch.int32 and aim.shape == (n, 2)
        assert act.data_ptr() == mgr.pvp_discrete_action_tensor().ptr, "zero copy"
        ti = mgr.train_interface()
        assert list(ti["inputs"]) == ["discrete", "aim", "resets", "sim_ctrl", "policy_assignments"]
        assert ti["outputs"]["self"].shape == (n, 43) and ti["outputs"]["fwd_lidar"].shape == (n, 2, 32, 4)
        assert ti["outputs"]["episode_results"].shape == (W, 30)
        mgr.init(); orc.init()
        stream = torch.cuda.Stream()
        self_obs = mgr.self_observation_tensor().to_torch()
        for s in range(25):
            d, a = S.random_actions(s, W)
            with torch.cuda.stream(stream):
                act.copy_(torch.from_numpy(d).cuda(), non_blocking=True)
                aim.copy_(torch.from_numpy(a).cuda(), non_blocking=True)
                mgr.step_async(stream.cuda_stream)
            stream.synchronize()
            orc.set_actions(d, a); orc.step()
        assert np.array_equal(self_obs.cpu().numpy().view(np.int32), orc.export("self_obs").view(np.int32))
        # gpuStreamStep: caller-owned device buffers in train-interface order
        lib, h = mgr._lib, mgr._h
        arr = C.POINTER(mp._NamedTensor)()
        cnt = C.c_uint32()
        lib.mpenv_train_interface(h, C.byref(arr), C.byref(cnt))
        bufs, keep = (C.c_void_p * cnt.value)(), {}
        for i in range(cnt.value):
            name = arr[i].name.decode()
            t = (ti["inputs"] if arr[i].is_input else ti["outputs"])[name]
            if name in ("agent_map", "unmasked_agent_map"):
                bufs[i] = None
                continue
            buf = torch.zeros(t.shape, dtype=torch.int32 if t.typestr == "<i4" else torch.float32, device="cuda")
            if arr[i].is_input:
                buf.copy_(t.to_torch())
            keep[name] = buf
            bufs[i] = buf.data_ptr()
        d, a = S.random_actions(25, W)
        keep["discrete"].copy_(torch.from_numpy(d).cuda()); keep["aim"].copy_(torch.from_numpy(a).cuda())
        torch.cuda.synchronize()
        lib.mpenv_stream_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        assert lib.mpenv_stream_step(h, C.c_void_p(stream.cuda_stream), bufs) == 0
        stream.synchronize()
        orc.set_actions(d, a); orc.step()
        def check_outputs(tag):
            for ours, theirs in (("self", "self_obs"), ("rewards", "reward"), ("fwd_lidar", "fwd_lidar"),
                                 ("opponents_last_known", "opponent_last_known_obs"), ("episode_results", "match_result")):
                x = keep[ours].cpu().numpy().reshape(orc.export(theirs).shape)
                assert np.array_equal(x.view(np.int32), orc.export(theirs).view(np.int32)), (tag, ours)
        check_outputs("mpenv_stream_step")
        # the XLA custom-call entry (bindings.cpp:149-158): void(stream, buffers, opaque, opaque_len),
        # the handle pointer carried in `opaque`
        d, a = S.random_actions(26, W)
        keep["discrete"].copy_(torch.from_numpy(d).cuda()); keep["aim"].copy_(torch.from_numpy(a).cuda())
        torch.cuda.synchronize()
        lib.mpenv_xla_step.restype = None
        lib.mpenv_xla_step.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
        opaque = C.string_at(C.byref(C.c_void_p(h.value)), 8)
        lib.mpenv_xla_step(C.c_void_p(stream.cuda_stream), bufs, opaque, 8)
        stream.synchronize()
        orc.set_actions(d, a); orc.step()
        check_outputs("mpenv_xla_step")
    finally:
        orc.close()
        del mgr


def test_pipelined_readback_equals_sync_read():
    """mpenv_read_tensors_pipelined: the snapshot taken after step t must survive step t+1."""
    import ctypes as C
    W = 16
    a_, b_ = S.GpuSim(W, auto_reset=True), S.GpuSim(W, auto_reset=True)
    try:
        L, h = a_.lib, a_.h
        L.mpenv_read_tensors_pipelined.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        L.mpenv_read_wait.argtypes = [C.c_void_p]
        L.mpenv_stream.restype = C.c_void_p
        L.mpenv_stream.argtypes = [C.c_void_p]
        names = ["self_obs", "fwd_lidar", "opponent_last_known_obs", "reward", "match_result"]
        bufs = {n: np.zeros((S.rows_of(S.EXPORT_SHAPES[n][0], W), S.EXPORT_SHAPES[n][1]), S.EXPORT_SHAPES[n][2])
                for n in names}
        ids = (C.c_uint32 * len(names))(*[S.EXPORT_IDS[n] for n in names])
        dst = (C.c_void_p * len(names))(*[bufs[n].ctypes.data for n in names])
        a_.init(); b_.init()
        prev = None
        for s in range(12):
            d, act = S.seek_policy(b_, s, W)
            a_.set_actions(d, act); b_.set_actions(d, act)
            assert L.mpenv_step_async(h, C.c_void_p(L.mpenv_stream(h))) == 0
            assert L.mpenv_read_wait(h) == 0
            if prev is not None:          # host buffers now hold step s-1, although step s is already queued
                for n in names:
                    assert np.array_equal(bufs[n].view(np.int32), prev[n].view(np.int32)), (s, n)
            assert L.mpenv_read_tensors_pipelined(h, ids, dst, len(names)) == 0
            b_.step()
            prev = {n: b_.export(n) for n in names}
        assert L.mpenv_read_wait(h) == 0
        for n in names:
            assert np.array_equal(bufs[n].view(np.int32), prev[n].view(np.int32)), n
    finally:
        a_.close(); b_.close()


def test_soak_256_worlds_grazing_wall():
    """256 worlds x 520 steps, seek / random policy phases as tools/soak.py: at step 462 world 165 holds the
    grazing-wall case of test_hostsim_parity.py::test_grazing_wall_bounded_sweep_regression (bounded sweeps
    must detect it and fall back to the reference's unbounded loop)."""
    W = 256
    gpu = S.GpuSim(W, auto_reset=True, full_team=False)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0)
    try:
        gpu.init(); orc.init()
        for s in range(520):
            d, a = S.seek_policy(orc, s, W) if (s // 100) % 2 == 0 else S.random_actions(s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a); sim.step()
            if s % 50 == 49 or 460 <= s <= 464:
                assert_same(gpu, orc, W, f"step {s}", full_team=False)
    finally:
        gpu.close(); orc.close()


def test_episode_end_with_random_actions_gpu():
    """See test_hostsim_parity.py::test_episode_end_with_random_actions: a whole 3000-step episode and its auto-reset."""
    W = 8
    gpu = S.GpuSim(W, auto_reset=True, full_team=False)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0)
    try:
        gpu.init(); orc.init()
        for s in range(3004):
            d, a = S.random_actions(s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a); sim.step()
            if s % 500 == 499 or s >= 2996:
                assert_same(gpu, orc, W, f"step {s}", full_team=False)
    finally:
        gpu.close(); orc.close()


@pytest.mark.parametrize("W", [1, 3, 37, 700])
def test_lidar_packets_vs_oracle(W):
    """k_lidar_packet (five 32-ray packets per 2-agent task, tasks drawn from a device-wide counter,
    lidar_packet.cuh) against the oracle: warps with no task, one task, an odd number of
    tasks and many of them."""
    gpu = S.GpuSim(W, auto_reset=True, full_team=False)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0)
    try:
        gpu.init(); orc.init()
        for step in range(40 if W <= 37 else 12):
            d, a = S.random_actions(step, W)
            for s in (gpu, orc):
                s.set_actions(d, a)
                s.step()
            for name in ("fwd_lidar", "rear_lidar"):
                assert np.array_equal(gpu.export(name).view(np.int32), orc.export(name).view(np.int32)), \
                    f"step {step}: {name}"
            if step % 10 == 9:       # the forked graph (no full-team tensors): every output, not only lidar
                assert_same(gpu, orc, W, f"step {step}", full_team=False)
    finally:
        gpu.close(); orc.close()


def test_large_map_fallback_kernels(monkeypatch):
    """MPENV_KERNELS=legacy: the per-thread k_los / k_lidar of step_kernels.cuh on the
    reference-format BVH -- the path mpenv_create takes by itself for a map whose BVH does
    not fit in shared memory (mpenv_capi.cu) -- against the oracle."""
    monkeypatch.setenv("MPENV_KERNELS", "legacy")
    W = 8
    gpu = S.GpuSim(W, auto_reset=True)
    monkeypatch.delenv("MPENV_KERNELS")
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0)
    try:
        gpu.init(); orc.init()
        for s in range(120):
            d, a = S.seek_policy(orc, s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a); sim.step()
            if s % 10 == 9:
                assert_same(gpu, orc, W, f"legacy step {s}")
    finally:
        gpu.close(); orc.close()


@pytest.mark.parametrize("pillars,los_warps,lidar_ctas", [(40, 24, 2), (110, 12, 1), (200, 0, 0), (400, 0, 0)])
def test_larger_maps(tmp_path, pillars, los_warps, lidar_ctas):
    """simple_map plus N pillars (12 triangles each): the kernels mpenv_create picks as the BVH outgrows
    shared memory -- two packet-lidar CTAs per SM and a 24-warp visibility CTA (732 triangles), one
    32-warp packet CTA and 12 visibility warps (1572), the per-thread kernels on the reference-format
    BVH in shared memory (2652) and in L2 (5052) -- each against the oracle on the same map."""
    scene = S.make_scene(str(tmp_path / "pillars"), extra_boxes=S.pillar_boxes(pillars))
    W = 8
    # without the full-team tensors the graphs fork (lidar beside visibility -> observations, mpenv_capi.cu):
    # covered on one packet and one per-thread kernel selection
    full_team = pillars in (40, 200)
    gpu = S.GpuSim(W, auto_reset=True, scene=scene, full_team=full_team)
    orc = S.Oracle(W, auto_reset=True, det=True, num_threads=0, scene=scene)
    try:
        assert gpu.kernel_variants() == (los_warps, lidar_ctas)
        gpu.init(); orc.init()
        assert_same(gpu, orc, W, "init", full_team=full_team)
        for s in range(100):
            d, a = S.seek_policy(orc, s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a); sim.step()
            if s % 10 == 9:
                assert_same(gpu, orc, W, f"{pillars} pillars, step {s}", full_team=full_team)
        assert gpu.damaged_guards() == 0
    finally:
        gpu.close(); orc.close()


@pytest.mark.parametrize("variant,flags", [("north", 0), ("south", 128), ("sky", 8)])
def test_zone_capture_defend_gpu(tmp_path, variant, flags):
    """Task::ZoneCaptureDefend on the synthetic 4-zone map: defender win at 8 points,
    attacker win on the first point, "all attackers have died" -- see test_hostsim_parity.py."""
    from test_hostsim_parity import zone_capture_defend_rollout
    outcomes = zone_capture_defend_rollout(
        lambda W, f, scene: S.GpuSim(W, sim_flags=f, auto_reset=True, scene=scene, task=4),
        str(tmp_path / "four_zones"), variant, flags, assert_same)
    expect = {"north": (False, 8), "south": (True, 1), "sky": (False, 0)}[variant]
    assert expect in outcomes, f"{variant}: outcomes {outcomes}"


def test_zone_capture_defend_needs_four_zones():
    with pytest.raises(RuntimeError, match="at least 4 zones"):
        S.GpuSim(2, task=4)


def test_subzones_gpu():
    """SimFlags::SubZones: subzoneSystem + subzoneRewardSystem, agents re-assigned to the
    sub-zones they stand in through the live policy tensor -- see test_hostsim_parity.py."""
    from test_hostsim_parity import subzones_rollout
    subzones_rollout(lambda W, f: S.GpuSim(W, sim_flags=f, auto_reset=True), assert_same)


def test_double_buffered_snapshots():
    """mpenv_snapshot_tensors_async / mpenv_read_snapshot_async in the bench's order: the snapshot of
    step s is queued while the transfer of step s-1 may still be running, with ONE set of host buffers."""
    import ctypes as C
    W = 16
    a_, b_ = S.GpuSim(W, auto_reset=True), S.GpuSim(W, auto_reset=True)
    try:
        L, h = a_.lib, a_.h
        L.mpenv_snapshot_tensors_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.mpenv_read_snapshot_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.mpenv_read_wait.argtypes = [C.c_void_p]
        L.mpenv_stream.restype = C.c_void_p
        L.mpenv_stream.argtypes = [C.c_void_p]
        names = ["self_obs", "fwd_lidar", "opponent_last_known_obs", "reward", "match_result"]
        bufs = {n: np.zeros((S.rows_of(S.EXPORT_SHAPES[n][0], W), S.EXPORT_SHAPES[n][1]), S.EXPORT_SHAPES[n][2])
                for n in names}
        ids = (C.c_uint32 * len(names))(*[S.EXPORT_IDS[n] for n in names])
        dst = (C.c_void_p * len(names))(*[bufs[n].ctypes.data for n in names])
        a_.init(); b_.init()
        prev = None
        for s in range(14):
            d, act = S.seek_policy(b_, s, W)
            a_.set_actions(d, act); b_.set_actions(d, act)
            assert L.mpenv_step_async(h, C.c_void_p(L.mpenv_stream(h))) == 0
            assert L.mpenv_snapshot_tensors_async(h, ids, len(names)) == 0
            assert L.mpenv_read_wait(h) == 0
            if prev is not None:          # host buffers hold step s-1 although step s and its snapshot are queued
                for n in names:
                    assert np.array_equal(bufs[n].view(np.int32), prev[n].view(np.int32)), (s, n)
            assert L.mpenv_read_snapshot_async(h, dst, len(names)) == 0
            b_.step()
            prev = {n: b_.export(n) for n in names}
        assert L.mpenv_read_wait(h) == 0
        for n in names:
            assert np.array_equal(bufs[n].view(np.int32), prev[n].view(np.int32)), n
        # a read without a pending snapshot is an error, not a stale copy
        assert L.mpenv_read_snapshot_async(h, dst, len(names)) != 0
    finally:
        a_.close(); b_.close()
