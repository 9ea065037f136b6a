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
        return orc.export("match_result")
    finally:
        gpu.close(); orc.close()


def random_policy(sim, s, W):
    return S.random_actions(s, W)


def test_random_actions_32_worlds_bit_exact():
    # BASELINE.json configs[0], shortened: 32 worlds, random actions
    rollout(32, 200, random_policy, check_every=20)


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
        # the first 32 worlds of the big batch equal a 32-world oracle run
        orc = S.Oracle(32, auto_reset=True, det=True, num_threads=0)
        orc.init()
        for s in range(20):
            d, a = S.random_actions(s, 32)
            orc.set_actions(d, a); orc.step()
        for name in ("self_obs", "fwd_lidar", "reward", "opponent_obs"):
            x, y = gpu.export(name), orc.export(name)
            assert np.array_equal(x[:len(y)].view(np.int32), y.view(np.int32)), name
        orc.close()
    finally:
        gpu.close()
