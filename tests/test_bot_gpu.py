"""SURVEY 8(f-3) on the B200: the scripted A* bot (policy id -1) fused into k_move,
against the reference's planAStarAISystem / applyBotActionsSystem in the oracle."""
import numpy as np
import pytest

import simlib as S

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("bots", ["all", "team_b"])
def test_bot_gpu_matches_oracle(bots):
    W, steps = 8, 300
    mask = np.ones((W, 12), bool)
    if bots == "team_b":
        mask[:, :6] = False
    gpu = S.GpuSim(W, auto_reset=True)
    orc = S.Oracle(W, auto_reset=True, det=True)
    try:
        gpu.init(); orc.init()
        pol = np.where(mask, -1, 0).astype(np.int32).reshape(-1, 1)
        for sim in (gpu, orc):
            sim.set_export("agent_policy", pol)
        for s in range(steps):
            d, a = S.seek_policy(orc, s, W)
            for sim in (gpu, orc):
                sim.set_actions(d, a)
                sim.step()
            for name in S.STEP_OUTPUTS + ["pvp_discrete_action", "pvp_aim_action"]:
                x, y = gpu.export(name), orc.export(name)
                assert np.array_equal(x.view(np.int32), y.view(np.int32)), f"step {s}: {name}"
            if s % 20 == 19:
                for w in range(W):
                    dd = S.diff_world_state(gpu.world_state(w), orc.world_state(w))
                    assert not dd, f"step {s} world {w}: {dd[:6]}"
        acts = orc.export("pvp_discrete_action").reshape(W, 12, 4)
        assert (acts[mask][:, 0] > 0).any(), "no bot ever moved"
    finally:
        gpu.close(); orc.close()
