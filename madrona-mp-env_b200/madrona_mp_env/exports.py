"""Export tables of the step path: the reference's ExportID order (src/sim.hpp:15-66)
and the per-row layout of every exported tensor (src/mgr.cpp:1965-2381), plus the
train-interface output list (src/mgr.cpp:2398-2431).  Shared by the package, bench.py
and the tests."""
import os

import numpy as np

AGENTS = 12
SCENE = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))),
                     "data", "simple_map")

# reference ExportID (src/sim.hpp:15-66)
EXPORT_IDS = dict(
    reset=0, world_curriculum=1, explore_action=2, pvp_discrete_action=3, pvp_aim_action=4,
    pvp_discrete_aim_action=5, reward=6, done=7, match_result=8, agent_policy=9,
    self_obs=10, teammate_obs=11, opponent_obs=12, opponent_last_known_obs=13,
    self_pos=14, teammate_pos=15, opponent_pos=16, opponent_last_known_pos=17,
    opponent_masks=18, fwd_lidar=19, rear_lidar=20, agent_map=21, unmasked_agent_map=22,
    hp=23, alive=24, magazine=25, full_team_actions=26, full_team_global=27,
    full_team_players=28, full_team_enemies=29, full_team_last_known_enemies=30,
    full_team_fwd_lidar=31, full_team_rear_lidar=32, full_team_reward=33, full_team_done=34,
    full_team_policy=35, event_log=36, packed_step_snapshot=37, filters_state_obs=38,
    reward_hyper_params=39,
)
# per-row float/int32 counts and dtype; rows: 'agent' (W*12), 'world' (W), 'team' (W*2)
EXPORT_SHAPES = dict(
    reset=("world", 1, np.int32), world_curriculum=("world", 1, np.int32),
    pvp_discrete_action=("agent", 4, np.int32), pvp_aim_action=("agent", 2, np.float32),
    pvp_discrete_aim_action=("agent", 2, np.int32), reward=("agent", 1, np.float32),
    done=("agent", 1, np.int32), match_result=("world", 30, np.int32),
    agent_policy=("agent", 1, np.int32), self_obs=("agent", 43, np.float32),
    teammate_obs=("agent", 5 * 32, np.float32), opponent_obs=("agent", 6 * 32, np.float32),
    opponent_last_known_obs=("agent", 6 * 32, np.float32), self_pos=("agent", 3, np.float32),
    teammate_pos=("agent", 15, np.float32), opponent_pos=("agent", 18, np.float32),
    opponent_last_known_pos=("agent", 18, np.float32), opponent_masks=("agent", 6, np.float32),
    fwd_lidar=("agent", 2 * 32 * 4, np.float32), rear_lidar=("agent", 2 * 8 * 4, np.float32),
    hp=("agent", 1, np.float32), alive=("agent", 1, np.float32), magazine=("agent", 2, np.int32),
    full_team_global=("team", 16, np.float32), full_team_players=("team", 6 * 28, np.float32),
    full_team_enemies=("team", 6 * 33, np.float32),
    full_team_last_known_enemies=("team", 6 * 24, np.float32),
    full_team_fwd_lidar=("team", 6 * 256, np.float32), full_team_rear_lidar=("team", 6 * 64, np.float32),
    full_team_reward=("team", 1, np.float32), full_team_done=("team", 1, np.int32),
    filters_state_obs=("agent", 1, np.float32), reward_hyper_params=("agent", 9, np.float32),
)
# the train-interface outputs (mgr.cpp:2398-2431) + exported state
STEP_OUTPUTS = [
    "self_obs", "teammate_obs", "opponent_obs", "opponent_last_known_obs", "self_pos",
    "teammate_pos", "opponent_pos", "opponent_last_known_pos", "opponent_masks", "fwd_lidar",
    "rear_lidar", "hp", "alive", "magazine", "filters_state_obs", "reward", "done", "match_result",
    "reward_hyper_params",
]
FULL_TEAM_OUTPUTS = [
    "full_team_global", "full_team_players", "full_team_enemies", "full_team_last_known_enemies",
    "full_team_fwd_lidar", "full_team_rear_lidar", "full_team_reward", "full_team_done",
]


def rows_of(kind, num_worlds):
    return {"agent": num_worlds * AGENTS, "world": num_worlds, "team": num_worlds * 2}[kind]

