"""CPU checks of the drop-in boundary (no GPU compute): libmpenv_b200.so loads and
exports every symbol include/mpenv.h declares, the host-side asset section works
without a device, error behaviour, and the Python mirror of the reference binding
(bindings.cpp:11-160) keeps its names and refuses to run without CUDA."""
import ctypes as C
import inspect
import os
import re

import numpy as np
import pytest

import simlib as S

HEADER = os.path.join(S.ROOT, "include", "mpenv.h")


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(S.PRODUCT_SO):
        import __graft_entry__ as g
        g.build()
    return S.load_product()


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?[A-Za-z_][A-Za-z0-9_]*\s*\**\s*\*?\s*(mpenv_[a-z0-9_]+)\s*\(", src, flags=re.M)
    return sorted(set(names))


def test_every_declared_symbol_is_exported(lib):
    names = declared_functions()
    assert len(names) >= 30, names
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/mpenv.h but not exported"


def test_header_compiles_as_plain_c(tmp_path):
    """The boundary is a C ABI: the header must be consumable by a C compiler."""
    src = tmp_path / "t.c"
    src.write_text('#include "mpenv.h"\nint main(void){ mpenv_config c; (void)c; return MPENV_NUM_STATS == 16 ? 0 : 1; }\n')
    import subprocess
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(S.ROOT, "include"), "-c",
                        str(src), "-o", str(tmp_path / "t.o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_map_section_runs_on_host(lib, built):
    """mpenv_map_* is host-only asset code (Impl::init's loaders + our BVH builder)."""
    nodes, verts, info = S.load_map(lib)
    assert info.num_tris == 252 and info.num_zones == 3
    assert (info.num_a_spawns, info.num_b_spawns, info.num_respawns) == (8, 8, 16)
    assert np.allclose(list(info.world_bounds), [-2007.875, -2007.875, -163.979, 2007.875, 2007.875, 623.423], atol=1e-2)
    # the same builder is what the hostsim harness and the oracle were given
    n2, v2, _ = S.load_map(C.CDLL(S.HOSTSIM_SO))
    assert np.array_equal(nodes, n2) and np.array_equal(verts, v2)
    # every leaf holds 1..2 triangles; children indices in range
    ch = nodes.view(np.int32).reshape(len(nodes), 16)[:, 11:15]
    tri_size = nodes[:, 16:20]
    for n in range(len(nodes)):
        for k in range(4):
            c = int(ch[n, k])
            if c == -1:
                continue
            if c < 0:
                assert 1 <= tri_size[n, k] <= 2 and (c & 0x7FFFFFFF) + tri_size[n, k] <= info.num_tris
            else:
                assert 0 < c < len(nodes)


def test_map_load_errors(lib, tmp_path):
    m = C.c_void_p()
    lib.mpenv_map_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.mpenv_map_last_error.restype = C.c_char_p
    assert lib.mpenv_map_load(str(tmp_path).encode(), C.byref(m)) == -2          # MPENV_ERR_IO
    assert b"collisions.bin" in lib.mpenv_map_last_error()
    (tmp_path / "collisions.bin").write_bytes(b"\x00" * 10)                      # truncated
    assert lib.mpenv_map_load(str(tmp_path).encode(), C.byref(m)) == -2


def test_create_argument_errors(lib):
    h = C.c_void_p()
    scene = S.SCENE.encode()
    bad = [
        S.MpenvConfig(0, 4, 10, 0, 0, 1, 6, 0, 1, scene, 0, 1, None, None, None, None, 0),            # TDM: unsupported task
        S.MpenvConfig(0, 4, 10, 0, 0, 2, 4, 0, 1, scene, 0, 1, None, None, None, None, 0),            # team size
        S.MpenvConfig(0, 0, 10, 0, 0, 2, 6, 0, 1, scene, 0, 1, None, None, None, None, 0),            # zero worlds
        S.MpenvConfig(0, 4, 10, 0, 1 << 6, 2, 6, 0, 1, scene, 0, 1, None, None, None, None, 0),       # HardcodedSpawns flag
    ]
    for cfg in bad:
        assert lib.mpenv_create(C.byref(cfg), C.byref(h)) == -1           # MPENV_ERR_INVALID
        assert lib.mpenv_last_error(None)


def _have_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_cuda(), reason="checks the no-device behaviour")
def test_no_cuda_fails_loudly(lib):
    """There is no CPU execution path: without a device create() must fail."""
    h = C.c_void_p()
    cfg = S.MpenvConfig(0, 4, 10, 0, 0, 2, 6, 0, 1, S.SCENE.encode(), 0, 1, None, None, None, None, 0)
    assert lib.mpenv_create(C.byref(cfg), C.byref(h)) == -3               # MPENV_ERR_CUDA
    assert b"no usable CUDA device" in lib.mpenv_last_error(None)
    import madrona_mp_env as mp
    with pytest.raises(RuntimeError):
        mp.SimManager(mp.ExecMode.CUDA, 0, 4, 10, False, 0, mp.Task.Zone, 6, 0, 1, S.SCENE)
    with pytest.raises(ValueError):
        mp.SimManager(mp.ExecMode.CPU, 0, 4, 10, False, 0, mp.Task.Zone, 6, 0, 1, S.SCENE)


def test_python_mirror_surface():
    """Names of bindings.cpp:14-36 (enums), :93-108 (ctor keywords), :109-146 (methods)."""
    import madrona_mp_env as mp
    assert [t.name for t in mp.Task] == ["Explore", "TDM", "Zone", "Turret", "ZoneCaptureDefend"]
    assert mp.SimFlags.SubZones == 1 << 11 and mp.SimFlags.StaggerStarts == 1 << 4
    params = list(inspect.signature(mp.SimManager.__init__).parameters)[1:17]
    assert params == ["exec_mode", "gpu_id", "num_worlds", "rand_seed", "auto_reset", "sim_flags", "task_type",
                      "team_size", "num_pbt_policies", "policy_history_size", "scene_path", "train_flank",
                      "replay_log_path", "record_log_path", "event_log_path", "curriculum_data_path"]
    for m in ("init", "step", "reset_tensor", "sim_control_tensor", "pvp_discrete_action_tensor",
              "pvp_aim_action_tensor", "pvp_discrete_aim_action_tensor", "reward_tensor", "done_tensor",
              "policy_assignment_tensor", "match_result_tensor", "self_observation_tensor",
              "filters_state_observation_tensor", "teammate_observations_tensor",
              "opponent_observations_tensor", "opponent_last_known_observations_tensor",
              "self_position_tensor", "teammate_position_observations_tensor",
              "opponent_position_observations_tensor", "opponent_last_known_position_observations_tensor",
              "opponent_masks_tensor", "fwd_lidar_tensor", "rear_lidar_tensor", "agent_map_tensor",
              "unmasked_agent_map_tensor", "hp_tensor", "magazine_tensor", "alive_tensor",
              "reward_hyper_params_tensor", "full_team_global_observations_tensor",
              "full_team_player_observations_tensor", "full_team_enemy_observations_tensor",
              "full_team_fwd_lidar_tensor", "full_team_rear_lidar_tensor", "train_interface"):
        assert callable(getattr(mp.SimManager, m)), m


def test_shard_offsets():
    from madrona_mp_env import dist as D
    assert [D.shard_world_offset(r, 8192) for r in range(8)] == [r * 8192 for r in range(8)]
    with pytest.raises(ValueError):
        D.shard_world_offset(-1, 8)
    with pytest.raises(ValueError):
        D.shard_world_offset(1 << 20, 8192)
