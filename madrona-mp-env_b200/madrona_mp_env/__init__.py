"""madrona_mp_env -- host-side mirror of the reference's Python binding
(reference src/bindings.cpp:11-160: module `madrona_mp_env`, class SimManager)
over the B200-native C ABI (include/mpenv.h, libmpenv_b200.so).

Same constructor keywords, method names and tensor getters as the reference,
so `scripts/jax_train.py`-style callers can switch module and keep going.
Tensors alias the simulator's device memory (zero copy): they expose
`__cuda_array_interface__`, `.to_torch()` and `__dlpack__`, as
madrona::py::Tensor does.  There is no CPU execution mode: constructing a
SimManager without a CUDA device raises.
"""
import ctypes as C
import enum
import os

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB_PATH = os.path.join(_PKG_DIR, "libmpenv_b200.so")


class Task(enum.IntEnum):           # bindings.cpp:14-20
    Explore = 0
    TDM = 1
    Zone = 2
    Turret = 3
    ZoneCaptureDefend = 4


class SimFlags(enum.IntFlag):       # bindings.cpp:22-36
    Default = 0
    SpawnInMiddle = 1 << 0
    RandomizeHPMagazine = 1 << 1
    NavmeshSpawn = 1 << 2
    NoRespawn = 1 << 3
    StaggerStarts = 1 << 4
    EnableCurriculum = 1 << 5
    HardcodedSpawns = 1 << 6
    RandomFlipTeams = 1 << 7
    StaticFlipTeams = 1 << 8
    FullTeamPolicy = 1 << 9
    SimEvalMode = 1 << 10
    SubZones = 1 << 11


class ExecMode(enum.IntEnum):       # madrona.ExecMode
    CPU = 0
    CUDA = 1


class _Config(C.Structure):
    _fields_ = [("gpu_id", C.c_int32), ("num_worlds", C.c_uint32), ("rand_seed", C.c_uint32),
                ("auto_reset", C.c_int32), ("sim_flags", C.c_uint32), ("task_type", C.c_uint32),
                ("team_size", C.c_uint32), ("num_pbt_policies", C.c_uint32),
                ("policy_history_size", C.c_uint32), ("scene_dir", C.c_char_p),
                ("world_id_offset", C.c_uint32), ("enable_full_team", C.c_uint32),
                ("event_log_dir", C.c_char_p), ("record_log_path", C.c_char_p),
                ("replay_log_path", C.c_char_p), ("curriculum_data_path", C.c_char_p),
                ("train_flank", C.c_uint32)]


class _TensorDesc(C.Structure):
    _fields_ = [("dev_ptr", C.c_void_p), ("dtype", C.c_int32), ("ndim", C.c_int32),
                ("dims", C.c_int64 * 4), ("gpu_id", C.c_int32)]


class _NamedTensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("export_id", C.c_uint32), ("is_input", C.c_int32)]


_lib = None


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise ImportError(f"{_LIB_PATH} not built (run __graft_entry__.build()); there is no fallback")
        L = C.CDLL(_LIB_PATH)
        L.mpenv_create.argtypes = [C.POINTER(_Config), C.POINTER(C.c_void_p)]
        for f in ("mpenv_destroy", "mpenv_init", "mpenv_step", "mpenv_sync"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.mpenv_step_async.argtypes = [C.c_void_p, C.c_void_p]
        L.mpenv_init_async.argtypes = [C.c_void_p, C.c_void_p]
        L.mpenv_tensor.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(_TensorDesc)]
        L.mpenv_last_error.restype = C.c_char_p
        L.mpenv_last_error.argtypes = [C.c_void_p]
        L.mpenv_train_interface.argtypes = [C.c_void_p, C.POINTER(C.POINTER(_NamedTensor)),
                                            C.POINTER(C.c_uint32)]
        L.mpenv_trigger_reset.argtypes = [C.c_void_p, C.c_int32]
        L.mpenv_episode_stats_async.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
        _lib = L
    return _lib


class Tensor:
    """madrona::py::Tensor stand-in: a typed view of simulator-owned device memory."""

    def __init__(self, ptr, dtype, shape, gpu_id, owner):
        self.ptr, self.shape, self.gpu_id = ptr, tuple(int(s) for s in shape), gpu_id
        self.typestr = "<i4" if dtype == 0 else "<f4"
        self._owner = owner          # keeps the SimManager (and its memory) alive

    @property
    def __cuda_array_interface__(self):
        return {"shape": self.shape, "typestr": self.typestr, "data": (self.ptr, False),
                "version": 3, "strides": None}

    def to_torch(self):
        import torch
        with torch.cuda.device(self.gpu_id):
            t = torch.as_tensor(self, device=f"cuda:{self.gpu_id}")
        t._mpenv_owner = self._owner
        return t

    def __dlpack__(self, stream=None):
        return self.to_torch().__dlpack__(stream=stream)

    def __dlpack_device__(self):
        return (2, self.gpu_id)      # kDLCUDA


# export ids (reference ExportID, src/sim.hpp:15-66)
_EX = dict(reset=0, world_curriculum=1, pvp_discrete_action=3, pvp_aim_action=4,
           pvp_discrete_aim_action=5, reward=6, done=7, match_result=8, agent_policy=9,
           self_obs=10, teammate_obs=11, opponent_obs=12, opponent_last_known_obs=13, self_pos=14,
           teammate_pos=15, opponent_pos=16, opponent_last_known_pos=17, opponent_masks=18,
           fwd_lidar=19, rear_lidar=20, agent_map=21, unmasked_agent_map=22, hp=23, alive=24,
           magazine=25, full_team_global=27, full_team_players=28, full_team_enemies=29,
           full_team_last_known_enemies=30, full_team_fwd_lidar=31, full_team_rear_lidar=32,
           full_team_reward=33, full_team_done=34, filters_state=38, reward_hyper_params=39,
           sim_control=1000)


class SimManager:
    """bindings.cpp:38-147.  exec_mode must be CUDA."""

    def __init__(self, exec_mode, gpu_id, num_worlds, rand_seed, auto_reset, sim_flags, task_type,
                 team_size, num_pbt_policies, policy_history_size, scene_path, train_flank=False,
                 replay_log_path=None, record_log_path=None, event_log_path=None,
                 curriculum_data_path=None, world_id_offset=0, enable_full_team=True):
        if int(exec_mode) != int(ExecMode.CUDA):
            raise ValueError("madrona_mp_env (B200): only ExecMode.CUDA exists; there is no CPU path")
        self._lib = _load()
        self._scene = os.fsencode(scene_path)
        # mgr.hpp:47-49; the C ABI copies the strings
        enc = lambda v: None if v is None else os.fsencode(v)
        cfg = _Config(int(gpu_id), int(num_worlds), int(rand_seed), int(bool(auto_reset)),
                      int(sim_flags), int(task_type), int(team_size), int(num_pbt_policies),
                      int(policy_history_size), self._scene, int(world_id_offset), int(enable_full_team),
                      enc(event_log_path), enc(record_log_path), enc(replay_log_path),
                      enc(curriculum_data_path), int(bool(train_flank)))
        self._h = C.c_void_p()
        rc = self._lib.mpenv_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            raise RuntimeError(f"SimManager: {self._lib.mpenv_last_error(None).decode()}")
        self.gpu_id, self.num_worlds = int(gpu_id), int(num_worlds)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.mpenv_destroy(h)
            self._h = None

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(self._lib.mpenv_last_error(self._h).decode())

    def init(self):
        self._check(self._lib.mpenv_init(self._h))

    def step(self):
        self._check(self._lib.mpenv_step(self._h))

    def step_async(self, cuda_stream):
        """Stream-ordered step (gpuStreamStep without the buffer copies)."""
        self._check(self._lib.mpenv_step_async(self._h, C.c_void_p(int(cuda_stream))))

    def init_async(self, cuda_stream):
        self._check(self._lib.mpenv_init_async(self._h, C.c_void_p(int(cuda_stream))))

    def sync(self):
        self._check(self._lib.mpenv_sync(self._h))

    def trigger_reset(self, world_idx):
        self._check(self._lib.mpenv_trigger_reset(self._h, int(world_idx)))

    def episode_stats(self, cuda_stream=0):
        """Device int64[16] of per-GPU episode statistics (input of the allreduce)."""
        out = C.c_void_p()
        self._check(self._lib.mpenv_episode_stats_async(self._h, C.c_void_p(int(cuda_stream)), C.byref(out)))
        import torch
        class _Raw:
            __cuda_array_interface__ = {"shape": (16,), "typestr": "<i8", "data": (out.value, False),
                                        "version": 3, "strides": None}
        t = torch.as_tensor(_Raw(), device=f"cuda:{self.gpu_id}")
        t._mpenv_owner = self
        return t

    def _tensor(self, key):
        d = _TensorDesc()
        self._check(self._lib.mpenv_tensor(self._h, _EX[key], C.byref(d)))
        return Tensor(d.dev_ptr, d.dtype, [d.dims[i] for i in range(d.ndim)], d.gpu_id, self)

    # tensor getters, bindings.cpp:112-146
    def reset_tensor(self): return self._tensor("reset")
    def sim_control_tensor(self): return self._tensor("sim_control")
    def pvp_discrete_action_tensor(self): return self._tensor("pvp_discrete_action")
    def pvp_aim_action_tensor(self): return self._tensor("pvp_aim_action")
    def pvp_discrete_aim_action_tensor(self): return self._tensor("pvp_discrete_aim_action")
    def reward_tensor(self): return self._tensor("reward")
    def done_tensor(self): return self._tensor("done")
    def policy_assignment_tensor(self): return self._tensor("agent_policy")
    def world_curriculum_tensor(self): return self._tensor("world_curriculum")
    def match_result_tensor(self): return self._tensor("match_result")
    def self_observation_tensor(self): return self._tensor("self_obs")
    def filters_state_observation_tensor(self): return self._tensor("filters_state")
    def teammate_observations_tensor(self): return self._tensor("teammate_obs")
    def opponent_observations_tensor(self): return self._tensor("opponent_obs")
    def opponent_last_known_observations_tensor(self): return self._tensor("opponent_last_known_obs")
    def self_position_tensor(self): return self._tensor("self_pos")
    def teammate_position_observations_tensor(self): return self._tensor("teammate_pos")
    def opponent_position_observations_tensor(self): return self._tensor("opponent_pos")
    def opponent_last_known_position_observations_tensor(self): return self._tensor("opponent_last_known_pos")
    def opponent_masks_tensor(self): return self._tensor("opponent_masks")
    def fwd_lidar_tensor(self): return self._tensor("fwd_lidar")
    def rear_lidar_tensor(self): return self._tensor("rear_lidar")
    def agent_map_tensor(self): return self._tensor("agent_map")
    def unmasked_agent_map_tensor(self): return self._tensor("agent_map")   # bindings.cpp:135
    def hp_tensor(self): return self._tensor("hp")
    def magazine_tensor(self): return self._tensor("magazine")
    def alive_tensor(self): return self._tensor("alive")
    def reward_hyper_params_tensor(self): return self._tensor("reward_hyper_params")
    def full_team_global_observations_tensor(self): return self._tensor("full_team_global")
    def full_team_player_observations_tensor(self): return self._tensor("full_team_players")
    def full_team_enemy_observations_tensor(self): return self._tensor("full_team_enemies")
    def full_team_last_known_enemy_observations_tensor(self): return self._tensor("full_team_last_known_enemies")
    def full_team_fwd_lidar_tensor(self): return self._tensor("full_team_fwd_lidar")
    def full_team_rear_lidar_tensor(self): return self._tensor("full_team_rear_lidar")
    def full_team_reward_tensor(self): return self._tensor("full_team_reward")
    def full_team_done_tensor(self): return self._tensor("full_team_done")

    def train_interface(self):
        """Manager::trainInterface (mgr.cpp:2383-2431) as {'inputs': {...}, 'outputs': {...}}."""
        arr = C.POINTER(_NamedTensor)()
        n = C.c_uint32()
        self._check(self._lib.mpenv_train_interface(self._h, C.byref(arr), C.byref(n)))
        out = {"inputs": {}, "outputs": {}}
        for i in range(n.value):
            d = _TensorDesc()
            self._check(self._lib.mpenv_tensor(self._h, arr[i].export_id, C.byref(d)))
            t = Tensor(d.dev_ptr, d.dtype, [d.dims[k] for k in range(d.ndim)], d.gpu_id, self)
            out["inputs" if arr[i].is_input else "outputs"][arr[i].name.decode()] = t
        return out
