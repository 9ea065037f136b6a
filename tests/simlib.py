"""ctypes plumbing shared by the tests, __graft_entry__.smoke() and bench.py's
CPU-baseline leg: the oracle (reference sim code under oracle/shim), the
hostsim harness (device stage functions compiled for the host, tests only) and
the product library (libmpenv_b200.so, CUDA only).

Nothing here is imported by the product package.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENE = os.path.join(ROOT, "data", "simple_map")
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libmpenv_ref.so")
REF_DET_SO = os.path.join(ORACLE_DIR, "_ref", "libmpenv_ref_det.so")
REF_FAST_SO = os.path.join(ORACLE_DIR, "_ref", "libmpenv_ref_fast.so")    # -O3 build of REF_SO: the timed CPU baseline
DET_MATH_SO = os.path.join(ORACLE_DIR, "_ref", "libdet_math.so")
HOSTSIM_SO = os.path.join(ROOT, "tests", "hostsim", "_build", "libhostsim.so")
PRODUCT_SO = os.path.join(ROOT, "madrona-mp-env_b200", "libmpenv_b200.so")
CSRC = os.path.join(ROOT, "madrona-mp-env_b200", "csrc")
sys.path.insert(0, os.path.join(ROOT, "madrona-mp-env_b200"))

AGENTS = 12

from madrona_mp_env.exports import (EXPORT_IDS, EXPORT_SHAPES, STEP_OUTPUTS, FULL_TEAM_OUTPUTS,  # noqa: E402,F401
                                     rows_of)


# ---------------------------------------------------------------- world state
class AgentState(C.Structure):
    _fields_ = [
        ("pos", C.c_float * 3), ("rot", C.c_float * 4), ("vel", C.c_float * 3),
        ("move_new_pos", C.c_float * 3), ("move_new_vel", C.c_float * 3), ("move_max_vel", C.c_float),
        ("aim_yaw", C.c_float), ("aim_pitch", C.c_float), ("aim_rot", C.c_float * 4),
        ("hp", C.c_float), ("alive", C.c_float), ("mag_bullets", C.c_int32), ("mag_reloading", C.c_int32),
        ("dmg", C.c_float * 6), ("pose_cur", C.c_int32), ("pose_tgt", C.c_int32),
        ("pose_transition", C.c_int32), ("rng_key", C.c_uint32 * 2), ("rng_count", C.c_uint32),
        ("landed_shot_on", C.c_int32), ("remaining_respawn_steps", C.c_int32),
        ("remaining_autoheal_steps", C.c_int32), ("successful_kill", C.c_int32),
        ("was_shot_count", C.c_int32), ("was_killed", C.c_int32), ("fired_shot_t", C.c_float),
        ("in_zone", C.c_int32), ("min_dist_to_zone", C.c_float), ("has_died", C.c_int32),
        ("reloaded_full_mag", C.c_int32), ("weapon_type", C.c_int32),
        ("crumb_total_penalty", C.c_float), ("crumb_steps_since_new", C.c_int32),
        ("crumb_last", C.c_int32), ("start_pos", C.c_float * 3), ("explore_new_cells", C.c_uint32),
        ("can_see", C.c_int32 * 6), ("team", C.c_int32), ("team_offset", C.c_int32),
        ("policy", C.c_int32), ("reward_coefs", C.c_float * 9),
        ("in_sub_zone", C.c_int32), ("min_dist_to_sub_zone", C.c_float),
    ]


class CrumbState(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("penalty", C.c_float), ("team", C.c_int32),
                ("team_offset", C.c_int32)]


class WorldState(C.Structure):
    _fields_ = [
        ("team_a", C.c_int32), ("cur_step", C.c_int32), ("is_finished", C.c_int32),
        ("enable_spawn_curriculum", C.c_int32), ("cur_curriculum_tier", C.c_uint32),
        ("cur_curriculum_spawn_idx", C.c_uint32),
        ("zone_cur", C.c_int32), ("zone_controlling_team", C.c_int32), ("zone_contested", C.c_int32),
        ("zone_captured", C.c_int32), ("zone_earned_point", C.c_int32),
        ("zone_steps_remaining", C.c_int32), ("zone_steps_until_point", C.c_int32),
        ("match_result", C.c_int32 * 30), ("zone_stats", C.c_int32 * 25),
        ("base_rng_key", C.c_uint32 * 2), ("base_rng_count", C.c_uint32),
        ("cur_episode_idx", C.c_uint32), ("world_episode_counter", C.c_uint32),
        ("world_curriculum", C.c_int32), ("reset_flag", C.c_int32),
        ("filters_active", C.c_uint64 * 2), ("filters_last_matches", (C.c_int32 * 3) * 2),
        ("filters_last_matched_step", C.c_int32 * 2),
        ("goal_min_dist", C.c_float * 2), ("goal_team_step_rewards", C.c_float * 2),
        ("team_rewards", C.c_float * 2),
        ("sub_zones", (C.c_int32 * 3) * 8),
        ("spawn_a_last_used", C.c_uint32 * 128), ("spawn_b_last_used", C.c_uint32 * 128),
        ("respawn_last_used", C.c_uint32 * 128),
        ("num_breadcrumbs", C.c_int32), ("breadcrumbs", CrumbState * 192),
        ("agents", AgentState * 12),
    ]


def struct_to_dict(s):
    out = {}
    for name, _ in s._fields_:
        v = getattr(s, name)
        if isinstance(v, C.Structure):
            out[name] = struct_to_dict(v)
        elif isinstance(v, C.Array):
            if issubclass(v._type_, C.Structure):
                out[name] = [struct_to_dict(x) for x in v]
            else:
                out[name] = np.ctypeslib.as_array(v).copy()
        else:
            out[name] = v
    return out


# Fields whose float values are compared bit-for-bit by their int32 view.
def diff_world_state(a, b, path="", float_tol=0.0, skip=()):
    """Returns a list of 'path: a != b' strings for differing leaves."""
    diffs = []
    for k in a:
        p = f"{path}.{k}" if path else k
        if any(p.endswith(s) for s in skip):
            continue
        va, vb = a[k], b[k]
        if isinstance(va, dict):
            diffs += diff_world_state(va, vb, p, float_tol, skip)
        elif isinstance(va, list):
            for i, (xa, xb) in enumerate(zip(va, vb)):
                diffs += diff_world_state(xa, xb, f"{p}[{i}]", float_tol, skip)
        elif isinstance(va, np.ndarray):
            if va.dtype.kind == "f":
                if float_tol == 0.0:
                    same = np.array_equal(va.view(np.int32), vb.view(np.int32)) or np.array_equal(va, vb)
                else:
                    same = np.allclose(va, vb, rtol=float_tol, atol=float_tol)
            else:
                same = np.array_equal(va, vb)
            if not same:
                diffs.append(f"{p}: {va} != {vb}")
        elif isinstance(va, float):
            if float_tol == 0.0:
                same = (va == vb) or (va != va and vb != vb)
            else:
                same = abs(va - vb) <= float_tol * max(1.0, abs(va), abs(vb))
            if not same:
                diffs.append(f"{p}: {va!r} != {vb!r}")
        else:
            if va != vb:
                diffs.append(f"{p}: {va} != {vb}")
    return diffs


# --------------------------------------------------------------------- builds
def _run(cmd, cwd=None):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"build failed: {' '.join(cmd)}\n{r.stdout[-4000:]}")


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in sources)


def _csrc_files():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [
        os.path.join(ROOT, "include", "mpenv.h"), os.path.join(ROOT, "include", "mpenv_state.h")]


def build_hostsim():
    src = [os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")] + _csrc_files()
    if _stale(HOSTSIM_SO, src):
        os.makedirs(os.path.dirname(HOSTSIM_SO), exist_ok=True)
        _run(["g++", "-std=c++17", "-O2", "-march=x86-64-v3", "-ffp-contract=off", "-fno-fast-math",
              "-fPIC", "-shared", "-o", HOSTSIM_SO,
              os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp"),
              os.path.join(CSRC, "assets.cpp"), os.path.join(CSRC, "map_capi.cpp"),
              os.path.join(CSRC, "navmesh.cpp")])
    return HOSTSIM_SO


def build_oracle():
    """Builds oracle/_ref/*.so when /root/reference is present (this container);
    on the GPU box the prebuilt files travel with the snapshot."""
    have = all(os.path.exists(p) for p in (REF_SO, REF_DET_SO, REF_FAST_SO, DET_MATH_SO))
    if os.path.isdir("/root/reference/src"):
        deps = [os.path.join(ORACLE_DIR, f) for f in ("ref_driver.cpp", "det_math_override.h",
                "det_math_probe.cpp", "shim/shim_core.cpp", "shim/madrona/shim_core.hpp",
                "shim/madrona/geo.hpp", "Makefile")] + [os.path.join(CSRC, "mp_math.h")]
        if not have or any(_stale(p, deps) for p in (REF_SO, REF_DET_SO, REF_FAST_SO, DET_MATH_SO)):
            _run(["make", "-j4"], cwd=ORACLE_DIR)
    elif not have:
        raise RuntimeError("oracle/_ref is missing and /root/reference is not available to build it")
    return REF_SO, REF_DET_SO


# ---------------------------------------------------------------------- map
def make_scene(dst_dir, zones=None, extra_boxes=None):
    """A copy of data/simple_map with its own zones.bin (map_importer.cpp:545-561: u32 n, AABB[n],
    f32 rotation[n]) and / or extra axis-aligned boxes appended to collisions.bin
    (map_importer.cpp:35-124, 233) as one more mesh each -- an affine copy of one of the map's own
    wall boxes, so vertex order and winding are the reference's.
    zones: [(xmin, ymin, zmin, xmax, ymax, zmax, rotation), ...];  extra_boxes: [(xmin, ymin, zmin, xmax, ymax, zmax), ...]."""
    import shutil
    import struct
    os.makedirs(dst_dir, exist_ok=True)
    for f in os.listdir(SCENE):
        shutil.copy(os.path.join(SCENE, f), os.path.join(dst_dir, f))
    if zones is not None:
        z = np.asarray(zones, np.float32).reshape(-1, 7)
        with open(os.path.join(dst_dir, "zones.bin"), "wb") as fh:
            fh.write(np.uint32(len(z)).tobytes())
            fh.write(np.ascontiguousarray(z[:, :6]).tobytes())
            fh.write(np.ascontiguousarray(z[:, 6]).tobytes())
    if extra_boxes:
        b = open(os.path.join(SCENE, "collisions.bin"), "rb").read()
        off = 24
        nmat, nname = struct.unpack("<QQ", b[off:off + 16]); off += 16
        head_end = off + nname + 4 * nmat
        off = head_end
        nmesh, nvert, ntri = struct.unpack("<QQQ", b[off:off + 24]); off += 24
        verts = np.frombuffer(b[off:off + 12 * nvert], np.float32).reshape(-1, 3); off += 12 * nvert
        idx = np.frombuffer(b[off:off + 12 * ntri], np.uint32).reshape(-1, 3); off += 12 * ntri
        mats = np.frombuffer(b[off:off + 4 * ntri], np.uint32); off += 4 * ntri
        meshes = np.frombuffer(b[off:off + 16 * nmesh], np.uint32).reshape(-1, 4)
        tv0, tnv, tt0, tnt = [int(x) for x in meshes[5]]              # template: an interior wall box
        tv = verts[tv0:tv0 + tnv]
        lo, hi = tv.min(axis=0), tv.max(axis=0)
        new_v, new_i, new_m, new_meshes = [verts], [idx], [mats], [meshes]
        nv, nt = int(nvert), int(ntri)
        for bx in extra_boxes:
            blo, bhi = np.asarray(bx[:3], np.float32), np.asarray(bx[3:], np.float32)
            new_v.append((blo + (tv - lo) / (hi - lo) * (bhi - blo)).astype(np.float32))
            new_i.append(idx[tt0:tt0 + tnt])                             # mesh-local indices
            new_m.append(mats[tt0:tt0 + tnt])
            new_meshes.append(np.array([[nv, tnv, nt, tnt]], np.uint32))
            nv += tnv; nt += tnt
        with open(os.path.join(dst_dir, "collisions.bin"), "wb") as fh:
            fh.write(b[:head_end])
            fh.write(struct.pack("<QQQ", nmesh + len(extra_boxes), nv, nt))
            for arr in (np.concatenate(new_v), np.concatenate(new_i), np.concatenate(new_m), np.concatenate(new_meshes)):
                fh.write(np.ascontiguousarray(arr).tobytes())
    return dst_dir


def pillar_boxes(n, seed=5):
    """n pillars (30 x 30 x 260, standing on the floor) scattered over the open parts of simple_map,
    away from the spawn rows at |y| ~ 1750."""
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < n:
        x, y = rng.uniform(-1800, 1800), rng.uniform(-1500, 1500)
        out.append((x, y, 0.4, x + 30.0, y + 30.0, 260.0))
    return out


# simple_map's three zones plus a fourth around the eastern platform, rotated 0.3 rad: the map
# Task::ZoneCaptureDefend needs (its objective is zone 3, sim.cpp:822-825)
FOUR_ZONES = [(-759.59906, -488.75443, -10.08213, 759.59906, 488.75443, 383.38632, 0.0),
              (-1953.6677, -1051.0848, -10.08213, -1114.1248, 1051.0848, 383.38632, 0.0),
              (1107.1403, -1051.0848, -10.08213, 1946.6832, 1051.0848, 383.38632, 0.0),
              (-600.0, 500.0, -10.0, 700.0, 1500.0, 380.0, 0.3)]


class MapInfo(C.Structure):
    _fields_ = [("num_nodes", C.c_uint32), ("num_tris", C.c_uint32), ("num_leaves", C.c_uint32),
                ("root_aabb", C.c_float * 6), ("world_bounds", C.c_float * 6),
                ("num_a_spawns", C.c_uint32), ("num_b_spawns", C.c_uint32),
                ("num_respawns", C.c_uint32), ("num_zones", C.c_uint32)]


def load_map(lib, scene=SCENE):
    """Returns (nodes u8[num_nodes,64], verts f32[num_tris*3,3], info) using the
    mpenv_map_* C ABI of `lib` (hostsim or the product library)."""
    lib.mpenv_map_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.mpenv_map_info.argtypes = [C.c_void_p, C.POINTER(MapInfo)]
    lib.mpenv_map_nodes.restype = C.c_void_p
    lib.mpenv_map_nodes.argtypes = [C.c_void_p]
    lib.mpenv_map_verts.restype = C.c_void_p
    lib.mpenv_map_verts.argtypes = [C.c_void_p]
    lib.mpenv_map_free.argtypes = [C.c_void_p]
    lib.mpenv_map_last_error.restype = C.c_char_p
    m = C.c_void_p()
    rc = lib.mpenv_map_load(scene.encode(), C.byref(m))
    if rc != 0:
        raise RuntimeError(f"mpenv_map_load failed ({rc}): {lib.mpenv_map_last_error().decode()}")
    info = MapInfo()
    lib.mpenv_map_info(m, C.byref(info))
    nodes = np.ctypeslib.as_array(C.cast(lib.mpenv_map_nodes(m), C.POINTER(C.c_uint8)),
                                  shape=(info.num_nodes, 64)).copy()
    verts = np.ctypeslib.as_array(C.cast(lib.mpenv_map_verts(m), C.POINTER(C.c_float)),
                                  shape=(info.num_tris * 3, 3)).copy()
    lib.mpenv_map_free(m)
    return nodes, verts, info


class NavmeshDesc(C.Structure):
    _fields_ = [("num_verts", C.c_uint32), ("num_tris", C.c_uint32), ("verts", C.c_void_p),
                ("tri_indices", C.c_void_p), ("tri_adjacency", C.c_void_p), ("astar_lookup", C.c_void_p)]


def load_navmesh(lib, scene=SCENE):
    """(verts f32[nv,3], tris u32[nt,3], adjacency i32[nt,3], lookup u32[nt,nt]) from the
    mpenv_map_navmesh C ABI of `lib`."""
    lib.mpenv_map_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.mpenv_map_navmesh.argtypes = [C.c_void_p, C.POINTER(NavmeshDesc)]
    lib.mpenv_map_free.argtypes = [C.c_void_p]
    lib.mpenv_map_last_error.restype = C.c_char_p
    m = C.c_void_p()
    rc = lib.mpenv_map_load(scene.encode(), C.byref(m))
    if rc != 0:
        raise RuntimeError(f"mpenv_map_load failed ({rc}): {lib.mpenv_map_last_error().decode()}")
    d = NavmeshDesc()
    rc = lib.mpenv_map_navmesh(m, C.byref(d))
    if rc != 0:
        raise RuntimeError(f"mpenv_map_navmesh failed ({rc}): {lib.mpenv_map_last_error().decode()}")
    arr = lambda ptr, ct, shape: np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=shape).copy()
    out = (arr(d.verts, C.c_float, (d.num_verts, 3)), arr(d.tri_indices, C.c_uint32, (d.num_tris, 3)),
           arr(d.tri_adjacency, C.c_int32, (d.num_tris, 3)), arr(d.astar_lookup, C.c_uint32, (d.num_tris, d.num_tris)))
    lib.mpenv_map_free(m)
    return out


# ------------------------------------------------------------------- oracle
class RefConfig(C.Structure):
    _fields_ = [("numWorlds", C.c_uint32), ("randSeed", C.c_uint32), ("autoReset", C.c_int32),
                ("simFlags", C.c_uint32), ("taskType", C.c_uint32), ("teamSize", C.c_uint32),
                ("numPBTPolicies", C.c_uint32), ("policyHistorySize", C.c_uint32),
                ("worldIDOffset", C.c_uint32), ("numThreads", C.c_uint32),
                ("enableEventLog", C.c_uint32), ("enableRecordLog", C.c_uint32),
                ("enableReplayLog", C.c_uint32), ("numCurriculumSnapshots", C.c_uint32),
                ("curriculumSnapshots", C.c_void_p), ("numNavVerts", C.c_uint32), ("numNavTris", C.c_uint32),
                ("navVerts", C.c_void_p), ("navTriIndices", C.c_void_p), ("navTriAdjacency", C.c_void_p),
                ("navLookup", C.c_void_p), ("trainFlank", C.c_uint32)]


# Log record layouts (reference types.hpp:574-634, 729-760; csrc/mp_defs.h).
EVENT_DT = np.dtype([("type", "<u4"), ("pad0", "<u4"), ("matchID", "<u8"), ("step", "<u4"),
                     ("b0", "u1"), ("b1", "u1"), ("h1", "<u2")])
PLAYER_SNAP_DT = np.dtype([("pos", "<i2", (3,)), ("yaw", "<i2"), ("pitch", "<i2"), ("magNumBullets", "u1"),
                           ("isReloading", "u1"), ("hp", "u1"), ("flags", "u1")])
SNAP_DT = np.dtype([("numEvents", "<u4"), ("eventMask", "<u4"), ("matchID", "<u8"), ("step", "<u2"),
                    ("curZone", "u1"), ("curZoneController", "i1"), ("zoneStepsRemaining", "<u2"),
                    ("stepsUntilPoint", "<u2"), ("players", PLAYER_SNAP_DT, (12,))])
AGENT_LOG_DT = np.dtype([("position", "<f4", (3,)), ("aim", "<f4", (6,)), ("hp", "<f4"), ("mag", "<i4", (2,)),
                         ("standState", "<i4", (3,)), ("shotAgentIdx", "<i4"), ("firedShotT", "<f4"),
                         ("wasKilled", "u1"), ("successfullKill", "u1"), ("pad", "u1", (2,))])
STEP_LOG_DT = np.dtype([("agentData", AGENT_LOG_DT, (12,)), ("curStep", "<i4")])
CURRICULUM_DT = np.dtype([("step", "<u2"), ("curZone", "u1"), ("curZoneController", "i1"),
                          ("zoneStepsRemaining", "<u2"), ("stepsUntilPoint", "<u2"),
                          ("players", PLAYER_SNAP_DT, (12,))])
assert EVENT_DT.itemsize == 24 and SNAP_DT.itemsize == 192 and STEP_LOG_DT.itemsize == 868
assert CURRICULUM_DT.itemsize == 176


def curriculum_from_step_snapshots(snaps):
    """CurriculumSnapshot records (types.hpp:816-819) from PackedStepSnapshot records: the
    body after the 16-byte {numEvents, eventMask, matchID} header (what
    build_trajectory_curriculum.cpp:9-60 reassembles from the analytics DB)."""
    raw = np.ascontiguousarray(snaps).view(np.uint8).reshape(len(snaps), SNAP_DT.itemsize)
    return np.ascontiguousarray(raw[:, 16:]).view(CURRICULUM_DT).reshape(len(snaps))

EV_CAPTURE, EV_RELOAD, EV_KILL, EV_PLAYER_SHOT = 1, 2, 4, 8


def canonical_events(ev):
    """Events with the bytes the reference leaves unspecified masked out: struct
    padding, and the upper half of the 4-byte payload union for the 2-byte
    payloads (Reload / Kill / PlayerShot)."""
    ev = ev.copy()
    ev["pad0"] = 0
    ev["h1"][ev["type"] != EV_CAPTURE] = 0
    return ev


def canonical_step_logs(sl):
    sl = sl.copy()
    sl["agentData"]["pad"] = 0
    return sl


class _SimBase:
    num_worlds = 0

    def event_log(self):
        """(events, step_snapshots) of the last step as structured arrays."""
        raise NotImplementedError

    def record_log(self):
        raise NotImplementedError

    def set_replay_log(self, step_logs):
        raise NotImplementedError

    def export(self, name):
        kind, cols, dt = EXPORT_SHAPES[name]
        arr = np.empty((rows_of(kind, self.num_worlds), cols), dtype=dt)
        self._read(EXPORT_IDS[name], arr)
        return arr

    def set_export(self, name, arr):
        kind, cols, dt = EXPORT_SHAPES[name]
        arr = np.ascontiguousarray(arr, dtype=dt).reshape(rows_of(kind, self.num_worlds), cols)
        self._write(EXPORT_IDS[name], arr)

    def set_actions(self, discrete, discrete_aim, aim=None):
        self.set_export("pvp_discrete_action", discrete)
        self.set_export("pvp_discrete_aim_action", discrete_aim)
        if aim is not None:
            self.set_export("pvp_aim_action", aim)

    def world_state(self, w):
        ws = WorldState()
        self._state(w, ws)
        return struct_to_dict(ws)


class Oracle(_SimBase):
    """The reference's own sim.cpp / utils.cpp / level_gen.cpp / mesh_bvh.inl under
    oracle/shim.  det=True routes sinf/cosf/logf/atan2f/asinf to mp_math.h."""

    def __init__(self, num_worlds, rand_seed=10, auto_reset=False, sim_flags=0, det=True,
                 world_id_offset=0, num_threads=1, scene=SCENE, bvh=None, event_log=False,
                 record_log=False, replay_log=False, curriculum=None, navmesh=True, train_flank=False,
                 fast=False, task=2):
        """fast=True (with det=False): the -O3 build of the libm oracle, for timing only."""
        build_oracle()
        assert not (fast and det), "the timed -O3 build exists for the libm arithmetic only"
        self.lib = C.CDLL(REF_DET_SO if det else (REF_FAST_SO if fast else REF_SO))
        L = self.lib
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.POINTER(RefConfig), C.c_char_p, C.c_void_p, C.c_uint32,
                                 C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        for f in ("ref_destroy", "ref_init", "ref_step"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.ref_read_export.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.ref_write_export.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.ref_get_world_state.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(WorldState)]
        L.ref_trace_ray.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        L.ref_trace_ray_brute.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_sphere_cast.restype = C.c_float
        L.ref_sphere_cast.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        L.ref_sphere_cast_brute.restype = C.c_float
        L.ref_sphere_cast_brute.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float]
        L.ref_set_sim_control.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_num_threads.restype = C.c_uint32
        L.ref_num_threads.argtypes = [C.c_void_p]
        if bvh is None:
            bvh = load_map(C.CDLL(build_hostsim()), scene)
        nodes, verts, info = bvh
        v20 = np.zeros((verts.shape[0], 5), dtype=np.float32)     # BVHVertex {pos, uv}
        v20[:, :3] = verts
        cfg = RefConfig(num_worlds, rand_seed, int(auto_reset), sim_flags, int(task), 6, 0, 1,
                        world_id_offset, num_threads, int(event_log), int(record_log), int(replay_log),
                        0 if curriculum is None else len(curriculum),
                        None if curriculum is None else np.ascontiguousarray(curriculum).ctypes.data)
        cfg.trainFlank = int(train_flank)
        if navmesh:
            nv, nt, na, nl = self._nav = load_navmesh(C.CDLL(build_hostsim()), scene)
            cfg.numNavVerts, cfg.numNavTris = len(nv), len(nt)
            cfg.navVerts, cfg.navTriIndices = nv.ctypes.data, nt.ctypes.data
            cfg.navTriAdjacency, cfg.navLookup = na.ctypes.data, nl.ctypes.data
        L.ref_read_event_log.restype = C.c_int64
        L.ref_read_event_log.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.ref_event_log_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_record_log.restype = C.c_void_p
        L.ref_record_log.argtypes = [C.c_void_p]
        L.ref_replay_log.restype = C.c_void_p
        L.ref_replay_log.argtypes = [C.c_void_p]
        assert L.ref_event_row_bytes(0) == EVENT_DT.itemsize and L.ref_event_row_bytes(1) == SNAP_DT.itemsize
        assert L.ref_step_log_bytes() == STEP_LOG_DT.itemsize
        root = np.array(list(info.root_aabb), dtype=np.float32)
        self.h = L.ref_create(C.byref(cfg), scene.encode(), nodes.ctypes.data, info.num_nodes,
                              v20.ctypes.data, verts.shape[0], info.num_leaves, root.ctypes.data)
        if not self.h:
            raise RuntimeError("ref_create failed")
        self.num_worlds = num_worlds

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    def init(self):
        self.lib.ref_init(self.h)

    def step(self):
        self.lib.ref_step(self.h)

    def num_threads(self):
        return self.lib.ref_num_threads(self.h)

    def event_log(self):
        ev = np.zeros(self.num_worlds * 32, EVENT_DT)
        sn = np.zeros(self.num_worlds, SNAP_DT)
        ne = self.lib.ref_read_event_log(self.h, 0, ev.ctypes.data, len(ev))
        ns = self.lib.ref_read_event_log(self.h, 1, sn.ctypes.data, len(sn))
        assert ne >= 0 and ns >= 0
        cnt = np.zeros(2, np.uint32)
        self.lib.ref_event_log_counts(self.h, cnt[0:].ctypes.data, cnt[1:].ctypes.data)
        assert (int(cnt[0]), int(cnt[1])) == (ne, ns), "EventLogGlobalState counters disagree with the tables"
        return ev[:ne], sn[:ns]

    def record_log(self):
        ptr = self.lib.ref_record_log(self.h)
        assert ptr
        buf = (C.c_char * (self.num_worlds * STEP_LOG_DT.itemsize)).from_address(ptr)
        return np.frombuffer(buf, STEP_LOG_DT).copy()

    def set_replay_log(self, step_logs):
        ptr = self.lib.ref_replay_log(self.h)
        assert ptr and step_logs.nbytes == self.num_worlds * STEP_LOG_DT.itemsize
        C.memmove(ptr, np.ascontiguousarray(step_logs).ctypes.data, step_logs.nbytes)

    def _read(self, eid, arr):
        rc = self.lib.ref_read_export(self.h, eid, arr.ctypes.data, arr.nbytes)
        assert rc == 0, f"ref_read_export({eid}) -> {rc}"

    def _write(self, eid, arr):
        rc = self.lib.ref_write_export(self.h, eid, arr.ctypes.data, arr.nbytes)
        assert rc == 0, f"ref_write_export({eid}) -> {rc}"

    def _state(self, w, ws):
        assert self.lib.ref_get_world_state(self.h, w, C.byref(ws)) == 0

    def set_sim_control(self, v):
        a = np.asarray(v, dtype=np.int32)
        self.lib.ref_set_sim_control(self.h, a.ctypes.data)

    def trace_ray(self, o, d):
        o = np.asarray(o, np.float32); d = np.asarray(d, np.float32)
        t = C.c_float()
        hit = self.lib.ref_trace_ray(self.h, o.ctypes.data, d.ctypes.data, np.float32(3.4028234663852886e38), C.byref(t))
        return hit, t.value

    def trace_ray_brute(self, o, d):
        o = np.asarray(o, np.float32); d = np.asarray(d, np.float32)
        t = C.c_float()
        hit = self.lib.ref_trace_ray_brute(self.h, o.ctypes.data, d.ctypes.data, C.byref(t))
        return hit, t.value

    def sphere_cast(self, o, d, r=15.0):
        o = np.asarray(o, np.float32); d = np.asarray(d, np.float32)
        n = np.zeros(3, np.float32)
        t = self.lib.ref_sphere_cast(self.h, o.ctypes.data, d.ctypes.data, r, n.ctypes.data)
        return t, n

    def sphere_cast_brute(self, o, d, r=15.0):
        o = np.asarray(o, np.float32); d = np.asarray(d, np.float32)
        return self.lib.ref_sphere_cast_brute(self.h, o.ctypes.data, d.ctypes.data, r)

    def trace_rays(self, o, d, brute=False):
        """Batched traceRay (mesh_bvh.inl:110-208) / brute force over all triangles: (hit i32[n], t f32[n])."""
        o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
        t = np.empty(len(o), np.float32); hit = np.empty(len(o), np.int32)
        self.lib.ref_trace_rays.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
        self.lib.ref_trace_rays(self.h, len(o), o.ctypes.data, d.ctypes.data, int(brute), t.ctypes.data, hit.ctypes.data)
        return hit, t

    def sphere_casts(self, o, d, r=15.0, brute=False):
        o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
        t = np.empty(len(o), np.float32); n = np.zeros((len(o), 3), np.float32)
        self.lib.ref_sphere_casts.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_float, C.c_int32,
                                              C.c_void_p, C.c_void_p]
        self.lib.ref_sphere_casts(self.h, len(o), o.ctypes.data, d.ctypes.data, r, int(brute), t.ctypes.data, n.ctypes.data)
        return t, n


# ------------------------------------------------------------------ hostsim
class HostSimConfig(C.Structure):
    _fields_ = [("numWorlds", C.c_uint32), ("randSeed", C.c_uint32), ("autoReset", C.c_int32),
                ("simFlags", C.c_uint32), ("worldIDOffset", C.c_uint32), ("enableFullTeam", C.c_uint32),
                ("enableEventLog", C.c_uint32), ("enableRecordLog", C.c_uint32),
                ("enableReplayLog", C.c_uint32), ("numCurriculumSnapshots", C.c_uint32),
                ("curriculumSnapshots", C.c_void_p), ("trainFlank", C.c_uint32), ("taskType", C.c_uint32)]


class HostSim(_SimBase):
    """Device stage functions compiled for the host (tests only)."""

    def __init__(self, num_worlds, rand_seed=10, auto_reset=False, sim_flags=0, world_id_offset=0,
                 full_team=True, scene=SCENE, event_log=False, record_log=False, replay_log=False,
                 curriculum=None, train_flank=False, task=2):
        self.lib = C.CDLL(build_hostsim())
        L = self.lib
        L.hostsim_create.restype = C.c_void_p
        L.hostsim_create.argtypes = [C.POINTER(HostSimConfig), C.c_char_p]
        for f in ("hostsim_destroy", "hostsim_init", "hostsim_step"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.hostsim_read_export.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.hostsim_write_export.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.hostsim_get_world_state.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(WorldState)]
        L.hostsim_trace_ray.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.hostsim_sphere_cast.restype = C.c_float
        L.hostsim_sphere_cast.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        cfg = HostSimConfig(num_worlds, rand_seed, int(auto_reset), sim_flags, world_id_offset,
                            int(full_team), int(event_log), int(record_log), int(replay_log),
                            0 if curriculum is None else len(curriculum),
                            None if curriculum is None else np.ascontiguousarray(curriculum).ctypes.data,
                            int(train_flank), int(task))
        L.hostsim_read_event_log.restype = C.c_int64
        L.hostsim_read_event_log.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.hostsim_record_log.restype = C.c_void_p
        L.hostsim_record_log.argtypes = [C.c_void_p]
        L.hostsim_replay_log.restype = C.c_void_p
        L.hostsim_replay_log.argtypes = [C.c_void_p]
        self.h = L.hostsim_create(C.byref(cfg), scene.encode())
        if not self.h:
            raise RuntimeError("hostsim_create failed")
        self.num_worlds = num_worlds

    def close(self):
        if self.h:
            self.lib.hostsim_destroy(self.h)
            self.h = None

    def init(self):
        self.lib.hostsim_init(self.h)

    def step(self):
        self.lib.hostsim_step(self.h)

    def event_log(self):
        ev = np.zeros(self.num_worlds * 32, EVENT_DT)
        sn = np.zeros(self.num_worlds, SNAP_DT)
        ne = self.lib.hostsim_read_event_log(self.h, 0, ev.ctypes.data, len(ev))
        ns = self.lib.hostsim_read_event_log(self.h, 1, sn.ctypes.data, len(sn))
        assert ne >= 0 and ns >= 0
        return ev[:ne], sn[:ns]

    def record_log(self):
        ptr = self.lib.hostsim_record_log(self.h)
        buf = (C.c_char * (self.num_worlds * STEP_LOG_DT.itemsize)).from_address(ptr)
        return np.frombuffer(buf, STEP_LOG_DT).copy()

    def set_replay_log(self, step_logs):
        ptr = self.lib.hostsim_replay_log(self.h)
        assert step_logs.nbytes == self.num_worlds * STEP_LOG_DT.itemsize
        C.memmove(ptr, np.ascontiguousarray(step_logs).ctypes.data, step_logs.nbytes)

    def _read(self, eid, arr):
        rc = self.lib.hostsim_read_export(self.h, eid, arr.ctypes.data, arr.nbytes)
        assert rc == 0, f"hostsim_read_export({eid}) -> {rc}"

    def _write(self, eid, arr):
        rc = self.lib.hostsim_write_export(self.h, eid, arr.ctypes.data, arr.nbytes)
        assert rc == 0, f"hostsim_write_export({eid}) -> {rc}"

    def _state(self, w, ws):
        assert self.lib.hostsim_get_world_state(self.h, w, C.byref(ws)) == 0

    def trace_ray(self, o, d):
        o = np.asarray(o, np.float32); d = np.asarray(d, np.float32)
        t = C.c_float()
        hit = self.lib.hostsim_trace_ray(self.h, o.ctypes.data, d.ctypes.data, C.byref(t))
        return hit, t.value

    def sphere_cast(self, o, d, r=15.0):
        o = np.asarray(o, np.float32); d = np.asarray(d, np.float32)
        n = np.zeros(3, np.float32)
        t = self.lib.hostsim_sphere_cast(self.h, o.ctypes.data, d.ctypes.data, r, n.ctypes.data)
        return t, n

    def trace_rays(self, o, d):
        o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
        t = np.empty(len(o), np.float32); hit = np.empty(len(o), np.int32)
        self.lib.hostsim_trace_rays.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.hostsim_trace_rays(self.h, len(o), o.ctypes.data, d.ctypes.data, t.ctypes.data, hit.ctypes.data)
        return hit, t

    def sphere_casts(self, o, d, r=15.0):
        o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
        t = np.empty(len(o), np.float32); n = np.zeros((len(o), 3), np.float32)
        self.lib.hostsim_sphere_casts.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_float,
                                                  C.c_void_p, C.c_void_p]
        self.lib.hostsim_sphere_casts(self.h, len(o), o.ctypes.data, d.ctypes.data, r, t.ctypes.data, n.ctypes.data)
        return t, n


# ------------------------------------------------------------------ actions
_ROT = (13, 15, 26, 6, 17, 29, 16, 24)


def threefry_split(a, b, idx, idx_upper=0):
    """Vectorised Threefry2x32-20 key split (same scheme as mp_math.h randSplit)."""
    a = np.asarray(a, np.uint32); b = np.asarray(b, np.uint32)
    ks = [a, b, np.uint32(0x1BD11BDA) ^ a ^ b]
    with np.errstate(over="ignore"):
        x0 = np.asarray(idx, np.uint32) + ks[0]
        x1 = np.asarray(idx_upper, np.uint32) + ks[1]

        def rounds(x0, x1, rs):
            for r in rs:
                x0 = x0 + x1
                x1 = (x1 << np.uint32(r)) | (x1 >> np.uint32(32 - r))
                x1 = x1 ^ x0
            return x0, x1
        x0, x1 = rounds(x0, x1, _ROT[:4]); x0 = x0 + ks[1]; x1 = x1 + ks[2] + np.uint32(1)
        x0, x1 = rounds(x0, x1, _ROT[4:]); x0 = x0 + ks[2]; x1 = x1 + ks[0] + np.uint32(2)
        x0, x1 = rounds(x0, x1, _ROT[:4]); x0 = x0 + ks[0]; x1 = x1 + ks[1] + np.uint32(3)
        x0, x1 = rounds(x0, x1, _ROT[4:]); x0 = x0 + ks[1]; x1 = x1 + ks[2] + np.uint32(4)
        x0, x1 = rounds(x0, x1, _ROT[:4])
        return x0 + ks[2], x1 + ks[0] + np.uint32(5)


def random_actions(step, num_worlds, world_id_offset=0, seed=1234, stress=False, magazine=None):
    """The synthetic action stream of SURVEY 8(d): a counter RNG keyed on (seed,
    step, global_world*12+agent), so every shard / implementation consumes the
    same stream.  Returns (discrete i32[N,4], discrete_aim i32[N,2])."""
    n = num_worlds * AGENTS
    gid = (np.arange(n, dtype=np.uint64) + np.uint64(world_id_offset) * np.uint64(AGENTS)).astype(np.uint32)
    k0, k1 = threefry_split(np.uint32(seed), np.uint32(0x9E3779B9), gid, np.uint32(step))
    r0, r1 = threefry_split(k0, k1, np.uint32(0))
    r2, r3 = threefry_split(k0, k1, np.uint32(1))
    u = lambda bits: (bits >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -24)
    move_amount = (r0 % np.uint32(3)).astype(np.int32)
    move_angle = ((r0 >> np.uint32(8)) % np.uint32(8)).astype(np.int32)
    uf = u(r1)
    fire = np.where(uf < 0.6, 0, np.where(uf < 0.95, 1, 2)).astype(np.int32)
    us = u(r2)
    stand = np.where(us < 0.8, 0, np.where(us < 0.9, 1, 2)).astype(np.int32)
    yaw = (r3 % np.uint32(13)).astype(np.int32)
    pitch = ((r3 >> np.uint32(8)) % np.uint32(7)).astype(np.int32)
    if stress:
        fire = np.ones(n, np.int32)
        if magazine is not None:
            fire = np.where(np.asarray(magazine).reshape(n, 2)[:, 0] == 0, 2, 1).astype(np.int32)
    discrete = np.stack([move_amount, move_angle, fire, stand], axis=1)
    discrete_aim = np.stack([yaw, pitch], axis=1)
    return discrete, discrete_aim


def seek_policy(sim, step, num_worlds, world_id_offset=0, seed=77):
    """A scripted policy that produces combat (the uniform random stream almost
    never brings the teams into contact): turn towards a visible opponent and
    fire, otherwise run towards the zone centre; a third of the decisions stay
    random.  Computed from `sim`'s exported observations, so feed the SAME
    arrays to every implementation under comparison."""
    n = num_worlds * AGENTS
    self_obs = sim.export("self_obs")
    opp = sim.export("opponent_obs").reshape(n, 6, 32)
    mag = sim.export("magazine")
    rnd_d, rnd_a = random_actions(step, num_worlds, world_id_offset, seed)
    yaw_amt = np.pi * np.array([0, 1 / 256, 1 / 128, 1 / 64, 1 / 32, 1 / 16, 1 / 8])
    pitch_amt = np.pi * np.array([0, 1 / 128, 1 / 64, 1 / 32])

    def bucket(delta, amounts, centre):
        k = np.searchsorted(amounts, np.abs(delta), side="right") - 1
        return (centre + np.sign(delta).astype(np.int64) * k).astype(np.int32)

    los = opp[:, :, 30] > 0
    dist = np.where(los, opp[:, :, 23], np.inf)
    tgt = np.argmin(dist, axis=1)
    has_tgt = los.any(axis=1)
    rows = np.arange(n)
    yaw_d = np.where(has_tgt, opp[rows, tgt, 24], self_obs[:, 31])
    pitch_d = np.where(has_tgt, opp[rows, tgt, 25], 0.0)
    discrete = np.zeros((n, 4), np.int32)
    discrete[:, 0] = np.where(has_tgt, 0, 2)                 # stop to shoot, else run
    discrete[:, 1] = 0                                       # forward
    discrete[:, 2] = np.where(mag[:, 0] == 0, 2, np.where(has_tgt, 1, 0))
    discrete[:, 3] = 0
    aim = np.stack([bucket(yaw_d, yaw_amt, 6), bucket(pitch_d, pitch_amt, 3)], axis=1)
    use_rnd = (rnd_d[:, 1] % 3 == 0) & ~has_tgt
    discrete = np.where(use_rnd[:, None], rnd_d, discrete)
    aim = np.where(use_rnd[:, None], rnd_a, aim)
    return discrete.astype(np.int32), aim.astype(np.int32)


# ------------------------------------------------------------------ product
class MpenvConfig(C.Structure):
    _fields_ = [("gpu_id", C.c_int32), ("num_worlds", C.c_uint32), ("rand_seed", C.c_uint32),
                ("auto_reset", C.c_int32), ("sim_flags", C.c_uint32), ("task_type", C.c_uint32),
                ("team_size", C.c_uint32), ("num_pbt_policies", C.c_uint32),
                ("policy_history_size", C.c_uint32), ("scene_dir", C.c_char_p),
                ("world_id_offset", C.c_uint32), ("enable_full_team", C.c_uint32),
                ("event_log_dir", C.c_char_p), ("record_log_path", C.c_char_p),
                ("replay_log_path", C.c_char_p), ("curriculum_data_path", C.c_char_p),
                ("train_flank", C.c_uint32)]


class TensorDesc(C.Structure):
    _fields_ = [("dev_ptr", C.c_void_p), ("dtype", C.c_int32), ("ndim", C.c_int32),
                ("dims", C.c_int64 * 4), ("gpu_id", C.c_int32)]


COUNT_SO = os.path.join(os.path.dirname(PRODUCT_SO), "libmpenv_b200_count.so")


def load_product(so=None):
    """Loads libmpenv_b200.so (built by __graft_entry__.build()); no fallback.
    so=COUNT_SO: the instrumented twin (ray census, never timed)."""
    so = so or PRODUCT_SO
    if not os.path.exists(so):
        raise RuntimeError(f"{so} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    L = C.CDLL(so)
    L.mpenv_ray_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
    L.mpenv_create.argtypes = [C.POINTER(MpenvConfig), C.POINTER(C.c_void_p)]
    for f in ("mpenv_destroy", "mpenv_init", "mpenv_step", "mpenv_sync"):
        getattr(L, f).argtypes = [C.c_void_p]
    L.mpenv_step_async.argtypes = [C.c_void_p, C.c_void_p]
    L.mpenv_init_async.argtypes = [C.c_void_p, C.c_void_p]
    L.mpenv_tensor.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(TensorDesc)]
    L.mpenv_write_tensor_async.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
    L.mpenv_read_tensor_async.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
    L.mpenv_get_world_state.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(WorldState)]
    L.mpenv_last_error.restype = C.c_char_p
    L.mpenv_last_error.argtypes = [C.c_void_p]
    L.mpenv_launches_per_step.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.mpenv_episode_stats_async.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
    L.mpenv_host_alloc.argtypes = [C.c_uint64, C.POINTER(C.c_void_p)]
    L.mpenv_host_free.argtypes = [C.c_void_p]
    L.mpenv_trigger_reset.argtypes = [C.c_void_p, C.c_int32]
    L.mpenv_event_log.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32),
                                  C.POINTER(C.c_void_p), C.POINTER(C.c_uint32)]
    L.mpenv_record_log.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32)]
    L.mpenv_set_replay_log.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    return L


class GpuSim(_SimBase):
    """The product: libmpenv_b200.so through its C ABI, host buffers in and out."""

    def __init__(self, num_worlds, rand_seed=10, auto_reset=False, sim_flags=0, world_id_offset=0,
                 full_team=True, scene=SCENE, gpu_id=0, event_log=False, record_log=False, replay_log=False,
                 curriculum_path=None, train_flank=False, instrumented=False, task=2):
        """instrumented: load libmpenv_b200_count.so (mpenv_ray_counts works there).
        event_log / record_log / replay_log: False = off, True = in memory only,
        a path = the reference's files (event_log: directory)."""
        self.lib = load_product(COUNT_SO if instrumented else None)
        self._scene = scene.encode()
        path = lambda v: None if v is False or v is None else (b"" if v is True else str(v).encode())
        self._log_paths = [path(event_log), path(record_log), path(replay_log),
                           None if curriculum_path is None else str(curriculum_path).encode()]
        cfg = MpenvConfig(gpu_id, num_worlds, rand_seed, int(auto_reset), sim_flags, int(task), 6, 0, 1,
                          self._scene, world_id_offset, int(full_team), *self._log_paths, int(train_flank))
        self.h = C.c_void_p()
        rc = self.lib.mpenv_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            raise RuntimeError(f"mpenv_create failed ({rc}): {self.lib.mpenv_last_error(None).decode()}")
        self.num_worlds = num_worlds

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed ({rc}): {self.lib.mpenv_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            self.lib.mpenv_destroy(self.h)
            self.h = None

    def init(self):
        self._check(self.lib.mpenv_init(self.h), "mpenv_init")

    def step(self):
        self._check(self.lib.mpenv_step(self.h), "mpenv_step")

    def _read(self, eid, arr):
        self._check(self.lib.mpenv_read_tensor_async(self.h, eid, arr.ctypes.data, arr.nbytes), "read")
        self._check(self.lib.mpenv_sync(self.h), "sync")

    def _write(self, eid, arr):
        self._check(self.lib.mpenv_write_tensor_async(self.h, eid, arr.ctypes.data, arr.nbytes), "write")
        self._check(self.lib.mpenv_sync(self.h), "sync")

    def _state(self, w, ws):
        self._check(self.lib.mpenv_get_world_state(self.h, w, C.byref(ws)), "get_world_state")

    def event_log(self):
        ev, sn = C.c_void_p(), C.c_void_p()
        ne, ns = C.c_uint32(), C.c_uint32()
        self._check(self.lib.mpenv_event_log(self.h, C.byref(ev), C.byref(ne), C.byref(sn), C.byref(ns)), "event_log")
        def view(ptr, n, dt):
            if n == 0:
                return np.zeros(0, dt)
            return np.frombuffer((C.c_char * (n * dt.itemsize)).from_address(ptr.value), dt).copy()
        return view(ev, ne.value, EVENT_DT), view(sn, ns.value, SNAP_DT)

    def record_log(self):
        ptr, n = C.c_void_p(), C.c_uint32()
        self._check(self.lib.mpenv_record_log(self.h, C.byref(ptr), C.byref(n)), "record_log")
        buf = (C.c_char * (n.value * STEP_LOG_DT.itemsize)).from_address(ptr.value)
        return np.frombuffer(buf, STEP_LOG_DT).copy()

    def set_replay_log(self, step_logs):
        a = np.ascontiguousarray(step_logs)
        self._check(self.lib.mpenv_set_replay_log(self.h, a.ctypes.data, a.nbytes), "set_replay_log")

    def ray_counts(self, reset=False):
        """uint64[4]: lidar, line-of-sight, hitscan rays and swept-sphere casts (instrumented build only)."""
        out = np.zeros(4, np.uint64)
        self._check(self.lib.mpenv_ray_counts(self.h, out.ctypes.data, int(reset)), "ray_counts")
        return out

    def launches(self):
        a, b = C.c_uint32(), C.c_uint32()
        self.lib.mpenv_launches_per_step(self.h, C.byref(a), C.byref(b))
        return a.value, b.value

    def kernel_variants(self):
        """(warps per CTA of the pooled visibility kernel, CTAs per SM of the packet lidar kernel);
        0 = the per-thread kernel (mpenv_kernel_variants)."""
        a, b = C.c_int32(), C.c_int32()
        self.lib.mpenv_kernel_variants.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        self._check(self.lib.mpenv_kernel_variants(self.h, C.byref(a), C.byref(b)), "kernel_variants")
        return a.value, b.value

    def damaged_guards(self):
        """Number of device buffers whose guard bands were written to (mpenv_check_guards)."""
        n = C.c_uint32()
        self.lib.mpenv_check_guards.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
        self._check(self.lib.mpenv_check_guards(self.h, C.byref(n)), "check_guards")
        return n.value

    def episode_stats(self):
        """int64[16] episode statistics of this shard (k_stats), synchronous."""
        host = np.empty(16, np.int64)
        self.lib.mpenv_episode_stats.argtypes = [C.c_void_p, C.c_void_p]
        self._check(self.lib.mpenv_episode_stats(self.h, host.ctypes.data), "episode_stats")
        return host


def episode_stats_host(sim):
    """numpy mirror of k_stats (step_kernels.cuh): the per-shard input of the
    multi-GPU allreduce, computed from any _SimBase's exports + world state."""
    W = sim.num_worlds
    done = sim.export("done").reshape(W, AGENTS)
    mr = sim.export("match_result")
    alive = sim.export("alive").reshape(W, AGENTS)
    reward = sim.export("reward").reshape(W, AGENTS)
    out = np.zeros(16, np.int64)
    out[0] = W
    fin = done[:, 0] != 0
    out[1] = fin.sum()
    for k in range(3):
        out[2 + k] = (fin & (mr[:, 0] == k)).sum()
    out[5:7] = mr[fin, 1:3].sum(axis=0)
    out[7:9] = mr[fin, 3:5].sum(axis=0)
    out[9] = (alive != 0).sum()
    out[10] = done.sum()
    out[11] = np.rint(reward.astype(np.float32) * np.float32(1.0e6)).astype(np.int64).sum()
    out[12] = sum(sim.world_state(w)["cur_step"] for w in range(W))
    return out
