/*
 * mpenv.h -- C ABI of libmpenv_b200.so: the B200-native replacement for the
 * per-step simulation path of MadronaMPEnv's Manager (reference src/mgr.hpp).
 *
 * Each entry point names the reference interface it replaces.  Plain pointers
 * and sizes only; no C++ or torch types cross this boundary.  One handle = one
 * GPU = one CUDA stream; calls on a handle are not re-entrant (the reference
 * Manager is single-threaded per instance, mgr.hpp:31-161).  Multi-GPU = one
 * handle per device, each created with its own world_id_offset.
 *
 * All functions return MPENV_OK (0) or a negative error code; the message is
 * available from mpenv_last_error().  (The reference FATALs / aborts on these
 * conditions, mgr.cpp:204-206, 1855.)
 */
#ifndef MPENV_H
#define MPENV_H

#include <stddef.h>
#include <stdint.h>
#include "mpenv_state.h"

#ifdef __cplusplus
extern "C" {
#endif

#define MPENV_OK 0
#define MPENV_ERR_INVALID (-1)   /* bad argument / unsupported configuration */
#define MPENV_ERR_IO (-2)        /* asset file missing or corrupt */
#define MPENV_ERR_CUDA (-3)      /* CUDA runtime error (no device, OOM, launch failure) */

/* Task (reference src/types.hpp:45-51); MPENV_TASK_ZONE and MPENV_TASK_ZONE_CAPTURE_DEFEND (needs a map with
 * >= 4 zones: the objective is pinned to zone 3, sim.cpp:822-825) are implemented. */
#define MPENV_TASK_EXPLORE 0
#define MPENV_TASK_TDM 1
#define MPENV_TASK_ZONE 2
#define MPENV_TASK_TURRET 3
#define MPENV_TASK_ZONE_CAPTURE_DEFEND 4

/* SimFlags (reference src/sim_flags.hpp:7-21). */
#define MPENV_FLAG_DEFAULT 0u
#define MPENV_FLAG_SPAWN_IN_MIDDLE (1u << 0)
#define MPENV_FLAG_RANDOMIZE_HP_MAGAZINE (1u << 1)
#define MPENV_FLAG_NAVMESH_SPAWN (1u << 2)
#define MPENV_FLAG_NO_RESPAWN (1u << 3)
#define MPENV_FLAG_STAGGER_STARTS (1u << 4)
#define MPENV_FLAG_ENABLE_CURRICULUM (1u << 5)
#define MPENV_FLAG_HARDCODED_SPAWNS (1u << 6)
#define MPENV_FLAG_RANDOM_FLIP_TEAMS (1u << 7)
#define MPENV_FLAG_STATIC_FLIP_TEAMS (1u << 8)
#define MPENV_FLAG_FULL_TEAM_POLICY (1u << 9)
#define MPENV_FLAG_SIM_EVAL_MODE (1u << 10)
#define MPENV_FLAG_SUBZONES (1u << 11)

/* Replaces Manager::Config (mgr.hpp:33-52) + MapConfig (mgr.hpp:15-23). */
typedef struct mpenv_config {
    int32_t gpu_id;              /* Config::gpuID */
    uint32_t num_worlds;         /* Config::numWorlds (worlds on THIS device) */
    uint32_t rand_seed;          /* Config::randSeed */
    int32_t auto_reset;          /* Config::autoReset */
    uint32_t sim_flags;          /* Config::simFlags */
    uint32_t task_type;          /* Config::taskType; MPENV_TASK_ZONE or MPENV_TASK_ZONE_CAPTURE_DEFEND */
    uint32_t team_size;          /* Config::teamSize; must be 6 */
    uint32_t num_pbt_policies;   /* Config::numPBTPolicies (carried, unused by the step) */
    uint32_t policy_history_size;/* Config::policyHistorySize (carried, unused) */
    const char *scene_dir;       /* directory holding collisions/spawns/zones.bin
                                    (bindings.cpp:58-76 derives the four file names from it) */
    uint32_t world_id_offset;    /* global id of local world 0: the episode RNG key is
                                    split_i(initKey, episode, GLOBAL world id) (sim.cpp:743-746),
                                    so shards of one big batch reproduce it world for world */
    uint32_t enable_full_team;   /* also fill the FullTeam* tensors every step (sim.cpp:3054-3301) */
    /* Logs (Config::eventLogPath / recordLogPath / replayLogPath, mgr.hpp:47-49).  NULL = off.
     * An EMPTY string enables the log in memory only (mpenv_event_log / mpenv_record_log /
     * mpenv_set_replay_log) without touching the file system.  With any of them on, the step
     * entry points synchronise with the host every step, as the reference's do
     * (CUDAImpl::saveEvents, mgr.cpp:417-470). */
    const char *event_log_dir;   /* directory receiving events.bin + steps.bin (mgr.cpp:218-231) */
    const char *record_log_path; /* StepLog[num_worlds] appended per step (mgr.cpp:253-257) */
    const char *replay_log_path; /* StepLog[num_worlds] consumed per step; replaces the gameplay
                                    systems (pvpReplayLogic, sim.cpp:5587-5604) */
    /* Config::curriculumDataPath (mgr.hpp:50, mgr.cpp:1424-1441): a file of CurriculumSnapshot
     * records (176 B = PackedMatchState + 12 PackedPlayerSnapshot, types.hpp:816-819; the body
     * of a steps.bin record after its 16-byte header).  Half of the non-eval episodes then
     * start from a random snapshot (level_gen.cpp:498-581).  NULL = off. */
    const char *curriculum_data_path;
    uint32_t train_flank;        /* Config::trainFlank (mgr.hpp:46): RewardMode::Flank, flankRewardSystem
                                    (sim.cpp:4211-4290) instead of zoneRewardSystem */
} mpenv_config;

typedef struct mpenv_handle mpenv_handle;

/* Tensor export ids = reference ExportID (src/sim.hpp:15-66). */
enum mpenv_export_id {
    MPENV_EXPORT_RESET = 0, MPENV_EXPORT_WORLD_CURRICULUM, MPENV_EXPORT_EXPLORE_ACTION,
    MPENV_EXPORT_PVP_DISCRETE_ACTION, MPENV_EXPORT_PVP_AIM_ACTION,
    MPENV_EXPORT_PVP_DISCRETE_AIM_ACTION, MPENV_EXPORT_REWARD, MPENV_EXPORT_DONE,
    MPENV_EXPORT_MATCH_RESULT, MPENV_EXPORT_AGENT_POLICY, MPENV_EXPORT_SELF_OBSERVATION,
    MPENV_EXPORT_TEAMMATE_OBSERVATIONS, MPENV_EXPORT_OPPONENT_OBSERVATIONS,
    MPENV_EXPORT_OPPONENT_LAST_KNOWN_OBSERVATIONS, MPENV_EXPORT_SELF_POSITION_OBSERVATION,
    MPENV_EXPORT_TEAMMATE_POSITION_OBSERVATIONS, MPENV_EXPORT_OPPONENT_POSITION_OBSERVATIONS,
    MPENV_EXPORT_OPPONENT_LAST_KNOWN_POSITION_OBSERVATIONS, MPENV_EXPORT_OPPONENT_MASKS,
    MPENV_EXPORT_FWD_LIDAR, MPENV_EXPORT_REAR_LIDAR, MPENV_EXPORT_AGENT_MAP,
    MPENV_EXPORT_UNMASKED_AGENT_MAP, MPENV_EXPORT_HP, MPENV_EXPORT_ALIVE, MPENV_EXPORT_MAGAZINE,
    MPENV_EXPORT_FULL_TEAM_ACTIONS, MPENV_EXPORT_FULL_TEAM_GLOBAL, MPENV_EXPORT_FULL_TEAM_PLAYERS,
    MPENV_EXPORT_FULL_TEAM_ENEMIES, MPENV_EXPORT_FULL_TEAM_LAST_KNOWN_ENEMIES,
    MPENV_EXPORT_FULL_TEAM_FWD_LIDAR, MPENV_EXPORT_FULL_TEAM_REAR_LIDAR,
    MPENV_EXPORT_FULL_TEAM_REWARD, MPENV_EXPORT_FULL_TEAM_DONE,
    MPENV_EXPORT_FULL_TEAM_POLICY_ASSIGNMENTS, MPENV_EXPORT_EVENT_LOG,
    MPENV_EXPORT_PACKED_STEP_SNAPSHOT, MPENV_EXPORT_FILTERS_STATE_OBSERVATION,
    MPENV_EXPORT_REWARD_HYPER_PARAMS, MPENV_NUM_EXPORTS,
    MPENV_EXPORT_SIM_CONTROL = 1000  /* Manager::simControlTensor (mgr.cpp:287-293) */
};

#define MPENV_DTYPE_I32 0
#define MPENV_DTYPE_F32 1

/* Replaces madrona::py::Tensor(ptr, dtype, dims, gpuID) (mgr.cpp:295-301):
 * a raw DEVICE pointer owned by the handle, stable for its lifetime, row-major,
 * leading dimension world-major then agent. */
typedef struct mpenv_tensor_desc {
    void *dev_ptr;
    int32_t dtype;
    int32_t ndim;
    int64_t dims[4];
    int32_t gpu_id;
} mpenv_tensor_desc;

/* Manager::Manager(cfg) (mgr.cpp:1914-1919 -> Impl::init mgr.cpp:1213): loads the
 * map, builds the BVH, allocates device state and constructs every world
 * (Sim::Sim, sim.cpp:5850).  Fails with MPENV_ERR_CUDA when no CUDA device is
 * usable -- there is no CPU execution path. */
int mpenv_create(const mpenv_config *cfg, mpenv_handle **out);
/* Manager::~Manager */
void mpenv_destroy(mpenv_handle *h);
/* Manager::init (mgr.cpp:1923-1946): force a reset of every world, run the Init
 * graph (reset + observations), return when results are visible. */
int mpenv_init(mpenv_handle *h);
/* Manager::step (mgr.cpp:1948-1951): one simulation step, synchronous. */
int mpenv_step(mpenv_handle *h);
/* Stream-ordered variants, no host synchronisation: the kernel sequence of
 * gpuStreamInit / gpuStreamStep (mgr.cpp:507-645) without the buffer copies.
 * `cuda_stream` is a cudaStream_t. */
int mpenv_init_async(mpenv_handle *h, void *cuda_stream);
int mpenv_step_async(mpenv_handle *h, void *cuda_stream);
/* Manager::gpuStreamInit / gpuStreamStep (mgr.cpp:507-645): copy the
 * train-interface inputs from `buffers` into the sim tensors, run, copy the
 * outputs back, all on `cuda_stream`.  Buffer order = mpenv_train_interface(). */
int mpenv_stream_init(mpenv_handle *h, void *cuda_stream, void **buffers);
int mpenv_stream_step(mpenv_handle *h, void *cuda_stream, void **buffers);
/* Debug aid (no reference counterpart): every device buffer of the handle is allocated between two
 * 256-byte guard bands; this counts the buffers whose bands were written to (0 = no kernel wrote
 * out of bounds since mpenv_create).  Synchronises the handle's stream. */
int mpenv_check_guards(mpenv_handle *h, uint32_t *num_damaged);
/* XLA GPU custom-call entry points for sim.jax() (bindings.cpp:149-158: madrona::py::JAXInterface
 * registers gpuStreamInit / gpuStreamStep as XLA custom calls).  Signature = XLA's original GPU
 * custom-call ABI, void(cudaStream_t, void **buffers, const char *opaque, size_t opaque_len);
 * `opaque` carries the 8 bytes of the mpenv_handle pointer (as the reference packs its Manager *),
 * `buffers` = the train-interface inputs then outputs, mpenv_train_interface() order.  They forward
 * to mpenv_stream_init / mpenv_stream_step; an error is printed and left in mpenv_last_error(). */
void mpenv_xla_init(void *cuda_stream, void **buffers, const char *opaque, size_t opaque_len);
void mpenv_xla_step(void *cuda_stream, void **buffers, const char *opaque, size_t opaque_len);
/* The handle's own cudaStream_t (the stream of mpenv_step / the *_tensor_async copies). */
void *mpenv_stream(mpenv_handle *h);
/* Waits for the handle's own stream. */
int mpenv_sync(mpenv_handle *h);
/* Manager::*Tensor() exporters (mgr.cpp:1965-2381). */
int mpenv_tensor(mpenv_handle *h, uint32_t export_id, mpenv_tensor_desc *out);

/* Host-buffer access to a tensor: the copies cudaCopyStepInputs /
 * cudaCopyStepOutputs and the viewer setters do (mgr.cpp:625-629, 2492-2565),
 * for callers that hold HOST memory.  Stream-ordered on the handle's stream
 * (follow with mpenv_sync); `bytes` must equal the tensor's size.  Pinned host
 * memory from mpenv_host_alloc makes the copies truly asynchronous. */
int mpenv_write_tensor_async(mpenv_handle *h, uint32_t export_id, const void *host_src, uint64_t bytes);
int mpenv_read_tensor_async(mpenv_handle *h, uint32_t export_id, void *host_dst, uint64_t bytes);
/* Pipelined read-back for host consumers: snapshots the `n` listed tensors
 * device->device on the handle's stream (ordered after the step that produced
 * them, so the NEXT step may overwrite the live tensors), then copies the
 * snapshot device->host on a second stream, overlapping the next step's
 * kernels with the PCIe transfer.  host_dst[i] must stay valid until
 * mpenv_read_wait() returns; a new call first waits (stream-ordered) for the
 * previous transfer.  Pinned host memory required for real overlap. */
int mpenv_read_tensors_pipelined(mpenv_handle *h, const uint32_t *export_ids, void *const *host_dst, uint32_t n);
/* The two halves of mpenv_read_tensors_pipelined, for a host loop that reuses ONE set of host
 * buffers: the snapshots are double-buffered on the device, so the snapshot of step t+1 can be
 * queued (right after mpenv_step_async) while step t's transfer is still running; the transfer
 * of step t+1 is then started, with mpenv_read_snapshot_async, once the consumer is done with the
 * host buffers (after mpenv_read_wait).  Keeps the PCIe link busy back to back.
 * mpenv_read_snapshot_async reads the tensors of the latest snapshot, in its order (n must match). */
int mpenv_snapshot_tensors_async(mpenv_handle *h, const uint32_t *export_ids, uint32_t n);
int mpenv_read_snapshot_async(mpenv_handle *h, void *const *host_dst, uint32_t n);
int mpenv_read_wait(mpenv_handle *h);
int mpenv_host_alloc(uint64_t bytes, void **out);
void mpenv_host_free(void *p);

/* Manager::trainInterface (mgr.cpp:2383-2431): the named input / output tensors
 * in the order the stream entry points expect their buffers. */
typedef struct mpenv_named_tensor {
    const char *name;
    uint32_t export_id;
    int32_t is_input;
} mpenv_named_tensor;
int mpenv_train_interface(mpenv_handle *h, const mpenv_named_tensor **out, uint32_t *count);

/* Manager::triggerReset / setPvPAction / setHP / setUniformAgentPolicy
 * (mgr.cpp:2484-2658): small host->device writes used by drivers and tests. */
int mpenv_trigger_reset(mpenv_handle *h, int32_t world_idx);
int mpenv_set_pvp_action(mpenv_handle *h, int32_t world_idx, int32_t agent_idx,
                         const int32_t discrete[4], const float aim[2], const int32_t discrete_aim[2]);
int mpenv_set_hp(mpenv_handle *h, int32_t world_idx, int32_t agent_idx, int32_t hp);
int mpenv_set_uniform_agent_policy(mpenv_handle *h, int32_t policy_idx);

/* Per-device reduction of episode statistics for the multi-GPU allreduce
 * (new).  Writes MPENV_NUM_STATS int64 values to a device buffer
 * owned by the handle and returns its pointer; stream-ordered. */
#define MPENV_NUM_STATS 16
int mpenv_episode_stats_async(mpenv_handle *h, void *cuda_stream, int64_t **dev_out);
/* Synchronous host variant: reduces on the handle's stream and copies the
 * MPENV_NUM_STATS values to `host_out`. */
int mpenv_episode_stats(mpenv_handle *h, int64_t *host_out);
/* The same reduction for callers that must not allocate (code under CUDA graph capture, serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Writes MPENV_NUM_STATS int64 values to `dev_out`, a DEVICE buffer the
 * caller owns, stream-ordered on `cuda_stream`; the call allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine).  Its scratch -- per-CTA partial sums and a ticket, 33 KB -- is a static
 * device array that every GPU loading the library holds until exit, so:
 *   - a handle with more than MPENV_STATIC_STATS_MAX_WORLDS worlds is rejected (MPENV_ERR_INVALID,
 *     mpenv_last_error() names the limit);
 *   - all calls on one GPU share the scratch: calls that overlap on the GPU (two streams, two handles,
 *     two processes do NOT share it) race; order them on one stream or with events.
 * The values are identical to mpenv_episode_stats_async's (integer sums). */
#define MPENV_STATIC_STATS_MAX_WORLDS 65536u
int mpenv_episode_stats_static(mpenv_handle *h, void *cuda_stream, int64_t *dev_out);

/* ---- logs ---------------------------------------------------
 * Event log: GameEvent (24 B: u32 type, u32 pad, u64 matchID, u32 step, 4 payload
 * bytes; types.hpp:729-760) and PackedStepSnapshot (192 B, types.hpp:629-634) records of
 * the LAST step, in host memory owned by the handle (valid until the next step).
 * Order: world-major; within a world the fire events in agent order (Reload, or
 * PlayerShot then Kill), then Capture -- the order of the reference's CPU executor
 * for one world; the reference's order ACROSS worlds is unspecified (atomic append). */
int mpenv_event_log(mpenv_handle *h, const void **events, uint32_t *num_events,
                    const void **step_states, uint32_t *num_step_states);
/* Record log: StepLog[num_worlds] (868 B each, types.hpp:574-589) of the last step. */
int mpenv_record_log(mpenv_handle *h, const void **step_logs, uint32_t *num_worlds);
/* Replay from memory (replay_log_path == ""): StepLog[num_worlds] for the NEXT step. */
int mpenv_set_replay_log(mpenv_handle *h, const void *step_logs, uint64_t bytes);

/* Test / debug surface: full per-world state (see mpenv_state.h). */
int mpenv_get_world_state(mpenv_handle *h, uint32_t world_idx, mpenv_world_state *out);
/* Number of kernel launches issued by one step / one init (for bench accounting). */
int mpenv_launches_per_step(mpenv_handle *h, uint32_t *step_launches, uint32_t *init_launches);
/* Which ray kernels the map's BVH size selected at create():
 *   los_pool_warps     warps per CTA of the warp-pooled visibility kernel, 0 = the per-thread kernel;
 *   lidar_packet_ctas  CTAs per SM of the ray-packet lidar kernel (2 or 1), 0 = the per-thread kernel. */
int mpenv_kernel_variants(mpenv_handle *h, int32_t *los_pool_warps, int32_t *lidar_packet_ctas);

/* Ray census of the INSTRUMENTED build (libmpenv_b200_count.so = same sources with
 * -DMPENV_COUNT_RAYS): cumulative counts since the last reset of
 * out[0] lidar rays, out[1] line-of-sight rays, out[2] hitscan rays, out[3] swept-sphere
 * casts, counted where the reference calls traceRayAgainstWorld / sphereCast.  The product
 * library returns MPENV_ERR_INVALID (its kernels carry no counters). */
int mpenv_ray_counts(mpenv_handle *h, uint64_t *out, int32_t reset);

/* Measurement aid: runs ONE step as individual launches on the handle's stream
 * with a CUDA event pair around each kernel; writes up to `cap` durations (ms)
 * and the matching kernel names.  Synchronous.  Same kernels and results as
 * mpenv_step. */
int mpenv_step_timed(mpenv_handle *h, float *kernel_ms, const char **kernel_names, uint32_t cap,
                     uint32_t *count);

const char *mpenv_last_error(mpenv_handle *h);   /* h may be NULL: last create() error */

/* ---- map section: host-side asset loading (Impl::init's loadMapData /
 * loadMapSpawnData / loadMapZones, mgr.cpp:1228-1236, 1360) ---------------- */
typedef struct mpenv_map mpenv_map;
typedef struct mpenv_map_info_t {
    uint32_t num_nodes, num_tris, num_leaves;
    float root_aabb[6];
    float world_bounds[6];
    uint32_t num_a_spawns, num_b_spawns, num_respawns, num_zones;
} mpenv_map_info_t;
int mpenv_map_load(const char *scene_dir, mpenv_map **out);
void mpenv_map_free(mpenv_map *m);
int mpenv_map_info(const mpenv_map *m, mpenv_map_info_t *info);
/* num_nodes x 64-byte nodes in the reference MeshBVH::Node layout (mesh_bvh.hpp:61-86) */
const void *mpenv_map_nodes(const mpenv_map *m);
/* (num_tris + 1) x 3 x 3 floats: de-indexed triangle positions in leaf order */
const float *mpenv_map_verts(const mpenv_map *m);
/* Navmesh + A* next-triangle table of the scripted bot (policy id -1): the data the
 * reference builds in Impl::init (mgr.cpp:1300-1330: importNavmesh, vertex de-duplication,
 * Navmesh::initFromPolygons, buildAStarLookup mgr.cpp:946-1211).  Pointers are owned by `m`. */
typedef struct mpenv_navmesh_t {
    uint32_t num_verts, num_tris;
    const float *verts;            /* num_verts x 3 */
    const uint32_t *tri_indices;   /* num_tris x 3 */
    const int32_t *tri_adjacency;  /* num_tris x 3, -1 = border */
    const uint32_t *astar_lookup;  /* num_tris x num_tris: next triangle [start][goal], 0xFFFFFFFF = none */
} mpenv_navmesh_t;
int mpenv_map_navmesh(const mpenv_map *m, mpenv_navmesh_t *out);
const char *mpenv_map_last_error(void);

#ifdef __cplusplus
}
#endif

#endif
