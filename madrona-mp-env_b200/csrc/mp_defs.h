// mp_defs.h -- shared plain-data definitions for the B200 MadronaMPEnv step.
//
// Everything here compiles both as CUDA device code (nvcc, sm_100a) and as
// plain host C++ (tests/hostsim builds the same stage functions for the CPU
// test-suite; the product library never runs them on the host).
//
// Row layouts of the exported tensors follow the reference component structs
// (reference src/types.hpp:173-193, 211-237, 246-252, 254-264, 275-423,
// 477-492) because those layouts ARE the tensor API (mgr.cpp:1965-2381).
#pragma once

#include <stdint.h>
#include <float.h>
#include <math.h>

#if defined(__CUDACC__)
#define MP_HD __host__ __device__ __forceinline__
#define MP_HD_NOINLINE __host__ __device__ __noinline__
#else
#define MP_HD inline
#define MP_HD_NOINLINE inline
#endif

namespace mpenv {

// ---- gameplay constants (reference src/consts.hpp:7-74) -------------------
constexpr int kTeamSize = 6;            // consts::maxTeamSize
constexpr int kNumTeams = 2;
constexpr int kAgents = 12;             // agents per world (Task::Zone, 6v6)
constexpr int kMaxZones = 5;
constexpr int kNumStepsPerZone = 600;
constexpr int kZonePointInterval = 20;
constexpr int kZoneWinPoints = 125;
constexpr int kPoseTransitionSpeed = 10;
constexpr float kAgentRadius = 15.f;
constexpr float kStandHeight = 65.f;
constexpr float kCrouchHeight = 47.f;
constexpr float kProneHeight = 30.f;
constexpr float kMaxRunVelocity = 400.f;
constexpr float kMaxWalkVelocity = 200.f;
constexpr float kMaxCrouchVelocity = 50.f;
constexpr float kMaxProneVelocity = 20.f;
constexpr float kDeaccelerateRate = 1000.f;
constexpr int kRespawnInvincibleSteps = 5;
constexpr int kAutohealDelaySteps = 150;
constexpr float kAutohealPerStep = 5.f;
constexpr int kEpisodeLen = 3000;
constexpr int kNumMoveAmountBuckets = 3;
constexpr int kNumMoveAngleBuckets = 8;
constexpr int kFwdLidarW = 32, kFwdLidarH = 2;
constexpr int kRearLidarW = 8, kRearLidarH = 2;
constexpr int kMaxWeaponTypes = 3;
constexpr float kDeltaT = 0.05f;
constexpr int kAimYawBuckets = 13;
constexpr int kAimPitchBuckets = 7;
constexpr float kPi = 3.14159265358979323846264338327950288f;

constexpr int kMaxSpawns = 128;         // SpawnUsageCounter::maxNumSpawns
constexpr int kExploreGridW = 81;       // ExploreTracker::gridWidth/Height
constexpr int kExploreGridMax = 40;
constexpr int kMaxGoalRegions = 10;
constexpr int kMaxCrumbs = 192;         // per-world breadcrumb pool

// SimFlags (reference src/sim_flags.hpp:7-21)
enum SimFlagBits : uint32_t {
    kFlagSpawnInMiddle = 1u << 0,
    kFlagRandomizeHPMagazine = 1u << 1,
    kFlagNavmeshSpawn = 1u << 2,
    kFlagNoRespawn = 1u << 3,
    kFlagStaggerStarts = 1u << 4,
    kFlagEnableCurriculum = 1u << 5,
    kFlagHardcodedSpawns = 1u << 6,
    kFlagRandomFlipTeams = 1u << 7,
    kFlagStaticFlipTeams = 1u << 8,
    kFlagFullTeamPolicy = 1u << 9,
    kFlagSimEvalMode = 1u << 10,
    kFlagSubZones = 1u << 11,
};

struct Vec2 { float x, y; };
struct Vec3 { float x, y, z; };
struct Quat { float w, x, y, z; };
struct AABB { Vec3 pMin, pMax; };

struct RandKey { uint32_t a, b; };
struct RNG { RandKey k; uint32_t count; };

// ---- per-agent state components (device arrays of numWorlds*12 rows) ------
struct Aim { float yaw, pitch; Quat rot; };                     // types.hpp:455
struct MoveState { Vec3 newPos; Vec3 newVel; float maxVel; };   // types.hpp:471
struct Magazine { int32_t numBullets; int32_t isReloading; };   // types.hpp:481
struct StandState { int32_t cur, tgt, transition; };            // types.hpp:221
struct CombatState {                                            // types.hpp:510
    RNG rng;
    int32_t landedShotOn;        // agent index in world, -1 = none
    int32_t remainingRespawnSteps;
    int32_t remainingStepsBeforeAutoheal;
    int32_t successfulKill;
    int32_t wasShotCount;
    int32_t wasKilled;
    float firedShotT;
    int32_t inZone;
    float minDistToZone;
    int32_t hasDiedDuringEpisode;
    int32_t reloadedFullMag;
    int32_t weaponType;
};
// SimFlags::SubZones state (types.hpp:791-796; CombatState::inSubZone / minDistToSubZone, types.hpp:510),
// kept apart from CombatState / WorldState so that the default step does not carry it.
constexpr int kNumSubZones = 8;                                 // sim.hpp:190
struct SubZoneAgentState { int32_t inSubZone; float minDistToSubZone; };
struct SubZoneState { int32_t curControllingTeam, isContested, isCaptured; };
struct CrumbAgentState {                                        // types.hpp:772
    float totalPenalty;
    uint32_t lastCrumbUID;       // 0 = Entity::none()
    int32_t stepsSinceNew;
};
struct Crumb {                                                  // types.hpp:766
    Vec3 pos;
    float penalty;
    int32_t agent;               // owner: team*6 + offset
    uint32_t uid;
};
struct RewardHyperParams {                                      // types.hpp:254
    float teamSpirit, shotScale, exploreScale, inZoneScale, zoneTeamContestScale,
          zoneTeamCtrlScale, zoneDistScale, zoneEarnedPointScale, breadcrumbScale;
};

// ---- action rows (exported, written by the caller) ------------------------
struct DiscreteAction { int32_t moveAmount, moveAngle, fire, stand; }; // :173
struct AimAction { float yaw, pitch; };                                 // :190
struct DiscreteAimAction { int32_t yaw, pitch; };                       // :180

// ---- observation rows (exported) ------------------------------------------
struct StandObs { float curS, curC, curP, tgtS, tgtC, tgtP, transition; };
struct ZoneObs {
    float centerX, centerY, centerZ, toCenterDist, toCenterYaw, toCenterPitch,
          myTeamControlling, enemyTeamControlling, isContested, isCaptured,
          stepsUntilPoint, stepsRemaining, id[4];
};
struct PlayerCommonObs {                                        // types.hpp:301
    float isValid, isAlive, globalX, globalY, globalZ, facingYaw, facingPitch,
          velocityX, velocityY, velocityZ, yawVelocity, pitchVelocity;
    StandObs stand;
    float inZone;
    float weaponTypeObs[kMaxWeaponTypes];
};
struct OtherPlayerCommonObs : PlayerCommonObs {                 // types.hpp:319
    float toPlayerDist, toPlayerYaw, toPlayerPitch, relativeFacingYaw,
          relativeFacingPitch;
};
struct CombatObs { float hp, magazine, isReloading, timeBeforeAutoheal; };
struct SelfObs : PlayerCommonObs { CombatObs combat; ZoneObs zone; };
struct TeammateObs : OtherPlayerCommonObs { CombatObs combat; };
struct OpponentObs : OtherPlayerCommonObs {
    float wasHit, firedShot, hasLOS, teamKnowsLocation;
};
struct NormPosObs { float x, y, z; };
struct LidarSample { float depth, isWall, isTeammate, isOpponent; };

static_assert(sizeof(SelfObs) == 43 * 4, "SelfObservation is 43 floats");
static_assert(sizeof(TeammateObs) == 32 * 4, "TeammateObservation is 32 floats");
static_assert(sizeof(OpponentObs) == 32 * 4, "OpponentObservation is 32 floats");

// Full-team rows (types.hpp:1048-1120)
struct FullTeamZoneObs {
    float centerX, centerY, centerZ, myTeamControlling, enemyTeamControlling,
          isContested, isCaptured, stepsUntilPoint, stepsRemaining, id[4];
};
struct FullTeamGlobalObs { float teamID[2]; float fractionMatchRemaining; FullTeamZoneObs zone; };
struct FullTeamCommonObs {
    float isValid; float id[kTeamSize]; float isAlive, globalX, globalY, globalZ,
          facingYaw, facingPitch, velocityX, velocityY, velocityZ;
    StandObs stand; float inZone;
};
struct FullTeamPlayerObs : FullTeamCommonObs { float hp, magazine, isReloading, timeBeforeAutoheal; };
struct FullTeamEnemyObs : FullTeamCommonObs {
    float wasHit, firedShot; float hasLOS[kTeamSize]; float teamKnowsLocation;
};
static_assert(sizeof(FullTeamGlobalObs) == 16 * 4, "global 16 floats");
static_assert(sizeof(FullTeamCommonObs) == 24 * 4, "common 24 floats");
static_assert(sizeof(FullTeamPlayerObs) == 28 * 4, "player 28 floats");
static_assert(sizeof(FullTeamEnemyObs) == 33 * 4, "enemy 33 floats");

// ---- per-world singletons -------------------------------------------------
struct MatchInfo {                                              // types.hpp:127
    int32_t teamA, curStep, isFinished, enableSpawnCurriculum;
    uint32_t curCurriculumTier, curCurriculumSpawnIdx;
};
struct ZoneState {                                              // types.hpp:563
    int32_t curZone, curControllingTeam, isContested, isCaptured, earnedPoint,
            zoneStepsRemaining, stepsUntilPoint;
};
struct ZoneStats { int32_t numSwaps, capturedSteps[2], contestedSteps, activeSteps; };
struct MatchResult {                                            // types.hpp:246
    int32_t winResult, teamTotalKills[2], teamObjectivePoints[2];
    ZoneStats zoneStats[kMaxZones];
};
static_assert(sizeof(MatchResult) == 30 * 4, "MatchResult is 30 x i32");

struct WorldState {
    MatchInfo match;
    ZoneState zone;
    ZoneStats zoneStats[kMaxZones];           // Sim::zoneStats
    RNG baseRNG;
    uint32_t curEpisodeIdx, worldEpisodeCounter;
    int32_t episodeCurriculum;                // Sim::episodeCurriculum
    uint32_t filtersActive[2];                // low 32 bits of the u64 mask
    int32_t filtersLastMatches[2][3];
    int32_t filtersLastMatchedStep[2];
    float goalMinDist[kMaxGoalRegions];
    float goalTeamStepRewards[2];
    float teamRewards[2];
    int32_t numCrumbs;
    uint32_t nextCrumbUID;
    uint32_t matchIDValid;                    // Sim::matchID != ~0 (set by the first initWorld after the ctor)
};

// ---- event log / step snapshots / record + replay log (types.hpp:574-634, 729-760)
enum EventType : uint32_t { kEvNone = 0, kEvCapture = 1, kEvReload = 2, kEvKill = 4, kEvPlayerShot = 8 };
struct GameEvent {                                              // types.hpp:729
    uint32_t type;
    uint32_t pad0;
    uint64_t matchID;
    uint32_t step;
    uint8_t b0, b1;              // capture: zoneIDX, captureTeam; reload: player, numBulletsAtReloadTime;
                                 // kill: killer, killed; playerShot: attacker, target
    uint16_t h1;                 // capture: inZoneMask
};
static_assert(sizeof(GameEvent) == 24, "GameEvent is 24 bytes");
struct PackedPlayerSnapshot {                                   // types.hpp:603
    int16_t pos[3]; int16_t yaw, pitch;
    uint8_t magNumBullets, isReloading, hp, flags;
};
struct PackedMatchState {                                       // types.hpp:621
    uint16_t step; uint8_t curZone; int8_t curZoneController;
    uint16_t zoneStepsRemaining, stepsUntilPoint;
};
struct PackedStepSnapshot {                                     // types.hpp:629
    uint32_t numEvents, eventMask;
    uint64_t matchID;
    PackedMatchState matchState;
    PackedPlayerSnapshot players[kAgents];
};
static_assert(sizeof(PackedPlayerSnapshot) == 14, "PackedPlayerSnapshot is 14 bytes");
static_assert(sizeof(PackedStepSnapshot) == 192, "PackedStepSnapshot is 192 bytes");
struct AgentLogData {                                           // types.hpp:574
    Vec3 position; Aim aim; float hp; Magazine mag; StandState standState;
    int32_t shotAgentIdx; float firedShotT;
    uint8_t wasKilled, successfullKill, pad[2];
};
struct StepLog { AgentLogData agentData[kAgents]; int32_t curStep; };   // types.hpp:586
static_assert(sizeof(AgentLogData) == 72, "AgentLogData is 72 bytes");
static_assert(sizeof(StepLog) == 868, "StepLog is 868 bytes");
struct CurriculumSnapshot {                                     // types.hpp:816
    PackedMatchState matchState;
    PackedPlayerSnapshot players[kAgents];
};
static_assert(sizeof(CurriculumSnapshot) == 176, "CurriculumSnapshot is 176 bytes");
constexpr int kEventSlotsPerAgent = 2;      // fireSystem logs Reload, or PlayerShot (+ Kill)
constexpr int kEventSlotsPerWorld = kAgents * kEventSlotsPerAgent + 1;   // + Capture

struct SpawnUsage {                                             // types.hpp:95
    uint32_t a[kMaxSpawns], b[kMaxSpawns], respawn[kMaxSpawns];
};

// ---- static assets (read-only on device) ----------------------------------
struct Spawn { AABB region; float yawMin, yawMax; };            // types.hpp:60
struct ZOBB { Vec3 pMin, pMax; float rotation; };               // types.hpp:785
struct GoalRegion {                                             // types.hpp:798
    ZOBB sub[3]; int32_t numSub; int32_t attackerTeam; float rewardStrength;
};
struct WeaponStats { int32_t magSize, reloadTime; float dmgPerBullet, accuracyScale; };

// MeshBVH::Node, reference src/mesh_bvh.hpp:61-86 (64 bytes).
struct BVHNode {
    float minX, minY, minZ;
    int8_t expX, expY, expZ;
    uint8_t internalNodes;
    uint8_t triSize[4];
    uint8_t qMinX[4], qMinY[4], qMinZ[4], qMaxX[4], qMaxY[4], qMaxZ[4];
    int32_t children[4];
    int32_t parentID;
};
static_assert(sizeof(BVHNode) == 64, "MeshBVH::Node is 64 bytes");
constexpr int32_t kBVHSentinel = (int32_t)0xFFFFFFFF;

struct BVHView {
    const BVHNode *nodes;
    const Vec3 *verts;      // 3 per triangle, de-indexed, leaf order
    AABB rootAABB;
    uint32_t numNodes, numTris;
};

// Navmesh + A* next-triangle table for the scripted bot (LevelData, types.hpp:136-143).
struct NavView {
    const Vec3 *verts;
    const uint32_t *tris;      // 3 per triangle
    const uint32_t *lookup;    // numTris x numTris
    uint32_t numTris;
};

constexpr int kVertGrid = 64;
struct MapConsts {
    AABB worldBounds;
    float maxDist;                      // Sim::maxDist (sim.cpp:5855)
    float frustum[4];                   // Sim::frustumData (sim.cpp:5869-5882)
    uint32_t numASpawns, numBSpawns, numRespawns, numZones;
    Spawn aSpawns[kMaxSpawns], bSpawns[kMaxSpawns], respawns[kMaxSpawns];
    AABB zoneBoxes[kMaxZones];
    float zoneRots[kMaxZones];
    GoalRegion goalRegions[2];
    int32_t numGoalRegions;
    ZOBB subZones[kNumSubZones];        // level_gen.cpp:279-322 (two from zones 1 / 2, six hard-coded)
    WeaponStats weapons[kMaxWeaponTypes];
    int32_t numWeaponTypes;
    // lidar direction tables: -cos(theta), sin(theta) per column
    // (sim.cpp:3347-3350; theta depends only on the column index)
    float fwdLidarX[kFwdLidarW], fwdLidarY[kFwdLidarW];
    float rearLidarX[kRearLidarW], rearLidarY[kRearLidarW];
    // Occupancy of the mesh vertices on a kVertGrid x kVertGrid grid over x, y (bit = some vertex in
    // the cell): what sphereSweepMayTouchMirroredVertex (mp_geom.h) looks up.
    float vertGridMinX, vertGridMinY, vertGridInvCellX, vertGridInvCellY;
    uint32_t vertGrid[kVertGrid * kVertGrid / 32];
};

struct SimConfig {
    uint32_t numWorlds;
    uint32_t worldIDOffset;             // global id of local world 0 (multi-GPU shards)
    uint32_t simFlags;
    int32_t autoReset;
    RandKey initRandKey;                // mgr.cpp:1736-1738
    int32_t eventLog;                   // TaskConfig::eventGlobalState != nullptr
    int32_t recordLog;                  // TaskConfig::recordLog != nullptr
    int32_t replayLog;                  // TaskConfig::replayLog != nullptr
    int32_t numCurriculumSnapshots;     // TrajectoryCurriculum::numSnapshots (mgr.cpp:1419-1441)
    int32_t rewardFlank;                // TaskConfig::rewardMode == RewardMode::Flank (mgr.cpp:1745-1750)
    int32_t taskCaptureDefend;          // TaskConfig::task == Task::ZoneCaptureDefend (else Task::Zone)
};

// TrainControl (types.hpp:158), live tensor (mgr.cpp:287-293)
struct TrainControl { int32_t evalMode, randomizeEpisodeLengthAfterReset, randomizeTeamSides; };

// All device pointers of one simulator instance.
struct SimPtrs {
    // per agent
    Vec3 *pos; Quat *rot; Vec3 *vel; MoveState *move; Aim *aim;
    float *hp; float *alive; Magazine *mag; float *dmg;   // dmg: [N][6]
    StandState *stand; CombatState *combat; CrumbAgentState *crumbAgent;
    SubZoneAgentState *subZoneAgent;     // SimFlags::SubZones only
    Vec3 *startPos; uint32_t *exploreGrid; uint32_t *exploreNew;
    uint8_t *canSee;                     // [N][8] (6 used)
    int32_t *policy; RewardHyperParams *rewardCoefs;
    DiscreteAction *actDiscrete; AimAction *actAim; DiscreteAimAction *actDiscreteAim;
    float *reward; int32_t *done;
    SelfObs *selfObs; TeammateObs *teammateObs; OpponentObs *opponentObs;
    OpponentObs *lastKnownObs;
    NormPosObs *selfPos; NormPosObs *teammatePos; NormPosObs *opponentPos;
    NormPosObs *lastKnownPos;
    float *masks; float *filtersObs;
    LidarSample *fwdLidar; LidarSample *rearLidar;
    // per world
    WorldState *world; SpawnUsage *spawnUsage; Crumb *crumbs;
    MatchResult *matchResult; int32_t *resetFlag; int32_t *worldCurriculum;
    SubZoneState *subZones;              // [W][8], SimFlags::SubZones only
    // per team (2 per world), full-team interface
    FullTeamGlobalObs *ftGlobal; FullTeamPlayerObs *ftPlayers; FullTeamEnemyObs *ftEnemies;
    FullTeamCommonObs *ftLastKnown; LidarSample *ftFwdLidar; LidarSample *ftRearLidar;
    float *ftReward; int32_t *ftDone;
    // logs (allocated only when enabled)
    GameEvent *evAgent;                  // [N][2]: this step's fireSystem events, type 0 = unused
    GameEvent *evWorld;                  // [W]: this step's Capture event
    PackedStepSnapshot *stepSnap;        // [W]
    StepLog *recordLog;                  // [W]
    StepLog *replayLog;                  // [W]
    // global
    const CurriculumSnapshot *curriculum;   // TrajectoryCurriculum::snapshots (not in the buffer table)
    TrainControl *trainCtrl;
};

}  // namespace mpenv
