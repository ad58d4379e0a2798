// Device-resident arena state ("arena blob") and the constants of the Rocket League simulation.
//
// One ArenaDev per arena, contiguous in HBM; the lane group that owns an arena moves it to/from shared memory with
// 16-byte coalesced accesses.  Units are Bullet units internally (1 bt = 50 uu) exactly like the reference
// (RocketSim/src/BulletLink.h:12-15); conversion to Unreal units happens where the reference converts
// (Car::GetState / Ball::GetState, Sim/Car/Car.cpp:10-17, Sim/Ball/Ball.cpp:27-33).
//
// Constants restate RocketSim/src/RLConst.h:10-437 and Sim/Car/CarConfig/CarConfig.cpp:20-86 (Octane).
#pragma once
#include <stddef.h>
#include "sim_math.cuh"

namespace ggl {

constexpr int kMaxCars = 6;
constexpr int kNumPads = 34;
constexpr int kNumActions = 90;
constexpr int kModeSoccar = 0, kModeHeatseeker = 1;     // ArenaDev::gameMode = ggl_game_mode (include/gigaenv.h)
constexpr int kMaxContacts = 36;        // contact rows per arena per tick: per-body windows + a shared overflow pool (sim_world.cuh);
                                        // a point that finds neither is dropped AND counted in ArenaDev::contactOverflow
constexpr int kMaxPool = 16;            // largest overflow pool of the three layouts (contact_pool_size)
constexpr int kMaxRewards = 16;
constexpr int kMaxTerminals = 4;

constexpr float kUU2BT = 1.f / 50.f;
constexpr float kBT2UU = 50.f;
constexpr float kTick = 1.f / 120.f;

namespace rl {   // RLConst.h
constexpr float GRAVITY_Z = -650.f;
constexpr float ARENA_EXTENT_X = 4096.f, ARENA_EXTENT_Y = 5120.f, ARENA_HEIGHT = 2048.f;
constexpr float CAR_MASS = 180.f, BALL_MASS = CAR_MASS / 6.f;
constexpr float CARBALL_FRICTION = 2.0f, CARBALL_RESTITUTION = 0.0f;
constexpr float CARWORLD_FRICTION = 0.3f, CARWORLD_RESTITUTION = 0.3f;
constexpr float CARCAR_FRICTION = 0.09f, CARCAR_RESTITUTION = 0.1f;
constexpr float BALL_REST_Z = 93.15f, BALL_MAX_ANG_SPEED = 6.f, BALL_DRAG = 0.03f, BALL_FRICTION = 0.35f, BALL_RESTITUTION = 0.6f;
constexpr float WORLD_FRICTION = 0.6f, WORLD_RESTITUTION = 0.3f;   // Arena.cpp:505-506
constexpr float CAR_MAX_SPEED = 2300.f, BALL_MAX_SPEED = 6000.f, BOOST_MAX = 100.f;
constexpr float BOOST_USED_PER_SECOND = BOOST_MAX / 3, BOOST_MIN_TIME = 0.1f;
constexpr float BOOST_ACCEL_GROUND = 2975 / 3.f, BOOST_ACCEL_AIR = 3175 / 3.f, BOOST_SPAWN_AMOUNT = BOOST_MAX / 3;
constexpr float CAR_MAX_ANG_SPEED = 5.5f;
constexpr float SUPERSONIC_START_SPEED = 2200.f, SUPERSONIC_MAINTAIN_MIN_SPEED = SUPERSONIC_START_SPEED - 100.f, SUPERSONIC_MAINTAIN_MAX_TIME = 1.f;
constexpr float POWERSLIDE_RISE_RATE = 5, POWERSLIDE_FALL_RATE = 2;
constexpr float THROTTLE_TORQUE_AMOUNT = CAR_MASS * 400.f, BRAKE_TORQUE_AMOUNT = CAR_MASS * (14.25f + (1.f / 3.f));
constexpr float STOPPING_FORWARD_VEL = 25.f, COASTING_BRAKE_FACTOR = 0.15f, BRAKING_NO_THROTTLE_SPEED_THRESH = 0.01f;
constexpr float THROTTLE_DEADZONE = 0.001f, THROTTLE_AIR_ACCEL = 200 / 3.f;
constexpr float JUMP_ACCEL = 4375.f / 3.f, JUMP_IMMEDIATE_FORCE = 875.f / 3.f, JUMP_MIN_TIME = 0.025f;
constexpr float JUMP_RESET_TIME_PAD = (1 / 40.f), JUMP_MAX_TIME = 0.2f, DOUBLEJUMP_MAX_DELAY = 1.25f;
constexpr float FLIP_Z_DAMP_120 = 0.35f, FLIP_Z_DAMP_START = 0.15f, FLIP_Z_DAMP_END = 0.21f, FLIP_TORQUE_TIME = 0.65f;
constexpr float FLIP_PITCHLOCK_EXTRA_TIME = 0.3f, FLIP_INITIAL_VEL_SCALE = 500.f, FLIP_TORQUE_X = 260.f, FLIP_TORQUE_Y = 224.f;
constexpr float FLIP_FORWARD_IMPULSE_MAX_SPEED_SCALE = 1.f, FLIP_SIDE_IMPULSE_MAX_SPEED_SCALE = 1.9f;
constexpr float FLIP_BACKWARD_IMPULSE_MAX_SPEED_SCALE = 2.5f, FLIP_BACKWARD_IMPULSE_SCALE_X = 16.f / 15.f;
constexpr float BALL_RADIUS = 91.25f, GOAL_THRESHOLD_Y = 5124.25f;
constexpr float CAR_TORQUE_SCALE = (float)(2 * 3.14159265358979323846 / (1 << 16) * 1000);
constexpr float CAR_AUTOFLIP_IMPULSE = 200, CAR_AUTOFLIP_TORQUE = 50, CAR_AUTOFLIP_TIME = 0.4f;
constexpr float CAR_AUTOFLIP_NORMZ_THRESH = 0.70710678118654752440f, CAR_AUTOFLIP_ROLL_THRESH = 2.8f;
constexpr float CAR_AUTOROLL_FORCE = 100, CAR_AUTOROLL_TORQUE = 80;
constexpr float BALL_CAR_EXTRA_IMPULSE_Z_SCALE = 0.35f, BALL_CAR_EXTRA_IMPULSE_FORWARD_SCALE = 0.65f, BALL_CAR_EXTRA_IMPULSE_MAXDELTAVEL = 4600.f;
constexpr float CAR_SPAWN_REST_Z = 17.f, CAR_RESPAWN_Z = 36.f, BUMP_COOLDOWN_TIME = 0.25f, BUMP_MIN_FORWARD_DIST = 64.5f, DEMO_RESPAWN_TIME = 3.f;
constexpr float SUSP_FORCE_SCALE_FRONT = 36.f - (1.f / 4.f), SUSP_FORCE_SCALE_BACK = 54.f + (1.f / 4.f) + (1.5f / 100.f);
constexpr float SUSP_STIFFNESS = 500.f, DAMP_COMPRESSION = 25.f, DAMP_RELAXATION = 40.f, MAX_SUSP_TRAVEL = 12.f, SUSP_SUBTRACTION = 0.05f;
constexpr float AIR_TORQUE_X = 130, AIR_TORQUE_Y = 95, AIR_TORQUE_Z = 400, AIR_DAMP_X = 30, AIR_DAMP_Y = 20, AIR_DAMP_Z = 50;
constexpr float PAD_CYL_HEIGHT = 95, PAD_CYL_RAD_BIG = 208, PAD_CYL_RAD_SMALL = 144, PAD_BOX_HEIGHT = 64, PAD_BOX_RAD_BIG = 160, PAD_BOX_RAD_SMALL = 120;
constexpr float PAD_COOLDOWN_BIG = 10, PAD_COOLDOWN_SMALL = 4, PAD_BOOST_BIG = 100, PAD_BOOST_SMALL = 12;
// Octane (CarConfig.cpp)
constexpr float HITBOX_X = 120.507f, HITBOX_Y = 86.6994f, HITBOX_Z = 38.6591f;
constexpr float HITBOX_OFF_X = 13.87566f, HITBOX_OFF_Y = 0.f, HITBOX_OFF_Z = 20.755f;
constexpr float FRONT_WHEEL_RAD = 12.50f, BACK_WHEEL_RAD = 15.00f, FRONT_SUS_REST = 38.755f, BACK_SUS_REST = 37.055f;
constexpr float FRONT_WHEEL_X = 51.25f, FRONT_WHEEL_Y = 25.90f, FRONT_WHEEL_Z = 20.755f;
constexpr float BACK_WHEEL_X = -33.75f, BACK_WHEEL_Y = 29.50f, BACK_WHEEL_Z = 20.755f;
constexpr float DODGE_DEADZONE = 0.5f;
// collision margins of the reference's Bullet shapes
constexpr float CONVEX_MARGIN = 0.04f;        // btBoxShape default (CONVEX_DISTANCE_MARGIN)
namespace hs {   // RLConst.h:151-174 namespace Heatseeker
constexpr float INITIAL_TARGET_SPEED = 2900, TARGET_SPEED_INCREMENT = 85, MIN_SPEEDUP_INTERVAL = 1, TARGET_Y = 5120, TARGET_Z = 320;
constexpr float HORIZONTAL_BLEND = 1.45f, VERTICAL_BLEND = 0.78f, SPEED_BLEND = 0.3f, MAX_SPEED = 4600;
constexpr float MAX_TURN_PITCH = (float)(7000 * 3.14159265358979323846 / (1 << 15));
constexpr float WALL_BOUNCE_CHANGE_Y_THRESH = 300, WALL_BOUNCE_CHANGE_Y_NORMAL = 0.5f, WALL_BOUNCE_FORCE_SCALE = 1 / 3.f, WALL_BOUNCE_UP_FRAC = 0.3f;
}
}  // namespace rl

struct Controls {
    float throttle, steer, pitch, yaw, roll;
    uint8_t jump, boost, handbrake, _pad;
};

enum GroundKind : int { GROUND_NONE = -1, GROUND_WORLD = 0, GROUND_BALL = 1, GROUND_CAR0 = 2 };

struct Wheel {
    // carried from tick to tick (btWheelInfoRL)
    float steerAngle, engineForce, brake, latFriction, longFriction, extraPushback;
    // per-tick scratch
    V3 hardPoint, contactPoint, contactNormal, latDir, impulse;
    float suspLen, suspRelVel, clippedInv, suspForce;
    int ground;   // GroundKind; GROUND_CAR0 + c for car index c
};

struct alignas(16) CarDev {
    // rigid body (Bullet units)
    V3 pos; M3 basis; V3 vel, angVel, force, torque; M3 invInertiaW; V3 velImpulseCache;
    V3 velSnap, angVelSnap;              // velocities at the start of the pre-tick phase (read by other cars' wheels)
    V3 aabbMin, aabbMax;                 // broadphase AABB of this tick
    // CarState (Sim/Car/Car.h:17-115)
    V3 flipRelTorque, worldContactNormal;
    float jumpTime, flipTime, airTime, airTimeSinceJump, boost, timeSpentBoosting, supersonicTime, handbrakeVal;
    float autoFlipTimer, autoFlipTorqueScale, carContactCooldown, demoRespawnTimer;
    V3 hitRelPos, hitBallPos, hitExtraVel;
    uint64_t hitTick, hitTickExtra, updateCounter;
    uint32_t carContactOtherId, id;
    uint8_t team, isOnGround, hasJumped, hasDoubleJumped, hasFlipped, isFlipping, isJumping, isSupersonic;
    uint8_t isAutoFlipping, worldContactHas, isDemoed, hitValid;
    uint8_t wheelsWithContact[4];
    uint8_t rbDisabled, ballHitNew, _padc[2];        // rigid body has DISABLE_SIMULATION | CF_NO_CONTACT_RESPONSE this tick (Car.cpp:69-83)
    Controls controls, lastControls;
    Wheel w[4];
};

struct BallDev {
    V3 pos, vel, angVel, force, torque, velImpulseCache, aabbMin, aabbMax;
    uint64_t updateCounter;
    uint32_t active;
    // BallState::HeatseekerInfo (Sim/Ball/Ball.h:23-32): which net the ball seeks (0 = none), its target speed, time since the last hit
    float hsTargetDir, hsTargetSpeed, hsTimeSinceHit;
};

struct PadDev {                  // 8 bytes: car ids are 1 .. kMaxCars
    float cooldown;
    uint8_t prevLockedCarId;
    uint8_t curLockedCarId;      // per tick (0 = none)
    uint8_t isActive, _p;
};

// One contact row handed from the narrow phase to the solver.
struct Contact {
    V3 normal;         // world, on B, pointing from B towards A
    V3 relA, relB;     // contact point relative to each body origin
    float dist, friction, restitution;
    int16_t bodyA, bodyB;    // 0 = ball, 1+c = car c, -1 = static world
    int16_t special, staticId;   // staticId: which static body a world row touches (plane_static_code, sim_world.cuh), 0 otherwise
};

// PlayerEventState slots (RLGymCPP Gamestates/Player.h), written by the event tracker and by the bump callback
enum { EV_GOAL = 0, EV_SAVE, EV_ASSIST, EV_SHOT, EV_SHOTPASS, EV_BUMP, EV_BUMPED, EV_DEMO, EV_DEMOED };

struct EnvPlayer {     // per players[] slot bookkeeping of the env layer (GameState / Player, RLGymCPP)
    float prevAction[8];
    float prevBoost;
    uint8_t prevIsFlipping, prevIsOnGround, events[9], _pad;
};

// Layout: [header .. contacts][cars].  cars[] is LAST so that a lane group stages only the first numCars cars of an arena in
// shared memory (smem slot = arena_slot_bytes(numCars)); contacts[] is per-tick scratch that lives in the slot and is only
// written back to HBM by the raw-tick kernel (tests read it through ggl_envset_get_contacts).
// narrow-phase result of one (car, mesh triangle) candidate that produced a contact point (see sim_boxtri.cuh)
constexpr int kMaxMeshCand = 24;        // candidate triangles per car per tick
constexpr int kMaxMeshHit = 6;          // stored narrow-phase hits per car; more fall back to the serial path (6 keeps the 2v2
                                        // slot at 7.7 KB = 7 one-warp blocks per SM, i.e. 4096 arenas in a single wave)
struct CarTriHit { V3 normalOnB, pointOnB; float distance; uint32_t cand; };

struct alignas(16) ArenaDev {
    uint64_t tickCount;
    int32_t numCars;
    uint32_t stepCost;                             // work estimate of the current env step (rows, triangle candidates ...), see env_kernels.cuh
    BallDev ball;
    PadDev pads[kNumPads];
    // GameEventTracker (Sim/GameEventTracker/GameEventTracker.h:78-86)
    float shotCooldown;
    uint8_t ballShot, ballShotGoalTeam, ballScoredLast, _pad1;
    uint64_t lastBallUpdateCount;
    // env layer (GameState.h:22-43, NoTouchCondition.h)
    uint64_t gsLastTickCount;
    int32_t lastTouchCarId;
    float noTouchTimer[kMaxTerminals];
    V3 prevBallVel;
    uint8_t terminal, slotCar[kMaxCars], _pad2;    // slotCar[s] = car index (carId-1) of players[s]
    EnvPlayer players[kMaxCars];                   // indexed by SLOT
    uint32_t rng;                                  // minstd_rand0 state (respawn slot / kickoff permutation)
    float stepDeltaTime;                           // GameState.deltaTime of the current env step
    uint8_t stepGoalScored, hasResetState, gameMode /* ggl_game_mode, set at create */, _pad3;
    // per-tick scratch
    int32_t numContacts;
    uint32_t bodiesInContact;                      // bit b: body b (0 ball, 1 + c car c) has a solver row this tick
    uint8_t cntWorld, ballAwake, cntBall[kMaxCars], cntRest[kMaxCars];
    uint8_t rayBoxMask, _pad4;                     // per tick: bit c = car c has a tracked box a suspension ray could meet (sim_tick.cuh prologue)
    uint8_t island[8];                             // island[b] = lowest body index of the contact island body b belongs to
    uint32_t wheelDefer;                           // per tick: bit 4c+i = wheel i of car c stands on a car pre-ticked before c
    uint32_t contactOverflow;                      // STICKY: contact points dropped since create because windows and pool were full
    alignas(16) Contact contacts[kMaxContacts];    // from here to cars[]: per-tick scratch that never needs to travel to HBM
    uint8_t contactOrder[kMaxContacts];           // solver row k = contacts[contactOrder[k]], k < numContacts (gather_contacts)
    alignas(4) uint8_t poolOwner[kMaxPool];        // who claimed overflow-pool slot j (POOL_* codes, sim_world.cuh)
    uint32_t poolCount;                            // overflow-pool slots claimed this tick (atomic: the bodies' lanes race for them)
    int8_t padHit[8];                              // per tick: the boost pad car c collides with (-1 = none), see car_check_pads
    uint16_t ballCand[24];                         // per tick: candidate triangles of the ball (collide_ball_mesh)
    // car-vs-mesh staging: candidates per car (gathered by the car's lane), hits found by all lanes of the group
    uint16_t meshCand[kMaxCars][kMaxMeshCand];
    uint8_t meshCandN[kMaxCars], meshHitOverflow[kMaxCars];
    uint32_t meshHitN[kMaxCars];
    CarTriHit meshHit[kMaxCars][kMaxMeshHit];
    alignas(16) CarDev cars[kMaxCars];
};
static_assert(sizeof(CarDev) % 16 == 0, "CarDev must be a multiple of 16 bytes (vector staging)");
static_assert(sizeof(ArenaDev) % 16 == 0, "ArenaDev must be a multiple of 16 bytes (vector staging)");
GGL_HD size_t arena_hdr_bytes() { return offsetof(ArenaDev, cars); }
GGL_HD size_t arena_persist_bytes() { return offsetof(ArenaDev, contacts); }     // header part that is state, not scratch
GGL_HD size_t arena_slot_bytes(int numCars) { return offsetof(ArenaDev, cars) + (size_t)numCars * sizeof(CarDev); }

// ------------------------------------------------------------------ static world (shared by all arenas, read-only)
// edgeAngle / flags = btTriangleInfo of Bullet's internal-edge utility (btTriangleInfoMap.h), generated on the host at mesh load
enum : uint32_t { TRI_CONVEX_V0V1 = 1, TRI_CONVEX_V1V2 = 2, TRI_CONVEX_V2V0 = 4, TRI_SWAP_V0V1 = 8, TRI_SWAP_V1V2 = 16, TRI_SWAP_V2V0 = 32,
                  TRI_HAS_INFO = 0x100, TRI_MESH_SHIFT = 16 };
constexpr int kMaxWorldMeshes = 28;           // + 4 planes = the 32 bits of WorldDev::bpCellStatics
struct Tri {
    V3 v0, v1, v2;
    float edgeAngle[3];      // V0V1, V1V2, V2V0 (2*pi = no neighbour)
    uint32_t flags;          // TRI_* | meshIndex << TRI_MESH_SHIFT
    // what every suspension ray computes first about a triangle (btTriangleRaycastCallback::processTriangle): the unnormalised
    // normal (v1 - v0) x (v2 - v0) and dot(v0, normal), evaluated once at mesh load with the same float operations
    V3 rayN;
    float rayD;
};
struct WorldDev {
    const Tri* tris;       // all mesh triangles (Bullet units)
    int numTris;
    // uniform grid over triangles
    const int* cellStart;  // [gx*gy*gz + 1]
    const int* cellTris;   // triangle indices per cell
    int gx, gy, gz;
    V3 gridMin;
    float cellSize, invCellSize;
    // a box of EMPTY grid cells around the middle of the arena (sim_convert.cuh: world_find_free_box): an AABB inside it touches
    // no triangle, so cars, ball and suspension rays in open play skip the grid walk (and its global loads) altogether
    V3 freeMin, freeMax;
    // The reference's broadphase registry (btRSBroadphase.cpp:97-180): per broadphase cell (7.4 bt = ArenaConfig::maxAABBLen, grid
    // over ArenaConfig::minPos..maxPos) the static bodies entered in it, as a bit set (bit s = static s: the meshes in creation
    // order, then floor, ceiling, left wall, right wall - Arena.cpp:1036-1101), and every static's broadphase box (6 floats each).
    // Together they say which (body, static) pairs - and so which manifolds, empty ones included - the reference creates in a
    // tick, which decides the order its solver sees the rows in (solver_order, sim_world.cuh). Global memory: read on the rare
    // ticks in which one island holds rows of more than one manifold.
    const uint32_t* bpCellStatics;
    const float* bpStaticBox;
    int bpX, bpY, bpZ, numStatics;
    V3 bpMin;
    float bpCell;
};
GGL_HD bool world_aabb_is_free(const WorldDev& W, V3 mn, V3 mx) {
    return mn.x > W.freeMin.x && mn.y > W.freeMin.y && mn.z > W.freeMin.z && mx.x < W.freeMax.x && mx.y < W.freeMax.y && mx.z < W.freeMax.z;
}

}  // namespace ggl
