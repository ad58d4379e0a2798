/* gigaenv.h — C ABI of libgigaenv_cuda.so, the B200-native rollout hot path.
 *
 * Every entry point replaces one piece of the reference's CPU path (paths relative to
 * /root/reference/GigaLearnCPP):
 *
 *   ggl_envset_create            RLGymCPP/src/RLGymCPP/EnvSet/EnvSet.cpp:46-111   (EnvSet ctor)
 *   ggl_envset_step_first_half   EnvSet.cpp:113-130   (StepFirstHalf: snapshot prev, Arena::Step(actionDelay))
 *   ggl_envset_step_second_half  EnvSet.cpp:132-273   (parse actions, Arena::Step(rest), events, GameState,
 *                                                       terminals, weighted rewards, obs, masks)
 *   ggl_envset_reset             EnvSet.cpp:275-354   (Reset / ResetArena for arenas with terminal != 0)
 *   ggl_envset_sync              EnvSet.h:107         (Sync)
 *   ggl_envset_views             EnvSet.h:35-60       (EnvState buffers; here device pointers)
 *   ggl_envset_set_state/get_state   RocketSim Car::SetState/GetState (Sim/Car/Car.cpp:10-36),
 *                                Ball::SetState/GetState (Sim/Ball/Ball.cpp:27-49), BoostPad::SetState
 *   ggl_arena_step               RocketSim Arena::Step (Sim/Arena/Arena.cpp:716-812) — raw ticks, for parity tests
 *   ggl_gae_compute              src/private/GigaLearnCPP/PPO/GAE.cpp:7-208  (GAE::Compute)
 *   ggl_action_table             RLGymCPP/src/RLGymCPP/ActionParsers/DefaultAction.cpp:3-89
 *   ggl_expbuf_plan_batches / ggl_expbuf_gather   src/private/GigaLearnCPP/PPO/ExperienceBuffer.cpp:100-163 / :15-98
 *   ggl_traj_*                                     src/public/GigaLearnCPP/Learner.cpp:504-547, 701-712, 823-863 (trajectory assembly)
 *   ggl_obsstat_*                                  src/private/GigaLearnCPP/Util/WelfordStat.h:69-243 (BatchedWelfordStat), Learner.cpp:680-694
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a
 * negative ggl_status; nothing throws across the boundary; ggl_last_error() gives the text.
 * All `d_` pointers are DEVICE pointers on the envset's device; `stream` is a cudaStream_t
 * passed as void* (NULL = legacy default stream).
 *
 * The host-side state structs (ggl_car_state, ggl_arena_state) are in Unreal units exactly
 * like RocketSim's CarState/BallState; the device keeps Bullet units (uu/50) internally and
 * converts with the same float operations the reference uses, so set→get round-trips match
 * the reference bit for bit.
 */
#ifndef GIGAENV_H
#define GIGAENV_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGL_MAX_CARS      6      /* 3v3 */
#define GGL_NUM_PADS      34     /* soccar boost pads, RLConst.h:192-253 */
#define GGL_NUM_ACTIONS   90     /* DefaultAction table */
#define GGL_MAX_REWARDS   16
#define GGL_MAX_TERMINALS 4
#define GGL_MAX_SETTER_CHILDREN 4 /* CombinedState children */
#define GGL_MAX_OBS_PLAYERS 6    /* upper bound of DefaultObsPadded::maxPlayers (players per team the row has room for) */

typedef enum ggl_status {
    GGL_OK = 0,
    GGL_ERR_INVALID = -1,       /* bad argument */
    GGL_ERR_CUDA = -2,          /* CUDA runtime error, see ggl_last_error */
    GGL_ERR_UNSUPPORTED = -3,   /* plugin/descriptor kind with no device implementation */
    GGL_ERR_NOMEM = -4
} ggl_status;

/* ---- plugin descriptors (device translations of the RLGC plugin objects) ------------- */

typedef enum ggl_obs_kind {
    GGL_OBS_DEFAULT = 0,    /* RLGC::DefaultObs   (ObsBuilders/DefaultObs.cpp:4-54)   51+19*P floats */
    GGL_OBS_ADVANCED = 1,   /* RLGC::AdvancedObs  (ObsBuilders/AdvancedObs.cpp:193-270) 51+29*P floats */
    GGL_OBS_DEFAULT_PADDED = 2 /* RLGC::DefaultObsPadded (ObsBuilders/DefaultObsPadded.cpp:4-68) 51+38*maxPlayers floats: self, then
                               maxPlayers-1 teammate rows and maxPlayers opponent rows, zero rows for absent players, each list
                               std::shuffle'd per call with the arena's stream (the reference: the calling thread's engine) */
} ggl_obs_kind;

typedef enum ggl_reward_kind {   /* Rewards/CommonRewards.h */
    GGL_REW_PLAYER_GOAL = 0, GGL_REW_ASSIST, GGL_REW_SHOT, GGL_REW_SHOT_PASS, GGL_REW_SAVE,
    GGL_REW_BUMP, GGL_REW_BUMPED_PENALTY, GGL_REW_DEMO, GGL_REW_DEMOED_PENALTY,   /* :7-28 event rewards */
    GGL_REW_GOAL,                   /* :32-44  params[0]=concedeScale */
    GGL_REW_VELOCITY,               /* :47-54  params[0]=isNegative */
    GGL_REW_VELOCITY_BALL_TO_GOAL,  /* :57-72  params[0]=ownGoal */
    GGL_REW_VELOCITY_PLAYER_TO_BALL,/* :75-82 */
    GGL_REW_FACE_BALL,              /* :85-91 */
    GGL_REW_TOUCH_BALL,             /* :93-98 */
    GGL_REW_SPEED,                  /* :100-105 */
    GGL_REW_WAVEDASH,               /* :107-119 */
    GGL_REW_PICKUP_BOOST,           /* :121-133 */
    GGL_REW_SAVE_BOOST,             /* :136-144 params[0]=exponent */
    GGL_REW_AIR,                    /* :147-152 */
    GGL_REW_TOUCH_ACCEL,            /* :157-178 */
    GGL_REW_STRONG_TOUCH,           /* :180-202 params[0]=minRewardedVel params[1]=maxRewardedVel (uu/s) */
    GGL_REW_KICKOFF_PROXIMITY_2V2,  /* Rewards/KickoffProximityReward2v2Enhanced.h:14-367 params = goerReward, cheaterReward,
                                       dynamicWeight, rotationPrepWeight (defaults 1.2, 0.6, 0.3, 0.2) */
    GGL_REW_KIND_COUNT
} ggl_reward_kind;

typedef struct ggl_reward_desc {
    int32_t kind;           /* ggl_reward_kind */
    float   weight;         /* WeightedReward::weight (Reward.h:84-90) */
    float   params[4];
    int32_t zero_sum;       /* wrapped in RLGC::ZeroSumReward: a pass-through in the EnvSet path */
} ggl_reward_desc;

typedef enum ggl_terminal_kind {
    GGL_TERM_GOAL_SCORE = 0,    /* TerminalConditions/GoalScoreCondition.h:8-14   NORMAL */
    GGL_TERM_NO_TOUCH = 1       /* TerminalConditions/NoTouchCondition.h:18-28    TRUNCATED, params[0]=maxTime */
} ggl_terminal_kind;

typedef struct ggl_terminal_desc {
    int32_t kind;
    float   params[2];
} ggl_terminal_desc;

typedef enum ggl_game_mode {     /* RocketSim GameMode (RocketSim.h); the device integrator has these two */
    GGL_MODE_SOCCAR = 0,
    GGL_MODE_HEATSEEKER = 1      /* soccar arena; the ball steers towards a goal (Sim/Ball/Ball.cpp:152-253), unlimited boost and
                                    respawns at 100 (MutatorConfig.cpp:29-33), 4 kickoff slots and a moving kickoff ball (Arena.cpp:132-200) */
} ggl_game_mode;

typedef enum ggl_state_setter_kind {
    GGL_SETTER_KICKOFF = 0,     /* StateSetters/KickoffState.h → Arena::ResetToRandomKickoff (Arena.cpp:112-216) */
    GGL_SETTER_INJECTED = 1,    /* reset states are supplied by the host through ggl_envset_set_reset_state */
    GGL_SETTER_FUZZED_KICKOFF = 2, /* StateSetters/FuzzedKickoffState.h:7-27: kickoff, then every car's position +- 0.1 uu per axis */
    GGL_SETTER_RANDOM = 3,      /* StateSetters/RandomState.cpp:11-62, flags = ggl_random_state_flags */
    GGL_SETTER_COMBINED = 4     /* StateSetters/CombinedState.h:10-46: one child per reset, drawn by weight */
} ggl_state_setter_kind;
typedef enum ggl_random_state_flags {   /* RandomState ctor arguments (RandomState.h:9-11) */
    GGL_RANDOM_BALL_SPEED = 1, GGL_RANDOM_CAR_SPEED = 2, GGL_RANDOM_CARS_ON_GROUND = 4
} ggl_random_state_flags;
typedef struct ggl_setter_child {  /* one (StateSetter*, weight) pair of CombinedState's ctor */
    int32_t kind;               /* GGL_SETTER_KICKOFF, _FUZZED_KICKOFF, _RANDOM or _INJECTED */
    int32_t flags;              /* GGL_SETTER_RANDOM: ggl_random_state_flags */
    float   weight;             /* >= 0, finite (CombinedState.h:25-27) */
} ggl_setter_child;

/* Arena collision meshes: the reference's .cmf layout (CollisionMeshFile.cpp:11-35):
 * int32 numTris, int32 numVerts, int32 tri[numTris][3], float vert[numVerts][3] (Bullet units). */
typedef struct ggl_mesh_blob {
    const uint8_t* data;
    size_t         size;
} ggl_mesh_blob;

typedef struct ggl_envset_desc {
    int32_t device;             /* CUDA device ordinal */
    int32_t num_arenas;
    int32_t players_per_team;   /* 1..3 (blue cars are added first, then orange: carId 1..P blue, P+1..2P orange) */
    int32_t tick_skip;          /* EnvSetConfig::tickSkip */
    int32_t action_delay;       /* EnvSetConfig::actionDelay */
    int32_t obs_kind;           /* ggl_obs_kind */
    int32_t num_rewards;
    ggl_reward_desc rewards[GGL_MAX_REWARDS];
    int32_t num_terminals;
    ggl_terminal_desc terminals[GGL_MAX_TERMINALS];
    int32_t state_setter;       /* ggl_state_setter_kind */
    int32_t num_meshes;
    const ggl_mesh_blob* meshes;
    uint64_t seed;              /* device RNG seed (kickoff permutation, respawn slot) */
    int32_t slot_order_reversed;/* default players[] slot order: 1 = reverse AddCar order, 0 = AddCar order. The reference's
                                   order is the iteration order of std::unordered_set<Car*> — it depends on heap
                                   addresses and differs per arena, so it is DATA: override with ggl_envset_set_slot_order */
    int32_t obs_max_players;    /* GGL_OBS_DEFAULT_PADDED: DefaultObsPadded::maxPlayers, players_per_team..GGL_MAX_OBS_PLAYERS */
    int32_t state_setter_flags; /* GGL_SETTER_RANDOM: ggl_random_state_flags */
    int32_t num_setter_children;/* GGL_SETTER_COMBINED: 1..GGL_MAX_SETTER_CHILDREN */
    ggl_setter_child setter_children[GGL_MAX_SETTER_CHILDREN];
    int32_t game_mode;          /* ggl_game_mode */
    float   obs_coefs[5];       /* GGL_OBS_DEFAULT / _DEFAULT_PADDED: the constructor's posCoef x, y, z, velCoef, angVelCoef
                                   (DefaultObs.h:9-17); all five 0 = the defaults 1/4096, 1/5120, 1/2044, 1/2300, 1/5.5 */
} ggl_envset_desc;

/* ---- host-visible state (mirrors RocketSim CarState/BallState/BoostPadState + hidden wheel state) ---- */

typedef struct ggl_controls {   /* Sim/CarControls.h */
    float throttle, steer, pitch, yaw, roll;
    uint8_t jump, boost, handbrake, _pad;
} ggl_controls;

typedef struct ggl_wheel_state { /* btWheelInfoRL fields that survive from one tick to the next */
    float steer_angle, engine_force, brake, lat_friction, long_friction, extra_pushback;
} ggl_wheel_state;

typedef struct ggl_car_state {   /* Sim/Car/Car.h:17-115 */
    float pos[3];
    float rot[9];                /* forward, right, up (RotMat, column vectors of the Bullet basis) */
    float vel[3];
    float ang_vel[3];
    float flip_rel_torque[3];
    float jump_time, flip_time, air_time, air_time_since_jump;
    float boost, time_spent_boosting, supersonic_time, handbrake_val;
    float auto_flip_timer, auto_flip_torque_scale;
    float world_contact_normal[3];
    float car_contact_cooldown;
    float demo_respawn_timer;
    float hit_rel_pos[3], hit_ball_pos[3], hit_extra_vel[3];   /* BallHitInfo */
    uint64_t hit_tick, hit_tick_extra;                         /* ~0 when never */
    uint32_t car_contact_other_id;
    uint32_t id;
    uint8_t team;
    uint8_t is_on_ground, has_jumped, has_double_jumped, has_flipped, is_flipping, is_jumping;
    uint8_t is_supersonic, is_auto_flipping, world_contact_has, is_demoed, hit_valid;
    uint8_t wheels_with_contact[4];
    ggl_controls last_controls;
    ggl_controls controls;
    ggl_wheel_state wheels[4];
    uint64_t update_counter;
    /* Hidden state like the wheels': the car's broadphase box as the LAST tick's collision detection left it (Bullet units). The
     * suspension rays of the next tick are tested against it before the car's shape (btCollisionWorld.cpp:913-945), and nothing
     * refreshes it in between - so it decides which rays can meet this car at all. All six 0 = unknown: the box of the current
     * transform stands in. min.x > max.x = no other car was within reach (the device does not track the box then). */
    float ray_aabb_min[3], ray_aabb_max[3];
} ggl_car_state;

typedef struct ggl_ball_state {  /* Sim/Ball/Ball.h:16-41 */
    float pos[3], vel[3], ang_vel[3];
    uint64_t update_counter;
    float hs_y_target_dir, hs_cur_target_speed, hs_time_since_hit;   /* BallState::HeatseekerInfo (:23-32); a default ball has 0, 2900, 0 */
    float _pad;
} ggl_ball_state;

typedef struct ggl_pad_state {   /* Sim/BoostPad/BoostPad.h:23-34, RocketSim pad order (6 big then 28 small) */
    uint8_t  is_active;
    uint8_t  _pad[3];
    float    cooldown;
    uint32_t prev_locked_car_id;
} ggl_pad_state;

typedef struct ggl_arena_state {
    uint64_t tick_count;
    int32_t  num_cars;
    int32_t  _pad;
    ggl_ball_state ball;
    ggl_car_state  cars[GGL_MAX_CARS];      /* AddCar order: index = carId-1 */
    ggl_pad_state  pads[GGL_NUM_PADS];
} ggl_arena_state;

/* ---- device views of EnvState (EnvSet.h:35-60) ---- */
typedef struct ggl_envset_views {
    int32_t num_arenas, num_players, obs_size, num_actions, players_per_arena;
    float*   d_obs;          /* [num_players, obs_size] row-major */
    uint8_t* d_action_masks; /* [num_players, num_actions] */
    float*   d_rewards;      /* [num_players] */
    uint8_t* d_terminals;    /* [num_arenas] 0 / 1 NORMAL / 2 TRUNCATED */
    int32_t* d_slot_car_id;  /* [num_players] carId of players[] slot */
    uint8_t* d_slot_team;    /* [num_players] */
    uint8_t* d_events;       /* [num_players, 9] PlayerEventState: goal,save,assist,shot,shotPass,bump,bumped,demo,demoed */
    float*   d_last_rewards; /* [num_arenas, num_rewards] raw (unweighted) reward of slot 0 per reward fn */
} ggl_envset_views;

typedef struct ggl_envset ggl_envset;

int  ggl_envset_create(const ggl_envset_desc* desc, ggl_envset** out);
int  ggl_envset_destroy(ggl_envset* env);
int  ggl_envset_step_first_half(ggl_envset* env, void* stream);
int  ggl_envset_step_second_half(ggl_envset* env, const int32_t* d_actions, void* stream);
/* One fused launch = Reset + StepFirstHalf + StepSecondHalf (the Learner loop of Learner.cpp:669-793 with
 * the action indices already on the device). */
int  ggl_envset_step(ggl_envset* env, const int32_t* d_actions, void* stream);
int  ggl_envset_reset(ggl_envset* env, void* stream);
/* EnvSet::ResetArena(index) called directly (EnvSet.h:108, EnvSet.cpp:275-331): resets `arena` whatever its terminal flag
 * (state setter, tracker and plugin resets, obs + mask rows rebuilt, terminal cleared); arena = -1 resets every arena. */
int  ggl_envset_reset_arena(ggl_envset* env, int arena, void* stream);
int  ggl_envset_sync(ggl_envset* env);
int  ggl_envset_views_get(ggl_envset* env, ggl_envset_views* out);
int  ggl_envset_set_state(ggl_envset* env, int arena, const ggl_arena_state* host_state);
int  ggl_envset_get_state(ggl_envset* env, int arena, ggl_arena_state* host_state);
/* test / debug: ordered solver contact rows of `arena`'s last tick, out[max][20] floats per row (bodyA, bodyB, normal xyz,
 * relA xyz, relB xyz, distance, friction, restitution, special); returns the row count (>= 0) or a negative ggl_status */
int  ggl_envset_get_contacts(ggl_envset* env, int arena, float* out, int max);
/* Contact points dropped since create because an arena's contact rows were exhausted (per-body windows + shared overflow pool,
 * 36 rows per arena and tick), summed over the arenas. The reference has no such limit (btPersistentManifold pool); a non-zero
 * value means some arena's solver saw fewer rows than the reference's would. Synchronises the device. */
int  ggl_envset_contact_overflow(ggl_envset* env, uint64_t* out_dropped);
/* players[] slot order of every arena: car_ids[num_players], car_ids[arena*P + slot] = carId (1-based AddCar order)
 * occupying GameState.players[slot] (GameState.cpp:82-86). Takes effect at the next reset/step. */
int  ggl_envset_set_slot_order(ggl_envset* env, const int32_t* h_car_ids);
/* The arena's random stream: one std::minstd_rand0 (= libstdc++'s std::default_random_engine, the reference's engine type,
 * Math.cpp:59-64) per arena, seeded from desc.seed, drawn from in the reference's order by everything on the path that is
 * random there: the respawn slot (Car.cpp:48), the kickoff shuffle (Arena.cpp:136), FuzzedKickoffState / RandomState /
 * CombinedState, DefaultObsPadded's two shuffles per row. The reference shares one time-seeded engine per worker thread; with
 * one arena per thread and the same engine state both sides draw the same numbers (how the parity tests pin these plugins).
 * state is the engine's raw state, 1..2147483646; arena = -1 in set_rng sets every arena. */
int  ggl_envset_set_rng(ggl_envset* env, int arena, uint32_t state);
int  ggl_envset_get_rng(ggl_envset* env, int arena, uint32_t* out_state);
/* state the next reset of `arena` will install when state_setter == GGL_SETTER_INJECTED */
int  ggl_envset_set_reset_state(ggl_envset* env, int arena, const ggl_arena_state* host_state);
/* raw Arena::Step(ticks) on every arena with the cars' current `controls` (no env bookkeeping) */
int  ggl_arena_step(ggl_envset* env, int ticks, void* stream);
/* device-side uniformly random mask-valid actions (bench / smoke helper; Philox-free LCG hash on (seed, step, slot)) */
int  ggl_envset_random_actions(ggl_envset* env, int32_t* d_actions, uint64_t seed, void* stream);

/* the same rule on HOST mask rows (policy stand-in for the host-buffer benchmark / examples; no device work) */
int  ggl_host_random_actions(const uint8_t* h_masks, int32_t* h_actions, int num_players, int num_actions, uint64_t seed);

/* Host-convenience wrappers through HOST buffers (the reference-facing call: EnvState vectors live on the host,
 * EnvSet.h:35-60). actions: int32[num_players]; outputs may be NULL to skip the copy. */
int  ggl_envset_step_host(ggl_envset* env, const int32_t* h_actions,
                          float* h_obs, uint8_t* h_masks, float* h_rewards, uint8_t* h_terminals);

/* StepSecondHalf with HOST action indices (EnvSet.cpp:132: `const IList& actionIndices`): H2D of the actions, the second-half
 * launch, and a wait for it. Outputs are then fetched with ggl_envset_read_host. */
int  ggl_envset_step_second_half_host(ggl_envset* env, const int32_t* h_actions);
/* copies the EnvState buffers to HOST memory (any pointer may be NULL to skip it) after waiting for pending work */
int  ggl_envset_read_host(ggl_envset* env, float* h_obs, uint8_t* h_masks, float* h_rewards, uint8_t* h_terminals);

/* The Learner's loop through HOST buffers, pipelined like the reference (Learner.cpp:669-793: Reset, StepFirstHalf(async),
 * policy inference on state.obs while the first half runs, Sync, StepSecondHalf):
 *   begin  = Reset + copy of the post-reset obs / action masks to the host + launch of the first half; returns as soon as
 *            obs and masks have landed, with the first half still running (it never writes obs or masks, EnvSet.cpp:113-130)
 *   finish = H2D of the action indices, second half, D2H of rewards / terminals (and, when h_final_obs is not NULL, of the
 *            obs rows as they stand after the step = the final obs of episodes that just ended); waits for completion.
 * Host pointers should be pinned for the copies to overlap. */
int  ggl_envset_begin_step_host(ggl_envset* env, float* h_obs, uint8_t* h_masks);
int  ggl_envset_finish_step_host(ggl_envset* env, const int32_t* h_actions, float* h_rewards, uint8_t* h_terminals, float* h_final_obs);

/* ---- DefaultAction ---- */
/* table: float[90*8] (throttle, steer, pitch, yaw, roll, jump, boost, handbrake) */
int  ggl_action_table(float* out_table, int* out_num_actions);

/* ---- GAE (GAE.cpp:7-208) ----
 * d_rews f32[L], d_terminals i8[L] (0 / 1 NORMAL / 2 TRUNCATED), d_val_preds f32[L],
 * d_trunc_val_preds f32[K] or NULL (then truncated steps bootstrap from val_preds[i+1] like non-terminal ones),
 * outputs f32[L]; d_out_clip_portion f32[1] (may be NULL).
 * d_workspace: ggl_gae_workspace_bytes(L) bytes of device scratch.
 * Returns GGL_ERR_INVALID if K != number of TRUNCATED entries is detected (checked asynchronously: the count
 * is written to d_out_trunc_count i32[1] when non-NULL, mirroring GAE.cpp:196-197). */
size_t ggl_gae_workspace_bytes(int64_t L);
int  ggl_gae_compute(const float* d_rews, const int8_t* d_terminals, const float* d_val_preds,
                     const float* d_trunc_val_preds, int64_t L, int64_t K,
                     float gamma, float lambda, float return_std, float clip_range,
                     float* d_out_advantages, float* d_out_returns, float* d_out_target_values,
                     float* d_out_clip_portion, int32_t* d_out_trunc_count,
                     void* d_workspace, void* stream);
/* ggl_gae_compute without a workspace argument and without any allocation. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it). The scan's scratch (tile look-back records and TRUNCATED counters) lives in a file-scope
 * `static __device__ __align__(256)` array of GGL_GAE_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit;
 * its device address is resolved once per device with cudaGetSymbolAddress. L > GGL_GAE_STATIC_MAX_L returns GGL_ERR_INVALID
 * with a message. The call makes no cudaMalloc / cudaFree / allocating library call; inputs and outputs are the caller's, same
 * meaning and alignment rules as ggl_gae_compute. All calls on one device share the array: the part that must start at zero
 * is cleared on `stream` at the start of every call, and two calls that overlap on the GPU (different streams, no ordering
 * between them) race on it - serialise them. */
#define GGL_GAE_STATIC_MAX_L          (1ll << 24)     /* elements per call (4096 tiles of 4096) */
#define GGL_GAE_STATIC_SCRATCH_BYTES  (128 * 1024)    /* per GPU, resident until the library is unloaded */
int  ggl_gae_compute_static(const float* d_rews, const int8_t* d_terminals, const float* d_val_preds,
                            const float* d_trunc_val_preds, int64_t L, int64_t K,
                            float gamma, float lambda, float return_std, float clip_range,
                            float* d_out_advantages, float* d_out_returns, float* d_out_target_values,
                            float* d_out_clip_portion, int32_t* d_out_trunc_count, void* stream);
/* Host-buffer wrapper (the reference-facing call: GAE::Compute takes CPU tensors). */
int  ggl_gae_compute_host(const float* h_rews, const int8_t* h_terminals, const float* h_val_preds,
                          const float* h_trunc_val_preds, int64_t L, int64_t K,
                          float gamma, float lambda, float return_std, float clip_range,
                          float* h_out_advantages, float* h_out_returns, float* h_out_target_values,
                          float* h_out_clip_portion);

/* ---- ExperienceBuffer batching (ExperienceBuffer.cpp:15-163) ----
 * The six experience tensors (ExperienceBuffer.h:8-15, set at Learner.cpp:975-980); a NULL member is an undefined tensor.
 * const-ness is nominal: the same struct describes the sources and the destinations. */
typedef struct ggl_expbuf_tensors {
    float*   states;         /* [rows, obs_size] */
    int32_t* actions;        /* [rows_actions] */
    float*   log_probs;      /* [rows] */
    float*   target_values;  /* [rows] */
    uint8_t* action_masks;   /* [rows, num_actions] */
    float*   advantages;     /* [rows] */
} ggl_expbuf_tensors;
/* GetAllBatchesShuffled's partition (:143-160) of a permuted stream of exp_size = min(rows of states, rows of actions) samples
 * into minibatches: fills out_starts/out_sizes (up to max_batches entries, either may be NULL) and returns the number of
 * batches (>= 0) or a negative ggl_status. Pure host arithmetic. */
int64_t ggl_expbuf_plan_batches(int64_t exp_size, int64_t batch_size, int overbatching, int64_t* out_starts, int64_t* out_sizes,
                                int64_t max_batches);
/* One fused gather of `count` samples: dst.X[i] = src.X[d_indices[i]] for every defined tensor, indices >= the row limit clamped
 * to the last row (:42-69), actions clamped to [0, max_action_index] when max_action_index >= 0 (:88-90). All pointers are
 * DEVICE pointers; d_indices is int64 like the reference's index vector and must be >= 0. Minibatch b of the epoch is the
 * view [start_b, start_b + size_b) of the destinations. */
int  ggl_expbuf_gather(const ggl_expbuf_tensors* src, int64_t rows_states, int64_t rows_actions, int32_t obs_size, int32_t num_actions,
                       const int64_t* d_indices, int64_t count, int64_t max_action_index, const ggl_expbuf_tensors* dst, void* stream);

/* ---- Trajectory assembly: the order of the flat experience stream (src/public/GigaLearnCPP/Learner.cpp:504-547, 701-712, 823-863) ----
 * The Learner keeps one growing buffer per player (`trajectories[p]`), appends it to `combinedTraj` in player order whenever that
 * player's episode ends (arena terminal, or forced TRUNCATED once the buffer holds maxEpisodeLength steps, :848-850) and stops
 * collecting at the first step after which combinedTraj holds >= tsPerItr timesteps (:665). Here both levels live on the device:
 * per-player buffers [num_players][max_episode_length] and the combined stream; a finished segment is moved by a device copy at
 * the step it finishes, so the combined tensors are at any time exactly what the reference's vectors would hold. */
typedef struct ggl_traj ggl_traj;
typedef struct ggl_traj_desc {
    int32_t device;
    int32_t num_players;          /* envSet->state.numPlayers */
    int32_t players_per_arena;    /* terminals are per arena (EnvSet.h:35-60); player p sits in arena p / players_per_arena */
    int32_t obs_size, num_actions;
    int32_t max_episode_length;   /* (int)(maxEpisodeDuration * 120 / tickSkip), Learner.cpp:550; also the per-player buffer depth */
    int64_t capacity_timesteps;   /* rows of the combined stream: >= tsPerItr + what one step can append */
    int64_t capacity_truncated;   /* rows of nextStates */
    int64_t capacity_player_rows; /* steps the per-player buffers hold IN TOTAL (a pool of 16-step pages shared by all players):
                                     0 = worst case, num_players x max_episode_length (22 GB at 16 384 players x 1 800 steps x
                                     obs 167); > 0 = the steps in flight one expects (players x episode length, with margin).
                                     A pool that runs out makes ggl_traj_counts fail - nothing is dropped silently */
} ggl_traj_desc;
typedef struct ggl_traj_views {   /* DEVICE pointers into the combined stream, valid until destroy; rows [0, length) are defined */
    float*   states;        /* [capacity_timesteps, obs_size]   Trajectory::states      */
    uint8_t* action_masks;  /* [capacity_timesteps, num_actions]            actionMasks */
    int32_t* actions;       /* [capacity_timesteps]                          actions     */
    float*   log_probs;     /* [capacity_timesteps]                          logProbs    */
    float*   rewards;       /* [capacity_timesteps]                          rewards     */
    int8_t*  terminals;     /* [capacity_timesteps]  0 / 1 NORMAL / 2 TRUNCATED          */
    float*   next_states;   /* [capacity_truncated, obs_size]   final obs of every TRUNCATED segment, in stream order (:855-858) */
} ggl_traj_views;
int  ggl_traj_create(const ggl_traj_desc* d, ggl_traj** out);
int  ggl_traj_destroy(ggl_traj* t);
int  ggl_traj_views_get(ggl_traj* t, ggl_traj_views* out);
/* Start of a collection iteration: combinedTraj.Clear() (:631). Per-player buffers keep their unfinished episodes. */
int  ggl_traj_begin_iteration(ggl_traj* t, void* stream);
/* Before the env step (:701-712): append every player's current obs row and mask row to its buffer. d_obs f32[num_players,
 * obs_size], d_masks u8[num_players, num_actions] are the envset's output buffers after Reset. */
int  ggl_traj_push_obs(ggl_traj* t, const float* d_obs, const uint8_t* d_masks, void* stream);
/* After the env step (:823-863): append action, reward, log prob and the terminal type (arena terminal, else TRUNCATED when the
 * buffer reached max_episode_length); move finished buffers to the combined stream in player order, push the final obs row of
 * TRUNCATED ones (d_obs_final = the envset's obs after the step) to next_states. d_terminals u8[num_arenas]. */
int  ggl_traj_push_result(ggl_traj* t, const int32_t* d_actions, const float* d_log_probs, const float* d_rewards,
                          const uint8_t* d_terminals, const float* d_obs_final, void* stream);
/* combinedTraj.Length() and the number of next_states rows after the pushes queued so far on `stream` (synchronises it).
 * GGL_ERR_INVALID if a push ran out of capacity (nothing is written past the end). */
int  ggl_traj_counts(ggl_traj* t, void* stream, int64_t* out_length, int64_t* out_truncated);
/* debug / tests: current length of every player's unfinished buffer (host int32[num_players]) */
int  ggl_traj_player_lengths(ggl_traj* t, int32_t* h_out);

/* ---- Running obs statistics + in-place normalisation (src/private/GigaLearnCPP/Util/WelfordStat.h:69-243 BatchedWelfordStat, as
 * used every collection step at src/public/GigaLearnCPP/Learner.cpp:680-694) ----
 * increment_rows = IncrementRow (:89-97) over d_obs[d_row_indices[k]] for k = 0..n-1 IN ORDER (the row choice, Math::RandInt on the
 * host, is an input); normalize = NormalizeInPlace (:131-216): no-op while count < 2, else obs[i][j] = obs[i][j] * inv[j] + neg[j]
 * with the clamped-mean / floored-std coefficients. Double-precision state, bit-exact against the reference. */
typedef struct ggl_obsstat ggl_obsstat;
int  ggl_obsstat_create(int device, int width, ggl_obsstat** out);
int  ggl_obsstat_destroy(ggl_obsstat* s);
int  ggl_obsstat_increment_rows(ggl_obsstat* s, const float* d_obs, const int32_t* d_row_indices, int32_t n, void* stream);
int  ggl_obsstat_normalize(ggl_obsstat* s, float* d_obs, int64_t num_rows, double mean_clamp, double min_std, void* stream);
/* runningMeans / runningVariances / count to and from the host (ToJSON / ReadFromJSON, :218-232) */
int  ggl_obsstat_get(ggl_obsstat* s, double* h_means, double* h_vars, int64_t* h_count);
int  ggl_obsstat_set(ggl_obsstat* s, const double* h_means, const double* h_vars, int64_t count);

/* Device-memory plumbing for host code that links this library and nothing of CUDA (the g++-only mirrors under host/):
 * cudaMalloc / cudaFree / cudaMemcpyAsync H2D (asynchronous on `stream`) / D2H (synchronises `stream`). */
int  ggl_dev_malloc(int device, size_t bytes, void** out);
int  ggl_dev_free(void* p);
int  ggl_dev_upload(void* d_dst, const void* h_src, size_t bytes, void* stream);
int  ggl_dev_download(void* h_dst, const void* d_src, size_t bytes, void* stream);

/* kernels launched by this library since load (bench.py's gpu_launches) */
uint64_t ggl_launch_count(void);
const char* ggl_last_error(void);
const char* ggl_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GIGAENV_H */
