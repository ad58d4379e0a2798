// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
// C-ABI harness around the UNMODIFIED reference (RocketSim + Bullet + RLGymCPP) compiled from
// /root/reference by oracle/Makefile into oracle/_ref/libref_oracle.so.  Only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs may load it.
//
// It drives the reference's own RLGC::EnvSet (EnvSet.cpp) with the same descriptor struct the
// CUDA library takes (include/gigaenv.h) and adds full state injection/extraction, including the
// hidden btVehicleRL wheel state that Car::SetState does not touch.
#include <RLGymCPP/EnvSet/EnvSet.h>
#include <RLGymCPP/ObsBuilders/DefaultObs.h>
#include <RLGymCPP/ObsBuilders/AdvancedObs.h>
#include <RLGymCPP/ActionParsers/DefaultAction.h>
#include <RLGymCPP/Rewards/CommonRewards.h>
#include <RLGymCPP/Rewards/ZeroSumReward.h>
#include <RLGymCPP/Rewards/KickoffProximityReward2v2Enhanced.h>
#include <RLGymCPP/TerminalConditions/GoalScoreCondition.h>
#include <RLGymCPP/TerminalConditions/NoTouchCondition.h>
#include <RLGymCPP/StateSetters/KickoffState.h>
#include <RLGymCPP/StateSetters/FuzzedKickoffState.h>
#include <RLGymCPP/StateSetters/RandomState.h>
#include <RLGymCPP/StateSetters/CombinedState.h>
#include <RLGymCPP/ObsBuilders/DefaultObsPadded.h>
#include <RLGymCPP/ThreadPool.h>

#include "gigaenv.h"

#include <chrono>
#include <random>

using namespace RLGC;

namespace {

struct RefEnv;

// StateSetter that installs a host-provided full arena state (SURVEY §7 hard part 3: state injection
// instead of re-implementing libstdc++'s std::shuffle on the device).
void ApplyArenaState(Arena* arena, const ggl_arena_state& st);

struct InjectedStateSetter : StateSetter {
    RefEnv* env;
    int arenaIdx;
    void ResetArena(Arena* arena) override;
};

struct RefEnv {
    EnvSet* envSet = nullptr;
    ggl_envset_desc desc;
    std::vector<ggl_arena_state> resetStates;
    std::vector<uint8_t> hasResetState;
    std::vector<std::vector<Car*>> carsById;   // [arena][carId-1]
};

void InjectedStateSetter::ResetArena(Arena* arena) {
    if (env && arenaIdx < (int)env->hasResetState.size() && env->hasResetState[arenaIdx])
        ApplyArenaState(arena, env->resetStates[arenaIdx]);
    else
        arena->ResetToRandomKickoff();
}

Vec V3(const float* f) { return Vec(f[0], f[1], f[2]); }
void Put3(float* o, const Vec& v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }

CarControls ToControls(const ggl_controls& c) {
    CarControls r;
    r.throttle = c.throttle; r.steer = c.steer; r.pitch = c.pitch; r.yaw = c.yaw; r.roll = c.roll;
    r.jump = c.jump; r.boost = c.boost; r.handbrake = c.handbrake;
    return r;
}
ggl_controls FromControls(const CarControls& c) {
    ggl_controls r = {};
    r.throttle = c.throttle; r.steer = c.steer; r.pitch = c.pitch; r.yaw = c.yaw; r.roll = c.roll;
    r.jump = c.jump; r.boost = c.boost; r.handbrake = c.handbrake;
    return r;
}

Car* CarById(Arena* arena, uint32_t id) {
    auto it = arena->_carIDMap.find(id);
    return it == arena->_carIDMap.end() ? nullptr : it->second;
}

void ApplyCarState(Car* car, const ggl_car_state& s) {
    CarState cs;
    cs.pos = V3(s.pos);
    cs.rotMat = RotMat(V3(s.rot), V3(s.rot + 3), V3(s.rot + 6));
    cs.vel = V3(s.vel);
    cs.angVel = V3(s.ang_vel);
    cs.isOnGround = s.is_on_ground;
    for (int i = 0; i < 4; i++) cs.wheelsWithContact[i] = s.wheels_with_contact[i];
    cs.hasJumped = s.has_jumped; cs.hasDoubleJumped = s.has_double_jumped; cs.hasFlipped = s.has_flipped;
    cs.flipRelTorque = V3(s.flip_rel_torque);
    cs.jumpTime = s.jump_time; cs.flipTime = s.flip_time;
    cs.isFlipping = s.is_flipping; cs.isJumping = s.is_jumping;
    cs.airTime = s.air_time; cs.airTimeSinceJump = s.air_time_since_jump;
    cs.boost = s.boost; cs.timeSpentBoosting = s.time_spent_boosting;
    cs.isSupersonic = s.is_supersonic; cs.supersonicTime = s.supersonic_time;
    cs.handbrakeVal = s.handbrake_val;
    cs.isAutoFlipping = s.is_auto_flipping; cs.autoFlipTimer = s.auto_flip_timer; cs.autoFlipTorqueScale = s.auto_flip_torque_scale;
    cs.worldContact.hasContact = s.world_contact_has; cs.worldContact.contactNormal = V3(s.world_contact_normal);
    cs.carContact.otherCarID = s.car_contact_other_id; cs.carContact.cooldownTimer = s.car_contact_cooldown;
    cs.isDemoed = s.is_demoed; cs.demoRespawnTimer = s.demo_respawn_timer;
    cs.ballHitInfo.isValid = s.hit_valid;
    cs.ballHitInfo.relativePosOnBall = V3(s.hit_rel_pos);
    cs.ballHitInfo.ballPos = V3(s.hit_ball_pos);
    cs.ballHitInfo.extraHitVel = V3(s.hit_extra_vel);
    cs.ballHitInfo.tickCountWhenHit = s.hit_tick;
    cs.ballHitInfo.tickCountWhenExtraImpulseApplied = s.hit_tick_extra;
    cs.lastControls = ToControls(s.last_controls);
    car->SetState(cs);
    car->_internalState.updateCounter = s.update_counter;
    car->controls = ToControls(s.controls);
    for (int i = 0; i < 4; i++) {
        auto& w = car->_bulletVehicle.m_wheelInfo[i];
        w.m_steerAngle = s.wheels[i].steer_angle;
        w.m_engineForce = s.wheels[i].engine_force;
        w.m_brake = s.wheels[i].brake;
        w.m_latFriction = s.wheels[i].lat_friction;
        w.m_longFriction = s.wheels[i].long_friction;
        w.m_extraPushback = s.wheels[i].extra_pushback;
    }
}

void ReadCarState(Car* car, ggl_car_state& s) {
    memset(&s, 0, sizeof(s));
    CarState cs = car->GetState();
    Put3(s.pos, cs.pos);
    Put3(s.rot, cs.rotMat.forward); Put3(s.rot + 3, cs.rotMat.right); Put3(s.rot + 6, cs.rotMat.up);
    Put3(s.vel, cs.vel); Put3(s.ang_vel, cs.angVel);
    s.is_on_ground = cs.isOnGround;
    for (int i = 0; i < 4; i++) s.wheels_with_contact[i] = cs.wheelsWithContact[i];
    s.has_jumped = cs.hasJumped; s.has_double_jumped = cs.hasDoubleJumped; s.has_flipped = cs.hasFlipped;
    Put3(s.flip_rel_torque, cs.flipRelTorque);
    s.jump_time = cs.jumpTime; s.flip_time = cs.flipTime;
    s.is_flipping = cs.isFlipping; s.is_jumping = cs.isJumping;
    s.air_time = cs.airTime; s.air_time_since_jump = cs.airTimeSinceJump;
    s.boost = cs.boost; s.time_spent_boosting = cs.timeSpentBoosting;
    s.is_supersonic = cs.isSupersonic; s.supersonic_time = cs.supersonicTime;
    s.handbrake_val = cs.handbrakeVal;
    s.is_auto_flipping = cs.isAutoFlipping; s.auto_flip_timer = cs.autoFlipTimer; s.auto_flip_torque_scale = cs.autoFlipTorqueScale;
    s.world_contact_has = cs.worldContact.hasContact; Put3(s.world_contact_normal, cs.worldContact.contactNormal);
    s.car_contact_other_id = cs.carContact.otherCarID; s.car_contact_cooldown = cs.carContact.cooldownTimer;
    s.is_demoed = cs.isDemoed; s.demo_respawn_timer = cs.demoRespawnTimer;
    s.hit_valid = cs.ballHitInfo.isValid;
    Put3(s.hit_rel_pos, cs.ballHitInfo.relativePosOnBall);
    Put3(s.hit_ball_pos, cs.ballHitInfo.ballPos);
    Put3(s.hit_extra_vel, cs.ballHitInfo.extraHitVel);
    s.hit_tick = cs.ballHitInfo.tickCountWhenHit;
    s.hit_tick_extra = cs.ballHitInfo.tickCountWhenExtraImpulseApplied;
    s.last_controls = FromControls(cs.lastControls);
    s.controls = FromControls(car->controls);
    s.update_counter = cs.updateCounter;
    if (btBroadphaseProxy* px = car->_rigidBody.getBroadphaseHandle()) {
        s.ray_aabb_min[0] = px->m_aabbMin.x(); s.ray_aabb_min[1] = px->m_aabbMin.y(); s.ray_aabb_min[2] = px->m_aabbMin.z();
        s.ray_aabb_max[0] = px->m_aabbMax.x(); s.ray_aabb_max[1] = px->m_aabbMax.y(); s.ray_aabb_max[2] = px->m_aabbMax.z();
    }
    s.id = car->id;
    s.team = (uint8_t)car->team;
    for (int i = 0; i < 4; i++) {
        auto& w = car->_bulletVehicle.m_wheelInfo[i];
        s.wheels[i].steer_angle = w.m_steerAngle;
        s.wheels[i].engine_force = w.m_engineForce;
        s.wheels[i].brake = w.m_brake;
        s.wheels[i].lat_friction = w.m_latFriction;
        s.wheels[i].long_friction = w.m_longFriction;
        s.wheels[i].extra_pushback = w.m_extraPushback;
    }
}

void ApplyArenaState(Arena* arena, const ggl_arena_state& st) {
    arena->tickCount = st.tick_count;
    BallState bs;
    bs.pos = V3(st.ball.pos); bs.vel = V3(st.ball.vel); bs.angVel = V3(st.ball.ang_vel);
    bs.hsInfo.yTargetDir = st.ball.hs_y_target_dir; bs.hsInfo.curTargetSpeed = st.ball.hs_cur_target_speed; bs.hsInfo.timeSinceHit = st.ball.hs_time_since_hit;
    arena->ball->SetState(bs);
    arena->ball->_internalState.updateCounter = st.ball.update_counter;
    for (int i = 0; i < st.num_cars; i++) {
        Car* car = CarById(arena, i + 1);
        if (car) {
            ApplyCarState(car, st.cars[i]);
            // the broadphase box the next tick's suspension rays are tested against (ggl_car_state::ray_aabb_*): the given one, or
            // - unknown / "nobody within reach" - the box of the transform just set
            btBroadphaseProxy* px = car->_rigidBody.getBroadphaseHandle();
            if (px) {
                const float* a = st.cars[i].ray_aabb_min; const float* b = st.cars[i].ray_aabb_max;
                bool zero = true;
                for (int k = 0; k < 3; k++) zero = zero && a[k] == 0.f && b[k] == 0.f;
                btVector3 mn(a[0], a[1], a[2]), mx(b[0], b[1], b[2]);
                if (zero || a[0] > b[0]) car->_rigidBody.getAabb(mn, mx);
                arena->_bulletWorld.getBroadphase()->setAabb(px, mn, mx, arena->_bulletWorld.getDispatcher());
            }
        }
    }
    for (int i = 0; i < GGL_NUM_PADS && i < (int)arena->_boostPads.size(); i++) {
        BoostPadState ps;
        ps.isActive = st.pads[i].is_active;
        ps.cooldown = st.pads[i].cooldown;
        ps.prevLockedCarID = st.pads[i].prev_locked_car_id;
        ps.curLockedCar = NULL;
        arena->_boostPads[i]->SetState(ps);
    }
}

void ReadArenaState(Arena* arena, ggl_arena_state& st) {
    memset(&st, 0, sizeof(st));
    st.tick_count = arena->tickCount;
    st.num_cars = (int)arena->_cars.size();
    BallState bs = arena->ball->GetState();
    Put3(st.ball.pos, bs.pos); Put3(st.ball.vel, bs.vel); Put3(st.ball.ang_vel, bs.angVel);
    st.ball.update_counter = bs.updateCounter;
    st.ball.hs_y_target_dir = bs.hsInfo.yTargetDir; st.ball.hs_cur_target_speed = bs.hsInfo.curTargetSpeed; st.ball.hs_time_since_hit = bs.hsInfo.timeSinceHit;
    for (int i = 0; i < st.num_cars && i < GGL_MAX_CARS; i++) {
        Car* car = CarById(arena, i + 1);
        if (car) ReadCarState(car, st.cars[i]);
    }
    for (int i = 0; i < GGL_NUM_PADS && i < (int)arena->_boostPads.size(); i++) {
        BoostPadState ps = arena->_boostPads[i]->GetState();
        st.pads[i].is_active = ps.isActive;
        st.pads[i].cooldown = ps.cooldown;
        st.pads[i].prev_locked_car_id = ps.prevLockedCarID;
    }
}

Reward* MakeReward(const ggl_reward_desc& d) {
    Reward* r = nullptr;
    switch (d.kind) {
    case GGL_REW_PLAYER_GOAL: r = new PlayerGoalReward(); break;
    case GGL_REW_ASSIST: r = new AssistReward(); break;
    case GGL_REW_SHOT: r = new ShotReward(); break;
    case GGL_REW_SHOT_PASS: r = new ShotPassReward(); break;
    case GGL_REW_SAVE: r = new SaveReward(); break;
    case GGL_REW_BUMP: r = new BumpReward(); break;
    case GGL_REW_BUMPED_PENALTY: r = new BumpedPenalty(); break;
    case GGL_REW_DEMO: r = new DemoReward(); break;
    case GGL_REW_DEMOED_PENALTY: r = new DemoedPenalty(); break;
    case GGL_REW_GOAL: r = new GoalReward(d.params[0]); break;
    case GGL_REW_VELOCITY: r = new VelocityReward(d.params[0] != 0); break;
    case GGL_REW_VELOCITY_BALL_TO_GOAL: r = new VelocityBallToGoalReward(d.params[0] != 0); break;
    case GGL_REW_VELOCITY_PLAYER_TO_BALL: r = new VelocityPlayerToBallReward(); break;
    case GGL_REW_FACE_BALL: r = new FaceBallReward(); break;
    case GGL_REW_TOUCH_BALL: r = new TouchBallReward(); break;
    case GGL_REW_SPEED: r = new SpeedReward(); break;
    case GGL_REW_WAVEDASH: r = new WavedashReward(); break;
    case GGL_REW_PICKUP_BOOST: r = new PickupBoostReward(); break;
    case GGL_REW_SAVE_BOOST: r = new SaveBoostReward(d.params[0]); break;
    case GGL_REW_AIR: r = new AirReward(); break;
    case GGL_REW_TOUCH_ACCEL: r = new TouchAccelReward(); break;
    case GGL_REW_STRONG_TOUCH: {
        auto* s = new StrongTouchReward();
        s->minRewardedVel = d.params[0]; s->maxRewardedVel = d.params[1];
        r = s; break;
    }
    case GGL_REW_KICKOFF_PROXIMITY_2V2: {
        auto* k = new KickoffProximityReward2v2Enhanced();
        k->goerReward = d.params[0]; k->cheaterReward = d.params[1]; k->dynamicWeight = d.params[2]; k->rotationPrepWeight = d.params[3];
        r = k; break;
    }
    default: return nullptr;
    }
    if (d.zero_sum) r = new ZeroSumReward(r, 0.5f);
    return r;
}

} // namespace

extern "C" {

// meshes: soccar .cmf blobs (CollisionMeshFile.cpp:11-35). Idempotent like RocketSim::InitFromMem.
int ref_init(const ggl_mesh_blob* meshes, int n) {
    try {
        std::map<GameMode, std::vector<FileData>> m;
        auto& v = m[GameMode::SOCCAR];
        for (int i = 0; i < n; i++) v.push_back(FileData(meshes[i].data, meshes[i].data + meshes[i].size));
        RocketSim::InitFromMem(m, true);
        return 0;
    } catch (std::exception& e) {
        return -1;
    }
}

// Replace the global pool (ThreadPool.h:9-16). n<=0 → hardware_concurrency.
void ref_set_threads(int n) {
    g_ThreadPool.WaitUntilDone();
    delete g_ThreadPool._tp;
    if (n <= 0) n = (int)std::thread::hardware_concurrency();
    g_ThreadPool._tp = new dp::thread_pool<>(n);
    g_ThreadPool._numThreads = g_ThreadPool._tp->size();
}
int ref_num_threads() { return g_ThreadPool._numThreads; }

// Seeds RocketSim's thread-local RNG (Math.cpp:59-64) on every pool worker and the calling thread.
// Deterministic only with ref_set_threads(1).
void ref_seed(uint64_t seed) {
    RocketSim::Math::GetRandEngine().seed(seed);
    int n = g_ThreadPool._numThreads;
    std::atomic<int> arrived{0};
    for (int i = 0; i < n; i++)
        g_ThreadPool.StartJobAsync([&arrived, seed, n]() {
            RocketSim::Math::GetRandEngine().seed(seed);
            arrived++;
            // hold the worker until every worker has been seeded once
            auto t0 = std::chrono::steady_clock::now();
            while (arrived.load() < n && std::chrono::steady_clock::now() - t0 < std::chrono::milliseconds(200))
                std::this_thread::yield();
        });
    g_ThreadPool.WaitUntilDone();
}

// Where RocketSim's thread-local engine (Math.cpp:59-64) stands: the NEXT value a copy of it would yield (the state itself is
// private; next = 16807 * state mod 2^31-1 identifies it). which = 0 the calling thread, 1 = the pool's worker (meaningful with
// ref_set_threads(1)). Setting it: ref_seed(state) - minstd_rand0::seed(s) installs s itself for 1 <= s <= 2147483646.
uint32_t ref_rng_peek(int which) {
    uint32_t out = 0;
    auto read = [&out]() { std::default_random_engine copy = RocketSim::Math::GetRandEngine(); out = (uint32_t)copy(); };
    if (which == 0) read();
    else { g_ThreadPool.StartJobAsync(read); g_ThreadPool.WaitUntilDone(); }
    return out;
}

void* ref_envset_create(const ggl_envset_desc* d) {
    try {
        auto* env = new RefEnv();
        env->desc = *d;
        env->resetStates.resize(d->num_arenas);
        env->hasResetState.assign(d->num_arenas, 0);
        EnvSetConfig cfg = {};
        cfg.numArenas = d->num_arenas;
        cfg.tickSkip = d->tick_skip;
        cfg.actionDelay = d->action_delay;
        cfg.saveRewards = true;
        cfg.shuffleRewardSampling = false;
        ggl_envset_desc desc = *d;
        cfg.envCreateFn = [desc, env](int index) {
            EnvCreateResult r = {};
            Arena* arena = Arena::Create(desc.game_mode == GGL_MODE_HEATSEEKER ? GameMode::HEATSEEKER : GameMode::SOCCAR);
            for (int i = 0; i < desc.players_per_team; i++) arena->AddCar(Team::BLUE);
            for (int i = 0; i < desc.players_per_team; i++) arena->AddCar(Team::ORANGE);
            r.arena = arena;
            for (int i = 0; i < desc.num_rewards; i++) {
                Reward* rw = MakeReward(desc.rewards[i]);
                if (!rw) throw std::runtime_error("bad reward kind");
                r.rewards.push_back(WeightedReward(rw, desc.rewards[i].weight));
            }
            for (int i = 0; i < desc.num_terminals; i++) {
                if (desc.terminals[i].kind == GGL_TERM_GOAL_SCORE) r.terminalConditions.push_back(new GoalScoreCondition());
                else r.terminalConditions.push_back(new NoTouchCondition(desc.terminals[i].params[0]));
            }
            bool customCoefs = false;
            for (int i = 0; i < 5; i++) customCoefs = customCoefs || desc.obs_coefs[i] != 0.f;
            const Vec pc(desc.obs_coefs[0], desc.obs_coefs[1], desc.obs_coefs[2]);
            if (desc.obs_kind == GGL_OBS_DEFAULT_PADDED)
                r.obsBuilder = customCoefs ? new DefaultObsPadded(desc.obs_max_players, pc, desc.obs_coefs[3], desc.obs_coefs[4]) : new DefaultObsPadded(desc.obs_max_players);
            else if (desc.obs_kind == GGL_OBS_ADVANCED) r.obsBuilder = new AdvancedObs();
            else r.obsBuilder = customCoefs ? new DefaultObs(pc, desc.obs_coefs[3], desc.obs_coefs[4]) : new DefaultObs();
            r.actionParser = new DefaultAction();
            auto makeSetter = [env, index](int kind, int flags) -> StateSetter* {
                if (kind == GGL_SETTER_INJECTED) {
                    auto* s = new InjectedStateSetter();
                    s->env = env; s->arenaIdx = index;
                    return s;
                }
                if (kind == GGL_SETTER_FUZZED_KICKOFF) return new FuzzedKickoffState();
                if (kind == GGL_SETTER_RANDOM)
                    return new RandomState((flags & GGL_RANDOM_BALL_SPEED) != 0, (flags & GGL_RANDOM_CAR_SPEED) != 0, (flags & GGL_RANDOM_CARS_ON_GROUND) != 0);
                return new KickoffState();
            };
            if (desc.state_setter == GGL_SETTER_COMBINED) {
                std::vector<std::pair<StateSetter*, float>> kids;
                for (int i = 0; i < desc.num_setter_children; i++)
                    kids.push_back({makeSetter(desc.setter_children[i].kind, desc.setter_children[i].flags), desc.setter_children[i].weight});
                r.stateSetter = new CombinedState(kids);
            } else {
                r.stateSetter = makeSetter(desc.state_setter, desc.state_setter_flags);
            }
            r.userInfo = nullptr;
            return r;
        };
        // EnvSet.cpp:52-89 pushes arenas in completion order while callbacks capture the creation index:
        // construct on a single worker so order == index.
        int prevThreads = g_ThreadPool._numThreads;
        if (prevThreads != 1) ref_set_threads(1);
        env->envSet = new EnvSet(cfg);
        if (prevThreads != 1) ref_set_threads(prevThreads);
        return env;
    } catch (std::exception& e) {
        fprintf(stderr, "ref_envset_create: %s\n", e.what());
        return nullptr;
    }
}

void ref_envset_destroy(void* p) {
    auto* env = (RefEnv*)p;
    if (!env) return;
    delete env->envSet;
    delete env;
}

void ref_envset_info(void* p, int* numPlayers, int* obsSize, int* numActions) {
    auto* e = ((RefEnv*)p)->envSet;
    *numPlayers = e->state.numPlayers; *obsSize = e->obsSize; *numActions = (int)e->state.actionMasks.size[1];
}

void ref_envset_step_first_half(void* p, int async) { ((RefEnv*)p)->envSet->StepFirstHalf(async != 0); }
void ref_envset_sync(void* p) { ((RefEnv*)p)->envSet->Sync(); }
void ref_envset_step_second_half(void* p, const int32_t* actions) {
    auto* e = ((RefEnv*)p)->envSet;
    IList a(actions, actions + e->state.numPlayers);
    e->StepSecondHalf(a, false);
}
void ref_envset_reset(void* p) { ((RefEnv*)p)->envSet->Reset(); }
void ref_envset_reset_arena(void* p, int arena) { ((RefEnv*)p)->envSet->ResetArena(arena); }

// Learner.cpp:669-793 order: Reset, StepFirstHalf(async), <inference>, Sync, StepSecondHalf.
void ref_envset_step(void* p, const int32_t* actions) {
    auto* e = ((RefEnv*)p)->envSet;
    e->Reset();
    e->StepFirstHalf(true);
    e->Sync();
    IList a(actions, actions + e->state.numPlayers);
    e->StepSecondHalf(a, false);
}

void ref_envset_copy_out(void* p, float* obs, uint8_t* masks, float* rewards, uint8_t* terminals) {
    auto* e = ((RefEnv*)p)->envSet;
    if (obs) memcpy(obs, e->state.obs.data.data(), e->state.obs.data.size() * sizeof(float));
    if (masks) memcpy(masks, e->state.actionMasks.data.data(), e->state.actionMasks.data.size());
    if (rewards) memcpy(rewards, e->state.rewards.data(), e->state.rewards.size() * sizeof(float));
    if (terminals) memcpy(terminals, e->state.terminals.data(), e->state.terminals.size());
}

// players[] slot order (unordered_set iteration)
void ref_envset_slot_info(void* p, int32_t* carIds, uint8_t* teams) {
    auto* e = ((RefEnv*)p)->envSet;
    int k = 0;
    for (size_t a = 0; a < e->arenas.size(); a++)
        for (auto& pl : e->state.gameStates[a].players) { carIds[k] = pl.carId; teams[k] = (uint8_t)pl.team; k++; }
}

void ref_envset_events(void* p, uint8_t* events) {
    auto* e = ((RefEnv*)p)->envSet;
    int k = 0;
    for (size_t a = 0; a < e->arenas.size(); a++)
        for (auto& pl : e->state.gameStates[a].players) {
            auto& ev = pl.eventState;
            uint8_t v[9] = { ev.goal, ev.save, ev.assist, ev.shot, ev.shotPass, ev.bump, ev.bumped, ev.demo, ev.demoed };
            memcpy(events + 9 * k, v, 9); k++;
        }
}

void ref_envset_last_rewards(void* p, float* out) {
    auto* env = (RefEnv*)p; auto* e = env->envSet;
    int nr = env->desc.num_rewards;
    for (size_t a = 0; a < e->arenas.size(); a++)
        for (int r = 0; r < nr; r++)
            out[a * nr + r] = (r < (int)e->state.lastRewards[a].size()) ? e->state.lastRewards[a][r] : 0.f;
}

void ref_envset_get_state(void* p, int arena, ggl_arena_state* st) { ReadArenaState(((RefEnv*)p)->envSet->arenas[arena], *st); }
void ref_envset_set_state(void* p, int arena, const ggl_arena_state* st) { ApplyArenaState(((RefEnv*)p)->envSet->arenas[arena], *st); }
void ref_envset_set_reset_state(void* p, int arena, const ggl_arena_state* st) {
    auto* env = (RefEnv*)p;
    env->resetStates[arena] = *st; env->hasResetState[arena] = 1;
}

// raw Arena::Step on one arena (arena<0: all, sequentially on the calling thread)
void ref_arena_step(void* p, int arena, int ticks) {
    auto* e = ((RefEnv*)p)->envSet;
    if (arena >= 0) e->arenas[arena]->Step(ticks);
    else for (auto* a : e->arenas) a->Step(ticks);
}

void ref_action_table(float* out, int* n) {
    DefaultAction da;
    *n = (int)da.actions.size();
    for (size_t i = 0; i < da.actions.size(); i++)
        for (int j = 0; j < 8; j++) out[i * 8 + j] = da.actions[i][j];
}

// Timed rollout with uniformly random mask-valid actions from std::mt19937(seed).
// Returns seconds for `steps` env steps after `warmup` untimed ones.
double ref_envset_bench(void* p, int steps, int warmup, uint32_t seed) {
    auto* e = ((RefEnv*)p)->envSet;
    std::mt19937 rng(seed);
    int np = e->state.numPlayers, na = (int)e->state.actionMasks.size[1];
    IList actions(np);
    auto pick = [&]() {
        const uint8_t* m = e->state.actionMasks.data.data();
        for (int i = 0; i < np; i++) {
            int cnt = 0;
            for (int a = 0; a < na; a++) cnt += m[i * na + a] != 0;
            int k = cnt ? (int)(rng() % cnt) : 0, sel = 0;
            for (int a = 0; a < na; a++) if (m[i * na + a]) { if (k-- == 0) { sel = a; break; } }
            actions[i] = sel;
        }
    };
    auto one = [&]() {
        e->Reset();
        e->StepFirstHalf(true);
        pick();
        e->Sync();
        e->StepSecondHalf(actions, false);
    };
    for (int i = 0; i < warmup; i++) one();
    auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) one();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}


// ---- debugging aids for the parity tests: contact manifolds of the last tick and shape constants ----
static int BodyCode(const btCollisionObject* o) {
    if (o->getUserIndex() == BT_USERINFO_TYPE_CAR) return (int)((Car*)o->getUserPointer())->id;
    if (o->getUserIndex() == BT_USERINFO_TYPE_BALL) return 0;
    return -1;
}
// out: [max][20] floats per contact point: bodyA, bodyB, normalOnB xyz, posA xyz, posB xyz, distance, friction,
// restitution, isSpecial, localA xyz, appliedImpulse.   Returns the number of points.
int ref_arena_contacts(void* p, int arena, float* out, int max) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[arena];
    auto* disp = a->_bulletWorld.getDispatcher();
    int n = 0;
    for (int i = 0; i < disp->getNumManifolds(); i++) {
        btPersistentManifold* m = disp->getManifoldByIndexInternal(i);
        for (int j = 0; j < m->getNumContacts() && n < max; j++) {
            const btManifoldPoint& cp = m->getContactPoint(j);
            float* o = out + 20 * n++;
            o[0] = (float)BodyCode(m->getBody0()); o[1] = (float)BodyCode(m->getBody1());
            o[2] = cp.m_normalWorldOnB.x(); o[3] = cp.m_normalWorldOnB.y(); o[4] = cp.m_normalWorldOnB.z();
            o[5] = cp.m_positionWorldOnA.x(); o[6] = cp.m_positionWorldOnA.y(); o[7] = cp.m_positionWorldOnA.z();
            o[8] = cp.m_positionWorldOnB.x(); o[9] = cp.m_positionWorldOnB.y(); o[10] = cp.m_positionWorldOnB.z();
            o[11] = cp.m_distance1; o[12] = cp.m_combinedFriction; o[13] = cp.m_combinedRestitution; o[14] = cp.m_isSpecial ? 1.f : 0.f;
            o[15] = cp.m_localPointA.x(); o[16] = cp.m_localPointA.y(); o[17] = cp.m_localPointA.z();
            o[18] = cp.m_appliedImpulse; o[19] = (float)i;
        }
    }
    return n;
}

// Every manifold of the dispatcher after the last tick, empty ones included, in array order: out[i] = {body0, body1, points}
int ref_arena_manifolds(void* p, int arena, int* out, int max) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[arena];
    auto* disp = a->_bulletWorld.getDispatcher();
    int n = 0;
    for (int i = 0; i < disp->getNumManifolds() && n < max; i++) {
        btPersistentManifold* m = disp->getManifoldByIndexInternal(i);
        auto code = [a](const btCollisionObject* o) -> int {
            const int c = BodyCode(o);
            if (c >= 0) return c;
            // a static body: -(100 + its index in the world's collision object array), planes 1000 further down
            const btCollisionObjectArray& arr = a->_bulletWorld.getCollisionObjectArray();
            int idx = 0;
            for (int q = 0; q < arr.size(); q++) if (arr[q] == o) { idx = q; break; }
            return -(100 + idx) - (o->getCollisionShape()->getShapeType() == STATIC_PLANE_PROXYTYPE ? 1000 : 0);
        };
        out[3 * n] = code(m->getBody0()); out[3 * n + 1] = code(m->getBody1()); out[3 * n + 2] = m->getNumContacts();
        n++;
    }
    return n;
}

// Broadphase boxes of the static bodies, in world-array order (the order of ref_arena_manifolds' static codes): out[i][6] = min
// xyz, max xyz in Bullet units. Returns their number.
int ref_arena_static_boxes(void* p, int arena, float* out, int max) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[arena];
    const btCollisionObjectArray& arr = a->_bulletWorld.getCollisionObjectArray();
    int n = 0;
    for (int q = 0; q < arr.size() && n < max; q++) {
        if (!arr[q]->isStaticObject()) continue;
        btBroadphaseProxy* px = arr[q]->getBroadphaseHandle();
        float* o = out + 6 * n++;
        o[0] = px->m_aabbMin.x(); o[1] = px->m_aabbMin.y(); o[2] = px->m_aabbMin.z();
        o[3] = px->m_aabbMax.x(); o[4] = px->m_aabbMax.y(); o[5] = px->m_aabbMax.z();
    }
    return n;
}

// btAlignedObjectArray::quickSort itself, on (key, index) pairs compared by key only like btPersistentManifoldSortPredicate
// compares island ids: idx[] comes back as the permutation the reference's sort applies to n manifolds with these keys.
struct KeyIdx { int key, idx; };
struct KeyLess { bool operator()(const KeyIdx& a, const KeyIdx& b) const { return a.key < b.key; } };
void ref_quicksort_keys(const int* keys, int* idx, int n) {
    btAlignedObjectArray<KeyIdx> arr;
    for (int i = 0; i < n; i++) { KeyIdx k; k.key = keys[i]; k.idx = i; arr.push_back(k); }
    arr.quickSort(KeyLess());
    for (int i = 0; i < n; i++) idx[i] = arr[i].idx;
}

// Broadphase box of the ball after the last tick (min xyz, max xyz, Bullet units)
void ref_arena_ball_box(void* p, int arena, float* o) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[arena];
    btBroadphaseProxy* px = a->ball->_rigidBody.getBroadphaseHandle();
    o[0] = px->m_aabbMin.x(); o[1] = px->m_aabbMin.y(); o[2] = px->m_aabbMin.z();
    o[3] = px->m_aabbMax.x(); o[4] = px->m_aabbMax.y(); o[5] = px->m_aabbMax.z();
}

// wheels of car `carId` after the last tick: out[4][12] = contact normal xyz, contact point xyz, suspension length,
// ground kind (-1 none, 0 static world, 1 ball, 2 + carId-1 car), impulse xyz, suspension relative velocity
void ref_arena_wheel_info(void* p, int arena, int carId, float* out) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[arena];
    Car* car = CarById(a, carId);
    for (int i = 0; i < 4; i++) {
        auto& w = car->_bulletVehicle.m_wheelInfo[i];
        float* o = out + 12 * i;
        o[0] = w.m_raycastInfo.m_contactNormalWS.x(); o[1] = w.m_raycastInfo.m_contactNormalWS.y(); o[2] = w.m_raycastInfo.m_contactNormalWS.z();
        o[3] = w.m_raycastInfo.m_contactPointWS.x(); o[4] = w.m_raycastInfo.m_contactPointWS.y(); o[5] = w.m_raycastInfo.m_contactPointWS.z();
        o[6] = w.m_raycastInfo.m_suspensionLength;
        float kind = -1;
        const btCollisionObject* g = (const btCollisionObject*)w.m_raycastInfo.m_groundObject;
        if (g) {
            if (g == &a->ball->_rigidBody) kind = 1;
            else if (g->isStaticObject()) kind = 0;
            else for (Car* c : a->_cars) if (g == &c->_rigidBody) kind = 2 + (float)(c->id - 1);
        }
        o[7] = kind;
        o[8] = w.m_impulse.x(); o[9] = w.m_impulse.y(); o[10] = w.m_impulse.z(); o[11] = w.m_suspensionRelativeVelocity;
    }
}

// The convex ray cast a suspension ray runs against a car's hitbox child (kind 0) or the ball's sphere (kind 1), with the
// reference's own shapes: btCollisionWorld::rayTestSingle on a temporary transform (pos, row-major basis). out = fraction,
// normal xyz (as ClosestRayResultCallback stores it); returns 1 on a hit.
int ref_convex_ray_cast(void* p, int kind, const float* from, const float* to, const float* pos, const float* basis, float* out) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[0];
    Car* car = *a->_cars.begin();
    btTransform tf;
    tf.setOrigin(btVector3(pos[0], pos[1], pos[2]));
    tf.setBasis(btMatrix3x3(basis[0], basis[1], basis[2], basis[3], basis[4], basis[5], basis[6], basis[7], basis[8]));
    btTransform rf, rt; rf.setIdentity(); rt.setIdentity();
    btVector3 f(from[0], from[1], from[2]), t(to[0], to[1], to[2]);
    rf.setOrigin(f); rt.setOrigin(t);
    btCollisionWorld::ClosestRayResultCallback cb(f, t, nullptr);
    btCollisionObject* obj = kind == 0 ? (btCollisionObject*)&car->_rigidBody : (btCollisionObject*)&a->ball->_rigidBody;
    btCollisionShape* shape = kind == 0 ? (btCollisionShape*)&car->_compoundShape : (btCollisionShape*)a->ball->_collisionShape;
    btCollisionWorld::rayTestSingle(rf, rt, obj, shape, tf, cb);
    if (!cb.hasHit()) return 0;
    out[0] = cb.m_closestHitFraction; out[1] = cb.m_hitNormalWorld.x(); out[2] = cb.m_hitNormalWorld.y(); out[3] = cb.m_hitNormalWorld.z();
    return 1;
}

// Triangle indices of mesh `mesh` in the order its quantized BVH stores the leaves (= the order queries report them).
// Returns the triangle count; out may be NULL.
int ref_mesh_leaf_order(int mesh, int* out, int max) {
    auto shapes = RocketSim::GetArenaCollisionShapes(GameMode::SOCCAR);
    if (mesh < 0 || mesh >= (int)shapes.size()) return -1;
    btOptimizedBvh* bvh = shapes[mesh]->getOptimizedBvh();
    const QuantizedNodeArray& nodes = bvh->getQuantizedNodeArray();
    // the array is sized 2 n; a tree of n leaves uses the first 2 n - 1 entries
    int n = 0;
    for (int i = 0; i < nodes.size() - 1; i++)
        if (nodes[i].isLeafNode()) { if (out && n < max) out[n] = nodes[i].getTriangleIndex(); n++; }
    return n;
}

// out[0..2] car invInertiaDiagLocal, [3] car invMass, [4..6] ball invInertiaDiagLocal, [7] ball invMass, [8] ball shape radius,
// [9] ball shape margin, [10] car compound contact-breaking threshold, [11] ball contact-breaking threshold,
// [12] box margin, [13..15] box half extents with margin, [16] solver erp2, [17] solver erp, [18] linearSlop, [19] restitution vel threshold
void ref_debug_constants(void* p, float* out) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[0];
    Car* car = *a->_cars.begin();
    btVector3 ci = car->_rigidBody.getInvInertiaDiagLocal();
    out[0] = ci.x(); out[1] = ci.y(); out[2] = ci.z(); out[3] = car->_rigidBody.getInvMass();
    btVector3 bi = a->ball->_rigidBody.getInvInertiaDiagLocal();
    out[4] = bi.x(); out[5] = bi.y(); out[6] = bi.z(); out[7] = a->ball->_rigidBody.getInvMass();
    out[8] = a->ball->GetRadiusBullet(); out[9] = a->ball->_collisionShape->getMargin();
    out[10] = car->_compoundShape.getContactBreakingThreshold(gContactBreakingThreshold);
    out[11] = a->ball->_collisionShape->getContactBreakingThreshold(gContactBreakingThreshold);
    out[12] = car->_childHitboxShape.getMargin();
    btVector3 he = car->_childHitboxShape.getHalfExtentsWithMargin();
    out[13] = he.x(); out[14] = he.y(); out[15] = he.z();
    auto& si = a->_bulletWorld.getSolverInfo();
    out[16] = si.m_erp2; out[17] = si.m_erp; out[18] = si.m_linearSlop; out[19] = si.m_restitutionVelocityThreshold;
    out[20] = (float)si.m_numIterations; out[21] = (float)si.m_solverMode; out[22] = si.m_splitImpulsePenetrationThreshold;
    out[23] = si.m_splitImpulseTurnErp; out[24] = (float)a->_bulletWorld.getDispatchInfo().m_useContinuous;
    out[25] = a->_worldCollisionRBAmount > 0 ? a->_worldCollisionBvhShapes[0].getMargin() : -1.f;
    out[26] = a->_worldCollisionRBAmount > 0 ? a->_worldCollisionRBs[0].getFriction() : -1.f;
    out[27] = (float)a->_worldCollisionRBAmount;
}


// only the cars' `controls` of a host state (leaves everything else untouched)
void ref_envset_set_controls(void* p, int arena, const ggl_arena_state* st) {
    Arena* a = ((RefEnv*)p)->envSet->arenas[arena];
    for (int i = 0; i < st->num_cars; i++) {
        Car* car = CarById(a, i + 1);
        if (car) car->controls = ToControls(st->cars[i].controls);
    }
}

} // extern "C"
