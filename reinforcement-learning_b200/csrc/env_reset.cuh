// Third part of the env layer: the step phases of EnvSet::StepFirstHalf / StepSecondHalf / ResetArena as functions a
// lane group can split (leader-only bookkeeping, per-player outputs). See env_logic.cuh for reference citations.
#pragma once
#include "env_step.cuh"

namespace ggl {

// ---- StepFirstHalf bookkeeping (EnvSet.cpp:115-123): prev snapshot + clear event flags
GGL_HD void env_begin_step(ArenaDev& A) {
    A.stepCost = 0;
    A.prevBallVel = A.ball.vel * kBT2UU;
    for (int s = 0; s < A.numCars; s++) {
        const CarDev& c = A.cars[A.slotCar[s]];
        EnvPlayer& ep = A.players[s];
        ep.prevBoost = c.boost; ep.prevIsFlipping = c.isFlipping; ep.prevIsOnGround = c.isOnGround;
        for (int e = 0; e < 9; e++) ep.events[e] = 0;
    }
}

// ---- StepSecondHalf: parse actions, set controls (EnvSet.cpp:146-153, Action.h:34-46)
GGL_HD void env_apply_actions(ArenaDev& A, const int32_t* actions /* this arena's slots */, const float* table) {
    for (int s = 0; s < A.numCars; s++) {
        int idx = actions[s];
        if (idx < 0) idx = 0;
        if (idx >= kNumActions) idx = kNumActions - 1;
        const float* a = table + 8 * idx;
        Controls k;
        k.throttle = a[0]; k.steer = a[1]; k.pitch = a[2]; k.yaw = a[3]; k.roll = a[4];
        k.jump = a[5] == 1.f; k.boost = a[6] == 1.f; k.handbrake = a[7] == 1.f; k._pad = 0;
        A.cars[A.slotCar[s]].controls = k;
        for (int i = 0; i < 8; i++) A.players[s].prevAction[i] = a[i];
    }
}

// ---- after the last tick: event tracker, GameState update, terminal conditions (EnvSet.cpp:158-181)
GGL_HD void env_after_ticks(ArenaDev& A, const EnvCfg& cfg) {
    event_tracker_update(A);
    uint64_t dt = A.tickCount - A.gsLastTickCount;
    int tickSkip = (int)dt;
    A.stepDeltaTime = tickSkip * (1.0f / 120.0f);
    bool anyTouch = false;
    for (int s = 0; s < A.numCars; s++) {
        const CarDev& c = A.cars[A.slotCar[s]];
        bool touched = c.hitValid && c.hitTick >= A.gsLastTickCount;
        if (touched) { A.lastTouchCarId = (int)c.id; anyTouch = true; }
    }
    A.stepGoalScored = ball_scored(A) ? 1 : 0;
    uint8_t term = 0;
    for (int t = 0; t < cfg.numTerminals; t++) {
        bool fired = false, trunc = false;
        if (cfg.terminals[t].kind == GGL_TERM_GOAL_SCORE) {
            fired = A.stepGoalScored != 0;
        } else {
            trunc = true;
            if (anyTouch) { A.noTouchTimer[t] = 0.f; fired = false; }
            else { A.noTouchTimer[t] += A.stepDeltaTime; fired = A.noTouchTimer[t] >= cfg.terminals[t].params[0]; }
        }
        if (fired) {
            uint8_t cur = trunc ? 2 : 1;
            if (term == 0) term = cur;
            else if (cur == 1) term = cur;
        }
    }
    A.terminal = term;
}

// ---- per player: weighted reward sum, obs row, mask row (EnvSet.cpp:183-268)
GGL_HD float env_player_reward(const ArenaDev& A, const EnvCfg& cfg, int slot, float* rawOut /* [numRewards] or null */) {
    const PlayerView p = player_view(A, A.slotCar[slot]);
    V3 bp = A.ball.pos * kBT2UU, bv = A.ball.vel * kBT2UU;
    float total = 0.f;
    for (int r = 0; r < cfg.numRewards; r++) {
        float v = reward_value(A, cfg.rewards[r], p, slot, A.stepGoalScored != 0, bp, bv);
        if (rawOut) rawOut[r] = v;
        total += v * cfg.rewards[r].weight;
    }
    return total;
}

GGL_HD void env_end_step(ArenaDev& A) { A.gsLastTickCount = A.tickCount; }

// ---- reset -------------------------------------------------------------------------------------------------
GGL_HD void car_set_spawn(CarDev& c, V3 posUU, float yaw) {
    car_reset_state_defaults(c);
    c.pos = posUU * kUU2BT;
    c.basis = m3_from_euler_ypr(yaw, -0.f, -0.f);
    car_update_inertia(c);
    c.aabbMin = c.aabbMax = v3(0, 0, 0);
}

// Arena::ResetToRandomKickoff (Arena.cpp:112-216): std::shuffle of the five spawn slots on the arena's stream (the draws
// libstdc++ makes, std_shuffle_nib), the i-th car of each team (in _cars = slot order) on slot order[i], orange mirrored.
GGL_HD void arena_reset_kickoff(ArenaDev& A) {
    const float sx[5] = {-2048, 2048, -256, 256, 0}, sy[5] = {-2560, -2560, -3840, -3840, -4608};
    const float q = 0.78539816339744830962f;
    const float syaw[5] = {q * 1, q * 3, q * 2, q * 2, q * 2};
    // heatseeker: four slots on one line (RLConst.h:306-311), only the first four entries of the order are shuffled and an
    // entry is clamped to the last slot (Arena.cpp:132-136, 159)
    const float hx[4] = {-1000, 1000, -2000, 2000};
    const bool heat = A.gameMode == kModeHeatseeker;
    const uint32_t order = std_shuffle_nib(A.rng, heat ? 4 : 5);
    int nb = 0, no = 0;
    for (int s = 0; s < A.numCars; s++) {
        CarDev& c = A.cars[A.slotCar[s]];
        int i = (c.team == 0) ? nb++ : no++;
        int k = (int)nib_get(order, i < 5 ? i : 4);
        if (heat && k > 3) k = 3;
        V3 pos = heat ? v3(hx[k], -4620.f, rl::CAR_SPAWN_REST_Z) : v3(sx[k], sy[k], rl::CAR_SPAWN_REST_Z);
        float yaw = heat ? q * 2 : syaw[k];
        if (c.team != 0) { pos = pos * v3(-1, -1, 1); yaw = (float)((double)yaw + 3.14159265358979323846); }
        car_set_spawn(c, pos, yaw);
    }
    BallDev& b = A.ball;
    b.pos = v3(0, 0, rl::BALL_REST_Z) * kUU2BT; b.vel = v3(0, 0, 0); b.angVel = v3(0, 0, 0);
    b.velImpulseCache = v3(0, 0, 0); b.force = v3(0, 0, 0); b.torque = v3(0, 0, 0); b.updateCounter = 0;
    b.hsTargetDir = 0.f; b.hsTargetSpeed = rl::hs::INITIAL_TARGET_SPEED; b.hsTimeSinceHit = 0.f;        // BallState() defaults (Ball.h:23-32)
    if (heat) {   // one more raw draw decides the side the ball starts on (Arena.cpp:196-200)
        const float sgn = (minstd_next(A.rng) % 2u) ? 1.f : -1.f;
        b.pos = (v3(-1000.f, -2220.f, 92.75f) * v3(1, sgn, 1)) * kUU2BT;
        b.vel = (v3(0.f, -65.f, 650.f) * v3(1, sgn, 1)) * kUU2BT;
    }
    for (int p = 0; p < kNumPads; p++) { A.pads[p].isActive = 1; A.pads[p].cooldown = 0.f; A.pads[p].prevLockedCarId = 0; A.pads[p].curLockedCarId = 0; }
}

// FuzzedKickoffState::ResetArena (FuzzedKickoffState.h:17-26): kickoff, then per car in _cars order GetState (position back in
// uu), += RandFloat(-0.1, 0.1) per axis x, y, z, SetState (Car.cpp:10-36: to Bullet units again, everything else as spawned)
GGL_HD void arena_reset_fuzzed_kickoff(ArenaDev& A) {
    arena_reset_kickoff(A);
    for (int s = 0; s < A.numCars; s++) {
        CarDev& c = A.cars[A.slotCar[s]];
        V3 p = c.pos * kBT2UU;
        p.x += rand_float(A.rng, -0.1f, 0.1f);
        p.y += rand_float(A.rng, -0.1f, 0.1f);
        p.z += rand_float(A.rng, -0.1f, 0.1f);
        c.pos = p * kUU2BT;
    }
}

// RLGC::Math::RandVec (RLGymCPP/Math.cpp:3-9): `Vec(RandFloat(x), RandFloat(y), RandFloat(z))`. The three calls are constructor
// arguments, whose evaluation order the language leaves open; the reference as built here (g++, x86-64) evaluates them last to
// first, so z is drawn first. The same holds for Angle(yaw, pitch, roll) and for the operands of `vec * RandFloat()` below
// (all pinned by tests/test_setters_cpu.py against the compiled reference).
GGL_HD V3 rand_vec(uint32_t& s, V3 lo, V3 hi) {
    V3 r;
    r.z = rand_float(s, lo.z, hi.z);
    r.y = rand_float(s, lo.y, hi.y);
    r.x = rand_float(s, lo.x, hi.x);
    return r;
}
GGL_HD V3 rand_norm_vec(uint32_t& s) { return rs_normalized(rand_vec(s, v3(-1, -1, -1), v3(1, 1, 1))); }

// RandomState::ResetArena (RandomState.cpp:11-62), draw for draw
GGL_HD void arena_reset_random(ArenaDev& A, int flags) {
    arena_reset_kickoff(A);
    const float X_MAX = 3500, Y_MAX = 4000, Z_MAX = 1820, CAR_Z_MIN = 150;
    const float PITCH_MAX = (float)(3.14159265358979323846 / 2), YAW_MAX = (float)3.14159265358979323846, ROLL_MAX = YAW_MAX, ANGVEL_MAX = 5.5f;
    uint32_t& g = A.rng;
    {
        BallDev& b = A.ball;
        V3 pos = rand_vec(g, v3(-X_MAX, -Y_MAX, 92.75f), v3(X_MAX, Y_MAX, Z_MAX));     // CommonValues::BALL_RADIUS
        V3 vel = v3(0, 0, 0), av = v3(0, 0, 0);
        if (flags & GGL_RANDOM_BALL_SPEED) {
            const float speed = rand_float(g, 0.f, 4000.f);       // `RandNormVec() * RandFloat(..)`: the right operand is drawn first
            vel = rand_norm_vec(g) * speed;
            av = rand_vec(g, v3(-4, -4, -4), v3(4, 4, 4));
        }
        b.pos = pos * kUU2BT; b.vel = vel * kUU2BT; b.angVel = av;
    }
    for (int s = 0; s < A.numCars; s++) {
        CarDev& c = A.cars[A.slotCar[s]];
        V3 pos = rand_vec(g, v3(-X_MAX, -Y_MAX, CAR_Z_MIN), v3(X_MAX, Y_MAX, Z_MAX));
        V3 vel = v3(0, 0, 0), av = v3(0, 0, 0);
        if (flags & GGL_RANDOM_CAR_SPEED) {
            (void)rand_vec(g, v3(-1, -1, -1), v3(1, 1, 1));              // randVelDir: drawn, never used (:41)
            const float speed = rand_float(g, 0.f, rl::CAR_MAX_SPEED);
            vel = rand_norm_vec(g) * speed;
            av = rand_norm_vec(g) * ANGVEL_MAX;
        }
        float roll = rand_float(g, -ROLL_MAX, ROLL_MAX);
        float pitch = rand_float(g, -PITCH_MAX, PITCH_MAX);
        const float yaw = rand_float(g, -YAW_MAX, YAW_MAX);
        const bool onGround = (flags & GGL_RANDOM_CARS_ON_GROUND) ? true : (rand_float(g, 0.f, 1.f) > 0.5f);
        if (onGround) { pos.z = 17.f; pitch = 0.f; roll = 0.f; vel.z = 0.f; av = v3(0, 0, 0); }
        const float boost = rand_float(g, 0.f, 100.f);
        car_reset_state_defaults(c);
        c.pos = pos * kUU2BT;
        c.basis = m3_from_euler_ypr(yaw, -pitch, -roll);              // Angle::ToRotMat (MathTypes.cpp:73-78)
        c.vel = vel * kUU2BT; c.angVel = av; c.boost = boost;
        car_update_inertia(c);
        c.aabbMin = c.aabbMax = v3(0, 0, 0);
    }
}

GGL_HD void arena_run_setter(ArenaDev& A, int kind, int flags) {
    if (kind == GGL_SETTER_FUZZED_KICKOFF) arena_reset_fuzzed_kickoff(A);
    else if (kind == GGL_SETTER_RANDOM) arena_reset_random(A, flags);
    else arena_reset_kickoff(A);
}

// The StateSetter call of EnvSet::ResetArena (EnvSet.cpp:276). `injected` = the host-supplied reset state of this arena or
// null. CombinedState (CombinedState.h:34-45): f = RandFloat(0, totalWeight), first child with f <= cumulative weight.
GGL_HD void arena_state_set(ArenaDev& A, const EnvCfg& cfg, const ArenaDev* injected) {
    int kind = cfg.stateSetter, flags = cfg.setterFlags;
    if (kind == GGL_SETTER_COMBINED) {
        const float f = rand_float(A.rng, 0.f, cfg.setterTotalWeight);
        int pick = cfg.numSetterChildren - 1;
        for (int i = 0; i < cfg.numSetterChildren; i++)
            if (f <= cfg.setterChildren[i].cumWeight) { pick = i; break; }
        kind = cfg.setterChildren[pick].kind; flags = cfg.setterChildren[pick].flags;
    }
    if (kind == GGL_SETTER_INJECTED && injected && injected->hasResetState) {
        const ArenaDev& R = *injected;
        A.tickCount = R.tickCount; A.ball = R.ball;
        for (int c = 0; c < A.numCars; c++) A.cars[c] = R.cars[c];
        for (int p = 0; p < kNumPads; p++) A.pads[p] = R.pads[p];
    } else {
        arena_run_setter(A, kind, flags);
    }
}

// EnvSet::ResetArena bookkeeping after the StateSetter ran (EnvSet.cpp:276-303)
GGL_HD void env_after_state_set(ArenaDev& A) {
    A.ballScoredLast = 0; A.ballShot = 0; A.shotCooldown = 0.f;        // GameEventTracker::ResetPersistentInfo
    A.lastTouchCarId = -1;
    for (int s = 0; s < A.numCars; s++) {
        const CarDev& c = A.cars[A.slotCar[s]];
        if (c.hitValid) A.lastTouchCarId = (int)c.id;                   // fresh GameState: every valid hit counts as "this step"
        EnvPlayer& ep = A.players[s];
        for (int i = 0; i < 8; i++) ep.prevAction[i] = 0.f;
        for (int e = 0; e < 9; e++) ep.events[e] = 0;
        ep.prevBoost = c.boost; ep.prevIsFlipping = c.isFlipping; ep.prevIsOnGround = c.isOnGround;
    }
    A.gsLastTickCount = A.tickCount;
    for (int t = 0; t < kMaxTerminals; t++) A.noTouchTimer[t] = 0.f;
    A.prevBallVel = A.ball.vel * kBT2UU;
    A.stepGoalScored = ball_scored(A) ? 1 : 0;
    A.stepDeltaTime = 0.f;
    A.terminal = 0;
}

}  // namespace ggl
