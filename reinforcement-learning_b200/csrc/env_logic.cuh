// Env layer on top of the arena tick: action table + masks, GameState bookkeeping, game events, terminal
// conditions, weighted rewards, observation rows and the kickoff reset — the per-arena body of RLGC::EnvSet.
//
// Reference behaviour restated (paths under GigaLearnCPP/RLGymCPP/src/RLGymCPP unless noted):
//   EnvSet/EnvSet.cpp:113-130      StepFirstHalf          -> env_first_half
//   EnvSet/EnvSet.cpp:132-273      StepSecondHalf         -> env_second_half
//   EnvSet/EnvSet.cpp:275-304      ResetArena             -> env_reset_arena
//   EnvSet/EnvSet.cpp:4-42         event callbacks -> PlayerEventState
//   ActionParsers/DefaultAction.cpp:3-118   action table, masks
//   Gamestates/GameState.cpp:53-131, Player.cpp:8-26
//   ObsBuilders/DefaultObs.cpp:4-54, AdvancedObs.cpp:108-270 (+ AdvancedObs.h:10-13 class coefficients)
//   Rewards/CommonRewards.h:7-203, TerminalConditions/*.h
//   RocketSim/src/Sim/GameEventTracker/GameEventTracker.cpp:5-165, Sim/Arena/Arena.cpp:112-216,827-975
#pragma once
#include "../../include/gigaenv.h"
#include "sim_tick.cuh"

namespace ggl {
static_assert(kModeSoccar == GGL_MODE_SOCCAR && kModeHeatseeker == GGL_MODE_HEATSEEKER, "game mode codes");


struct RewardCfg {
    int kind;
    float weight;
    float params[4];
};
struct TerminalCfg {
    int kind;
    float params[2];
};
struct ActionBits {         // DefaultAction's ground / air / boost / jump action sets as bit sets over the 90 actions
    uint32_t ground[3], air[3], boost[3], jump[3];
};
struct SetterCfg {
    int kind, flags;
    float cumWeight;        // CombinedState::cumulativeWeights (CombinedState.h:29-31)
};
struct EnvCfg {
    int numArenas, playersPerTeam, numCars, tickSkip, actionDelay, obsKind, obsSize;
    int numRewards, numTerminals, stateSetter;
    int obsMaxPlayers, setterFlags, numSetterChildren, gameMode;
    float setterTotalWeight;
    float obsCoef[5];           // DefaultObs / DefaultObsPadded: posCoef x y z, velCoef, angVelCoef
    SetterCfg setterChildren[GGL_MAX_SETTER_CHILDREN];
    ActionBits actionBits;
    RewardCfg rewards[kMaxRewards];
    TerminalCfg terminals[kMaxTerminals];
};

// floats per obs row (DefaultObs.cpp:4-54, AdvancedObs.cpp:193-270, DefaultObsPadded.cpp:4-68)
inline int env_obs_size(int kind, int cars, int maxPlayers) {
    if (kind == GGL_OBS_DEFAULT_PADDED) return 51 + 19 * 2 * maxPlayers;
    return 51 + (kind == GGL_OBS_ADVANCED ? 29 : 19) * cars;
}
// obs / state-setter part of the descriptor -> EnvCfg (shared by the product and the host compile of the parity tests);
// returns a message for a descriptor with no device implementation, nullptr when fine
inline const char* env_cfg_plugins(EnvCfg& c, const ggl_envset_desc* d) {
    c.obsKind = d->obs_kind; c.obsMaxPlayers = 0;
    if (d->obs_kind == GGL_OBS_DEFAULT_PADDED) {
        if (d->obs_max_players < d->players_per_team || d->obs_max_players > GGL_MAX_OBS_PLAYERS)
            return "DefaultObsPadded: obs_max_players must be players_per_team..GGL_MAX_OBS_PLAYERS (DefaultObsPadded.cpp:40-44)";
        c.obsMaxPlayers = d->obs_max_players;
    } else if (d->obs_kind != GGL_OBS_DEFAULT && d->obs_kind != GGL_OBS_ADVANCED) {
        return "unsupported obs kind";
    }
    c.obsSize = env_obs_size(d->obs_kind, 2 * d->players_per_team, c.obsMaxPlayers);
    {
        bool any = false;
        for (int i = 0; i < 5; i++) any = any || d->obs_coefs[i] != 0.f;
        const float dflt[5] = {1 / 4096.f, 1 / 5120.f, 1 / 2044.f, 1 / 2300.f, 1 / 5.5f};      // CommonValues.h: SIDE_WALL_X, BACK_WALL_Y, CEILING_Z, CAR_MAX_SPEED, CAR_MAX_ANG_VEL
        for (int i = 0; i < 5; i++) c.obsCoef[i] = any ? d->obs_coefs[i] : dflt[i];
        if (any && d->obs_kind == GGL_OBS_ADVANCED) return "AdvancedObs has no coefficient arguments (AdvancedObs.h:10-13)";
    }
    if (d->game_mode != GGL_MODE_SOCCAR && d->game_mode != GGL_MODE_HEATSEEKER) return "unsupported game mode (soccar and heatseeker have a device integrator)";
    c.gameMode = d->game_mode;
    c.stateSetter = d->state_setter; c.setterFlags = d->state_setter_flags; c.numSetterChildren = 0; c.setterTotalWeight = 0.f;
    if (d->state_setter < GGL_SETTER_KICKOFF || d->state_setter > GGL_SETTER_COMBINED) return "unsupported state setter kind";
    if (d->state_setter == GGL_SETTER_COMBINED) {
        if (d->num_setter_children < 1 || d->num_setter_children > GGL_MAX_SETTER_CHILDREN)
            return "CombinedState: 1..GGL_MAX_SETTER_CHILDREN children";
        c.numSetterChildren = d->num_setter_children;
        for (int i = 0; i < d->num_setter_children; i++) {
            const ggl_setter_child& ch = d->setter_children[i];
            if (ch.kind < GGL_SETTER_KICKOFF || ch.kind >= GGL_SETTER_COMBINED) return "CombinedState: unsupported child kind";
            if (!(ch.weight >= 0.f) || ch.weight > 3.0e38f) return "CombinedState: negative or invalid weight (CombinedState.h:25-27)";
            c.setterTotalWeight += ch.weight;
            c.setterChildren[i].kind = ch.kind; c.setterChildren[i].flags = ch.flags; c.setterChildren[i].cumWeight = c.setterTotalWeight;
        }
    }
    return nullptr;
}

// ---------------------------------------------------------------------------------------------- DefaultAction
// 24 ground actions then 66 air actions; index -> (throttle, steer, pitch, yaw, roll, jump, boost, handbrake)
GGL_HD void action_from_index(int idx, float a[8]) {
    int k = 0;
    const float RF[3] = {-1.f, 0.f, 1.f};
    for (int t = 0; t < 3; t++)
        for (int s = 0; s < 3; s++)
            for (int b = 0; b < 2; b++)
                for (int h = 0; h < 2; h++) {
                    if (b == 1 && RF[t] != 1.f) continue;
                    if (k == idx) {
                        a[0] = RF[t]; a[1] = RF[s]; a[2] = 0.f; a[3] = RF[s]; a[4] = 0.f; a[5] = 0.f; a[6] = (float)b; a[7] = (float)h;
                        return;
                    }
                    k++;
                }
    for (int p = 0; p < 3; p++)
        for (int y = 0; y < 3; y++)
            for (int r = 0; r < 3; r++)
                for (int j = 0; j < 2; j++)
                    for (int b = 0; b < 2; b++) {
                        if (j == 1 && RF[y] != 0.f) continue;
                        if (RF[p] == RF[r] && RF[r] == (float)j && j == 0) continue;
                        if (k == idx) {
                            float hb = (j == 1) && (RF[p] != 0.f || RF[y] != 0.f || RF[r] != 0.f) ? 1.f : 0.f;
                            a[0] = (float)b; a[1] = RF[y]; a[2] = RF[p]; a[3] = RF[y]; a[4] = RF[r]; a[5] = (float)j; a[6] = (float)b; a[7] = hb;
                            return;
                        }
                        k++;
                    }
    for (int i = 0; i < 8; i++) a[i] = 0.f;
}

constexpr int kNumGroundActions = 24;

GGL_HD bool car_has_flip_or_jump(const CarDev& c) {
    return c.isOnGround || (!c.hasFlipped && !c.hasDoubleJumped && c.airTimeSinceJump < rl::DOUBLEJUMP_MAX_DELAY);
}

// DefaultAction::GetActionMask with the masks the constructor derives from the table. The four sets the
// constructor builds (ground, air, boost, jump actions) are 90-bit constants of the table: computed once (action_bits_from_table,
// host side, at create) instead of per player and step from 450 table loads.
GGL_HD void action_bits_from_table(const float* table /* [90*8] */, ActionBits& B) {
    for (int w = 0; w < 3; w++) B.ground[w] = B.air[w] = B.boost[w] = B.jump[w] = 0u;
    for (int i = 0; i < kNumActions; i++) {
        const float* a = table + 8 * i;
        const bool jump = a[5] != 0.f, boost = a[6] != 0.f;
        const bool ground = i < kNumGroundActions;
        bool air = (i > kNumGroundActions && !jump);
        if (i < kNumGroundActions && a[0] == a[6] && ((a[3] != 0.f) == (a[7] != 0.f))) air = true;
        const uint32_t bit = 1u << (i & 31);
        if (ground) B.ground[i >> 5] |= bit;
        if (air) B.air[i >> 5] |= bit;
        if (boost) B.boost[i >> 5] |= bit;
        if (jump) B.jump[i >> 5] |= bit;
    }
}
GGL_HD void action_mask_row(const CarDev& c, const ActionBits& B, uint8_t* out /* 2-byte aligned: rows are 90 bytes */) {
    const bool onGround = c.isOnGround != 0;
    const bool noBoost = c.boost == 0.f;
    const bool turtled = c.worldContactHas && c.worldContactNormal.z > 0.9f;
    const bool jumpOk = car_has_flip_or_jump(c) || turtled;
    uint16_t* o = reinterpret_cast<uint16_t*>(out);
    for (int w = 0; w < 3; w++) {
        uint32_t m = onGround ? B.ground[w] : B.air[w];
        if (noBoost) m &= ~B.boost[w];
        if (jumpOk) m |= B.jump[w];
        const int n = w < 2 ? 16 : (kNumActions - 64) / 2;          // byte pairs of this word: 16, 16, 13
        for (int k = 0; k < n; k++) o[16 * w + k] = (uint16_t)(((m >> (2 * k)) & 1u) | (((m >> (2 * k + 1)) & 1u) << 8));
    }
}

// ---------------------------------------------------------------------------------------------- game events
GGL_HD bool ball_scored(const ArenaDev& A) {
    float y = A.ball.pos.y * kBT2UU;
    return fabsf(y) > (rl::GOAL_THRESHOLD_Y + rl::BALL_RADIUS);
}

// Arena::IsBallProbablyGoingIn for soccar (Arena.cpp:827-860)
GGL_HD bool ball_probably_going_in(const ArenaDev& A, float maxTime, float extraMargin, int* goalTeam) {
    // heatseeker: the reference has no prediction for this mode and raises a fatal error as soon as the tracker asks for one
    // (Arena.cpp:867-870, i.e. on the first fast ball); here the tracker simply sees no shot, so shot / save events stay off
    if (A.gameMode == kModeHeatseeker) return false;
    V3 p = A.ball.pos * kBT2UU, v = A.ball.vel * kBT2UU;
    if (fabsf(v.y) < kEps) return false;
    float sgn = fsgn(v.y);
    float goalY = rl::GOAL_THRESHOLD_Y * sgn;
    float dist = fabsf(p.y - goalY);
    float t = dist / fabsf(v.y);
    if (t > maxTime) return false;
    V3 g = v3(0, 0, rl::GRAVITY_Z);
    V3 ex = p + (v * t) + (g * t * t) / 2.f;
    float margin = rl::BALL_RADIUS * 0.1f + extraMargin;
    if (ex.z > 642.775f + margin) return false;
    if (fabsf(ex.x) > 892.755f + margin) return false;
    if (goalTeam) *goalTeam = (sgn < 0.f) ? 0 : 1;
    return true;
}

// GetShooterPasser (GameEventTracker.cpp:5-46): cars visited in players[] slot order; returns car indices or -1
GGL_HD bool shooter_passer(const ArenaDev& A, int team, int& shooter, bool findPasser, int& passer, uint64_t maxShooterTicks,
                           uint64_t maxPasserTicks) {
    shooter = passer = -1;
    for (int s = 0; s < A.numCars; s++) {
        int ci = A.slotCar[s];
        const CarDev& c = A.cars[ci];
        if (c.team != team || !c.hitValid) continue;
        if (c.hitTick + maxShooterTicks >= A.tickCount) {
            if (shooter < 0 || c.hitTick > A.cars[shooter].hitTick) shooter = ci;
        }
    }
    if (shooter >= 0 && findPasser) {
        uint64_t shootTick = A.cars[shooter].hitTick;
        for (int s = 0; s < A.numCars; s++) {
            int ci = A.slotCar[s];
            const CarDev& c = A.cars[ci];
            if (c.team != team || !c.hitValid || ci == shooter) continue;
            if (c.hitTick + maxPasserTicks >= shootTick) {
                if (passer < 0 || c.hitTick > A.cars[passer].hitTick) passer = ci;
            }
        }
    }
    return shooter >= 0;
}

GGL_HD int slot_of_car(const ArenaDev& A, int ci) {
    for (int s = 0; s < A.numCars; s++)
        if (A.slotCar[s] == ci) return s;
    return 0;
}
GGL_HD void set_event(ArenaDev& A, int ci, int ev) {
    if (ci >= 0) A.players[slot_of_car(A, ci)].events[ev] = 1;
}

// GameEventTracker::Update with the default config (shotMinSpeed 1750, shotTouchMinDelay .3, shotEventCooldown 1,
// shotMinScoreTime 2, goalMaxTouchTime 4, passMaxTouchTime 2)
GGL_HD void event_tracker_update(ArenaDev& A) {
    bool scored = ball_scored(A);
    const float tickrate = 120.f;
    uint64_t cnt = A.ball.updateCounter;
    if (cnt > A.lastBallUpdateCount) {
        uint64_t deltaTicks = cnt - A.lastBallUpdateCount;
        float deltaTime = deltaTicks * kTick;
        if (scored && !A.ballScoredLast) {
            int sh, pa;
            int team = (-A.ball.pos.y < 0.f) ? 0 : 1;
            if (shooter_passer(A, team, sh, true, pa, (uint64_t)(4.0f * tickrate), (uint64_t)(2.0f * tickrate))) {
                set_event(A, sh, EV_GOAL);
                set_event(A, pa, EV_ASSIST);
            }
        } else {
            if (!A.ballShot) {
                if (A.shotCooldown > 0.f) {
                    A.shotCooldown = fmaxf(A.shotCooldown - deltaTime, 0.f);
                } else {
                    float sp2 = len2(A.ball.vel * kBT2UU);
                    if (sp2 >= 1750.f * 1750.f) {
                        int goalTeam = 0;
                        if (ball_probably_going_in(A, 2.0f, 0.f, &goalTeam)) {
                            int shooterTeam = goalTeam == 0 ? 1 : 0;
                            uint64_t minDelay = (uint64_t)(0.3f * tickrate);
                            int sh, pa;
                            if (shooter_passer(A, shooterTeam, sh, true, pa, deltaTicks + minDelay, (uint64_t)(2.0f * tickrate))) {
                                uint64_t since = A.tickCount - A.cars[sh].hitTick;
                                if (since >= minDelay) {
                                    A.ballShot = 1;
                                    A.ballShotGoalTeam = (uint8_t)goalTeam;
                                    A.shotCooldown = 1.0f;
                                    set_event(A, sh, EV_SHOT);
                                    set_event(A, pa, EV_SHOTPASS);
                                }
                            }
                        }
                    }
                }
            } else {
                if (!ball_probably_going_in(A, 2.0f, 0.f, nullptr)) {
                    int saver, unused;
                    if (shooter_passer(A, A.ballShotGoalTeam, saver, false, unused, deltaTicks, 0)) set_event(A, saver, EV_SAVE);
                    A.ballShot = 0;
                }
            }
        }
    } else if (cnt == A.lastBallUpdateCount) {
        return;
    } else {
        A.ballScoredLast = 0; A.ballShot = 0; A.shotCooldown = 0.f;
    }
    A.ballScoredLast = scored;
    A.lastBallUpdateCount = cnt;
}

}  // namespace ggl
