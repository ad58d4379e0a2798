// Second part of the env layer: per-step GameState snapshot, terminals, rewards, observation rows, kickoff reset.
// See env_logic.cuh for the reference citations.
#pragma once
#include "env_logic.cuh"

namespace ggl {

// What the reward / obs code needs of one player, in Unreal units like RLGC::Player (: CarState)
struct PlayerView {
    V3 pos, vel, angVel, fwd, right, up;
    float boost;
    int team, carId;
    bool onGround, hasFlipOrJump, demoed, hasJumped, touchedStep;
};

GGL_HD PlayerView player_view(const ArenaDev& A, int ci) {
    const CarDev& c = A.cars[ci];
    PlayerView p;
    p.pos = c.pos * kBT2UU; p.vel = c.vel * kBT2UU; p.angVel = c.angVel;
    const M3 rb = car_reported_basis(c);
    p.fwd = col(rb, 0); p.right = col(rb, 1); p.up = col(rb, 2);
    p.boost = c.boost; p.team = c.team; p.carId = (int)c.id;
    p.onGround = c.isOnGround; p.hasFlipOrJump = car_has_flip_or_jump(c); p.demoed = c.isDemoed; p.hasJumped = c.hasJumped;
    p.touchedStep = c.hitValid && c.hitTick >= A.gsLastTickCount;
    return p;
}

// KickoffProximityReward2v2Enhanced (RG/Rewards/KickoffProximityReward2v2Enhanced.h:14-367), statement by statement.
// params: [0] goerReward, [1] cheaterReward (unused by the reference too), [2] dynamicWeight (unused), [3] rotationPrepWeight.
// `state.players` is walked in slot order; the teammate is the LAST other player of the team the loop meets (:64-69).
GGL_HD float rs_clampf(float v, float lo, float hi) { float a = (v > lo) ? v : lo; return (a < hi) ? a : hi; }   // RS_CLAMP = RS_MIN(RS_MAX(v, lo), hi)
GGL_HD float reward_kickoff_proximity_2v2(const ArenaDev& A, const RewardCfg& r, const PlayerView& p, V3 ballPos, V3 ballVel) {
    // IsKickoffActive :48-58
    if (!(len(ballVel) < 2.f && ballPos.z < 150.f && len(v3(ballPos.x, ballPos.y, 0.f)) < 50.f)) return 0.f;
    // AnalyzeTeamState :60-96
    bool hasTeammate = false;
    PlayerView mate = p;
    float mateDist = 0.f, closestOpp = 3.402823466e+38f, secondOpp = 3.402823466e+38f, totalOppSpeed = 0.f;
    V3 oppCom = v3(0, 0, 0);
    int oppCount = 0;
    for (int s = 0; s < A.numCars; s++) {
        const PlayerView q = player_view(A, A.slotCar[s]);
        if (q.team == p.team && q.carId != p.carId) {
            mate = q; hasTeammate = true; mateDist = len(q.pos - ballPos);
        } else if (q.team != p.team) {
            const float d = len(q.pos - ballPos);
            totalOppSpeed += len(q.vel);
            oppCount++;
            if (d < closestOpp) { secondOpp = closestOpp; closestOpp = d; }
            else if (d < secondOpp) secondOpp = d;
            oppCom = oppCom + q.pos;
        }
    }
    if (oppCount > 0) oppCom = oppCom / (float)oppCount;
    (void)secondOpp; (void)totalOppSpeed;
    if (!hasTeammate) return 0.f;
    const float playerDist = len(p.pos - ballPos);
    // DeterminePlayerRole :98-122 (+ CalculateSpawnAdvantage :124-133)
    const float distanceScore = (playerDist < mateDist) ? 0.4f : 0.f;
    const V3 playerToBall = rs_normalized(ballPos - p.pos);
    const V3 mateToBall = rs_normalized(ballPos - mate.pos);
    const float playerVelToBall = dot(p.vel, playerToBall);
    const float mateVelToBall = dot(mate.vel, mateToBall);
    const float speedScore = (playerVelToBall > mateVelToBall) ? 0.3f : 0.f;
    const float boostScore = (p.boost > mate.boost + 10.f) ? 0.2f : 0.f;
    const float angP = atan2f(p.pos.y - ballPos.y, p.pos.x - ballPos.x), angM = atan2f(mate.pos.y - ballPos.y, mate.pos.x - ballPos.x);
    const float spawnScore = ((fabsf(angP - angM) > (3.14159f / 3.f)) ? 1.f : 0.f) * 0.1f;
    const float totalScore = distanceScore + speedScore + boostScore + spawnScore;
    if (totalScore >= 0.5f) {
        // CalculateGoerReward :135-161
        const float goer = r.params[0];
        const float baseReward = (playerDist < closestOpp) ? goer : -goer * 0.5f;
        const float speedBonus = rs_clampf(playerVelToBall / 2300.f, -0.3f, 0.3f);
        float boostEfficiency = 0.f;
        if (p.boost > 50.f && playerDist > 1000.f) boostEfficiency = 0.1f;
        else if (p.boost < 20.f && playerDist > 800.f) boostEfficiency = -0.15f;
        const float approach = dot(playerToBall, rs_normalized(p.vel));
        const float angleBonus = ((0.f > approach) ? 0.f : approach) * 0.2f;
        return rs_clampf(baseReward + speedBonus + boostEfficiency + angleBonus, -1.5f, 1.5f);
    }
    // CalculateCheaterReward :163-188
    const V3 ownGoal = (p.team == 0) ? v3(0.f, -6000.f, 642.775f / 2) : v3(0.f, 6000.f, 642.775f / 2);
    // CalculateDynamicIdealPosition :190-228
    const V3 fieldCenter = v3(0.f, 0.f, 100.f);
    const V3 centerMul = v3(fieldCenter.x * 1.3f, fieldCenter.y * 1.3f, fieldCenter.z * 1.3f);
    const V3 baseIdeal = (ownGoal + centerMul) * 0.5f;
    V3 threat = rs_normalized(oppCom - ownGoal);
    threat = v3(threat.x * 200.f, threat.y * 200.f, threat.z * 200.f);
    V3 mateOffset = v3(0.f, 0.f, 0.f);
    if (len(mate.pos - fieldCenter) > 1500.f) {
        const V3 dir = rs_normalized(mate.pos - baseIdeal);
        mateOffset = v3(dir.x * 300.f, dir.y * 300.f, dir.z * 300.f);
    }
    const V3 threatAdj = v3(threat.x * 0.3f, threat.y * 0.3f, threat.z * 0.3f);
    const V3 mateAdj = v3(mateOffset.x * 0.2f, mateOffset.y * 0.2f, mateOffset.z * 0.2f);
    V3 ideal = (baseIdeal + threatAdj) + mateAdj;
    ideal.x = rs_clampf(ideal.x, -3000.f, 3000.f);
    ideal.y = rs_clampf(ideal.y, -4000.f, 4000.f);
    ideal.z = (ideal.z > 17.f) ? ideal.z : 17.f;
    const float distToIdeal = len(p.pos - ideal);
    // CalculatePositioningReward :230-254
    float positioning;
    if (distToIdeal <= 600.f) positioning = 0.5f * (1.f - (distToIdeal / 600.f));
    else if (distToIdeal <= 1200.f) { const float ratio = (distToIdeal - 600.f) / (1200.f - 600.f); positioning = 0.5f * (1.f - ratio) * 0.7f; }
    else if (distToIdeal <= 2000.f) { const float ratio = (distToIdeal - 1200.f) / (2000.f - 1200.f); positioning = -0.1f * ratio; }
    else positioning = -0.3f;
    // CalculateStrategicBoostReward :256-286: only the six large pads (z > 72) of CommonValues::BOOST_LOCATIONS, in table order
    float bestBoost = 0.f;
    for (int i = 0; i < 6; i++) {
        const V3 bpz = v3((i & 1) ? ((i == 3) ? 3584.f : 3072.f) : ((i == 2) ? -3584.f : -3072.f), (i < 2) ? -4096.f : ((i < 4) ? 0.f : 4096.f), 73.f);
        const float distToBoost = len(p.pos - bpz);
        const float accessibility = 1.f - rs_clampf(distToBoost / 1500.f, 0.f, 1.f);
        // CalculateBoostStrategicValue :288-300
        const float distToBall = len(bpz - ballPos);
        const bool corner = fabsf(bpz.x) > 2500.f && fabsf(bpz.y) > 3500.f;
        const float baseValue = corner ? 0.8f : 0.6f;
        const float proximity = 1.f - rs_clampf(distToBall / 3000.f, 0.f, 1.f);
        const float strategic = baseValue * (0.3f + proximity * 0.7f);
        // CalculateBoostDenialValue :302-307
        const float deny = rs_clampf(1.f - (len(oppCom - bpz) / 2000.f), 0.f, 0.3f);
        const float total = accessibility * (strategic + deny);
        bestBoost = (bestBoost > total) ? bestBoost : total;
    }
    float boostLevel = 1.f;
    if (p.boost < 30.f) boostLevel = 1.5f; else if (p.boost > 80.f) boostLevel = 0.5f;
    const float boostReward = (bestBoost * boostLevel) * 0.25f;
    // CalculateRotationPreparation :309-325 (+ CalculateOptimalSupportPosition :327-342)
    const V3 mateToGoal = rs_normalized(ownGoal - mate.pos);
    const V3 perp = rs_normalized(v3(-mateToGoal.y, mateToGoal.x, 0.f));
    const V3 goalOff = v3(mateToGoal.x * 800.f, mateToGoal.y * 800.f, mateToGoal.z * 800.f);
    const V3 perpOff = v3(perp.x * 600.f, perp.y * 600.f, perp.z * 600.f);
    const V3 support = (mate.pos + goalOff) + perpOff;
    const float supportReadiness = 1.f - rs_clampf(len(p.pos - support) / 1000.f, 0.f, 1.f);
    const float align0 = dot(rs_normalized(p.vel), rs_normalized(support - p.pos));
    const float velAlign = (0.f > align0) ? 0.f : align0;
    const float rotation = (supportReadiness * 0.7f + velAlign * 0.3f) * r.params[3];
    // CalculateOpponentAwareness :344-353
    const float awareAngle = dot(rs_normalized(oppCom - p.pos), playerToBall);
    const float awareness = rs_clampf(awareAngle * 0.5f + 0.5f, 0.f, 1.f) * 0.1f;
    // CalculateDynamicCampingPenalty :355-366
    const float distToGoal = len(p.pos - ownGoal);
    float minDist = 800.f;
    if (len(ballPos - ownGoal) < 2000.f) minDist *= 0.7f;
    float camping = 0.f;
    if (distToGoal < minDist) camping = -0.4f * (1.f - (distToGoal / minDist));
    camping *= 0.05f;
    const float totalReward = positioning + boostReward + rotation + awareness + camping;
    return rs_clampf(totalReward, -0.8f, 0.8f);
}

// one reward function for one player (CommonRewards.h); `slot` indexes A.players
GGL_HD float reward_value(const ArenaDev& A, const RewardCfg& r, const PlayerView& p, int slot, bool goalScored, V3 ballPos, V3 ballVel) {
    const EnvPlayer& ep = A.players[slot];
    switch (r.kind) {
    case GGL_REW_PLAYER_GOAL: return (float)ep.events[EV_GOAL];
    case GGL_REW_ASSIST: return (float)ep.events[EV_ASSIST];
    case GGL_REW_SHOT: return (float)ep.events[EV_SHOT];
    case GGL_REW_SHOT_PASS: return (float)ep.events[EV_SHOTPASS];
    case GGL_REW_SAVE: return (float)ep.events[EV_SAVE];
    case GGL_REW_BUMP: return (float)ep.events[EV_BUMP];
    case GGL_REW_BUMPED_PENALTY: return -(float)ep.events[EV_BUMPED];
    case GGL_REW_DEMO: return (float)ep.events[EV_DEMO];
    case GGL_REW_DEMOED_PENALTY: return -(float)ep.events[EV_DEMOED];
    case GGL_REW_GOAL: {
        if (!goalScored) return 0.f;
        int ballTeam = ballPos.y < 0.f ? 0 : 1;
        return (p.team != ballTeam) ? 1.f : r.params[0];
    }
    case GGL_REW_VELOCITY: return len(p.vel) / 2300.f * (1 - 2 * (r.params[0] != 0.f ? 1 : 0));
    case GGL_REW_VELOCITY_BALL_TO_GOAL: {
        bool targetOrange = p.team == 0;
        if (r.params[0] != 0.f) targetOrange = !targetOrange;
        V3 target = targetOrange ? v3(0, 6000.f, 642.775f / 2) : v3(0, -6000.f, 642.775f / 2);
        V3 dir = rs_normalized(target - ballPos);
        return dot(dir, ballVel / 6000.f);
    }
    case GGL_REW_VELOCITY_PLAYER_TO_BALL: {
        V3 dir = rs_normalized(ballPos - p.pos);
        return dot(dir, p.vel / 2300.f);
    }
    case GGL_REW_FACE_BALL: return dot(p.fwd, rs_normalized(ballPos - p.pos));
    case GGL_REW_TOUCH_BALL: return p.touchedStep ? 1.f : 0.f;
    case GGL_REW_SPEED: return len(p.vel) / 2300.f;
    case GGL_REW_WAVEDASH: return (p.onGround && (ep.prevIsFlipping && !ep.prevIsOnGround)) ? 1.f : 0.f;
    case GGL_REW_PICKUP_BOOST:
        return (p.boost > ep.prevBoost) ? sqrtf(p.boost / 100.f) - sqrtf(ep.prevBoost / 100.f) : 0.f;
    case GGL_REW_SAVE_BOOST: return fclamp(powf(p.boost / 100, r.params[0]), 0.f, 1.f);
    case GGL_REW_AIR: return p.onGround ? 0.f : 1.f;
    case GGL_REW_TOUCH_ACCEL: {
        if (!p.touchedStep) return 0.f;
        const float maxSp = 110.f * (250.f / 9.f);
        float prevFrac = fminf(1.f, len(A.prevBallVel) / maxSp), curFrac = fminf(1.f, len(ballVel) / maxSp);
        return curFrac > prevFrac ? (curFrac - prevFrac) : 0.f;
    }
    case GGL_REW_STRONG_TOUCH: {
        if (!p.touchedStep) return 0.f;
        float hit = len(ballVel - A.prevBallVel);
        if (hit < r.params[0]) return 0.f;
        return fminf(1.f, hit / r.params[1]);
    }
    case GGL_REW_KICKOFF_PROXIMITY_2V2: return reward_kickoff_proximity_2v2(A, r, p, ballPos, ballVel);
    default: return 0.f;
    }
}

// pads in RLGym order (CommonValues::BOOST_LOCATIONS) -> RocketSim pad index (GameState.cpp:7-51 boostPadIndexMap)
GGL_HD int pad_index_map(int i) {
    const uint8_t map[34] = {6, 7, 8, 4, 5, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 0, 19, 20, 1, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 2, 3, 31, 32, 33};
    return map[i];
}

GGL_HD void write_player_default(float*& o, const PlayerView& p, bool inv, const float* coef /* EnvCfg::obsCoef */) {
    const float pcx = coef[0], pcy = coef[1], pcz = coef[2], vc = coef[3], ac = coef[4];
    o[0] = (inv ? -p.pos.x : p.pos.x) * pcx; o[1] = (inv ? -p.pos.y : p.pos.y) * pcy; o[2] = p.pos.z * pcz;
    o[3] = inv ? -p.fwd.x : p.fwd.x; o[4] = inv ? -p.fwd.y : p.fwd.y; o[5] = p.fwd.z;
    o[6] = inv ? -p.up.x : p.up.x; o[7] = inv ? -p.up.y : p.up.y; o[8] = p.up.z;
    o[9] = (inv ? -p.vel.x : p.vel.x) * vc; o[10] = (inv ? -p.vel.y : p.vel.y) * vc; o[11] = p.vel.z * vc;
    o[12] = (inv ? -p.angVel.x : p.angVel.x) * ac; o[13] = (inv ? -p.angVel.y : p.angVel.y) * ac; o[14] = p.angVel.z * ac;
    o[15] = p.boost / 100; o[16] = p.onGround ? 1.f : 0.f; o[17] = p.hasFlipOrJump ? 1.f : 0.f; o[18] = p.demoed ? 1.f : 0.f;
    o += 19;
}

GGL_HD V3 inv_xy(V3 v, bool inv) { return inv ? v3(-v.x, -v.y, v.z) : v; }

GGL_HD void write_player_advanced(float*& o, const PlayerView& pl, bool inv, V3 ballPos, V3 ballVel) {
    const float PC = 1.0f / 2300.0f, VC = 1.0f / 2300.0f, AC = 1.0f / 5.5f;
    V3 pos = inv_xy(pl.pos, inv), vel = inv_xy(pl.vel, inv), av = inv_xy(pl.angVel, inv);
    V3 f = inv_xy(pl.fwd, inv), r = inv_xy(pl.right, inv), u = inv_xy(pl.up, inv);
    o[0] = pos.x * PC; o[1] = pos.y * PC; o[2] = pos.z * PC;
    o[3] = f.x; o[4] = f.y; o[5] = f.z;
    o[6] = u.x; o[7] = u.y; o[8] = u.z;
    o[9] = vel.x * VC; o[10] = vel.y * VC; o[11] = vel.z * VC;
    o[12] = av.x * AC; o[13] = av.y * AC; o[14] = av.z * AC;
    o[15] = (f.x * av.x + f.y * av.y + f.z * av.z) * AC;
    o[16] = (r.x * av.x + r.y * av.y + r.z * av.z) * AC;
    o[17] = (u.x * av.x + u.y * av.y + u.z * av.z) * AC;
    V3 rb = ballPos - pos;
    o[18] = (f.x * rb.x + f.y * rb.y + f.z * rb.z) * PC;
    o[19] = (r.x * rb.x + r.y * rb.y + r.z * rb.z) * PC;
    o[20] = (u.x * rb.x + u.y * rb.y + u.z * rb.z) * PC;
    V3 rv = ballVel - vel;
    o[21] = (f.x * rv.x + f.y * rv.y + f.z * rv.z) * VC;
    o[22] = (r.x * rv.x + r.y * rv.y + r.z * rv.z) * VC;
    o[23] = (u.x * rv.x + u.y * rv.y + u.z * rv.z) * VC;
    o[24] = pl.boost * 0.01f; o[25] = pl.onGround ? 1.f : 0.f; o[26] = pl.hasFlipOrJump ? 1.f : 0.f;
    o[27] = pl.demoed ? 1.f : 0.f; o[28] = pl.hasJumped ? 1.f : 0.f;
    o += 29;
}

// DefaultObsPadded's two std::shuffle calls for ONE row (DefaultObsPadded.cpp:59-60): teammates (maxPlayers-1 entries) first,
// then opponents (maxPlayers entries); item k of a list before the shuffle = k-th such player in players[] order, zero rows after
GGL_HD void padded_obs_draw(uint32_t& rng, int maxPlayers, uint32_t& permMates, uint32_t& permOpps) {
    permMates = std_shuffle_nib(rng, maxPlayers - 1);
    permOpps = std_shuffle_nib(rng, maxPlayers);
}
// The rows of an arena are built one after the other in players[] order on one engine (EnvSet.cpp:257-268), so row `slot` sees
// the stream after the draws of rows 0..slot-1: replayed here (a handful of integer steps) so that every row's lane can work
// on its own; returns the stream after row `slot`.
GGL_HD uint32_t padded_obs_draws_upto(uint32_t rng, int maxPlayers, int slot, uint32_t& permMates, uint32_t& permOpps) {
    for (int s = 0; s <= slot; s++) padded_obs_draw(rng, maxPlayers, permMates, permOpps);
    return rng;
}

// One observation row for players[slot]. permMates / permOpps: DefaultObsPadded only.
GGL_HD void build_obs_row(const ArenaDev& A, const EnvCfg& cfg, int slot, float* out, uint32_t permMates = 0, uint32_t permOpps = 0) {
    const int ci = A.slotCar[slot];
    const PlayerView me = player_view(A, ci);
    const bool inv = me.team == 1;
    V3 bp = inv_xy(A.ball.pos * kBT2UU, inv), bv = inv_xy(A.ball.vel * kBT2UU, inv), bw = inv_xy(A.ball.angVel, inv);
    float* o = out;
    if (cfg.obsKind == GGL_OBS_ADVANCED) {
        const float PC = 1 / 5000.f, VC = 1 / 2300.f, AC = 1 / 3.f;      // class constants bind in BuildObs
        o[0] = bp.x * PC; o[1] = bp.y * PC; o[2] = bp.z * PC; o[3] = bv.x * VC; o[4] = bv.y * VC; o[5] = bv.z * VC;
        o[6] = bw.x * AC; o[7] = bw.y * AC; o[8] = bw.z * AC;
    } else {
        const float* k = cfg.obsCoef;
        o[0] = bp.x * k[0]; o[1] = bp.y * k[1]; o[2] = bp.z * k[2];
        o[3] = bv.x * k[3]; o[4] = bv.y * k[3]; o[5] = bv.z * k[3];
        o[6] = bw.x * k[4]; o[7] = bw.y * k[4]; o[8] = bw.z * k[4];
    }
    o += 9;
    for (int i = 0; i < 8; i++) o[i] = A.players[slot].prevAction[i];
    o += 8;
    for (int i = 0; i < kNumPads; i++) {
        // boostPads[i] = pad map[i]; boostPadsInv[i] = pad map[33-i]; timers likewise; GetBoostPadTimers is swapped
        const PadDev& pd = A.pads[pad_index_map(inv ? 33 - i : i)];
        if (cfg.obsKind == GGL_OBS_ADVANCED) {
            const PadDev& tm = A.pads[pad_index_map(inv ? i : 33 - i)];
            o[i] = pd.isActive ? 1.0f : 1.0f / (1.0f + tm.cooldown);
        } else {
            o[i] = pd.isActive ? 1.f : 0.f;
        }
    }
    o += kNumPads;
    if (cfg.obsKind == GGL_OBS_ADVANCED) write_player_advanced(o, me, inv, bp, bv);
    else write_player_default(o, me, inv, cfg.obsCoef);
    if (cfg.obsKind == GGL_OBS_DEFAULT_PADDED) {
        // teammates then opponents, each list padded with zero rows and shuffled: list[j] = item nibble j of the permutation
        const int nm = cfg.obsMaxPlayers - 1, no = cfg.obsMaxPlayers;
        for (int i = 0; i < 19 * (nm + no); i++) o[i] = 0.f;
        int km = 0, ko = 0;
        for (int s = 0; s < A.numCars; s++) {
            const int oc = A.slotCar[s];
            if (oc == ci) continue;
            const bool mate = A.cars[oc].team == me.team;
            const int k = mate ? km++ : ko++;
            const uint32_t perm = mate ? permMates : permOpps;
            int j = 0;
            while (j < 7 && (int)nib_get(perm, j) != k) j++;
            float* dst = o + 19 * (mate ? j : nm + j);
            write_player_default(dst, player_view(A, oc), inv, cfg.obsCoef);
        }
        return;
    }
    for (int pass = 0; pass < 2; pass++)
        for (int s = 0; s < A.numCars; s++) {
            int oc = A.slotCar[s];
            if (oc == ci) continue;
            const bool mate = A.cars[oc].team == me.team;
            if (mate != (pass == 0)) continue;
            PlayerView ov = player_view(A, oc);
            if (cfg.obsKind == GGL_OBS_ADVANCED) write_player_advanced(o, ov, inv, bp, bv);
            else write_player_default(o, ov, inv, cfg.obsCoef);
        }
}

}  // namespace ggl
