// One 120 Hz arena tick and its phase structure (Arena::Step, RocketSim/src/Sim/Arena/Arena.cpp:716-812).
//
// The tick is written as PHASES so that the lanes of the group that owns an arena can share the work:
//   phase_cars_pre   per car   : Car::_PreTickUpdate (rays, friction, control logic, suspension, boost)
//   world phases     per body / leader : gravity, damping, contacts, solver, integration (see world_forces below)
//   phase_cars_post  per car   : Car::_PostTickUpdate + _FinishPhysicsTick
//   phase_finish     leader    : boost pad pickup, ball finish, tick counter
// `Exec` tells a lane which cars it owns and how to synchronise the group (host build: one lane, no-op sync).
#pragma once
#include "sim_world.cuh"
#include "sim_boxbox.cuh"
#include "sim_boxtri.cuh"

namespace ggl {

// block-level meeting points of the tick (bit k = after phase k, see ExecGroup::phase_sync): rays, contacts, solver, end
#ifndef GGL_PSYNC_MASK
#define GGL_PSYNC_MASK (4 | 32 | 128 | 2048)
#endif

// weights of the per-tick work estimate (contact rows, triangle candidates, cars on their wheels); the experiment build reads
// them from a device symbol set through GGL_COST_W="a,b,c" at create
#if defined(GGL_COST_SWEEP) && defined(__CUDACC__)
__device__ int g_costW[3] = {4, 6, 1};
#endif
#if defined(GGL_COST_SWEEP) && defined(__CUDA_ARCH__)
#define GGL_COST_W(i) g_costW[i]
#else
#define GGL_COST_W(i) ((i) == 0 ? 4 : (i) == 1 ? 6 : 1)
#endif

struct ExecSerial {
    GGL_HD int lane() const { return 0; }
    GGL_HD int lanes() const { return 1; }
    GGL_HD void sync() const {}
    GGL_HD void phase_sync() const {}
};

// Boost pad table, RocketSim order: 6 big pads then 28 small ones (RLConst.h:192-253), Unreal units.
#if defined(__CUDACC__)
#define GGL_TABLE static __device__ __constant__
#else
#define GGL_TABLE static const
#endif
GGL_TABLE float kPadPosUU[kNumPads][3] = {
    {-3584.f, 0.f, 73.f}, {3584.f, 0.f, 73.f}, {-3072.f, 4096.f, 73.f}, {3072.f, 4096.f, 73.f}, {-3072.f, -4096.f, 73.f}, {3072.f, -4096.f, 73.f},
    {0.f, -4240.f, 70.f}, {-1792.f, -4184.f, 70.f}, {1792.f, -4184.f, 70.f}, {-940.f, -3308.f, 70.f}, {940.f, -3308.f, 70.f}, {0.f, -2816.f, 70.f},
    {-3584.f, -2484.f, 70.f}, {3584.f, -2484.f, 70.f}, {-1788.f, -2300.f, 70.f}, {1788.f, -2300.f, 70.f}, {-2048.f, -1036.f, 70.f}, {0.f, -1024.f, 70.f},
    {2048.f, -1036.f, 70.f}, {-1024.f, 0.f, 70.f}, {1024.f, 0.f, 70.f}, {-2048.f, 1036.f, 70.f}, {0.f, 1024.f, 70.f}, {2048.f, 1036.f, 70.f},
    {-1788.f, 2300.f, 70.f}, {1788.f, 2300.f, 70.f}, {-3584.f, 2484.f, 70.f}, {3584.f, 2484.f, 70.f}, {0.f, 2816.f, 70.f}, {-940.f, 3308.f, 70.f},
    {940.f, 3308.f, 70.f}, {-1792.f, 4184.f, 70.f}, {1792.f, 4184.f, 70.f}, {0.f, 4240.f, 70.f}};
GGL_HD V3 pad_pos_uu(int i) {
#if defined(__CUDACC__) && !defined(__CUDA_ARCH__)
    (void)i; return v3(0, 0, 0);          // host pass of nvcc: the table is device-resident; the host build of the simulator is g++
#else
    return v3(kPadPosUU[i][0], kPadPosUU[i][1], kPadPosUU[i][2]);
#endif
}

// world AABB of the car's compound shape (btCompoundShape::getAabb over the child box AABB)
GGL_HD void car_aabb(const CarDev& c, V3& mn, V3& mx) {
    const BoxGeom bg = car_box();
    V3 off = car_hitbox_offset();
    V3 lmin = off - bg.full, lmax = off + bg.full;
    V3 lhe = (lmax - lmin) * 0.5f, lc = (lmax + lmin) * 0.5f;
    V3 centre = c.pos + mul(c.basis, lc);
    V3 ext = v3(dot(vabs(c.basis.r0), lhe), dot(vabs(c.basis.r1), lhe), dot(vabs(c.basis.r2), lhe));
    mn = centre - ext; mx = centre + ext;
}

// BoostPadGrid: 8 x 10 cells of 1024 uu, at most one pad per cell (BoostPadGrid.cpp:27-38 on the pad table above): pad index
// of cell (x, y) or -1
GGL_TABLE signed char kCellPad[8][10] = {
    {-1, -1, 12, -1, -1, 0, -1, 26, -1, -1}, {-1, 4, -1, -1, -1, -1, -1, -1, -1, 2}, {7, -1, 14, 16, -1, -1, 21, 24, -1, 31},
    {-1, 9, -1, -1, -1, 19, -1, -1, 29, -1}, {6, 10, 11, -1, 17, -1, 22, 28, 30, 33}, {8, -1, 15, -1, -1, 20, -1, 25, -1, 32},
    {-1, -1, -1, 18, -1, -1, 23, -1, -1, -1}, {-1, 5, 13, -1, -1, 1, -1, 27, -1, 3}};
GGL_HD int cell_pad(int x, int y) {
#if defined(__CUDACC__) && !defined(__CUDA_ARCH__)
    (void)x; (void)y; return -1;          // host pass of nvcc (see pad_pos_uu)
#else
    return kCellPad[x][y];
#endif
}

// BoostPadGrid::CheckCollision (BoostPadGrid.cpp:5-25) for one car: the pads of the 3 x 3 cells around the car, each through
// BoostPad::_CheckCollide (Sim/BoostPad/BoostPad.cpp:62-86). A car can be inside at most one pad (cylinders of radius <= 208 uu
// and lock boxes of half width <= 160 uu around pads that are >= 1000 uu apart), so the result is ONE pad index per car
// (A.padHit); the locks themselves are written afterwards in players[] slot order, because the reference visits its cars one
// after the other and a later car overwrites the lock of an earlier one (pads_apply_locks).
GGL_HD void car_check_pads(ArenaDev& A, int ci) {
    const CarDev& car = A.cars[ci];
    A.padHit[ci] = -1;
    if (car.isDemoed || car.boost >= 100.f) return;
    const V3 posUU = car.pos * kBT2UU;
    if (posUU.z > rl::PAD_CYL_HEIGHT + 250.f) return;
    const int ix = (int)(posUU.x / 1024 + 4), iy = (int)(posUU.y / 1024 + 5);
    int lx0 = ix - 1 < 0 ? 0 : ix - 1, lx1 = ix + 1 > 7 ? 7 : ix + 1, ly0 = iy - 1 < 0 ? 0 : iy - 1, ly1 = iy + 1 > 9 ? 9 : iy + 1;
    {   // cull (exact): a pad lies inside its own cell, and neither test below can pass for a pad further than 300 uu from the
        // car along x or y (cylinder radius <= 208; lock box half width 160 + the car's AABB reach, < 100 uu for the hitbox
        // half diagonal 75 plus its offset 25) - so only the cells that the car's 300 uu surroundings touch can matter
        const float R = 300.f;
        const int cx0 = (int)((posUU.x - R) / 1024 + 4), cx1 = (int)((posUU.x + R) / 1024 + 4);
        const int cy0 = (int)((posUU.y - R) / 1024 + 5), cy1 = (int)((posUU.y + R) / 1024 + 5);
        if (cx0 > lx0) lx0 = cx0;
        if (cx1 < lx1) lx1 = cx1;
        if (cy0 > ly0) ly0 = cy0;
        if (cy1 < ly1) ly1 = cy1;
    }
    for (int i = lx0; i <= lx1; i++)
        for (int j = ly0; j <= ly1; j++) {
            const int p = cell_pad(i, j);
            if (p < 0) continue;
            const V3 pp = pad_pos_uu(p);
            const bool big = p < 6;
            const V3 pbt = pp * kUU2BT;
            const PadDev& pd = A.pads[p];
            bool colliding = false;
            if (pd.prevLockedCarId == car.id) {
                V3 amin, amax;
                car_aabb(car, amin, amax);
                const float boxRad = (big ? rl::PAD_BOX_RAD_BIG : rl::PAD_BOX_RAD_SMALL) * kUU2BT;
                const V3 bmin = pbt - v3(boxRad, boxRad, 0), bmax = pbt + v3(boxRad, boxRad, rl::PAD_BOX_HEIGHT * kUU2BT);
                colliding = (bmax.x > amin.x && bmax.y > amin.y && bmax.z > amin.z) && (bmin.x < amax.x && bmin.y < amax.y && bmin.z < amax.z);
            } else {
                const float rad = (big ? rl::PAD_CYL_RAD_BIG : rl::PAD_CYL_RAD_SMALL) * kUU2BT;
                const float dx = car.pos.x - pbt.x, dy = car.pos.y - pbt.y;
                if (dx * dx + dy * dy < rad * rad) colliding = fabsf(car.pos.z - pbt.z) < (rl::PAD_CYL_HEIGHT * kUU2BT);
            }
            if (colliding) A.padHit[ci] = (int8_t)p;
        }
}
GGL_HD void pads_apply_locks(ArenaDev& A) {
    for (int s = 0; s < A.numCars; s++) {
        const int c = A.slotCar[s];
        if (A.padHit[c] >= 0) A.pads[A.padHit[c]].curLockedCarId = (uint8_t)A.cars[c].id;
    }
}

// boost pads pre-tick (BoostPad.cpp:51-59), one pad
GGL_HD void pad_pre_tick(PadDev& pd) {
    if (pd.cooldown > 0.f) pd.cooldown = fmaxf(pd.cooldown - kTick, 0.f);
    pd.isActive = (pd.cooldown == 0.f);
    pd.curLockedCarId = 0;
}

// The rigid-body part of a tick, split so that body b (0 = ball, 1 + c = car c) is handled by lane b % lanes:
//   world_forces(b)   gravity + damping                                    per body
//   collide_body(b)   contact generation into the body's windows           per body   (sim_world.cuh)
//   gather_contacts   compaction to solver order + ordered callbacks       leader
//   solve_island      PGS over one contact island                          per island | body_free_velocity(b) for bodies without rows
//   world_integrate(b) transform integration, clear forces                 per body
GGL_HD void world_forces(ArenaDev& A, int b) {
    const V3 g = v3(0, 0, rl::GRAVITY_Z * kUU2BT);
    if (b == 0) {
        if (A.gameMode == kModeHeatseeker) ball_heatseeker_pre_tick(A.ball);      // Ball::_PreTickUpdate, before the world steps (Arena.cpp:778)
        if (A.ball.active) A.ball.force += g * (1.f / kBallInvMass);
        A.ball.vel *= powf(1.f - rl::BALL_DRAG, kTick);
        A.ball.angVel *= powf(1.f - 0.f, kTick);
    } else if (!A.cars[b - 1].rbDisabled) {
        A.cars[b - 1].force += g * (1.f / (1.f / rl::CAR_MASS));
    }
}

GGL_HD void world_integrate(ArenaDev& A, int b) {
    if (b == 0) {
        if (A.ballAwake) A.ball.pos = A.ball.pos + A.ball.vel * kTick;
        A.ball.force = v3(0, 0, 0); A.ball.torque = v3(0, 0, 0);
    } else {
        CarDev& c = A.cars[b - 1];
        if (!c.rbDisabled) {
            V3 np; M3 nb;
            integrate_transform(c.pos, c.basis, c.vel, c.angVel, kTick, np, nb);
            c.pos = np; c.basis = nb;
            car_update_inertia(c);
        }
        c.force = v3(0, 0, 0); c.torque = v3(0, 0, 0);
    }
}

// BoostPad::_PostTickUpdate (BoostPad.cpp:88-105) of one pad. Pads are independent here: a car locks at most one pad per tick
// (car_check_pads), so two pads never feed the same car.
GGL_HD void pad_post_tick(ArenaDev& A, int p) {
    PadDev& pd = A.pads[p];
    uint32_t locked = 0;
    if (pd.curLockedCarId) {
        locked = pd.curLockedCarId;
        if (pd.isActive) {
            CarDev& c = A.cars[locked - 1];
            c.boost = fminf(c.boost + (p < 6 ? rl::PAD_BOOST_BIG : rl::PAD_BOOST_SMALL), rl::BOOST_MAX);
            pd.isActive = 0;
            pd.cooldown = p < 6 ? rl::PAD_COOLDOWN_BIG : rl::PAD_COOLDOWN_SMALL;
        }
    }
    pd.prevLockedCarId = (uint8_t)locked;
}

GGL_HD void phase_finish(ArenaDev& A) {
    // Ball::_FinishPhysicsTick
    BallDev& b = A.ball;
    if (!is_zero(b.velImpulseCache)) { b.vel += b.velImpulseCache; b.velImpulseCache = v3(0, 0, 0); }
    const float maxV = rl::BALL_MAX_SPEED * kUU2BT;
    if (len2(b.vel) > maxV * maxV) b.vel = normalized(b.vel) * maxV;
    if (len2(b.angVel) > rl::BALL_MAX_ANG_SPEED * rl::BALL_MAX_ANG_SPEED) b.angVel = normalized(b.angVel) * rl::BALL_MAX_ANG_SPEED;
    b.updateCounter++;
    A.tickCount++;
}

template <class Exec>
GGL_HD void arena_tick(ArenaDev& A, const WorldDev& W, const Exec& ex) {
    GGL_PH_BEGIN();
    if (ex.lane() == 0) {
        A.ball.active = (len2(A.ball.vel) == 0.f && len2(A.ball.angVel) == 0.f) ? 0u : 1u;
        A.wheelDefer = 0;
        A.poolCount = 0;
        for (int j = 0; j < kMaxPool / 4; j++) reinterpret_cast<uint32_t*>(A.poolOwner)[j] = 0xFFFFFFFFu;   // unclaimed (ContactSink::at)
    }
    for (int c = ex.lane(); c < A.numCars; c += ex.lanes()) { A.cars[c].velSnap = A.cars[c].vel; A.cars[c].angVelSnap = A.cars[c].angVel; }
    if (A.numCars > 0)
        for (int p = ex.lane(); p < kNumPads; p += ex.lanes()) pad_pre_tick(A.pads[p]);
    ex.sync();
    GGL_PH(0);
    for (int c = ex.lane(); c < A.numCars; c += ex.lanes()) car_pre_tick_a(A, c);
    ex.sync();
    if (ex.lane() == 0) {       // after the respawns of car_pre_tick_a (a respawned car has the all-zero box = "use the current transform")
        uint32_t m = 0;
        for (int c = 0; c < A.numCars; c++)
            if (!(A.cars[c].aabbMin.x > A.cars[c].aabbMax.x)) m |= 1u << c;      // not the empty box car_update_ray_aabb leaves for far cars
        A.rayBoxMask = (uint8_t)m;
    }
    ex.sync();
    if (GGL_PSYNC_MASK & 2) ex.phase_sync();
    GGL_PH(1);
    // the front wheels of all cars first, then the rear wheels (which never steer): with 8 lanes and 4 cars a round of the loop is
    // all front or all rear, and the rear round skips the steering frame as a whole
    for (int t = ex.lane(); t < 4 * A.numCars; t += ex.lanes()) {
        const int rear = t >= 2 * A.numCars ? 1 : 0, u = t - 2 * A.numCars * rear;
        const int car = u >> 1, wheel = (u & 1) + 2 * rear;
        if (!A.cars[car].isDemoed) wheel_first(A, W, car, wheel);
    }
    ex.sync();
    if (GGL_PSYNC_MASK & 4) ex.phase_sync();
    GGL_PH(2);
    if (A.wheelDefer == 0) {
        for (int c = ex.lane(); c < A.numCars; c += ex.lanes()) car_pre_tick_c(A, c);
    } else {
        // some wheel stands on a car that the reference pre-ticks earlier (wheel_first): the rest of the pre-tick runs in the
        // reference's car order, the flagged wheels reading the velocities the earlier cars have by then
        for (int s = 0; s < A.numCars; s++) {
            const int c = A.slotCar[s];
            if (c % ex.lanes() == ex.lane()) {
                for (int i = 0; i < 4; i++)
                    if (A.wheelDefer & (1u << (c * 4 + i))) wheel_friction(A, c, i, true);
                car_pre_tick_c(A, c);
            }
            ex.sync();
        }
    }
    ex.sync();
    if (GGL_PSYNC_MASK & 8) ex.phase_sync();
    GGL_PH(3);
    const int nb = 1 + A.numCars;
    for (int b = ex.lane(); b < nb; b += ex.lanes()) world_forces(A, b);
    ex.sync();      // the car-ball callback reads the damped ball velocity
    GGL_PH(4);
    // car vs mesh in three stages: candidates per car, narrow phase (GJK/EPA, the expensive part) of ALL (car, triangle)
    // pairs of the arena spread over the lanes of the group, then the order-dependent manifold updates per car
    for (int c = ex.lane(); c < A.numCars; c += ex.lanes()) car_mesh_gather(A, W, c);
    ex.sync();
    {
        int total = 0;
        for (int c = 0; c < A.numCars; c++) total += A.meshCandN[c];
        for (int p = ex.lane(); p < total; p += ex.lanes()) {
            int c = 0, k = p;
            while (k >= (int)A.meshCandN[c]) { k -= A.meshCandN[c]; c++; }
            car_mesh_narrow(A, W, c, k, ex.lane());
        }
    }
    ex.sync();
    for (int b = ex.lane(); b < nb; b += ex.lanes()) collide_body(A, W, b);
    ex.sync();
    if (GGL_PSYNC_MASK & 32) ex.phase_sync();
    GGL_PH(5);
    if (ex.lane() == 0) {
        int any = A.cntWorld;
        for (int c = 0; c < A.numCars; c++) any += A.cntBall[c] + A.cntRest[c];
        // work estimate of this tick for the launch geometry of the NEXT env step (env_kernels.cuh: arena_of): contact rows
        // (narrow phase + 10 solver iterations each), triangle candidates (a GJK/EPA each), wheels on something
        int cost = GGL_COST_W(0) * any;
        for (int c = 0; c < A.numCars; c++) cost += GGL_COST_W(1) * A.meshCandN[c] + GGL_COST_W(2) * A.cars[c].isOnGround;
        A.stepCost += (uint32_t)cost;
        if (any) gather_contacts(A, W);
        else {      // nothing touched anything this tick (88 % of the arena-ticks of the benchmark's rollout)
            A.numContacts = 0; A.bodiesInContact = 0; A.ballAwake = A.ball.active ? 1 : 0;
            for (int b = 0; b < 8; b++) A.island[b] = (uint8_t)b;
        }
    }
    ex.sync();
    if (GGL_PSYNC_MASK & 64) ex.phase_sync();
    GGL_PH(6);
    // one lane per contact island (its lowest body index); bodies without a row take the free-body path on their own lane
    for (int b = ex.lane(); b < nb; b += ex.lanes()) {
        if (!(A.bodiesInContact & (1u << b))) body_free_velocity(A, b);
        else if (A.island[b] == b) solve_island(A, b);
    }
    ex.sync();
    if (GGL_PSYNC_MASK & 128) ex.phase_sync();
    GGL_PH(7);
    for (int b = ex.lane(); b < nb; b += ex.lanes()) world_integrate(A, b);
    ex.sync();
    if (GGL_PSYNC_MASK & 256) ex.phase_sync();
    GGL_PH(8);
    for (int c = ex.lane(); c < A.numCars; c += ex.lanes()) car_post_tick(A.cars[c]);
    ex.sync();
    if (GGL_PSYNC_MASK & 512) ex.phase_sync();
    GGL_PH(9);
    for (int c = ex.lane(); c < A.numCars; c += ex.lanes()) car_check_pads(A, c);
    ex.sync();
    if (ex.lane() == 0) pads_apply_locks(A);
    ex.sync();
    if (A.numCars > 0)
        for (int p = ex.lane(); p < kNumPads; p += ex.lanes()) pad_post_tick(A, p);
    if (GGL_PSYNC_MASK & 1024) ex.phase_sync();
    GGL_PH(10);
    if (ex.lane() == 0) phase_finish(A);
    ex.sync();
    if (GGL_PSYNC_MASK & 2048) ex.phase_sync();
    GGL_PH(11);
}

}  // namespace ggl
