// Per-car part of one 120 Hz tick: suspension ray casts, wheel friction, and the car control state machine
// (throttle / steer / jump / flip / air control / boost / auto-flip / auto-roll).
//
// Reference behaviour restated (paths under GigaLearnCPP/RLGymCPP/RocketSim/src):
//   Sim/Car/Car.cpp:58-131        Car::_PreTickUpdate           -> car_pre_tick
//   Sim/Car/Car.cpp:330-475       _UpdateWheels                  -> update_wheels
//   Sim/Car/Car.cpp:477-505       _UpdateBoost                   -> update_boost
//   Sim/Car/Car.cpp:507-554       _UpdateJump                    -> update_jump
//   Sim/Car/Car.cpp:556-641       _UpdateAirTorque               -> update_air_torque
//   Sim/Car/Car.cpp:643-761       _UpdateDoubleJumpOrFlip        -> update_double_jump_or_flip
//   Sim/Car/Car.cpp:763-798       _UpdateAutoFlip                -> update_auto_flip
//   Sim/Car/Car.cpp:800-833       _UpdateAutoRoll                -> update_auto_roll
//   Sim/Car/Car.cpp:133-193       _PostTickUpdate/_FinishPhysicsTick -> car_post_tick
//   Sim/btVehicleRL/btVehicleRL.cpp:64-92,118-212,277-402  wheel transform, ray cast, suspension, friction
//   libsrc/bullet3-3.24/BulletDynamics/ConstraintSolver/btContactConstraint.cpp:60-150 resolveSingleCollision/Bilateral
//   libsrc/bullet3-3.24/BulletCollision/NarrowPhaseCollision/btRaycastCallback.cpp:34-110 ray vs triangle
#pragma once
#include "sim_gjk.cuh"
#include "sim_types.cuh"

namespace ggl {

// ---------------------------------------------------------------------------------------------- curves
// LinearPieceCurve::GetOutput (Math/Math.cpp:5-35) over the tables of RLConst.h:340-437
// The tables are spelled out as nested comparisons on literals: as arrays handed to a generic evaluator they ended up on the
// thread's stack (ncu: local loads on the steering curve, every car every tick). curve_seg is the interpolation of one piece
// exactly as GetOutput writes it: range and difference of the piece's end points, then in - x0 over the range.
GGL_HD float curve_seg(float in, float x0, float x1, float y0, float y1) {
    float range = x1 - x0;
    float diff = y1 - y0;
    float f = (in - x0) / range;
    return y0 + diff * f;
}
#define GGL_CURVE2(name, X0, X1, Y0, Y1) \
    GGL_HD float name(float in) { if (in <= (X0)) return (Y0); if ((X1) > in) return curve_seg(in, X0, X1, Y0, Y1); return (Y1); }
#define GGL_CURVE3(name, X0, X1, X2, Y0, Y1, Y2)                                                  \
    GGL_HD float name(float in) {                                                                \
        if (in <= (X0)) return (Y0);                                                             \
        if ((X1) > in) return curve_seg(in, X0, X1, Y0, Y1);                                     \
        if ((X2) > in) return curve_seg(in, X1, X2, Y1, Y2);                                     \
        return (Y2);                                                                             \
    }
GGL_HD float curve_steer_angle(float in) {
    if (in <= 0.f) return 0.53356f;
    if (500.f > in) return curve_seg(in, 0.f, 500.f, 0.53356f, 0.31930f);
    if (1000.f > in) return curve_seg(in, 500.f, 1000.f, 0.31930f, 0.18203f);
    if (1500.f > in) return curve_seg(in, 1000.f, 1500.f, 0.18203f, 0.10570f);
    if (1750.f > in) return curve_seg(in, 1500.f, 1750.f, 0.10570f, 0.08507f);
    if (3000.f > in) return curve_seg(in, 1750.f, 3000.f, 0.08507f, 0.03454f);
    return 0.03454f;
}
GGL_CURVE2(curve_powerslide_steer, 0.f, 2500.f, 0.39235f, 0.12610f)
GGL_CURVE3(curve_drive_torque, 0.f, 1400.f, 1410.f, 1.0f, 0.1f, 0.0f)
GGL_CURVE3(curve_non_sticky, 0.f, 0.7075f, 1.f, 0.1f, 0.5f, 1.0f)
GGL_CURVE2(curve_lat_friction, 0.f, 1.f, 1.0f, 0.2f)
GGL_CURVE2(curve_handbrake_long, 0.f, 1.f, 0.5f, 0.9f)
GGL_HD float curve_ball_extra_impulse(float in) {
    if (in <= 0.f) return 0.65f;
    if (500.f > in) return curve_seg(in, 0.f, 500.f, 0.65f, 0.65f);
    if (2300.f > in) return curve_seg(in, 500.f, 2300.f, 0.65f, 0.55f);
    if (4600.f > in) return curve_seg(in, 2300.f, 4600.f, 0.55f, 0.30f);
    return 0.30f;
}
GGL_CURVE3(curve_bump_ground, 0.f, 1400.f, 2200.f, (5.f / 6.f), 1100.f, 1530.f)
GGL_CURVE3(curve_bump_air, 0.f, 1400.f, 2200.f, (5.f / 6.f), 1390.f, 1945.f)
GGL_CURVE3(curve_bump_up, 0.f, 1400.f, 2200.f, (2.f / 6.f), 278.f, 417.f)

// ---------------------------------------------------------------------------------------------- car constants
GGL_HD V3 car_half_extents() { return v3(rl::HITBOX_X * kUU2BT / 2, rl::HITBOX_Y * kUU2BT / 2, rl::HITBOX_Z * kUU2BT / 2); }
GGL_HD V3 car_hitbox_offset() { return v3(rl::HITBOX_OFF_X * kUU2BT, rl::HITBOX_OFF_Y * kUU2BT, rl::HITBOX_OFF_Z * kUU2BT); }
// btBoxShape::calculateLocalInertia on the half extents WITH margin, then 1/x (btRigidBody::setMassProps). The reference box
// stores (he - 0.04) and later shrinks its margin to 10% of the smallest half extent (setSafeMargin), so the extents with
// margin are slightly smaller than the nominal hitbox.
GGL_HD V3 car_box_full_extents() {
    V3 he = car_half_extents();
    V3 core = he - v3(rl::CONVEX_MARGIN, rl::CONVEX_MARGIN, rl::CONVEX_MARGIN);
    float mn = fminf(he.x, fminf(he.y, he.z));
    float safe = 0.1f * mn;
    float margin = safe < rl::CONVEX_MARGIN ? safe : rl::CONVEX_MARGIN;
    return core + v3(margin, margin, margin);
}
GGL_HD V3 car_inv_inertia_local() {
    V3 he = car_box_full_extents();
    float lx = 2.f * he.x, ly = 2.f * he.y, lz = 2.f * he.z;
    V3 in = v3(rl::CAR_MASS / 12.f * (ly * ly + lz * lz), rl::CAR_MASS / 12.f * (lx * lx + lz * lz), rl::CAR_MASS / 12.f * (lx * lx + ly * ly));
    return v3(1.f / in.x, 1.f / in.y, 1.f / in.z);
}
GGL_HD void car_update_inertia(CarDev& c) {
    c.invInertiaW = mul(scaled(c.basis, car_inv_inertia_local()), transpose(c.basis));
}
// World AABB of the car's compound shape (btCompoundShape::getAabb over the child box's local AABB, no compound margin)
GGL_HD void car_world_aabb(V3 pos, const M3& basis, V3& mn, V3& mx) {
    const V3 he = car_box_full_extents(), off = car_hitbox_offset();
    const V3 lmin = off - he, lmax = off + he;
    const V3 lhe = (lmax - lmin) * 0.5f, lc = (lmax + lmin) * 0.5f;
    const V3 centre = pos + mul(basis, lc);
    const V3 ext = v3(dot(vabs(basis.r0), lhe), dot(vabs(basis.r1), lhe), dot(vabs(basis.r2), lhe));
    mn = centre - ext; mx = centre + ext;
}
// The broadphase AABB a suspension ray of the NEXT tick is tested against (btSingleRayCallback::process,
// btCollisionWorld.cpp:913-945): btCollisionWorld::updateSingleAabb (:135-175) runs inside this tick's collision detection, on
// the transform the tick started with, grown by the contact breaking threshold 0.02 and - continuous collision detection
// being on - merged with the AABB at the predicted transform (btDiscreteDynamicsWorld::predictUnconstraintMotion). Nobody
// refreshes it before the next tick's rays. Only kept when another car is close enough to be reached by a ray next tick
// (6 bt = 300 uu between centres; cars close at most 92 uu per tick); otherwise the box is left empty.
GGL_HD void car_update_ray_aabb(ArenaDev& A, int ci) {
    CarDev& c = A.cars[ci];
    bool near = false;
    for (int j = 0; j < A.numCars; j++)
        if (j != ci && len2(A.cars[j].pos - c.pos) < 36.f) near = true;
    if (!near) { c.aabbMin = v3(1e30f, 1e30f, 1e30f); c.aabbMax = v3(-1e30f, -1e30f, -1e30f); return; }
    const V3 thr = v3(0.02f, 0.02f, 0.02f);
    V3 mn, mx, mn2, mx2, np; M3 nb;
    car_world_aabb(c.pos, c.basis, mn, mx);
    integrate_transform(c.pos, c.basis, c.vel, c.angVel, kTick, np, nb);
    car_world_aabb(np, nb, mn2, mx2);
    c.aabbMin = vmin(mn - thr, mn2 - thr); c.aabbMax = vmax(mx + thr, mx2 + thr);
}
// What Car::GetState reports as rotMat: the reference refreshes _internalState.rotMat at the start of _PostTickUpdate and
// returns early from it while the car is demolished (Car.cpp:133-140), so from the tick of the demolition until the respawn
// the reported orientation is the one the car had when that tick began, while the rigid body is still solved and moved in
// that tick. The frozen matrix is parked in wheel 0's per-tick scratch, which nobody touches while isDemoed is set.
GGL_HD void car_freeze_reported_basis(CarDev& c) { c.w[0].hardPoint = c.basis.r0; c.w[0].contactPoint = c.basis.r1; c.w[0].contactNormal = c.basis.r2; }
GGL_HD M3 car_reported_basis(const CarDev& c) {
    if (!c.isDemoed) return c.basis;
    M3 m; m.r0 = c.w[0].hardPoint; m.r1 = c.w[0].contactPoint; m.r2 = c.w[0].contactNormal;
    return m;
}
GGL_HD V3 car_fwd(const CarDev& c) { return col(c.basis, 0); }
GGL_HD V3 car_right(const CarDev& c) { return col(c.basis, 1); }
GGL_HD V3 car_up(const CarDev& c) { return col(c.basis, 2); }
GGL_HD V3 vel_at(const V3& vel, const V3& angVel, V3 rel) { return vel + cross(angVel, rel); }
GGL_HD void car_apply_central_impulse(CarDev& c, V3 imp) { c.vel += imp * (1.f / rl::CAR_MASS); }
GGL_HD void car_apply_impulse(CarDev& c, V3 imp, V3 rel) {
    car_apply_central_impulse(c, imp);
    c.angVel += mul(c.invInertiaW, cross(rel, imp));
}
// btRigidBody::computeImpulseDenominator
GGL_HD float impulse_denominator(float invMass, const M3& invInertiaW, V3 r0, V3 n) {
    V3 c0 = cross(r0, n);
    V3 vec = cross(tmul(invInertiaW, c0), r0);
    return invMass + dot(n, vec);
}

struct WheelGeom {
    V3 conn;            // chassis connection point (car space, bt)
    float restLen, radius, forceScale;
};
GGL_HD WheelGeom wheel_geom(int i) {
    const bool front = i < 2;
    const bool left = (i % 2) != 0;
    WheelGeom g;
    float x = front ? rl::FRONT_WHEEL_X : rl::BACK_WHEEL_X;
    float y = front ? rl::FRONT_WHEEL_Y : rl::BACK_WHEEL_Y;
    float z = front ? rl::FRONT_WHEEL_Z : rl::BACK_WHEEL_Z;
    if (left) y *= -1;
    g.conn = v3(x, y, z) * kUU2BT;
    float rest = front ? rl::FRONT_SUS_REST : rl::BACK_SUS_REST;
    rest -= rl::MAX_SUSP_TRAVEL;
    g.restLen = rest * kUU2BT;
    g.radius = (front ? rl::FRONT_WHEEL_RAD : rl::BACK_WHEEL_RAD) * kUU2BT;
    g.forceScale = front ? rl::SUSP_FORCE_SCALE_FRONT : rl::SUSP_FORCE_SCALE_BACK;
    return g;
}
constexpr float kSuspTravel = ((rl::MAX_SUSP_TRAVEL * kUU2BT) * 100) / 100;

// ---------------------------------------------------------------------------------------------- ray casting
struct RayHit {
    float t;
    V3 normal;
    int ground;   // GroundKind
};

GGL_HD void ray_vs_plane(V3 from, V3 to, V3 n, float c, RayHit& h) {
    float da = dot(n, from) - c;
    float db = dot(n, to) - c;
    if (da * db >= 0.f) return;
    float t = da / (da - db);
    if (t < h.t) {
        h.t = t;
        h.normal = (da <= 0.f) ? -n : n;
        h.ground = GROUND_WORLD;
    }
}

GGL_HD void ray_vs_tri(V3 from, V3 to, const Tri& tr, RayHit& h) {
    const V3 tn = tr.rayN;
    const float dist = tr.rayD;
    float da = dot(tn, from) - dist;
    float db = dot(tn, to) - dist;
    if (da * db >= 0.f) return;
    float t = da / (da - db);
    if (t < h.t) {
        float tol = len2(tn) * -0.0001f;
        float s = 1.f - t;
        V3 p = v3(s * from.x + t * to.x, s * from.y + t * to.y, s * from.z + t * to.z);
        V3 v0p = tr.v0 - p, v1p = tr.v1 - p;
        if (dot(cross(v0p, v1p), tn) >= tol) {
            V3 v2p = tr.v2 - p;
            if (dot(cross(v1p, v2p), tn) >= tol && dot(cross(v2p, v0p), tn) >= tol) {
                V3 nn = normalized(tn);
                h.t = t;
                h.normal = (da <= 0.f) ? -nn : nn;
                h.ground = GROUND_WORLD;
            }
        }
    }
}

// btSubsimplexConvexCast::calcTimeOfImpact (BT/BulletCollision/NarrowPhaseCollision/btSubSimplexConvexCast.cpp:31-160) for the
// zero-radius point shape swept from -> to against a resting convex shape B given by its world-space support mapping: what
// btCollisionWorld::rayTestSingleInternal (btCollisionWorld.cpp:267-312) runs for every convex shape a suspension ray meets
// (the ball's sphere, the box child of a car's compound). The hit normal is the simplex's closest vector at the last advance
// of lambda, NOT the geometric face normal, and the simplex is never re-centred when lambda moves, so the iteration is
// restated step by step (32 iterations, epsilon 1e-4 on the squared distance, btConvexCast.h:28-29).
template <class SupB>
GGL_HD bool subsimplex_point_cast(V3 from, V3 to, SupB supB, float& fraction, V3& normal) {
    Voronoi S;
    vs_reset(S);
    const V3 r = to - from;
    const V3 originB = supB.origin;
    float lambda = 0.f;
    V3 ipA = from;
    V3 supA = from;
    V3 sB = supB(r);
    V3 v = supA - sB;
    int maxIter = 32;
    V3 n = v3(0, 0, 0);
    float dist2 = len2(v);
    while (dist2 > 0.0001f && maxIter--) {
        supA = ipA;
        sB = supB(v);
        V3 w = supA - sB;
        const float VdotW = dot(v, w);
        if (lambda > 1.f) return false;
        if (VdotW > 0.f) {
            const float VdotR = dot(v, r);
            if (VdotR >= -(kEps * kEps)) return false;
            lambda = lambda - VdotW / VdotR;
            const float s1 = 1.f - lambda;          // btVector3::setInterpolate3, applied to BOTH transforms: B does not
            ipA = v3(s1 * from.x + lambda * to.x, s1 * from.y + lambda * to.y, s1 * from.z + lambda * to.z);   // move, but
            supB.origin = v3(s1 * originB.x + lambda * originB.x, s1 * originB.y + lambda * originB.y,         // s*o + l*o
                             s1 * originB.z + lambda * originB.z);                                             // rounds
            w = supA - sB;
            n = v;
        }
        if (!vs_in_simplex(S, w)) vs_add(S, w, supA, sB);
        vs_update(S);
        v = S.cachedV;
        dist2 = S.cachedValid ? len2(v) : 0.f;
    }
    fraction = lambda;
    if (len2(n) >= kEps * kEps) normal = normalized(n);
    else normal = v3(0, 0, 0);
    if (dot(normal, r) >= -0.f) return false;
    return true;
}

struct SupSphere {
    V3 origin; float radius;
    GGL_HD V3 operator()(V3 d) const {       // btSphereShape::localGetSupportingVertex (btSphereShape.cpp:36-48); identity basis (the
        V3 nrm;                              // ball never rotates: btRigidBody m_noRot)
        if (len2(d) < kEps * kEps) nrm = normalized(v3(-1.f, -1.f, -1.f));
        else nrm = normalized(d);
        return radius * nrm + origin;
    }
};
struct SupBox {
    V3 origin; M3 basis; V3 he;
    GGL_HD V3 operator()(V3 d) const {       // btBoxShape::localGetSupportingVertex on the child's world transform
        return mul(basis, box_support(he, tmul(basis, d))) + origin;
    }
};

// out of line: a wheel ray meets a ball or another car's AABB in well under 1 % of the ticks, and the iteration is ~4 K
// instructions that would otherwise sit in the middle of the per-tick path of an instruction-fetch bound kernel
GGL_HDN bool cast_point_vs_sphere(V3 from, V3 to, V3 center, float radius, float& fraction, V3& normal) {
    SupSphere sup; sup.origin = center; sup.radius = radius;
    return subsimplex_point_cast(from, to, sup, fraction, normal);
}
GGL_HDN bool cast_point_vs_box(V3 from, V3 to, V3 origin, const M3& basis, V3 he, float& fraction, V3& normal) {
    SupBox sup; sup.basis = basis; sup.origin = origin; sup.he = he;
    return subsimplex_point_cast(from, to, sup, fraction, normal);
}

// rayTestSingleInternal's bookkeeping around the cast (btCollisionWorld.cpp:288-311) + ClosestRayResultCallback
GGL_HD void ray_take_cast_hit(bool hit, float fraction, V3 normal, int kind, RayHit& h) {
    if (!hit) return;
    if (!(len2(normal) > 0.0001f)) return;
    if (fraction < h.t) {
        h.t = fraction;
        h.normal = normalized(normal);
        h.ground = kind;
    }
}

GGL_HD void ray_vs_sphere(V3 from, V3 to, V3 center, float radius, int kind, RayHit& h) {
    // cull (conservative): the cast stops within 0.01 of the surface, so a segment that stays 0.02 clear of it cannot hit
    {
        V3 d = to - from, m = from - center;
        const float rr = radius + 0.02f;
        float a = dot(d, d), b = dot(m, d), c = dot(m, m) - rr * rr;
        if (a <= 0.f) return;
        if (c > 0.f && b > 0.f) return;
        if (c > 0.f && b * b - a * c < 0.f) return;
        if (c > 0.f && -b > a + sqrtf(b * b - a * c)) return;        // first crossing beyond the end of the segment
    }
    float fr; V3 nrm;
    const bool hit = cast_point_vs_sphere(from, to, center, radius, fr, nrm);
    ray_take_cast_hit(hit, fr, nrm, kind, h);
}

GGL_HD void ray_vs_car_box(V3 from, V3 to, const CarDev& oc, int kind, RayHit& h) {
    // The reference casts against every car whose broadphase AABB the segment crosses (btSingleRayCallback::process,
    // btCollisionWorld.cpp:913-945; the car's compound has no AABB tree, Car.cpp:198). The cast itself can report a hit for a
    // segment that misses the box (a degenerate simplex ends the iteration early), so this test must be the reference's:
    // the AABB of the LAST tick's collision detection (car_update_ray_aabb). A state that was just injected has none yet
    // (all zero): the AABB of the current transform stands in.
    {
        if (oc.aabbMin.x > oc.aabbMax.x) return;        // the usual case: no other car was within reach last tick (car_update_ray_aabb)
        V3 mn = oc.aabbMin, mx = oc.aabbMax;
        if (is_zero(mn) && is_zero(mx)) {
            car_world_aabb(oc.pos, oc.basis, mn, mx);
            mn = mn - v3(0.02f, 0.02f, 0.02f); mx = mx + v3(0.02f, 0.02f, 0.02f);
        }
        if (mn.x > mx.x) return;
        // the usual case first: the segment's own box misses the car's (then the slab test below fails on that axis too)
        if (fminf(from.x, to.x) > mx.x || fmaxf(from.x, to.x) < mn.x || fminf(from.y, to.y) > mx.y || fmaxf(from.y, to.y) < mn.y ||
            fminf(from.z, to.z) > mx.z || fmaxf(from.z, to.z) < mn.z) return;
        const V3 d = to - from;
        float t0 = 0.f, t1 = 1.f;
        for (int i = 0; i < 3; i++) {
            const float oi = comp(from, i), di = comp(d, i), lo = comp(mn, i), hi = comp(mx, i);
            if (fabsf(di) < 1e-12f) {
                if (oi < lo || oi > hi) return;
            } else {
                const float inv = 1.f / di;
                float ta = (lo - oi) * inv, tb = (hi - oi) * inv;
                if (ta > tb) { const float tmp = ta; ta = tb; tb = tmp; }
                if (ta > t0) t0 = ta;
                if (tb < t1) t1 = tb;
                if (t0 > t1) return;
            }
        }
    }
    float fr; V3 nrm;
    const bool hit = cast_point_vs_box(from, to, oc.pos + mul(oc.basis, car_hitbox_offset()), oc.basis, car_box_full_extents(), fr, nrm);
    ray_take_cast_hit(hit, fr, nrm, kind, h);
}

GGL_HD int grid_coord(float v, float mn, float inv, int n) {
    int c = (int)floorf((v - mn) * inv);
    return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

// Closest hit of the segment from->to against the arena (4 planes + mesh), the ball and the other cars.
GGL_HD void ray_cast_world(const ArenaDev& A, const WorldDev& W, int selfCar, V3 from, V3 to, RayHit& h) {
    h.t = 1.f; h.ground = GROUND_NONE; h.normal = v3(0, 0, 0);
    const float ex = rl::ARENA_EXTENT_X * kUU2BT, hz = rl::ARENA_HEIGHT * kUU2BT;
    // static planes (Arena.cpp:1060-1101): floor, ceiling, -x wall, +x wall
    ray_vs_plane(from, to, v3(0, 0, 1), 0.f, h);
    ray_vs_plane(from - v3(0, 0, hz), to - v3(0, 0, hz), v3(0, 0, -1), 0.f, h);
    ray_vs_plane(from - v3(-ex, 0, hz / 2), to - v3(-ex, 0, hz / 2), v3(1, 0, 0), 0.f, h);
    ray_vs_plane(from - v3(ex, 0, hz / 2), to - v3(ex, 0, hz / 2), v3(-1, 0, 0), 0.f, h);
    // mesh triangles through the uniform grid
    V3 rmn = vmin(from, to), rmx = vmax(from, to);
    if (W.numTris > 0 && !world_aabb_is_free(W, rmn, rmx)) {
        V3 mn = rmn, mx = rmx;
        int x0 = grid_coord(mn.x, W.gridMin.x, W.invCellSize, W.gx), x1 = grid_coord(mx.x, W.gridMin.x, W.invCellSize, W.gx);
        int y0 = grid_coord(mn.y, W.gridMin.y, W.invCellSize, W.gy), y1 = grid_coord(mx.y, W.gridMin.y, W.invCellSize, W.gy);
        int z0 = grid_coord(mn.z, W.gridMin.z, W.invCellSize, W.gz), z1 = grid_coord(mx.z, W.gridMin.z, W.invCellSize, W.gz);
        for (int z = z0; z <= z1; z++)
            for (int y = y0; y <= y1; y++)
                for (int x = x0; x <= x1; x++) {
                    int cell = (z * W.gy + y) * W.gx + x;
                    for (int k = W.cellStart[cell]; k < W.cellStart[cell + 1]; k++) ray_vs_tri(from, to, W.tris[W.cellTris[k]], h);
                }
    }
    // dynamic bodies
    ray_vs_sphere(from, to, A.ball.pos, rl::BALL_RADIUS * kUU2BT, GROUND_BALL, h);
    for (uint32_t m = A.rayBoxMask & ~(1u << selfCar); m; m &= m - 1) {      // cars with a tracked box, in index order (usually none)
#if defined(__CUDA_ARCH__)
        const int c = __ffs((int)m) - 1;
#else
        const int c = __builtin_ctz(m);
#endif
        ray_vs_car_box(from, to, A.cars[c], GROUND_CAR0 + c, h);
    }
    // a hit on a body without contact response (a demolished car) reports nothing at all
    // (btDefaultVehicleRaycaster::castRay, BulletDynamics/Vehicle/btDefaultVehicleRaycaster.cpp)
    if (h.ground >= GROUND_CAR0 && A.cars[h.ground - GROUND_CAR0].isDemoed) h.ground = GROUND_NONE;
}

// ---------------------------------------------------------------------------------------------- vehicle
// updateVehicleFirst for ONE wheel: ray cast + suspension contact info (btVehicleRL.cpp:64-92,118-212), then that wheel's
// part of calcFrictionImpulses (:313-387). Wheels are independent of each other here, so the 4 x P wheel tasks of an arena
// are spread over the lanes of the group.
GGL_HD void wheel_friction(ArenaDev& A, int ci, int i, bool otherCurrent);
GGL_HD void wheel_first(ArenaDev& A, const WorldDev& W, int ci, int i) {
    CarDev& c = A.cars[ci];
    const V3 up = car_up(c), right = car_right(c);
    {
        Wheel& w = c.w[i];
        const WheelGeom g = wheel_geom(i);
        // wheel frame with LAST tick's steer angle
        if (w.steerAngle == 0.f) {
            w.latDir = right;           // the rotation by angle 0 is the identity matrix (up to the sign of zeros): rear wheels always
        } else {
            Q4 q = quat_axis_angle(up, w.steerAngle);
            w.latDir = mul(m3_from_quat(q), right);
        }
        w.hardPoint = c.pos + mul(c.basis, g.conn);
        const V3 wheelDir = -up;
        const float rayLen = g.restLen + kSuspTravel + g.radius - rl::SUSP_SUBTRACTION;
        const V3 source = w.hardPoint;
        const V3 target = source + wheelDir * rayLen;
        w.contactPoint = target;
        RayHit h;
        ray_cast_world(A, W, ci, source, target, h);
        w.ground = h.ground;
        if (h.ground != GROUND_NONE) {
            float s = 1.f - h.t;
            w.contactPoint = v3(s * source.x + h.t * target.x, s * source.y + h.t * target.y, s * source.z + h.t * target.z);
            w.contactNormal = normalized(h.normal);
            float traceLen = dot(w.hardPoint - w.contactPoint, up);
            w.suspLen = fclamp(traceLen - g.radius, g.restLen - kSuspTravel, g.restLen + kSuspTravel);
            float denom = dot(w.contactNormal, up);
            V3 relpos = w.contactPoint - c.pos;
            V3 velAt = vel_at(c.vel, c.angVel, relpos);
            float projVel = dot(w.contactNormal, velAt);
            if (denom > 0.1f) {
                float inv = 1.f / denom;
                w.suspRelVel = projVel * inv;
                w.clippedInv = inv;
            } else {
                w.suspRelVel = 0.f;
                w.clippedInv = 10.f;
            }
            if (h.ground == GROUND_WORLD) {
                float thresh = (g.restLen + g.radius) - rl::SUSP_SUBTRACTION;
                if (traceLen < thresh) {
                    float delta = traceLen - thresh;
                    // resolveSingleCollision(chassis, static body, hit, normal, distance = delta, applyImpulses = false)
                    float relVel = dot(w.contactNormal, velAt);
                    float posErr = 0.2f * -delta / kTick;
                    float velErr = -(1.0f + 0.f) * relVel;
                    float denom0 = impulse_denominator(1.f / rl::CAR_MASS, c.invInertiaW, w.contactPoint - c.pos, w.contactNormal);
                    float jacInv = 1.f / (denom0 + 0.f);
                    float imp = posErr * jacInv + velErr * jacInv;
                    imp = 0.f > imp ? 0.f : imp;
                    w.extraPushback = imp / 4;
                }
            }
        } else {
            w.suspLen = g.restLen + kSuspTravel;
            w.suspRelVel = 0.f;
            w.contactNormal = -wheelDir;
            w.clippedInv = 1.f;
            w.extraPushback = 0.f;
        }
    }
    // A wheel standing on ANOTHER CAR reads that car's velocity in its friction impulse. The reference pre-ticks the cars one
    // after the other in its car-set order (Arena.cpp:752-764 = players[] slot order), so a car that comes LATER sees the
    // velocity the earlier car has AFTER its own pre-tick (wheel and jump impulses are applied to the body at once,
    // btVehicleRL.cpp:390-402). Those wheels are flagged here and finished in slot order by the tick (sim_tick.cuh); every
    // other wheel (world, ball, a car that comes later in the order) is independent of the other cars' pre-ticks.
    if (c.w[i].ground >= GROUND_CAR0) {
        const int o = c.w[i].ground - GROUND_CAR0;
        int sc = 0, so = 0;
        for (int s = 0; s < A.numCars; s++) { if (A.slotCar[s] == ci) sc = s; if (A.slotCar[s] == o) so = s; }
        if (so < sc) {
#if defined(__CUDA_ARCH__)
            atomicOr(&A.wheelDefer, 1u << (ci * 4 + i));
#else
            A.wheelDefer |= 1u << (ci * 4 + i);
#endif
            return;
        }
    }
    wheel_friction(A, ci, i, false);
}

// one wheel's part of calcFrictionImpulses (btVehicleRL.cpp:313-387) with LAST tick's engine force / brake / friction factors;
// `otherCurrent`: a car the wheel stands on is read with its CURRENT velocity (it was pre-ticked before this car) instead
// of the velocity it had when the tick began
GGL_HD void wheel_friction(ArenaDev& A, int ci, int i, bool otherCurrent) {
    CarDev& c = A.cars[ci];
    const float frictionScale = rl::CAR_MASS / 3;
    {
        Wheel& w = c.w[i];
        if (w.ground == GROUND_NONE) { w.impulse = v3(0, 0, 0); return; }
        V3 axle = w.latDir;
        V3 n = w.contactNormal;
        float proj = dot(axle, n);
        axle -= n * proj;
        axle = safe_normalized(axle);
        V3 fwd = safe_normalized(cross(n, axle));
        // the body the wheel stands on
        V3 gPos = v3(0, 0, 0), gVel = v3(0, 0, 0), gAng = v3(0, 0, 0);
        float gInvMass = 0.f; M3 gInvI; gInvI.r0 = gInvI.r1 = gInvI.r2 = v3(0, 0, 0);
        if (w.ground == GROUND_BALL) {
            gPos = A.ball.pos; gVel = A.ball.vel; gAng = A.ball.angVel; gInvMass = 1.f / rl::BALL_MASS;
            float r = rl::BALL_RADIUS * kUU2BT;
            float ii = 1.f / (0.4f * rl::BALL_MASS * r * r);
            gInvI.r0 = v3(ii, 0, 0); gInvI.r1 = v3(0, ii, 0); gInvI.r2 = v3(0, 0, ii);
        } else if (w.ground >= GROUND_CAR0) {
            const CarDev& oc = A.cars[w.ground - GROUND_CAR0];
            gPos = oc.pos; gVel = otherCurrent ? oc.vel : oc.velSnap; gAng = otherCurrent ? oc.angVel : oc.angVelSnap; gInvMass = 1.f / rl::CAR_MASS; gInvI = oc.invInertiaW;
        }
        // resolveSingleBilateral(chassis, cp, ground, cp, 0, axle) -> side impulse
        V3 rel1 = w.contactPoint - c.pos, rel2 = w.contactPoint - gPos;
        V3 v1 = vel_at(c.vel, c.angVel, rel1), v2 = vel_at(gVel, gAng, rel2);
        V3 vrel = v1 - v2;
        float diag = impulse_denominator(1.f / rl::CAR_MASS, c.invInertiaW, rel1, axle) + impulse_denominator(gInvMass, gInvI, rel2, axle);
        float sideImpulse = -0.2f * dot(axle, vrel) * (1.f / diag);
        float rolling;
        if (w.engineForce == 0.f) {
            if (w.brake != 0.f) {
                V3 carRel = w.contactPoint - c.pos;
                V3 cv = vel_at(c.vel, c.angVel, carRel) - vel_at(gVel, gAng, carRel);
                float rv = dot(cv, fwd);
                rolling = fclamp(-rv * 113.73963f, -w.brake, w.brake);
            } else {
                rolling = 0.f;
            }
        } else {
            rolling = -w.engineForce / frictionScale;
        }
        V3 total = (fwd * rolling * w.longFriction) + (axle * sideImpulse * w.latFriction);
        w.impulse = total * frictionScale;
    }
}

GGL_HD V3 upwards_dir_from_wheels(const CarDev& c) {
    V3 sum = v3(0, 0, 0);
    for (int i = 0; i < 4; i++)
        if (c.w[i].ground != GROUND_NONE) sum += c.w[i].contactNormal;
    if (is_zero(sum)) return car_up(c);
    return safe_normalized(sum);
}

GGL_HD void vehicle_second(CarDev& c) {
    float force[4];
    for (int i = 0; i < 4; i++) {
        Wheel& w = c.w[i];
        const WheelGeom g = wheel_geom(i);
        if (w.ground != GROUND_NONE) {
            float f = (g.restLen - w.suspLen) * rl::SUSP_STIFFNESS * w.clippedInv;
            float damp = (w.suspRelVel < 0.f) ? rl::DAMP_COMPRESSION : rl::DAMP_RELAXATION;
            f = f - (damp * w.suspRelVel);
            f *= g.forceScale;
            if (f < 0.f) f = 0.f;
            force[i] = f;
        } else {
            force[i] = 0.f;
        }
        w.suspForce = force[i];
    }
    for (int i = 0; i < 4; i++) {
        Wheel& w = c.w[i];
        if (force[i] != 0.f) {
            V3 off = w.contactPoint - c.pos;
            float scale = (force[i] * kTick) + w.extraPushback;
            car_apply_impulse(c, w.contactNormal * scale, off);
        }
    }
    // applyFrictionImpulses
    V3 up = car_up(c);
    for (int i = 0; i < 4; i++) {
        Wheel& w = c.w[i];
        if (!is_zero(w.impulse)) {
            V3 off = w.contactPoint - c.pos;
            float upDot = dot(up, off);
            V3 rel = off - up * upDot;
            car_apply_impulse(c, w.impulse * kTick, rel);
        }
    }
}

// ---------------------------------------------------------------------------------------------- car logic
GGL_HD void update_wheels(CarDev& c, int numWheels, float fwdSpeedUU) {
    const float absSpeed = fabsf(fwdSpeedUU);
    bool worldContact = false;
    for (int i = 0; i < 4; i++) worldContact |= (c.w[i].ground == GROUND_WORLD);
    if (c.controls.handbrake) c.handbrakeVal += rl::POWERSLIDE_RISE_RATE * kTick;
    else c.handbrakeVal -= rl::POWERSLIDE_FALL_RATE * kTick;
    c.handbrakeVal = fclamp(c.handbrakeVal, 0.f, 1.f);

    float realThrottle = c.controls.throttle;
    float realBrake = 0.f;
    if (c.controls.boost && c.boost > 0.f) realThrottle = 1.f;
    {
        float driveScale = curve_drive_torque(absSpeed);
        float engineThrottle = realThrottle;
        if (!c.controls.handbrake) {
            float absThrottle = fabsf(realThrottle);
            if (absThrottle >= rl::THROTTLE_DEADZONE) {
                if (absSpeed > rl::STOPPING_FORWARD_VEL && fsgn(realThrottle) != fsgn(fwdSpeedUU)) {
                    realBrake = 1.f;
                    if (absSpeed > rl::BRAKING_NO_THROTTLE_SPEED_THRESH) engineThrottle = 0.f;
                }
            } else {
                engineThrottle = 0.f;
                realBrake = (absSpeed < rl::STOPPING_FORWARD_VEL) ? 1.f : rl::COASTING_BRAKE_FACTOR;
            }
        }
        if (numWheels < 3) driveScale /= 4;
        float engineForce = engineThrottle * (rl::THROTTLE_TORQUE_AMOUNT * kUU2BT) * driveScale;
        float brakeForce = realBrake * (rl::BRAKE_TORQUE_AMOUNT * kUU2BT);
        for (int i = 0; i < 4; i++) { c.w[i].engineForce = engineForce; c.w[i].brake = brakeForce; }
    }
    {
        float steer = curve_steer_angle(absSpeed);
        if (c.handbrakeVal != 0.f) steer += (curve_powerslide_steer(absSpeed) - steer) * c.handbrakeVal;
        steer *= c.controls.steer;
        c.w[0].steerAngle = steer;
        c.w[1].steerAngle = steer;
    }
    for (int i = 0; i < 4; i++) {
        Wheel& w = c.w[i];
        if (w.ground == GROUND_NONE) continue;
        V3 latDir = w.latDir;
        V3 longDir = cross(latDir, w.contactNormal);
        float input = 0.f;
        V3 delta = w.hardPoint - c.pos;
        V3 cv = (cross(c.angVel, delta) + c.vel) * kBT2UU;
        float base = fabsf(dot(cv, latDir));
        if (base > 5.f) input = base / (fabsf(dot(cv, longDir)) + base);
        float lat = curve_lat_friction(input);
        float lng = 1.f;   // LONG_FRICTION_CURVE is empty -> default output 1
        if (c.handbrakeVal != 0.f) {
            float hb = c.handbrakeVal;
            lat *= (0.1f - 1.f) * hb + 1.f;     // HANDBRAKE_LAT_FRICTION_FACTOR_CURVE is the constant 0.1
            lng *= (curve_handbrake_long(input) - 1.f) * hb + 1.f;
        } else {
            lng = 1.f;
        }
        if (realThrottle == 0.f) {
            float ns = curve_non_sticky(w.contactNormal.z);
            lat *= ns;
            lng *= ns;
        }
        w.latFriction = lat;
        w.longFriction = lng;
    }
    if (worldContact) {
        V3 upDir = upwards_dir_from_wheels(c);
        bool fullStick = (realThrottle != 0.f) || (absSpeed > rl::STOPPING_FORWARD_VEL);
        float stick = 0.5f;
        if (fullStick) stick += 1.f - fabsf(upDir.z);
        c.force += upDir * stick * (rl::GRAVITY_Z * kUU2BT) * rl::CAR_MASS;
    }
}

GGL_HD void update_boost(CarDev& c, float usedPerSecond /* MutatorConfig::boostUsedPerSecond: 0 in heatseeker */) {
    if (c.timeSpentBoosting > 0.f) {
        if (!c.controls.boost && c.timeSpentBoosting >= rl::BOOST_MIN_TIME) c.timeSpentBoosting = 0.f;
        else c.timeSpentBoosting += kTick;
    } else if (c.controls.boost) {
        c.timeSpentBoosting = kTick;
    }
    if (c.boost > 0.f && c.timeSpentBoosting > 0.f) {
        c.boost = fmaxf(c.boost - usedPerSecond * kTick, 0.f);
        float accel = (c.isOnGround ? rl::BOOST_ACCEL_GROUND : rl::BOOST_ACCEL_AIR) * kUU2BT;
        c.force += accel * car_fwd(c) * rl::CAR_MASS;
    }
    c.boost = fminf(c.boost, rl::BOOST_MAX);
}

GGL_HD void update_jump(CarDev& c, bool jumpPressed) {
    if (c.isOnGround && !c.isJumping) {
        if (!(c.hasJumped && c.jumpTime < rl::JUMP_MIN_TIME + rl::JUMP_RESET_TIME_PAD)) {
            c.hasJumped = 0;
            c.jumpTime = 0.f;
        }
    }
    if (c.isJumping) {
        c.isJumping = (c.jumpTime < rl::JUMP_MIN_TIME || (c.controls.jump && c.jumpTime < rl::JUMP_MAX_TIME)) ? 1 : 0;
    } else if (c.isOnGround && jumpPressed) {
        c.isJumping = 1;
        c.jumpTime = 0.f;
        V3 f = car_up(c) * rl::JUMP_IMMEDIATE_FORCE * kUU2BT * rl::CAR_MASS;
        car_apply_central_impulse(c, f);
    }
    if (c.isJumping) {
        c.hasJumped = 1;
        V3 f = car_up(c) * rl::JUMP_ACCEL;
        if (c.jumpTime < rl::JUMP_MIN_TIME) f *= 0.62f;
        c.force += f * kUU2BT * rl::CAR_MASS;
    }
    if (c.isJumping || c.hasJumped) c.jumpTime += kTick;
}

GGL_HD void update_air_torque(CarDev& c, bool updateAirControl) {
    V3 dirPitch = -car_right(c), dirYaw = car_up(c), dirRoll = -car_fwd(c);
    bool doAir = false;
    if (c.isFlipping) c.isFlipping = (c.hasFlipped && c.flipTime < rl::FLIP_TORQUE_TIME) ? 1 : 0;
    if (c.isFlipping) {
        V3 rel = c.flipRelTorque;
        if (!is_zero(c.flipRelTorque)) {
            float pitchScale = 1.f;
            if (rel.y != 0.f && c.controls.pitch != 0.f) {
                if (fsgn(rel.y) == fsgn(c.controls.pitch)) {
                    pitchScale = 1.f - fminf(fabsf(c.controls.pitch), 1.f);
                    doAir = true;
                }
            }
            rel.y *= pitchScale;
            V3 dodge = rel * v3(rl::FLIP_TORQUE_X, rl::FLIP_TORQUE_Y, 0.f);
            M3 m = mul(inverse(c.invInertiaW), c.basis);
            c.torque += mul(m, dodge);
        } else {
            doAir = true;
        }
    } else {
        doAir = true;
    }
    doAir = doAir && !c.isAutoFlipping;
    doAir = doAir && updateAirControl;
    if (doAir) {
        float pitchTorqueScale = 1.f;
        V3 torque;
        if (c.controls.pitch != 0.f || c.controls.yaw != 0.f || c.controls.roll != 0.f) {
            if (c.isFlipping) pitchTorqueScale = 0.f;
            else if (c.hasFlipped && c.flipTime < rl::FLIP_TORQUE_TIME + rl::FLIP_PITCHLOCK_EXTRA_TIME) pitchTorqueScale = 0.f;
            torque = (c.controls.pitch * dirPitch * pitchTorqueScale * rl::AIR_TORQUE_X) +
                     (c.controls.yaw * dirYaw * rl::AIR_TORQUE_Y) +
                     (c.controls.roll * dirRoll * rl::AIR_TORQUE_Z);
        } else {
            torque = v3(0, 0, 0);
        }
        V3 av = c.angVel;
        float dampPitch = dot(dirPitch, av) * rl::AIR_DAMP_X * (1.f - fabsf(c.controls.pitch * pitchTorqueScale));
        float dampYaw = dot(dirYaw, av) * rl::AIR_DAMP_Y * (1.f - fabsf(c.controls.yaw));
        float dampRoll = dot(dirRoll, av) * rl::AIR_DAMP_Z;
        V3 damping = (dirYaw * dampYaw) + (dirPitch * dampPitch) + (dirRoll * dampRoll);
        c.torque += mul(inverse(c.invInertiaW), torque - damping) * rl::CAR_TORQUE_SCALE;
    }
    if (c.controls.throttle != 0.f) c.force += car_fwd(c) * c.controls.throttle * rl::THROTTLE_AIR_ACCEL * kUU2BT * rl::CAR_MASS;
}

GGL_HD void update_double_jump_or_flip(CarDev& c, bool jumpPressed, float fwdSpeedUU) {
    if (c.isOnGround) {
        c.hasDoubleJumped = 0; c.hasFlipped = 0;
        c.airTime = 0.f; c.airTimeSinceJump = 0.f; c.flipTime = 0.f;
    } else {
        c.airTime += kTick;
        if (c.hasJumped && !c.isJumping) c.airTimeSinceJump += kTick;
        else c.airTimeSinceJump = 0.f;
        if (jumpPressed && c.airTimeSinceJump < rl::DOUBLEJUMP_MAX_DELAY) {
            float mag = fabsf(c.controls.yaw) + fabsf(c.controls.pitch) + fabsf(c.controls.roll);
            bool isFlipInput = mag >= rl::DODGE_DEADZONE;
            bool canUse = (!c.hasDoubleJumped && !c.hasFlipped);
            if (c.isAutoFlipping) canUse = false;
            if (canUse) {
                if (isFlipInput) {
                    c.flipTime = 0.f; c.hasFlipped = 1; c.isFlipping = 1;
                    float ratio = fabsf(fwdSpeedUU) / rl::CAR_MAX_SPEED;
                    V3 dd = v3(-c.controls.pitch, c.controls.yaw + c.controls.roll, 0.f);
                    if (fabsf(c.controls.yaw + c.controls.roll) < 0.1f && fabsf(c.controls.pitch) < 0.1f) dd = v3(0, 0, 0);
                    else dd = safe_normalized(dd);
                    c.flipRelTorque = v3(-dd.y, dd.x, 0.f);
                    if (fabsf(dd.x) < 0.1f) dd.x = 0.f;
                    if (fabsf(dd.y) < 0.1f) dd.y = 0.f;
                    if (!(len2(dd) < kEps * kEps)) {
                        bool backwards;
                        if (fabsf(fwdSpeedUU) < 100.0f) backwards = dd.x < 0.0f;
                        else backwards = (dd.x >= 0.0f) != (fwdSpeedUU >= 0.0f);
                        V3 dv = dd * rl::FLIP_INITIAL_VEL_SCALE;
                        float maxScaleX = backwards ? rl::FLIP_BACKWARD_IMPULSE_MAX_SPEED_SCALE : rl::FLIP_FORWARD_IMPULSE_MAX_SPEED_SCALE;
                        dv.x *= ((maxScaleX - 1.f) * ratio) + 1.f;
                        dv.y *= ((rl::FLIP_SIDE_IMPULSE_MAX_SPEED_SCALE - 1.f) * ratio) + 1.f;
                        if (backwards) dv.x *= rl::FLIP_BACKWARD_IMPULSE_SCALE_X;
                        V3 f = car_fwd(c);
                        float ang = atan2f(f.y, f.x);
                        V3 xd = v3(cosf(ang), -sinf(ang), 0.f), yd = v3(sinf(ang), cosf(ang), 0.f);
                        V3 fin = v3(dot(dv, xd), dot(dv, yd), 0.f);
                        car_apply_central_impulse(c, fin * kUU2BT * rl::CAR_MASS);
                    }
                } else {
                    V3 f = car_up(c) * rl::JUMP_IMMEDIATE_FORCE * kUU2BT * rl::CAR_MASS;
                    car_apply_central_impulse(c, f);
                    c.hasDoubleJumped = 1;
                }
            }
        }
    }
    if (c.isFlipping) {
        c.flipTime += kTick;
        if (c.flipTime <= rl::FLIP_TORQUE_TIME) {
            if (c.flipTime >= rl::FLIP_Z_DAMP_START && (c.vel.z < 0.f || c.flipTime < rl::FLIP_Z_DAMP_END))
                c.vel.z *= powf(1.f - rl::FLIP_Z_DAMP_120, kTick / (1 / 120.f));
        }
    } else if (c.hasFlipped) {
        c.flipTime += kTick;
    }
}

GGL_HD void update_auto_flip(CarDev& c, bool jumpPressed) {
    if (jumpPressed && c.worldContactHas && c.worldContactNormal.z > rl::CAR_AUTOFLIP_NORMZ_THRESH) {
        float yaw, pitch, roll;
        euler_ypr_from_m3(c.basis, yaw, pitch, roll);
        roll *= -1.f;
        float absRoll = fabsf(roll);
        if (absRoll > rl::CAR_AUTOFLIP_ROLL_THRESH) {
            c.autoFlipTimer = rl::CAR_AUTOFLIP_TIME * (absRoll / kPi);
            c.autoFlipTorqueScale = (roll > 0.f) ? 1.f : -1.f;
            c.isAutoFlipping = 1;
            car_apply_central_impulse(c, -car_up(c) * rl::CAR_AUTOFLIP_IMPULSE * kUU2BT * rl::CAR_MASS);
        }
    }
    if (c.isAutoFlipping) {
        if (c.autoFlipTimer <= 0.f) {
            c.isAutoFlipping = 0;
            c.autoFlipTimer = 0.f;
        } else {
            c.angVel += car_fwd(c) * rl::CAR_AUTOFLIP_TORQUE * c.autoFlipTorqueScale * kTick;
            c.autoFlipTimer -= kTick;
        }
    }
}

GGL_HD void update_auto_roll(CarDev& c, int numWheels) {
    V3 groundUp = (numWheels > 0) ? upwards_dir_from_wheels(c) : c.worldContactNormal;
    V3 groundDown = -groundUp;
    V3 f = car_fwd(c), r = car_right(c);
    V3 crossRight = cross(groundUp, f);
    V3 crossFwd = cross(groundDown, crossRight);
    float rightFactor = 1.f - fclamp(dot(r, crossRight), 0.f, 1.f);
    float fwdFactor = 1.f - fclamp(dot(f, crossFwd), 0.f, 1.f);
    V3 torqueDirRight = f * (dot(r, groundUp) >= 0.f ? -1.f : 1.f);
    V3 torqueDirFwd = r * (dot(f, groundUp) >= 0.f ? 1.f : -1.f);
    V3 tr = torqueDirRight * rightFactor, tf = torqueDirFwd * fwdFactor;
    c.force += groundDown * rl::CAR_AUTOROLL_FORCE * kUU2BT * rl::CAR_MASS;
    c.torque += mul(inverse(c.invInertiaW), tf + tr) * rl::CAR_AUTOROLL_TORQUE;
}

// minstd_rand0 (libstdc++ std::default_random_engine): x <- 16807 x mod (2^31 - 1)
GGL_HD uint32_t minstd_next(uint32_t& s) {
    uint64_t p = (uint64_t)s * 16807ull;
    s = (uint32_t)(p % 2147483647ull);
    return s;
}

// The libstdc++ / RocketSim pieces the reference draws through, restated so that an arena's stream yields exactly what the
// reference's thread engine would from the same state (pinned against the compiled reference, tests/test_setters_cpu.py):
//  * std::uniform_int_distribution<unsigned long>{0, n-1} over minstd_rand0: the engine's range 2^31-3 is neither 2^32-1 nor
//    2^64-1, so bits/uniform_int_dist.h takes its "2 divisions" downscaling branch (scale, reject the tail, divide);
//  * std::shuffle (bits/stl_algo.h): one leading swap when the count is even, then two swap positions per draw;
//  * RocketSim::Math::RandFloat (Math.cpp:54-57): min + (engine() / (float)engine.max()) * (max - min).
GGL_HD uint32_t std_uniform_below(uint32_t& s, uint32_t n) {
    const uint32_t scaling = 2147483645u / n, past = n * scaling;
    uint32_t r;
    do r = minstd_next(s) - 1u; while (r >= past);
    return r / scaling;
}
// a permutation of up to 8 items, 4 bits each (item at position i = nibble i): no indexed local array, no stack
GGL_HD uint32_t nib_get(uint32_t p, int i) { return (p >> (4 * i)) & 15u; }
GGL_HD uint32_t nib_swap(uint32_t p, int i, int j) {
    const uint32_t a = nib_get(p, i), b = nib_get(p, j);
    p &= ~((15u << (4 * i)) | (15u << (4 * j)));
    return p | (b << (4 * i)) | (a << (4 * j));
}
GGL_HD uint32_t std_shuffle_nib(uint32_t& s, int n) {
    uint32_t p = 0x76543210u;
    if (n < 2) return p;
    int i = 1;
    if ((n & 1) == 0) { p = nib_swap(p, i, (int)std_uniform_below(s, 2u)); i++; }
    while (i < n) {
        const uint32_t sr = (uint32_t)i + 1u;
        const uint32_t x = std_uniform_below(s, sr * (sr + 1u));
        p = nib_swap(p, i, (int)(x / (sr + 1u))); i++;
        p = nib_swap(p, i, (int)(x % (sr + 1u))); i++;
    }
    return p;
}
GGL_HD float rand_float(uint32_t& s, float lo, float hi) { return lo + ((float)minstd_next(s) / 2147483648.0f) * (hi - lo); }

// Car::SetState(CarState()) part of Car::Respawn (Car.cpp:43-56, 20-36)
GGL_HD void car_reset_state_defaults(CarDev& c) {
    c.vel = v3(0, 0, 0); c.angVel = v3(0, 0, 0); c.velImpulseCache = v3(0, 0, 0);
    c.flipRelTorque = v3(0, 0, 0); c.worldContactNormal = v3(0, 0, 0);
    c.jumpTime = c.flipTime = c.airTime = c.airTimeSinceJump = 0.f;
    c.boost = rl::BOOST_SPAWN_AMOUNT; c.timeSpentBoosting = 0.f; c.supersonicTime = 0.f; c.handbrakeVal = 0.f;
    c.autoFlipTimer = 0.f; c.autoFlipTorqueScale = 0.f; c.carContactCooldown = 0.f; c.demoRespawnTimer = 0.f;
    c.hitRelPos = c.hitBallPos = c.hitExtraVel = v3(0, 0, 0);
    c.hitTick = ~0ull; c.hitTickExtra = ~0ull; c.updateCounter = 0;
    c.carContactOtherId = 0;
    c.isOnGround = 1; c.hasJumped = c.hasDoubleJumped = c.hasFlipped = c.isFlipping = c.isJumping = 0;
    c.isSupersonic = c.isAutoFlipping = c.worldContactHas = c.isDemoed = c.hitValid = 0;
    for (int i = 0; i < 4; i++) c.wheelsWithContact[i] = 0;
    Controls z; z.throttle = z.steer = z.pitch = z.yaw = z.roll = 0.f; z.jump = z.boost = z.handbrake = z._pad = 0;
    c.lastControls = z;
}

GGL_HD void car_respawn(ArenaDev& A, CarDev& c) {
    const float rx[4] = {-2304, -2688, 2304, 2688};
    int idx = (int)(minstd_next(A.rng) % 4u);
    car_reset_state_defaults(c);
    float sy = (c.team == 0) ? 1.f : -1.f;
    c.pos = v3(rx[idx], -4608.f * sy, rl::CAR_RESPAWN_Z) * kUU2BT;
    float yaw = (float)(3.14159265358979323846 / 2) + (c.team == 0 ? 0.f : kPi);
    c.basis = m3_from_euler_ypr(yaw, -0.f, -0.f);
    if (A.gameMode == kModeHeatseeker) c.boost = 100.f;         // MutatorConfig::carSpawnBoostAmount (MutatorConfig.cpp:29-33, Car.cpp:72)
    car_update_inertia(c);
    c.aabbMin = c.aabbMax = v3(0, 0, 0);
}

GGL_HD void clamp_controls(Controls& k) {
    k.throttle = fclamp(k.throttle, -1.f, 1.f); k.steer = fclamp(k.steer, -1.f, 1.f);
    k.pitch = fclamp(k.pitch, -1.f, 1.f); k.yaw = fclamp(k.yaw, -1.f, 1.f); k.roll = fclamp(k.roll, -1.f, 1.f);
}

// Car::_PreTickUpdate in three lane-group phases: (a) per car: controls, demolition timer / respawn; (b) per wheel:
// wheel_first; (c) per car: everything after updateVehicleFirst.
GGL_HD void car_pre_tick_a(ArenaDev& A, int ci) {
    CarDev& c = A.cars[ci];
    clamp_controls(c.controls);
    // the rigid-body flags follow isDemoed as it was on ENTRY: the tick a car respawns its body is still disabled
    c.rbDisabled = c.isDemoed;
    if (c.isDemoed) {
        c.demoRespawnTimer = fmaxf(c.demoRespawnTimer - kTick, 0.f);
        if (c.demoRespawnTimer == 0.f) car_respawn(A, c);
    }
}

GGL_HD void car_pre_tick_c(ArenaDev& A, int ci) {
    CarDev& c = A.cars[ci];
    if (c.isDemoed) return;
    bool jumpPressed = c.controls.jump && !c.lastControls.jump;
    int numWheels = 0;
    for (int i = 0; i < 4; i++) {
        uint8_t in = c.w[i].ground != GROUND_NONE;
        c.wheelsWithContact[i] = in;
        numWheels += in;
    }
    c.isOnGround = numWheels >= 3;
    float fwdSpeedUU = dot(c.vel, car_fwd(c)) * kBT2UU;
    update_wheels(c, numWheels, fwdSpeedUU);
    if (numWheels < 3) update_air_torque(c, numWheels == 0);
    else c.isFlipping = 0;
    update_jump(c, jumpPressed);
    update_auto_flip(c, jumpPressed);
    update_double_jump_or_flip(c, jumpPressed, fwdSpeedUU);
    if (c.controls.throttle != 0.f && ((numWheels > 0 && numWheels < 4) || c.worldContactHas)) update_auto_roll(c, numWheels);
    c.worldContactHas = 0;
    vehicle_second(c);
    update_boost(c, A.gameMode == kModeHeatseeker ? 0.f : rl::BOOST_USED_PER_SECOND);
}

GGL_HD void car_post_tick(CarDev& c) {
    if (c.isDemoed) return;
    // _PostTickUpdate
    float speed2 = len2(c.vel * kBT2UU);
    if (c.isSupersonic && c.supersonicTime < rl::SUPERSONIC_MAINTAIN_MAX_TIME)
        c.isSupersonic = speed2 >= rl::SUPERSONIC_MAINTAIN_MIN_SPEED * rl::SUPERSONIC_MAINTAIN_MIN_SPEED;
    else
        c.isSupersonic = speed2 >= rl::SUPERSONIC_START_SPEED * rl::SUPERSONIC_START_SPEED;
    if (c.isSupersonic) c.supersonicTime += kTick; else c.supersonicTime = 0.f;
    if (c.carContactCooldown > 0.f) c.carContactCooldown = fmaxf(c.carContactCooldown - kTick, 0.f);
    c.lastControls = c.controls;
    // _FinishPhysicsTick
    if (!is_zero(c.velImpulseCache)) {
        c.vel += c.velImpulseCache;
        c.velImpulseCache = v3(0, 0, 0);
    }
    const float maxV = rl::CAR_MAX_SPEED * kUU2BT;
    if (len2(c.vel) > maxV * maxV) c.vel = normalized(c.vel) * maxV;
    if (len2(c.angVel) > rl::CAR_MAX_ANG_SPEED * rl::CAR_MAX_ANG_SPEED) c.angVel = normalized(c.angVel) * rl::CAR_MAX_ANG_SPEED;
    c.updateCounter++;
}

}  // namespace ggl
