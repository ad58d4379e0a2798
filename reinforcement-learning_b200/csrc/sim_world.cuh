// The rigid-body part of one tick: gravity, damping, contact generation for the fixed scene
// "<= 6 boxes + 1 sphere + 4 planes + triangle meshes", the game-logic contact callbacks, a projected Gauss-Seidel
// solver with split-impulse penetration recovery, and transform integration.
//
// It is a fixed-function restatement of what the reference obtains from its vendored Bullet 3.24 (paths under
// GigaLearnCPP/RLGymCPP/RocketSim/libsrc/bullet3-3.24 unless noted):
//   BulletDynamics/Dynamics/btDiscreteDynamicsWorld.cpp:325-431,985-1044   step / predict / integrate
//   BulletCollision/CollisionDispatch/btConvexPlaneCollisionAlgorithm.cpp:92-131 convex vs plane (deepest support vertex)
//   BulletCollision/CollisionDispatch/SphereTriangleDetector.cpp:28-236     sphere vs triangle
//   BulletCollision/NarrowPhaseCollision/btGjkPairDetector.cpp              closest points box/sphere, box/triangle (margins)
//   BulletCollision/CollisionDispatch/btBoxBoxDetector.cpp:267              box vs box SAT
//   BulletCollision/CollisionDispatch/btManifoldResult.cpp:56-77,109-206    material combine + contact callback
//   BulletCollision/NarrowPhaseCollision/btPersistentManifold.cpp:110-227   4-point manifold reduction
//   BulletDynamics/ConstraintSolver/btSequentialImpulseConstraintSolver.cpp:795-990,1071-1211,1601-1830,1878-1904
//   RocketSim/src/Sim/Arena/Arena.cpp:218-427                                contact callbacks (car-ball, car-car, car-world, ball-world)
#pragma once
#include "sim_car.cuh"
#include "sim_prof.cuh"
#include "sim_edge.cuh"

namespace ggl {

constexpr float kBallRadiusBT = rl::BALL_RADIUS * kUU2BT;
constexpr float kBallCBT = (rl::BALL_RADIUS * kUU2BT + 0.08f) * 0.02f;     // contact breaking threshold of the ball shape
constexpr float kBallInvMass = 1.f / rl::BALL_MASS;

// box shape of the car as the reference builds it: implicit half extents = he - 0.04, margin limited to 10% of the
// smallest half extent (btBoxShape ctor + setSafeMargin)
struct BoxGeom {
    V3 core;        // half extents without margin
    float margin;
    V3 full;        // core + margin
};
GGL_HD BoxGeom car_box() {
    BoxGeom b;
    V3 he = car_half_extents();
    b.core = he - v3(rl::CONVEX_MARGIN, rl::CONVEX_MARGIN, rl::CONVEX_MARGIN);
    float mn = fminf(he.x, fminf(he.y, he.z));
    float safe = 0.1f * mn;
    b.margin = safe < rl::CONVEX_MARGIN ? safe : rl::CONVEX_MARGIN;
    b.full = b.core + v3(b.margin, b.margin, b.margin);
    return b;
}
// contact breaking threshold of the car's compound shape: 0.02 * (bounding sphere radius + |centre|) of its local AABB
GGL_HD float car_cbt() {
    BoxGeom b = car_box();
    V3 off = car_hitbox_offset();
    V3 mn = off - b.full, mx = off + b.full;
    float radius = len(mx - mn) * 0.5f;
    V3 centre = (mn + mx) * 0.5f;
    return (radius + len(centre)) * 0.02f;
}

GGL_HD float ball_inv_inertia() { return 1.f / (0.4f * rl::BALL_MASS * kBallRadiusBT * kBallRadiusBT); }

// ---------------------------------------------------------------------------------------------- contact emission
// Where a body's contact generation writes: first a private WINDOW of A.contacts (so that the lanes of a group can generate
// the contacts of different bodies concurrently), then slots claimed one by one from the arena's shared overflow POOL at the
// end of A.contacts. Layout of A.contacts for P cars (contact_car_window / contact_pool_base):
//   [0, 4)                 ball-world rows
//   [4 + i R, 4 + (i+1) R) car i's world + car-car rows, R = 8 / 4 / 3 for 1v1 / 2v2 / 3v3
//   [4 + P R, kMaxContacts) the pool: 16 / 16 / 14 slots; a car's single car-ball row always lives here
// gather_contacts turns windows + pool into the solver's row order (A.contactOrder) without moving a row. A point that finds
// neither a window row nor a pool slot is dropped and counted in A.contactOverflow (sticky; tests and smoke assert it is 0).
enum : uint8_t { POOL_CAR_REST = 0x00, POOL_BALL_WORLD = 0x20, POOL_CAR_BALL = 0x40 };      // | car index
GGL_HD int contact_car_window(int numCars) { return numCars <= 2 ? 8 : (numCars <= 4 ? 4 : 3); }
GGL_HD int contact_pool_base(int numCars) { return 4 + numCars * contact_car_window(numCars); }
GGL_HD int contact_pool_size(int numCars) { return kMaxContacts - contact_pool_base(numCars); }

struct ContactSink {
    Contact* buf;           // the body's window
    int n, cap;             // logical rows so far (window first, then pool), window size
    ArenaDev* A;            // for the pool
    uint8_t owner;          // POOL_* | car index
    // logical row k: window row k, or the (k - cap)-th pool slot this sink claimed (slots are claimed in increasing order by
    // one lane, so the j-th slot with this owner tag is the j-th claim; the pool path is rare, the search is 16 bytes)
    GGL_HD Contact& at(int k) {
        if (k < cap) return buf[k];
        int j = k - cap, slot = 0;
        for (;; slot++)
            if (A->poolOwner[slot] == owner && j-- == 0) break;
        return A->contacts[contact_pool_base(A->numCars) + slot];
    }
};
GGL_HD ContactSink make_sink(ArenaDev& A, Contact* window, int cap, uint8_t owner) {
    ContactSink S; S.buf = window; S.n = 0; S.cap = cap; S.A = &A; S.owner = owner;
    return S;
}
// one more logical row: the next window row or a freshly claimed pool slot
GGL_HD bool sink_grow(ContactSink& S) {
    if (S.n < S.cap) { S.n++; return true; }
    const int size = contact_pool_size(S.A->numCars);
#if defined(__CUDA_ARCH__)
    const uint32_t slot = atomicAdd(&S.A->poolCount, 1u);
#else
    const uint32_t slot = S.A->poolCount++;
#endif
    if ((int)slot >= size) {
#if defined(__CUDA_ARCH__)
        atomicAdd(&S.A->contactOverflow, 1u);
#else
        S.A->contactOverflow++;
#endif
        return false;
    }
    S.A->poolOwner[slot] = S.owner;
    S.n++;
    return true;
}
struct ManifoldCtx {     // <= 4 points per body pair, rebuilt each tick
    int first;           // index of the manifold's first row in the sink
    int count;
    V3 localA[4];
};

GGL_HD float area3(V3 p, V3 q, V3 r, V3 s) { return len2(cross(p - q, s - r)); }

// Adds a point to the manifold (replacement policy of btPersistentManifold::sortCachedPoints). Returns the row index in the sink.
GGL_HD int manifold_add(ContactSink& S, ManifoldCtx& m, const Contact& c, V3 localA) {
    int idx;
    if (m.count < 4) {
        if (!sink_grow(S)) return -1;
        idx = m.count++;
    } else {
        int deepest = -1;
        float maxPen = c.dist;
        for (int i = 0; i < 4; i++)
            if (S.at(m.first + i).dist < maxPen) { deepest = i; maxPen = S.at(m.first + i).dist; }
        float r0 = 0, r1 = 0, r2 = 0, r3 = 0;
        if (deepest != 0) r0 = area3(localA, m.localA[1], m.localA[2], m.localA[3]);
        if (deepest != 1) r1 = area3(localA, m.localA[0], m.localA[2], m.localA[3]);
        if (deepest != 2) r2 = len2(cross(localA - m.localA[0], m.localA[3] - m.localA[1]));
        if (deepest != 3) r3 = len2(cross(localA - m.localA[0], m.localA[2] - m.localA[1]));
        float best = fabsf(r0); idx = 0;
        if (fabsf(r1) > best) { best = fabsf(r1); idx = 1; }
        if (fabsf(r2) > best) { best = fabsf(r2); idx = 2; }
        if (fabsf(r3) > best) { best = fabsf(r3); idx = 3; }
    }
    S.at(m.first + idx) = c;
    m.localA[idx] = localA;
    return m.first + idx;
}

GGL_HD ManifoldCtx manifold_begin(const ContactSink& S) { ManifoldCtx m; m.first = S.n; m.count = 0; return m; }

// ---- GameMode::HEATSEEKER: the ball steers towards a goal (Sim/Ball/Ball.cpp:152-253) ----
// Angle helpers of RocketSim (Math/MathTypes/MathTypes.cpp:80-120, Math/Math.cpp:66-98), yaw / pitch only (roll stays 0)
GGL_HD float wrap_normalize(float val, float minmax) {
    float r = fmodf(val, minmax * 2);
    if (r > minmax) r -= minmax * 2;
    else if (r < -minmax) r += minmax * 2;
    return r;
}
GGL_HD void angle_from_vec(V3 f, float& yaw, float& pitch) {
    const float eps = 1.1920929e-7f;
    if (fabsf(f.y) > eps || fabsf(f.x) > eps) {
        yaw = atan2f(f.y, f.x);
        const float d2 = sqrtf(f.x * f.x + f.y * f.y);
        pitch = -atan2f(-f.z, d2);
    } else {
        yaw = 0.f;
        pitch = f.z > eps ? (float)(3.14159265358979323846 / 2) : (f.z < -eps ? (float)(-3.14159265358979323846 / 2) : 0.f);
    }
}
// Ball::_PreTickUpdate for HEATSEEKER (:153-201)
GGL_HD void ball_heatseeker_pre_tick(BallDev& b) {
    const float dirY = b.hsTargetDir;
    if (dirY == 0.f) return;
    const float kPiF2 = (float)3.14159265358979323846, kHalf = (float)(3.14159265358979323846 / 2);
    const V3 pos = b.pos * kBT2UU, vel = b.vel * kBT2UU;
    float vy, vp, gy, gp;
    angle_from_vec(vel, vy, vp);
    angle_from_vec(v3(0.f, rl::hs::TARGET_Y * dirY, rl::hs::TARGET_Z) - pos, gy, gp);
    // deltaAngle = angleToGoal - velAngle = velAngle.GetDeltaTo(angleToGoal), NormalizeFix'd
    const float dy = wrap_normalize(gy - vy, kPiF2), dp = wrap_normalize(gp - vp, kHalf);
    const float l2 = vel.x * vel.x + vel.y * vel.y + vel.z * vel.z;
    const float curSpeed = l2 > 0.f ? sqrtf(l2) : 0.f;                      // Vec::Length
    const float speedRatio = curSpeed / rl::hs::MAX_SPEED;
    const float base = speedRatio * kTick;
    float ny = vy + dy * base * rl::hs::HORIZONTAL_BLEND;
    float np = vp + dp * base * rl::hs::VERTICAL_BLEND;
    ny = wrap_normalize(ny, kPiF2); np = wrap_normalize(np, kHalf);
    np = np < -rl::hs::MAX_TURN_PITCH ? -rl::hs::MAX_TURN_PITCH : (np > rl::hs::MAX_TURN_PITCH ? rl::hs::MAX_TURN_PITCH : np);   // RS_CLAMP = max(min, min(v, max))
    {   // Math::RoundAngleUE3
        const float TO_INTS = (float)((double)(float)(1 << 15) / 3.14159265358979323846);
        const float BACK = (1.f / TO_INTS) * 4.f;
        const int ry = ((int)(ny * TO_INTS) >> 2) & 0x3FFF, rp = ((int)(np * TO_INTS) >> 2) & 0x3FFF;
        ny = ry * BACK; np = rp * BACK;
    }
    const float newSpeed = curSpeed + ((b.hsTargetSpeed - curSpeed) * rl::hs::SPEED_BLEND);
    const float cp = cosf(-np), cy = cosf(ny), sy = sinf(ny), sp = sinf(-np);   // Angle::GetForwardVec
    const V3 dir = v3(cp * cy, cp * sy, -sp);
    b.vel = (dir * newSpeed) * kUU2BT;
    b.hsTimeSinceHit += kTick;
}
// Ball::_OnHit (:203-214)
GGL_HD void ball_heatseeker_on_hit(BallDev& b, int carTeam) {
    const bool canIncrease = (b.hsTimeSinceHit > rl::hs::MIN_SPEEDUP_INTERVAL) || (b.hsTargetDir == 0.f);
    const float newDir = carTeam == 0 ? 1.f : -1.f;
    if (canIncrease && newDir != b.hsTargetDir) {
        b.hsTimeSinceHit = 0.f;
        b.hsTargetSpeed = fminf(b.hsTargetSpeed + rl::hs::TARGET_SPEED_INCREMENT, rl::hs::MAX_SPEED);
    }
    b.hsTargetDir = newDir;
}
// Ball::_OnWorldCollision (:216-246): called for every ball-world contact point as it is added, with the contact normal
GGL_HD void ball_heatseeker_on_world(BallDev& b, V3 normal) {
    if (b.hsTargetDir == 0.f) return;
    const V3 pos = b.pos * kBT2UU;
    const float relNormalY = normal.y * b.hsTargetDir, relY = pos.y * b.hsTargetDir;
    if (relNormalY <= -rl::hs::WALL_BOUNCE_CHANGE_Y_NORMAL && relY >= rl::ARENA_EXTENT_Y - rl::hs::WALL_BOUNCE_CHANGE_Y_THRESH) {
        b.hsTargetDir *= -1.f;
        const V3 vel = b.vel * kBT2UU;
        const V3 goal = v3(0.f, rl::hs::TARGET_Y * b.hsTargetDir, rl::hs::TARGET_Z);
        const V3 toGoal = rs_normalized(goal - pos);
        const V3 bounceDir = toGoal * (1 - rl::hs::WALL_BOUNCE_UP_FRAC) + v3(0, 0, 1) * rl::hs::WALL_BOUNCE_UP_FRAC;
        const float l2 = vel.x * vel.x + vel.y * vel.y + vel.z * vel.z;
        const V3 impulse = bounceDir * (l2 > 0.f ? sqrtf(l2) : 0.f) * rl::hs::WALL_BOUNCE_FORCE_SCALE;
        b.velImpulseCache += impulse * kUU2BT;
    }
}

// ---------------------------------------------------------------------------------------------- narrow phase
// ball vs the 4 planes and the mesh; every contact is flagged special (Arena.cpp:265-273)
// Contact::staticId of a row against the static world says which static body it touches (the reference has one manifold per body
// pair; solver_order needs to tell them apart): 1 + mesh for an arena mesh, -1 .. -4 for floor, ceiling, left wall, right wall.
GGL_HD int16_t plane_static_code(V3 n) { return (int16_t)(n.z > 0.f ? -1 : (n.z < 0.f ? -2 : (n.x > 0.f ? -3 : -4))); }

GGL_HD void collide_ball_plane(ArenaDev& A, ContactSink& S, V3 n, V3 origin) {
    const BallDev& b = A.ball;
    V3 vtx = (-n) * kBallRadiusBT;            // support point of the sphere towards the plane
    V3 inPlane = (b.pos - origin) + vtx;
    float dist = dot(n, inPlane);
    if (dist < kBallCBT) {
        V3 proj = inPlane - dist * n;
        V3 pOnB = origin + proj;
        ManifoldCtx m = manifold_begin(S);
        Contact c;
        c.normal = n; c.dist = dist;
        V3 pA = pOnB + n * dist;
        c.relA = pA - b.pos; c.relB = pOnB - origin;
        c.friction = rl::BALL_FRICTION < rl::WORLD_FRICTION ? rl::BALL_FRICTION : rl::WORLD_FRICTION;
        c.restitution = rl::BALL_RESTITUTION > rl::WORLD_RESTITUTION ? rl::BALL_RESTITUTION : rl::WORLD_RESTITUTION;
        c.bodyA = 0; c.bodyB = -1; c.special = 1; c.staticId = plane_static_code(n);
        if (manifold_add(S, m, c, pA - b.pos) >= 0 && A.gameMode == kModeHeatseeker) ball_heatseeker_on_world(A.ball, n);
    }
}

GGL_HD V3 closest_point_triangle(V3 p, V3 a, V3 b, V3 c) {
    V3 ab = b - a, ac = c - a, ap = p - a;
    float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) return a;
    V3 bp = p - b;
    float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) return b;
    V3 cp = p - c;
    float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.f && d5 <= d6) return c;
    float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { float v = d1 / (d1 - d3); return a + v * ab; }
    float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { float v = d2 / (d2 - d6); return a + v * ac; }
    float va = d3 * d6 - d5 * d4;
    if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) { float v = (d4 - d3) / ((d4 - d3) + (d5 - d6)); return b + v * (c - b); }
    float denom = 1.f / (va + vb + vc);
    float v = vb * denom, w = vc * denom;
    return a + v * ab + w * ac;
}

GGL_HD bool sphere_triangle(V3 centre, const Tri& t, float radius, float cbt, V3& point, V3& normalOut, float& depth) {
    float rt = radius + cbt;
    V3 normal = cross(t.v1 - t.v0, t.v2 - t.v0);
    float l2 = len2(normal);
    bool has = false;
    V3 cp = v3(0, 0, 0);
    if (l2 >= kEps * kEps) {
        normal = normal / sqrtf(l2);
        float dplane = dot(centre - t.v0, normal);
        if (dplane < 0.f) { dplane *= -1.f; normal = normal * -1.f; }
        if (dplane < rt) {
            // barycentric point-in-triangle test
            V3 u = t.v1 - t.v0, v = t.v2 - t.v0, n = cross(u, v), w = centre - t.v0;
            float nl = dot(n, n);
            float gamma = dot(cross(u, w), n) / nl, beta = dot(cross(w, v), n) / nl, alpha = 1 - gamma - beta;
            if (0 <= alpha && alpha <= 1 && 0 <= beta && beta <= 1 && 0 <= gamma && gamma <= 1) {
                has = true;
                cp = centre - normal * dplane;
            } else {
                V3 near = closest_point_triangle(centre, t.v0, t.v1, t.v2);
                if (len2(near - centre) < rt * rt) { has = true; cp = near; }
            }
        }
    }
    if (has) {
        V3 toC = centre - cp;
        float d2 = len2(toC);
        if (d2 < rt * rt) {
            if (d2 > kEps) {
                float d = sqrtf(d2);
                normalOut = normalized(toC);
                point = cp;
                depth = -(radius - d);
            } else {
                normalOut = normal; point = cp; depth = -radius;
            }
            return true;
        }
    }
    return false;
}

GGL_HD bool tri_overlaps_aabb(const Tri& t, V3 mn, V3 mx) {
    if (fminf(fminf(t.v0.x, t.v1.x), t.v2.x) > mx.x) return false;
    if (fmaxf(fmaxf(t.v0.x, t.v1.x), t.v2.x) < mn.x) return false;
    if (fminf(fminf(t.v0.z, t.v1.z), t.v2.z) > mx.z) return false;
    if (fmaxf(fmaxf(t.v0.z, t.v1.z), t.v2.z) < mn.z) return false;
    if (fminf(fminf(t.v0.y, t.v1.y), t.v2.y) > mx.y) return false;
    if (fmaxf(fmaxf(t.v0.y, t.v1.y), t.v2.y) < mn.y) return false;
    return true;
}

GGL_HD void collide_ball_mesh(ArenaDev& A, ContactSink& S, const WorldDev& W) {
    if (W.numTris <= 0) return;
    const BallDev& b = A.ball;
    float ext = kBallRadiusBT + 0.08f;
    V3 mn = b.pos - v3(ext, ext, ext), mx = b.pos + v3(ext, ext, ext);
    if (world_aabb_is_free(W, mn, mx)) return;          // open play: every grid cell the ball's box touches is empty
    int x0 = grid_coord(mn.x, W.gridMin.x, W.invCellSize, W.gx), x1 = grid_coord(mx.x, W.gridMin.x, W.invCellSize, W.gx);
    int y0 = grid_coord(mn.y, W.gridMin.y, W.invCellSize, W.gy), y1 = grid_coord(mx.y, W.gridMin.y, W.invCellSize, W.gy);
    int z0 = grid_coord(mn.z, W.gridMin.z, W.invCellSize, W.gz), z1 = grid_coord(mx.z, W.gridMin.z, W.invCellSize, W.gz);
    // candidate triangles in (mesh, triangle) index order; one manifold per mesh body (the list lives in the arena's scratch: a
    // stack array with run-time indices would be local memory, and this is a one-lane insertion sort)
    uint16_t* cand = A.ballCand; int nc = 0;
    for (int z = z0; z <= z1; z++)
        for (int y = y0; y <= y1; y++)
            for (int x = x0; x <= x1; x++) {
                int cell = (z * W.gy + y) * W.gx + x;
                for (int k = W.cellStart[cell]; k < W.cellStart[cell + 1]; k++) {
                    int ti = W.cellTris[k];
                    int pos = 0; bool dup = false;
                    while (pos < nc && (int)cand[pos] <= ti) { dup |= (int)cand[pos] == ti; pos++; }
                    if (dup || nc >= 24) continue;
                    // cull (exact): a triangle whose box misses the ball's box is further than radius + 0.08 away, and
                    // sphere_triangle reports nothing beyond radius + kBallCBT (0.038)
                    if (!tri_overlaps_aabb(W.tris[ti], mn, mx)) continue;
                    for (int q = nc; q > pos; q--) cand[q] = cand[q - 1];
                    cand[pos] = (uint16_t)ti; nc++;
                }
            }
    int curMesh = -1;
    ManifoldCtx m = manifold_begin(S);
    for (int k = 0; k < nc; k++) {
        const Tri& t = W.tris[cand[k]];
        int mesh = (int)(t.flags >> TRI_MESH_SHIFT);
        if (mesh != curMesh) { m = manifold_begin(S); curMesh = mesh; }
        V3 p, n; float depth;
        if (sphere_triangle(b.pos, t, kBallRadiusBT, kBallCBT, p, n, depth)) {
            if (depth > kBallCBT) continue;
            Contact c;
            c.normal = n; c.dist = depth;
            V3 pA = p + n * depth;
            c.relA = pA - b.pos; c.relB = p;
            c.friction = rl::BALL_FRICTION < rl::WORLD_FRICTION ? rl::BALL_FRICTION : rl::WORLD_FRICTION;
            c.restitution = rl::BALL_RESTITUTION > rl::WORLD_RESTITUTION ? rl::BALL_RESTITUTION : rl::WORLD_RESTITUTION;
            c.bodyA = 0; c.bodyB = -1; c.special = 1; c.staticId = (int16_t)(1 + mesh);
            int row = manifold_add(S, m, c, pA - b.pos);
            if (row < 0) continue;
            if (A.gameMode == kModeHeatseeker) ball_heatseeker_on_world(A.ball, n);     // before the internal-edge adjustment (Arena.cpp:268-277)
            adjust_internal_edge(t, n, p, pA, depth);
            S.at(row).normal = n;
            S.at(row).relB = p;
        }
    }
}

// car hitbox vs plane: the single deepest support vertex (perturbation iterations are off in the reference)
GGL_HD void collide_car_plane(ArenaDev& A, ContactSink& S, int ci, V3 n, V3 origin, V3 boxPos /* car.pos + basis * hitbox offset */) {
    CarDev& car = A.cars[ci];
    const BoxGeom bg = car_box();
    // cull (exact): the support vertex is never further inside than the box centre minus the half diagonal
    if (dot(n, boxPos - origin) - len(bg.full) * 1.0001f > car_cbt()) return;
    V3 dirLocal = tmul(car.basis, -n);
    V3 vtx = v3(dirLocal.x >= 0.f ? bg.full.x : -bg.full.x, dirLocal.y >= 0.f ? bg.full.y : -bg.full.y, dirLocal.z >= 0.f ? bg.full.z : -bg.full.z);
    V3 inPlane = (boxPos - origin) + mul(car.basis, vtx);
    float dist = dot(n, inPlane);
    if (dist < car_cbt()) {
        V3 pOnB = origin + (inPlane - dist * n);
        V3 pA = pOnB + n * dist;
        ManifoldCtx m = manifold_begin(S);
        Contact c;
        c.normal = n; c.dist = dist;
        c.relA = pA - car.pos; c.relB = pOnB - origin;
        c.friction = rl::CARWORLD_FRICTION; c.restitution = rl::CARWORLD_RESTITUTION;
        c.bodyA = (int16_t)(1 + ci); c.bodyB = -1; c.special = 0; c.staticId = plane_static_code(n);
        if (manifold_add(S, m, c, tmul(car.basis, pA - car.pos)) >= 0) {
            car.worldContactHas = 1;
            car.worldContactNormal = n;
        }
    }
}

// car hitbox (core box + margin) vs ball (point + radius): closest points, GJK margin convention
GGL_HD void collide_car_ball(ArenaDev& A, ContactSink& S, int ci) {
    CarDev& car = A.cars[ci];
    BallDev& ball = A.ball;
    const BoxGeom bg = car_box();
    V3 boxPos = car.pos + mul(car.basis, car_hitbox_offset());
    V3 cl = tmul(car.basis, ball.pos - boxPos);
    V3 q = v3(fclamp(cl.x, -bg.core.x, bg.core.x), fclamp(cl.y, -bg.core.y, bg.core.y), fclamp(cl.z, -bg.core.z, bg.core.z));
    V3 vl = q - cl;                         // from the sphere centre to the box core (B -> A)
    float d2 = len2(vl);
    float margins = bg.margin + kBallRadiusBT;
    float maxd = margins + kBallCBT;
    V3 nl; float dist;
    if (d2 > kEps * kEps) {
        if (d2 > maxd * maxd) return;
        float d = sqrtf(d2);
        nl = vl * (1.f / d);
        dist = d - margins;
    } else {
        // centre inside the core box: axis of least penetration
        float px = bg.core.x - fabsf(cl.x), py = bg.core.y - fabsf(cl.y), pz = bg.core.z - fabsf(cl.z);
        if (px <= py && px <= pz) { nl = v3(cl.x >= 0.f ? -1.f : 1.f, 0, 0); dist = -px - margins; }
        else if (py <= pz) { nl = v3(0, cl.y >= 0.f ? -1.f : 1.f, 0); dist = -py - margins; }
        else { nl = v3(0, 0, cl.z >= 0.f ? -1.f : 1.f); dist = -pz - margins; }
    }
    if (dist > kBallCBT) return;
    V3 n = mul(car.basis, nl);
    V3 pOnB = ball.pos + n * kBallRadiusBT;
    V3 pA = pOnB + n * dist;
    ManifoldCtx m = manifold_begin(S);
    Contact c;
    c.normal = n; c.dist = dist;
    c.relA = pA - car.pos; c.relB = pOnB - ball.pos;
    c.friction = rl::CARBALL_FRICTION; c.restitution = rl::CARBALL_RESTITUTION;
    c.bodyA = (int16_t)(1 + ci); c.bodyB = 0; c.special = 0; c.staticId = 0;
    if (manifold_add(S, m, c, tmul(car.basis, pA - car.pos)) < 0) return;

    // Arena::_BtCallback_OnCarBallCollision (Arena.cpp:283-334)
    V3 carPosUU = car.pos * kBT2UU, carVelUU = car.vel * kBT2UU;
    V3 ballPosUU = ball.pos * kBT2UU, ballVelUU = ball.vel * kBT2UU;
    car.hitValid = 1;
    car.hitRelPos = c.relB * kBT2UU;
    car.hitTick = A.tickCount;
    car.hitBallPos = ballPosUU;
    car.hitExtraVel = v3(0, 0, 0);
    if ((A.tickCount > car.hitTickExtra + 1) || (car.hitTickExtra > A.tickCount)) {
        car.hitTickExtra = A.tickCount;
    } else {
        return;
    }
    car.ballHitNew = 2;             // Ball::_OnHit runs whenever the extra impulse was not skipped (Arena.cpp:333), in car order (gather_contacts)
    V3 fwd = car_fwd(car);
    V3 relPos = ballPosUU - carPosUU, relVel = ballVelUU - carVelUU;
    float relSpeed = fminf(len(relVel), rl::BALL_CAR_EXTRA_IMPULSE_MAXDELTAVEL);
    if (relSpeed > 0.f) {
        V3 hitDir = safe_normalized(relPos * v3(1, 1, rl::BALL_CAR_EXTRA_IMPULSE_Z_SCALE));
        V3 adj = fwd * dot(hitDir, fwd) * (1 - rl::BALL_CAR_EXTRA_IMPULSE_FORWARD_SCALE);
        hitDir = safe_normalized(hitDir - adj);
        V3 added = (hitDir * relSpeed) * curve_ball_extra_impulse(relSpeed) * 1.f;
        car.hitExtraVel = added;
        car.ballHitNew = 3;         // bit 0: ball.velImpulseCache += added, done in car order by gather_contacts
    }
}

// ---------------------------------------------------------------------------------------------- solver
struct SolverBody {
    V3 dv, dw, push, turn, extF, extT, vel, angVel;
    float invMass;
    M3 invI;
    V3 pos;
    int active;
};

// One solver row, 20 floats = 80 bytes: exactly the per-tick scratch of one wheel (Wheel::hardPoint .. Wheel::ground), which is
// dead between the end of the car pre-tick and the next tick's ray casts. That is where the rows of an arena live while its
// islands are solved (row_slot): like the solver bodies above, they are walked by one lane per island, and on a per-thread
// stack every float of a row would be a sector of its own. n2 of a two-body row is -n1 and is not stored; limits, body indices
// and the friction coefficient come from the contact (still in shared memory) or are arguments.
struct Row {
    V3 n1, rxn1, angA, rxn2, angB;
    float jacInv, rhs, rhsPen, applied, appliedPush;
};
static_assert(sizeof(Row) == 80, "a solver row must fit the scratch part of a Wheel");
static_assert(offsetof(Wheel, ground) + sizeof(int) - offsetof(Wheel, hardPoint) == sizeof(Row), "Wheel scratch layout changed");

GGL_HD void sb_apply(SolverBody& s, V3 lin, V3 ang, float mag) { s.dv += lin * mag; s.dw += ang * mag; }
GGL_HD void sb_push(SolverBody& s, V3 lin, V3 ang, float mag) { s.push += lin * mag; s.turn += ang * mag; }

GGL_HD void setup_contact_row(Row& r, const SolverBody* sb, int a, int b, V3 n, V3 rel1, V3 rel2, float dist, float restitutionCoef) {
    const float invDt = 1.f / kTick;
    const SolverBody& A = sb[a];
    V3 t0 = cross(rel1, n);
    r.angA = mul(A.invI, t0);
    V3 t1 = cross(rel2, n);
    r.angB = (b >= 0) ? mul(sb[b].invI, -t1) : v3(0, 0, 0);
    float denom0 = A.invMass + dot(n, cross(r.angA, rel1));
    float denom1 = 0.f;
    if (b >= 0) denom1 = sb[b].invMass + dot(n, cross(-r.angB, rel2));
    r.jacInv = 1.f / (denom0 + denom1 + 0.f);
    r.n1 = n; r.rxn1 = t0;
    V3 n2;
    if (b >= 0) { n2 = -n; r.rxn2 = -t1; } else { n2 = v3(0, 0, 0); r.rxn2 = v3(0, 0, 0); }
    float penetration = dist + 0.f;
    V3 vel1 = vel_at(A.vel, A.angVel, rel1);
    V3 vel2 = (b >= 0) ? vel_at(sb[b].vel, sb[b].angVel, rel2) : v3(0, 0, 0);
    float relVel0 = dot(n, vel1 - vel2);
    float restitution = (fabsf(relVel0) < 0.2f) ? 0.f : restitutionCoef * -relVel0;
    if (restitution <= 0.f) restitution = 0.f;
    r.applied = 0.f; r.appliedPush = 0.f;
    V3 zero = v3(0, 0, 0);
    float v1n = dot(r.n1, A.vel + A.extF) + dot(r.rxn1, A.angVel + A.extT);
    float v2n = (b >= 0) ? dot(n2, sb[b].vel + sb[b].extF) + dot(r.rxn2, sb[b].angVel + sb[b].extT) : dot(n2, zero);
    float relVel = v1n + v2n;
    float posErr = 0.f;
    float velErr = restitution - relVel;
    if (penetration > 0.f) posErr = 0.f;
    else posErr = -penetration * 0.8f * invDt;
    r.rhs = velErr * r.jacInv;
    r.rhsPen = posErr * r.jacInv;
}

GGL_HD void setup_friction_row(Row& r, const SolverBody* sb, int a, int b, V3 axis, V3 rel1, V3 rel2) {
    const SolverBody& A = sb[a];
    r.n1 = axis;
    r.rxn1 = cross(rel1, axis);
    r.angA = mul(A.invI, r.rxn1);
    V3 n2;
    if (b >= 0) {
        n2 = -axis;
        r.rxn2 = cross(rel2, n2);
        r.angB = mul(sb[b].invI, r.rxn2);
    } else {
        n2 = r.rxn2 = r.angB = v3(0, 0, 0);
    }
    float denom0 = A.invMass + dot(axis, cross(r.angA, rel1));
    float denom1 = (b >= 0) ? sb[b].invMass + dot(axis, cross(-r.angB, rel2)) : 0.f;
    r.jacInv = 1.f / (denom0 + denom1);
    float v1n = dot(r.n1, A.vel + A.extF) + dot(r.rxn1, A.angVel);
    float v2n = (b >= 0) ? dot(n2, sb[b].vel + sb[b].extF) + dot(r.rxn2, sb[b].angVel) : 0.f;
    float relVel = v1n + v2n;
    r.rhs = (0.f - relVel) * r.jacInv;
    r.rhsPen = 0.f;
    r.applied = 0.f; r.appliedPush = 0.f;
}

GGL_HD void friction_dir(const SolverBody* sb, int a, int b, V3 n, V3 rel1, V3 rel2, V3& dir) {
    // btSolverBody::getVelocityInLocalPointNoDelta: velocities including the external force/torque impulses
    V3 vel1 = vel_at(sb[a].vel + sb[a].extF, sb[a].angVel + sb[a].extT, rel1);
    V3 vel2 = (b >= 0) ? vel_at(sb[b].vel + sb[b].extF, sb[b].angVel + sb[b].extT, rel2) : v3(0, 0, 0);
    V3 vel = vel1 - vel2;
    float rv = dot(n, vel);
    dir = vel - n * rv;
    float l2 = len2(dir);
    if (l2 > kEps) {
        dir *= 1.f / sqrtf(l2);
    } else {
        V3 q;
        plane_space(n, dir, q);
    }
}

// returns the impulse change it applied
GGL_HD float resolve_row(Row& c, SolverBody* sb, int a, int b, float lower, float upper, bool lowerOnly) {
    float delta = c.rhs - c.applied * 0.f;
    SolverBody& A = sb[a];
    const V3 n2 = -c.n1;
    float d1 = dot(c.n1, A.dv) + dot(c.rxn1, A.dw);
    float d2 = (b >= 0) ? dot(n2, sb[b].dv) + dot(c.rxn2, sb[b].dw) : 0.f;
    delta -= d1 * c.jacInv;
    delta -= d2 * c.jacInv;
    float sum = c.applied + delta;
    if (sum < lower) { delta = lower - c.applied; c.applied = lower; }
    else if (!lowerOnly && sum > upper) { delta = upper - c.applied; c.applied = upper; }
    else c.applied = sum;
    sb_apply(A, c.n1 * A.invMass, c.angA, delta);
    if (b >= 0) sb_apply(sb[b], n2 * sb[b].invMass, c.angB, delta);
    return delta;
}

GGL_HD void resolve_push(Row& c, SolverBody* sb, int a, int b) {
    if (c.rhsPen == 0.f) return;
    float delta = c.rhsPen - c.appliedPush * 0.f;
    SolverBody& A = sb[a];
    const V3 n2 = -c.n1;
    float d1 = dot(c.n1, A.push) + dot(c.rxn1, A.turn);
    float d2 = (b >= 0) ? dot(n2, sb[b].push) + dot(c.rxn2, sb[b].turn) : 0.f;
    delta -= d1 * c.jacInv;
    delta -= d2 * c.jacInv;
    float sum = c.appliedPush + delta;
    if (sum < 0.f) { delta = 0.f - c.appliedPush; c.appliedPush = 0.f; }
    else c.appliedPush = sum;
    sb_push(A, c.n1 * A.invMass, c.angA, delta);
    if (b >= 0) sb_push(sb[b], n2 * sb[b].invMass, c.angB, delta);
}

constexpr int kMaxRows = kMaxContacts + 1;

// Solves the contact rows of ONE island (the bodies whose island label is `rep`, A.island[] from gather_contacts) and writes
// their velocities (and push-corrected positions) back. Projected Gauss-Seidel is sequential over rows, but a row only reads
// and writes the solver bodies it connects, so islands that share no dynamic body are independent: solving them on
// different lanes, each in the reference's row order restricted to the island, gives bit-identical results to the single
// sequential sweep of btSequentialImpulseConstraintSolver (the host build does the same thing island after island).
//
// Where the data lives. The sweep is one dependent chain through dv / dw of the bodies a row connects, walked by ONE lane per
// island; on a per-thread stack every float of that chain is a sector of its own (words of one lane's stack sit 128 bytes
// apart) in a 28 KB L1 that 224 threads share, i.e. a local-memory round trip per step. So:
//   * the solver bodies of an arena live in its shared-memory scratch, the car-vs-mesh hit staging, which is dead once contact
//     generation is over (bodies belong to exactly one island, so the lanes of different islands never share an entry);
//   * the rows live in the wheels' per-tick scratch (row 2k = contact k of the arena's solver order, 2k + 1 = its friction
//     row, the averaged ball-world pair last; row s sits in wheel s & 3 of car s >> 2) as long as the arena has no more
//     contacts than 2 x cars and no demolished car (a demolished car parks its reported orientation in wheel 0); otherwise
//     (pile-ups) on the calling lane's stack. The friction row's two unused words carry the row's bodies and coefficient.
static_assert(sizeof(SolverBody) * (1 + kMaxCars) <= sizeof(CarTriHit) * kMaxCars * kMaxMeshHit, "solver bodies must fit the mesh-hit staging");

template <bool SMEM>
struct RowStore {
    ArenaDev& A;
    Row lrows[SMEM ? 1 : 2 * kMaxRows];
    GGL_HD explicit RowStore(ArenaDev& a) : A(a) {}
    GGL_HD Row& at(int s) {
        if (SMEM) return *reinterpret_cast<Row*>(&A.cars[s >> 2].w[s & 3].hardPoint);
        return lrows[s];
    }
};

template <bool SMEM>
GGL_HD void solve_island_rows(ArenaDev& A, SolverBody* sb, int rep) {
    RowStore<SMEM> R(A);
    const int nc = A.numContacts;
    // the island's rows as a packed list of contact indices (6 bits each would not fit 37 entries: the stack path keeps a byte
    // list, the shared-memory path has at most 13 entries of 5 bits)
    unsigned long long list = 0;
    uint8_t llist[SMEM ? 1 : kMaxRows];
    int nr = 0;
    auto put = [&](int k) { if (SMEM) list |= (unsigned long long)k << (5 * nr); else llist[nr] = (uint8_t)k; nr++; };
    auto get = [&](int j) -> int { return SMEM ? (int)((list >> (5 * j)) & 31ull) : (int)llist[j]; };
    int nSpecial = 0; V3 spNormal = v3(0, 0, 0); float spDist = 0.f, spFriction = 0.f, spRest = 0.f;
    for (int k = 0; k < nc; k++) {
        const Contact& c = A.contacts[A.contactOrder[k]];
        const int a = c.bodyA, b = c.bodyB;
        if (A.island[a] != rep) continue;
        Row& N = R.at(2 * k);
        Row& F = R.at(2 * k + 1);
        setup_contact_row(N, sb, a, b, c.normal, c.relA, c.relB, c.dist, c.restitution);
        if (c.special) {
            nSpecial++; spFriction = c.friction; spRest = c.restitution; spNormal += c.normal; spDist += len(c.relA);
        }
        V3 dir;
        friction_dir(sb, a, b, c.normal, c.relA, c.relB, dir);
        setup_friction_row(F, sb, a, b, dir, c.relA, c.relB);
        F.rhsPen = c.friction;                                                                  // spare words of the friction row:
        F.appliedPush = __int_as_float_portable((a & 0xff) | ((b & 0xff) << 8) | (c.special ? 0x10000 : 0));   // coefficient, bodies, flag
        put(k);
    }
    if (nSpecial > 0) {
        // averaged ball-world contact against a fixed body (convertContactSpecial)
        float distance = spDist / nSpecial;
        V3 normal = spNormal / (float)nSpecial;
        V3 rel1 = normal * -distance, rel2 = v3(0, 0, 0);
        Row& N = R.at(2 * nc);
        Row& F = R.at(2 * nc + 1);
        setup_contact_row(N, sb, 0, -1, normal, rel1, rel2, distance, spRest);
        V3 dir;
        friction_dir(sb, 0, -1, normal, rel1, rel2, dir);
        setup_friction_row(F, sb, 0, -1, dir, rel1, rel2);
        F.rhsPen = spFriction;
        F.appliedPush = __int_as_float_portable(0 | (0xff << 8));
        put(nc);
    }
    if (nr == 0) return;
    for (int it = 0; it < 10; it++) {
        bool any = false;
        for (int j = 0; j < nr; j++) {
            const int k = get(j);
            Row& N = R.at(2 * k);
            const int meta = __float_as_int_portable(R.at(2 * k + 1).appliedPush);
            any |= N.rhsPen != 0.f;
            resolve_push(N, sb, meta & 0xff, (int)(int8_t)((meta >> 8) & 0xff));
        }
        if (!any) break;
    }
    // The reference always runs its 10 iterations (RocketSim's solveGroupCacheFriendlyIterations has no residual test). An
    // iteration in which EVERY row's impulse change is exactly zero leaves every applied impulse and every body's delta velocity
    // as it found them, so all later iterations would repeat it: stopping there changes nothing. Islands with one or two
    // active rows reach that fixed point after a few iterations; rows that only sit inside the breaking threshold at once.
    for (int it = 0; it < 10; it++) {
        bool moved = false;
        for (int j = 0; j < nr; j++) {
            const int k = get(j);
            const int meta = __float_as_int_portable(R.at(2 * k + 1).appliedPush);
            if (meta & 0x10000) continue;
            moved |= resolve_row(R.at(2 * k), sb, meta & 0xff, (int)(int8_t)((meta >> 8) & 0xff), 0.f, 1e10f, true) != 0.f;
        }
        for (int j = 0; j < nr; j++) {
            const int k = get(j);
            Row& F = R.at(2 * k + 1);
            const float tot = R.at(2 * k).applied;
            if (tot > 0.f) {
                const int meta = __float_as_int_portable(F.appliedPush);
                const float mu = F.rhsPen;
                moved |= resolve_row(F, sb, meta & 0xff, (int)(int8_t)((meta >> 8) & 0xff), -(mu * tot), mu * tot, false) != 0.f;
            }
        }
        if (!moved) break;
    }
}

GGL_HDN void solve_island_rows_on_stack(ArenaDev& A, SolverBody* sb, int rep) { solve_island_rows<false>(A, sb, rep); }   // pile-ups: out of line

GGL_HDN void solve_island(ArenaDev& A, int rep) {
    SolverBody* sb = reinterpret_cast<SolverBody*>(&A.meshHit[0][0]);
    const int nb = 1 + A.numCars;
    uint32_t inContact = 0;                            // bodies of this island that have a row
    for (int b = 0; b < nb; b++)
        if (A.island[b] == rep && (A.bodiesInContact & (1u << b))) inContact |= 1u << b;
    // ball
    if (inContact & 1u) {
        SolverBody& s = sb[0];
        s.dv = s.dw = s.push = s.turn = v3(0, 0, 0);
        s.active = (int)A.ballAwake;                   // the solver sees the ball as active when an active car touches it
        s.invMass = kBallInvMass;
        float ii = ball_inv_inertia();
        s.invI.r0 = v3(ii, 0, 0); s.invI.r1 = v3(0, ii, 0); s.invI.r2 = v3(0, 0, ii);
        s.vel = A.ball.vel; s.angVel = A.ball.angVel; s.pos = A.ball.pos;
        s.extF = A.ball.force * kBallInvMass * kTick;
        s.extT = tmul(s.invI, A.ball.torque) * kTick;
    }
    bool anyDemoed = false;
    for (int i = 0; i < A.numCars; i++) {
        anyDemoed |= A.cars[i].isDemoed != 0;
        if (!(inContact & (2u << i))) continue;
        SolverBody& s = sb[1 + i];
        const CarDev& c = A.cars[i];
        s.dv = s.dw = s.push = s.turn = v3(0, 0, 0);
        s.active = !c.rbDisabled;
        s.invMass = 1.f / rl::CAR_MASS;
        s.invI = c.invInertiaW;
        s.vel = c.vel; s.angVel = c.angVel; s.pos = c.pos;
        s.extF = c.force * s.invMass * kTick;
        s.extT = tmul(c.invInertiaW, c.torque) * kTick;
    }
    // (numContacts + the averaged pair) x 2 rows against 4 wheel slots per car; 5-bit entries in the packed list
    if (!anyDemoed && A.numContacts + 1 <= 2 * A.numCars && A.numContacts < 31) solve_island_rows<true>(A, sb, rep);
    else solve_island_rows_on_stack(A, sb, rep);
    // write back
    for (int i = 0; i < nb; i++) {
        if (!(inContact & (1u << i))) continue;
        SolverBody& s = sb[i];
        if (!s.active) continue;
        V3 v = (s.vel + s.dv) + s.extF;
        V3 w = (s.angVel + s.dw) + s.extT;
        bool pushed = !(is_zero(s.push) && is_zero(s.turn));
        if (i == 0) {
            A.ball.vel = v; A.ball.angVel = w;
            if (pushed) A.ball.pos = s.pos + s.push * kTick;
        } else {
            CarDev& c = A.cars[i - 1];
            c.vel = v; c.angVel = w;
            if (pushed) {
                V3 np; M3 nbasis;
                integrate_transform(c.pos, c.basis, s.push, s.turn * 0.1f, kTick, np, nbasis);
                c.pos = np; c.basis = nbasis;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------- world step
// defined in sim_boxbox.cuh / sim_boxtri.cuh
GGL_HDN void collide_car_car(ArenaDev& A, ContactSink& S, int lo, int hi);
// cull in front of the out-of-line pair test (exact: boxes inside disjoint bounding spheres are disjoint, and btBoxBoxDetector
// has no margin; a disabled body's manifold is never solved)
GGL_HD bool car_car_may_touch(const CarDev& a, const CarDev& b) {
    if (a.rbDisabled || b.rbDisabled) return false;
    {   // the usual case first: the cars' origins alone are too far apart (the box centres sit |offset| from them)
        const float far = 2.0002f * len(car_box().full) + 2.f * len(car_hitbox_offset()) + 0.01f;
        if (len2(a.pos - b.pos) > far * far) return false;
    }
    const V3 pa = a.pos + mul(a.basis, car_hitbox_offset()), pb = b.pos + mul(b.basis, car_hitbox_offset());
    const float r2 = 2.0002f * len(car_box().full);
    return !(len2(pa - pb) > r2 * r2);
}
GGL_HDN void car_mesh_emit(ArenaDev& A, ContactSink& S, const WorldDev& W, int ci);
GGL_HD void car_car_callback(ArenaDev& A, int i1, int i2, V3 localA, V3 localB);

// Velocity update of a body that no contact row touches: what the solver's write-back does with zero delta impulses.
GGL_HD void body_free_velocity(ArenaDev& A, int b) {
    if (b == 0) {
        if (!A.ball.active) return;
        const float ii = ball_inv_inertia();
        M3 invI; invI.r0 = v3(ii, 0, 0); invI.r1 = v3(0, ii, 0); invI.r2 = v3(0, 0, ii);
        A.ball.vel = (A.ball.vel + v3(0, 0, 0)) + A.ball.force * kBallInvMass * kTick;
        A.ball.angVel = (A.ball.angVel + v3(0, 0, 0)) + tmul(invI, A.ball.torque) * kTick;
    } else {
        CarDev& c = A.cars[b - 1];
        if (c.rbDisabled) return;
        const float invMass = 1.f / rl::CAR_MASS;
        c.vel = (c.vel + v3(0, 0, 0)) + c.force * invMass * kTick;
        c.angVel = (c.angVel + v3(0, 0, 0)) + tmul(c.invInertiaW, c.torque) * kTick;
    }
}

// All contacts body `b` (0 = ball, 1 + c = car c) is the first body of, written into that body's windows.
GGL_HDN void collide_body(ArenaDev& A, const WorldDev& W, int b) {
    const float ex = rl::ARENA_EXTENT_X * kUU2BT, hz = rl::ARENA_HEIGHT * kUU2BT;
    if (b == 0) {
        ContactSink S = make_sink(A, A.contacts, 4, POOL_BALL_WORLD);
        if (A.ball.active) {
            GGL_SUB_BEGIN();
            collide_ball_mesh(A, S, W);
            GGL_SUB(4);
            collide_ball_plane(A, S, v3(0, 0, 1), v3(0, 0, 0));
            collide_ball_plane(A, S, v3(0, 0, -1), v3(0, 0, hz));
            collide_ball_plane(A, S, v3(1, 0, 0), v3(-ex, 0, hz / 2));
            collide_ball_plane(A, S, v3(-1, 0, 0), v3(ex, 0, hz / 2));
        }
        A.cntWorld = (uint8_t)S.n;
        return;
    }
    const int i = b - 1, P = A.numCars, R = contact_car_window(P);
    A.cars[i].ballHitNew = 0;
    ContactSink SB = make_sink(A, A.contacts, 0, (uint8_t)(POOL_CAR_BALL | i));
    ContactSink S = make_sink(A, A.contacts + 4 + i * R, R, (uint8_t)(POOL_CAR_REST | i));
    if (!A.cars[i].rbDisabled) {
        car_update_ray_aabb(A, i);
        GGL_SUB_BEGIN();
        collide_car_ball(A, SB, i);
        GGL_SUB(0);
        if (A.meshCandN[i]) car_mesh_emit(A, S, W, i);      // candidates and narrow phase ran in the two stages before (sim_tick.cuh)
        GGL_SUB(1);
        {
            // the four arena planes are axis-aligned: dot(n, boxPos - origin) is one component difference (the products with the
            // normal's 0 and 1 components add nothing), so the cull of collide_car_plane is evaluated here without the call
            const V3 boxPos = A.cars[i].pos + mul(A.cars[i].basis, car_hitbox_offset());
            const float lim = len(car_box().full) * 1.0001f, cbt = car_cbt();
            if (!((boxPos.z - 0.f) - lim > cbt)) collide_car_plane(A, S, i, v3(0, 0, 1), v3(0, 0, 0), boxPos);
            if (!(-(boxPos.z - hz) - lim > cbt)) collide_car_plane(A, S, i, v3(0, 0, -1), v3(0, 0, hz), boxPos);
            if (!((boxPos.x - (-ex)) - lim > cbt)) collide_car_plane(A, S, i, v3(1, 0, 0), v3(-ex, 0, hz / 2), boxPos);
            if (!(-(boxPos.x - ex) - lim > cbt)) collide_car_plane(A, S, i, v3(-1, 0, 0), v3(ex, 0, hz / 2), boxPos);
        }
        GGL_SUB(2);
        for (int j = i + 1; j < P; j++)
            if (car_car_may_touch(A.cars[i], A.cars[j])) collide_car_car(A, S, i, j);
        GGL_SUB(3);
    }
    A.cntBall[i] = (uint8_t)SB.n;
    A.cntRest[i] = (uint8_t)S.n;
}

// ---------------------------------------------------------------------------------------------- the reference's row order
// The order in which the reference's solver sees the contact rows of one island is not the order they were found in. Bullet
// hands the solver the dispatcher's manifold array sorted by island id with btAlignedObjectArray::quickSort
// (btSimulationIslandManager.cpp:370-384, predicate :177-184): an unstable sort, run over ALL manifolds of the tick - the ones
// without a contact point included - so the permutation of the manifolds that share an island id depends on how many manifolds
// there are and where they sit. Projected Gauss-Seidel is order-dependent, so whenever one island holds rows of two or more
// manifolds (a car that touches the ball while its hitbox rests on the floor ...) the order has to be reproduced:
//   1. broadphase boxes of the ball and the cars (btCollisionWorld::updateSingleAabb: the shape's box at the tick's transform,
//      grown by 0.02, merged with the box at the predicted transform) and their grid cell (btRSBroadphase.h:90-108, the cell of
//      the box's minimum corner);
//   2. the pair list of btRSBroadphase::calculateOverlappingPairs (:393-470; every pair is dropped and found again each tick):
//      per dynamic body in handle order (ball, then cars by id), the statics entered in its cell whose box overlaps, in creation
//      order, then the dynamic bodies entered in its cell (= whose cell is a neighbour) whose box overlaps and that have no pair
//      with it yet. The reference walks that last list in the order the bodies ARRIVED in the cell (history); handle order is
//      used here, which can only differ when a body has pairs with two or more other dynamic bodies;
//   3. the manifold array in dispatch order = pair order: the ball's manifolds with meshes and planes are made with the pair's
//      algorithm (constructors of btConvexConcave / btConvexPlaneCollisionAlgorithm); every manifold of a car - with a static,
//      the ball or another car - only if the hitbox's own box at the tick's transform overlaps the other shape's own box
//      (the child test of btCompoundCollisionAlgorithm.cpp:128-139; thresholds measured on the reference);
//   4. island ids: union-find over the dynamic-dynamic pairs in pair order, unite(p, q) hanging p's root under q's
//      (btUnionFind.h:76-98 with USE_PATH_COMPRESSION), element index = ball 0, cars 1.. by id; a manifold's id is that of its
//      dynamic body (:167-174); manifolds of a car without contact response (demolished) never reach the sort
//      (btCollisionDispatcher::needsResponse);
//   5. the quickSort itself (middle pivot, Hoare partition that swaps equal keys), on the array of ids;
//   6. rows re-emitted manifold by manifold, rows of one manifold keeping their order.
// If a row's manifold is not in the predicted array (the registry is an approximation at cell borders) the order is left alone.
constexpr int kMaxOrderManifolds = 48;
struct OrderManifold { int8_t dyn, other; };    // other: >= 0 dynamic body, < 0: -(1 + static)
// Working set of solver_order. It lives in the arena's car-vs-mesh hit staging (shared memory, dead between contact generation
// and the solver's bodies): as per-thread arrays every word would be a cold local-memory line of its own for the one lane that
// runs this.
struct OrderScratch {
    V3 mn[1 + kMaxCars], mx[1 + kMaxCars], tmn[1 + kMaxCars], tmx[1 + kMaxCars];     // proxy boxes, the shapes' own boxes
    OrderManifold mf[kMaxOrderManifolds];
    uint8_t key[kMaxOrderManifolds], idx[kMaxOrderManifolds], stLo[kMaxOrderManifolds], stHi[kMaxOrderManifolds];
    uint8_t rowM[kMaxContacts], out[kMaxContacts];
    uint8_t cell[1 + kMaxCars][4];
    uint8_t parent[8];
    int8_t first[8];
};
static_assert(sizeof(OrderScratch) <= sizeof(CarTriHit) * kMaxCars * kMaxMeshHit, "solver_order's working set must fit the mesh-hit staging");
GGL_HD OrderScratch& order_scratch(ArenaDev& A) { return *reinterpret_cast<OrderScratch*>(&A.meshHit[0][0]); }
#if defined(GGL_ORDER_DEBUG) && !defined(__CUDA_ARCH__)
// host test builds: the last predicted manifold array (tests/hostsim)
struct OrderDebug { int n; OrderManifold mf[kMaxOrderManifolds]; uint8_t key[kMaxOrderManifolds]; int reordered, unknownRow, always, off; V3 mn[1 + kMaxCars], mx[1 + kMaxCars]; };
inline OrderDebug& order_debug() { static thread_local OrderDebug d; return d; }
#endif

GGL_HD bool order_boxes_overlap(V3 amn, V3 amx, V3 bmn, V3 bmx) {
    return !(amn.x > bmx.x || amx.x < bmn.x || amn.y > bmx.y || amx.y < bmn.y || amn.z > bmx.z || amx.z < bmn.z);
}
// btAlignedObjectArray::quickSort restated on an array of keys: middle element as pivot, Hoare partition that also swaps equal
// keys, then both parts. The parts are disjoint, so the order in which they are taken up (the reference recurses, here a stack
// of pending ranges) does not change the result.
GGL_HD void order_sort_keys(uint8_t* key, uint8_t* idx, int n, uint8_t* stLo, uint8_t* stHi) {
    int sp = 0;
    stLo[0] = 0; stHi[0] = (uint8_t)(n - 1); sp = 1;
#pragma unroll 1
    while (sp > 0) {
        sp--;
        const int lo = stLo[sp], hi = stHi[sp];
        int i = lo, j = hi;
        const uint8_t x = key[(lo + hi) / 2];
#pragma unroll 1
        do {
#pragma unroll 1
            while (key[i] < x) i++;
#pragma unroll 1
            while (x < key[j]) j--;
            if (i <= j) {
                const uint8_t tk = key[i]; key[i] = key[j]; key[j] = tk;
                const uint8_t ti = idx[i]; idx[i] = idx[j]; idx[j] = ti;
                i++; j--;
            }
        } while (i <= j);
        if (lo < j) { stLo[sp] = (uint8_t)lo; stHi[sp] = (uint8_t)j; sp++; }
        if (i < hi) { stLo[sp] = (uint8_t)i; stHi[sp] = (uint8_t)hi; sp++; }
    }
}
// Builds the manifold array of steps 1-4 into mf[] / key[]; returns its length (0: too many manifolds to say). Written for
// SIZE: this is cold code in a kernel that is bound by instruction fetch - every 128-byte line of it is a miss that the whole
// block waits for at the next barrier (a first version of 34 KB cost 5 % of the env step while running in 0.08 % of the
// arena-ticks, tools/gpu_r2ay*.sh) - so loops are not unrolled, and a car's proxy box is the one car_update_ray_aabb left this
// tick when there is one (another car within 300 uu), else its own box moved by vel * dt: the turn of the predicted transform
// (<= 0.07 bt at the corners) is left out, which can only misplace a pair whose boxes are that close to touching.
GGL_HDN int order_predict_manifolds(const ArenaDev& A, const WorldDev& W, OrderScratch& S) {
    const int nb = 1 + A.numCars;           // body b has the b-th broadphase handle: ball, then cars by id (= index + 1)
    V3* mn = S.mn; V3* mx = S.mx; V3* tmn = S.tmn; V3* tmx = S.tmx;
    uint8_t (*cell)[4] = S.cell;
    OrderManifold* mf = S.mf;
    uint8_t* key = S.key;
    const V3 thr = v3(0.02f, 0.02f, 0.02f);
#pragma unroll 1
    for (int b = 0; b < nb; b++) {
        V3 a0, a1, step;
        bool tracked = false;
        if (b == 0) {
            const float re = kBallRadiusBT + 0.08f;        // btSphereShape::getAabb of the reference's Bullet adds 0.08 (btSphereShape.cpp:55)
            a0 = A.ball.pos - v3(re, re, re); a1 = A.ball.pos + v3(re, re, re);
            step = A.ball.vel * kTick;
        } else {
            const CarDev& c = A.cars[b - 1];
            car_world_aabb(c.pos, c.basis, a0, a1);
            step = c.vel * kTick;
            if (!(c.aabbMin.x > c.aabbMax.x) && !c.rbDisabled) { mn[b] = c.aabbMin; mx[b] = c.aabbMax; tracked = true; }
        }
        tmn[b] = a0; tmx[b] = a1;
        if (!tracked) { mn[b] = vmin(a0, a0 + step) - thr; mx[b] = vmax(a1, a1 + step) + thr; }
        const int ci = (int)((mn[b].x - W.bpMin.x) / W.bpCell), cj = (int)((mn[b].y - W.bpMin.y) / W.bpCell), ck = (int)((mn[b].z - W.bpMin.z) / W.bpCell);
        cell[b][0] = (uint8_t)(ci < 0 ? 0 : (ci >= W.bpX ? W.bpX - 1 : ci));
        cell[b][1] = (uint8_t)(cj < 0 ? 0 : (cj >= W.bpY ? W.bpY - 1 : cj));
        cell[b][2] = (uint8_t)(ck < 0 ? 0 : (ck >= W.bpZ ? W.bpZ - 1 : ck));
    }
#if defined(GGL_ORDER_DEBUG) && !defined(__CUDA_ARCH__)
    for (int b = 0; b < nb; b++) { order_debug().mn[b] = mn[b]; order_debug().mx[b] = mx[b]; }
#endif
    uint8_t* parent = S.parent;
#pragma unroll 1
    for (int r = 0; r < nb; r++) parent[r] = (uint8_t)r;
    int n = 0;
#pragma unroll 1
    for (int p = 0; p < nb; p++) {
        const bool response = p == 0 || !A.cars[p - 1].rbDisabled;
        uint32_t bits = W.bpCellStatics[(cell[p][0] * W.bpY + cell[p][1]) * W.bpZ + cell[p][2]];
        if (p == 0 && !A.ball.active) bits = 0;      // a sleeping ball against a static (asleep by construction): no algorithm (btCollisionDispatcher::needsCollision)
        if (!response) bits = 0;                      // a demolished car's manifolds never reach the sort
#pragma unroll 1
        while (bits) {
#if defined(__CUDA_ARCH__)
            const int s = __ffs((int)bits) - 1;
#else
            const int s = __builtin_ctz(bits);
#endif
            bits &= bits - 1;
            // bpStaticBox holds the statics' own boxes; their proxies are 0.02 larger. The ball's manifold is made with the pair
            // (proxy boxes overlap); a car's, by the compound algorithm, only if the hitbox's tight box at the tick's transform
            // overlaps the static's own box (btCompoundCollisionAlgorithm.cpp:128-139; measured on the reference: a car box
            // 0.1996 above the floor has a (car, floor) manifold, one 0.2046 above has none, whatever its velocity)
            const float* bx = W.bpStaticBox + 6 * s;
            const V3 smn = v3(bx[0], bx[1], bx[2]), smx = v3(bx[3], bx[4], bx[5]);
            const V3 lo = p == 0 ? mn[0] - thr : tmn[p], hi = p == 0 ? mx[0] + thr : tmx[p];
            if (!order_boxes_overlap(lo, hi, smn, smx)) continue;
            if (n >= kMaxOrderManifolds) return 0;
            mf[n].dyn = (int8_t)p; mf[n].other = (int8_t)(-(1 + s)); n++;
        }
        // a body with a lower handle found its pair with p when it was walked (the conditions are symmetric)
#pragma unroll 1
        for (int q = p + 1; q < nb; q++) {
            if (abs((int)cell[q][0] - (int)cell[p][0]) > 1 || abs((int)cell[q][1] - (int)cell[p][1]) > 1 || abs((int)cell[q][2] - (int)cell[p][2]) > 1) continue;
            if (!order_boxes_overlap(mn[p], mx[p], mn[q], mx[q])) continue;
            if (!response || A.cars[q - 1].rbDisabled) continue;     // no island merge, and no manifold that reaches the sort
            int i = p, j = q;
            while (parent[i] != i) i = parent[i];
            while (parent[j] != j) j = parent[j];
            if (i != j) parent[i] = (uint8_t)j;
            if (!order_boxes_overlap(tmn[q], tmx[q], tmn[p], tmx[p])) continue;      // compound child test on the bodies' own boxes
            if (n >= kMaxOrderManifolds) return 0;
            mf[n].dyn = (int8_t)p; mf[n].other = (int8_t)q; n++;
        }
    }
#pragma unroll 1
    for (int k = 0; k < n; k++) {
        int i = mf[k].dyn;
        while (parent[i] != i) i = parent[i];
        key[k] = (uint8_t)i;
    }
    return n;
}

// Leader: turns windows + pool into the solver's row order (A.contactOrder, nothing is moved), runs the order-dependent
// contact callbacks (ball extra impulse in car order, car-car bump/demo logic per point), the ball wake-up rule, and marks the
// bodies that have a row. Row order = the reference's (pair-cache insertion order): the ball's world rows, the
// car-ball rows in car order, then every car's world + car-car rows; within a body, window rows before pool rows, pool rows in
// claim order (a body's lane claims them one after the other, so claim order is its emission order). That is the order of the
// reference's dispatcher array; when an island holds rows of more than one manifold the solver's order is another one, and
// solver_order (above) re-emits the list.
GGL_HD int pool_next(const ArenaDev& A, uint8_t owner, int& cursor) {
    const int used = (int)A.poolCount < contact_pool_size(A.numCars) ? (int)A.poolCount : contact_pool_size(A.numCars);
    while (cursor < used && A.poolOwner[cursor] != owner) cursor++;
    return cursor < used ? contact_pool_base(A.numCars) + cursor++ : -1;
}

// The manifold a row belongs to, as an OrderManifold (the (ball, car) manifold is listed under the ball)
GGL_HD OrderManifold order_row_manifold(const Contact& c, int planeBase) {
    OrderManifold m;
    if (c.bodyB < 0) {
        const int s = c.staticId > 0 ? c.staticId - 1 : planeBase + (-c.staticId - 1);
        m.dyn = (int8_t)c.bodyA; m.other = (int8_t)(-(1 + s));
    } else if (c.bodyB == 0) { m.dyn = 0; m.other = (int8_t)c.bodyA; }
    else { m.dyn = (int8_t)c.bodyA; m.other = (int8_t)c.bodyB; }
    return m;
}

// Steps 5 and 6: re-emits A.contactOrder in the reference's order. Leader, after the islands are labelled.
GGL_HDN void solver_order(ArenaDev& A, const WorldDev& W) {
    OrderScratch& S = order_scratch(A);
    {   // only if two rows of ONE island belong to different manifolds
        bool mixed = false;
        int8_t* first = S.first;
#pragma unroll 1
        for (int b = 0; b < 8; b++) first[b] = -1;
#pragma unroll 1
        for (int k = 0; k < A.numContacts; k++) {
            const Contact& c = A.contacts[A.contactOrder[k]];
            const int isl = A.island[c.bodyA];
            if (first[isl] < 0) { first[isl] = (int8_t)k; continue; }
            const Contact& f = A.contacts[A.contactOrder[first[isl]]];
            mixed |= f.bodyA != c.bodyA || f.bodyB != c.bodyB || f.staticId != c.staticId;
        }
#if defined(GGL_ORDER_DEBUG) && !defined(__CUDA_ARCH__)
        if (order_debug().always) mixed = true;
#endif
#if defined(GGL_ORDER_COUNT) && defined(__CUDA_ARCH__)      // experiment: calls in the high half, runs in the low half of the overflow counter
        atomicAdd(&A.contactOverflow, mixed ? 0x10001u : 0x10000u);
#endif
        if (!mixed) return;
    }
#if !defined(GGL_ORDER_COST)
#define GGL_ORDER_COST 24u
#endif
    A.stepCost += GGL_ORDER_COST;      // work estimate for the next step's block balancing (sim_tick.cuh): about three contact rows' worth, on the leader alone
    OrderManifold* mf = S.mf;
    uint8_t* key = S.key; uint8_t* idx = S.idx; uint8_t* rowM = S.rowM; uint8_t* out = S.out;
    const int n = order_predict_manifolds(A, W, S);
#if defined(GGL_ORDER_DEBUG) && !defined(__CUDA_ARCH__)
    OrderDebug& D = order_debug();
    D.n = n; D.reordered = 0; D.unknownRow = 0;
#pragma unroll 1
    for (int k = 0; k < n; k++) { D.mf[k] = mf[k]; D.key[k] = key[k]; }
#endif
    if (n < 2) return;
    const int planeBase = W.numStatics - 4;
#pragma unroll 1
    for (int k = 0; k < A.numContacts; k++) {
        const OrderManifold r = order_row_manifold(A.contacts[A.contactOrder[k]], planeBase);
        int found = -1;
#pragma unroll 1
        for (int q = 0; q < n; q++)
            if ((mf[q].dyn == r.dyn && mf[q].other == r.other) || (r.other > 0 && mf[q].dyn == r.other && mf[q].other == r.dyn)) { found = q; break; }
        if (found < 0) {
#if defined(GGL_ORDER_DEBUG) && !defined(__CUDA_ARCH__)
            D.unknownRow = 1;
#endif
            return;
        }
        rowM[k] = (uint8_t)found;
    }
#pragma unroll 1
    for (int q = 0; q < n; q++) idx[q] = (uint8_t)q;
    order_sort_keys(key, idx, n, S.stLo, S.stHi);
    int w = 0;
#pragma unroll 1
    for (int q = 0; q < n; q++)
#pragma unroll 1
        for (int k = 0; k < A.numContacts; k++)
            if (rowM[k] == idx[q]) out[w++] = A.contactOrder[k];
#if defined(GGL_ORDER_DEBUG) && !defined(__CUDA_ARCH__)
#pragma unroll 1
    for (int k = 0; k < w; k++) D.reordered |= out[k] != A.contactOrder[k];
#endif
#pragma unroll 1
    for (int k = 0; k < w; k++) A.contactOrder[k] = out[k];
}

GGL_HDN void gather_contacts(ArenaDev& A, const WorldDev& W) {
    const int P = A.numCars, R = contact_car_window(P);
    {   // nothing touched anything this tick (the usual case: 88 % of the arena-ticks of the benchmark's rollout)
        int any = A.cntWorld;
        for (int i = 0; i < P; i++) any += A.cntBall[i] + A.cntRest[i];
        if (any == 0) {
            A.numContacts = 0; A.bodiesInContact = 0; A.ballAwake = A.ball.active ? 1 : 0;
            for (int b = 0; b < 8; b++) A.island[b] = (uint8_t)b;
            return;
        }
    }
    int n = 0;
    {
        int cur = 0;
        for (int k = 0; k < A.cntWorld; k++) {
            const int row = k < 4 ? k : pool_next(A, POOL_BALL_WORLD, cur);
            if (row >= 0) A.contactOrder[n++] = (uint8_t)row;
        }
    }
    for (int i = 0; i < P; i++)
        if (A.cntBall[i]) {
            int cur = 0;
            const int row = pool_next(A, (uint8_t)(POOL_CAR_BALL | i), cur);
            if (row >= 0) A.contactOrder[n++] = (uint8_t)row;
            if (A.cars[i].ballHitNew & 1) A.ball.velImpulseCache += A.cars[i].hitExtraVel * kUU2BT;
            if ((A.cars[i].ballHitNew & 2) && A.gameMode == kModeHeatseeker) ball_heatseeker_on_hit(A.ball, A.cars[i].team);
        }
    for (int i = 0; i < P; i++) {
        int cur = 0;
        for (int k = 0; k < A.cntRest[i]; k++) {
            const int row = k < R ? 4 + i * R + k : pool_next(A, (uint8_t)(POOL_CAR_REST | i), cur);
            if (row < 0) continue;
            A.contactOrder[n++] = (uint8_t)row;
            const Contact& c = A.contacts[row];
            if (c.bodyA > 0 && c.bodyB > 0) {
                const CarDev& a = A.cars[c.bodyA - 1];
                const CarDev& b = A.cars[c.bodyB - 1];
                car_car_callback(A, c.bodyA - 1, c.bodyB - 1, tmul(a.basis, c.relA), tmul(b.basis, c.relB));
            }
        }
    }
    A.numContacts = n;
    bool ballAwake = A.ball.active != 0;
    if (!ballAwake) {
        for (int k = 0; k < n; k++)
            if (A.contacts[A.contactOrder[k]].bodyB == 0) ballAwake = true;      // touched by an active car: the island wakes it up
        if (!ballAwake) {
            int w = 0;
            for (int k = 0; k < n; k++)
                if (A.contacts[A.contactOrder[k]].bodyA != 0) A.contactOrder[w++] = A.contactOrder[k];
            A.numContacts = w;
        }
    }
    A.ballAwake = ballAwake ? 1 : 0;
    uint32_t mask = 0;
    // Can one island hold rows of two manifolds (solver_order)? Only if a body touches two different statics (its world rows are
    // contiguous) or a row joins two dynamic bodies and is not alone. Decided here, in the loop that reads the rows anyway:
    // this is the leader's serial path of every tick with a contact, and an instruction costs ~24 cycles at this kernel's
    // issue rate (a 400-instruction test here measured 4 % of the env step, tools/gpu_r2ay.sh).
    bool twoStatics = false, anyDyn = false;
    int prevA = -1, prevS = 0;
    for (int k = 0; k < A.numContacts; k++) {
        const Contact& c = A.contacts[A.contactOrder[k]];
        mask |= 1u << c.bodyA;
        if (c.bodyB >= 0) { mask |= 1u << c.bodyB; anyDyn = true; }
        else { twoStatics |= c.bodyA == prevA && c.staticId != prevS; prevA = c.bodyA; prevS = c.staticId; }
    }
    bool several = twoStatics || (anyDyn && A.numContacts >= 2);
    A.bodiesInContact = mask;
    // contact islands: bodies connected through rows between two dynamic bodies (car-ball, car-car) share a label
    for (int b = 0; b < 8; b++) A.island[b] = (uint8_t)b;
    for (int pass = 0; pass < kMaxCars; pass++) {
        bool changed = false;
        for (int k = 0; k < A.numContacts; k++) {
            const Contact& c = A.contacts[A.contactOrder[k]];
            if (c.bodyB < 0) continue;
            const uint8_t la = A.island[c.bodyA], lb = A.island[c.bodyB];
            if (la != lb) { const uint8_t m = la < lb ? la : lb; A.island[c.bodyA] = m; A.island[c.bodyB] = m; changed = true; }
        }
        if (!changed) break;
    }
    // rows of more than one manifold: if two of them share an island, the reference's order is not the order of discovery
#if defined(GGL_ORDER_DEBUG) && !defined(__CUDA_ARCH__)
    order_debug().n = -1;
    if (order_debug().always) several = true;
    if (order_debug().off) several = false;
#endif
#if !defined(GGL_NO_SOLVER_ORDER)       // experiment builds only (tools/gpu_r2ay.sh: what the reordering costs)
    if (several && W.bpCellStatics) solver_order(A, W);
#endif
}

}  // namespace ggl
