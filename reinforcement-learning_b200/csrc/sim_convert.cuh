// Host-side conversion between the C-ABI state structs (Unreal units, include/gigaenv.h) and the device arena blob
// (Bullet units), plus the triangle-grid builder for the arena meshes.  The float operations are the ones the
// reference performs in Car::SetState/GetState (Sim/Car/Car.cpp:10-36) and Ball::SetState/GetState
// (Sim/Ball/Ball.cpp:27-49): pos * UU_TO_BT on the way in, * BT_TO_UU on the way out.
#pragma once
#include <algorithm>
#include <string.h>
#include <vector>
#include "../../include/gigaenv.h"
#include "sim_tick.cuh"

namespace ggl {

inline V3 v3p(const float* f) { return v3(f[0], f[1], f[2]); }
inline void put3(float* o, V3 v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }

inline Controls controls_from_host(const ggl_controls& k) {
    Controls c;
    c.throttle = k.throttle; c.steer = k.steer; c.pitch = k.pitch; c.yaw = k.yaw; c.roll = k.roll;
    c.jump = k.jump; c.boost = k.boost; c.handbrake = k.handbrake; c._pad = 0;
    return c;
}
inline ggl_controls controls_to_host(const Controls& c) {
    ggl_controls k;
    k.throttle = c.throttle; k.steer = c.steer; k.pitch = c.pitch; k.yaw = c.yaw; k.roll = c.roll;
    k.jump = c.jump; k.boost = c.boost; k.handbrake = c.handbrake; k._pad = 0;
    return k;
}

inline void car_from_host(const ggl_car_state& s, CarDev& c, int index, int numCars) {
    memset(&c, 0, sizeof(c));
    c.pos = v3p(s.pos) * kUU2BT;
    // host rot = forward, right, up (columns of the basis)
    V3 f = v3p(s.rot), r = v3p(s.rot + 3), u = v3p(s.rot + 6);
    c.basis.r0 = v3(f.x, r.x, u.x); c.basis.r1 = v3(f.y, r.y, u.y); c.basis.r2 = v3(f.z, r.z, u.z);
    c.vel = v3p(s.vel) * kUU2BT;
    c.angVel = v3p(s.ang_vel);
    car_update_inertia(c);
    c.flipRelTorque = v3p(s.flip_rel_torque);
    c.worldContactNormal = v3p(s.world_contact_normal);
    c.jumpTime = s.jump_time; c.flipTime = s.flip_time; c.airTime = s.air_time; c.airTimeSinceJump = s.air_time_since_jump;
    c.boost = s.boost; c.timeSpentBoosting = s.time_spent_boosting; c.supersonicTime = s.supersonic_time;
    c.handbrakeVal = s.handbrake_val; c.autoFlipTimer = s.auto_flip_timer; c.autoFlipTorqueScale = s.auto_flip_torque_scale;
    c.carContactCooldown = s.car_contact_cooldown; c.demoRespawnTimer = s.demo_respawn_timer;
    c.hitRelPos = v3p(s.hit_rel_pos); c.hitBallPos = v3p(s.hit_ball_pos); c.hitExtraVel = v3p(s.hit_extra_vel);
    c.hitTick = s.hit_tick; c.hitTickExtra = s.hit_tick_extra; c.updateCounter = s.update_counter;
    c.aabbMin = v3p(s.ray_aabb_min); c.aabbMax = v3p(s.ray_aabb_max);       // all zero = unknown (ray_vs_car_box)
    c.carContactOtherId = s.car_contact_other_id;
    c.id = (uint32_t)(index + 1);
    c.team = (index < numCars / 2) ? 0 : 1;
    c.isOnGround = s.is_on_ground; c.hasJumped = s.has_jumped; c.hasDoubleJumped = s.has_double_jumped;
    c.hasFlipped = s.has_flipped; c.isFlipping = s.is_flipping; c.isJumping = s.is_jumping;
    c.isSupersonic = s.is_supersonic; c.isAutoFlipping = s.is_auto_flipping; c.worldContactHas = s.world_contact_has;
    c.isDemoed = s.is_demoed; c.hitValid = s.hit_valid;
    if (c.isDemoed) car_freeze_reported_basis(c);      // an injected demolished car reports the orientation it was given
    for (int i = 0; i < 4; i++) c.wheelsWithContact[i] = s.wheels_with_contact[i];
    c.controls = controls_from_host(s.controls);
    c.lastControls = controls_from_host(s.last_controls);
    for (int i = 0; i < 4; i++) {
        Wheel& w = c.w[i];
        w.steerAngle = s.wheels[i].steer_angle; w.engineForce = s.wheels[i].engine_force; w.brake = s.wheels[i].brake;
        w.latFriction = s.wheels[i].lat_friction; w.longFriction = s.wheels[i].long_friction;
        w.extraPushback = s.wheels[i].extra_pushback;
        w.ground = GROUND_NONE;
    }
}

inline void car_to_host(const CarDev& c, ggl_car_state& s) {
    memset(&s, 0, sizeof(s));
    put3(s.pos, c.pos * kBT2UU);
    const M3 rb = car_reported_basis(c);
    put3(s.rot, col(rb, 0)); put3(s.rot + 3, col(rb, 1)); put3(s.rot + 6, col(rb, 2));
    put3(s.vel, c.vel * kBT2UU);
    put3(s.ang_vel, c.angVel);
    put3(s.flip_rel_torque, c.flipRelTorque);
    put3(s.world_contact_normal, c.worldContactNormal);
    s.jump_time = c.jumpTime; s.flip_time = c.flipTime; s.air_time = c.airTime; s.air_time_since_jump = c.airTimeSinceJump;
    s.boost = c.boost; s.time_spent_boosting = c.timeSpentBoosting; s.supersonic_time = c.supersonicTime;
    s.handbrake_val = c.handbrakeVal; s.auto_flip_timer = c.autoFlipTimer; s.auto_flip_torque_scale = c.autoFlipTorqueScale;
    s.car_contact_cooldown = c.carContactCooldown; s.demo_respawn_timer = c.demoRespawnTimer;
    put3(s.hit_rel_pos, c.hitRelPos); put3(s.hit_ball_pos, c.hitBallPos); put3(s.hit_extra_vel, c.hitExtraVel);
    s.hit_tick = c.hitTick; s.hit_tick_extra = c.hitTickExtra; s.update_counter = c.updateCounter;
    put3(s.ray_aabb_min, c.aabbMin); put3(s.ray_aabb_max, c.aabbMax);
    s.car_contact_other_id = c.carContactOtherId; s.id = c.id; s.team = c.team;
    s.is_on_ground = c.isOnGround; s.has_jumped = c.hasJumped; s.has_double_jumped = c.hasDoubleJumped;
    s.has_flipped = c.hasFlipped; s.is_flipping = c.isFlipping; s.is_jumping = c.isJumping;
    s.is_supersonic = c.isSupersonic; s.is_auto_flipping = c.isAutoFlipping; s.world_contact_has = c.worldContactHas;
    s.is_demoed = c.isDemoed; s.hit_valid = c.hitValid;
    for (int i = 0; i < 4; i++) s.wheels_with_contact[i] = c.wheelsWithContact[i];
    s.controls = controls_to_host(c.controls);
    s.last_controls = controls_to_host(c.lastControls);
    for (int i = 0; i < 4; i++) {
        const Wheel& w = c.w[i];
        s.wheels[i].steer_angle = w.steerAngle; s.wheels[i].engine_force = w.engineForce; s.wheels[i].brake = w.brake;
        s.wheels[i].lat_friction = w.latFriction; s.wheels[i].long_friction = w.longFriction;
        s.wheels[i].extra_pushback = w.extraPushback;
    }
}

// Physical state only; env-layer bookkeeping (GameState, trackers) is left untouched.
inline void arena_phys_from_host(const ggl_arena_state& st, ArenaDev& A) {
    A.tickCount = st.tick_count;
    A.numCars = st.num_cars;
    BallDev& b = A.ball;
    memset(&b, 0, sizeof(b));
    b.pos = v3p(st.ball.pos) * kUU2BT;
    b.vel = v3p(st.ball.vel) * kUU2BT;
    b.angVel = v3p(st.ball.ang_vel);
    b.updateCounter = st.ball.update_counter;
    b.hsTargetDir = st.ball.hs_y_target_dir; b.hsTargetSpeed = st.ball.hs_cur_target_speed; b.hsTimeSinceHit = st.ball.hs_time_since_hit;
    for (int i = 0; i < st.num_cars && i < kMaxCars; i++) car_from_host(st.cars[i], A.cars[i], i, st.num_cars);
    for (int p = 0; p < kNumPads; p++) {
        A.pads[p].isActive = st.pads[p].is_active;
        A.pads[p].cooldown = st.pads[p].cooldown;
        A.pads[p].prevLockedCarId = (uint8_t)st.pads[p].prev_locked_car_id;
        A.pads[p].curLockedCarId = 0;
    }
}

inline void arena_phys_to_host(const ArenaDev& A, ggl_arena_state& st) {
    memset(&st, 0, sizeof(st));
    st.tick_count = A.tickCount;
    st.num_cars = A.numCars;
    put3(st.ball.pos, A.ball.pos * kBT2UU);
    put3(st.ball.vel, A.ball.vel * kBT2UU);
    put3(st.ball.ang_vel, A.ball.angVel);
    st.ball.update_counter = A.ball.updateCounter;
    st.ball.hs_y_target_dir = A.ball.hsTargetDir; st.ball.hs_cur_target_speed = A.ball.hsTargetSpeed; st.ball.hs_time_since_hit = A.ball.hsTimeSinceHit;
    for (int i = 0; i < A.numCars && i < kMaxCars; i++) car_to_host(A.cars[i], st.cars[i]);
    for (int p = 0; p < kNumPads; p++) {
        st.pads[p].is_active = A.pads[p].isActive;
        st.pads[p].cooldown = A.pads[p].cooldown;
        st.pads[p].prev_locked_car_id = A.pads[p].prevLockedCarId;
    }
}

// ---------------------------------------------------------------------------------------------- arena meshes
struct WorldHost {
    std::vector<Tri> tris;
    std::vector<int> cellStart, cellTris, meshStart;   // meshStart[m] = one past the last triangle of mesh m
    std::vector<uint32_t> bpCellStatics;               // WorldDev::bpCellStatics / bpStaticBox
    std::vector<float> bpStaticBox;
    int bpX = 0, bpY = 0, bpZ = 0, numStatics = 0;
    V3 bpMin = v3(0, 0, 0);
    float bpCell = 7.4f;
    int numMeshes = 0;
    int gx = 1, gy = 1, gz = 1;
    V3 gridMin = v3(0, 0, 0);
    float cellSize = 1.f;
    V3 freeMin = v3(0, 0, 0), freeMax = v3(0, 0, 0);      // empty-cell box (WorldDev::freeMin / freeMax); empty itself when there is none
};

// .cmf layout (CollisionMeshFile.cpp:11-35): int32 numTris, int32 numVerts, int32 tri[numTris][3], float vert[numVerts][3]
inline bool world_add_cmf(WorldHost& W, const uint8_t* data, size_t size) {
    if (size < 8) return false;
    int32_t nt, nv;
    memcpy(&nt, data, 4); memcpy(&nv, data + 4, 4);
    if (nt < 0 || nv < 0 || size < 8 + (size_t)nt * 12 + (size_t)nv * 12) return false;
    const int32_t* idx = (const int32_t*)(data + 8);
    const float* vert = (const float*)(data + 8 + (size_t)nt * 12);
    for (int i = 0; i < nt; i++) {
        int a = idx[3 * i], b = idx[3 * i + 1], c = idx[3 * i + 2];
        if (a < 0 || b < 0 || c < 0 || a >= nv || b >= nv || c >= nv) return false;
        Tri t;
        t.v0 = v3p(vert + 3 * a); t.v1 = v3p(vert + 3 * b); t.v2 = v3p(vert + 3 * c);
        t.edgeAngle[0] = t.edgeAngle[1] = t.edgeAngle[2] = 6.283185307179586232f;      // btTriangleInfo defaults (SIMD_2_PI, flags 0)
        t.flags = (uint32_t)W.numMeshes << TRI_MESH_SHIFT;
        W.tris.push_back(t);
    }
    W.meshStart.push_back((int)W.tris.size());
    W.numMeshes++;
    return true;
}

// btGenerateInternalEdgeInfo + btConnectivityProcessor::processTriangle (btInternalEdgeUtility.cpp:47-354): for every triangle,
// every other triangle of the SAME mesh sharing exactly two vertices (by position, threshold 1e-4^2 on the squared distance)
// defines the dihedral angle / convexity of the shared edge. The reference visits the neighbours through the mesh BVH; the
// visiting order only matters for non-manifold edges (last neighbour wins), index order is used here.
inline V3 tri_calc_normal(V3 a, V3 b, V3 c) { return normalized(cross(b - a, c - a)); }
inline void world_build_edge_info(WorldHost& W) {
    const float eqThresh = 0.0001f * 0.0001f, planarEps = 0.0001f;
    for (int mesh = 0; mesh < W.numMeshes; mesh++) {
        const int lo = mesh == 0 ? 0 : W.meshStart[mesh - 1], hi = W.meshStart[mesh];
        for (int ia = lo; ia < hi; ia++) {
            Tri& A = W.tris[ia];
            const V3 va[3] = {A.v0, A.v1, A.v2};
            V3 amin = vmin(va[0], vmin(va[1], va[2])), amax = vmax(va[0], vmax(va[1], va[2]));
            for (int ib = lo; ib < hi; ib++) {
                if (ib == ia) continue;
                const Tri& B = W.tris[ib];
                const V3 vb[3] = {B.v0, B.v1, B.v2};
                V3 bmin = vmin(vb[0], vmin(vb[1], vb[2])), bmax = vmax(vb[0], vmax(vb[1], vb[2]));
                if (bmin.x > amax.x || bmax.x < amin.x || bmin.y > amax.y || bmax.y < amin.y || bmin.z > amax.z || bmax.z < amin.z) continue;
                if (len2(cross(vb[1] - vb[0], vb[2] - vb[0])) < eqThresh) continue;
                if (len2(cross(va[1] - va[0], va[2] - va[0])) < eqThresh) continue;
                int numshared = 0, sA[3] = {-1, -1, -1}, sB[3] = {-1, -1, -1};
                bool dup = false;
                for (int i = 0; i < 3 && !dup; i++)
                    for (int j = 0; j < 3; j++)
                        if (len2(va[i] - vb[j]) < eqThresh) {
                            sA[numshared] = i; sB[numshared] = j; numshared++;
                            if (numshared >= 3) { dup = true; break; }
                        }
                if (dup || numshared != 2) continue;
                if (sA[0] == 0 && sA[1] == 2) { sA[0] = 2; sA[1] = 0; int t = sB[1]; sB[1] = sB[0]; sB[0] = t; }
                A.flags |= TRI_HAS_INFO;
                const int sumA = sA[0] + sA[1], otherA = 3 - sumA, otherB = 3 - (sB[0] + sB[1]);
                V3 edge = va[sA[1]] - va[sA[0]];
                V3 normalA = tri_calc_normal(va[0], va[1], va[2]);
                V3 normalB = tri_calc_normal(vb[sB[1]], vb[sB[0]], vb[otherB]);
                edge = normalized(edge);
                V3 edgeCrossA = normalized(cross(edge, normalA));
                if (dot(edgeCrossA, va[otherA] - va[sA[0]]) < 0.f) edgeCrossA = edgeCrossA * -1.f;
                V3 edgeCrossB = normalized(cross(edge, normalB));
                if (dot(edgeCrossB, vb[otherB] - vb[sB[0]]) < 0.f) edgeCrossB = edgeCrossB * -1.f;
                float ang4 = 0.f, corrected = 0.f;
                bool isConvex = false;
                V3 calcEdge = cross(edgeCrossA, edgeCrossB);
                if (!(len2(calcEdge) < planarEps)) {
                    calcEdge = normalized(calcEdge);
                    V3 calcNormalA = normalized(cross(calcEdge, edgeCrossA));
                    float angle2 = edge_get_angle(calcNormalA, edgeCrossA, edgeCrossB);
                    ang4 = kPi - angle2;
                    isConvex = dot(normalA, edgeCrossB) < 0.f;
                    corrected = isConvex ? ang4 : -ang4;
                }
                int e; V3 ev;
                if (sumA == 1) { e = 0; ev = va[0] - va[1]; }
                else if (sumA == 2) { e = 2; ev = va[2] - va[0]; }
                else { e = 1; ev = va[1] - va[2]; }
                V3 computedB = quat_rotate(quat_axis_angle(ev, -corrected), normalA);
                if (dot(computedB, normalB) < 0.f) A.flags |= (TRI_SWAP_V0V1 << e);
                A.edgeAngle[e] = -corrected;
                if (isConvex) A.flags |= (TRI_CONVEX_V0V1 << e);
            }
        }
    }
}

// The order in which the reference meets the triangles of ONE mesh: a box-vs-mesh or sphere-vs-mesh query walks the mesh's
// quantized BVH (btBvhTriangleMeshShape(mesh, useQuantizedAabbCompression = true), RocketSim.cpp:167) and reports leaves in
// the order they have in the node array, i.e. the order the leaf array is left in by the recursive partition of
// btQuantizedBvh::buildTree (BT/BulletCollision/BroadphaseCollision/btQuantizedBvh.cpp:116-305) on the quantized triangle
// boxes built by btOptimizedBvh::build (BT/BulletCollision/CollisionShapes/btOptimizedBvh.cpp:100-170, quantisation
// btQuantizedBvh.cpp:76-105 + btQuantizedBvh.h:331-358). That order decides the row order inside a multi-triangle manifold and
// which contact normal a car's worldContact keeps (the last one), so the triangles of every mesh are STORED in it: index order
// on the device is then the reference's visiting order. Returns the permutation (new position -> old index within the mesh).
struct BvhOrder {
    V3 bmin, bmax, quant;
    std::vector<uint16_t> qmin, qmax;     // [leaf][3]
    std::vector<int> tri;                 // leaf -> triangle index within the mesh
    void quantize(uint16_t* out, V3 p, bool isMax) const {
        V3 v = v3((p.x - bmin.x) * quant.x, (p.y - bmin.y) * quant.y, (p.z - bmin.z) * quant.z);
        if (isMax) {
            out[0] = (uint16_t)((uint16_t)(v.x + 1.f) | 1); out[1] = (uint16_t)((uint16_t)(v.y + 1.f) | 1); out[2] = (uint16_t)((uint16_t)(v.z + 1.f) | 1);
        } else {
            out[0] = (uint16_t)((uint16_t)(v.x) & 0xfffe); out[1] = (uint16_t)((uint16_t)(v.y) & 0xfffe); out[2] = (uint16_t)((uint16_t)(v.z) & 0xfffe);
        }
    }
    V3 unquantize(const uint16_t* q) const {
        V3 v = v3((float)q[0] / quant.x, (float)q[1] / quant.y, (float)q[2] / quant.z);
        return v + bmin;
    }
    void set_quantization(V3 aabbMin, V3 aabbMax) {
        const V3 clamp = v3(1.f, 1.f, 1.f), k = v3(65533.f, 65533.f, 65533.f);
        bmin = aabbMin - clamp; bmax = aabbMax + clamp;
        V3 size = bmax - bmin;
        quant = v3(k.x / size.x, k.y / size.y, k.z / size.z);
        uint16_t q[3];
        quantize(q, bmin, false);
        V3 v = unquantize(q);
        bmin = vmin(bmin, v - clamp);
        size = bmax - bmin; quant = v3(k.x / size.x, k.y / size.y, k.z / size.z);
        quantize(q, bmax, true);
        v = unquantize(q);
        bmax = vmax(bmax, v + clamp);
        size = bmax - bmin; quant = v3(k.x / size.x, k.y / size.y, k.z / size.z);
    }
    V3 centre(int i) const { return (unquantize(&qmax[3 * i]) + unquantize(&qmin[3 * i])) * 0.5f; }
    void swap_leaves(int a, int b) {
        for (int k = 0; k < 3; k++) { std::swap(qmin[3 * a + k], qmin[3 * b + k]); std::swap(qmax[3 * a + k], qmax[3 * b + k]); }
        std::swap(tri[a], tri[b]);
    }
    void build(int start, int end) {
        const int n = end - start;
        if (n == 1) return;
        // calcSplittingAxis
        V3 means = v3(0, 0, 0), var = v3(0, 0, 0);
        for (int i = start; i < end; i++) means += centre(i);
        means *= (1.f / (float)n);
        for (int i = start; i < end; i++) { V3 d = centre(i) - means; var += v3(d.x * d.x, d.y * d.y, d.z * d.z); }
        var *= (1.f / ((float)n - 1.f));
        const int axis = var.x < var.y ? (var.y < var.z ? 2 : 1) : (var.x < var.z ? 2 : 0);        // btVector3::maxAxis
        // sortAndCalcSplittingIndex
        V3 m2 = v3(0, 0, 0);
        for (int i = start; i < end; i++) m2 += centre(i);
        m2 *= (1.f / (float)n);
        const float splitValue = comp(m2, axis);
        int split = start;
        for (int i = start; i < end; i++)
            if (comp(centre(i), axis) > splitValue) { swap_leaves(i, split); split++; }
        const int range = n / 3;
        if (split <= start + range || split >= end - 1 - range) split = start + (n >> 1);
        build(start, split);
        build(split, end);
    }
};

inline std::vector<int> bvh_leaf_order(const Tri* t, int n) {
    BvhOrder B;
    // btTriangleMeshShape::recalcLocalAabb (btTriangleMeshShape.cpp:58-70): vertex extents, mesh margin 0
    V3 mn = v3(1e18f, 1e18f, 1e18f), mx = v3(-1e18f, -1e18f, -1e18f);
    for (int i = 0; i < n; i++) {
        mn = vmin(mn, vmin(t[i].v0, vmin(t[i].v1, t[i].v2)));
        mx = vmax(mx, vmax(t[i].v0, vmax(t[i].v1, t[i].v2)));
    }
    B.set_quantization(mn, mx);
    B.qmin.resize(3 * n); B.qmax.resize(3 * n); B.tri.resize(n);
    for (int i = 0; i < n; i++) {
        V3 a = vmin(t[i].v0, vmin(t[i].v1, t[i].v2)), b = vmax(t[i].v0, vmax(t[i].v1, t[i].v2));
        if (b.x - a.x < 0.002f) { b.x = b.x + 0.001f; a.x = a.x - 0.001f; }
        if (b.y - a.y < 0.002f) { b.y = b.y + 0.001f; a.y = a.y - 0.001f; }
        if (b.z - a.z < 0.002f) { b.z = b.z + 0.001f; a.z = a.z - 0.001f; }
        B.quantize(&B.qmin[3 * i], a, false);
        B.quantize(&B.qmax[3 * i], b, true);
        B.tri[i] = i;
    }
    if (n > 0) B.build(0, n);
    return B.tri;
}

inline void world_sort_bvh_order(WorldHost& W) {
    for (int mesh = 0; mesh < W.numMeshes; mesh++) {
        const int lo = mesh == 0 ? 0 : W.meshStart[mesh - 1], hi = W.meshStart[mesh];
        if (hi - lo < 2) continue;
        const std::vector<int> order = bvh_leaf_order(W.tris.data() + lo, hi - lo);
        std::vector<Tri> tmp(W.tris.begin() + lo, W.tris.begin() + hi);
        for (int k = 0; k < hi - lo; k++) W.tris[lo + k] = tmp[order[k]];
    }
}

inline void world_find_free_box(WorldHost& W);
inline void world_build_broadphase_registry(WorldHost& W);

// Uniform grid (cell = 5 bt = 250 uu) over the triangles; a triangle is listed in every cell its AABB (+0.1) touches.
inline void world_build_grid(WorldHost& W) {
    world_build_edge_info(W);
    world_sort_bvh_order(W);
    for (Tri& t : W.tris) {
        const V3 v10 = t.v1 - t.v0, v20 = t.v2 - t.v0;
        t.rayN = cross(v10, v20);
        t.rayD = dot(t.v0, t.rayN);
    }
    W.cellSize = 5.f;
    W.gridMin = v3(-4500.f * kUU2BT, -6200.f * kUU2BT, -100.f * kUU2BT);
    V3 gmax = v3(4500.f * kUU2BT, 6200.f * kUU2BT, 2200.f * kUU2BT);
    W.gx = (int)ceilf((gmax.x - W.gridMin.x) / W.cellSize);
    W.gy = (int)ceilf((gmax.y - W.gridMin.y) / W.cellSize);
    W.gz = (int)ceilf((gmax.z - W.gridMin.z) / W.cellSize);
    const int nc = W.gx * W.gy * W.gz;
    std::vector<std::vector<int>> cells(nc);
    const float inv = 1.f / W.cellSize;
    for (size_t i = 0; i < W.tris.size(); i++) {
        const Tri& t = W.tris[i];
        V3 mn = vmin(t.v0, vmin(t.v1, t.v2)) - v3(0.1f, 0.1f, 0.1f), mx = vmax(t.v0, vmax(t.v1, t.v2)) + v3(0.1f, 0.1f, 0.1f);
        int x0 = grid_coord(mn.x, W.gridMin.x, inv, W.gx), x1 = grid_coord(mx.x, W.gridMin.x, inv, W.gx);
        int y0 = grid_coord(mn.y, W.gridMin.y, inv, W.gy), y1 = grid_coord(mx.y, W.gridMin.y, inv, W.gy);
        int z0 = grid_coord(mn.z, W.gridMin.z, inv, W.gz), z1 = grid_coord(mx.z, W.gridMin.z, inv, W.gz);
        for (int z = z0; z <= z1; z++)
            for (int y = y0; y <= y1; y++)
                for (int x = x0; x <= x1; x++) cells[(z * W.gy + y) * W.gx + x].push_back((int)i);
    }
    W.cellStart.assign(nc + 1, 0);
    W.cellTris.clear();
    for (int c = 0; c < nc; c++) {
        W.cellStart[c] = (int)W.cellTris.size();
        W.cellTris.insert(W.cellTris.end(), cells[c].begin(), cells[c].end());
    }
    W.cellStart[nc] = (int)W.cellTris.size();
    world_find_free_box(W);
    world_build_broadphase_registry(W);
}

// Largest box of empty grid cells grown greedily from the cell above the centre spot (face by face, as long as the new slab of
// cells is empty). Grid cells list every triangle whose AABB (+0.1) touches them, so nothing inside the box is near a triangle.
inline void world_find_free_box(WorldHost& W) {
    auto empty = [&](int x, int y, int z) { const int c = (z * W.gy + y) * W.gx + x; return W.cellStart[c] == W.cellStart[c + 1]; };
    const float inv = 1.f / W.cellSize;
    int lo[3] = {grid_coord(0.f, W.gridMin.x, inv, W.gx), grid_coord(0.f, W.gridMin.y, inv, W.gy), grid_coord(20.f, W.gridMin.z, inv, W.gz)};
    int hi[3] = {lo[0], lo[1], lo[2]};
    W.freeMin = W.freeMax = v3(0, 0, 0);
    if (W.tris.empty() || !empty(lo[0], lo[1], lo[2])) return;
    const int dim[3] = {W.gx, W.gy, W.gz};
    bool grew = true;
    while (grew) {
        grew = false;
        for (int face = 0; face < 6; face++) {
            const int ax = face >> 1, up = face & 1;
            const int layer = up ? hi[ax] + 1 : lo[ax] - 1;
            if (layer < 0 || layer >= dim[ax]) continue;
            bool ok = true;
            int a[3];
            for (a[0] = (ax == 0 ? layer : lo[0]); ok && a[0] <= (ax == 0 ? layer : hi[0]); a[0]++)
                for (a[1] = (ax == 1 ? layer : lo[1]); ok && a[1] <= (ax == 1 ? layer : hi[1]); a[1]++)
                    for (a[2] = (ax == 2 ? layer : lo[2]); ok && a[2] <= (ax == 2 ? layer : hi[2]); a[2]++) ok = empty(a[0], a[1], a[2]);
            if (ok) { if (up) hi[ax] = layer; else lo[ax] = layer; grew = true; }
        }
    }
    // grid_coord clamps: a coordinate outside the grid maps to the border cell, so the box may only be open-ended where the
    // border cells themselves are empty (they are part of the box then); keep it finite and simple: cell bounds as they are
    W.freeMin = v3(W.gridMin.x + lo[0] * W.cellSize, W.gridMin.y + lo[1] * W.cellSize, W.gridMin.z + lo[2] * W.cellSize);
    W.freeMax = v3(W.gridMin.x + (hi[0] + 1) * W.cellSize, W.gridMin.y + (hi[1] + 1) * W.cellSize, W.gridMin.z + (hi[2] + 1) * W.cellSize);
}

// btRSBroadphase's static registry (see WorldDev::bpCellStatics). Grid: ArenaConfig.h:22-29 (minPos, maxPos, maxAABBLen 370 uu),
// cells per axis = ceil(range / cell) (btRSBroadphase.cpp:78-81). A static is entered, for every cell of its box's cell range
// (a mesh: only the cells one of its triangles reaches), into that cell and its neighbours (:97-180; "reaches" = the triangle's
// box overlaps the cell - the reference asks its quantized BVH, which is a little more generous at cell borders). Boxes: a mesh's
// is its vertex box (the arena meshes have no collision margin), a plane's the half space of btStaticPlaneShape.cpp:40-54 (0.2
// past the plane); the proxies are these grown by the contact breaking threshold 0.02 (btCollisionWorld::updateSingleAabb),
// checked against the reference's (oracle/ref_harness.cpp: ref_arena_static_boxes, tools/manifold_order_check.py).
// bpStaticBox keeps the shapes' own boxes.
inline void world_build_broadphase_registry(WorldHost& W) {
    W.bpMin = v3(-4500.f, -6000.f, 0.f) * kUU2BT;
    const V3 mx = v3(4500.f, 6000.f, 2500.f) * kUU2BT;
    W.bpCell = 370.f * kUU2BT;
    auto cells = [&](float range) { int n = (int)ceilf(range / W.bpCell); return n < 1 ? 1 : n; };
    W.bpX = cells(mx.x - W.bpMin.x); W.bpY = cells(mx.y - W.bpMin.y); W.bpZ = cells(mx.z - W.bpMin.z);
    const int nc = W.bpX * W.bpY * W.bpZ, M = W.numMeshes < kMaxWorldMeshes ? W.numMeshes : kMaxWorldMeshes;
    W.numStatics = M + 4;
    std::vector<uint32_t> own(nc, 0u);
    std::vector<V3> lo(W.numStatics, v3(1e30f, 1e30f, 1e30f)), hi(W.numStatics, v3(-1e30f, -1e30f, -1e30f));
    auto coord = [&](float v, float mn, int n) { int c = (int)((v - mn) / W.bpCell); return c < 0 ? 0 : (c >= n ? n - 1 : c); };
    auto enter = [&](V3 a, V3 b, int s) {
        for (int i = coord(a.x, W.bpMin.x, W.bpX); i <= coord(b.x, W.bpMin.x, W.bpX); i++)
            for (int j = coord(a.y, W.bpMin.y, W.bpY); j <= coord(b.y, W.bpMin.y, W.bpY); j++)
                for (int k = coord(a.z, W.bpMin.z, W.bpZ); k <= coord(b.z, W.bpMin.z, W.bpZ); k++) own[(i * W.bpY + j) * W.bpZ + k] |= 1u << s;
    };
    for (const Tri& t : W.tris) {
        const int m = (int)(t.flags >> TRI_MESH_SHIFT);
        if (m >= M) continue;
        const V3 a = vmin(t.v0, vmin(t.v1, t.v2)), b = vmax(t.v0, vmax(t.v1, t.v2));
        lo[m] = vmin(lo[m], a); hi[m] = vmax(hi[m], b);
        enter(a, b, m);
    }
    const float grow = 0.02f, big = 1e18f;
    const float ex = rl::ARENA_EXTENT_X * kUU2BT, hz = rl::ARENA_HEIGHT * kUU2BT, off = 0.2f;
    lo[M + 0] = v3(-big, -big, -big); hi[M + 0] = v3(big, big, 0.f + off);            // floor
    lo[M + 1] = v3(-big, -big, hz - off); hi[M + 1] = v3(big, big, big);             // ceiling
    lo[M + 2] = v3(-big, -big, -big); hi[M + 2] = v3(-ex + off, big, big);           // left wall (normal +x)
    lo[M + 3] = v3(ex - off, -big, -big); hi[M + 3] = v3(big, big, big);             // right wall
    for (int p = 0; p < 4; p++) enter(lo[M + p] - v3(grow, grow, grow), vmin(hi[M + p] + v3(grow, grow, grow), mx), M + p);
    W.bpCellStatics.assign(nc, 0u);
    for (int i = 0; i < W.bpX; i++)
        for (int j = 0; j < W.bpY; j++)
            for (int k = 0; k < W.bpZ; k++) {
                uint32_t bits = 0;
                for (int a = -1; a <= 1; a++)
                    for (int b = -1; b <= 1; b++)
                        for (int c = -1; c <= 1; c++) {
                            const int x = i + a, y = j + b, z = k + c;
                            if (x < 0 || y < 0 || z < 0 || x >= W.bpX || y >= W.bpY || z >= W.bpZ) continue;
                            bits |= own[(x * W.bpY + y) * W.bpZ + z];
                        }
                W.bpCellStatics[(i * W.bpY + j) * W.bpZ + k] = bits;
            }
    W.bpStaticBox.resize(6 * W.numStatics);
    for (int s2 = 0; s2 < W.numStatics; s2++) {
        float* o = &W.bpStaticBox[6 * s2];
        o[0] = lo[s2].x; o[1] = lo[s2].y; o[2] = lo[s2].z; o[3] = hi[s2].x; o[4] = hi[s2].y; o[5] = hi[s2].z;
    }
}

inline WorldDev world_view(const WorldHost& W) {
    WorldDev d;
    d.tris = W.tris.data(); d.numTris = (int)W.tris.size();
    d.cellStart = W.cellStart.data(); d.cellTris = W.cellTris.data();
    d.gx = W.gx; d.gy = W.gy; d.gz = W.gz; d.gridMin = W.gridMin; d.cellSize = W.cellSize; d.invCellSize = 1.f / W.cellSize;
    d.freeMin = W.freeMin; d.freeMax = W.freeMax;
    d.bpCellStatics = W.bpCellStatics.empty() ? nullptr : W.bpCellStatics.data();
    d.bpStaticBox = W.bpStaticBox.empty() ? nullptr : W.bpStaticBox.data();
    d.bpX = W.bpX; d.bpY = W.bpY; d.bpZ = W.bpZ; d.numStatics = W.numStatics; d.bpMin = W.bpMin; d.bpCell = W.bpCell;
    return d;
}

}  // namespace ggl
