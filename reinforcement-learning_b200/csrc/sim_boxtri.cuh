// Car hitbox vs arena mesh: per-triangle GJK/EPA closest points + Bullet's internal-edge normal correction.
//
// Reference path (BT = RocketSim/libsrc/bullet3-3.24/BulletCollision):
//   BT/CollisionDispatch/btCompoundCollisionAlgorithm.cpp:229      compound child (the hitbox box) vs the mesh
//   BT/CollisionDispatch/btConvexConcaveCollisionAlgorithm.cpp:71-187,216-248  triangle callback: AABB test, normal early-out
//   BT/CollisionDispatch/btConvexConvexAlgorithm.cpp:508            box vs btTriangleShape -> one gjkPairDetector.getClosestPoints
//   BT/NarrowPhaseCollision/btGjkPairDetector.cpp:686-959           (restated in box_tri_closest below)
//   BT/CollisionDispatch/btManifoldResult.cpp:109-206               addContactPoint + contact-added callback
//   BT/CollisionDispatch/btInternalEdgeUtility.cpp:414-798          btAdjustInternalEdgeContacts (adjust_internal_edge)
//   RocketSim/src/Sim/Arena/Arena.cpp:218-281,420-427               callback: car-world friction/restitution, worldContact
#pragma once
#include "sim_world.cuh"
#include "sim_epa.cuh"

namespace ggl {

struct ClosestOut { V3 normalOnB, pointOnB; float distance; };

// btGjkPairDetector::getClosestPointsNonVirtual for (box core, marginA) vs (triangle, margin 0).
// `s` is given in world coordinates; the detector re-centres both transforms by positionOffset like the reference.
// `vs` = working storage of the simplex solver, provided by the caller: its vertex arrays are indexed at run time, so as a
// stack object they are local memory; the lane-parallel narrow phase hands in a slice of the arena's shared-memory scratch.
GGL_HDN bool box_tri_closest(const BoxTri& world, float maxDistSq, ClosestOut& out, Voronoi& vs) {
    BoxTri s = world;
    const V3 positionOffset = (world.posA + world.posB) * 0.5f;
    s.posA = world.posA - positionOffset;
    s.posB = world.posB - positionOffset;
    const float marginA = s.marginA, marginB = 0.f;
    float distance = 0.f;
    V3 normalInB = v3(0, 0, 0), pointOnA = v3(0, 0, 0), pointOnB = v3(0, 0, 0);
    V3 axis = v3(0, 1, 0);
    bool isValid = false, checkSimplex = false;
    int degenerate = 0, curIter = 0;
    float squaredDistance = 1e18f;
    const float margin = marginA + marginB;
    vs_reset(vs);
    while (true) {
        V3 sepInA = tmul(world.basisA, -axis);
        V3 sepInB = axis;
        V3 pInA = box_support(s.he, sepInA);
        V3 qInB = tri_support(s, sepInB);
        V3 pWorld = mul(s.basisA, pInA) + s.posA;
        V3 qWorld = qInB + s.posB;
        V3 w = pWorld - qWorld;
        float delta = dot(axis, w);
        if (delta > 0.f && delta * delta > squaredDistance * maxDistSq) { degenerate = 10; checkSimplex = true; break; }
        if (vs_in_simplex(vs, w)) { degenerate = 1; checkSimplex = true; break; }
        float f0 = squaredDistance - delta, f1 = squaredDistance * 1.0e-6f;
        if (f0 <= f1) { degenerate = (f0 <= 0.f) ? 2 : 11; checkSimplex = true; break; }
        vs_add(vs, w, pWorld, qWorld);
        bool ok = vs_update(vs);
        V3 newAxis = vs.cachedV;
        if (!ok) { degenerate = 3; checkSimplex = true; break; }
        if (len2(newAxis) < 1.0e-6f) { axis = newAxis; degenerate = 6; checkSimplex = true; break; }
        float prev = squaredDistance;
        squaredDistance = len2(newAxis);
        if (prev - squaredDistance <= kEps * prev) { checkSimplex = true; degenerate = 12; break; }
        axis = newAxis;
        if (curIter++ > 1000) break;
        if (vs.n == 4) { degenerate = 13; break; }
    }
    if (checkSimplex) {
        vs_update(vs);
        pointOnA = vs.cachedP1; pointOnB = vs.cachedP2;
        normalInB = axis;
        float lenSqr = len2(axis);
        if (lenSqr < 0.0001f) degenerate = 5;
        if (lenSqr > kEps * kEps) {
            float rlen = 1.f / sqrtf(lenSqr);
            normalInB *= rlen;
            float sd = sqrtf(squaredDistance);
            pointOnA -= axis * (marginA / sd);
            pointOnB += axis * (marginB / sd);
            distance = (1.f / rlen) - margin;
            isValid = true;
        }
    }
    bool catchDegenerate = degenerate && ((distance + margin) < 0.01f);
    if (!isValid || catchDegenerate) {
        // btGjkEpaPenetrationDepthSolver::calcPenDepth over its guess directions
        // btVector3::safeNormalize (LinearMath/btVector3.h:287-300)
        V3 ab = s.posB - s.posA;
        float l2ab = len2(ab);
        V3 g0 = (l2ab >= kEps * kEps) ? ab / sqrtf(l2ab) : v3(1, 0, 0);
        V3 ba = s.posA - s.posB;
        float l2ba = len2(ba);
        V3 g1 = (l2ba >= kEps * kEps) ? ba / sqrtf(l2ba) : v3(1, 0, 0);
        const V3 guesses[9] = {g0, g1, v3(0, 0, 1), v3(0, 1, 0), v3(1, 0, 0), v3(1, 1, 0), v3(1, 1, 1), v3(0, 1, 1), v3(1, 0, 1)};
        bool isValid2 = false, haveDist = false;
        PenResult pr; pr.witA = pr.witB = pr.normal = v3(0, 0, 0); pr.distance = 0.f;
        V3 sepAxis = v3(0, 0, 0);
        for (int i = 0; i < 9; i++) {
            if (epa_penetration(s, guesses[i], pr)) { isValid2 = true; sepAxis = pr.normal; break; }
            if (epa_distance(s, guesses[i], pr)) { haveDist = true; sepAxis = pr.normal; break; }
        }
        if (!isValid2 && !haveDist) { pr.witA = pr.witB = v3(0, 0, 0); sepAxis = v3(0, 0, 0); }
        if (isValid2) {
            V3 tmpN = pr.witB - pr.witA;
            float lenSqr = len2(tmpN);
            if (lenSqr <= kEps * kEps) { tmpN = sepAxis; lenSqr = len2(sepAxis); }
            if (lenSqr > kEps * kEps) {
                tmpN = tmpN / sqrtf(lenSqr);
                float d2 = -len(pr.witA - pr.witB);
                if (!isValid || d2 < distance) {
                    distance = d2; pointOnA = pr.witA; pointOnB = pr.witB; normalInB = tmpN; isValid = true;
                }
            }
        } else if (len2(sepAxis) > 0.f) {
            float d2 = len(pr.witA - pr.witB) - margin;
            if (!isValid || d2 < distance) {
                distance = d2;
                pointOnA = pr.witA - sepAxis * marginA;
                pointOnB = pr.witB + sepAxis * marginB;
                normalInB = normalized(sepAxis);
                isValid = true;
            }
        }
    }
    if (isValid && (distance < 0.f || distance * distance < maxDistSq)) {
        // m_fixContactNormalDirection: normal must point from B's AABB centre towards A's
        V3 extA = v3(dot(vabs(s.basisA.r0), s.he + v3(marginA, marginA, marginA)), dot(vabs(s.basisA.r1), s.he + v3(marginA, marginA, marginA)),
                     dot(vabs(s.basisA.r2), s.he + v3(marginA, marginA, marginA)));
        V3 aMin = s.posA - extA, aMax = s.posA + extA;
        V3 posAc = (aMax + aMin) * 0.5f;
        V3 bMin = vmin(s.t0, vmin(s.t1, s.t2)) + s.posB, bMax = vmax(s.t0, vmax(s.t1, s.t2)) + s.posB;
        V3 posBc = (bMin + bMax) * 0.5f;
        if (dot(posAc - posBc, normalInB) < 0.f) normalInB = normalInB * -1.f;
        out.normalOnB = normalInB;
        out.pointOnB = pointOnB + positionOffset;
        out.distance = distance;
        return true;
    }
    return false;
}

// ---------------------------------------------------------------------------------------------- car vs mesh
// Narrow phase of one (car, triangle) pair: pure function of the car's transform and the triangle. False = no contact point.
GGL_HDN bool car_tri_narrow(const CarDev& car, const Tri& t, V3 boxPos, const BoxGeom& bg, float cbt, ClosestOut& o, Voronoi& vs) {
    // early out: hitbox fully on one side of the triangle plane, further away than the breaking threshold
    V3 tn = normalized(cross(t.v1 - t.v0, t.v2 - t.v0));
    const M3 invB = inverse(car.basis);
    for (int side = 0; side < 2; side++) {
        V3 lp = box_support(bg.full, mul(invB, tn));
        V3 wp = mul(car.basis, lp) + boxPos;
        float dist = dot(tn, t.v0) - dot(tn, wp);
        if (dist > cbt) return false;
        tn = tn * -1.f;
    }
    BoxTri s;
    s.basisA = car.basis; s.posA = boxPos; s.he = bg.core; s.marginA = bg.margin;
    s.t0 = t.v0; s.t1 = t.v1; s.t2 = t.v2; s.posB = v3(0, 0, 0);
    float md = bg.margin + 0.f + cbt;
    if (!box_tri_closest(s, md * md, o, vs)) return false;
    if (o.distance > cbt) return false;
    return true;
}

// Adds the contact point of a narrow-phase hit to the (car, mesh) manifold `m` (order-dependent: replacement policy)
GGL_HD void car_tri_emit(ArenaDev& A, ContactSink& S, ManifoldCtx& m, int ci, const Tri& t, const ClosestOut& o) {
    CarDev& car = A.cars[ci];
    V3 pA = o.pointOnB + o.normalOnB * o.distance;
    Contact c;
    c.normal = o.normalOnB; c.dist = o.distance;
    c.relA = pA - car.pos; c.relB = o.pointOnB;
    c.friction = rl::CARWORLD_FRICTION; c.restitution = rl::CARWORLD_RESTITUTION;
    c.bodyA = (int16_t)(1 + ci); c.bodyB = -1; c.special = 0; c.staticId = (int16_t)(1 + (int)(t.flags >> TRI_MESH_SHIFT));
    int row = manifold_add(S, m, c, tmul(car.basis, pA - car.pos));
    if (row < 0) return;
    car.worldContactHas = 1;
    car.worldContactNormal = o.normalOnB;
    V3 n = o.normalOnB, pB = o.pointOnB;
    adjust_internal_edge(t, n, pB, pA, o.distance);
    S.at(row).normal = n;
    S.at(row).relB = pB;
}

GGL_HD V3 car_box_pos(const CarDev& car) { return car.pos + mul(car.basis, car_hitbox_offset()); }

// Stage 1 (the car's lane): candidate triangles of car `ci`, de-duplicated and in (mesh, triangle) index order -- one
// manifold per mesh body, visited in the reference's order.
GGL_HDN void car_mesh_gather_cells(ArenaDev& A, const WorldDev& W, int ci, V3 mn, V3 mx);
GGL_HD void car_mesh_gather(ArenaDev& A, const WorldDev& W, int ci) {
    A.meshCandN[ci] = 0; A.meshHitN[ci] = 0; A.meshHitOverflow[ci] = 0;
    if (W.numTris <= 0 || A.cars[ci].rbDisabled) return;
    const CarDev& car = A.cars[ci];
    const BoxGeom bg = car_box();
    V3 boxPos = car_box_pos(car);
    V3 ext = v3(dot(vabs(car.basis.r0), bg.full), dot(vabs(car.basis.r1), bg.full), dot(vabs(car.basis.r2), bg.full));
    V3 mn = boxPos - ext, mx = boxPos + ext;
    if (world_aabb_is_free(W, mn, mx)) return;          // open play: every grid cell the box touches is empty
    car_mesh_gather_cells(A, W, ci, mn, mx);
}
GGL_HDN void car_mesh_gather_cells(ArenaDev& A, const WorldDev& W, int ci, V3 mn, V3 mx) {
    int x0 = grid_coord(mn.x, W.gridMin.x, W.invCellSize, W.gx), x1 = grid_coord(mx.x, W.gridMin.x, W.invCellSize, W.gx);
    int y0 = grid_coord(mn.y, W.gridMin.y, W.invCellSize, W.gy), y1 = grid_coord(mx.y, W.gridMin.y, W.invCellSize, W.gy);
    int z0 = grid_coord(mn.z, W.gridMin.z, W.invCellSize, W.gz), z1 = grid_coord(mx.z, W.gridMin.z, W.invCellSize, W.gz);
    uint16_t* cand = A.meshCand[ci];
    int nc = 0;
    for (int z = z0; z <= z1; z++)
        for (int y = y0; y <= y1; y++)
            for (int x = x0; x <= x1; x++) {
                int cell = (z * W.gy + y) * W.gx + x;
                for (int k = W.cellStart[cell]; k < W.cellStart[cell + 1]; k++) {
                    int ti = W.cellTris[k];
                    if (!tri_overlaps_aabb(W.tris[ti], mn, mx)) continue;
                    int pos = 0; bool dup = false;
                    while (pos < nc && (int)cand[pos] <= ti) { dup |= (int)cand[pos] == ti; pos++; }
                    if (dup || nc >= kMaxMeshCand) continue;
                    for (int q = nc; q > pos; q--) cand[q] = cand[q - 1];
                    cand[pos] = (uint16_t)ti; nc++;
                }
            }
    A.meshCandN[ci] = (uint8_t)nc;
}

// Stage 2 (any lane): narrow phase of candidate `k` of car `ci`; a hit is appended to the car's hit list (any order).
// The simplex solver's storage: lane l < 8 of the arena's group uses bytes [256 l, 256 l + 256) of the contact rows (+ the head of
// the row-order list behind them), which nobody needs before contact generation starts; further lanes (16-lane builds) fall
// back to their stack.
static_assert(8 * sizeof(Voronoi) <= sizeof(Contact) * kMaxContacts + kMaxContacts, "simplex scratch must fit contacts[] + contactOrder[]");
GGL_HD void car_mesh_narrow(ArenaDev& A, const WorldDev& W, int ci, int k, int lane) {
    const CarDev& car = A.cars[ci];
    ClosestOut o;
    Voronoi onStack;
    Voronoi& vs = lane < 8 ? reinterpret_cast<Voronoi*>(A.contacts)[lane] : onStack;
    if (!car_tri_narrow(car, W.tris[A.meshCand[ci][k]], car_box_pos(car), car_box(), car_cbt(), o, vs)) return;
#if defined(__CUDA_ARCH__)
    const uint32_t slot = atomicAdd(&A.meshHitN[ci], 1u);
#else
    const uint32_t slot = A.meshHitN[ci]++;
#endif
    if (slot >= (uint32_t)kMaxMeshHit) { A.meshHitOverflow[ci] = 1; return; }
    CarTriHit& h = A.meshHit[ci][slot];
    h.normalOnB = o.normalOnB; h.pointOnB = o.pointOnB; h.distance = o.distance; h.cand = (uint32_t)k;
}

// Stage 3 (the car's lane): contact points into the car's manifolds in candidate order.
GGL_HDN void car_mesh_emit(ArenaDev& A, ContactSink& S, const WorldDev& W, int ci) {
    const int nc = A.meshCandN[ci];
    if (nc == 0) return;
    int curMesh = -1;
    ManifoldCtx m = manifold_begin(S);
    if (A.meshHitOverflow[ci]) {
        // more hits than the staging holds: redo this car serially (same results, candidate order)
        const CarDev& car = A.cars[ci];
        for (int k = 0; k < nc; k++) {
            const Tri& t = W.tris[A.meshCand[ci][k]];
            int mesh = (int)(t.flags >> TRI_MESH_SHIFT);
            if (mesh != curMesh) { m = manifold_begin(S); curMesh = mesh; }
            ClosestOut o;
            Voronoi vs;        // serial redo while contact rows are being written: on the stack
            if (car_tri_narrow(car, t, car_box_pos(car), car_box(), car_cbt(), o, vs)) car_tri_emit(A, S, m, ci, t, o);
        }
        return;
    }
    const int nh = (int)A.meshHitN[ci];
    uint32_t done = 0;
    for (int n = 0; n < nh; n++) {
        int best = -1;                                   // next hit in candidate order
        for (int q = 0; q < nh; q++)
            if (!(done & (1u << q)) && (best < 0 || A.meshHit[ci][q].cand < A.meshHit[ci][best].cand)) best = q;
        done |= 1u << best;
        const CarTriHit& h = A.meshHit[ci][best];
        const Tri& t = W.tris[A.meshCand[ci][h.cand]];
        int mesh = (int)(t.flags >> TRI_MESH_SHIFT);
        if (mesh != curMesh) { m = manifold_begin(S); curMesh = mesh; }
        ClosestOut o; o.normalOnB = h.normalOnB; o.pointOnB = h.pointOnB; o.distance = h.distance;
        car_tri_emit(A, S, m, ci, t, o);
    }
}

}  // namespace ggl
