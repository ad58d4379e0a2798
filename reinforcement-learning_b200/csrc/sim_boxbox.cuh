// Car hitbox vs car hitbox: separating-axis box-box contact generation (up to 4 points) and the bump / demolition
// game logic that rides on the contact callback.
//
// Reference path (BT = RocketSim/libsrc/bullet3-3.24/BulletCollision/CollisionDispatch):
//   BT/btCompoundCompoundCollisionAlgorithm.cpp:284    child box vs child box, shared manifold of the car pair
//   BT/btBoxBoxCollisionAlgorithm.cpp:55-83            detector on the half extents WITH margin, unlimited distance
//   BT/btBoxBoxDetector.cpp:77-262                     helpers (line closest approach, rect/quad clipping, point culling)
//   BT/btBoxBoxDetector.cpp:267-730                    dBoxBox2: 15-axis SAT, edge-edge point or face clipping, <= 4 points
//   RocketSim/src/Sim/Arena/Arena.cpp:336-418          _BtCallback_OnCarCarCollision (friction 0.09 / restitution 0.1, bump, demo)
//   RocketSim/src/Sim/Car/Car.cpp:38-41                Car::Demolish
#pragma once
#include "sim_world.cuh"
#ifdef GGL_DEBUG_BOXBOX
#include <stdio.h>
#define GGL_DBG(...) printf(__VA_ARGS__)
#else
#define GGL_DBG(...)
#endif

namespace ggl {

GGL_HD V3 selv(V3 a, V3 b, V3 c, int i) { return i == 0 ? a : (i == 1 ? b : c); }

GGL_HD void line_closest_approach(V3 pa, V3 ua, V3 pb, V3 ub, float& alpha, float& beta) {
    V3 p = pb - pa;
    float uaub = dot(ua, ub), q1 = dot(ua, p), q2 = -dot(ub, p);
    float d = 1 - uaub * uaub;
    if (d <= 0.0001f) { alpha = 0.f; beta = 0.f; }
    else { d = 1.f / d; alpha = (q1 + uaub * q2) * d; beta = (uaub * q1 + q2) * d; }
}

// intersection polygon of the rectangle [-h0,h0]x[-h1,h1] with the quad p (4 x,y pairs); returns the vertex count (<= 8)
GGL_HD int intersect_rect_quad(const float h[2], const float p[8], float ret[16]) {
    int nq = 4, nr = 0;
    float bufA[16], bufB[16];
    for (int i = 0; i < 8; i++) bufA[i] = p[i];
    float* q = bufA;
    float* r = bufB;
    bool done = false;
    for (int dir = 0; dir <= 1 && !done; dir++) {
        for (int sign = -1; sign <= 1 && !done; sign += 2) {
            float* pq = q;
            float* pr = r;
            nr = 0;
            for (int i = nq; i > 0; i--) {
                if (sign * pq[dir] < h[dir]) {
                    pr[0] = pq[0]; pr[1] = pq[1]; pr += 2; nr++;
                    if (nr & 8) { q = r; done = true; break; }
                }
                float* nextq = (i > 1) ? pq + 2 : q;
                if ((sign * pq[dir] < h[dir]) ^ (sign * nextq[dir] < h[dir])) {
                    pr[1 - dir] = pq[1 - dir] + (nextq[1 - dir] - pq[1 - dir]) / (nextq[dir] - pq[dir]) * (sign * h[dir] - pq[dir]);
                    pr[dir] = sign * h[dir];
                    pr += 2; nr++;
                    if (nr & 8) { q = r; done = true; break; }
                }
                pq += 2;
            }
            if (done) break;
            q = r;
            r = (q == bufB) ? bufA : bufB;
            nq = nr;
        }
    }
    for (int i = 0; i < nr * 2; i++) ret[i] = q[i];
    return nr;
}

// picks m of the n points p (x,y pairs), spread by angle around the centroid, always keeping point i0
GGL_HD void cull_points(int n, const float p[], int m, int i0, int iret[]) {
    const float kPiF = 3.14159265f;
    float a, cx, cy, q;
    if (n == 1) { cx = p[0]; cy = p[1]; }
    else if (n == 2) { cx = 0.5f * (p[0] + p[2]); cy = 0.5f * (p[1] + p[3]); }
    else {
        a = 0; cx = 0; cy = 0;
        for (int i = 0; i < n - 1; i++) {
            q = p[i * 2] * p[i * 2 + 3] - p[i * 2 + 2] * p[i * 2 + 1];
            a += q;
            cx += q * (p[i * 2] + p[i * 2 + 2]);
            cy += q * (p[i * 2 + 1] + p[i * 2 + 3]);
        }
        q = p[n * 2 - 2] * p[1] - p[0] * p[n * 2 - 1];
        if (fabsf(a + q) > kEps) a = 1.f / (3.0f * (a + q)); else a = 1e18f;
        cx = a * (cx + q * (p[n * 2 - 2] + p[0]));
        cy = a * (cy + q * (p[n * 2 - 1] + p[1]));
    }
    float ang[8];
    int avail[8];
    for (int i = 0; i < n; i++) { ang[i] = atan2f(p[i * 2 + 1] - cy, p[i * 2] - cx); avail[i] = 1; }
    avail[i0] = 0;
    iret[0] = i0;
    int w = 1;
    for (int j = 1; j < m; j++) {
        a = (float)j * (2 * kPiF / m) + ang[i0];
        if (a > kPiF) a -= 2 * kPiF;
        float maxdiff = 1e9f, diff;
        iret[w] = i0;
        for (int i = 0; i < n; i++)
            if (avail[i]) {
                diff = fabsf(ang[i] - a);
                if (diff > kPiF) diff = 2 * kPiF - diff;
                if (diff < maxdiff) { maxdiff = diff; iret[w] = i; }
            }
        avail[iret[w]] = 0;
        w++;
    }
}

struct BoxBoxOut { int n; V3 normalOnB; V3 point[4]; float depth[4]; };   // depth = signed distance (negative = penetrating)

// dBoxBox2 + btBoxBoxDetector::getClosestPoints output convention (normal on B = -normal, distance = -depth)
GGL_HDN void box_box(V3 p1, const M3& R1, V3 A, V3 p2, const M3& R2, V3 B, BoxBoxOut& out) {
    out.n = 0;
    const float fudge = 1.05f;
    const V3 a1c[3] = {col(R1, 0), col(R1, 1), col(R1, 2)}, b1c[3] = {col(R2, 0), col(R2, 1), col(R2, 2)};
    const float Ah[3] = {A.x, A.y, A.z}, Bh[3] = {B.x, B.y, B.z};
    V3 p = p2 - p1;
    float pp[3] = {dot(a1c[0], p), dot(a1c[1], p), dot(a1c[2], p)};
    float R[3][3], Q[3][3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) { R[i][j] = dot(a1c[i], b1c[j]); Q[i][j] = fabsf(R[i][j]); }
    float s = -1e30f, s2, l;
    int invert = 0, code = 0;
    V3 normalC = v3(0, 0, 0);
    int normalR = -1;      // 0..2 column of R1, 3..5 column of R2
#define GGL_TST(expr1, expr2, nr, cc) { float e1 = (expr1); s2 = fabsf(e1) - (expr2); if (s2 > 0.f) return; \
        if (s2 > s) { s = s2; normalR = (nr); invert = (e1 < 0.f); code = (cc); } }
    GGL_TST(pp[0], (Ah[0] + Bh[0] * Q[0][0] + Bh[1] * Q[0][1] + Bh[2] * Q[0][2]), 0, 1);
    GGL_TST(pp[1], (Ah[1] + Bh[0] * Q[1][0] + Bh[1] * Q[1][1] + Bh[2] * Q[1][2]), 1, 2);
    GGL_TST(pp[2], (Ah[2] + Bh[0] * Q[2][0] + Bh[1] * Q[2][1] + Bh[2] * Q[2][2]), 2, 3);
    GGL_TST(dot(b1c[0], p), (Ah[0] * Q[0][0] + Ah[1] * Q[1][0] + Ah[2] * Q[2][0] + Bh[0]), 3, 4);
    GGL_TST(dot(b1c[1], p), (Ah[0] * Q[0][1] + Ah[1] * Q[1][1] + Ah[2] * Q[2][1] + Bh[1]), 4, 5);
    GGL_TST(dot(b1c[2], p), (Ah[0] * Q[0][2] + Ah[1] * Q[1][2] + Ah[2] * Q[2][2] + Bh[2]), 5, 6);
#undef GGL_TST
#define GGL_TST(expr1, expr2, n1, n2, n3, cc) { float e1 = (expr1); s2 = fabsf(e1) - (expr2); if (s2 > kEps) return; \
        l = sqrtf((n1) * (n1) + (n2) * (n2) + (n3) * (n3)); \
        if (l > kEps) { s2 /= l; if (s2 * fudge > s) { s = s2; normalR = -1; normalC = v3((n1) / l, (n2) / l, (n3) / l); invert = (e1 < 0.f); code = (cc); } } }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) Q[i][j] += 1.0e-5f;
    GGL_TST(pp[2] * R[1][0] - pp[1] * R[2][0], (Ah[1] * Q[2][0] + Ah[2] * Q[1][0] + Bh[1] * Q[0][2] + Bh[2] * Q[0][1]), 0.f, -R[2][0], R[1][0], 7);
    GGL_TST(pp[2] * R[1][1] - pp[1] * R[2][1], (Ah[1] * Q[2][1] + Ah[2] * Q[1][1] + Bh[0] * Q[0][2] + Bh[2] * Q[0][0]), 0.f, -R[2][1], R[1][1], 8);
    GGL_TST(pp[2] * R[1][2] - pp[1] * R[2][2], (Ah[1] * Q[2][2] + Ah[2] * Q[1][2] + Bh[0] * Q[0][1] + Bh[1] * Q[0][0]), 0.f, -R[2][2], R[1][2], 9);
    GGL_TST(pp[0] * R[2][0] - pp[2] * R[0][0], (Ah[0] * Q[2][0] + Ah[2] * Q[0][0] + Bh[1] * Q[1][2] + Bh[2] * Q[1][1]), R[2][0], 0.f, -R[0][0], 10);
    GGL_TST(pp[0] * R[2][1] - pp[2] * R[0][1], (Ah[0] * Q[2][1] + Ah[2] * Q[0][1] + Bh[0] * Q[1][2] + Bh[2] * Q[1][0]), R[2][1], 0.f, -R[0][1], 11);
    GGL_TST(pp[0] * R[2][2] - pp[2] * R[0][2], (Ah[0] * Q[2][2] + Ah[2] * Q[0][2] + Bh[0] * Q[1][1] + Bh[1] * Q[1][0]), R[2][2], 0.f, -R[0][2], 12);
    GGL_TST(pp[1] * R[0][0] - pp[0] * R[1][0], (Ah[0] * Q[1][0] + Ah[1] * Q[0][0] + Bh[1] * Q[2][2] + Bh[2] * Q[2][1]), -R[1][0], R[0][0], 0.f, 13);
    GGL_TST(pp[1] * R[0][1] - pp[0] * R[1][1], (Ah[0] * Q[1][1] + Ah[1] * Q[0][1] + Bh[0] * Q[2][2] + Bh[2] * Q[2][0]), -R[1][1], R[0][1], 0.f, 14);
    GGL_TST(pp[1] * R[0][2] - pp[0] * R[1][2], (Ah[0] * Q[1][2] + Ah[1] * Q[0][2] + Bh[0] * Q[2][1] + Bh[1] * Q[2][0]), -R[1][2], R[0][2], 0.f, 15);
#undef GGL_TST
    if (!code) return;
    V3 normal;
    if (normalR >= 0) normal = normalR < 3 ? selv(a1c[0], a1c[1], a1c[2], normalR) : selv(b1c[0], b1c[1], b1c[2], normalR - 3);
    else normal = mul(R1, normalC);
    if (invert) normal = -normal;
    const float depth = -s;
    out.normalOnB = -normal;
    if (code > 6) {
        V3 pa = p1;
        for (int j = 0; j < 3; j++) {
            float sign = (dot(normal, a1c[j]) > 0.f) ? 1.f : -1.f;
            pa += a1c[j] * (sign * Ah[j]);
        }
        V3 pb = p2;
        for (int j = 0; j < 3; j++) {
            float sign = (dot(normal, b1c[j]) > 0.f) ? -1.f : 1.f;
            pb += b1c[j] * (sign * Bh[j]);
        }
        V3 ua = selv(a1c[0], a1c[1], a1c[2], (code - 7) / 3), ub = selv(b1c[0], b1c[1], b1c[2], (code - 7) % 3);
        float alpha, beta;
        line_closest_approach(pa, ua, pb, ub, alpha, beta);
        pb += ub * beta;
        out.n = 1; out.point[0] = pb; out.depth[0] = -depth;
        return;
    }
    // face-something contact: reference face on box a (code 1..3) or box b (4..6), incident face on the other box.
    // Axes / extents are picked with selects (selv/comp), never by indexing small local arrays with a run-time index:
    // nvcc 12.9 -O3 kept the constant-indexed elements in registers and read a stale stack copy for the dynamic index.
    const bool refA = code <= 3;
    const V3 Ra0 = refA ? a1c[0] : b1c[0], Ra1 = refA ? a1c[1] : b1c[1], Ra2 = refA ? a1c[2] : b1c[2];
    const V3 Rb0 = refA ? b1c[0] : a1c[0], Rb1 = refA ? b1c[1] : a1c[1], Rb2 = refA ? b1c[2] : a1c[2];
    const V3 SaV = refA ? A : B, SbV = refA ? B : A;
    const V3 pa = refA ? p1 : p2, pb = refA ? p2 : p1;
    V3 normal2 = refA ? normal : -normal;
    const V3 nrV = v3(dot(Rb0, normal2), dot(Rb1, normal2), dot(Rb2, normal2));
    const V3 anr = vabs(nrV);
    int lanr, a1, a2;
    if (anr.y > anr.x) {
        if (anr.y > anr.z) { a1 = 0; lanr = 1; a2 = 2; } else { a1 = 0; a2 = 1; lanr = 2; }
    } else {
        if (anr.x > anr.z) { lanr = 0; a1 = 1; a2 = 2; } else { a1 = 0; a2 = 1; lanr = 2; }
    }
    const V3 RbL = selv(Rb0, Rb1, Rb2, lanr), RbA1 = selv(Rb0, Rb1, Rb2, a1), RbA2 = selv(Rb0, Rb1, Rb2, a2);
    const float SbL = comp(SbV, lanr), SbA1 = comp(SbV, a1), SbA2 = comp(SbV, a2);
    V3 center;
    if (comp(nrV, lanr) < 0.f) center = pb - pa + RbL * SbL;
    else center = pb - pa - RbL * SbL;
    int codeN = refA ? code - 1 : code - 4, code1, code2;
    if (codeN == 0) { code1 = 1; code2 = 2; } else if (codeN == 1) { code1 = 0; code2 = 2; } else { code1 = 0; code2 = 1; }
    const V3 RaC1 = selv(Ra0, Ra1, Ra2, code1), RaC2 = selv(Ra0, Ra1, Ra2, code2);
    const float SaN = comp(SaV, codeN);
    float quad[8];
    float c1 = dot(center, RaC1), c2 = dot(center, RaC2);
    float m11 = dot(RaC1, RbA1), m12 = dot(RaC1, RbA2), m21 = dot(RaC2, RbA1), m22 = dot(RaC2, RbA2);
    {
        float k1 = m11 * SbA1, k2 = m21 * SbA1, k3 = m12 * SbA2, k4 = m22 * SbA2;
        quad[0] = c1 - k1 - k3; quad[1] = c2 - k2 - k4;
        quad[2] = c1 - k1 + k3; quad[3] = c2 - k2 + k4;
        quad[4] = c1 + k1 + k3; quad[5] = c2 + k2 + k4;
        quad[6] = c1 + k1 - k3; quad[7] = c2 + k2 - k4;
    }
    float rect[2] = {comp(SaV, code1), comp(SaV, code2)};
    float ret[16];
    int n = intersect_rect_quad(rect, quad, ret);
    if (n < 1) return;
    V3 point[8];
    float dep[8];
    float det1 = 1.f / (m11 * m22 - m12 * m21);
    m11 *= det1; m12 *= det1; m21 *= det1; m22 *= det1;
    int cnum = 0;
    for (int j = 0; j < n; j++) {
        float k1 = m22 * (ret[j * 2] - c1) - m12 * (ret[j * 2 + 1] - c2);
        float k2 = -m21 * (ret[j * 2] - c1) + m11 * (ret[j * 2 + 1] - c2);
        point[cnum] = center + RbA1 * k1 + RbA2 * k2;
        dep[cnum] = SaN - dot(normal2, point[cnum]);
        if (dep[cnum] >= 0.f) { ret[cnum * 2] = ret[j * 2]; ret[cnum * 2 + 1] = ret[j * 2 + 1]; cnum++; }
    }
    if (cnum < 1) return;
    int maxc = 4;
    if (maxc > cnum) maxc = cnum;
    if (maxc < 1) maxc = 1;
    if (cnum <= maxc) {
        for (int j = 0; j < cnum; j++) {
            out.point[j] = (code < 4) ? point[j] + pa : point[j] + pa - normal * dep[j];
            out.depth[j] = -dep[j];
        }
        out.n = cnum;
    } else {
        int i1 = 0;
        float maxdepth = dep[0];
        for (int i = 1; i < cnum; i++)
            if (dep[i] > maxdepth) { maxdepth = dep[i]; i1 = i; }
        int iret[8];
        cull_points(cnum, ret, maxc, i1, iret);
        for (int j = 0; j < maxc; j++) {
            V3 pw = point[iret[j]] + pa;
            out.point[j] = (code < 4) ? pw : pw - normal * dep[iret[j]];
            out.depth[j] = -dep[iret[j]];
        }
        out.n = maxc;
    }
}

// Arena::_BtCallback_OnCarCarCollision for one manifold point (localA / localB = contact in each car's frame, bt units)
GGL_HD void car_car_callback(ArenaDev& A, int i1, int i2, V3 localA, V3 localB) {
    for (int pass = 0; pass < 2; pass++) {
        const bool swapped = pass == 1;
        CarDev& c1 = A.cars[swapped ? i2 : i1];
        CarDev& c2 = A.cars[swapped ? i1 : i2];
        if (c1.isDemoed || c2.isDemoed) return;
        if (c1.carContactOtherId == c2.id && c1.carContactCooldown > 0.f) continue;
        V3 pos1 = c1.pos * kBT2UU, pos2 = c2.pos * kBT2UU, vel1 = c1.vel * kBT2UU, vel2 = c2.vel * kBT2UU;
        V3 deltaPos = pos2 - pos1;
        if (dot(vel1, deltaPos) > 0.f) {
            V3 velDir = rs_normalized(vel1), dirTo = rs_normalized(deltaPos);
            float speedTowards = dot(vel1, dirTo), otherAway = dot(vel2, velDir);
            if (speedTowards > otherAway) {
                V3 lp = swapped ? localB : localA;
                if (lp.x * kBT2UU > rl::BUMP_MIN_FORWARD_DIST) {
                    bool isDemo = c1.isSupersonic && (c1.team != c2.team);
                    if (isDemo) {
                        if (!c2.isDemoed) car_freeze_reported_basis(c2);
                        c2.isDemoed = 1; c2.demoRespawnTimer = rl::DEMO_RESPAWN_TIME;
                    } else {
                        bool groundHit = c2.isOnGround;
                        float base = groundHit ? curve_bump_ground(speedTowards) : curve_bump_air(speedTowards);
                        V3 hitUp = c2.isOnGround ? car_up(c2) : v3(0, 0, 1);
                        V3 imp = velDir * base + hitUp * curve_bump_up(speedTowards) * 1.f;
                        c2.velImpulseCache += imp * kUU2BT;
                    }
                    c1.carContactOtherId = c2.id;
                    c1.carContactCooldown = rl::BUMP_COOLDOWN_TIME;
                    // _carBumpCallback -> RLGymCPP EnvSet.cpp:31-42 _BumpCallback: PlayerEventState of both players
                    if (c1.team != c2.team)
                        for (int sl = 0; sl < A.numCars; sl++) {
                            const uint32_t id = A.cars[A.slotCar[sl]].id;
                            if (id == c1.id) { A.players[sl].events[EV_BUMP] = 1; if (isDemo) A.players[sl].events[EV_DEMO] = 1; }
                            if (id == c2.id) { A.players[sl].events[EV_BUMPED] = 1; if (isDemo) A.players[sl].events[EV_DEMOED] = 1; }
                        }
                }
            }
        }
    }
}

// Pair (lo, hi) with lo < hi in AddCar order. The reference's manifold has body0 = the car added LATER (observed on the
// compiled reference for 1v1/2v2/3v3: pairs come out as (2,1), (3,2), (4,2), (4,3) ...), so A = cars[hi], B = cars[lo].
GGL_HDN void collide_car_car(ArenaDev& A, ContactSink& S, int lo, int hi) {
    const int i = hi, j = lo;
    CarDev& a = A.cars[i];
    CarDev& b = A.cars[j];
    if (a.rbDisabled || b.rbDisabled) return;      // CF_NO_CONTACT_RESPONSE: the manifold is never solved (needsResponse)
    const BoxGeom bg = car_box();
    V3 pa = a.pos + mul(a.basis, car_hitbox_offset()), pb = b.pos + mul(b.basis, car_hitbox_offset());
    // cull (exact): boxes inside disjoint bounding spheres are disjoint, so the SAT below would find a separating axis
    // and report no point (btBoxBoxDetector has no margin)
    {
        const float r2 = 2.0002f * len(bg.full);
        if (len2(pa - pb) > r2 * r2) return;
    }
    // child AABB overlap test of the compound-compound algorithm
    V3 ea = v3(dot(vabs(a.basis.r0), bg.full), dot(vabs(a.basis.r1), bg.full), dot(vabs(a.basis.r2), bg.full));
    V3 eb = v3(dot(vabs(b.basis.r0), bg.full), dot(vabs(b.basis.r1), bg.full), dot(vabs(b.basis.r2), bg.full));
    V3 amin = pa - ea, amax = pa + ea, bmin = pb - eb, bmax = pb + eb;
    if (amin.x > bmax.x || amax.x < bmin.x || amin.y > bmax.y || amax.y < bmin.y || amin.z > bmax.z || amax.z < bmin.z) return;
    BoxBoxOut o;
    box_box(pa, a.basis, bg.full, pb, b.basis, bg.full, o);
    if (o.n == 0) return;
    const float cbt = car_cbt();
    ManifoldCtx m = manifold_begin(S);
    for (int k = 0; k < o.n; k++) {
        if (o.depth[k] > cbt) continue;
        V3 pB = o.point[k];
        V3 pA = pB + o.normalOnB * o.depth[k];
        Contact c;
        c.normal = o.normalOnB; c.dist = o.depth[k];
        c.relA = pA - a.pos; c.relB = pB - b.pos;
        c.friction = rl::CARCAR_FRICTION; c.restitution = rl::CARCAR_RESTITUTION;
        c.bodyA = (int16_t)(1 + i); c.bodyB = (int16_t)(1 + j); c.special = 0; c.staticId = 0;
        // the bump / demo callback of this point runs in gather_contacts (leader, solver order): it writes to BOTH cars
        manifold_add(S, m, c, tmul(a.basis, pA - a.pos));
    }
}

}  // namespace ggl
