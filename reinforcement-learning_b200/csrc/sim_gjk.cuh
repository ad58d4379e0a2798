// Closest points / penetration depth between the car hitbox (box core + spherical margin) and one mesh triangle.
//
// Restates, for exactly this shape pair, what the reference gets from its vendored Bullet 3.24
// (paths under GigaLearnCPP/RLGymCPP/RocketSim/libsrc/bullet3-3.24/BulletCollision/NarrowPhaseCollision):
//   btVoronoiSimplexSolver.cpp:28-577     simplex bookkeeping, closest point on segment / triangle / tetrahedron
//   btGjkPairDetector.cpp:686-959         GJK loop on the margin-less shapes, margin handling, degenerate cases, EPA hand-off
//   btGjkEpaPenetrationDepthSolver.cpp:23-79, btGjkEpa2.cpp:60-1010   second GJK + expanding polytope on the shapes WITH margin
//   ../CollisionShapes/btConvexShape.cpp:100-215   support mappings of box and triangle
// The box is body A (world transform basisA/posA), the triangle is body B with identity transform (arena meshes are static
// and un-transformed). Output convention is Bullet's: normal on B pointing from B to A, point on B, signed distance.
//
// Deviation from the reference, stated: the EPA pools hold 64 vertices / 128 faces instead of 128 / 256 (thread-local
// memory); a box-triangle polytope that needs more than 64 support points ends with OutOfVertices earlier than Bullet's would.
#pragma once
#include "sim_math.cuh"

namespace ggl {

struct BoxTri {
    M3 basisA; V3 posA;      // box transform (already re-centred by the pair detector's positionOffset)
    V3 he;                   // box core half extents (btBoxShape implicit dimensions)
    float marginA;           // box collision margin
    V3 t0, t1, t2;           // triangle vertices in B's frame
    V3 posB;                 // translation of B (only the re-centring offset; rotation is identity)
};

GGL_HD V3 box_support(V3 he, V3 d) { return v3(d.x >= 0.f ? he.x : -he.x, d.y >= 0.f ? he.y : -he.y, d.z >= 0.f ? he.z : -he.z); }
GGL_HD V3 tri_support(const BoxTri& s, V3 d) {
    float d0 = dot(d, s.t0), d1 = dot(d, s.t1), d2 = dot(d, s.t2);
    int ax = d0 < d1 ? (d1 < d2 ? 2 : 1) : (d0 < d2 ? 2 : 0);     // btVector3::maxAxis
    return ax == 0 ? s.t0 : (ax == 1 ? s.t1 : s.t2);
}
// localGetSupportVertexNonVirtual: normalised direction, support + margin * direction
GGL_HD V3 norm_dir_for_margin(V3 d) {
    if (len2(d) < kEps * kEps) d = v3(-1.f, -1.f, -1.f);
    return d / len(d);
}

// ------------------------------------------------------------------------------------------- Voronoi simplex solver
struct Voronoi {
    V3 w[5], p[5], q[5];
    int n;
    V3 lastW, cachedP1, cachedP2, cachedV;
    float bc[4];
    bool usedA, usedB, usedC, usedD, degenerate, cachedValid, needsUpdate;
};

GGL_HD void vs_reset(Voronoi& s) {
    s.cachedValid = false; s.n = 0; s.needsUpdate = true;
    s.lastW = v3(1e18f, 1e18f, 1e18f);
    s.degenerate = false; s.bc[0] = s.bc[1] = s.bc[2] = s.bc[3] = 0.f; s.usedA = s.usedB = s.usedC = s.usedD = false;
    s.cachedP1 = s.cachedP2 = s.cachedV = v3(0, 0, 0);
}
GGL_HD void vs_add(Voronoi& s, V3 w, V3 p, V3 q) {
    s.lastW = w; s.needsUpdate = true;
    s.w[s.n] = w; s.p[s.n] = p; s.q[s.n] = q; s.n++;
}
GGL_HD void vs_remove(Voronoi& s, int i) {
    s.n--;
    s.w[i] = s.w[s.n]; s.p[i] = s.p[s.n]; s.q[i] = s.q[s.n];
}
GGL_HD void vs_reduce(Voronoi& s) {
    if (s.n >= 4 && !s.usedD) vs_remove(s, 3);
    if (s.n >= 3 && !s.usedC) vs_remove(s, 2);
    if (s.n >= 2 && !s.usedB) vs_remove(s, 1);
    if (s.n >= 1 && !s.usedA) vs_remove(s, 0);
}
GGL_HD bool vs_in_simplex(const Voronoi& s, V3 w) {
    bool found = false;
    for (int i = 0; i < s.n; i++)
        if (len2(s.w[i] - w) <= 0.0001f) { found = true; break; }
    if (w.x == s.lastW.x && w.y == s.lastW.y && w.z == s.lastW.z) return true;
    return found;
}

struct SubRes { V3 closest; bool uA, uB, uC, uD; float bc[4]; bool degenerate; };
GGL_HD void sr_bc(SubRes& r, float a, float b, float c, float d) { r.bc[0] = a; r.bc[1] = b; r.bc[2] = c; r.bc[3] = d; }

GGL_HD void vs_closest_tri(V3 p, V3 a, V3 b, V3 c, SubRes& r) {
    r.uA = r.uB = r.uC = r.uD = false;
    V3 ab = b - a, ac = c - a, ap = p - a;
    float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) { r.closest = a; r.uA = true; sr_bc(r, 1, 0, 0, 0); return; }
    V3 bp = p - b;
    float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) { r.closest = b; r.uB = true; sr_bc(r, 0, 1, 0, 0); return; }
    float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) {
        float v = d1 / (d1 - d3);
        r.closest = a + v * ab; r.uA = r.uB = true; sr_bc(r, 1 - v, v, 0, 0); return;
    }
    V3 cp = p - c;
    float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.f && d5 <= d6) { r.closest = c; r.uC = true; sr_bc(r, 0, 0, 1, 0); return; }
    float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) {
        float w = d2 / (d2 - d6);
        r.closest = a + w * ac; r.uA = r.uC = true; sr_bc(r, 1 - w, 0, w, 0); return;
    }
    float va = d3 * d6 - d5 * d4;
    if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
        float w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        r.closest = b + w * (c - b); r.uB = r.uC = true; sr_bc(r, 0, 1 - w, w, 0); return;
    }
    float denom = 1.0f / (va + vb + vc);
    float v = vb * denom, w = vc * denom;
    r.closest = a + ab * v + ac * w; r.uA = r.uB = r.uC = true; sr_bc(r, 1 - v - w, v, w, 0);
}

GGL_HD int vs_outside_plane(V3 p, V3 a, V3 b, V3 c, V3 d) {
    V3 n = cross(b - a, c - a);
    float signp = dot(p - a, n), signd = dot(d - a, n);
    if (signd * signd < (1e-4f * 1e-4f)) return -1;
    return signp * signd < 0.f ? 1 : 0;
}

// returns hasSeparation; on false, r.degenerate tells the two failure kinds apart
GGL_HD bool vs_closest_tetra(V3 p, V3 a, V3 b, V3 c, V3 d, SubRes& fin) {
    SubRes t;
    fin.closest = p; fin.uA = fin.uB = fin.uC = fin.uD = true; fin.degenerate = false;
    int oABC = vs_outside_plane(p, a, b, c, d), oACD = vs_outside_plane(p, a, c, d, b);
    int oADB = vs_outside_plane(p, a, d, b, c), oBDC = vs_outside_plane(p, b, d, c, a);
    if (oABC < 0 || oACD < 0 || oADB < 0 || oBDC < 0) { fin.degenerate = true; return false; }
    if (!oABC && !oACD && !oADB && !oBDC) return false;
    float best = 3.402823466e+38f;
    if (oABC) {
        vs_closest_tri(p, a, b, c, t);
        float sq = dot(t.closest - p, t.closest - p);
        if (sq < best) {
            best = sq; fin.closest = t.closest;
            fin.uA = t.uA; fin.uB = t.uB; fin.uC = t.uC; fin.uD = false;
            sr_bc(fin, t.bc[0], t.bc[1], t.bc[2], 0);
        }
    }
    if (oACD) {
        vs_closest_tri(p, a, c, d, t);
        float sq = dot(t.closest - p, t.closest - p);
        if (sq < best) {
            best = sq; fin.closest = t.closest;
            fin.uA = t.uA; fin.uB = false; fin.uC = t.uB; fin.uD = t.uC;
            sr_bc(fin, t.bc[0], 0, t.bc[1], t.bc[2]);
        }
    }
    if (oADB) {
        vs_closest_tri(p, a, d, b, t);
        float sq = dot(t.closest - p, t.closest - p);
        if (sq < best) {
            best = sq; fin.closest = t.closest;
            fin.uA = t.uA; fin.uB = t.uC; fin.uC = false; fin.uD = t.uB;
            sr_bc(fin, t.bc[0], t.bc[2], 0, t.bc[1]);
        }
    }
    if (oBDC) {
        vs_closest_tri(p, b, d, c, t);
        float sq = dot(t.closest - p, t.closest - p);
        if (sq < best) {
            best = sq; fin.closest = t.closest;
            fin.uA = false; fin.uB = t.uA; fin.uC = t.uC; fin.uD = t.uB;
            sr_bc(fin, 0, t.bc[0], t.bc[2], t.bc[1]);
        }
    }
    return true;
}

GGL_HD bool vs_bc_valid(const float* bc) { return bc[0] >= 0.f && bc[1] >= 0.f && bc[2] >= 0.f && bc[3] >= 0.f; }

GGL_HD bool vs_update(Voronoi& s) {
    if (!s.needsUpdate) return s.cachedValid;
    s.degenerate = false; s.bc[0] = s.bc[1] = s.bc[2] = s.bc[3] = 0.f; s.usedA = s.usedB = s.usedC = s.usedD = false;
    s.needsUpdate = false;
    const V3 o = v3(0, 0, 0);
    switch (s.n) {
    case 0: s.cachedValid = false; break;
    case 1:
        s.cachedP1 = s.p[0]; s.cachedP2 = s.q[0]; s.cachedV = s.cachedP1 - s.cachedP2;
        s.bc[0] = 1.f; s.cachedValid = true;
        break;
    case 2: {
        V3 from = s.w[0], to = s.w[1];
        V3 diff = o - from, v = to - from;
        float t = dot(v, diff);
        if (t > 0.f) {
            float dvv = dot(v, v);
            if (t < dvv) { t /= dvv; diff -= t * v; s.usedA = s.usedB = true; }
            else { t = 1.f; diff -= v; s.usedB = true; }
        } else { t = 0.f; s.usedA = true; }
        s.bc[0] = 1 - t; s.bc[1] = t; s.bc[2] = s.bc[3] = 0.f;
        s.cachedP1 = s.p[0] + t * (s.p[1] - s.p[0]);
        s.cachedP2 = s.q[0] + t * (s.q[1] - s.q[0]);
        s.cachedV = s.cachedP1 - s.cachedP2;
        vs_reduce(s);
        s.cachedValid = vs_bc_valid(s.bc);
        break;
    }
    case 3: {
        SubRes r; r.degenerate = false;
        vs_closest_tri(o, s.w[0], s.w[1], s.w[2], r);
        for (int i = 0; i < 4; i++) s.bc[i] = r.bc[i];
        s.usedA = r.uA; s.usedB = r.uB; s.usedC = r.uC; s.usedD = r.uD;
        s.cachedP1 = s.p[0] * s.bc[0] + s.p[1] * s.bc[1] + s.p[2] * s.bc[2];
        s.cachedP2 = s.q[0] * s.bc[0] + s.q[1] * s.bc[1] + s.q[2] * s.bc[2];
        s.cachedV = s.cachedP1 - s.cachedP2;
        vs_reduce(s);
        s.cachedValid = vs_bc_valid(s.bc);
        break;
    }
    case 4: {
        SubRes r;
        bool sep = vs_closest_tetra(o, s.w[0], s.w[1], s.w[2], s.w[3], r);
        if (sep) {
            for (int i = 0; i < 4; i++) s.bc[i] = r.bc[i];
            s.usedA = r.uA; s.usedB = r.uB; s.usedC = r.uC; s.usedD = r.uD;
            s.cachedP1 = s.p[0] * s.bc[0] + s.p[1] * s.bc[1] + s.p[2] * s.bc[2] + s.p[3] * s.bc[3];
            s.cachedP2 = s.q[0] * s.bc[0] + s.q[1] * s.bc[1] + s.q[2] * s.bc[2] + s.q[3] * s.bc[3];
            s.cachedV = s.cachedP1 - s.cachedP2;
            vs_reduce(s);
            s.cachedValid = vs_bc_valid(s.bc);
        } else {
            s.usedA = s.usedB = s.usedC = s.usedD = true;
            if (r.degenerate) { s.degenerate = true; s.cachedValid = false; }
            else { s.cachedValid = true; s.cachedV = o; }
        }
        break;
    }
    default: s.cachedValid = false;
    }
    return s.cachedValid;
}

}  // namespace ggl
