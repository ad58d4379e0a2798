// Penetration depth of box (with margin) vs triangle: the gjkepa2 GJK + expanding-polytope pair
// (BT/BulletCollision/NarrowPhaseCollision/btGjkEpa2.cpp:60-1010, driver btGjkEpaPenetrationDepthSolver.cpp:23-79).
// Pointer-linked simplices / face lists of the reference become index-linked pools so the state lives in thread-local
// memory; list orders (LIFO stock and hull lists) are kept because findbest() breaks ties by list order.
#pragma once
#include "sim_gjk.cuh"

namespace ggl {

constexpr int kEpaMaxVerts = 64;
constexpr int kEpaMaxFaces = 2 * kEpaMaxVerts;
constexpr float kGjk2Accuracy = 0.0001f, kGjk2MinDist = 0.0001f, kGjk2DupEps = 0.0001f;
constexpr float kEpaAccuracy = 0.0001f, kEpaPlaneEps = 0.00001f;
constexpr uint8_t kNil = 0xFF;

struct SV { V3 d, w; };

// MinkowskiDiff in A's local frame (btGjkEpa2.cpp:904-921): toshape1 = B^T A = A (B has identity rotation),
// toshape0 = A^-1 B
struct Mink {
    const BoxTri* s;
    bool margins;
    V3 originBinA;     // (posB - posA) expressed in A's frame
};
GGL_HD Mink mink_make(const BoxTri& s, bool margins) {
    Mink m; m.s = &s; m.margins = margins; m.originBinA = tmul(s.basisA, s.posB - s.posA);
    return m;
}
GGL_HD V3 mink_support0(const Mink& m, V3 d) {
    if (m.margins) { V3 n = norm_dir_for_margin(d); return box_support(m.s->he, n) + m.s->marginA * n; }
    return box_support(m.s->he, d);
}
GGL_HD V3 mink_support1(const Mink& m, V3 d) {
    V3 dB = mul(m.s->basisA, d);
    V3 sup;
    if (m.margins) { V3 n = norm_dir_for_margin(dB); sup = tri_support(*m.s, n) + 0.f * n; }
    else sup = tri_support(*m.s, dB);
    return tmul(m.s->basisA, sup) + m.originBinA;
}
GGL_HD V3 mink_support(const Mink& m, V3 d) { return mink_support0(m, d) - mink_support1(m, -d); }

struct Gjk2 {
    Mink shape;
    V3 ray;
    float distance;
    SV store[4];
    uint8_t freeIdx[4]; int nfree;
    uint8_t sc[2][4]; float sp[2][4]; int srank[2];   // two simplices of store indices
    int current;
    int status;       // 0 Valid, 1 Inside, 2 Failed
};

GGL_HD void g2_getsupport(const Gjk2& g, V3 d, SV& sv) { sv.d = d / len(d); sv.w = mink_support(g.shape, sv.d); }
GGL_HD void g2_append(Gjk2& g, int si, V3 v) {
    g.sp[si][g.srank[si]] = 0.f;
    uint8_t idx = g.freeIdx[--g.nfree];
    g.sc[si][g.srank[si]] = idx;
    g2_getsupport(g, v, g.store[idx]);
    g.srank[si]++;
}
GGL_HD void g2_remove(Gjk2& g, int si) { g.freeIdx[g.nfree++] = g.sc[si][--g.srank[si]]; }

GGL_HD float det3(V3 a, V3 b, V3 c) {
    return (a.y * b.z * c.x + a.z * b.x * c.y - a.x * b.z * c.y - a.y * b.x * c.z + a.x * b.y * c.z - a.z * b.y * c.x);
}

GGL_HD float proj2(V3 a, V3 b, float* w, unsigned& m) {
    V3 d = b - a;
    float l = len2(d);
    if (l > 0.f) {
        float t = l > 0.f ? -dot(a, d) / l : 0.f;
        if (t >= 1.f) { w[0] = 0; w[1] = 1; m = 2; return len2(b); }
        else if (t <= 0.f) { w[0] = 1; w[1] = 0; m = 1; return len2(a); }
        else { w[0] = 1 - (w[1] = t); m = 3; return len2(a + d * t); }
    }
    return -1.f;
}
GGL_HD float proj3(V3 a, V3 b, V3 c, float* w, unsigned& m) {
    const int imd3[3] = {1, 2, 0};
    V3 vt[3] = {a, b, c};
    V3 dl[3] = {a - b, b - c, c - a};
    V3 n = cross(dl[0], dl[1]);
    float l = len2(n);
    if (l > 0.f) {
        float mindist = -1.f;
        float subw[2] = {0.f, 0.f};
        unsigned subm = 0;
        for (int i = 0; i < 3; i++) {
            if (dot(vt[i], cross(dl[i], n)) > 0.f) {
                int j = imd3[i];
                float subd = proj2(vt[i], vt[j], subw, subm);
                if (mindist < 0.f || subd < mindist) {
                    mindist = subd;
                    m = ((subm & 1) ? 1u << i : 0u) + ((subm & 2) ? 1u << j : 0u);
                    w[i] = subw[0]; w[j] = subw[1]; w[imd3[j]] = 0.f;
                }
            }
        }
        if (mindist < 0.f) {
            float d = dot(a, n);
            float s = sqrtf(l);
            V3 p = n * (d / l);
            mindist = len2(p);
            m = 7;
            w[0] = len(cross(dl[1], b - p)) / s;
            w[1] = len(cross(dl[2], c - p)) / s;
            w[2] = 1 - (w[0] + w[1]);
        }
        return mindist;
    }
    return -1.f;
}
GGL_HD float proj4(V3 a, V3 b, V3 c, V3 d, float* w, unsigned& m) {
    const int imd3[3] = {1, 2, 0};
    V3 vt[4] = {a, b, c, d};
    V3 dl[3] = {a - d, b - d, c - d};
    float vl = det3(dl[0], dl[1], dl[2]);
    bool ng = (vl * dot(a, cross(b - c, a - b))) <= 0.f;
    if (ng && fabsf(vl) > 0.f) {
        float mindist = -1.f;
        float subw[3] = {0.f, 0.f, 0.f};
        unsigned subm = 0;
        for (int i = 0; i < 3; i++) {
            int j = imd3[i];
            float s = vl * dot(d, cross(dl[i], dl[j]));
            if (s > 0.f) {
                float subd = proj3(vt[i], vt[j], d, subw, subm);
                if (mindist < 0.f || subd < mindist) {
                    mindist = subd;
                    m = ((subm & 1) ? 1u << i : 0u) + ((subm & 2) ? 1u << j : 0u) + ((subm & 4) ? 8u : 0u);
                    w[i] = subw[0]; w[j] = subw[1]; w[imd3[j]] = 0.f; w[3] = subw[2];
                }
            }
        }
        if (mindist < 0.f) {
            mindist = 0.f; m = 15;
            w[0] = det3(c, b, d) / vl; w[1] = det3(a, c, d) / vl; w[2] = det3(b, a, d) / vl;
            w[3] = 1 - (w[0] + w[1] + w[2]);
        }
        return mindist;
    }
    return -1.f;
}

GGL_HDN int g2_evaluate(Gjk2& g, const Mink& shape, V3 guess) {
    int iterations = 0;
    float sqdist = 0.f, alpha = 0.f;
    V3 lastw[4];
    int clastw = 0;
    for (int i = 0; i < 4; i++) g.freeIdx[i] = (uint8_t)i;
    g.nfree = 4; g.current = 0; g.status = 0; g.shape = shape; g.distance = 0.f;
    g.srank[0] = 0; g.srank[1] = 0;
    g.ray = guess;
    float sqrl = len2(g.ray);
    g2_append(g, 0, sqrl > 0.f ? -g.ray : v3(1, 0, 0));
    g.sp[0][0] = 1.f;
    g.ray = g.store[g.sc[0][0]].w;
    sqdist = sqrl;
    lastw[0] = lastw[1] = lastw[2] = lastw[3] = g.ray;
    do {
        int next = 1 - g.current, cs = g.current;
        float rl = len(g.ray);
        if (rl < kGjk2MinDist) { g.status = 1; break; }
        g2_append(g, cs, -g.ray);
        V3 w = g.store[g.sc[cs][g.srank[cs] - 1]].w;
        bool found = false;
        for (int i = 0; i < 4; i++)
            if (len2(w - lastw[i]) < kGjk2DupEps) { found = true; break; }
        if (found) { g2_remove(g, cs); break; }
        lastw[clastw = (clastw + 1) & 3] = w;
        float omega = dot(g.ray, w) / rl;
        alpha = fmaxf(omega, alpha);
        if (((rl - alpha) - (kGjk2Accuracy * rl)) <= 0.f) { g2_remove(g, cs); break; }
        float weights[4];
        unsigned mask = 0;
        const uint8_t* c = g.sc[cs];
        switch (g.srank[cs]) {
        case 2: sqdist = proj2(g.store[c[0]].w, g.store[c[1]].w, weights, mask); break;
        case 3: sqdist = proj3(g.store[c[0]].w, g.store[c[1]].w, g.store[c[2]].w, weights, mask); break;
        case 4: sqdist = proj4(g.store[c[0]].w, g.store[c[1]].w, g.store[c[2]].w, g.store[c[3]].w, weights, mask); break;
        }
        if (sqdist >= 0.f) {
            g.srank[next] = 0;
            g.ray = v3(0, 0, 0);
            g.current = next;
            for (int i = 0, ni = g.srank[cs]; i < ni; i++) {
                if (mask & (1u << i)) {
                    g.sc[next][g.srank[next]] = c[i];
                    g.sp[next][g.srank[next]++] = weights[i];
                    g.ray += g.store[c[i]].w * weights[i];
                } else {
                    g.freeIdx[g.nfree++] = c[i];
                }
            }
            if (mask == 15) g.status = 1;
        } else {
            g2_remove(g, cs);
            break;
        }
        g.status = ((++iterations) < 128) ? g.status : 2;
    } while (g.status == 0);
    if (g.status == 0) g.distance = len(g.ray);
    else if (g.status == 1) g.distance = 0.f;
    return g.status;
}

// EncloseOrigin (btGjkEpa2.cpp:287-352) on the current simplex, unrolled by rank instead of recursive
GGL_HD bool g2_enclose4(Gjk2& g) {
    const int si = g.current; const uint8_t* c = g.sc[si];
    return fabsf(det3(g.store[c[0]].w - g.store[c[3]].w, g.store[c[1]].w - g.store[c[3]].w, g.store[c[2]].w - g.store[c[3]].w)) > 0.f;
}
GGL_HD bool g2_enclose3(Gjk2& g) {
    const int si = g.current; const uint8_t* c = g.sc[si];
    V3 n = cross(g.store[c[1]].w - g.store[c[0]].w, g.store[c[2]].w - g.store[c[0]].w);
    if (len2(n) > 0.f) {
        g2_append(g, si, n);
        if (g2_enclose4(g)) return true;
        g2_remove(g, si);
        g2_append(g, si, -n);
        if (g2_enclose4(g)) return true;
        g2_remove(g, si);
    }
    return false;
}
GGL_HD bool g2_enclose2(Gjk2& g) {
    const int si = g.current; const uint8_t* c = g.sc[si];
    V3 d = g.store[c[1]].w - g.store[c[0]].w;
    for (int i = 0; i < 3; i++) {
        V3 axis = v3(0, 0, 0);
        set_comp(axis, i, 1.f);
        V3 p = cross(d, axis);
        if (len2(p) > 0.f) {
            g2_append(g, si, p);
            if (g2_enclose3(g)) return true;
            g2_remove(g, si);
            g2_append(g, si, -p);
            if (g2_enclose3(g)) return true;
            g2_remove(g, si);
        }
    }
    return false;
}
GGL_HD bool g2_enclose(Gjk2& g) {
    switch (g.srank[g.current]) {
    case 2: return g2_enclose2(g);
    case 3: return g2_enclose3(g);
    case 4: return g2_enclose4(g);
    }
    return false;     // rank 1 is never passed in (EPA::Evaluate tests rank > 1 first)
}


// ------------------------------------------------------------------------------------------- EPA
struct EpaFace {
    V3 n; float d;
    uint8_t c[3], f[3], l[2], e[3], pass;
};
struct Epa {
    int status;        // 0 Valid 1 Touching 2 Degenerated 3 NonConvex 4 InvalidHull 5 OutOfFaces 6 OutOfVertices 7 AccuracyReached 8 FallBack 9 Failed
    V3 normal; float depth;
    SV sv[kEpaMaxVerts + 4];          // [0..3] = the GJK simplex, [4..] = vertices added by the expansion
    EpaFace fc[kEpaMaxFaces];
    int nextsv;
    uint8_t hullRoot, stockRoot; int hullCount, stockCount;
    uint8_t resC[3]; float resP[3]; int resRank;
};
struct EpaHorizon { uint8_t cf, ff; int nf; };

GGL_HD void epa_bind(Epa& E, uint8_t fa, int ea, uint8_t fb, int eb) {
    E.fc[fa].e[ea] = (uint8_t)eb; E.fc[fa].f[ea] = fb;
    E.fc[fb].e[eb] = (uint8_t)ea; E.fc[fb].f[eb] = fa;
}
GGL_HD void epa_append(Epa& E, uint8_t& root, int& count, uint8_t face) {
    E.fc[face].l[0] = kNil; E.fc[face].l[1] = root;
    if (root != kNil) E.fc[root].l[0] = face;
    root = face; ++count;
}
GGL_HD void epa_remove(Epa& E, uint8_t& root, int& count, uint8_t face) {
    EpaFace& f = E.fc[face];
    if (f.l[1] != kNil) E.fc[f.l[1]].l[0] = f.l[0];
    if (f.l[0] != kNil) E.fc[f.l[0]].l[1] = f.l[1];
    if (face == root) root = f.l[1];
    --count;
}
GGL_HD bool epa_edgedist(const Epa& E, const EpaFace& face, uint8_t ia, uint8_t ib, float& dist) {
    V3 a = E.sv[ia].w, b = E.sv[ib].w;
    V3 ba = b - a;
    V3 n_ab = cross(ba, face.n);
    float a_dot_nab = dot(a, n_ab);
    if (a_dot_nab < 0.f) {
        float ba_l2 = len2(ba), a_dot_ba = dot(a, ba), b_dot_ba = dot(b, ba);
        if (a_dot_ba > 0.f) dist = len(a);
        else if (b_dot_ba < 0.f) dist = len(b);
        else {
            float a_dot_b = dot(a, b);
            dist = sqrtf(fmaxf((len2(a) * len2(b) - a_dot_b * a_dot_b) / ba_l2, 0.f));
        }
        return true;
    }
    return false;
}
GGL_HD uint8_t epa_newface(Epa& E, uint8_t a, uint8_t b, uint8_t c, bool forced) {
    if (E.stockRoot != kNil) {
        uint8_t fi = E.stockRoot;
        epa_remove(E, E.stockRoot, E.stockCount, fi);
        epa_append(E, E.hullRoot, E.hullCount, fi);
        EpaFace& face = E.fc[fi];
        face.pass = 0; face.c[0] = a; face.c[1] = b; face.c[2] = c;
        face.n = cross(E.sv[b].w - E.sv[a].w, E.sv[c].w - E.sv[a].w);
        float l = len(face.n);
        bool v = l > kEpaAccuracy;
        if (v) {
            if (!(epa_edgedist(E, face, a, b, face.d) || epa_edgedist(E, face, b, c, face.d) || epa_edgedist(E, face, c, a, face.d)))
                face.d = dot(E.sv[a].w, face.n) / l;
            face.n = face.n / l;
            if (forced || face.d >= -kEpaPlaneEps) return fi;
            E.status = 3;
        } else {
            E.status = 2;
        }
        epa_remove(E, E.hullRoot, E.hullCount, fi);
        epa_append(E, E.stockRoot, E.stockCount, fi);
        return kNil;
    }
    E.status = 5;
    return kNil;
}
GGL_HD uint8_t epa_findbest(const Epa& E) {
    uint8_t minf = E.hullRoot;
    float mind = E.fc[minf].d * E.fc[minf].d;
    for (uint8_t f = E.fc[minf].l[1]; f != kNil; f = E.fc[f].l[1]) {
        float sqd = E.fc[f].d * E.fc[f].d;
        if (sqd < mind) { minf = f; mind = sqd; }
    }
    return minf;
}

// EPA::expand (btGjkEpa2.cpp:866-900), recursion replaced by an explicit frame stack
GGL_HD bool epa_expand(Epa& E, unsigned pass, uint8_t w, uint8_t f0, int e0, EpaHorizon& hz) {
    const int i1m3[3] = {1, 2, 0}, i2m3[3] = {2, 0, 1};
    struct Frame { uint8_t f, e, stage; };
    Frame st[kEpaMaxFaces];
    int sp = 0;
    st[0].f = f0; st[0].e = (uint8_t)e0; st[0].stage = 0;
    bool ret = false;
    while (sp >= 0) {
        Frame& fr = st[sp];
        EpaFace& F = E.fc[fr.f];
        if (fr.stage == 0) {
            if (F.pass == (uint8_t)pass) { ret = false; sp--; continue; }
            const int e1 = i1m3[fr.e];
            if ((dot(F.n, E.sv[w].w) - F.d) < -kEpaPlaneEps) {
                uint8_t nf = epa_newface(E, F.c[e1], F.c[fr.e], w, false);
                if (nf != kNil) {
                    epa_bind(E, nf, 0, fr.f, fr.e);
                    if (hz.cf != kNil) epa_bind(E, hz.cf, 1, nf, 2); else hz.ff = nf;
                    hz.cf = nf; ++hz.nf;
                    ret = true;
                } else ret = false;
                sp--; continue;
            }
            F.pass = (uint8_t)pass;
            fr.stage = 1;
            if (sp + 1 >= kEpaMaxFaces) { ret = false; sp--; continue; }
            st[sp + 1].f = F.f[e1]; st[sp + 1].e = F.e[e1]; st[sp + 1].stage = 0;
            sp++;
        } else if (fr.stage == 1) {
            if (!ret) { sp--; continue; }     // && short-circuit: second expand not evaluated
            const int e2 = i2m3[fr.e];
            fr.stage = 2;
            st[sp + 1].f = F.f[e2]; st[sp + 1].e = F.e[e2]; st[sp + 1].stage = 0;
            sp++;
        } else {
            if (ret) {
                epa_remove(E, E.hullRoot, E.hullCount, fr.f);
                epa_append(E, E.stockRoot, E.stockCount, fr.f);
            }
            sp--;
        }
    }
    return ret;
}

GGL_HDN int epa_evaluate(Epa& E, Gjk2& g, V3 guess) {
    E.status = 9; E.normal = v3(0, 0, 0); E.depth = 0.f; E.nextsv = 0;
    E.hullRoot = kNil; E.stockRoot = kNil; E.hullCount = 0; E.stockCount = 0;
    for (int i = 0; i < kEpaMaxFaces; i++) epa_append(E, E.stockRoot, E.stockCount, (uint8_t)(kEpaMaxFaces - i - 1));
    const int si = g.current;
    if (g.srank[si] > 1 && g2_enclose(g)) {
        E.status = 0; E.nextsv = 0;
        // the simplex (now a tetrahedron) becomes pool vertices 0..3
        uint8_t c[4] = {0, 1, 2, 3};
        for (int i = 0; i < 4; i++) E.sv[i] = g.store[g.sc[si][i]];
        if (det3(E.sv[0].w - E.sv[3].w, E.sv[1].w - E.sv[3].w, E.sv[2].w - E.sv[3].w) < 0.f) {
            SV t = E.sv[0]; E.sv[0] = E.sv[1]; E.sv[1] = t;
        }
        uint8_t tetra[4] = {epa_newface(E, c[0], c[1], c[2], true), epa_newface(E, c[1], c[0], c[3], true),
                            epa_newface(E, c[2], c[1], c[3], true), epa_newface(E, c[0], c[2], c[3], true)};
        if (E.hullCount == 4) {
            uint8_t best = epa_findbest(E);
            EpaFace outer = E.fc[best];
            unsigned pass = 0;
            epa_bind(E, tetra[0], 0, tetra[1], 0);
            epa_bind(E, tetra[0], 1, tetra[2], 0);
            epa_bind(E, tetra[0], 2, tetra[3], 0);
            epa_bind(E, tetra[1], 1, tetra[3], 2);
            epa_bind(E, tetra[1], 2, tetra[2], 1);
            epa_bind(E, tetra[2], 2, tetra[3], 1);
            E.status = 0;
            for (int it = 0; it < 255; ++it) {
                if (E.nextsv < kEpaMaxVerts) {
                    EpaHorizon hz; hz.cf = kNil; hz.ff = kNil; hz.nf = 0;
                    uint8_t w = (uint8_t)(4 + E.nextsv++);
                    bool valid = true;
                    E.fc[best].pass = (uint8_t)(++pass);
                    g2_getsupport(g, E.fc[best].n, E.sv[w]);
                    float wdist = dot(E.fc[best].n, E.sv[w].w) - E.fc[best].d;
                    if (wdist > kEpaAccuracy) {
                        for (int j = 0; j < 3 && valid; ++j) valid &= epa_expand(E, pass, w, E.fc[best].f[j], E.fc[best].e[j], hz);
                        if (valid && hz.nf >= 3) {
                            epa_bind(E, hz.cf, 1, hz.ff, 2);
                            epa_remove(E, E.hullRoot, E.hullCount, best);
                            epa_append(E, E.stockRoot, E.stockCount, best);
                            best = epa_findbest(E);
                            outer = E.fc[best];
                        } else { E.status = 4; break; }
                    } else { E.status = 7; break; }
                } else { E.status = 6; break; }
            }
            V3 proj = outer.n * outer.d;
            E.normal = outer.n; E.depth = outer.d; E.resRank = 3;
            E.resC[0] = outer.c[0]; E.resC[1] = outer.c[1]; E.resC[2] = outer.c[2];
            E.resP[0] = len(cross(E.sv[outer.c[1]].w - proj, E.sv[outer.c[2]].w - proj));
            E.resP[1] = len(cross(E.sv[outer.c[2]].w - proj, E.sv[outer.c[0]].w - proj));
            E.resP[2] = len(cross(E.sv[outer.c[0]].w - proj, E.sv[outer.c[1]].w - proj));
            float sum = E.resP[0] + E.resP[1] + E.resP[2];
            E.resP[0] /= sum; E.resP[1] /= sum; E.resP[2] /= sum;
            return E.status;
        }
    }
    E.status = 8;
    E.normal = -guess;
    float nl = len(E.normal);
    if (nl > 0.f) E.normal = E.normal / nl; else E.normal = v3(1, 0, 0);
    E.depth = 0.f; E.resRank = 1;
    E.sv[0] = g.store[g.sc[g.current][0]];
    E.resC[0] = 0; E.resP[0] = 1.f;
    return E.status;
}

struct PenResult { V3 witA, witB, normal; float distance; };

// btGjkEpaSolver2::Penetration (usemargins = true) / ::Distance; witnesses in the (re-centred) world frame of the pair detector
GGL_HDN bool epa_penetration(const BoxTri& s, V3 guess, PenResult& r) {
    Mink shape = mink_make(s, true);
    Gjk2 g;
    int st = g2_evaluate(g, shape, -guess);
    if (st == 1) {
        Epa E;
        int es = epa_evaluate(E, g, -guess);
        if (es != 9) {
            V3 w0 = v3(0, 0, 0);
            for (int i = 0; i < E.resRank; i++) w0 += mink_support0(shape, E.sv[E.resC[i]].d) * E.resP[i];
            r.witA = mul(s.basisA, w0) + s.posA;
            r.witB = mul(s.basisA, w0 - E.normal * E.depth) + s.posA;
            r.normal = -E.normal; r.distance = -E.depth;
            return true;
        }
    }
    return false;
}
GGL_HDN bool epa_distance(const BoxTri& s, V3 guess, PenResult& r) {
    Mink shape = mink_make(s, false);
    Gjk2 g;
    int st = g2_evaluate(g, shape, guess);
    if (st == 0) {
        V3 w0 = v3(0, 0, 0), w1 = v3(0, 0, 0);
        const int si = g.current;
        for (int i = 0; i < g.srank[si]; i++) {
            float p = g.sp[si][i];
            w0 += mink_support0(shape, g.store[g.sc[si][i]].d) * p;
            w1 += mink_support1(shape, -g.store[g.sc[si][i]].d) * p;
        }
        r.witA = mul(s.basisA, w0) + s.posA;
        r.witB = mul(s.basisA, w1) + s.posA;
        r.normal = w0 - w1;
        r.distance = len(r.normal);
        r.normal = r.normal / (r.distance > kGjk2MinDist ? r.distance : 1.f);
        return true;
    }
    return false;
}

}  // namespace ggl
