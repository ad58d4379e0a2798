// Bullet's internal-edge contact correction for triangle-mesh contacts
// (BT/BulletCollision/CollisionDispatch/btInternalEdgeUtility.cpp:28-46 helpers, :414-798 btAdjustInternalEdgeContacts,
//  applied to every contact by RocketSim/src/Sim/Arena/Arena.cpp:275-279). Per-triangle edge info is generated on the
// host at mesh load (sim_convert.cuh: world_build_edge_info, restating btInternalEdgeUtility.cpp:47-296,297-354).
#pragma once
#include "sim_types.cuh"

namespace ggl {

// ---------------------------------------------------------------------------------------------- internal edges
GGL_HD V3 nearest_on_segment(V3 p, V3 a, V3 b) {
    V3 d = b - a;
    if (len2(d) < kEps * kEps) return a;       // btVector3::fuzzyZero
    float t = dot(p - a, d) / dot(d, d);
    if (t < 0.f) t = 0.f; else if (t > 1.f) t = 1.f;
    return a + d * t;
}
GGL_HD float edge_get_angle(V3 edgeA, V3 normalA, V3 normalB) { return atan2f(dot(normalB, edgeA), dot(normalB, normalA)); }
GGL_HD V3 quat_rotate(Q4 q, V3 v) {
    // btQuaternion q2 = q * v; q2 *= q.inverse()
    Q4 a;
    a.x = q.w * v.x + q.y * v.z - q.z * v.y;
    a.y = q.w * v.y + q.z * v.x - q.x * v.z;
    a.z = q.w * v.z + q.x * v.y - q.y * v.x;
    a.w = -q.x * v.x - q.y * v.y - q.z * v.z;
    Q4 i; i.x = -q.x; i.y = -q.y; i.z = -q.z; i.w = q.w;
    Q4 r = qmul(a, i);
    return v3(r.x, r.y, r.z);
}
GGL_HD bool clamp_normal(V3 edge, V3 triNormal, V3 localN, float correctedEdgeAngle, V3& clamped) {
    V3 edgeCross = normalized(cross(edge, triNormal));
    float curAngle = edge_get_angle(edgeCross, triNormal, localN);
    if ((correctedEdgeAngle < 0.f && curAngle < correctedEdgeAngle) || (correctedEdgeAngle >= 0.f && curAngle > correctedEdgeAngle)) {
        float diff = correctedEdgeAngle - curAngle;
        clamped = mul(m3_from_quat(quat_axis_angle(edge, diff)), localN);
        return true;
    }
    return false;
}

// Adjusts a freshly added contact against triangle `t` in place (normal, point on B). `posA` = world point on A.
GGL_HD void adjust_internal_edge(const Tri& t, V3& normalOnB, V3& pointOnB, V3 posA, float dist) {
    if (!(t.flags & TRI_HAS_INFO)) return;
    const float kMaxEdgeAngle = 6.283185307179586232f, kEdgeDist = 0.1f, kConvexEps = 0.f;
    V3 v[3] = {t.v0, t.v1, t.v2};
    V3 triNormal = normalized(cross(t.v1 - t.v0, t.v2 - t.v0));
    const V3 contact = pointOnB;
    V3 localN = normalized(normalOnB);
    int best = -1;
    float bestDist = 1e18f;
    for (int e = 0; e < 3; e++) {
        if (fabsf(t.edgeAngle[e]) < kMaxEdgeAngle) {
            float l = len(contact - nearest_on_segment(contact, v[e], v[(e + 1) % 3]));
            if (l < bestDist) { best = e; bestDist = l; }
        }
    }
    if (best < 0) return;
    bool isNearEdge = false;
    int concaveHits = 0;
    {
        const int e = best;
        const float ang = t.edgeAngle[e];
        float l = len(contact - nearest_on_segment(contact, v[e], v[(e + 1) % 3]));
        if (l < kEdgeDist) {
            V3 edge = v[e] - v[(e + 1) % 3];
            isNearEdge = true;
            if (ang == 0.f) {
                concaveHits++;
            } else {
                bool convex = (t.flags & (TRI_CONVEX_V0V1 << e)) != 0;
                float swap = convex ? 1.f : -1.f;
                V3 nA = swap * triNormal;
                V3 computedB = quat_rotate(quat_axis_angle(edge, ang), triNormal);
                if (t.flags & (TRI_SWAP_V0V1 << e)) computedB = computedB * -1.f;
                V3 nB = swap * computedB;
                float NdotA = dot(localN, nA), NdotB = dot(localN, nB);
                bool backFacing = (NdotA < kConvexEps) && (NdotB < kConvexEps);
                if (backFacing) {
                    concaveHits++;
                } else {
                    V3 cl;
                    V3 ln = (e == 0) ? localN : normalOnB;       // edges 1 and 2 re-read the un-normalised contact normal
                    if (clamp_normal(edge, swap * triNormal, ln, ang, cl)) {
                        if (dot(cl, triNormal) > 0.f) {
                            normalOnB = cl;
                            pointOnB = posA - normalOnB * dist;
                        }
                    }
                }
            }
        }
    }
    if (isNearEdge && concaveHits > 0) {
        float d = dot(triNormal, localN);
        if (d < 0.f) return;
        normalOnB = triNormal;
        pointOnB = posA - normalOnB * dist;
    }
}

}  // namespace ggl
