// Small float3 / 3x3 math used by the arena simulator. Written so that every scalar expression rounds like the
// reference's float math when the TU is compiled with -fmad=false (nvcc) / -ffp-contract=off (host debug build).
// GGL_HD lets the same header compile for the device (product) and for the host lock-step debugger used by the
// CPU test-suite (tests/hostsim — test infrastructure, never loaded by the product library).
#pragma once
#include <string.h>
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define GGL_HD __host__ __device__ __forceinline__
#define GGL_HDN __host__ __device__ __noinline__
#else
#define GGL_HD inline
#define GGL_HDN inline
#endif

namespace ggl {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kHalfPi = 1.57079632679489661923f;
constexpr float kEps = 1.1920928955078125e-7f;   // FLT_EPSILON (SIMD_EPSILON)
constexpr float kSqrt12 = 0.7071067811865475244f;

struct V3 {
    float x, y, z;
};

GGL_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
GGL_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
GGL_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
GGL_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
GGL_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
GGL_HD V3 operator*(float s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
GGL_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
GGL_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
GGL_HD V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
GGL_HD V3& operator-=(V3& a, V3 b) { a = a - b; return a; }
GGL_HD V3& operator*=(V3& a, float s) { a = a * s; return a; }
GGL_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// bit casts that compile on both sides (the host build of the simulator is g++)
GGL_HD float __int_as_float_portable(int i) { float f; memcpy(&f, &i, 4); return f; }
GGL_HD int __float_as_int_portable(float f) { int i; memcpy(&i, &f, 4); return i; }
GGL_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
GGL_HD float len2(V3 a) { return dot(a, a); }
GGL_HD float len(V3 a) { return sqrtf(len2(a)); }
GGL_HD bool is_zero(V3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }
GGL_HD float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
GGL_HD void set_comp(V3& a, int i, float v) { if (i == 0) a.x = v; else if (i == 1) a.y = v; else a.z = v; }
GGL_HD float fclamp(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
GGL_HD float fsgn(float v) { return (float)((0.f < v) - (v < 0.f)); }   // RS_SGN
GGL_HD V3 vmin(V3 a, V3 b) { return v3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
GGL_HD V3 vmax(V3 a, V3 b) { return v3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
GGL_HD V3 vabs(V3 a) { return v3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }

// btVector3::normalized(): v / length  (Bullet multiplies by the reciprocal)
GGL_HD V3 normalized(V3 a) { return a * (1.f / len(a)); }
// btVector3::safeNormalize(): scale by max |component| first, zero -> (1,0,0)
GGL_HD V3 safe_normalized(V3 a) {
    float l2 = len2(a);
    if (l2 >= kEps * kEps) return a * (1.f / sqrtf(l2));
    return v3(1.f, 0.f, 0.f);
}
// RocketSim Vec::Normalized(): true division, zero vector when tiny
GGL_HD V3 rs_normalized(V3 a) {
    float l2 = len2(a);
    float l = l2 > 0.f ? sqrtf(l2) : 0.f;
    if (l > kEps * kEps) return a / l;
    return v3(0.f, 0.f, 0.f);
}

// Row-major 3x3 (same convention as a rotation "basis": columns are the body's forward / right / up axes).
struct M3 {
    V3 r0, r1, r2;
};
GGL_HD M3 m3_identity() { M3 m; m.r0 = v3(1, 0, 0); m.r1 = v3(0, 1, 0); m.r2 = v3(0, 0, 1); return m; }
GGL_HD V3 mul(const M3& m, V3 v) { return v3(dot(m.r0, v), dot(m.r1, v), dot(m.r2, v)); }
// v * M  ==  M^T v
GGL_HD V3 tmul(const M3& m, V3 v) {
    return v3(m.r0.x * v.x + m.r1.x * v.y + m.r2.x * v.z,
              m.r0.y * v.x + m.r1.y * v.y + m.r2.y * v.z,
              m.r0.z * v.x + m.r1.z * v.y + m.r2.z * v.z);
}
GGL_HD V3 col(const M3& m, int i) { return v3(comp(m.r0, i), comp(m.r1, i), comp(m.r2, i)); }
GGL_HD M3 transpose(const M3& m) { M3 t; t.r0 = col(m, 0); t.r1 = col(m, 1); t.r2 = col(m, 2); return t; }
GGL_HD M3 mul(const M3& a, const M3& b) {
    M3 t = transpose(b), r;
    r.r0 = v3(dot(a.r0, t.r0), dot(a.r0, t.r1), dot(a.r0, t.r2));
    r.r1 = v3(dot(a.r1, t.r0), dot(a.r1, t.r1), dot(a.r1, t.r2));
    r.r2 = v3(dot(a.r2, t.r0), dot(a.r2, t.r1), dot(a.r2, t.r2));
    return r;
}
// columns scaled by s (btMatrix3x3::scaled)
GGL_HD M3 scaled(const M3& m, V3 s) { M3 r; r.r0 = m.r0 * s; r.r1 = m.r1 * s; r.r2 = m.r2 * s; return r; }
GGL_HD M3 inverse(const M3& m) {
    V3 co = v3(m.r1.y * m.r2.z - m.r1.z * m.r2.y, m.r1.z * m.r2.x - m.r1.x * m.r2.z, m.r1.x * m.r2.y - m.r1.y * m.r2.x);
    float det = dot(m.r0, co);
    float s = 1.f / det;
    M3 r;
    r.r0 = v3(co.x * s, (m.r0.z * m.r2.y - m.r0.y * m.r2.z) * s, (m.r0.y * m.r1.z - m.r0.z * m.r1.y) * s);
    r.r1 = v3(co.y * s, (m.r0.x * m.r2.z - m.r0.z * m.r2.x) * s, (m.r0.z * m.r1.x - m.r0.x * m.r1.z) * s);
    r.r2 = v3(co.z * s, (m.r0.y * m.r2.x - m.r0.x * m.r2.y) * s, (m.r0.x * m.r1.y - m.r0.y * m.r1.x) * s);
    return r;
}

struct Q4 {
    float x, y, z, w;
};
GGL_HD Q4 quat_axis_angle(V3 axis, float angle) {
    float d = len(axis);
    float s = sinf(angle * 0.5f) / d;
    Q4 q; q.x = axis.x * s; q.y = axis.y * s; q.z = axis.z * s; q.w = cosf(angle * 0.5f);
    return q;
}
GGL_HD Q4 qmul(Q4 a, Q4 b) {
    Q4 q;
    q.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    q.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
    q.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
    q.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    return q;
}
GGL_HD float qlen2(Q4 q) { return q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w; }
GGL_HD M3 m3_from_quat(Q4 q) {
    float d = qlen2(q);
    float s = 2.f / d;
    float xs = q.x * s, ys = q.y * s, zs = q.z * s;
    float wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
    float xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
    float yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
    M3 m;
    m.r0 = v3(1.f - (yy + zz), xy - wz, xz + wy);
    m.r1 = v3(xy + wz, 1.f - (xx + zz), yz - wx);
    m.r2 = v3(xz - wy, yz + wx, 1.f - (xx + yy));
    return m;
}
// btMatrix3x3::getRotation. The three cases of the non-positive-trace branch are written out (Bullet indexes m_el[i][j] with
// run-time i, j, k): a dynamically indexed temporary would live on the thread's stack, and this runs for every car every tick.
GGL_HD Q4 quat_from_m3(const M3& m) {
    float trace = m.r0.x + m.r1.y + m.r2.z;
    Q4 q;
    if (trace > 0.f) {
        float s = sqrtf(trace + 1.f);
        q.w = s * 0.5f;
        s = 0.5f / s;
        q.x = (m.r2.y - m.r1.z) * s;
        q.y = (m.r0.z - m.r2.x) * s;
        q.z = (m.r1.x - m.r0.y) * s;
    } else {
        const int i = m.r0.x < m.r1.y ? (m.r1.y < m.r2.z ? 2 : 1) : (m.r0.x < m.r2.z ? 2 : 0);
        if (i == 0) {           // j = 1, k = 2
            float s = sqrtf(m.r0.x - m.r1.y - m.r2.z + 1.f);
            q.x = s * 0.5f;
            s = 0.5f / s;
            q.w = (m.r2.y - m.r1.z) * s;
            q.y = (m.r1.x + m.r0.y) * s;
            q.z = (m.r2.x + m.r0.z) * s;
        } else if (i == 1) {    // j = 2, k = 0
            float s = sqrtf(m.r1.y - m.r2.z - m.r0.x + 1.f);
            q.y = s * 0.5f;
            s = 0.5f / s;
            q.w = (m.r0.z - m.r2.x) * s;
            q.z = (m.r2.y + m.r1.z) * s;
            q.x = (m.r0.y + m.r1.x) * s;
        } else {                // j = 0, k = 1
            float s = sqrtf(m.r2.z - m.r0.x - m.r1.y + 1.f);
            q.z = s * 0.5f;
            s = 0.5f / s;
            q.w = (m.r1.x - m.r0.y) * s;
            q.x = (m.r0.z + m.r2.x) * s;
            q.y = (m.r1.z + m.r2.y) * s;
        }
    }
    return q;
}

// yaw/pitch/roll <-> matrix, Bullet's setEulerYPR / getEulerYPR conventions
GGL_HD M3 m3_from_euler_ypr(float yaw, float pitch, float roll) {
    float ci = cosf(roll), cj = cosf(pitch), ch = cosf(yaw);
    float si = sinf(roll), sj = sinf(pitch), sh = sinf(yaw);
    float cc = ci * ch, cs = ci * sh, sc = si * ch, ss = si * sh;
    M3 m;
    m.r0 = v3(cj * ch, sj * sc - cs, sj * cc + ss);
    m.r1 = v3(cj * sh, sj * ss + cc, sj * cs - sc);
    m.r2 = v3(-sj, cj * si, cj * ci);
    return m;
}
GGL_HD void euler_ypr_from_m3(const M3& m, float& yaw, float& pitch, float& roll) {
    yaw = atan2f(m.r1.x, m.r0.x);
    pitch = asinf(-m.r2.x);
    roll = atan2f(m.r2.y, m.r2.z);
    if (fabsf(pitch) == kHalfPi) {
        if (yaw > 0) yaw -= kPi; else yaw += kPi;
        if (roll > 0) roll -= kPi; else roll += kPi;
    }
}

// two tangents of a unit normal (btPlaneSpace1)
GGL_HD void plane_space(V3 n, V3& p, V3& q) {
    if (fabsf(n.z) > kSqrt12) {
        float a = n.y * n.y + n.z * n.z;
        float k = 1.f / sqrtf(a);
        p = v3(0.f, -n.z * k, n.y * k);
        q = v3(a * k, -n.x * p.z, n.x * p.y);
    } else {
        float a = n.x * n.x + n.y * n.y;
        float k = 1.f / sqrtf(a);
        p = v3(-n.y * k, n.x * k, 0.f);
        q = v3(-n.z * p.y, n.z * p.x, a * k);
    }
}

// Rigid transform integration with the exponential map (angle per step limited to pi/4), as the reference's
// btTransformUtil::integrateTransform (LinearMath/btTransformUtil.h:37-90).
GGL_HD void integrate_transform(V3 pos, const M3& basis, V3 linvel, V3 angvel, float dt, V3& out_pos, M3& out_basis) {
    out_pos = pos + linvel * dt;
    float a2 = len2(angvel);
    float fa = 0.f;
    if (a2 > kEps) fa = sqrtf(a2);
    const float thresh = 0.5f * kHalfPi;
    if (fa * dt > thresh) fa = thresh / dt;
    V3 axis;
    if (fa < 0.001f) {
        axis = angvel * (0.5f * dt - (dt * dt * dt) * 0.020833333333f * fa * fa);
    } else {
        axis = angvel * (sinf(0.5f * fa * dt) / fa);
    }
    Q4 dorn; dorn.x = axis.x; dorn.y = axis.y; dorn.z = axis.z; dorn.w = cosf(fa * dt * 0.5f);
    Q4 orn0 = quat_from_m3(basis);
    Q4 pred = qmul(dorn, orn0);
    float l2 = qlen2(pred);
    if (l2 > kEps) {
        float inv = 1.f / sqrtf(l2);
        pred.x *= inv; pred.y *= inv; pred.z *= inv; pred.w *= inv;
    }
    if (qlen2(pred) > kEps) out_basis = m3_from_quat(pred);
    else out_basis = basis;
}

}  // namespace ggl
