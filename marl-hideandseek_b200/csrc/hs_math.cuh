// Vector / quaternion / AABB arithmetic for the hide-and-seek kernels.
//
// Every expression keeps the reference's operand order and rounding points
// (M/include/madrona/math.inl) because level generation must be bit-exact and
// the physics is compared against the reference CPU backend: the library is
// compiled with -fmad=false, IEEE sqrt/div, and no reassociation, so a
// faithfully ordered expression produces the same float as g++ -ffp-contract=off.
#pragma once

#include <cfloat>
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define HS_HD __host__ __device__ __forceinline__
#else
#define HS_HD inline
#endif

namespace hs {

struct Vec3 { float x, y, z; };
struct Quat { float w, x, y, z; };
struct AABB { Vec3 lo, hi; };
struct Plane { Vec3 n; float d; };

HS_HD Vec3 v3(float x, float y, float z) { return Vec3{x, y, z}; }
HS_HD Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
HS_HD Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
HS_HD Vec3 operator-(Vec3 a) { return {-a.x, -a.y, -a.z}; }
HS_HD Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
HS_HD Vec3 operator*(float s, Vec3 a) { return {a.x * s, a.y * s, a.z * s}; }
// math.inl:421 — division multiplies by the rounded reciprocal.
HS_HD Vec3 operator/(Vec3 a, float s) { float inv = 1.f / s; return a * inv; }
HS_HD Vec3 &operator+=(Vec3 &a, Vec3 b) { a = a + b; return a; }
HS_HD Vec3 &operator-=(Vec3 &a, Vec3 b) { a = a - b; return a; }

HS_HD float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HS_HD Vec3 cross(Vec3 a, Vec3 b)
{
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
HS_HD float length2(Vec3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
HS_HD float length(Vec3 a) { return sqrtf(length2(a)); }
// CPU path of Vector3::invLength (math.inl:279-287): 1/sqrt, not rsqrt.
HS_HD Vec3 normalize(Vec3 a) { return a * (1.f / length(a)); }
HS_HD Vec3 mulDiag(Vec3 d, Vec3 v) { return {d.x * v.x, d.y * v.y, d.z * v.z}; }
HS_HD Vec3 vmin(Vec3 a, Vec3 b) { return {fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)}; }
HS_HD Vec3 vmax(Vec3 a, Vec3 b) { return {fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)}; }
HS_HD float comp(Vec3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

HS_HD Quat operator+(Quat a, Quat b) { return {a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z}; }
HS_HD Quat operator-(Quat a, Quat b) { return {a.w - b.w, a.x - b.x, a.y - b.y, a.z - b.z}; }
HS_HD Quat operator*(Quat a, Quat b)
{
    return {
        (a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z),
        (a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y),
        (a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x),
        (a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w),
    };
}
HS_HD Quat qinv(Quat q) { return {q.w, -q.x, -q.y, -q.z}; }
HS_HD Quat qnormalize(Quat q)
{
    float inv = 1.f / sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    return {q.w * inv, q.x * inv, q.y * inv, q.z * inv};
}
HS_HD Quat fromAngularVec(Vec3 v) { return {0.f, v.x, v.y, v.z}; }
// Quat::rotateVec (math.inl:717-726)
HS_HD Vec3 rotate(Quat q, Vec3 v)
{
    Vec3 pure = {q.x, q.y, q.z};
    Vec3 pxv = cross(pure, v);
    Vec3 pxpxv = cross(pure, pxv);
    return v + 2.f * ((pxv * q.w) + pxpxv);
}

struct Mat3 { Vec3 c0, c1, c2; };  // column major, as the reference

HS_HD Vec3 operator*(const Mat3 &m, Vec3 v) { return m.c0 * v.x + m.c1 * v.y + m.c2 * v.z; }

// Mat3x3::fromQuat (math.inl:1070-1099)
HS_HD Mat3 mat3FromQuat(Quat r)
{
    float x2 = r.x * r.x, y2 = r.y * r.y, z2 = r.z * r.z;
    float xz = r.x * r.z, xy = r.x * r.y, yz = r.y * r.z;
    float wx = r.w * r.x, wy = r.w * r.y, wz = r.w * r.z;
    return {
        {1.f - 2.f * (y2 + z2), 2.f * (xy + wz), 2.f * (xz - wy)},
        {2.f * (xy - wz), 1.f - 2.f * (x2 + z2), 2.f * (yz + wx)},
        {2.f * (xz + wy), 2.f * (yz - wx), 1.f - 2.f * (x2 + y2)},
    };
}

// Mat3x3::fromRS (math.inl:1101-1132)
HS_HD Mat3 mat3FromRS(Quat r, Vec3 s)
{
    float x2 = r.x * r.x, y2 = r.y * r.y, z2 = r.z * r.z;
    float xz = r.x * r.z, xy = r.x * r.y, yz = r.y * r.z;
    float wx = r.w * r.x, wy = r.w * r.y, wz = r.w * r.z;
    Vec3 ds = {2.f * s.x, 2.f * s.y, 2.f * s.z};
    return {
        {s.x - ds.x * (y2 + z2), ds.x * (xy + wz), ds.x * (xz - wy)},
        {ds.y * (xy - wz), s.y - ds.y * (x2 + z2), ds.y * (yz + wx)},
        {ds.z * (xz + wy), ds.z * (yz - wx), s.z - ds.z * (x2 + y2)},
    };
}

// AABB::applyTRS (math.inl:1737-1769), RTCD p.86
HS_HD AABB aabbApplyTRS(AABB b, Vec3 t, Quat r, Vec3 s)
{
    Mat3 m = mat3FromRS(r, s);
    float lo[3] = {b.lo.x, b.lo.y, b.lo.z};
    float hi[3] = {b.hi.x, b.hi.y, b.hi.z};
    float tr[3] = {t.x, t.y, t.z};
    // m[j][i]: column j, component i
    float mm[3][3] = {{m.c0.x, m.c0.y, m.c0.z}, {m.c1.x, m.c1.y, m.c1.z}, {m.c2.x, m.c2.y, m.c2.z}};
    float olo[3], ohi[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        olo[i] = ohi[i] = tr[i];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            float e = mm[j][i] * lo[j];
            float f = mm[j][i] * hi[j];
            if (e < f) { olo[i] += e; ohi[i] += f; }
            else { olo[i] += f; ohi[i] += e; }
        }
    }
    return {{olo[0], olo[1], olo[2]}, {ohi[0], ohi[1], ohi[2]}};
}

// AABB::overlaps — strict (math.inl:1601-1609)
HS_HD bool aabbOverlaps(const AABB &a, const AABB &b)
{
    return a.lo.x < b.hi.x && b.lo.x < a.hi.x &&
           a.lo.y < b.hi.y && b.lo.y < a.hi.y &&
           a.lo.z < b.hi.z && b.lo.z < a.hi.z;
}

// AABB::intersects — non-strict SAT (math.inl:1611-1622)
HS_HD bool aabbIntersects(const AABB &a, const AABB &b)
{
    bool xs = a.hi.x < b.lo.x || a.lo.x > b.hi.x;
    bool ys = a.hi.y < b.lo.y || a.lo.y > b.hi.y;
    bool zs = a.hi.z < b.lo.z || a.lo.z > b.hi.z;
    return !xs && !ys && !zs;
}

// AABB::rayIntersects (math.inl:1670-1696): slab test, inv_d may hold inf/NaN.
HS_HD bool aabbRayIntersects(Vec3 lo, Vec3 hi, Vec3 o, Vec3 inv_d, float t_min, float t_max)
{
    Vec3 tl = mulDiag(inv_d, lo - o);
    Vec3 tu = mulDiag(inv_d, hi - o);
    Vec3 mn = vmin(tl, tu);
    Vec3 mx = vmax(tl, tu);
    float t_box_min = fmaxf(mn.x, fmaxf(mn.y, fmaxf(mn.z, t_min)));
    float t_box_max = fminf(mx.x, fminf(mx.y, fminf(mx.z, t_max)));
    return t_box_min <= t_box_max;
}

}  // namespace hs
