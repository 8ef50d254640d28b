// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  (see orc_math.h header)
// CPU restatement of the reference's sphere: running-error interval arithmetic (util/math.h:818-1070, the CPU
// variants of the directed-rounding helpers util/float.h:199-300), Transform applied to Point3fi / Vector3fi
// (util/transform.h:133-176, 272-306), Sphere::BasicIntersect (shapes.h:147-233) and
// Sphere::InteractionFromIntersection (shapes.h:241-292) followed by Transform::operator()(SurfaceInteraction)
// (util/transform.cpp:229-261).  Full spheres only (zMin=-r, zMax=r, phiMax=360), object space = translate(center):
// the matrices below are written out in full so the arithmetic is the reference's general 4x4 form.
#pragma once
#include <algorithm>

#include "orc_math.h"

namespace orc {

inline float AddRoundUp(float a, float b) { return NextFloatUp(a + b); }
inline float AddRoundDown(float a, float b) { return NextFloatDown(a + b); }
inline float SubRoundUp(float a, float b) { return AddRoundUp(a, -b); }
inline float SubRoundDown(float a, float b) { return AddRoundDown(a, -b); }
inline float MulRoundUp(float a, float b) { return NextFloatUp(a * b); }
inline float MulRoundDown(float a, float b) { return NextFloatDown(a * b); }
inline float DivRoundUp(float a, float b) { return NextFloatUp(a / b); }
inline float DivRoundDown(float a, float b) { return NextFloatDown(a / b); }
inline float SqrtRoundUp(float a) { return NextFloatUp(std::sqrt(a)); }
inline float SqrtRoundDown(float a) { return std::max<float>(0, NextFloatDown(std::sqrt(a))); }

struct Interval {
    float low = 0, high = 0;
    Interval() = default;
    explicit Interval(float v) : low(v), high(v) {}
    Interval(float l, float h) : low(std::min(l, h)), high(std::max(l, h)) {}
    static Interval FromValueAndError(float v, float err) {
        Interval i;
        if (err == 0) i.low = i.high = v;
        else { i.low = SubRoundDown(v, err); i.high = AddRoundUp(v, err); }
        return i;
    }
    float Midpoint() const { return (low + high) / 2; }
    float Width() const { return high - low; }
    bool operator==(const Interval &o) const { return low == o.low && high == o.high; }
};
inline bool InRange(float v, Interval i) { return v >= i.low && v <= i.high; }
inline Interval operator+(Interval a, Interval b) { return {AddRoundDown(a.low, b.low), AddRoundUp(a.high, b.high)}; }
inline Interval operator-(Interval a, Interval b) { return {SubRoundDown(a.low, b.high), SubRoundUp(a.high, b.low)}; }
inline Interval operator*(Interval a, Interval b) {
    float lp[4] = {MulRoundDown(a.low, b.low), MulRoundDown(a.high, b.low), MulRoundDown(a.low, b.high), MulRoundDown(a.high, b.high)};
    float hp[4] = {MulRoundUp(a.low, b.low), MulRoundUp(a.high, b.low), MulRoundUp(a.low, b.high), MulRoundUp(a.high, b.high)};
    return {std::min({lp[0], lp[1], lp[2], lp[3]}), std::max({hp[0], hp[1], hp[2], hp[3]})};
}
inline Interval operator/(Interval a, Interval b) {
    if (InRange(0, b)) return Interval(-Infinity, Infinity);
    float lq[4] = {DivRoundDown(a.low, b.low), DivRoundDown(a.high, b.low), DivRoundDown(a.low, b.high), DivRoundDown(a.high, b.high)};
    float hq[4] = {DivRoundUp(a.low, b.low), DivRoundUp(a.high, b.low), DivRoundUp(a.low, b.high), DivRoundUp(a.high, b.high)};
    return {std::min({lq[0], lq[1], lq[2], lq[3]}), std::max({hq[0], hq[1], hq[2], hq[3]})};
}
inline Interval operator*(float f, Interval i) {
    if (f > 0) return Interval(MulRoundDown(f, i.low), MulRoundUp(f, i.high));
    return Interval(MulRoundDown(f, i.high), MulRoundUp(f, i.low));
}
inline Interval Sqr(Interval i) {
    float alow = std::abs(i.low), ahigh = std::abs(i.high);
    if (alow > ahigh) std::swap(alow, ahigh);
    if (InRange(0, i)) return Interval(0, MulRoundUp(ahigh, ahigh));
    return Interval(MulRoundDown(alow, alow), MulRoundUp(ahigh, ahigh));
}
inline Interval Sqrt(Interval i) { return {SqrtRoundDown(i.low), SqrtRoundUp(i.high)}; }

struct V3i {  // Point3fi / Vector3fi
    Interval x, y, z;
    V3 Mid() const { return V3(x.Midpoint(), y.Midpoint(), z.Midpoint()); }
    V3 Error() const { return V3(x.Width() / 2, y.Width() / 2, z.Width() / 2); }
    bool IsExact() const { return x.Width() == 0 && y.Width() == 0 && z.Width() == 0; }
    static V3i FromValueAndError(V3 v, V3 e) {
        return {Interval::FromValueAndError(v.x, e.x), Interval::FromValueAndError(v.y, e.y), Interval::FromValueAndError(v.z, e.z)};
    }
    static V3i Exact(V3 v) { return {Interval(v.x), Interval(v.y), Interval(v.z)}; }
};

struct Mat34 { float m[3][4]; };   // rows 0..2 of an affine Transform (row 3 = 0 0 0 1 => wp == 1)
inline Mat34 TranslateM(V3 d) { return Mat34{{{1, 0, 0, d.x}, {0, 1, 0, d.y}, {0, 0, 1, d.z}}}; }

// util/transform.h:133-176 Transform::operator()(Point3fi)
inline V3i XfPoint(const Mat34 &M, const V3i &p) {
    const float(*m)[4] = M.m;
    float x = p.x.Midpoint(), y = p.y.Midpoint(), z = p.z.Midpoint();
    V3 out, err;
    V3 inErr = p.Error();
    bool exact = p.IsExact();
    for (int r = 0; r < 3; ++r) {
        out[r] = (m[r][0] * x + m[r][1] * y) + (m[r][2] * z + m[r][3]);
        float g = gamma_(3) * (std::abs(m[r][0] * x) + std::abs(m[r][1] * y) + std::abs(m[r][2] * z) + std::abs(m[r][3]));
        if (exact) err[r] = g;
        else err[r] = (gamma_(3) + 1) * (std::abs(m[r][0]) * inErr.x + std::abs(m[r][1]) * inErr.y + std::abs(m[r][2]) * inErr.z) + g;
    }
    return V3i::FromValueAndError(out, err);
}
// util/transform.h:272-306 Transform::operator()(Vector3fi), exact input
inline V3i XfVectorExact(const Mat34 &M, V3 v) {
    const float(*m)[4] = M.m;
    V3 out, err;
    for (int r = 0; r < 3; ++r) {
        err[r] = gamma_(3) * (std::abs(m[r][0] * v.x) + std::abs(m[r][1] * v.y) + std::abs(m[r][2] * v.z));
        out[r] = m[r][0] * v.x + m[r][1] * v.y + m[r][2] * v.z;
    }
    return V3i::FromValueAndError(out, err);
}

// shapes.h:147-233 Sphere::BasicIntersect for a full sphere.  Returns tHit and the refined object-space hit point.
inline bool IntersectSphere(V3 ro, V3 rd, float tMax, V3 center, float radius, float *tHit, V3 *pObj) {
    Mat34 objectFromRender = TranslateM(-center);
    V3i oi = XfPoint(objectFromRender, V3i::Exact(ro));
    V3i di = XfVectorExact(objectFromRender, rd);
    Interval a = Sqr(di.x) + Sqr(di.y) + Sqr(di.z);
    Interval b = 2 * (di.x * oi.x + di.y * oi.y + di.z * oi.z);
    Interval c = Sqr(oi.x) + Sqr(oi.y) + Sqr(oi.z) - Sqr(Interval(radius));
    Interval s = b / (2 * a);
    V3i v{oi.x - s * di.x, oi.y - s * di.y, oi.z - s * di.z};
    Interval length = Sqrt(Sqr(v.x) + Sqr(v.y) + Sqr(v.z));
    Interval discrim = 4 * a * (Interval(radius) + length) * (Interval(radius) - length);
    if (discrim.low < 0) return false;
    Interval rootDiscrim = Sqrt(discrim);
    Interval q;
    if (b.Midpoint() < 0) q = -.5f * (b - rootDiscrim);
    else q = -.5f * (b + rootDiscrim);
    Interval t0 = q / a, t1 = c / q;
    if (t0.low > t1.low) std::swap(t0, t1);
    if (t0.high > tMax || t1.low <= 0) return false;
    Interval tShapeHit = t0;
    if (tShapeHit.low <= 0) {
        tShapeHit = t1;
        if (tShapeHit.high > tMax) return false;
    }
    V3 pHit = oi.Mid() + tShapeHit.Midpoint() * di.Mid();
    pHit = pHit * (radius / Length(pHit - V3(0, 0, 0)));
    if (pHit.x == 0 && pHit.y == 0) pHit.x = 1e-5f * radius;
    // phi = atan2(y, x) (+2 Pi) <= phiMax = Radians(360) always holds for a full sphere: no clipping re-test
    *tHit = tShapeHit.Midpoint();
    *pObj = pHit;
    return true;
}

// shapes.h:241-292 + transform.cpp:229-261: render-space pi (as midpoint + error), n, dpdu of a full-sphere hit.
inline void SphereInteraction(V3 pHit, V3 center, float radius, bool flipNormal, V3 *p, V3 *pError, V3 *n, V3 *dpduOut) {
    const float phiMax = 6.28318548f;               // Radians(Clamp(360, 0, 360))
    const float thetaZMin = 3.14159274f, thetaZMax = 0;   // acos(-1), acos(1)
    float cosTheta = pHit.z / radius;
    float zRadius = std::sqrt(Sqr(pHit.x) + Sqr(pHit.y));
    float cosPhi = pHit.x / zRadius, sinPhi = pHit.y / zRadius;
    V3 dpdu(-phiMax * pHit.y, phiMax * pHit.x, 0);
    float sinTheta = SafeSqrt(1 - Sqr(cosTheta));
    V3 dpdv = (thetaZMax - thetaZMin) * V3(pHit.z * cosPhi, pHit.z * sinPhi, -radius * sinTheta);
    V3 nObj = Normalize(Cross(dpdu, dpdv));
    if (flipNormal) nObj = -nObj;
    V3 pErrObj = gamma_(5) * Abs(pHit);
    Mat34 renderFromObject = TranslateM(center);
    V3i pi = XfPoint(renderFromObject, V3i::FromValueAndError(pHit, pErrObj));
    *p = pi.Mid(); *pError = pi.Error();
    // Normal3f through mInv^T = identity, then Normalize; Vector3f dpdu through m = identity rotation
    *n = Normalize(V3(nObj.x * 1.f + nObj.y * 0.f + nObj.z * 0.f, nObj.x * 0.f + nObj.y * 1.f + nObj.z * 0.f, nObj.x * 0.f + nObj.y * 0.f + nObj.z * 1.f));
    *dpduOut = dpdu;
}

}  // namespace orc
