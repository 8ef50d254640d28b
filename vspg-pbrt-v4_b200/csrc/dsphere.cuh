// Full spheres on the device: Sphere::BasicIntersect (shapes.h:147-233) with the reference's running-error interval
// arithmetic (util/math.h:818-1070; the CPU flavour of the directed roundings, float.h:199-300: round-to-nearest then
// one ulp outwards, which is what the reference's CPU integrator computes) and Sphere::InteractionFromIntersection
// (shapes.h:241-292) pushed through renderFromObject (transform.cpp:229-261).  Object space is translate(center), so
// the 4x4 products of Transform::operator()(Point3fi / Vector3fi) (transform.h:133-176, 272-306) reduce to one add
// per coordinate: the `0 * y` terms they drop are exact zeros for finite inputs.
#pragma once
#include "dmath.cuh"

namespace vspg {

struct Ival {
    float lo, hi;
    VD Ival() : lo(0), hi(0) {}
    VD explicit Ival(float v) : lo(v), hi(v) {}
    VD Ival(float a, float b) : lo(fminf(a, b)), hi(fmaxf(a, b)) {}
    VD float mid() const { return (lo + hi) / 2; }
    VD float half_width() const { return (hi - lo) / 2; }
};
VD Ival ival_from_error(float v, float err) {   // Interval::FromValueAndError
    if (err == 0) return Ival(v);
    Ival i; i.lo = next_down(v - err); i.hi = next_up(v + err);
    return i;
}
VD bool ival_has_zero(Ival i) { return 0.f >= i.lo && 0.f <= i.hi; }
VD Ival operator+(Ival a, Ival b) { return Ival(next_down(a.lo + b.lo), next_up(a.hi + b.hi)); }
VD Ival operator-(Ival a, Ival b) { return Ival(next_down(a.lo - b.hi), next_up(a.hi - b.lo)); }
VD Ival operator*(Ival a, Ival b) {
    float p0 = a.lo * b.lo, p1 = a.hi * b.lo, p2 = a.lo * b.hi, p3 = a.hi * b.hi;
    // NextFloatDown / NextFloatUp are monotone, so rounding the extreme product equals the extreme of the rounded ones
    return Ival(next_down(fminf(fminf(p0, p1), fminf(p2, p3))), next_up(fmaxf(fmaxf(p0, p1), fmaxf(p2, p3))));
}
VD Ival operator/(Ival a, Ival b) {
    if (ival_has_zero(b)) return Ival(-inff(), inff());
    float q0 = a.lo / b.lo, q1 = a.hi / b.lo, q2 = a.lo / b.hi, q3 = a.hi / b.hi;
    return Ival(next_down(fminf(fminf(q0, q1), fminf(q2, q3))), next_up(fmaxf(fmaxf(q0, q1), fmaxf(q2, q3))));
}
VD Ival ival_scale(float f, Ival i) {   // operator*(Float, Interval)
    if (f > 0) return Ival(next_down(f * i.lo), next_up(f * i.hi));
    return Ival(next_down(f * i.hi), next_up(f * i.lo));
}
VD Ival ival_sqr(Ival i) {
    float alow = fabsf(i.lo), ahigh = fabsf(i.hi);
    if (alow > ahigh) { float t = alow; alow = ahigh; ahigh = t; }
    if (ival_has_zero(i)) return Ival(0.f, next_up(ahigh * ahigh));
    return Ival(next_down(alow * alow), next_up(ahigh * ahigh));
}
VD Ival ival_sqrt(Ival i) { return Ival(fmaxf(0.f, next_down(sqrtf(i.lo))), next_up(sqrtf(i.hi))); }

// shapes.h:147-233 for zMin = -r, zMax = r, phiMax = 360: phi <= phiMax always holds (Radians(360) rounds to the
// same float as 2*Pi), so the clipping re-test never fires and atan2 is not needed.
__device__ __noinline__ bool intersect_sphere(V3 ro, V3 rd, float tMax, V3 center, float radius, float *tHit, V3 *pObj) {
    const float g3 = gammaN(3);
    Ival ox = ival_from_error(ro.x + -center.x, g3 * (fabsf(ro.x) + fabsf(center.x)));
    Ival oy = ival_from_error(ro.y + -center.y, g3 * (fabsf(ro.y) + fabsf(center.y)));
    Ival oz = ival_from_error(ro.z + -center.z, g3 * (fabsf(ro.z) + fabsf(center.z)));
    Ival dx = ival_from_error(rd.x, g3 * fabsf(rd.x)), dy = ival_from_error(rd.y, g3 * fabsf(rd.y)), dz = ival_from_error(rd.z, g3 * fabsf(rd.z));
    Ival a = ival_sqr(dx) + ival_sqr(dy) + ival_sqr(dz);
    Ival b = ival_scale(2.f, dx * ox + dy * oy + dz * oz);
    Ival R(radius);
    Ival c = ival_sqr(ox) + ival_sqr(oy) + ival_sqr(oz) - ival_sqr(R);
    Ival s = b / ival_scale(2.f, a);
    Ival vx = ox - s * dx, vy = oy - s * dy, vz = oz - s * dz;
    Ival len = ival_sqrt(ival_sqr(vx) + ival_sqr(vy) + ival_sqr(vz));
    Ival discrim = ival_scale(4.f, a) * (R + len) * (R - len);
    if (discrim.lo < 0) return false;
    Ival root = ival_sqrt(discrim);
    Ival q = b.mid() < 0 ? ival_scale(-.5f, b - root) : ival_scale(-.5f, b + root);
    Ival t0 = q / a, t1 = c / q;
    if (t0.lo > t1.lo) { Ival t = t0; t0 = t1; t1 = t; }
    if (t0.hi > tMax || t1.lo <= 0) return false;
    Ival tShape = t0;
    if (tShape.lo <= 0) {
        tShape = t1;
        if (tShape.hi > tMax) return false;
    }
    V3 pHit = V3(ox.mid(), oy.mid(), oz.mid()) + tShape.mid() * V3(dx.mid(), dy.mid(), dz.mid());
    pHit = pHit * (radius / length(pHit));
    if (pHit.x == 0 && pHit.y == 0) pHit.x = 1e-5f * radius;
    *tHit = tShape.mid();
    *pObj = pHit;
    return true;
}

// shapes.h:241-292 + transform.cpp:229-261 (+ SurfaceInteraction ctor, interaction.h): render-space point (midpoint
// and half-width of the transformed Point3fi), geometric normal, dpdu.
__device__ __noinline__ void sphere_interaction(V3 pHit, V3 center, float radius, bool flipNormal, V3 *p, V3 *pError, V3 *n, V3 *dpduOut) {
    const float phiMax = 6.28318548f;      // Radians(360)
    const float dTheta = 0.f - 3.14159274f;   // thetaZMax - thetaZMin = acos(1) - acos(-1)
    float cosTheta = pHit.z / radius;
    float zRadius = sqrtf(sqr(pHit.x) + sqr(pHit.y));
    float cosPhi = pHit.x / zRadius, sinPhi = pHit.y / zRadius;
    V3 dpdu(-phiMax * pHit.y, phiMax * pHit.x, 0);
    float sinTheta = safe_sqrt(1 - sqr(cosTheta));
    V3 dpdv = dTheta * V3(pHit.z * cosPhi, pHit.z * sinPhi, -radius * sinTheta);
    V3 nObj = normalize(cross(dpdu, dpdv));
    if (flipNormal) nObj = -nObj;
    const float g3 = gammaN(3), g5 = gammaN(5);
    float c[3] = {center.x, center.y, center.z}, ph[3] = {pHit.x, pHit.y, pHit.z}, po[3], pe[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        Ival in = ival_from_error(ph[k], g5 * fabsf(ph[k]));
        float x = in.mid(), e = in.half_width();
        // a zero half-width on all three axes (IsExact) cannot happen: |pHit| = radius > 0
        float err = (g3 + 1) * e + g3 * (fabsf(x) + fabsf(c[k]));
        Ival out = ival_from_error(x + c[k], err);
        po[k] = out.mid(); pe[k] = out.half_width();
    }
    *p = V3(po[0], po[1], po[2]); *pError = V3(pe[0], pe[1], pe[2]);
    *n = normalize(nObj);
    *dpduOut = dpdu;
}

}  // namespace vspg
