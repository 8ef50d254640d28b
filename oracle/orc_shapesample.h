// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  (see orc_math.h header)
// CPU restatement of the solid-angle sampling helpers BilinearPatch::Sample(ctx) / PDF(ctx, wi) use for rectangular
// patches (shapes.cpp:1254-1368): SphericalQuadArea (util/vecmath.h:1646-1664), SampleSphericalRectangle
// (util/sampling.cpp:163-220), InvertSphericalRectangleSample (util/sampling.cpp:222-345).
#pragma once
#include "orc_lightbvh.h"   // AngleBetween

namespace orc {

constexpr float MinSphericalSampleAreaQuad = 1e-4f;   // shapes.h:1538 (BilinearPatch)

inline float SphericalQuadArea(V3 a, V3 b, V3 c, V3 d) {
    V3 axb = Cross(a, b), bxc = Cross(b, c), cxd = Cross(c, d), dxa = Cross(d, a);
    if (LengthSquared(axb) == 0 || LengthSquared(bxc) == 0 || LengthSquared(cxd) == 0 || LengthSquared(dxa) == 0) return 0;
    axb = Normalize(axb); bxc = Normalize(bxc); cxd = Normalize(cxd); dxa = Normalize(dxa);
    float alpha = AngleBetween(dxa, -axb), beta = AngleBetween(axb, -bxc);
    float gamma = AngleBetween(bxc, -cxd), delta = AngleBetween(cxd, -dxa);
    return std::abs(alpha + beta + gamma + delta - 2 * Pi);
}

struct RectFrame { V3 x, y, z; };   // Frame::FromXY (vecmath.h:1864), ToLocal / FromLocal (:1905-1915)
inline V3 ToLocal(const RectFrame &R, V3 v) { return V3(Dot(v, R.x), Dot(v, R.y), Dot(v, R.z)); }
inline V3 FromLocal(const RectFrame &R, V3 v) { return v.x * R.x + v.y * R.y + v.z * R.z; }

inline V3 SampleSphericalRectangle(V3 pRef, V3 s, V3 ex, V3 ey, const float *u, float *pdf) {
    float exl = Length(ex), eyl = Length(ey);
    RectFrame R; R.x = ex / exl; R.y = ey / eyl; R.z = Cross(R.x, R.y);
    V3 dLocal = ToLocal(R, s - pRef);
    float z0 = dLocal.z;
    if (z0 > 0) { R.z = -R.z; z0 *= -1; }
    float x0 = dLocal.x, y0 = dLocal.y;
    float x1 = x0 + exl, y1 = y0 + eyl;
    V3 v00(x0, y0, z0), v01(x0, y1, z0), v10(x1, y0, z0), v11(x1, y1, z0);
    V3 n0 = Normalize(Cross(v00, v10)), n1 = Normalize(Cross(v10, v11));
    V3 n2 = Normalize(Cross(v11, v01)), n3 = Normalize(Cross(v01, v00));
    float g0 = AngleBetween(-n0, n1), g1 = AngleBetween(-n1, n2);
    float g2 = AngleBetween(-n2, n3), g3 = AngleBetween(-n3, n0);
    float solidAngle = g0 + g1 + g2 + g3 - 2 * Pi;
    if (solidAngle <= 0) { *pdf = 0; return s + u[0] * ex + u[1] * ey; }
    *pdf = std::max<float>(0, 1 / solidAngle);
    if (solidAngle < 1e-3) return s + u[0] * ex + u[1] * ey;
    float b0 = n0.z, b1 = n2.z;
    float au = u[0] * (g0 + g1 - 2 * Pi) + (u[0] - 1) * (g2 + g3);
    float fu = (vspg_cosf(au) * b0 - b1) / vspg_sinf(au);
    float cu = std::copysign(1 / std::sqrt(Sqr(fu) + Sqr(b0)), fu);
    cu = Clamp(cu, -OneMinusEpsilon, OneMinusEpsilon);
    float xu = -(cu * z0) / SafeSqrt(1 - Sqr(cu));
    xu = Clamp(xu, x0, x1);
    float dd = std::sqrt(Sqr(xu) + Sqr(z0));
    float h0 = y0 / std::sqrt(Sqr(dd) + Sqr(y0));
    float h1 = y1 / std::sqrt(Sqr(dd) + Sqr(y1));
    float hv = h0 + u[1] * (h1 - h0), hvsq = Sqr(hv);
    float yv = (hvsq < 1 - 1e-6f) ? (hv * dd) / std::sqrt(1 - hvsq) : y1;
    return pRef + FromLocal(R, V3(xu, yv, z0));
}

inline void InvertSphericalRectangleSample(V3 pRef, V3 s, V3 ex, V3 ey, V3 pRect, float *uOut) {
    float exl = Length(ex), eyl = Length(ey);
    RectFrame R; R.x = ex / exl; R.y = ey / eyl; R.z = Cross(R.x, R.y);
    V3 d = s - pRef;
    V3 dLocal = ToLocal(R, d);
    float z0 = dLocal.z;
    if (z0 > 0) { R.z = -R.z; z0 *= -1; }
    float z0sq = Sqr(z0);
    float x0 = dLocal.x, y0 = dLocal.y;
    float x1 = x0 + exl, y1 = y0 + eyl;
    float y0sq = Sqr(y0), y1sq = Sqr(y1);
    V3 v00(x0, y0, z0), v01(x0, y1, z0), v10(x1, y0, z0), v11(x1, y1, z0);
    V3 n0 = Normalize(Cross(v00, v10)), n1 = Normalize(Cross(v10, v11));
    V3 n2 = Normalize(Cross(v11, v01)), n3 = Normalize(Cross(v01, v00));
    float g0 = AngleBetween(-n0, n1), g1 = AngleBetween(-n1, n2);
    float g2 = AngleBetween(-n2, n3), g3 = AngleBetween(-n3, n0);
    float b0 = n0.z, b1 = n2.z, b0sq = Sqr(b0);
    float solidAngle = (float)(double(g0) + double(g1) + double(g2) + double(g3) - 2. * Pi);
    if (solidAngle < 1e-3) {
        V3 pq = pRect - s;
        uOut[0] = Dot(pq, ex) / LengthSquared(ex); uOut[1] = Dot(pq, ey) / LengthSquared(ey);
        return;
    }
    V3 v = ToLocal(R, pRect - pRef);
    float xu = v.x, yv = v.y;
    xu = Clamp(xu, x0, x1);
    if (xu == 0) xu = 1e-10;
    float invcusq = 1 + z0sq / Sqr(xu);
    float fusq = invcusq - b0sq;
    float fu = std::copysign(std::sqrt(fusq), xu);
    float sqrt_ = SafeSqrt(DifferenceOfProducts(b0, b0, b1, b1) + fusq);
    float au = vspg_atan2f(-(b1 * fu) - std::copysign(b0 * sqrt_, fu * b0), b0 * b1 - sqrt_ * std::abs(fu));
    if (au > 0) au -= 2 * Pi;
    if (fu == 0) au = Pi;
    float u0 = (au + g2 + g3) / solidAngle;
    float ddsq = Sqr(xu) + z0sq;
    float dd = std::sqrt(ddsq);
    float h0 = y0 / std::sqrt(ddsq + y0sq);
    float h1 = y1 / std::sqrt(ddsq + y1sq);
    float yvsq = Sqr(yv);
    float root = std::abs(h0 - h1) * std::sqrt(yvsq * (ddsq + yvsq)) / (ddsq + yvsq);
    float u1[2] = {(DifferenceOfProducts(h0, h0, h0, h1) - root) / Sqr(h0 - h1), (DifferenceOfProducts(h0, h0, h0, h1) + root) / Sqr(h0 - h1)};
    float hv[2] = {Lerp(u1[0], h0, h1), Lerp(u1[1], h0, h1)};
    float hvsq[2] = {Sqr(hv[0]), Sqr(hv[1])};
    float yz[2] = {(hv[0] * dd) / std::sqrt(1 - hvsq[0]), (hv[1] * dd) / std::sqrt(1 - hvsq[1])};
    uOut[0] = Clamp(u0, 0, 1);
    uOut[1] = (std::abs(yz[0] - yv) < std::abs(yz[1] - yv)) ? u1[0] : u1[1];
}

// ---------------------------------------------------------------------------------------------------- triangles
constexpr float MinSphericalSampleAreaTri = 3e-4f, MaxSphericalSampleAreaTri = 6.22f;   // shapes.h:1190-1191

inline V3 GramSchmidt(V3 v, V3 w) { return v - Dot(v, w) * w; }   // vecmath.h:994
inline float SphericalTriangleArea(V3 a, V3 b, V3 c) {            // vecmath.h:1638-1641
    return std::abs(2 * vspg_atan2f(Dot(a, Cross(b, c)), 1 + Dot(a, b) + Dot(a, c) + Dot(b, c)));
}
inline void SampleUniformTriangle(const float *u, float *b) {     // sampling.h:173-183
    float b0, b1;
    if (u[0] < u[1]) { b0 = u[0] / 2; b1 = u[1] - b0; }
    else { b1 = u[1] / 2; b0 = u[0] - b1; }
    b[0] = b0; b[1] = b1; b[2] = 1 - b0 - b1;
}
// util/sampling.cpp:28-107.  Returns false for the reference's empty return (all-zero barycentrics, pdf 0).
inline bool SampleSphericalTriangle(const V3 *v, V3 p, const float *u, float *bary, float *pdf) {
    *pdf = 0;
    V3 a = Normalize(v[0] - p), b = Normalize(v[1] - p), c = Normalize(v[2] - p);
    V3 n_ab = Cross(a, b), n_bc = Cross(b, c), n_ca = Cross(c, a);
    if (LengthSquared(n_ab) == 0 || LengthSquared(n_bc) == 0 || LengthSquared(n_ca) == 0) return false;
    n_ab = Normalize(n_ab); n_bc = Normalize(n_bc); n_ca = Normalize(n_ca);
    float alpha = AngleBetween(n_ab, -n_ca), beta = AngleBetween(n_bc, -n_ab), gamma = AngleBetween(n_ca, -n_bc);
    float A_pi = alpha + beta + gamma;
    float Ap_pi = Lerp(u[0], Pi, A_pi);
    float A = A_pi - Pi;
    *pdf = (A <= 0) ? 0 : 1 / A;
    float cosAlpha = vspg_cosf(alpha), sinAlpha = vspg_sinf(alpha);
    float sinPhi = vspg_sinf(Ap_pi) * cosAlpha - vspg_cosf(Ap_pi) * sinAlpha;
    float cosPhi = vspg_cosf(Ap_pi) * cosAlpha + vspg_sinf(Ap_pi) * sinAlpha;
    float k1 = cosPhi + cosAlpha;
    float k2 = sinPhi - sinAlpha * Dot(a, b);
    float cosBp = (k2 + (DifferenceOfProducts(k2, cosPhi, k1, sinPhi)) * cosAlpha) / ((SumOfProducts(k2, sinPhi, k1, cosPhi)) * sinAlpha);
    cosBp = Clamp(cosBp, -1, 1);
    float sinBp = SafeSqrt(1 - Sqr(cosBp));
    V3 cp = cosBp * a + sinBp * Normalize(GramSchmidt(c, a));
    float cosTheta = 1 - u[1] * (1 - Dot(cp, b));
    float sinTheta = SafeSqrt(1 - Sqr(cosTheta));
    V3 w = cosTheta * b + sinTheta * Normalize(GramSchmidt(cp, b));
    V3 e1 = v[1] - v[0], e2 = v[2] - v[0];
    V3 s1 = Cross(w, e2);
    float divisor = Dot(s1, e1);
    if (divisor == 0) { bary[0] = bary[1] = bary[2] = 1.f / 3.f; return true; }
    float invDivisor = 1 / divisor;
    V3 s = p - v[0];
    float b1 = Dot(s, s1) * invDivisor;
    float b2 = Dot(w, Cross(s, e1)) * invDivisor;
    b1 = Clamp(b1, 0, 1);
    b2 = Clamp(b2, 0, 1);
    if (b1 + b2 > 1) {   // as written in the reference: b2 is divided by the sum with the already rescaled b1
        b1 /= b1 + b2;
        b2 /= b1 + b2;
    }
    bary[0] = 1 - b1 - b2; bary[1] = b1; bary[2] = b2;
    return true;
}
// util/sampling.cpp:110-161
inline void InvertSphericalTriangleSample(const V3 *v, V3 p, V3 w, float *uo) {
    uo[0] = uo[1] = 0;
    V3 a = Normalize(v[0] - p), b = Normalize(v[1] - p), c = Normalize(v[2] - p);
    V3 n_ab = Cross(a, b), n_bc = Cross(b, c), n_ca = Cross(c, a);
    if (LengthSquared(n_ab) == 0 || LengthSquared(n_bc) == 0 || LengthSquared(n_ca) == 0) return;
    n_ab = Normalize(n_ab); n_bc = Normalize(n_bc); n_ca = Normalize(n_ca);
    float alpha = AngleBetween(n_ab, -n_ca), beta = AngleBetween(n_bc, -n_ab), gamma = AngleBetween(n_ca, -n_bc);
    V3 cp = Normalize(Cross(Cross(b, w), Cross(c, a)));
    if (Dot(cp, a + c) < 0) cp = -cp;
    float u0;
    if (Dot(a, cp) > 0.99999847691f) u0 = 0;
    else {
        V3 n_cpb = Cross(cp, b), n_acp = Cross(a, cp);
        if (LengthSquared(n_cpb) == 0 || LengthSquared(n_acp) == 0) { uo[0] = uo[1] = 0.5f; return; }
        n_cpb = Normalize(n_cpb); n_acp = Normalize(n_acp);
        float Ap = alpha + AngleBetween(n_ab, n_cpb) + AngleBetween(n_acp, -n_cpb) - Pi;
        float A = alpha + beta + gamma - Pi;
        u0 = Ap / A;
    }
    float u1 = (1 - Dot(w, b)) / (1 - Dot(cp, b));
    uo[0] = Clamp(u0, 0, 1); uo[1] = Clamp(u1, 0, 1);
}

}  // namespace orc
