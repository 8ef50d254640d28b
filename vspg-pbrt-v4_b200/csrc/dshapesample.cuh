// Solid-angle sampling of rectangular bilinear patches as BilinearPatch::Sample(ctx, u) / PDF(ctx, wi) do it
// (shapes.cpp:1254-1368): SphericalQuadArea (util/vecmath.h:1646-1664), SampleSphericalRectangle
// (util/sampling.cpp:163-220) and its inverse (util/sampling.cpp:222-345, needed for the cos-theta warp's pdf).
// The routines are out of line: they run once per next-event sample and must not add to the register budget of the
// lane-group scatter kernels that call them.
#pragma once
#include "dmath.cuh"

namespace vspg {

constexpr float kMinSphericalSampleAreaQuad = 1e-4f;   // shapes.h:1538

VD float safe_asin(float x) { return vspg_asinf(fminf(1.f, fmaxf(-1.f, x))); }
VD float angle_between(V3 v1, V3 v2) {   // util/vecmath.h:972-977
    if (dot(v1, v2) < 0) return kPi - 2 * safe_asin(length(v1 + v2) / 2);
    return 2 * safe_asin(length(v2 - v1) / 2);
}

__device__ __noinline__ float spherical_quad_area(V3 a, V3 b, V3 c, V3 d) {
    V3 axb = cross(a, b), bxc = cross(b, c), cxd = cross(c, d), dxa = cross(d, a);
    if (length_sq(axb) == 0 || length_sq(bxc) == 0 || length_sq(cxd) == 0 || length_sq(dxa) == 0) return 0;
    axb = normalize(axb); bxc = normalize(bxc); cxd = normalize(cxd); dxa = normalize(dxa);
    float alpha = angle_between(dxa, -axb), beta = angle_between(axb, -bxc);
    float gamma = angle_between(bxc, -cxd), delta = angle_between(cxd, -dxa);
    return fabsf(alpha + beta + gamma + delta - 2 * kPi);
}

// The local frame and the four internal angles both directions of the mapping start from.
struct SphRect {
    V3 rx, ry, rz;
    float x0, y0, z0, x1, y1;
    float g0, g1, g2, g3, b0, b1;
};
VD SphRect sph_rect_init(V3 pRef, V3 s, V3 ex, V3 ey) {
    SphRect q;
    float exl = length(ex), eyl = length(ey);
    q.rx = ex / exl; q.ry = ey / eyl; q.rz = cross(q.rx, q.ry);
    V3 d = s - pRef;
    V3 dLocal(dot(d, q.rx), dot(d, q.ry), dot(d, q.rz));
    q.z0 = dLocal.z;
    if (q.z0 > 0) { q.rz = -q.rz; q.z0 *= -1; }
    q.x0 = dLocal.x; q.y0 = dLocal.y;
    q.x1 = q.x0 + exl; q.y1 = q.y0 + eyl;
    V3 v00(q.x0, q.y0, q.z0), v01(q.x0, q.y1, q.z0), v10(q.x1, q.y0, q.z0), v11(q.x1, q.y1, q.z0);
    V3 n0 = normalize(cross(v00, v10)), n1 = normalize(cross(v10, v11));
    V3 n2 = normalize(cross(v11, v01)), n3 = normalize(cross(v01, v00));
    q.g0 = angle_between(-n0, n1); q.g1 = angle_between(-n1, n2);
    q.g2 = angle_between(-n2, n3); q.g3 = angle_between(-n3, n0);
    q.b0 = n0.z; q.b1 = n2.z;
    return q;
}

__device__ __noinline__ V3 sample_spherical_rectangle(V3 pRef, V3 s, V3 ex, V3 ey, float u0, float u1, float *pdf) {
    SphRect q = sph_rect_init(pRef, s, ex, ey);
    float solidAngle = q.g0 + q.g1 + q.g2 + q.g3 - 2 * kPi;
    if (solidAngle <= 0) { *pdf = 0; return s + u0 * ex + u1 * ey; }
    *pdf = fmaxf(0.f, 1 / solidAngle);
    if (solidAngle < 0.001f) return s + u0 * ex + u1 * ey;   // `< 1e-3` in double: the same set of floats
    float au = u0 * (q.g0 + q.g1 - 2 * kPi) + (u0 - 1) * (q.g2 + q.g3);
    float fu = (vspg_cosf(au) * q.b0 - q.b1) / vspg_sinf(au);
    float cu = copysignf(1 / sqrtf(sqr(fu) + sqr(q.b0)), fu);
    cu = clampf(cu, -kOneMinusEpsilon, kOneMinusEpsilon);
    float xu = -(cu * q.z0) / safe_sqrt(1 - sqr(cu));
    xu = clampf(xu, q.x0, q.x1);
    float dd = sqrtf(sqr(xu) + sqr(q.z0));
    float h0 = q.y0 / sqrtf(sqr(dd) + sqr(q.y0));
    float h1 = q.y1 / sqrtf(sqr(dd) + sqr(q.y1));
    float hv = h0 + u1 * (h1 - h0), hvsq = sqr(hv);
    float yv = (hvsq < 1 - 1e-6f) ? (hv * dd) / sqrtf(1 - hvsq) : q.y1;
    return pRef + (xu * q.rx + yv * q.ry + q.z0 * q.rz);
}

__device__ __noinline__ void invert_spherical_rectangle_sample(V3 pRef, V3 s, V3 ex, V3 ey, V3 pRect, float *uo0, float *uo1) {
    SphRect q = sph_rect_init(pRef, s, ex, ey);
    float z0sq = sqr(q.z0), y0sq = sqr(q.y0), y1sq = sqr(q.y1), b0sq = sqr(q.b0);
    float solidAngle = (float)((double)q.g0 + (double)q.g1 + (double)q.g2 + (double)q.g3 - 2. * kPi);
    if (solidAngle < 0.001f) {
        V3 pq = pRect - s;
        *uo0 = dot(pq, ex) / length_sq(ex); *uo1 = dot(pq, ey) / length_sq(ey);
        return;
    }
    V3 w = pRect - pRef;
    float xu = dot(w, q.rx), yv = dot(w, q.ry);
    xu = clampf(xu, q.x0, q.x1);
    if (xu == 0) xu = 1e-10f;
    float invcusq = 1 + z0sq / sqr(xu);
    float fusq = invcusq - b0sq;
    float fu = copysignf(sqrtf(fusq), xu);
    float root = safe_sqrt(diff_of_products(q.b0, q.b0, q.b1, q.b1) + fusq);
    float au = vspg_atan2f(-(q.b1 * fu) - copysignf(q.b0 * root, fu * q.b0), q.b0 * q.b1 - root * fabsf(fu));
    if (au > 0) au -= 2 * kPi;
    if (fu == 0) au = kPi;
    float u0 = (au + q.g2 + q.g3) / solidAngle;
    float ddsq = sqr(xu) + z0sq;
    float dd = sqrtf(ddsq);
    float h0 = q.y0 / sqrtf(ddsq + y0sq);
    float h1 = q.y1 / sqrtf(ddsq + y1sq);
    float yvsq = sqr(yv);
    float term = fabsf(h0 - h1) * sqrtf(yvsq * (ddsq + yvsq)) / (ddsq + yvsq);
    float dop = diff_of_products(h0, h0, h0, h1);
    float u1a = (dop - term) / sqr(h0 - h1), u1b = (dop + term) / sqr(h0 - h1);
    float hva = lerpf(u1a, h0, h1), hvb = lerpf(u1b, h0, h1);
    float yza = (hva * dd) / sqrtf(1 - sqr(hva)), yzb = (hvb * dd) / sqrtf(1 - sqr(hvb));
    *uo0 = clampf(u0, 0.f, 1.f);
    *uo1 = fabsf(yza - yv) < fabsf(yzb - yv) ? u1a : u1b;
}

// ---------------------------------------------------------------------------------------------------- triangles
constexpr float kMinSphericalSampleAreaTri = 3e-4f, kMaxSphericalSampleAreaTri = 6.22f;   // shapes.h:1190-1191

VD V3 gram_schmidt(V3 v, V3 w) { return v - dot(v, w) * w; }
VD float spherical_triangle_area(V3 a, V3 b, V3 c) {              // util/vecmath.h:1638-1641
    return fabsf(2 * vspg_atan2f(dot(a, cross(b, c)), 1 + dot(a, b) + dot(a, c) + dot(b, c)));
}

// util/sampling.cpp:28-107; false = the reference's empty return
__device__ __noinline__ bool sample_spherical_triangle(V3 v0, V3 v1, V3 v2, V3 p, float u0, float u1, float *bary, float *pdf) {
    *pdf = 0;
    V3 a = normalize(v0 - p), b = normalize(v1 - p), c = normalize(v2 - p);
    V3 n_ab = cross(a, b), n_bc = cross(b, c), n_ca = cross(c, a);
    if (length_sq(n_ab) == 0 || length_sq(n_bc) == 0 || length_sq(n_ca) == 0) return false;
    n_ab = normalize(n_ab); n_bc = normalize(n_bc); n_ca = normalize(n_ca);
    float alpha = angle_between(n_ab, -n_ca), beta = angle_between(n_bc, -n_ab), gamma = angle_between(n_ca, -n_bc);
    float A_pi = alpha + beta + gamma;
    float Ap_pi = lerpf(u0, kPi, A_pi);
    float A = A_pi - kPi;
    *pdf = (A <= 0) ? 0 : 1 / A;
    float cosAlpha = vspg_cosf(alpha), sinAlpha = vspg_sinf(alpha);
    float sinPhi = vspg_sinf(Ap_pi) * cosAlpha - vspg_cosf(Ap_pi) * sinAlpha;
    float cosPhi = vspg_cosf(Ap_pi) * cosAlpha + vspg_sinf(Ap_pi) * sinAlpha;
    float k1 = cosPhi + cosAlpha;
    float k2 = sinPhi - sinAlpha * dot(a, b);
    float cosBp = (k2 + (diff_of_products(k2, cosPhi, k1, sinPhi)) * cosAlpha) / ((sum_of_products(k2, sinPhi, k1, cosPhi)) * sinAlpha);
    cosBp = clampf(cosBp, -1.f, 1.f);
    float sinBp = safe_sqrt(1 - sqr(cosBp));
    V3 cp = cosBp * a + sinBp * normalize(gram_schmidt(c, a));
    float cosTheta = 1 - u1 * (1 - dot(cp, b));
    float sinTheta = safe_sqrt(1 - sqr(cosTheta));
    V3 w = cosTheta * b + sinTheta * normalize(gram_schmidt(cp, b));
    V3 e1 = v1 - v0, e2 = v2 - v0;
    V3 s1 = cross(w, e2);
    float divisor = dot(s1, e1);
    if (divisor == 0) { bary[0] = bary[1] = bary[2] = 1.f / 3.f; return true; }
    float invDivisor = 1 / divisor;
    V3 s = p - v0;
    float b1 = dot(s, s1) * invDivisor;
    float b2 = dot(w, cross(s, e1)) * invDivisor;
    b1 = clampf(b1, 0.f, 1.f);
    b2 = clampf(b2, 0.f, 1.f);
    if (b1 + b2 > 1) {   // as the reference writes it: b2 is divided by the sum with the already rescaled b1
        b1 /= b1 + b2;
        b2 /= b1 + b2;
    }
    bary[0] = 1 - b1 - b2; bary[1] = b1; bary[2] = b2;
    return true;
}

// util/sampling.cpp:110-161
__device__ __noinline__ void invert_spherical_triangle_sample(V3 v0, V3 v1, V3 v2, V3 p, V3 w, float *uo0, float *uo1) {
    *uo0 = *uo1 = 0;
    V3 a = normalize(v0 - p), b = normalize(v1 - p), c = normalize(v2 - p);
    V3 n_ab = cross(a, b), n_bc = cross(b, c), n_ca = cross(c, a);
    if (length_sq(n_ab) == 0 || length_sq(n_bc) == 0 || length_sq(n_ca) == 0) return;
    n_ab = normalize(n_ab); n_bc = normalize(n_bc); n_ca = normalize(n_ca);
    float alpha = angle_between(n_ab, -n_ca), beta = angle_between(n_bc, -n_ab), gamma = angle_between(n_ca, -n_bc);
    V3 cp = normalize(cross(cross(b, w), cross(c, a)));
    if (dot(cp, a + c) < 0) cp = -cp;
    float u0;
    if (dot(a, cp) > 0.99999847691f) u0 = 0;
    else {
        V3 n_cpb = cross(cp, b), n_acp = cross(a, cp);
        if (length_sq(n_cpb) == 0 || length_sq(n_acp) == 0) { *uo0 = *uo1 = 0.5f; return; }
        n_cpb = normalize(n_cpb); n_acp = normalize(n_acp);
        float Ap = alpha + angle_between(n_ab, n_cpb) + angle_between(n_acp, -n_cpb) - kPi;
        float A = alpha + beta + gamma - kPi;
        u0 = Ap / A;
    }
    float u1 = (1 - dot(w, b)) / (1 - dot(cp, b));
    *uo0 = clampf(u0, 0.f, 1.f); *uo1 = clampf(u1, 0.f, 1.f);
}

}  // namespace vspg
