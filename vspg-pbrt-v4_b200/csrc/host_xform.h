// Single-precision transform algebra of the scene front end, reproducing the reference's results bit for bit.
//
// The reference keeps every transformation as a pair (m, mInv) of float 4x4 matrices and builds them with
// error-compensated float arithmetic (TwoProd / TwoSum dot products, FMA-based differences of products):
//   matrix product            util/math.h:1475-1483   (4-term compensated inner product per element)
//   4x4 inverse               util/math.h:1571-1623   (Laplace expansion over 2x2 sub-determinants)
//   Translate / Scale / Rotate / LookAt / Perspective  util/transform.cpp:21-131, util/transform.h:220-247
//   point / vector application  util/transform.h:286-330 (plain float, left to right, w divide only when w != 1)
// Camera rays, render-space vertices and medium transforms are only identical to the reference's if these are
// reproduced operation by operation, so this header restates them on plain arrays.  Host code only; compiled with
// -ffp-contract=off so a*b+c stays two roundings unless std::fma is written out.
#pragma once
#include <cmath>
#include "../../include/vspg_libm.h"   // sin / cos / tan of the CTM and the projection: the path's portable libm

namespace xf {

struct CF { float v, err; };   // value + running error (util/math.h:191-203)
inline CF two_prod(float a, float b) { float ab = a * b; return {ab, std::fma(a, b, -ab)}; }
inline CF two_sum(float a, float b) { float s = a + b, d = s - a; return {s, (a - (s - d)) + (b - d)}; }
inline CF inner2(CF ab, CF tail) { CF s = two_sum(ab.v, tail.v); return {s.v, ab.err + (tail.err + s.err)}; }
// InnerProduct(a0,b0, a1,b1, ...): the recursion folds from the LAST pair backwards (util/math.h:596-602)
inline float inner3(float a0, float b0, float a1, float b1, float a2, float b2) {
    CF r = inner2(two_prod(a0, b0), inner2(two_prod(a1, b1), two_prod(a2, b2)));
    return r.v + r.err;
}
inline float inner4(float a0, float b0, float a1, float b1, float a2, float b2, float a3, float b3) {
    CF r = inner2(two_prod(a0, b0), inner2(two_prod(a1, b1), inner2(two_prod(a2, b2), two_prod(a3, b3))));
    return r.v + r.err;
}
inline float inner6(float a0, float b0, float a1, float b1, float a2, float b2, float a3, float b3, float a4, float b4, float a5, float b5) {
    CF r = inner2(two_prod(a0, b0), inner2(two_prod(a1, b1), inner2(two_prod(a2, b2), inner2(two_prod(a3, b3), inner2(two_prod(a4, b4), two_prod(a5, b5))))));
    return r.v + r.err;
}
inline float diff_of_products(float a, float b, float c, float d) {   // util/math.h:568-575
    float cd = c * d;
    float dop = std::fma(a, b, -cd);
    float err = std::fma(-c, d, cd);
    return dop + err;
}

struct M4f {
    float m[4][4];
    static M4f identity() { M4f r; for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) r.m[i][j] = i == j ? 1.f : 0.f; return r; }
};
inline M4f mul(const M4f &a, const M4f &b) {
    M4f r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            r.m[i][j] = inner4(a.m[i][0], b.m[0][j], a.m[i][1], b.m[1][j], a.m[i][2], b.m[2][j], a.m[i][3], b.m[3][j]);
    return r;
}
inline M4f transpose(const M4f &a) { M4f r; for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) r.m[i][j] = a.m[j][i]; return r; }
// false when singular (the reference then fills mInv with NaN or exits, depending on the caller)
inline bool inverse(const M4f &A, M4f *out) {
    const float(*m)[4] = A.m;
    float s0 = diff_of_products(m[0][0], m[1][1], m[1][0], m[0][1]);
    float s1 = diff_of_products(m[0][0], m[1][2], m[1][0], m[0][2]);
    float s2 = diff_of_products(m[0][0], m[1][3], m[1][0], m[0][3]);
    float s3 = diff_of_products(m[0][1], m[1][2], m[1][1], m[0][2]);
    float s4 = diff_of_products(m[0][1], m[1][3], m[1][1], m[0][3]);
    float s5 = diff_of_products(m[0][2], m[1][3], m[1][2], m[0][3]);
    float c0 = diff_of_products(m[2][0], m[3][1], m[3][0], m[2][1]);
    float c1 = diff_of_products(m[2][0], m[3][2], m[3][0], m[2][2]);
    float c2 = diff_of_products(m[2][0], m[3][3], m[3][0], m[2][3]);
    float c3 = diff_of_products(m[2][1], m[3][2], m[3][1], m[2][2]);
    float c4 = diff_of_products(m[2][1], m[3][3], m[3][1], m[2][3]);
    float c5 = diff_of_products(m[2][2], m[3][3], m[3][2], m[2][3]);
    float det = inner6(s0, c5, -s1, c4, s2, c3, s3, c2, s5, c0, -s4, c1);
    if (det == 0) return false;
    float s = 1 / det;
    float (*o)[4] = out->m;
    o[0][0] = s * inner3(m[1][1], c5, m[1][3], c3, -m[1][2], c4);
    o[0][1] = s * inner3(-m[0][1], c5, m[0][2], c4, -m[0][3], c3);
    o[0][2] = s * inner3(m[3][1], s5, m[3][3], s3, -m[3][2], s4);
    o[0][3] = s * inner3(-m[2][1], s5, m[2][2], s4, -m[2][3], s3);
    o[1][0] = s * inner3(-m[1][0], c5, m[1][2], c2, -m[1][3], c1);
    o[1][1] = s * inner3(m[0][0], c5, m[0][3], c1, -m[0][2], c2);
    o[1][2] = s * inner3(-m[3][0], s5, m[3][2], s2, -m[3][3], s1);
    o[1][3] = s * inner3(m[2][0], s5, m[2][3], s1, -m[2][2], s2);
    o[2][0] = s * inner3(m[1][0], c4, m[1][3], c0, -m[1][1], c2);
    o[2][1] = s * inner3(-m[0][0], c4, m[0][1], c2, -m[0][3], c0);
    o[2][2] = s * inner3(m[3][0], s4, m[3][3], s0, -m[3][1], s2);
    o[2][3] = s * inner3(-m[2][0], s4, m[2][1], s2, -m[2][3], s0);
    o[3][0] = s * inner3(-m[1][0], c3, m[1][1], c1, -m[1][2], c0);
    o[3][1] = s * inner3(m[0][0], c3, m[0][2], c0, -m[0][1], c1);
    o[3][2] = s * inner3(-m[3][0], s3, m[3][1], s1, -m[3][2], s0);
    o[3][3] = s * inner3(m[2][0], s3, m[2][2], s0, -m[2][1], s1);
    return true;
}

// A transformation and its inverse carried together (util/transform.h:26-66).
struct XF {
    M4f m = M4f::identity(), inv = M4f::identity();
    XF() = default;
    XF(const M4f &a, const M4f &ai) : m(a), inv(ai) {}
    explicit XF(const M4f &a) : m(a) {   // Transform(const SquareMatrix<4>&): inverse by Inverse(m), NaN when singular
        if (!inverse(a, &inv)) for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) inv.m[i][j] = NAN;
    }
};
inline XF operator*(const XF &a, const XF &b) { return XF(mul(a.m, b.m), mul(b.inv, a.inv)); }   // transform.cpp:141-143
inline XF inverse(const XF &a) { return XF(a.inv, a.m); }
inline XF transpose(const XF &a) { return XF(transpose(a.m), transpose(a.inv)); }

inline XF translate(float x, float y, float z) {
    XF r; r.m.m[0][3] = x; r.m.m[1][3] = y; r.m.m[2][3] = z; r.inv.m[0][3] = -x; r.inv.m[1][3] = -y; r.inv.m[2][3] = -z; return r;
}
inline XF scale(float x, float y, float z) {
    XF r; r.m.m[0][0] = x; r.m.m[1][1] = y; r.m.m[2][2] = z; r.inv.m[0][0] = 1 / x; r.inv.m[1][1] = 1 / y; r.inv.m[2][2] = 1 / z; return r;
}
inline float radians(float deg) { return (3.14159265358979323846f / 180) * deg; }
inline float length3(float x, float y, float z) { return std::sqrt(x * x + y * y + z * z); }   // Length = sqrt(LengthSquared), sums left to right
inline bool rotate(float deg, float ax, float ay, float az, XF *out) {   // util/transform.h:220-247
    float l = length3(ax, ay, az);
    if (l == 0) return false;
    float x = ax / l, y = ay / l, z = az / l;
    float s = vspg_sinf(radians(deg)), c = vspg_cosf(radians(deg));
    M4f r = M4f::identity();
    r.m[0][0] = x * x + (1 - x * x) * c; r.m[0][1] = x * y * (1 - c) - z * s; r.m[0][2] = x * z * (1 - c) + y * s;
    r.m[1][0] = x * y * (1 - c) + z * s; r.m[1][1] = y * y + (1 - y * y) * c; r.m[1][2] = y * z * (1 - c) - x * s;
    r.m[2][0] = x * z * (1 - c) - y * s; r.m[2][1] = y * z * (1 - c) + x * s; r.m[2][2] = z * z + (1 - z * z) * c;
    *out = XF(r, transpose(r));
    return true;
}
inline void cross3(const float a[3], const float b[3], float o[3]) {   // util/vecmath.h Cross: differences of products
    o[0] = diff_of_products(a[1], b[2], a[2], b[1]);
    o[1] = diff_of_products(a[2], b[0], a[0], b[2]);
    o[2] = diff_of_products(a[0], b[1], a[1], b[0]);
}
inline void normalize3(float v[3]) { float l = length3(v[0], v[1], v[2]); v[0] /= l; v[1] /= l; v[2] /= l; }
// LookAt (util/transform.cpp:81-113): returns cameraFromWorld with inverse worldFromCamera; false when up is parallel to the view
inline bool look_at(const float pos[3], const float look[3], const float up_in[3], XF *out) {
    M4f w = M4f::identity();
    w.m[0][3] = pos[0]; w.m[1][3] = pos[1]; w.m[2][3] = pos[2];
    float dir[3] = {look[0] - pos[0], look[1] - pos[1], look[2] - pos[2]};
    normalize3(dir);
    float up[3] = {up_in[0], up_in[1], up_in[2]};
    normalize3(up);
    float right[3]; cross3(up, dir, right);
    if (length3(right[0], right[1], right[2]) == 0) return false;
    normalize3(right);
    float nu[3]; cross3(dir, right, nu);
    for (int i = 0; i < 3; i++) { w.m[i][0] = right[i]; w.m[i][1] = nu[i]; w.m[i][2] = dir[i]; }
    M4f cfw;
    if (!inverse(w, &cfw)) return false;
    *out = XF(cfw, w);
    return true;
}
inline XF perspective(float fov, float n, float f) {   // util/transform.cpp:119-131
    M4f p = M4f::identity();
    p.m[2][2] = f / (f - n); p.m[2][3] = -f * n / (f - n); p.m[3][2] = 1; p.m[3][3] = 0;
    float inv_tan = 1 / vspg_tanf(radians(fov) / 2);
    return scale(inv_tan, inv_tan, 1) * XF(p);
}

// Transform::operator()(Point3f) / (Vector3f) (util/transform.h:286-330)
inline void apply_point(const M4f &t, const float p[3], float o[3]) {
    float x = p[0], y = p[1], z = p[2];
    float xp = t.m[0][0] * x + t.m[0][1] * y + t.m[0][2] * z + t.m[0][3];
    float yp = t.m[1][0] * x + t.m[1][1] * y + t.m[1][2] * z + t.m[1][3];
    float zp = t.m[2][0] * x + t.m[2][1] * y + t.m[2][2] * z + t.m[2][3];
    float wp = t.m[3][0] * x + t.m[3][1] * y + t.m[3][2] * z + t.m[3][3];
    if (wp == 1) { o[0] = xp; o[1] = yp; o[2] = zp; }
    else { o[0] = xp / wp; o[1] = yp / wp; o[2] = zp / wp; }
}
inline void apply_vector(const M4f &t, const float v[3], float o[3]) {
    float x = v[0], y = v[1], z = v[2];
    o[0] = t.m[0][0] * x + t.m[0][1] * y + t.m[0][2] * z;
    o[1] = t.m[1][0] * x + t.m[1][1] * y + t.m[1][2] * z;
    o[2] = t.m[2][0] * x + t.m[2][1] * y + t.m[2][2] * z;
}
inline float det3(const M4f &t) {   // Transform::SwapsHandedness: Determinant(SquareMatrix<3>) (util/math.h:1437-1446)
    float minor12 = diff_of_products(t.m[1][1], t.m[2][2], t.m[1][2], t.m[2][1]);
    float minor02 = diff_of_products(t.m[1][0], t.m[2][2], t.m[1][2], t.m[2][0]);
    float minor01 = diff_of_products(t.m[1][0], t.m[2][1], t.m[1][1], t.m[2][0]);
    return std::fma(t.m[0][2], minor01, diff_of_products(t.m[0][0], minor12, t.m[0][1], minor02));
}

}  // namespace xf
