// Surfaces on the device: bilinear-patch intersection with the reference's arithmetic, a stackless
// (threaded, skip-link) BVH walk, surface interaction construction and ray spawning.
#pragma once
#include "dmedia.cuh"
#include "dsphere.cuh"

namespace vspg {

struct Hit { float t, u, v; int prim; };   // prim < 0: miss

struct SurfaceIntr {
    V3 p, pError, n, ns, dpdu, wo;
    float u, v;
    int material, areaLight;
    bool hasInterface;
    int mediumInside, mediumOutside, rayMedium;
    VD int get_medium(V3 w) const {  // interaction.h:117-121
        if (hasInterface) return dotn(w, n) > 0 ? mediumOutside : mediumInside;
        return rayMedium;
    }
};

// ray.h:75-92
VD V3 offset_ray_origin(V3 p, V3 pError, V3 n, V3 w) {
    float d = dotn(vabs(n), pError);
    V3 offset = d * n;
    if (dotn(w, n) < 0) offset = -offset;
    V3 po = p + offset;
    if (offset.x > 0) po.x = next_up(po.x); else if (offset.x < 0) po.x = next_down(po.x);
    if (offset.y > 0) po.y = next_up(po.y); else if (offset.y < 0) po.y = next_down(po.y);
    if (offset.z > 0) po.z = next_up(po.z); else if (offset.z < 0) po.z = next_down(po.z);
    return po;
}

// util/math.h:613-636
VD bool quadratic(float a, float b, float c, float *t0, float *t1) {
    if (a == 0) {
        if (b == 0) return false;
        *t0 = *t1 = -c / b;
        return true;
    }
    float discrim = diff_of_products(b, b, 4 * a, c);
    if (discrim < 0) return false;
    float rootDiscrim = sqrtf(discrim);
    float q = -0.5f * (b + copysignf(rootDiscrim, b));
    *t0 = q / a; *t1 = c / q;
    if (*t0 > *t1) { float t = *t0; *t0 = *t1; *t1 = t; }
    return true;
}
// util/math.h SquareMatrix<3> Determinant
VD float det3(float m00, float m01, float m02, float m10, float m11, float m12, float m20, float m21, float m22) {
    float minor12 = diff_of_products(m11, m22, m12, m21);
    float minor02 = diff_of_products(m10, m22, m12, m20);
    float minor01 = diff_of_products(m10, m21, m11, m20);
    return fmaf(m02, minor01, diff_of_products(m00, minor12, m01, minor02));
}
// shapes.h:1279-1345 IntersectBilinearPatch.  The candidate (t,u,v) does not depend on tMax except for the final
// accept test, so the caller does the `t < tMax` comparison (this lets the BVH order break ties by primitive id).
VD bool intersect_bilinear_patch(V3 ro, V3 rd, V3 p00, V3 p10, V3 p01, V3 p11, float *uo, float *vo, float *to) {
    float a = dot(cross(p10 - p00, p01 - p11), rd);
    float c = dot(cross(p00 - ro, rd), p01 - p00);
    float b = dot(cross(p10 - ro, rd), p11 - p10) - (a + c);
    float u1, u2;
    if (!quadratic(a, b, c, &u1, &u2)) return false;
    float eps = gammaN(30) * (max_comp(vabs(ro)) + max_comp(vabs(rd)) + max_comp(vabs(p00)) + max_comp(vabs(p10)) +
                              max_comp(vabs(p01)) + max_comp(vabs(p11)));
    float t = inff(), u = 0, v = 0;
    bool found = false;
    if (0 <= u1 && u1 <= 1) {
        V3 uo_ = lerpv(u1, p00, p10);
        V3 ud = lerpv(u1, p01, p11) - uo_;
        V3 deltao = uo_ - ro;
        V3 perp = cross(rd, ud);
        float p2 = length_sq(perp);
        float v1 = det3(deltao.x, rd.x, perp.x, deltao.y, rd.y, perp.y, deltao.z, rd.z, perp.z);
        float t1 = det3(deltao.x, ud.x, perp.x, deltao.y, ud.y, perp.y, deltao.z, ud.z, perp.z);
        if (t1 > p2 * eps && 0 <= v1 && v1 <= p2) { u = u1; v = v1 / p2; t = t1 / p2; found = true; }
    }
    if (0 <= u2 && u2 <= 1 && u2 != u1) {
        V3 uo_ = lerpv(u2, p00, p10);
        V3 ud = lerpv(u2, p01, p11) - uo_;
        V3 deltao = uo_ - ro;
        V3 perp = cross(rd, ud);
        float p2 = length_sq(perp);
        float v2 = det3(deltao.x, rd.x, perp.x, deltao.y, rd.y, perp.y, deltao.z, rd.z, perp.z);
        float t2 = det3(deltao.x, ud.x, perp.x, deltao.y, ud.y, perp.y, deltao.z, ud.z, perp.z);
        t2 /= p2;
        if (0 <= v2 && v2 <= p2 && t > t2 && t2 > eps) { t = t2; u = u2; v = v2 / p2; found = true; }
    }
    if (!found) return false;
    *uo = u; *vo = v; *to = t;
    return true;
}


// shapes.cpp:168-263 IntersectTriangle (watertight: permute, shear, edge functions with the double-precision
// fallback at edges, conservative t > deltaT).  Returns barycentrics and t; the caller compares t with its closest hit.
VD bool intersect_triangle(V3 ro, V3 rd, float tMax, V3 p0, V3 p1, V3 p2, float *b0o, float *b1o, float *b2o, float *to) {
    if (length_sq(cross(p2 - p0, p1 - p0)) == 0) return false;
    V3 p0t = p0 - ro, p1t = p1 - ro, p2t = p2 - ro;
    V3 ad = vabs(rd);
    int kz = (ad.x > ad.y) ? ((ad.x > ad.z) ? 0 : 2) : ((ad.y > ad.z) ? 1 : 2);   // util/vecmath.h MaxComponentIndex
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    V3 d(rd[kx], rd[ky], rd[kz]);
    p0t = V3(p0t[kx], p0t[ky], p0t[kz]); p1t = V3(p1t[kx], p1t[ky], p1t[kz]); p2t = V3(p2t[kx], p2t[ky], p2t[kz]);
    float Sx = -d.x / d.z, Sy = -d.y / d.z, Sz = 1 / d.z;
    p0t.x += Sx * p0t.z; p0t.y += Sy * p0t.z;
    p1t.x += Sx * p1t.z; p1t.y += Sy * p1t.z;
    p2t.x += Sx * p2t.z; p2t.y += Sy * p2t.z;
    float e0 = diff_of_products(p1t.x, p2t.y, p1t.y, p2t.x);
    float e1 = diff_of_products(p2t.x, p0t.y, p2t.y, p0t.x);
    float e2 = diff_of_products(p0t.x, p1t.y, p0t.y, p1t.x);
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double p2txp1ty = (double)p2t.x * (double)p1t.y, p2typ1tx = (double)p2t.y * (double)p1t.x;
        e0 = (float)(p2typ1tx - p2txp1ty);
        double p0txp2ty = (double)p0t.x * (double)p2t.y, p0typ2tx = (double)p0t.y * (double)p2t.x;
        e1 = (float)(p0typ2tx - p0txp2ty);
        double p1txp0ty = (double)p1t.x * (double)p0t.y, p1typ0tx = (double)p1t.y * (double)p0t.x;
        e2 = (float)(p1typ0tx - p1txp0ty);
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = e0 + e1 + e2;
    if (det == 0) return false;
    p0t.z *= Sz; p1t.z *= Sz; p2t.z *= Sz;
    float tScaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0 && (tScaled >= 0 || tScaled < tMax * det)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > tMax * det)) return false;
    float invDet = 1 / det;
    float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    float t = tScaled * invDet;
    float maxZt = max_comp(vabs(V3(p0t.z, p1t.z, p2t.z)));
    float deltaZ = gammaN(3) * maxZt;
    float maxXt = max_comp(vabs(V3(p0t.x, p1t.x, p2t.x)));
    float maxYt = max_comp(vabs(V3(p0t.y, p1t.y, p2t.y)));
    float deltaX = gammaN(5) * (maxXt + maxZt);
    float deltaY = gammaN(5) * (maxYt + maxZt);
    float deltaE = 2 * (gammaN(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = max_comp(vabs(V3(e0, e1, e2)));
    float deltaT = 3 * (gammaN(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * fabsf(invDet);
    if (t <= deltaT) return false;
    *b0o = b0; *b1o = b1; *b2o = b2; *to = t;
    return true;
}

// Closest hit: stackless walk over the DFS-ordered threaded BVH.  Box test is the conservative slab test of
// util/vecmath.h:1574 (tFar widened by 1+2*gamma(3)) so no hit of the brute-force reference order is lost;
// ties in t go to the lower primitive index (= first primitive in scene order, as a linear scan would keep).
// KINDS: primitive kinds the scene holds (bit 0 quads, bit 1 triangles, bit 2 spheres; host: vspg_ctx::prim_kinds).  The
// BASELINE configs are quad-only; compiling the triangle / sphere branches out of their kernels took a third off the
// tracking kernels' code, whose top stall was instruction fetch (profiles/r02n).
constexpr int kKindsAll = 7, kKindsQuads = 1;
template <int KINDS = kKindsAll>
VD Hit scene_intersect(const DScene &sc, V3 ro, V3 rd, float tMax, unsigned &nodesVisited) {
    Hit h; h.t = tMax; h.prim = -1; h.u = h.v = 0;
    const float rayTMax = tMax;
    V3 inv(1 / rd.x, 1 / rd.y, 1 / rd.z);
    int node = 0;
    const int n = sc.n_bvh;
    while (node < n) {
        const float4 *np = reinterpret_cast<const float4 *>(&sc.bvh[node]);
        float4 a = __ldg(np), b = __ldg(np + 1);
        int4 c = __ldg(reinterpret_cast<const int4 *>(np + 2));
        nodesVisited++;
        float t0 = 0, t1 = h.t;
        bool hitBox = true;
        {
            float tn = (a.x - ro.x) * inv.x, tf = (b.x - ro.x) * inv.x;
            if (tn > tf) { float t = tn; tn = tf; tf = t; }
            tf *= 1 + 2 * gammaN(3);
            t0 = tn > t0 ? tn : t0; t1 = tf < t1 ? tf : t1;
            if (t0 > t1) hitBox = false;
            tn = (a.y - ro.y) * inv.y; tf = (b.y - ro.y) * inv.y;
            if (tn > tf) { float t = tn; tn = tf; tf = t; }
            tf *= 1 + 2 * gammaN(3);
            t0 = tn > t0 ? tn : t0; t1 = tf < t1 ? tf : t1;
            if (t0 > t1) hitBox = false;
            tn = (a.z - ro.z) * inv.z; tf = (b.z - ro.z) * inv.z;
            if (tn > tf) { float t = tn; tn = tf; tf = t; }
            tf *= 1 + 2 * gammaN(3);
            t0 = tn > t0 ? tn : t0; t1 = tf < t1 ? tf : t1;
            if (t0 > t1) hitBox = false;
        }
        int skip = __float_as_int(a.w), primBegin = __float_as_int(b.w), primCount = c.x;
        if (!hitBox) { node = skip; continue; }
        if (primBegin >= 0) {
            for (int i = 0; i < primCount; i++) {
                int q = __ldg(&sc.prim_ids[primBegin + i]);
                const DQuad &Q = sc.quads[q];
                float u, v, t, b2;
                if ((KINDS & 4) && Q.kind == PRIM_SPHERE) {   // tested against the ray's own tMax like triangles (both roots are compared with it)
                    V3 pObj;
                    if (intersect_sphere(ro, rd, rayTMax, ld3(Q.p00), Q.p10[0], &t, &pObj) && (t < h.t || (t == h.t && h.prim >= 0 && q < h.prim))) {
                        h.t = t; h.u = 0; h.v = 0; h.prim = q;
                    }
                    continue;
                }
                bool got = (KINDS & 2) && Q.kind == PRIM_TRIANGLE ? intersect_triangle(ro, rd, rayTMax, ld3(Q.p00), ld3(Q.p10), ld3(Q.p01), &u, &v, &b2, &t)
                                                   : intersect_bilinear_patch(ro, rd, ld3(Q.p00), ld3(Q.p10), ld3(Q.p01), ld3(Q.p11), &u, &v, &t);
                if (got && (t < h.t || (t == h.t && h.prim >= 0 && q < h.prim))) { h.t = t; h.u = u; h.v = v; h.prim = q; }
            }
            node = skip;
        } else node = node + 1;
    }
    return h;
}

// shapes.h:1398-1500 BilinearPatch::InteractionFromIntersection (no mesh uv / normals) + interaction.h:218-229
// Triangles: shapes.h:884-945 Triangle::InteractionFromIntersection with the default (0,0),(1,0),(1,1) uvs and no
// mesh normals; the barycentrics are recomputed from the ray (the hit record carries t, b0, b1 only).
template <int KINDS = kKindsAll>
VD SurfaceIntr make_intr(const DScene &sc, const Hit &h, V3 rayO, V3 rayD, int rayMedium) {
    SurfaceIntr s;
    const DQuad &q = sc.quads[h.prim];
    V3 p00 = ld3(q.p00), p10 = ld3(q.p10), p01 = ld3(q.p01), p11 = ld3(q.p11);
    if ((KINDS & 4) && q.kind == PRIM_SPHERE) {   // the hit record carries t only: the quadric hit point is recomputed from the ray
        float t; V3 pObj;
        intersect_sphere(rayO, rayD, inff(), p00, p10.x, &t, &pObj);
        sphere_interaction(pObj, p00, p10.x, q.flip_normal != 0, &s.p, &s.pError, &s.n, &s.dpdu);
        s.ns = s.n;
        s.wo = normalize(-rayD);
        s.u = s.v = 0;   // (phi, theta) parametrisation only feeds textures, which this path does not have
        s.material = q.material; s.areaLight = q.area_light;
        s.hasInterface = q.has_interface && (q.medium_inside != q.medium_outside);
        s.mediumInside = q.medium_inside; s.mediumOutside = q.medium_outside; s.rayMedium = rayMedium;
        return s;
    }
    if ((KINDS & 2) && q.kind == PRIM_TRIANGLE) {
        V3 p0 = p00, p1 = p10, p2 = p01;
        float b0 = h.u, b1 = h.v, b2 = 0, t;
        intersect_triangle(rayO, rayD, inff(), p0, p1, p2, &b0, &b1, &b2, &t);
        V3 dp02 = p0 - p2, dp12 = p1 - p2;
        // duv02 = (-1,-1), duv12 = (0,-1), determinant 1: DifferenceOfProducts collapses to one rounding per component
        V3 dpdu(fmaf(-1.f, dp02.x, dp12.x), fmaf(-1.f, dp02.y, dp12.y), fmaf(-1.f, dp02.z, dp12.z));
        V3 dpdv = -dp12;
        if (length_sq(cross(dpdu, dpdv)) == 0) {
            V3 ng = cross(p2 - p0, p1 - p0);
            coordinate_system(normalize(ng), &dpdu, &dpdv);
        }
        point_interval(b0 * p0 + b1 * p1 + b2 * p2, gammaN(7) * (vabs(b0 * p0) + vabs(b1 * p1) + vabs(b2 * p2)), &s.p, &s.pError);   // Point3fi(pHit, pError), shapes.h:929-935
        s.n = normalize(cross(dp02, dp12));
        if (q.flip_normal) s.n = -s.n;
        s.ns = s.n; s.dpdu = dpdu;
        s.wo = normalize(-rayD);
        s.u = b0 * 0.f + b1 * 1.f + b2 * 1.f; s.v = b0 * 0.f + b1 * 0.f + b2 * 1.f;
        s.material = q.material; s.areaLight = q.area_light;
        s.hasInterface = q.has_interface && (q.medium_inside != q.medium_outside);
        s.mediumInside = q.medium_inside; s.mediumOutside = q.medium_outside; s.rayMedium = rayMedium;
        return s;
    }
    V3 pLerp = lerpv(h.u, lerpv(h.v, p00, p01), lerpv(h.v, p10, p11));
    V3 dpdu = lerpv(h.v, p10, p11) - lerpv(h.v, p00, p01);
    V3 dpdv = lerpv(h.u, p01, p11) - lerpv(h.u, p00, p10);
    V3 pAbsSum = vabs(p00) + vabs(p01) + vabs(p10) + vabs(p11);
    point_interval(pLerp, gammaN(6) * pAbsSum, &s.p, &s.pError);   // Point3fi(p, pError), shapes.h:1464-1472
    s.n = normalize(cross(dpdu, dpdv));
    if (q.flip_normal) s.n = -s.n;
    s.ns = s.n; s.dpdu = dpdu;
    s.wo = normalize(-rayD);
    s.u = h.u; s.v = h.v;
    s.material = q.material; s.areaLight = q.area_light;
    s.hasInterface = q.has_interface && (q.medium_inside != q.medium_outside);
    s.mediumInside = q.medium_inside; s.mediumOutside = q.medium_outside; s.rayMedium = rayMedium;
    return s;
}

// lights.h:492-515 DiffuseAreaLight::L
VD V3 area_L(const DLight &l, V3 n, V3 w) {
    if (!l.two_sided && dotn(n, w) < 0) return V3();
    return ld3(l.L);
}

}  // namespace vspg
