// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  (see orc_math.h header)
// CPU restatement of GuidedVolPathVSPGIntegrator's render pass: camera sample, Li, SampleDistance,
// SampleLd, guided BSDF / phase sampling, film.  Paths relative to /root/reference/src/pbrt/;
// "vspg.cpp" = cpu/guidedvolpathvspgintegrator.cpp.
#pragma once
#include <atomic>
#include <cmath>
#include <thread>
#include <vector>

#include "orc_guiding.h"
#include "orc_media.h"
#include "orc_lightbvh.h"
#include "orc_sphere.h"
#include "orc_shapesample.h"

namespace orc {

struct Hit {
    bool valid = false;
    float t = Infinity;
    int kind = 0, index = -1;  // 0 quad, 1 triangle, 2 sphere
    float u = 0, v = 0;
    float b[3] = {0, 0, 0};  // triangle barycentrics (TriangleIntersection)
    V3 pObj;                 // sphere: refined object-space hit point (QuadricIntersection)
};

// Surface interaction data the integrator consumes (interaction.h SurfaceInteraction subset)
struct SurfaceIntr {
    V3 p, pError, n, ns, dpdu, wo;
    float uv[2];
    int material = -1, areaLight = -1;
    bool hasInterface = false;
    int mediumInside = -1, mediumOutside = -1;
    int rayMedium = -1;
    int GetMedium(V3 w) const {  // interaction.h:117-121
        if (hasInterface) return DotN(w, n) > 0 ? mediumOutside : mediumInside;
        return rayMedium;
    }
};

// ray.h:75-92
inline V3 OffsetRayOrigin(V3 p, V3 pError, V3 n, V3 w) {
    float d = DotN(Abs(n), pError);
    V3 offset = d * n;
    if (DotN(w, n) < 0) offset = -offset;
    V3 po = p + offset;
    for (int i = 0; i < 3; ++i) {
        if (offset[i] > 0) po[i] = NextFloatUp(po[i]);
        else if (offset[i] < 0) po[i] = NextFloatDown(po[i]);
    }
    return po;
}

// util/math.h:613-636
inline bool Quadratic(float a, float b, float c, float *t0, float *t1) {
    if (a == 0) {
        if (b == 0) return false;
        *t0 = *t1 = -c / b;
        return true;
    }
    float discrim = DifferenceOfProducts(b, b, 4 * a, c);
    if (discrim < 0) return false;
    float rootDiscrim = std::sqrt(discrim);
    float q = -0.5f * (b + std::copysign(rootDiscrim, b));
    *t0 = q / a; *t1 = c / q;
    if (*t0 > *t1) std::swap(*t0, *t1);
    return true;
}
// util/math.h SquareMatrix<3> Determinant: m00*DoP(m11,m22,m12,m21) - m01*DoP(m10,m22,m12,m20) + m02*DoP(m10,m21,m11,m20)
inline float Det3(float m00, float m01, float m02, float m10, float m11, float m12, float m20, float m21, float m22) {
    float minor12 = DifferenceOfProducts(m11, m22, m12, m21);
    float minor02 = DifferenceOfProducts(m10, m22, m12, m20);
    float minor01 = DifferenceOfProducts(m10, m21, m11, m20);
    return FMA(m02, minor01, DifferenceOfProducts(m00, minor12, m01, minor02));
}
// shapes.h:1279-1345 IntersectBilinearPatch
inline bool IntersectBilinearPatch(V3 ro, V3 rd, float tMax, V3 p00, V3 p10, V3 p01, V3 p11, float *uo, float *vo, float *to) {
    float a = Dot(Cross(p10 - p00, p01 - p11), rd);
    float c = Dot(Cross(p00 - ro, rd), p01 - p00);
    float b = Dot(Cross(p10 - ro, rd), p11 - p10) - (a + c);
    float u1, u2;
    if (!Quadratic(a, b, c, &u1, &u2)) return false;
    float eps = gamma_(30) * (MaxComponentValue(Abs(ro)) + MaxComponentValue(Abs(rd)) + MaxComponentValue(Abs(p00)) +
                              MaxComponentValue(Abs(p10)) + MaxComponentValue(Abs(p01)) + MaxComponentValue(Abs(p11)));
    float t = tMax, u = 0, v = 0;
    if (0 <= u1 && u1 <= 1) {
        V3 uo_ = LerpV(u1, p00, p10);
        V3 ud = LerpV(u1, p01, p11) - uo_;
        V3 deltao = uo_ - ro;
        V3 perp = Cross(rd, ud);
        float p2 = LengthSquared(perp);
        float v1 = Det3(deltao.x, rd.x, perp.x, deltao.y, rd.y, perp.y, deltao.z, rd.z, perp.z);
        float t1 = Det3(deltao.x, ud.x, perp.x, deltao.y, ud.y, perp.y, deltao.z, ud.z, perp.z);
        if (t1 > p2 * eps && 0 <= v1 && v1 <= p2) { u = u1; v = v1 / p2; t = t1 / p2; }
    }
    if (0 <= u2 && u2 <= 1 && u2 != u1) {
        V3 uo_ = LerpV(u2, p00, p10);
        V3 ud = LerpV(u2, p01, p11) - uo_;
        V3 deltao = uo_ - ro;
        V3 perp = Cross(rd, ud);
        float p2 = LengthSquared(perp);
        float v2 = Det3(deltao.x, rd.x, perp.x, deltao.y, rd.y, perp.y, deltao.z, rd.z, perp.z);
        float t2 = Det3(deltao.x, ud.x, perp.x, deltao.y, ud.y, perp.y, deltao.z, ud.z, perp.z);
        t2 /= p2;
        if (0 <= v2 && v2 <= p2 && t > t2 && t2 > eps) { t = t2; u = u2; v = v2 / p2; }
    }
    if (t >= tMax) return false;
    *uo = u; *vo = v; *to = t;
    return true;
}

// shapes.cpp:168-263 IntersectTriangle
inline bool IntersectTriangle(V3 ro, V3 rd, float tMax, V3 p0, V3 p1, V3 p2, float *b, float *to) {
    if (LengthSquared(Cross(p2 - p0, p1 - p0)) == 0) return false;
    V3 p0t = p0 - ro, p1t = p1 - ro, p2t = p2 - ro;
    V3 ad = Abs(rd);
    int kz = (ad.x > ad.y) ? ((ad.x > ad.z) ? 0 : 2) : ((ad.y > ad.z) ? 1 : 2);
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    V3 d(rd[kx], rd[ky], rd[kz]);
    p0t = V3(p0t[kx], p0t[ky], p0t[kz]); p1t = V3(p1t[kx], p1t[ky], p1t[kz]); p2t = V3(p2t[kx], p2t[ky], p2t[kz]);
    float Sx = -d.x / d.z, Sy = -d.y / d.z, Sz = 1 / d.z;
    p0t.x += Sx * p0t.z; p0t.y += Sy * p0t.z;
    p1t.x += Sx * p1t.z; p1t.y += Sy * p1t.z;
    p2t.x += Sx * p2t.z; p2t.y += Sy * p2t.z;
    float e0 = DifferenceOfProducts(p1t.x, p2t.y, p1t.y, p2t.x);
    float e1 = DifferenceOfProducts(p2t.x, p0t.y, p2t.y, p0t.x);
    float e2 = DifferenceOfProducts(p0t.x, p1t.y, p0t.y, p1t.x);
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
    float maxZt = MaxComponentValue(Abs(V3(p0t.z, p1t.z, p2t.z)));
    float deltaZ = gamma_(3) * maxZt;
    float maxXt = MaxComponentValue(Abs(V3(p0t.x, p1t.x, p2t.x)));
    float maxYt = MaxComponentValue(Abs(V3(p0t.y, p1t.y, p2t.y)));
    float deltaX = gamma_(5) * (maxXt + maxZt);
    float deltaY = gamma_(5) * (maxYt + maxZt);
    float deltaE = 2 * (gamma_(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = MaxComponentValue(Abs(V3(e0, e1, e2)));
    float deltaT = 3 * (gamma_(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * std::abs(invDet);
    if (t <= deltaT) return false;
    b[0] = b0; b[1] = b1; b[2] = b2; *to = t;
    return true;
}

struct LightSampleCtx { V3 p, pError, n, ns; };

// filters.h:26-45, filters.cpp:133-147 FilterSampler over a GaussianFilter (filters.h:80-118), with
// util/sampling.h:603-790 PiecewiseConstant1D / 2D and util/math.h:508-519 FindInterval
struct Piecewise1D {
    std::vector<float> func, cdf; float mn = 0, mx = 1, funcInt = 0;
    void init(const float *f, int n, float mn_, float mx_) {
        func.assign(f, f + n); cdf.assign(n + 1, 0.f); mn = mn_; mx = mx_;
        for (float &v : func) v = std::abs(v);
        cdf[0] = 0;
        for (int i = 1; i < n + 1; ++i) cdf[i] = cdf[i - 1] + func[i - 1] * (mx - mn) / n;
        funcInt = cdf[n];
        if (funcInt == 0) for (int i = 1; i < n + 1; ++i) cdf[i] = float(i) / float(n);
        else for (int i = 1; i < n + 1; ++i) cdf[i] /= funcInt;
    }
    float Sample(float u, float *pdf, int *offset) const {
        long sz = (long)cdf.size(), size = sz - 2, first = 1;
        while (size > 0) {
            long half = size >> 1, middle = first + half;
            bool pr = cdf[middle] <= u;
            first = pr ? middle + 1 : first;
            size = pr ? size - (half + 1) : half;
        }
        int o = (int)std::min(std::max(first - 1, 0l), sz - 2);
        *offset = o;
        float du = u - cdf[o];
        if (cdf[o + 1] - cdf[o] > 0) du /= cdf[o + 1] - cdf[o];
        *pdf = (funcInt > 0) ? func[o] / funcInt : 0;
        return Lerp((o + du) / (float)func.size(), mn, mx);
    }
};
struct GaussianFilterSampler {
    int nx = 0, ny = 0; std::vector<float> f; std::vector<Piecewise1D> cond; Piecewise1D marginal;
    static float Gaussian(float x, float mu, float sigma) {   // util/math.h:477-480
        return 1 / std::sqrt(2 * Pi * sigma * sigma) * FastExp(-Sqr(x - mu) / (2 * sigma * sigma));
    }
    void init(float rx, float ry, float sigma) {
        nx = int(32 * rx); ny = int(32 * ry);
        float expX = Gaussian(rx, 0, sigma), expY = Gaussian(ry, 0, sigma);
        f.assign((size_t)nx * ny, 0.f);
        for (int y = 0; y < ny; ++y)
            for (int x = 0; x < nx; ++x) {
                float px = Lerp((x + 0.5f) / nx, -rx, rx), py = Lerp((y + 0.5f) / ny, -ry, ry);   // Bounds2f::Lerp
                f[(size_t)y * nx + x] = std::max<float>(0, Gaussian(px, 0, sigma) - expX) * std::max<float>(0, Gaussian(py, 0, sigma) - expY);
            }
        cond.resize(ny);
        std::vector<float> marg(ny);
        for (int v = 0; v < ny; ++v) { cond[v].init(&f[(size_t)v * nx], nx, -rx, rx); marg[v] = cond[v].funcInt; }
        marginal.init(marg.data(), ny, -ry, ry);
    }
    void Sample(const float *u, float *p, float *weight) const {
        float pdfs[2]; int uv[2];
        float d1 = marginal.Sample(u[1], &pdfs[1], &uv[1]);
        float d0 = cond[uv[1]].Sample(u[0], &pdfs[0], &uv[0]);
        p[0] = d0; p[1] = d1;
        *weight = f[(size_t)uv[1] * nx + uv[0]] / (pdfs[0] * pdfs[1]);
    }
};

struct Scene {
    const vspg_scene_desc *d = nullptr;
    const vspg_guiding_desc *g = nullptr;
    vspg_integrator_params ip;
    std::vector<Medium> media;
    std::vector<float> quadArea, triArea, sphereArea;
    std::vector<float> aliasP, aliasQ; std::vector<int> aliasIdx;  // PowerLightSampler AliasTable bins (p, q, alias)
    BVHLightSampler lightBvh;     // "bvh" sampler (the reference's default)
    GaussianFilterSampler gaussFilter;   // PixelFilter "gaussian" (pbrt's default filter)
    V3 sceneCenter; float sceneRadius = 0;
    int seed = 0;
    bool ownMajorant = false;

    void init(const vspg_scene_desc *desc, const vspg_guiding_desc *gd, const vspg_integrator_params *p, bool ownMaj) {
        d = desc; g = gd; ip = *p; ownMajorant = ownMaj;
        media.resize(d->n_media);
        for (int i = 0; i < d->n_media; i++) media[i].init(&d->media[i], ownMaj);
        quadArea.resize(d->n_quads);
        V3 bmin(Infinity, Infinity, Infinity), bmax(-Infinity, -Infinity, -Infinity);
        for (int i = 0; i < d->n_quads; i++) {
            const vspg_quad &q = d->quads[i];
            V3 p00(q.p00), p10(q.p10), p01(q.p01), p11(q.p11);
            quadArea[i] = QuadArea(p00, p10, p01, p11);
            for (V3 p : {p00, p10, p01, p11})
                for (int a = 0; a < 3; a++) { bmin[a] = std::min(bmin[a], p[a]); bmax[a] = std::max(bmax[a], p[a]); }
        }
        triArea.resize(d->n_triangles); sphereArea.resize(d->n_spheres);
        for (int i = 0; i < d->n_spheres; i++) sphereArea[i] = 6.28318548f * d->spheres[i].radius * (d->spheres[i].radius - -d->spheres[i].radius);  // shapes.h:295
        for (int i = 0; i < d->n_triangles; i++) {
            const vspg_triangle &t = d->triangles[i];
            triArea[i] = 0.5f * Length(Cross(V3(t.p1) - V3(t.p0), V3(t.p2) - V3(t.p0)));   // shapes.h:853-860
            for (V3 p : {V3(t.p0), V3(t.p1), V3(t.p2)})
                for (int a = 0; a < 3; a++) { bmin[a] = std::min(bmin[a], p[a]); bmax[a] = std::max(bmax[a], p[a]); }
        }
        for (int i = 0; i < d->n_spheres; i++) {   // shapes.cpp:33-36 Sphere::Bounds through translate(center)
            const vspg_sphere &sp = d->spheres[i];
            for (int a = 0; a < 3; a++) { bmin[a] = std::min(bmin[a], -sp.radius + sp.center[a]); bmax[a] = std::max(bmax[a], sp.radius + sp.center[a]); }
        }
        // Bounds3::BoundingSphere: center = (pMin+pMax)/2, radius = Inside(center) ? Distance(center, pMax) : 0
        sceneCenter = (bmin + bmax) / 2;
        sceneRadius = (d->n_quads + d->n_triangles + d->n_spheres) ? Distance(sceneCenter, bmax) : 0;
        // PowerLightSampler (lightsamplers.cpp:76-96): power = (Phi(lambda) / lambda.PDF()).Average() with
        // lambda = SampledWavelengths::SampleVisible(0.5f) (spectrum.h:369-386; sampling.h:163-171), then an AliasTable
        // (sampling.cpp:563-618)
        {
            S3 pdf;
            for (int i = 0; i < 3; ++i) {
                float up = 0.5f + float(i) / 3;
                if (up > 1) up -= 1;
                float lambda = 538 - 138.888889f * vspg_atanhf(0.85691062f - 1.82750197f * up);
                pdf[i] = (lambda < 360 || lambda > 830) ? 0.f : 0.0039398042f / Sqr(vspg_coshf(0.0072f * (lambda - 538)));
            }
            int n = d->n_lights;
            std::vector<float> lightPower;
            for (int i = 0; i < n; i++) {
                S3 phi = LightPhi(i), r;
                for (int c = 0; c < 3; ++c) r[c] = (pdf[c] != 0) ? phi[c] / pdf[c] : 0.f;
                lightPower.push_back(r.Average());
            }
            float acc = 0.f;
            for (float v : lightPower) acc += v;
            if (acc == 0.f) for (float &v : lightPower) v = 1.f;
            double dsum = 0.;
            for (float v : lightPower) dsum += v;
            float sum = (float)dsum;   // Float sum = std::accumulate(weights.begin(), weights.end(), 0.)
            aliasP.assign(n, 0.f); aliasQ.assign(n, 1.f); aliasIdx.assign(n, -1);
            for (int i = 0; i < n; i++) aliasP[i] = lightPower[i] / sum;
            struct Outcome { float pHat; int index; };
            std::vector<Outcome> under, over;
            for (int i = 0; i < n; i++) {
                float pHat = aliasP[i] * n;
                if (pHat < 1) under.push_back(Outcome{pHat, i}); else over.push_back(Outcome{pHat, i});
            }
            while (!under.empty() && !over.empty()) {
                Outcome un = under.back(), ov = over.back();
                under.pop_back(); over.pop_back();
                aliasQ[un.index] = un.pHat; aliasIdx[un.index] = ov.index;
                float pExcess = un.pHat + ov.pHat - 1;
                if (pExcess < 1) under.push_back(Outcome{pExcess, ov.index}); else over.push_back(Outcome{pExcess, ov.index});
            }
            while (!over.empty()) { aliasQ[over.back().index] = 1; aliasIdx[over.back().index] = -1; over.pop_back(); }
            while (!under.empty()) { aliasQ[under.back().index] = 1; aliasIdx[under.back().index] = -1; under.pop_back(); }
        }
        lightBvh.Build(d->lights, d->n_lights, d->quads, quadArea, d->triangles, &triArea, d->spheres, &sphereArea);
        if (d->camera.filter_type == 1) gaussFilter.init(d->camera.filter_radius[0], d->camera.filter_radius[1], d->camera.filter_sigma);
    }
    static bool IsRectangle(V3 p00, V3 p10, V3 p01, V3 p11) {  // shapes.h BilinearPatch::IsRectangle
        if (p00 == p01 || p01 == p11 || p11 == p10 || p10 == p00) return false;
        V3 n = Normalize(Cross(p10 - p00, p01 - p00));
        if (AbsDotN(Normalize(p11 - p00), n) > 1e-5f) return false;
        V3 pCenter = (p00 + p01 + p10 + p11) / 4;
        float d2[4] = {DistanceSquared(p00, pCenter), DistanceSquared(p01, pCenter), DistanceSquared(p10, pCenter), DistanceSquared(p11, pCenter)};
        for (int i = 1; i < 4; ++i)
            if (std::abs(d2[i] - d2[0]) / d2[0] > 1e-4f) return false;
        return true;
    }
    static float QuadArea(V3 p00, V3 p10, V3 p01, V3 p11) {  // shapes.cpp:1038-1068
        if (IsRectangle(p00, p10, p01, p11)) return Distance(p00, p01) * Distance(p00, p10);
        constexpr int na = 3;
        V3 p[na + 1][na + 1];
        for (int i = 0; i <= na; ++i) {
            float u = float(i) / float(na);
            for (int j = 0; j <= na; ++j) {
                float v = float(j) / float(na);
                p[i][j] = LerpV(u, LerpV(v, p00, p01), LerpV(v, p10, p11));
            }
        }
        float area = 0;
        for (int i = 0; i < na; ++i)
            for (int j = 0; j < na; ++j) area += 0.5f * Length(Cross(p[i + 1][j + 1] - p[i][j], p[i + 1][j] - p[i][j + 1]));
        return area;
    }
    float ShapeArea(const vspg_light &l) const { return l.shape_kind == 2 ? sphereArea[l.shape_index] : l.shape_kind == 1 ? triArea[l.shape_index] : quadArea[l.shape_index]; }
    static S3 LightL(const vspg_light &l) { return l.scale * S3(l.L); }   // scale * Lemit->Sample(lambda), RGB build: Sample = rgb (spectrum.h:786-803)
    S3 LightPhi(int i) const {  // lights.cpp:196-198, 251-253, 826-842, 1038-1040
        const vspg_light &l = d->lights[i];
        S3 L(l.L);
        switch (l.type) {
        case VSPG_LIGHT_DIFFUSE_AREA: return Pi * (l.two_sided ? 2 : 1) * ShapeArea(l) * (L * l.scale);
        case VSPG_LIGHT_POINT: return 4 * Pi * l.scale * L;
        case VSPG_LIGHT_SPOT: return l.scale * L * 2 * Pi * ((1 - l.cos_falloff_start) + (l.cos_falloff_start - l.cos_falloff_end) / 2);   // lights.cpp:1452-1455
        case VSPG_LIGHT_DISTANT: return l.scale * L * Pi * Sqr(sceneRadius);
        default: return 4 * Pi * Pi * Sqr(sceneRadius) * l.scale * L;
        }
    }

    // Closest hit by brute force over all primitives with the reference's intersection arithmetic
    // (the product walks its own BVH; the closest hit is the same hit).
    Hit Intersect(V3 ro, V3 rd, float tMax) const {
        Hit h;
        float best = tMax;
        for (int i = 0; i < d->n_quads; i++) {
            const vspg_quad &q = d->quads[i];
            float u, v, t;
            if (IntersectBilinearPatch(ro, rd, best, V3(q.p00), V3(q.p10), V3(q.p01), V3(q.p11), &u, &v, &t)) {
                best = t; h.valid = true; h.t = t; h.kind = 0; h.index = i; h.u = u; h.v = v;
            }
        }
        // triangles: tested against the ray's own tMax, then compared with the closest hit so far (closest-hit
        // semantics of the aggregate; the product's BVH does the same)
        for (int i = 0; i < d->n_triangles; i++) {
            const vspg_triangle &q = d->triangles[i];
            float b[3], t;
            if (IntersectTriangle(ro, rd, tMax, V3(q.p0), V3(q.p1), V3(q.p2), b, &t) && t < best) {
                best = t; h.valid = true; h.t = t; h.kind = 1; h.index = i; h.u = b[0]; h.v = b[1];
                h.b[0] = b[0]; h.b[1] = b[1]; h.b[2] = b[2];
            }
        }
        // spheres: same convention (Sphere::BasicIntersect tests both roots against tMax)
        for (int i = 0; i < d->n_spheres; i++) {
            const vspg_sphere &q = d->spheres[i];
            float t; V3 pObj;
            if (IntersectSphere(ro, rd, tMax, V3(q.center), q.radius, &t, &pObj) && t < best) {
                best = t; h.valid = true; h.t = t; h.kind = 2; h.index = i; h.u = h.v = 0; h.pObj = pObj;
            }
        }
        return h;
    }
    // One shape's own Intersect (Shape::Intersect(ray, tMax)), as Shape::PDF(ctx, wi) calls it
    Hit IntersectPrim(int kind, int index, V3 ro, V3 rd, float tMax) const {
        Hit h;
        if (kind == 0) {
            const vspg_quad &q = d->quads[index];
            float u, v, t;
            if (IntersectBilinearPatch(ro, rd, tMax, V3(q.p00), V3(q.p10), V3(q.p01), V3(q.p11), &u, &v, &t)) { h.valid = true; h.t = t; h.kind = 0; h.index = index; h.u = u; h.v = v; }
        } else if (kind == 1) {
            const vspg_triangle &q = d->triangles[index];
            float b[3], t;
            if (IntersectTriangle(ro, rd, tMax, V3(q.p0), V3(q.p1), V3(q.p2), b, &t)) { h.valid = true; h.t = t; h.kind = 1; h.index = index; h.u = b[0]; h.v = b[1]; h.b[0] = b[0]; h.b[1] = b[1]; h.b[2] = b[2]; }
        } else {
            const vspg_sphere &q = d->spheres[index];
            float t; V3 pObj;
            if (IntersectSphere(ro, rd, tMax, V3(q.center), q.radius, &t, &pObj)) { h.valid = true; h.t = t; h.kind = 2; h.index = index; h.pObj = pObj; }
        }
        return h;
    }
    // shapes.h:1398-1500 BilinearPatch::InteractionFromIntersection (no mesh uv / normals) + primitive.cpp
    // GeometricPrimitive::Intersect -> SetIntersectionProperties (interaction.h:218-229)
    int HitMaterial(const Hit &h) const {
        return h.kind == 2 ? d->spheres[h.index].material : h.kind == 1 ? d->triangles[h.index].material : d->quads[h.index].material;
    }
    SurfaceIntr MakeIntr(const Hit &h, V3 rayD, int rayMedium) const {
        SurfaceIntr s;
        if (h.kind == 2) {  // shapes.h:241-292 Sphere::InteractionFromIntersection (uv unused by the diffuse path: left 0)
            const vspg_sphere &q = d->spheres[h.index];
            SphereInteraction(h.pObj, V3(q.center), q.radius, q.flip_normal != 0, &s.p, &s.pError, &s.n, &s.dpdu);
            s.ns = s.n;
            s.wo = Normalize(-rayD);
            s.uv[0] = s.uv[1] = 0;
            s.material = q.material; s.areaLight = q.area_light;
            s.hasInterface = q.has_interface && (q.medium_inside != q.medium_outside);
            s.mediumInside = q.medium_inside; s.mediumOutside = q.medium_outside; s.rayMedium = rayMedium;
            return s;
        }
        if (h.kind == 1) {  // shapes.h:884-945 Triangle::InteractionFromIntersection, default uvs, no mesh normals
            const vspg_triangle &q = d->triangles[h.index];
            V3 p0(q.p0), p1(q.p1), p2(q.p2);
            V3 dp02 = p0 - p2, dp12 = p1 - p2;
            const float duv02[2] = {-1, -1}, duv12[2] = {0, -1};
            float determinant = DifferenceOfProducts(duv02[0], duv12[1], duv02[1], duv12[0]);
            float invdet = 1 / determinant;
            V3 dpdu = V3(DifferenceOfProducts(duv12[1], dp02.x, duv02[1], dp12.x), DifferenceOfProducts(duv12[1], dp02.y, duv02[1], dp12.y),
                         DifferenceOfProducts(duv12[1], dp02.z, duv02[1], dp12.z)) * invdet;
            V3 dpdv = V3(DifferenceOfProducts(duv02[0], dp12.x, duv12[0], dp02.x), DifferenceOfProducts(duv02[0], dp12.y, duv12[0], dp02.y),
                         DifferenceOfProducts(duv02[0], dp12.z, duv12[0], dp02.z)) * invdet;
            if (LengthSquared(Cross(dpdu, dpdv)) == 0) {
                V3 ng = Cross(p2 - p0, p1 - p0);
                CoordinateSystem(Normalize(ng), &dpdu, &dpdv);
            }
            // Point3fi(pHit, pError) (shapes.h:929-935): the interval is rounded outwards, p() is its midpoint and Error() half its width
            V3i pi = V3i::FromValueAndError(h.b[0] * p0 + h.b[1] * p1 + h.b[2] * p2, gamma_(7) * (Abs(h.b[0] * p0) + Abs(h.b[1] * p1) + Abs(h.b[2] * p2)));
            s.p = pi.Mid(); s.pError = pi.Error();
            s.n = Normalize(Cross(dp02, dp12));
            if (q.flip_normal) s.n = -s.n;
            s.ns = s.n; s.dpdu = dpdu;
            s.wo = Normalize(-rayD);
            s.uv[0] = h.b[1] + h.b[2]; s.uv[1] = h.b[2];
            s.material = q.material; s.areaLight = q.area_light;
            s.hasInterface = q.has_interface && (q.medium_inside != q.medium_outside);
            s.mediumInside = q.medium_inside; s.mediumOutside = q.medium_outside; s.rayMedium = rayMedium;
            return s;
        }
        const vspg_quad &q = d->quads[h.index];
        V3 p00(q.p00), p10(q.p10), p01(q.p01), p11(q.p11);
        V3 pLerp = LerpV(h.u, LerpV(h.v, p00, p01), LerpV(h.v, p10, p11));
        V3 dpdu = LerpV(h.v, p10, p11) - LerpV(h.v, p00, p01);
        V3 dpdv = LerpV(h.u, p01, p11) - LerpV(h.u, p00, p10);
        V3 pAbsSum = Abs(p00) + Abs(p01) + Abs(p10) + Abs(p11);
        V3i pi = V3i::FromValueAndError(pLerp, gamma_(6) * pAbsSum);   // shapes.h:1464-1472 Point3fi(p, pError): outward-rounded interval
        s.p = pi.Mid(); s.pError = pi.Error();
        s.n = Normalize(Cross(dpdu, dpdv));
        if (q.flip_normal) s.n = -s.n;
        s.ns = s.n; s.dpdu = dpdu;
        s.wo = Normalize(-rayD);
        s.uv[0] = h.u; s.uv[1] = h.v;
        s.material = q.material; s.areaLight = q.area_light;
        s.hasInterface = q.has_interface && (q.medium_inside != q.medium_outside);
        s.mediumInside = q.medium_inside; s.mediumOutside = q.medium_outside; s.rayMedium = rayMedium;
        return s;
    }
    // lights.h:492-515 DiffuseAreaLight::L
    S3 AreaL(int light, V3 n, V3 w) const {
        const vspg_light &l = d->lights[light];
        if (!l.two_sided && DotN(n, w) < 0) return S3(0.f);
        return Scene::LightL(l);
    }
};

// cpu/guiding.h:57-358 GuidedBSDF for the DiffuseBxDF
struct GuidedBSDF {
    const Scene *sc; Sampler *sampler;
    bool enableGuiding, enableScatterGuiding, ris;
    float guidingProbability = 0.5f;
    bool useGuiding = false, useScatterGuiding = false;
    SamplingDistribution dist;
    Frame frame; S3 R; bool hasR = false;
    int lookups = 0;
    void setBSDF(V3 ns, V3 dpdus, S3 refl) { frame = Frame::FromXZ(Normalize(dpdus), ns); R = refl; hasR = (bool)R; }
    bool IsNonSpecular() const { return hasR; }  // DiffuseReflection flag only when R != 0 (bxdfs.h:82-84)
    S3 f(V3 woR, V3 wiR) const {  // bsdf.h:41-47 + bxdfs.h:39-43
        V3 wi = frame.ToLocal(wiR), wo = frame.ToLocal(woR);
        if (wo.z == 0) return S3(0.f);
        if (!(wo.z * wi.z > 0)) return S3(0.f);
        return R * InvPi;
    }
    float bsdfPDF(V3 woR, V3 wiR) const {
        V3 wo = frame.ToLocal(woR), wi = frame.ToLocal(wiR);
        if (wo.z == 0) return 0;
        if (!(wo.z * wi.z > 0)) return 0;
        return CosineHemispherePDF(std::abs(wi.z));
    }
    struct Sample { bool valid = false; S3 f; V3 wi; float pdf = 0, bsdfPdf = 0, misPdf = 0; };
    Sample bsdfSample(V3 woR, const float *u2) const {  // bsdf.h:58-80 + bxdfs.h:46-59
        Sample s;
        V3 wo = frame.ToLocal(woR);
        if (wo.z == 0 || !hasR) return s;
        V3 wi = SampleCosineHemisphere(u2);
        if (wo.z < 0) wi.z *= -1;
        float pdf = CosineHemispherePDF(std::abs(wi.z));
        S3 fv = R * InvPi;
        if (!fv || pdf == 0 || wi.z == 0) return s;
        s.valid = true; s.f = fv; s.wi = frame.FromLocal(wi); s.pdf = s.bsdfPdf = s.misPdf = pdf;
        return s;
    }
    // guiding.h:83-109
    bool init(V3 p, V3 rayD, V3 ns, float &rand) {
        bool success = false;
        if (IsNonSpecular() && sc->g && sc->g->surface) {
            int leaf;
            int r = field_get_region(sc->g->surface, p, &rand, &leaf);
            if (leaf >= 0) lookups++;
            if (r >= 0 && sc->g->surface->regions[r].valid) {
                dist.init(&sc->g->surface->regions[r], p);
                dist.region = r;
                V3 n = ns;
                if (DotN(-rayD, ns) < 0.f) n = -n;
                dist.applyCosineProduct(n);
                success = true;
            }
        }
        useGuiding = enableGuiding ? success : false;
        useScatterGuiding = enableScatterGuiding ? success : false;
        return success;
    }
    float PDF(V3 wo, V3 wi) const {  // guiding.h:271-289
        float b = bsdfPDF(wo, wi);
        if (useGuiding) {
            float gp = dist.pdf(wi);
            if (!ris) return ((1.0f - guidingProbability) * b) + (guidingProbability * gp);
            return (0.5f * b) + (0.5f * gp);
        }
        return b;
    }
    float VolumeScatterProbability(V3 wi) const { return useScatterGuiding ? dist.vsp(wi) : -1.f; }  // :295-305
    Sample Sample_f(V3 wo, float u, const float *u2) {  // :120-269
        if (!ris) {
            Sample bs; bool sampleBSDF = true;
            if (useGuiding) {
                if (guidingProbability > u) { u /= guidingProbability; sampleBSDF = false; }
                else { u -= guidingProbability; u /= (1.0f - guidingProbability); sampleBSDF = true; }
            }
            if (sampleBSDF) {
                bs = bsdfSample(wo, u2);
                if (bs.valid && useGuiding) {
                    float gp = dist.pdf(bs.wi);
                    bs.bsdfPdf = bs.pdf;
                    bs.pdf = ((1.0f - guidingProbability) * bs.pdf) + (guidingProbability * gp);
                    bs.misPdf = bs.pdf;
                }
            } else {
                V3 wi; float gp = dist.samplePdf(u2[0], u2[1], &wi);
                S3 fv = f(wo, wi); float bp = bsdfPDF(wo, wi);
                if (bp > 0.f) {
                    float pdf = ((1.0f - guidingProbability) * bp) + (guidingProbability * gp);
                    bs.valid = true; bs.f = fv; bs.wi = wi; bs.pdf = pdf; bs.bsdfPdf = bp; bs.misPdf = pdf;
                }
            }
            return bs;
        }
        if (!useGuiding) return bsdfSample(wo, u2);
        const float uniformIncomingRadiancePDF = ((1.0f / (4.0f * M_PI)));
        struct R_ { float bsdfPDF = 0, guidingPDF = 0, misPDF = 0, liPDF = 0; S3 f; V3 wi; float w = 0; } rs[2];
        Sample bs0 = bsdfSample(wo, u2);
        if (bs0.valid) {
            rs[0].f = bs0.f; rs[0].bsdfPDF = bs0.pdf; rs[0].wi = bs0.wi;
            rs[0].guidingPDF = dist.pdf(bs0.wi); rs[0].liPDF = dist.pdfLi(bs0.wi);
            rs[0].misPDF = 0.5f * (rs[0].bsdfPDF + rs[0].guidingPDF);
        }
        float s2[2]; sampler->Get2D(s2);
        V3 wi1; rs[1].guidingPDF = dist.samplePdf(s2[0], s2[1], &wi1);
        rs[1].liPDF = dist.pdfLi(wi1);
        rs[1].f = f(wo, wi1); rs[1].bsdfPDF = bsdfPDF(wo, wi1); rs[1].wi = wi1;
        rs[1].misPDF = 0.5f * (rs[1].bsdfPDF + rs[1].guidingPDF);
        float sumW = 0.f; int nS = 0;
        for (int i = 0; i < 2; i++)
            if (rs[i].bsdfPDF > 0.f) {
                rs[i].w = (rs[i].bsdfPDF * ((1.0f - guidingProbability) * uniformIncomingRadiancePDF + guidingProbability * rs[i].liPDF));
                rs[i].w /= rs[i].misPDF;
                sumW += rs[i].w; nS++;
            }
        Sample bs;
        if (nS == 0 || sumW <= 0.f) return bs;
        int idx = 0;
        float s1 = sumW * sampler->Get1D();
        float sumRis = 0.f;
        for (int i = 0; i < 2; i++) { sumRis += rs[i].w; if (s1 <= sumRis) { idx = i; break; } }
        float pdf = (rs[idx].w * rs[idx].misPDF) * (float(2) / sumW);
        bs.valid = true; bs.f = rs[idx].f; bs.wi = rs[idx].wi; bs.pdf = pdf; bs.bsdfPdf = rs[idx].bsdfPDF; bs.misPdf = rs[idx].misPDF;
        return bs;
    }
};

// cpu/guiding.h:360-638 GuidedPhaseFunction over HGPhaseFunction (media.h:44-74)
struct GuidedPhase {
    const Scene *sc; Sampler *sampler;
    bool enableGuiding, enableScatterGuiding, ris;
    float guidingProbability = 0.5f;
    bool useGuiding = false, useScatterGuiding = false;
    SamplingDistribution dist;
    float g = 0;
    int lookups = 0;
    float p(V3 wo, V3 wi) const { return HenyeyGreenstein(Dot(wo, wi), g); }
    struct Sample { bool valid = false; float p = 0; V3 wi; float pdf = 0, phasePdf = 0, misPdf = 0; };
    Sample phaseSample(V3 wo, const float *u) const {
        Sample s; float pdf;
        s.wi = SampleHenyeyGreenstein(wo, g, u, &pdf);
        s.valid = true; s.p = pdf; s.pdf = s.phasePdf = s.misPdf = pdf;
        return s;
    }
    bool init(float g_, V3 pos, V3 wo, float &rand) {  // :383-398 (wo argument = ray.d, vspg.cpp:810)
        g = g_;
        bool success = false;
        if (sc->g && sc->g->volume) {
            int leaf;
            int r = field_get_region(sc->g->volume, pos, &rand, &leaf);
            if (leaf >= 0) lookups++;
            if (r >= 0 && sc->g->volume->regions[r].valid) {
                dist.init(&sc->g->volume->regions[r], pos);
                dist.region = r;
                dist.applyHGProduct(wo, g);
                success = true;
            }
        }
        useGuiding = enableGuiding ? success : false;
        useScatterGuiding = enableScatterGuiding ? success : false;
        return success;
    }
    float PDF(V3 wo, V3 wi) const {
        float ph = p(wo, wi);
        if (useGuiding) {
            float gp = dist.pdf(wi);
            if (!ris) return ((1.0f - guidingProbability) * ph) + (guidingProbability * gp);
            return (0.5f * ph) + (0.5f * gp);
        }
        return ph;
    }
    float VolumeScatterProbability(V3 wi) const { return useScatterGuiding ? dist.vsp(wi) : -1.f; }
    Sample Sample_p(V3 wo, const float *uin) {
        float u[2] = {uin[0], uin[1]};
        if (!ris) {
            Sample ps; bool samplePhase = true;
            if (useGuiding) {
                if (guidingProbability > u[0]) { u[0] /= guidingProbability; samplePhase = false; }
                else { u[0] -= guidingProbability; u[0] /= (1.0f - guidingProbability); samplePhase = true; }
            }
            if (samplePhase) {
                ps = phaseSample(wo, u);
                if (ps.valid && useGuiding) {
                    float gp = dist.pdf(ps.wi);
                    ps.phasePdf = ps.pdf;
                    ps.pdf = ((1.0f - guidingProbability) * ps.pdf) + (guidingProbability * gp);
                    ps.misPdf = ps.pdf;
                }
            } else {
                V3 wi; float gp = dist.samplePdf(u[0], u[1], &wi);
                float pv = p(wo, wi); float ph = pv;
                if (ph > 0.f) {
                    float pdf = ((1.0f - guidingProbability) * ph) + (guidingProbability * gp);
                    ps.valid = true; ps.p = pv; ps.wi = wi; ps.pdf = pdf; ps.phasePdf = ph; ps.misPdf = pdf;
                }
            }
            return ps;
        }
        if (!useGuiding) return phaseSample(wo, u);
        const float uniformIncomingRadiancePDF = ((1.0f / (4.0f * M_PI)));
        struct R_ { float phasePDF = 0, guidingPDF = 0, misPDF = 0, liPDF = 0, p = 0; V3 wi; float w = 0; } rs[2];
        Sample ps0 = phaseSample(wo, u);
        rs[0].p = ps0.p; rs[0].phasePDF = ps0.pdf; rs[0].wi = ps0.wi;
        rs[0].guidingPDF = dist.pdf(ps0.wi); rs[0].liPDF = dist.pdfLi(ps0.wi);
        rs[0].misPDF = 0.5f * (rs[0].phasePDF + rs[0].guidingPDF);
        float s2[2]; sampler->Get2D(s2);
        V3 wi1; rs[1].guidingPDF = dist.samplePdf(s2[0], s2[1], &wi1);
        rs[1].liPDF = dist.pdfLi(wi1);
        rs[1].p = p(wo, wi1); rs[1].phasePDF = rs[1].p; rs[1].wi = wi1;
        rs[1].misPDF = 0.5f * (rs[1].phasePDF + rs[1].guidingPDF);
        float sumW = 0.f; int nS = 0;
        for (int i = 0; i < 2; i++)
            if (rs[i].phasePDF > 0.f) {
                rs[i].w = (rs[i].phasePDF * ((1.0f - guidingProbability) * uniformIncomingRadiancePDF + guidingProbability * rs[i].liPDF));
                rs[i].w /= rs[i].misPDF;
                sumW += rs[i].w; nS++;
            }
        Sample ps;
        if (nS == 0 || sumW <= 0.f) return ps;
        int idx = 0;
        float s1 = sumW * sampler->Get1D();
        float sumRis = 0.f;
        for (int i = 0; i < 2; i++) { sumRis += rs[i].w; if (s1 <= sumRis) { idx = i; break; } }
        float pdf = (rs[idx].w * rs[idx].misPDF) * (float(2) / sumW);
        ps.valid = true; ps.p = rs[idx].p; ps.wi = rs[idx].wi; ps.pdf = pdf; ps.phasePdf = rs[idx].phasePDF; ps.misPdf = rs[idx].misPDF;
        return ps;
    }
};

struct Counters { long paths = 0, queries = 0, dda = 0, lookups = 0, shadow = 0; };

struct Integrator {
    const Scene *sc;
    Counters cnt;
    // storeTrBuffer (vspg.cpp:190-193, 727-728; cpu/trbuffer.h:40-45): per-pixel running mean of the depth-0 ratio-tracking
    // transmittance estimate.  Shared by the worker threads (each pixel is rendered by one thread, samples in order).
    // storeISGBuffer: records of 8 floats per pixel {surfSpp, volSpp, surfContrib rgb, volContrib rgb}
    // (openpgl/imagespace/ImageSpaceGuidingBuffer.h:744-797 addSample with VSP_USE_PVOL_EST), the estimate of the last
    // Update() (null until the first one), and the event type of the path being traced
    float *isgRec = nullptr; const float *isgTrained = nullptr; bool isgVariance = false, firstVolume = false;
    void IsgAddSample(int px, int py, S3 L) {
        float *r = isgRec + 8 * ((size_t)py * sc->d->camera.width + px);
        float *spp = r + (firstVolume ? 1 : 0), *c = r + (firstVolume ? 5 : 2);
        *spp += 1;
        float alpha = 1.f / *spp;
        for (int k = 0; k < 3; k++) {
            float q = isgVariance ? L[k] * L[k] : L[k];
            c[k] = (1.f - alpha) * c[k] + alpha * q;
        }
    }
    // :632-742 update(), denoiser = identity
    static void IsgUpdate(const float *rec, size_t npix, bool variance, float *vsp) {
        for (size_t i = 0; i < npix; i++) {
            const float *r = rec + 8 * i;
            float pVol = r[1] / (r[0] + r[1]);
            pVol = std::max(0.f, std::min(1.f, pVol));
            float sv[3], vv[3];
            for (int k = 0; k < 3; k++) {
                float sm = r[2 + k], vm = r[5 + k];
                if (variance) { sv[k] = sm > 0.f ? (1.f - pVol) * std::sqrt(sm) : 0.f; vv[k] = vm > 0.f ? pVol * std::sqrt(vm) : 0.f; }
                else { sv[k] = (1.f - pVol) * sm; vv[k] = pVol * vm; }
            }
            float c[3] = {sv[0] + vv[0], sv[1] + vv[1], sv[2] + vv[2]};
            float cs = std::max(c[0], std::max(c[1], c[2])), vs = std::max(vv[0], std::max(vv[1], vv[2]));
            vsp[i] = cs > 0.f ? vs / cs : -1.f;
        }
    }
    float *trBuf = nullptr; int *trSpp = nullptr;
    void TrAddSample(int px, int py, S3 tr) {
        int pixIdx = py * sc->d->camera.width + px;
        trSpp[pixIdx] += 1;
        float alpha = 1.f / trSpp[pixIdx];
        for (int k = 0; k < 3; k++) trBuf[3 * pixIdx + k] = (1.f - alpha) * trBuf[3 * pixIdx + k] + alpha * tr[k];
    }

    // ---------------------------------------------------------------- light sampling
    // ctxP / ctxNs: LightSampleContext::p() and ::ns of the vertex (ns = 0 in a medium)
    int SampleLight(float u, float *pmf, V3 ctxP, V3 ctxNs) const {  // lightsamplers.h:33-38 (uniform) / power pmf via CDF walk / bvh
        int n = sc->d->n_lights;
        if (n == 0) return -1;
        if (sc->ip.light_sampler == VSPG_LIGHTSAMPLER_BVH) return sc->lightBvh.Sample(ctxP, ctxNs, u, pmf);
        if (sc->ip.light_sampler == VSPG_LIGHTSAMPLER_UNIFORM) {
            int li = std::min<int>((int)(u * n), n - 1);
            *pmf = 1.f / n;
            return li;
        }
        // AliasTable::Sample (sampling.cpp:620-645)
        int offset = std::min<int>(u * n, n - 1);
        float up = std::min<float>(u * n - offset, OneMinusEpsilon);
        if (up < sc->aliasQ[offset]) { *pmf = sc->aliasP[offset]; return offset; }
        int alias = sc->aliasIdx[offset];
        *pmf = sc->aliasP[alias];
        return alias;
    }
    float LightPMF(int light, const LightSampleCtx &ctx) const {
        int n = sc->d->n_lights;
        if (n == 0) return 0;
        if (sc->ip.light_sampler == VSPG_LIGHTSAMPLER_BVH) return sc->lightBvh.PMF(ctx.p, ctx.ns, light);
        if (sc->ip.light_sampler == VSPG_LIGHTSAMPLER_UNIFORM) return 1.f / n;
        return sc->aliasP[light];
    }
    // shapes.h:1053-1125 Triangle::Sample(ctx, u) (and :1014-1048 Sample(u) for the area-sampling fallback)
    static bool SampleTriangle(const vspg_triangle &q, float area, const LightSampleCtx &ctx, const float *u, V3 *pOut, V3 *pErr, V3 *nOut, float *pdfOut) {
        V3 p0(q.p0), p1(q.p1), p2(q.p2);
        V3 n = Normalize(Cross(p1 - p0, p2 - p0));
        if (q.flip_normal) n = -n;
        float solidAngle = SphericalTriangleArea(Normalize(p0 - ctx.p), Normalize(p1 - ctx.p), Normalize(p2 - ctx.p));
        float b[3];
        if (solidAngle < MinSphericalSampleAreaTri || solidAngle > MaxSphericalSampleAreaTri) {
            SampleUniformTriangle(u, b);
            // Interaction(Point3fi(p, pError), ...) (shapes.h:1040-1046): outward-rounded interval, p() = midpoint
            V3i pi = V3i::FromValueAndError(b[0] * p0 + b[1] * p1 + b[2] * p2, gamma_(6) * (Abs(b[0] * p0) + Abs(b[1] * p1) + Abs((1 - b[0] - b[1]) * p2)));
            V3 p = pi.Mid(); *pErr = pi.Error();
            V3 wi = p - ctx.p;
            if (LengthSquared(wi) == 0) return false;
            wi = Normalize(wi);
            float pdf = (1 / area) / (AbsDotN(n, -wi) / DistanceSquared(ctx.p, p));
            if (std::isinf(pdf)) return false;
            *pOut = p; *nOut = n; *pdfOut = pdf;
            return true;
        }
        float pdf = 1, us[2] = {u[0], u[1]};
        if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
            V3 wi[3] = {Normalize(p0 - ctx.p), Normalize(p1 - ctx.p), Normalize(p2 - ctx.p)};
            float w[4] = {std::max<float>(0.01, AbsDotN(ctx.ns, wi[1])), std::max<float>(0.01, AbsDotN(ctx.ns, wi[1])),
                          std::max<float>(0.01, AbsDotN(ctx.ns, wi[0])), std::max<float>(0.01, AbsDotN(ctx.ns, wi[2]))};
            float uw[2];
            SampleBilinear(us, w, uw);
            us[0] = uw[0]; us[1] = uw[1];
            pdf = BilinearPDF(us, w);
        }
        float triPDF;
        V3 v[3] = {p0, p1, p2};
        if (!SampleSphericalTriangle(v, ctx.p, us, b, &triPDF)) { b[0] = b[1] = b[2] = 0; }
        if (triPDF == 0) return false;
        pdf *= triPDF;
        V3i pi = V3i::FromValueAndError(b[0] * p0 + b[1] * p1 + b[2] * p2, gamma_(6) * (Abs(b[0] * p0) + Abs(b[1] * p1) + Abs((1 - b[0] - b[1]) * p2)));   // shapes.h:1117-1123
        *pErr = pi.Error();
        *pOut = pi.Mid(); *nOut = n; *pdfOut = pdf;
        return true;
    }
    // shapes.h:300-361 Sphere::Sample(ctx, u) (and shapes.cpp:38-59 Sample(u) for reference points inside the sphere).
    // ctx.pi is exact here, so ctx.OffsetRayOrigin(pCenter) == ctx.p().
    static bool SampleSphere(const vspg_sphere &q, float area, const LightSampleCtx &ctx, const float *u, V3 *pOut, V3 *pErr, V3 *nOut, float *pdfOut) {
        V3 pCenter(q.center); float radius = q.radius;
        if (DistanceSquared(ctx.p, pCenter) <= Sqr(radius)) {
            V3 pObj = radius * SampleUniformSphere(u);
            pObj = pObj * (radius / Length(pObj - V3(0, 0, 0)));
            V3 pObjError = gamma_(5) * Abs(pObj);
            V3 n = Normalize(pObj);
            if (q.flip_normal) n = -n;
            V3i pi = XfPoint(TranslateM(pCenter), V3i::FromValueAndError(pObj, pObjError));
            V3 p = pi.Mid();
            V3 wi = p - ctx.p;
            if (LengthSquared(wi) == 0) return false;
            wi = Normalize(wi);
            float pdf = (1 / area) / (AbsDotN(n, -wi) / DistanceSquared(ctx.p, p));
            if (std::isinf(pdf)) return false;
            *pOut = p; *pErr = pi.Error(); *nOut = n; *pdfOut = pdf;
            return true;
        }
        float sinThetaMax = radius / Distance(ctx.p, pCenter);
        float sin2ThetaMax = Sqr(sinThetaMax);
        float cosThetaMax = SafeSqrt(1 - sin2ThetaMax);
        float oneMinusCosThetaMax = 1 - cosThetaMax;
        float cosTheta = (cosThetaMax - 1) * u[0] + 1;
        float sin2Theta = 1 - Sqr(cosTheta);
        if (sin2ThetaMax < 0.00068523f) {
            sin2Theta = sin2ThetaMax * u[0];
            cosTheta = std::sqrt(1 - sin2Theta);
            oneMinusCosThetaMax = sin2ThetaMax / 2;
        }
        float cosAlpha = sin2Theta / sinThetaMax + cosTheta * SafeSqrt(1 - sin2Theta / Sqr(sinThetaMax));
        float sinAlpha = SafeSqrt(1 - Sqr(cosAlpha));
        float phi = u[1] * 2 * Pi;
        V3 w = SphericalDirection(sinAlpha, cosAlpha, phi);
        V3 fz = Normalize(pCenter - ctx.p), fx, fy;
        CoordinateSystem(fz, &fx, &fy);
        V3 mw = -w;
        V3 n = mw.x * fx + mw.y * fy + mw.z * fz;
        V3 p = pCenter + radius * n;
        if (q.flip_normal) n = -n;
        V3i pi = V3i::FromValueAndError(p, gamma_(5) * Abs(p));   // shapes.h:352-360 Interaction(Point3fi(p, pError), ...)
        *pErr = pi.Error();
        *pOut = pi.Mid(); *nOut = n; *pdfOut = 1 / (2 * Pi * oneMinusCosThetaMax);
        return true;
    }
    struct LiSample { bool valid = false; S3 L; V3 wi; float pdf = 0; V3 pLight, pLightErr, nLight; bool delta = false; };
    // shapes.cpp:1155-1214 Sample(u) (area sampling) + :1254-1280 conversion to solid angle; lights.cpp:796-818
    LiSample SampleLi(int light, const LightSampleCtx &ctx, const float *u) const {
        LiSample ls;
        const vspg_light &l = sc->d->lights[light];
        if (l.type == VSPG_LIGHT_POINT) {  // lights.h:243-250
            V3 p(l.pos);
            ls.wi = Normalize(p - ctx.p);
            ls.L = Scene::LightL(l) / DistanceSquared(p, ctx.p);
            ls.pdf = 1; ls.pLight = p; ls.delta = true; ls.valid = true;
            return ls;
        }
        if (l.type == VSPG_LIGHT_SPOT) {  // lights.h:852-866; I(): lights.cpp:1448-1451; SmoothStep: util/math.h:268-274
            V3 p(l.pos);
            V3 wi = Normalize(p - ctx.p);
            V3 v = -wi;
            const float *m = l.light_from_render;   // Transform::ApplyInverse(Vector3f), util/transform.h:401-406
            V3 wLight = Normalize(V3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z));
            float a = l.cos_falloff_end, b = l.cos_falloff_start, x = wLight.z, ss;
            if (a == b) ss = x < a ? 0.f : 1.f;
            else {
                float t = (x - a) / (b - a);
                t = t < 0 ? 0.f : t > 1 ? 1.f : t;
                ss = t * t * (3 - 2 * t);
            }
            S3 Li = ss * l.scale * S3(l.L) / DistanceSquared(p, ctx.p);
            if (!Li) return ls;
            ls.wi = wi; ls.L = Li; ls.pdf = 1; ls.pLight = p; ls.delta = true; ls.valid = true;
            return ls;
        }
        if (l.type == VSPG_LIGHT_DISTANT) {  // lights.h:316-323
            V3 wi = Normalize(V3(l.pos));
            ls.wi = wi; ls.L = Scene::LightL(l); ls.pdf = 1; ls.delta = true; ls.valid = true;
            ls.pLight = ctx.p + wi * (2 * sc->sceneRadius);
            return ls;
        }
        if (l.type == VSPG_LIGHT_UNIFORM_INFINITE) return ls;  // allowIncompletePDF = true (lights.cpp:1022-1023)
        if (l.shape_kind == 1 || l.shape_kind == 2) {
            V3 p, pErr, n; float pdf = 0;
            bool ok = l.shape_kind == 1 ? SampleTriangle(sc->d->triangles[l.shape_index], sc->triArea[l.shape_index], ctx, u, &p, &pErr, &n, &pdf)
                                        : SampleSphere(sc->d->spheres[l.shape_index], sc->sphereArea[l.shape_index], ctx, u, &p, &pErr, &n, &pdf);
            if (!ok || pdf == 0 || LengthSquared(p - ctx.p) == 0) return ls;   // lights.cpp:803
            V3 wi = Normalize(p - ctx.p);
            S3 Le = sc->AreaL(light, n, -wi);
            if (!Le) return ls;
            ls.valid = true; ls.L = Le; ls.wi = wi; ls.pdf = pdf; ls.pLight = p; ls.pLightErr = pErr; ls.nLight = n;
            return ls;
        }
        const vspg_quad &q = sc->d->quads[l.shape_index];
        V3 p00(q.p00), p10(q.p10), p01(q.p01), p11(q.p11);
        float pdf = 1; float uv[2];
        // shapes.cpp:1254-1327 BilinearPatch::Sample(ctx, u): rectangles that subtend enough solid angle are sampled
        // as spherical rectangles (with the cos-theta bilinear warp at surface reference points)
        V3 v00 = Normalize(p00 - ctx.p), v10 = Normalize(p10 - ctx.p), v01 = Normalize(p01 - ctx.p), v11 = Normalize(p11 - ctx.p);
        if (Scene::IsRectangle(p00, p10, p01, p11) && !(SphericalQuadArea(v00, v10, v11, v01) <= MinSphericalSampleAreaQuad)) {
            float us[2] = {u[0], u[1]};
            if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
                float w[4] = {std::max<float>(0.01, AbsDotN(v00, ctx.ns)), std::max<float>(0.01, AbsDotN(v10, ctx.ns)),
                              std::max<float>(0.01, AbsDotN(v01, ctx.ns)), std::max<float>(0.01, AbsDotN(v11, ctx.ns))};
                float uw[2];
                SampleBilinear(us, w, uw);
                us[0] = uw[0]; us[1] = uw[1];
                pdf *= BilinearPDF(us, w);
            }
            V3 eu = p10 - p00, ev = p01 - p00;
            float quadPDF;
            V3 p = SampleSphericalRectangle(ctx.p, p00, eu, ev, us, &quadPDF);
            pdf *= quadPDF;
            V3 n = Normalize(Cross(eu, ev));
            if (q.flip_normal) n = -n;
            // DiffuseAreaLight::SampleLi (lights.cpp:796-818): Interaction(p, n, time, st) has no error bounds
            if (pdf == 0 || LengthSquared(p - ctx.p) == 0) return ls;
            V3 wi = Normalize(p - ctx.p);
            S3 Le = sc->AreaL(light, n, -wi);
            if (!Le) return ls;
            ls.valid = true; ls.L = Le; ls.wi = wi; ls.pdf = pdf; ls.pLight = p; ls.pLightErr = V3(); ls.nLight = n;
            return ls;
        }
        if (!Scene::IsRectangle(p00, p10, p01, p11)) {
            float w[4] = {Length(Cross(p10 - p00, p01 - p00)), Length(Cross(p10 - p00, p11 - p10)),
                          Length(Cross(p01 - p00, p11 - p01)), Length(Cross(p11 - p10, p11 - p01))};
            SampleBilinear(u, w, uv);
            pdf = BilinearPDF(uv, w);
        } else { uv[0] = u[0]; uv[1] = u[1]; }
        V3 pu0 = LerpV(uv[1], p00, p01), pu1 = LerpV(uv[1], p10, p11);
        V3 pLerp = LerpV(uv[0], pu0, pu1);
        V3 dpdu = pu1 - pu0;
        V3 dpdv = LerpV(uv[0], p01, p11) - LerpV(uv[0], p00, p10);
        if (LengthSquared(dpdu) == 0 || LengthSquared(dpdv) == 0) return ls;
        V3 n = Normalize(Cross(dpdu, dpdv));
        if (q.flip_normal) n = -n;
        V3i pi = V3i::FromValueAndError(pLerp, gamma_(6) * (Abs(p00) + Abs(p01) + Abs(p10) + Abs(p11)));   // shapes.cpp:1200-1213
        V3 p = pi.Mid(), pError = pi.Error();
        float pdfA = pdf / Length(Cross(dpdu, dpdv));
        V3 wi = p - ctx.p;
        if (LengthSquared(wi) == 0) return ls;
        wi = Normalize(wi);
        float pdfW = pdfA / (AbsDotN(n, -wi) / DistanceSquared(ctx.p, p));
        if (std::isinf(pdfW)) return ls;
        if (pdfW == 0) return ls;
        S3 Le = sc->AreaL(light, n, -wi);
        if (!Le) return ls;
        ls.valid = true; ls.L = Le; ls.wi = wi; ls.pdf = pdfW; ls.pLight = p; ls.pLightErr = pError; ls.nLight = n;
        return ls;
    }
    // lights.cpp:820-824 -> shapes.cpp:1329-1347 (area-sampling branch), evaluated at the actual hit
    // Light::PDF_Li -> Shape::PDF(ctx, wi): every shape first intersects ITSELF with ctx.SpawnRay(wi) (shapes.cpp:1337-1341,
    // shapes.h:1133-1137, 366-370) -- not the path's own hit, whose ray may have been re-spawned at interfaces on the way.
    float AreaPDF_Li(int light, const LightSampleCtx &ctx, V3 wi, const SurfaceIntr &) const {
        const vspg_light &l = sc->d->lights[light];
        Hit sh = sc->IntersectPrim(l.shape_kind, l.shape_index, OffsetRayOrigin(ctx.p, ctx.pError, ctx.n, wi), wi, Infinity);
        if (!sh.valid) return 0;
        const SurfaceIntr hit = sc->MakeIntr(sh, wi, -1);
        if (l.shape_kind == 1) {   // shapes.h:1128-1173 Triangle::PDF(ctx, wi)
            const vspg_triangle &t = sc->d->triangles[l.shape_index];
            V3 p0(t.p0), p1(t.p1), p2(t.p2);
            float solidAngle = SphericalTriangleArea(Normalize(p0 - ctx.p), Normalize(p1 - ctx.p), Normalize(p2 - ctx.p));
            if (solidAngle < MinSphericalSampleAreaTri || solidAngle > MaxSphericalSampleAreaTri) {
                float r = (1 / sc->triArea[l.shape_index]) / (AbsDotN(hit.n, -wi) / DistanceSquared(ctx.p, hit.p));
                return std::isinf(r) ? 0 : r;
            }
            float pdf = 1 / solidAngle;
            if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
                V3 v[3] = {p0, p1, p2};
                float uu[2];
                InvertSphericalTriangleSample(v, ctx.p, wi, uu);
                V3 wv[3] = {Normalize(p0 - ctx.p), Normalize(p1 - ctx.p), Normalize(p2 - ctx.p)};
                float w[4] = {std::max<float>(0.01, AbsDotN(ctx.ns, wv[1])), std::max<float>(0.01, AbsDotN(ctx.ns, wv[1])),
                              std::max<float>(0.01, AbsDotN(ctx.ns, wv[0])), std::max<float>(0.01, AbsDotN(ctx.ns, wv[2]))};
                pdf *= BilinearPDF(uu, w);
            }
            return pdf;
        }
        if (l.shape_kind == 2) {   // shapes.h:364-392 Sphere::PDF(ctx, wi)
            const vspg_sphere &sp = sc->d->spheres[l.shape_index];
            V3 pCenter(sp.center);
            if (DistanceSquared(ctx.p, pCenter) <= Sqr(sp.radius)) {
                float r = (1 / sc->sphereArea[l.shape_index]) / (AbsDotN(hit.n, -wi) / DistanceSquared(ctx.p, hit.p));
                return std::isinf(r) ? 0 : r;
            }
            float sin2ThetaMax = sp.radius * sp.radius / DistanceSquared(ctx.p, pCenter);
            float cosThetaMax = SafeSqrt(1 - sin2ThetaMax);
            float oneMinusCosThetaMax = 1 - cosThetaMax;
            if (sin2ThetaMax < 0.00068523f) oneMinusCosThetaMax = sin2ThetaMax / 2;
            return 1 / (2 * Pi * oneMinusCosThetaMax);
        }
        const vspg_quad &q = sc->d->quads[l.shape_index];
        V3 p00(q.p00), p10(q.p10), p01(q.p01), p11(q.p11);
        float pdf = 1;
        V3 v00 = Normalize(p00 - ctx.p), v10 = Normalize(p10 - ctx.p), v01 = Normalize(p01 - ctx.p), v11 = Normalize(p11 - ctx.p);
        if (Scene::IsRectangle(p00, p10, p01, p11) && !(SphericalQuadArea(v00, v10, v11, v01) <= MinSphericalSampleAreaQuad)) {
            // shapes.cpp:1352-1368 PDF for a sample in the spherical rectangle
            float sp = 1 / SphericalQuadArea(v00, v10, v11, v01);
            if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
                float w[4] = {std::max<float>(0.01, AbsDotN(v00, ctx.ns)), std::max<float>(0.01, AbsDotN(v10, ctx.ns)),
                              std::max<float>(0.01, AbsDotN(v01, ctx.ns)), std::max<float>(0.01, AbsDotN(v11, ctx.ns))};
                float uu[2];
                InvertSphericalRectangleSample(ctx.p, p00, p10 - p00, p01 - p00, hit.p, uu);
                return BilinearPDF(uu, w) * sp;
            }
            return sp;
        }
        if (!Scene::IsRectangle(p00, p10, p01, p11)) {
            float w[4] = {Length(Cross(p10 - p00, p01 - p00)), Length(Cross(p10 - p00, p11 - p10)),
                          Length(Cross(p01 - p00, p11 - p01)), Length(Cross(p11 - p10, p11 - p01))};
            pdf = BilinearPDF(hit.uv, w);
        }
        V3 pu0 = LerpV(hit.uv[1], p00, p01), pu1 = LerpV(hit.uv[1], p10, p11);
        V3 dpdu = pu1 - pu0;
        V3 dpdv = LerpV(hit.uv[0], p01, p11) - LerpV(hit.uv[0], p00, p10);
        float pdfA = pdf / Length(Cross(dpdu, dpdv));
        float r = pdfA * (DistanceSquared(ctx.p, hit.p) / AbsDotN(hit.n, -wi));
        return std::isinf(r) ? 0 : r;
    }

    // ---------------------------------------------------------------- SampleLd  vspg.cpp:1136-1252
    S3 SampleLd(bool surface, const SurfaceIntr *si, V3 pMed, V3 woMed, int medium, GuidedBSDF *gb, GuidedPhase *gp, Sampler &sampler,
                S3 r_p, int ch) {
        LightSampleCtx ctx;
        V3 p, pErr, n, wo;
        if (surface) {
            ctx.p = si->p; ctx.pError = si->pError; ctx.n = si->n; ctx.ns = si->ns;
            // reflective-only BSDF: ctx.pi = intr.OffsetRayOrigin(intr.wo)  (vspg.cpp:1147-1149)
            ctx.p = OffsetRayOrigin(si->p, si->pError, si->n, si->wo); ctx.pError = V3();
            p = si->p; pErr = si->pError; n = si->n; wo = si->wo;
        } else {
            ctx.p = pMed; p = pMed; wo = woMed;
        }
        float u = sampler.Get1D();
        float pmf = 0;
        int light = SampleLight(u, &pmf, ctx.p, ctx.ns);
        float uLight[2]; sampler.Get2D(uLight);
        if (light < 0) return S3(0.f);
        LiSample ls = SampleLi(light, ctx, uLight);
        if (getenv("ORC_TRACE")) fprintf(stderr, "SampleLd surface=%d ctx.p=%.9g %.9g %.9g n=%.9g %.9g %.9g ns=%.9g %.9g %.9g light=%d u=%.9g %.9g valid=%d pdf=%.9g wi=%.9g %.9g %.9g pL=%.9g %.9g %.9g L=%g %g %g\n", (int)surface, ctx.p.x, ctx.p.y, ctx.p.z, ctx.n.x, ctx.n.y, ctx.n.z, ctx.ns.x, ctx.ns.y, ctx.ns.z, light, uLight[0], uLight[1], (int)ls.valid, ls.pdf, ls.wi.x, ls.wi.y, ls.wi.z, ls.pLight.x, ls.pLight.y, ls.pLight.z, ls.L[0], ls.L[1], ls.L[2]);
        if (!ls.valid || !ls.L || ls.pdf == 0) return S3(0.f);
        float p_l = pmf * ls.pdf;
        float scatterPDF; S3 f_hat;
        V3 wi = ls.wi;
        if (surface) { f_hat = gb->f(wo, wi) * AbsDotN(wi, si->ns); scatterPDF = 1.0f * gb->PDF(wo, wi); }
        else { f_hat = S3(gp->p(wo, wi)); scatterPDF = 1.0f * gp->PDF(wo, wi); }
        if (!f_hat) return S3(0.f);
        // Ray lightRay = intr.SpawnRayTo(ls->pLight)   (interaction.h:105-109 -> ray.h:104-109, both ends offset)
        auto spawnTo = [&](V3 pf_, V3 pfErr, V3 nf, int curMediumIfNoIface, const SurfaceIntr *from, Ray *out) {
            V3 pf = OffsetRayOrigin(pf_, pfErr, nf, ls.pLight - pf_);
            V3 pt = OffsetRayOrigin(ls.pLight, ls.pLightErr, ls.nLight, pf - ls.pLight);
            out->o = pf; out->d = pt - pf;
            out->medium = from ? from->GetMedium(out->d) : curMediumIfNoIface;
        };
        Ray lightRay;
        spawnTo(p, pErr, n, medium, surface ? si : nullptr, &lightRay);
        S3 T_ray(1.f), r_l(1.f), r_u(1.f);
        RNG rng(HashV3(lightRay.o), HashV3(lightRay.d));
        cnt.shadow++;
        while (lightRay.d != V3(0, 0, 0)) {
            Hit h = sc->Intersect(lightRay.o, lightRay.d, 1 - ShadowEpsilon);
            if (getenv("ORC_TRACE")) fprintf(stderr, "  lightRay o=%.9g %.9g %.9g d=%.9g %.9g %.9g medium=%d hit=%d t=%.9g mat=%d f_hat=%g scatterPDF=%g\n", lightRay.o.x, lightRay.o.y, lightRay.o.z, lightRay.d.x, lightRay.d.y, lightRay.d.z, lightRay.medium, (int)h.valid, h.valid ? h.t : -1.f, h.valid ? sc->HitMaterial(h) : -9, f_hat[0], scatterPDF);
            if (h.valid && sc->HitMaterial(h) >= 0) return S3(0.f);
            if (lightRay.medium >= 0) {
                float tMax = h.valid ? h.t : (1 - ShadowEpsilon);
                float uu = rng.Uniform();
                TrackStats st;
                S3 T_maj = SampleT_maj(sc->media[lightRay.medium], lightRay, tMax, uu, rng, ch, &st,
                                       [&](V3, const MediumProps &mp, S3 sigma_maj, S3 T_maj) {
                                           S3 sigma_n = ClampZero(sigma_maj - mp.sigma_a - mp.sigma_s);
                                           float pdf = T_maj[ch] * sigma_maj[ch];
                                           T_ray *= T_maj * sigma_n / pdf;
                                           r_l *= T_maj * sigma_maj / pdf;
                                           r_u *= T_maj * sigma_n / pdf;
                                           S3 Tr = T_ray / (r_l + r_u).Average();
                                           if (Tr.MaxComponentValue() < 0.05f) {
                                               float q = 0.75f;
                                               if (rng.Uniform() < q) T_ray = S3(0.f);
                                               else T_ray /= 1 - q;
                                           }
                                           if (!T_ray) return false;
                                           return true;
                                       });
                cnt.queries += st.queries; cnt.dda += st.dda;
                T_ray *= T_maj / T_maj[ch];
                r_l *= T_maj / T_maj[ch];
                r_u *= T_maj / T_maj[ch];
            }
            if (!T_ray) return S3(0.f);
            if (!h.valid) break;
            SurfaceIntr hs = sc->MakeIntr(h, lightRay.d, lightRay.medium);
            spawnTo(hs.p, hs.pError, hs.n, lightRay.medium, &hs, &lightRay);
        }
        r_l *= r_p * p_l;
        r_u *= r_p * scatterPDF;
        if (ls.delta) return f_hat * T_ray * ls.L / r_l.Average();
        else return f_hat * T_ray * ls.L / (r_l + r_u).Average();
    }

    struct PathState {
        S3 L{0.f}, beta{1.f}, r_u{1.f}, r_l{1.f};
        bool specularBounce = false, anyNonSpecularBounces = false, lastVertexVolume = false;
        int depth = 0;
        float etaScale = 1, rr_correction = 1;
        LightSampleCtx prevCtx;
        Ray ray;
    };

    // One distance-sampling decision (vspg.cpp:637-800 up to the chosen event).  Returns through `ev`.
    struct DistanceEvent { bool scattered = false, terminated = false; V3 p; MediumProps mp; int queries = 0, dda = 0; };

    // vspg.cpp:804-875 / 988-1058: the scatter-vertex part shared by both routines
    void ScatterVertex(PathState &ps, int px, int py, V3 p, const MediumProps &mp, Sampler &sampler, GuidedPhase &gphase, int ch,
                       bool &scattered, bool &terminated) {
        if (ps.beta && ps.r_u) {
            float v = sampler.Get1D();
            int before = gphase.lookups;
            gphase.init(mp.g, p, ps.ray.d, v);
            cnt.lookups += gphase.lookups - before;
            float survivalProb = 1.0f;
            if (ps.depth > sc->ip.min_rr_depth) {
                S3 rrw = (ps.beta / ps.r_u.Average()) * ps.rr_correction;
                survivalProb = ps.specularBounce ? 0.95f : std::max(0.f, std::min(1.f, rrw.MaxComponentValue()));
            }
            if (sc->ip.use_nee) {
                S3 Ld = SampleLd(false, nullptr, p, -ps.ray.d, ps.ray.medium, nullptr, &gphase, sampler, ps.r_u, ch);
                ps.L += ps.beta * Ld;
            }
            if (survivalProb < 1 && ps.depth > sc->ip.min_rr_depth) {
                float q = std::max<float>(0, 1 - survivalProb);
                if (sampler.Get1D() < q) { terminated = true; return; }
                ps.beta /= 1 - q;
            }
            float u[2]; sampler.Get2D(u);
            GuidedPhase::Sample s = gphase.Sample_p(-ps.ray.d, u);
            if (!s.valid || s.pdf == 0) { terminated = true; return; }
            float phaseFunctionWeight = s.p / s.pdf;
            ps.beta *= phaseFunctionWeight;
            ps.r_l = ps.r_u / s.pdf;
            ps.prevCtx = LightSampleCtx{p, V3(), V3(), V3()};
            scattered = true;
            ps.ray.o = p; ps.ray.d = s.wi;
            ps.specularBounce = false; ps.anyNonSpecularBounces = true;
            ps.lastVertexVolume = true;
        }
    }

    // vspg.cpp:637-1096.  `probe`: stop right after the event decision (used by the per-kernel parity hook).
    void SampleDistance(PathState &ps, int px, int py, float tMax, Sampler &sampler, RNG &rng, int ch, GuidedBSDF &gbsdf,
                        GuidedPhase &gphase, bool &scattered, bool &terminated, bool probe, const float *vspOverride,
                        DistanceEvent *ev) {
        const vspg_integrator_params &ip = sc->ip;
        const Medium &med = sc->media[ps.ray.medium];
        bool guideScatterDecision = false;
        float vsp = -1.f;
        auto validate = [&](float v) { guideScatterDecision = !(std::isnan(v) || v < 0.f || v > 1.f); return v; };
        if (ps.depth == 0) {
            if (ip.vsp_guiding && ip.vsp_primary_guiding) {  // vspg.cpp:1098-1112
                if (isgRec) vsp = isgTrained ? isgTrained[(size_t)py * sc->d->camera.width + px] : 0.5f;   // trained on the fly: 0.5 until IsReady()
                else if (sc->g && sc->g->isg_vsp) vsp = sc->g->isg_vsp[(size_t)py * sc->g->isg_width + px];
                else vsp = 0.5f;
                validate(vsp);
            }
        } else if (ip.vsp_guiding && ip.vsp_secondary_guiding) {
            if (vspOverride) vsp = *vspOverride;
            else vsp = ps.lastVertexVolume ? gphase.VolumeScatterProbability(ps.ray.d) : gbsdf.VolumeScatterProbability(ps.ray.d);
            validate(vsp);
        }
        if (guideScatterDecision) vsp = std::max(std::min(vsp, 0.999f), 0.001f);
        TrackStats st;
        bool use_resampling = !ip.vsp_sampling_nds && !med.IsHomogeneous();
        if (use_resampling) {
            float volumeRatioCompensated = 0, majorantScale = 1;
            float weightSum = 0;
            S3 trRatioEst(1.f), beta_rs(1.f), r_u_rs(1.f);
            struct Cand { V3 p; MediumProps mp; float wi = 0, sigmaTTrEst = 0; S3 num{0.f}, den{0.f}; } sel;
            S3 T_maj = SampleT_maj_Resampling(
                med, ps.ray, tMax, sampler.Get1D(), rng, ch, guideScatterDecision, vsp, volumeRatioCompensated, majorantScale, &st,
                [&](V3 p, const MediumProps &mp, S3 sigma_maj, S3 T_maj) {
                    S3 sigma_t = mp.sigma_s + mp.sigma_a;
                    S3 sigma_n = ClampZero(sigma_maj - sigma_t);
                    float wi = (sigma_t / sigma_maj * trRatioEst)[ch];
                    if (wi > 0) {
                        weightSum += wi;
                        if (sampler.Get1D() < wi / weightSum) {
                            float pdf = T_maj[ch] * sigma_t[ch];
                            sel.p = p; sel.mp = mp; sel.wi = wi; sel.sigmaTTrEst = wi;
                            sel.num = beta_rs * T_maj * mp.sigma_s / pdf;
                            sel.den = r_u_rs * T_maj * sigma_t / pdf;
                        }
                    }
                    float pdf = T_maj[ch] * sigma_n[ch];
                    beta_rs *= T_maj * sigma_n / pdf;
                    r_u_rs *= T_maj * sigma_n / pdf;
                    trRatioEst *= sigma_n / sigma_maj;
                    return true;
                });
            beta_rs *= T_maj / T_maj[ch];
            r_u_rs *= T_maj / T_maj[ch];
            cnt.queries += st.queries; cnt.dda += st.dda;
            if (ev) { ev->queries = (int)st.queries; ev->dda = (int)st.dda; }
            if (ps.depth == 0 && trBuf) TrAddSample(px, py, trRatioEst);
            float trS = trRatioEst[ch];
            Cand surf; surf.p = ps.ray.at(tMax); surf.wi = trS; surf.sigmaTTrEst = trS; surf.num = beta_rs; surf.den = r_u_rs;
            if (guideScatterDecision && trS < 1 && trS > 0 && weightSum > 0) {
                float volRatio = volumeRatioCompensated * ip.vsp_mis_ratio + (1 - trS) * (1 - ip.vsp_mis_ratio);
                float surfRatio = 1 - volRatio;
                surf.wi = surfRatio / volRatio * weightSum;
            }
            weightSum += surf.wi;
            bool selectSurface = false;
            if (weightSum == 0) return;
            else if (sampler.Get1D() < surf.wi / weightSum) { sel = surf; selectSurface = true; }
            float rf = weightSum * sel.sigmaTTrEst / sel.wi;
            if (selectSurface) {
                ps.beta *= sel.num * rf;
                ps.r_u *= sel.den;
                if (ps.beta.HasNaNs() || ps.r_u.HasNaNs() || YIsInf(ps.beta) || YIsInf(ps.r_u)) { terminated = true; return; }
            } else {
                if (ps.depth == 0) firstVolume = true;   // vspg.cpp:776-783 isgbSample.SetSurfaceEvent(false)
                if (ps.depth++ >= ip.max_depth) { terminated = true; return; }
                ps.beta *= sel.num * rf;
                ps.r_u *= sel.den;
                if (ps.beta.HasNaNs() || ps.r_u.HasNaNs() || YIsInf(ps.beta) || YIsInf(ps.r_u)) { terminated = true; return; }
                if (ev) { ev->scattered = true; ev->p = sel.p; ev->mp = sel.mp; }
                if (probe) { scattered = true; return; }
                ScatterVertex(ps, px, py, sel.p, sel.mp, sampler, gphase, ch, scattered, terminated);
            }
        } else {
            S3 beta_factor(1.f), r_u_factor(1.f);
            bool trLoaded = sc->g && sc->g->tr_rgb;
            S3 T_maj = SampleT_maj_OpticalDepthSpace(
                med, ps.ray, tMax, sampler.Get1D(), rng, ch, guideScatterDecision, vsp, ip.vsp_mis_ratio, ip.vsp_sampling_nds != 0,
                beta_factor, r_u_factor, &st, [&](V3 p, const MediumProps &mp, S3 sigma_maj, S3 T_maj, bool activateNDS) {
                    if (!ps.beta) { terminated = true; return false; }
                    if (ps.depth < ip.max_depth && mp.Le) {   // :895-906 emission from the medium at the tentative collision
                        float pdf = sigma_maj[ch] * T_maj[ch];
                        S3 betap = ps.beta * T_maj / pdf;
                        S3 r_e = ps.r_u * sigma_maj * T_maj / pdf;
                        if (r_e) ps.L += betap * mp.sigma_a * mp.Le / r_e.Average();
                    }
                    S3 sigma_t = mp.sigma_s + mp.sigma_a;
                    float pScatter = sigma_t[ch] / sigma_maj[ch];
                    bool NDS_plus = false;
                    if (ps.depth == 0 && ip.vsp_sampling_nds && ip.collision_probability_bias && trLoaded && activateNDS) {
                        NDS_plus = true;
                        float trEstCache = sc->g->tr_rgb[3 * ((size_t)py * sc->g->tr_width + px) + ch];
                        float gamma = 1 + trEstCache;
                        pScatter = vspg_powf(pScatter, 1 / gamma);
                    }
                    float pNull = std::max<float>(0, 1 - pScatter);
                    float um = rng.Uniform();
                    // SampleDiscrete({pScatter, pNull}, um)  util/sampling.h:79-112
                    int mode;
                    {
                        float w[2] = {pScatter, pNull};
                        float sumW = w[0] + w[1];
                        float up = um * sumW;
                        if (up == sumW) up = NextFloatDown(up);
                        int offset = 0; float sum = 0;
                        while (sum + w[offset] <= up) { sum += w[offset++]; if (offset > 1) { offset = 1; break; } }
                        mode = offset;
                    }
                    if (mode == 0) {
                        if (ps.depth == 0) firstVolume = true;   // vspg.cpp:950-957
                        if (ps.depth++ >= ip.max_depth) { terminated = true; return false; }
                        float pdf = T_maj[ch] * sigma_t[ch];
                        ps.beta *= T_maj * mp.sigma_s / pdf;
                        ps.r_u *= T_maj * sigma_t / pdf;
                        if (NDS_plus) ps.r_u *= sigma_maj * pScatter / sigma_t;
                        ps.beta *= beta_factor;
                        ps.r_u *= r_u_factor;
                        if (ev) { ev->scattered = true; ev->p = p; ev->mp = mp; }
                        if (probe) { scattered = true; return false; }
                        ScatterVertex(ps, px, py, p, mp, sampler, gphase, ch, scattered, terminated);
                        return false;
                    } else {
                        S3 sigma_n = ClampZero(sigma_maj - mp.sigma_a - mp.sigma_s);
                        float pdf = T_maj[ch] * sigma_n[ch];
                        ps.beta *= T_maj * sigma_n / pdf;
                        if (pdf == 0) ps.beta = S3(0.f);
                        ps.r_u *= T_maj * sigma_n / pdf;
                        if (NDS_plus) ps.r_u *= sigma_maj * (1 - pScatter) / sigma_n;
                        ps.r_l *= T_maj * sigma_maj / pdf;
                        return (bool)ps.beta && (bool)ps.r_u;
                    }
                });
            bool multiply_T_maj = !(scattered || terminated || !ps.beta || !ps.r_u);
            if (probe && ev && ev->scattered) multiply_T_maj = false;
            if (multiply_T_maj) {
                ps.beta *= T_maj / T_maj[ch];
                ps.r_u *= T_maj / T_maj[ch];
                ps.r_l *= T_maj / T_maj[ch];
                ps.beta *= beta_factor;
                ps.r_u *= r_u_factor;
                ps.r_l *= r_u_factor;
            }
            cnt.queries += st.queries; cnt.dda += st.dda;
            if (ev) { ev->queries = (int)st.queries; ev->dda = (int)st.dda; }
        }
    }

    // vspg.cpp:262-635
    S3 Li(int px, int py, Ray ray, Sampler &sampler, int ch) {
        const vspg_integrator_params &ip = sc->ip;
        GuidedBSDF gbsdf; gbsdf.sc = sc; gbsdf.sampler = &sampler; gbsdf.enableGuiding = ip.surface_guiding;
        gbsdf.enableScatterGuiding = ip.vsp_secondary_guiding; gbsdf.ris = ip.surface_guiding_ris;
        GuidedPhase gphase; gphase.sc = sc; gphase.sampler = &sampler; gphase.enableGuiding = ip.volume_guiding;
        gphase.enableScatterGuiding = ip.vsp_secondary_guiding; gphase.ris = ip.volume_guiding_ris;
        PathState ps; ps.ray = ray;
        while (true) {
            Hit h = sc->Intersect(ps.ray.o, ps.ray.d, Infinity);
            float tMax = h.valid ? h.t : Infinity;
            if (ps.ray.medium >= 0 && !std::isinf(tMax)) {
                bool scattered = false, terminated = false;
                uint64_t hash0 = HashFloat1(sampler.Get1D());
                uint64_t hash1 = HashFloat1(sampler.Get1D());
                RNG rng(hash0, hash1);
                SampleDistance(ps, px, py, tMax, sampler, rng, ch, gbsdf, gphase, scattered, terminated, false, nullptr, nullptr);
                if (terminated || !ps.beta || !ps.r_u) break;
                if (scattered) continue;
            }
            if (!h.valid) {
                for (int i = 0; i < sc->d->n_lights; i++) {
                    const vspg_light &l = sc->d->lights[i];
                    // infiniteLights holds LightType::Infinite AND LightType::DeltaDirection lights in this fork (cpu/integrators.h:77-81);
                    // DistantLight::Le returns scale * Lemit (lights.h:290-293) and only counts at depth 0 (vspg.cpp:356-358)
                    if (l.type != VSPG_LIGHT_UNIFORM_INFINITE && l.type != VSPG_LIGHT_DISTANT) continue;
                    S3 Le = Scene::LightL(l);
                    if (l.type == VSPG_LIGHT_DISTANT && ps.depth != 0) Le = S3(0.f);
                    if (ps.depth == 0 || ps.specularBounce) ps.L += ps.beta * Le / ps.r_u.Average();
                    else {
                        float lightPDF = LightPMF(i, ps.prevCtx) * 0.f;  // PDF_Li(..., allowIncompletePDF=true) = 0 (lights.cpp:1031-1036)
                        ps.r_l *= lightPDF;
                        float w_b = ip.use_nee ? 1.0f / (ps.r_u + ps.r_l).Average() : 1.f;
                        ps.L += ps.beta * w_b * Le;
                    }
                }
                break;
            }
            SurfaceIntr isect = sc->MakeIntr(h, ps.ray.d, ps.ray.medium);
            if (isect.areaLight >= 0) {
                S3 Le = sc->AreaL(isect.areaLight, isect.n, -ps.ray.d);
                if (Le) {
                    if (ps.depth == 0 || ps.specularBounce) ps.L += ps.beta * Le / ps.r_u.Average();
                    else {
                        float lightPDF = LightPMF(isect.areaLight, ps.prevCtx) * AreaPDF_Li(isect.areaLight, ps.prevCtx, ps.ray.d, isect);
                        ps.r_l *= lightPDF;
                        float w_l = ip.use_nee ? 1.0f / (ps.r_u + ps.r_l).Average() : 1.0f;
                        ps.L += ps.beta * w_l * Le;
                    }
                }
            }
            if (isect.material < 0) {  // interface: SkipIntersection (interaction.cpp:91-97)
                V3 d = ps.ray.d;
                ps.ray.o = OffsetRayOrigin(isect.p, isect.pError, isect.n, d);
                ps.ray.medium = isect.GetMedium(d);
                continue;
            }
            if (ps.depth++ >= ip.max_depth) break;
            const vspg_material &mat = sc->d->materials[isect.material];
            gbsdf.setBSDF(isect.ns, isect.dpdu, S3(mat.reflectance));
            float v = sampler.Get1D();
            int before = gbsdf.lookups;
            gbsdf.init(ps.ray.o + h.t * ps.ray.d, ps.ray.d, isect.ns, v);
            cnt.lookups += gbsdf.lookups - before;
            if (ip.use_nee && gbsdf.IsNonSpecular()) {
                S3 Ld = SampleLd(true, &isect, V3(), V3(), ps.ray.medium, &gbsdf, nullptr, sampler, ps.r_u, ch);
                ps.L += ps.beta * Ld;
            }
            ps.prevCtx = LightSampleCtx{isect.p, isect.pError, isect.n, isect.ns};
            V3 wo = -ps.ray.d;
            float u = sampler.Get1D();
            float u2[2]; sampler.Get2D(u2);
            GuidedBSDF::Sample bs = gbsdf.Sample_f(wo, u, u2);
            if (!bs.valid) break;
            ps.lastVertexVolume = false;
            ps.rr_correction *= bs.pdf / bs.bsdfPdf;
            S3 bsdfWeight = bs.f * AbsDotN(bs.wi, isect.ns) / bs.pdf;
            ps.beta *= bsdfWeight;
            ps.r_l = ps.r_u / bs.misPdf;
            ps.specularBounce = false; ps.anyNonSpecularBounces = true;
            ps.ray.o = OffsetRayOrigin(isect.p, isect.pError, isect.n, bs.wi);
            ps.ray.d = bs.wi;
            ps.ray.medium = isect.GetMedium(bs.wi);
            if (!ps.beta) break;
            float survivalProb = 1.f;
            if (ps.depth > ip.min_rr_depth) {
                S3 rrw = (ps.beta / ps.r_u.Average()) * ps.rr_correction * ps.etaScale;
                survivalProb = ps.specularBounce ? 0.95f : std::max(0.f, std::min(1.f, rrw.MaxComponentValue()));
            }
            if (survivalProb < 1 && ps.depth > ip.min_rr_depth) {
                float q = std::max<float>(0, 1 - survivalProb);
                if (sampler.Get1D() < q) break;
                ps.beta /= 1 - q;
            }
        }
        return ps.L;
    }

    // cpu/integrators.cpp:272-334 + samplers.h:797-814 + cameras.cpp:410-437
    bool CameraRay(int px, int py, Sampler &sampler, Ray *ray, int *ch, float *filterWeight) const {
        const vspg_camera &c = sc->d->camera;
        float lu = sampler.Get1D();
        *ch = std::min<int>((int)std::floor(lu * 3), 2);  // util/spectrum.h:335-339
        float uf[2]; sampler.Get2D(uf);
        // BoxFilter::Sample (filters.h): p = Lerp(u, -radius, radius), weight 1
        float fx = Lerp(uf[0], -c.filter_radius[0], c.filter_radius[0]), fy = Lerp(uf[1], -c.filter_radius[1], c.filter_radius[1]);
        *filterWeight = 1.f;
        if (c.filter_type == 1) { float fp[2]; sc->gaussFilter.Sample(uf, fp, filterWeight); fx = fp[0]; fy = fp[1]; }
        float pFilmX = px + fx + 0.5f, pFilmY = py + fy + 0.5f;
        (void)sampler.Get1D();  // time
        float ul[2]; sampler.Get2D(ul);
        // cameraFromRaster(pFilm): util/transform.h operator()(Point3) with w divide
        const float(*m)[4] = c.camera_from_raster.m;
        float x = pFilmX, y = pFilmY, z = 0;
        float xp = m[0][0] * x + m[0][1] * y + m[0][2] * z + m[0][3];
        float yp = m[1][0] * x + m[1][1] * y + m[1][2] * z + m[1][3];
        float zp = m[2][0] * x + m[2][1] * y + m[2][2] * z + m[2][3];
        float wp = m[3][0] * x + m[3][1] * y + m[3][2] * z + m[3][3];
        V3 pCamera = (wp == 1) ? V3(xp, yp, zp) : V3(xp, yp, zp) / wp;
        V3 o(0, 0, 0), d = Normalize(pCamera);
        if (c.lens_radius > 0) {
            float pl[2]; SampleUniformDiskConcentric(ul, pl);
            pl[0] *= c.lens_radius; pl[1] *= c.lens_radius;
            float ft = c.focal_distance / d.z;
            V3 pFocus = o + d * ft;
            o = V3(pl[0], pl[1], 0);
            d = Normalize(pFocus - o);
        }
        // CameraTransform::RenderFromCamera(ray): point and vector transforms (no error offsetting: Transform::operator()(Ray)
        // does offset the origin by its error bound; the camera origin (0,0,0) maps with zero error for pinhole cameras)
        const float(*r)[4] = c.render_from_camera.m;
        V3 ro((r[0][0] * o.x + r[0][1] * o.y) + (r[0][2] * o.z + r[0][3]), (r[1][0] * o.x + r[1][1] * o.y) + (r[1][2] * o.z + r[1][3]),
              (r[2][0] * o.x + r[2][1] * o.y) + (r[2][2] * o.z + r[2][3]));
        V3 rd(r[0][0] * d.x + r[0][1] * d.y + r[0][2] * d.z, r[1][0] * d.x + r[1][1] * d.y + r[1][2] * d.z,
              r[2][0] * d.x + r[2][1] * d.y + r[2][2] * d.z);
        ray->o = ro; ray->d = rd; ray->medium = c.medium;
        return true;
    }

    void RenderPixelSample(int px, int py, int sampleIndex, double *film) {
        Sampler sampler;
        sampler.StartPixelSample(px, py, sc->seed, sampleIndex);
        Ray ray; int ch; float fw;
        CameraRay(px, py, sampler, &ray, &ch, &fw);
        firstVolume = false;
        S3 L = Li(px, py, ray, sampler, ch);
        if (isgRec) IsgAddSample(px, py, L);   // vspg.cpp:613-622, inside Li: before the film's NaN guard
        if (L.HasNaNs() || YIsInf(L)) L = S3(0.f);
        cnt.paths++;
        // film.h:251-267 RGBFilm::AddSample
        const vspg_camera &c = sc->d->camera;
        float rgb[3] = {c.imaging_ratio * L[0], c.imaging_ratio * L[1], c.imaging_ratio * L[2]};
        float m = std::max({rgb[0], rgb[1], rgb[2]});
        if (m > c.max_component_value) for (float &v : rgb) v *= c.max_component_value / m;
        double *pix = film + 4 * ((size_t)py * c.width + px);
        for (int k = 0; k < 3; k++) pix[k] += fw * rgb[k];
        pix[3] += fw;
    }
};

}  // namespace orc
