// TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's BVHLightSampler -- construction, Sample, PMF.
//   src/pbrt/lightsamplers.h:100-404, lightsamplers.cpp:107-236   (CompactLightBounds, LightBVHNode, BVHLightSampler)
//   src/pbrt/lights.h:104-152 (LightBounds, Union), lights.cpp:200-205, 845-864 (light Bounds())
//   src/pbrt/util/vecmath.cpp:56-83 (DirectionCone Union), util/vecmath.h:1733-1828 (OctahedralVector, BoundSubtendedDirections)
//   src/pbrt/shapes.cpp:1070-1126 (BilinearPatch Bounds / NormalBounds)
// Nodes keep the QUANTISED CompactLightBounds fields and decode them at every Importance() call, as the reference does.
#pragma once
#include <algorithm>
#include <cstdint>
#include <utility>
#include <vector>

#include "orc_math.h"   // LBox below = pbrt Bounds3f with the empty-box default (orc_media.h has its own Bounds3)

namespace orc {

struct LBox {
    V3 pMin = V3(Infinity, Infinity, Infinity), pMax = V3(-Infinity, -Infinity, -Infinity);
    V3 Diagonal() const { return pMax - pMin; }
    float SurfaceArea() const { V3 d = Diagonal(); return 2 * (d.x * d.y + d.x * d.z + d.y * d.z); }
};
inline LBox Union(const LBox &a, const LBox &b) {
    LBox r;
    for (int k = 0; k < 3; k++) { r.pMin[k] = std::min(a.pMin[k], b.pMin[k]); r.pMax[k] = std::max(a.pMax[k], b.pMax[k]); }
    return r;
}
inline LBox Union(const LBox &a, V3 p) { LBox b; b.pMin = b.pMax = p; return Union(a, b); }

struct DirectionCone {
    V3 w; float cosTheta = Infinity;
    bool IsEmpty() const { return cosTheta == Infinity; }
    static DirectionCone EntireSphere() { DirectionCone c; c.w = V3(0, 0, 1); c.cosTheta = -1; return c; }
};
inline float SafeACos_(float x) { return vspg_acosf(std::min(1.f, std::max(-1.f, x))); }
inline float AngleBetween(V3 v1, V3 v2) {  // vecmath.h:972-977
    if (Dot(v1, v2) < 0) return Pi - 2 * SafeASin(Length(v1 + v2) / 2);
    return 2 * SafeASin(Length(v2 - v1) / 2);
}
// util/transform.cpp Rotate(theta degrees, axis) applied to a vector
inline V3 RotateDeg(float thetaDeg, V3 axis, V3 v) {
    float th = (Pi / 180) * thetaDeg;
    float sinTheta = vspg_sinf(th), cosTheta = vspg_cosf(th);
    V3 a = Normalize(axis);
    float m00 = a.x * a.x + (1 - a.x * a.x) * cosTheta, m01 = a.x * a.y * (1 - cosTheta) - a.z * sinTheta, m02 = a.x * a.z * (1 - cosTheta) + a.y * sinTheta;
    float m10 = a.x * a.y * (1 - cosTheta) + a.z * sinTheta, m11 = a.y * a.y + (1 - a.y * a.y) * cosTheta, m12 = a.y * a.z * (1 - cosTheta) - a.x * sinTheta;
    float m20 = a.x * a.z * (1 - cosTheta) - a.y * sinTheta, m21 = a.y * a.z * (1 - cosTheta) + a.x * sinTheta, m22 = a.z * a.z + (1 - a.z * a.z) * cosTheta;
    return V3(m00 * v.x + m01 * v.y + m02 * v.z, m10 * v.x + m11 * v.y + m12 * v.z, m20 * v.x + m21 * v.y + m22 * v.z);
}
inline DirectionCone Union(const DirectionCone &a, const DirectionCone &b) {  // vecmath.cpp:56-83
    if (a.IsEmpty()) return b;
    if (b.IsEmpty()) return a;
    float theta_a = SafeACos_(a.cosTheta), theta_b = SafeACos_(b.cosTheta);
    float theta_d = AngleBetween(a.w, b.w);
    if (std::min(theta_d + theta_b, Pi) <= theta_a) return a;
    if (std::min(theta_d + theta_a, Pi) <= theta_b) return b;
    float theta_o = (theta_a + theta_d + theta_b) / 2;
    if (theta_o >= Pi) return DirectionCone::EntireSphere();
    float theta_r = theta_o - theta_a;
    V3 wr = Cross(a.w, b.w);
    if (LengthSquared(wr) == 0) return DirectionCone::EntireSphere();
    V3 w = RotateDeg((180 / Pi) * theta_r, wr, a.w);
    DirectionCone r; r.w = Normalize(w); r.cosTheta = vspg_cosf(theta_o);
    return r;
}

struct LightBounds {  // lights.h:104-135
    LBox bounds; float phi = 0; V3 w; float cosTheta_o = 1, cosTheta_e = 1; bool twoSided = false;
    V3 Centroid() const { return (bounds.pMin + bounds.pMax) / 2; }
};
inline LightBounds Union(const LightBounds &a, const LightBounds &b) {  // lights.h:137-152
    if (a.phi == 0) return b;
    if (b.phi == 0) return a;
    DirectionCone ca; ca.w = Normalize(a.w); ca.cosTheta = a.cosTheta_o;
    DirectionCone cb; cb.w = Normalize(b.w); cb.cosTheta = b.cosTheta_o;
    DirectionCone cone = Union(ca, cb);
    LightBounds r;
    r.bounds = Union(a.bounds, b.bounds); r.w = Normalize(cone.w); r.phi = a.phi + b.phi;
    r.cosTheta_o = cone.cosTheta; r.cosTheta_e = std::min(a.cosTheta_e, b.cosTheta_e); r.twoSided = a.twoSided | b.twoSided;
    return r;
}

// lightsamplers.h:100-228
struct CompactLightBounds {
    uint16_t wx = 0, wy = 0;            // OctahedralVector
    float phi = 0;
    unsigned qCosTheta_o = 0, qCosTheta_e = 0; bool twoSided = false;
    uint16_t qb[2][3] = {{0, 0, 0}, {0, 0, 0}};

    static float Sign(float v) { return std::copysign(1.f, v); }
    static uint16_t Encode(float f) { return (uint16_t)std::round(std::min(1.f, std::max(0.f, (f + 1) / 2)) * 65535.f); }
    static float QuantizeBounds(float c, float mn, float mx) {
        if (mn == mx) return 0;
        return 65535.f * std::min(1.f, std::max(0.f, (c - mn) / (mx - mn)));
    }
    CompactLightBounds() = default;
    CompactLightBounds(const LightBounds &lb, const LBox &allb) {
        V3 v = Normalize(lb.w);
        v = v / (std::abs(v.x) + std::abs(v.y) + std::abs(v.z));   // vecmath.h:1738-1748
        if (v.z >= 0) { wx = Encode(v.x); wy = Encode(v.y); }
        else { wx = Encode((1 - std::abs(v.y)) * Sign(v.x)); wy = Encode((1 - std::abs(v.x)) * Sign(v.y)); }
        phi = lb.phi;
        qCosTheta_o = (unsigned)std::floor(32767.f * ((lb.cosTheta_o + 1) / 2));
        qCosTheta_e = (unsigned)std::floor(32767.f * ((lb.cosTheta_e + 1) / 2));
        twoSided = lb.twoSided;
        for (int c = 0; c < 3; ++c) {
            qb[0][c] = (uint16_t)std::floor(QuantizeBounds(lb.bounds.pMin[c], allb.pMin[c], allb.pMax[c]));
            qb[1][c] = (uint16_t)std::ceil(QuantizeBounds(lb.bounds.pMax[c], allb.pMin[c], allb.pMax[c]));
        }
    }
    V3 W() const {   // vecmath.h:1751-1764
        V3 v;
        v.x = -1 + 2 * (wx / 65535.f);
        v.y = -1 + 2 * (wy / 65535.f);
        v.z = 1 - (std::abs(v.x) + std::abs(v.y));
        if (v.z < 0) { float xo = v.x; v.x = (1 - std::abs(v.y)) * Sign(xo); v.y = (1 - std::abs(xo)) * Sign(v.y); }
        return Normalize(v);
    }
    float CosTheta_o() const { return 2 * (qCosTheta_o / 32767.f) - 1; }
    float CosTheta_e() const { return 2 * (qCosTheta_e / 32767.f) - 1; }
    LBox Bounds(const LBox &allb) const {
        LBox b;
        for (int c = 0; c < 3; c++) {
            b.pMin[c] = Lerp(qb[0][c] / 65535.f, allb.pMin[c], allb.pMax[c]);
            b.pMax[c] = Lerp(qb[1][c] / 65535.f, allb.pMin[c], allb.pMax[c]);
        }
        return b;
    }
    float Importance(V3 p, V3 n, const LBox &allb) const {   // lightsamplers.h:140-201
        LBox bounds = Bounds(allb);
        float cosTheta_o = CosTheta_o(), cosTheta_e = CosTheta_e();
        V3 pc = (bounds.pMin + bounds.pMax) / 2;
        float d2 = DistanceSquared(p, pc);
        d2 = std::max(d2, Length(bounds.Diagonal()) / 2);
        auto cosSubClamped = [](float sinTheta_a, float cosTheta_a, float sinTheta_b, float cosTheta_b) -> float {
            if (cosTheta_a > cosTheta_b) return 1;
            return cosTheta_a * cosTheta_b + sinTheta_a * sinTheta_b;
        };
        auto sinSubClamped = [](float sinTheta_a, float cosTheta_a, float sinTheta_b, float cosTheta_b) -> float {
            if (cosTheta_a > cosTheta_b) return 0;
            return sinTheta_a * cosTheta_b - cosTheta_a * sinTheta_b;
        };
        V3 wi = Normalize(p - pc);
        float cosTheta_w = Dot(W(), wi);
        if (twoSided) cosTheta_w = std::abs(cosTheta_w);
        float sinTheta_w = SafeSqrt(1 - Sqr(cosTheta_w));
        // BoundSubtendedDirections(bounds, p).cosTheta   vecmath.h:1815-1828
        float cosTheta_b;
        {
            V3 pCenter = (bounds.pMin + bounds.pMax) / 2;
            bool inside = pCenter.x >= bounds.pMin.x && pCenter.x <= bounds.pMax.x && pCenter.y >= bounds.pMin.y && pCenter.y <= bounds.pMax.y &&
                          pCenter.z >= bounds.pMin.z && pCenter.z <= bounds.pMax.z;
            float radius = inside ? Distance(pCenter, bounds.pMax) : 0;
            if (DistanceSquared(p, pCenter) < Sqr(radius)) cosTheta_b = -1;
            else { float sin2ThetaMax = Sqr(radius) / DistanceSquared(pCenter, p); cosTheta_b = SafeSqrt(1 - sin2ThetaMax); }
        }
        float sinTheta_b = SafeSqrt(1 - Sqr(cosTheta_b));
        float sinTheta_o = SafeSqrt(1 - Sqr(cosTheta_o));
        float cosTheta_x = cosSubClamped(sinTheta_w, cosTheta_w, sinTheta_o, cosTheta_o);
        float sinTheta_x = sinSubClamped(sinTheta_w, cosTheta_w, sinTheta_o, cosTheta_o);
        float cosThetap = cosSubClamped(sinTheta_x, cosTheta_x, sinTheta_b, cosTheta_b);
        if (cosThetap <= cosTheta_e) return 0;
        float importance = phi * cosThetap / d2;
        if (n != V3(0, 0, 0)) {
            float cosTheta_i = AbsDotN(wi, n);
            float sinTheta_i = SafeSqrt(1 - Sqr(cosTheta_i));
            float cosThetap_i = cosSubClamped(sinTheta_i, cosTheta_i, sinTheta_b, cosTheta_b);
            importance *= cosThetap_i;
        }
        return std::max<float>(importance, 0);
    }
};
struct LightBVHNode { CompactLightBounds lightBounds; unsigned childOrLightIndex = 0; bool isLeaf = false; };

// lightsamplers.h:258-404, lightsamplers.cpp:107-236
struct BVHLightSampler {
    std::vector<LightBVHNode> nodes;
    std::vector<int> infiniteLights;       // lights whose Bounds() is empty (distant, infinite), in light order
    std::vector<uint32_t> lightToBitTrail; // per light
    std::vector<char> hasKey;              // per light: in the tree
    LBox allLightBounds;
    typedef std::pair<int, LightBounds> Item;

    static float EvaluateCost(const LightBounds &b, const LBox &bounds, int dim) {   // lightsamplers.h:383-396
        if (b.phi == 0) return 0;   // an empty side; the reference reads uninitialised cosines here (see host_lightbvh.h)
        float theta_o = vspg_acosf(b.cosTheta_o), theta_e = vspg_acosf(b.cosTheta_e);
        float theta_w = std::min(theta_o + theta_e, Pi);
        float sinTheta_o = SafeSqrt(1 - Sqr(b.cosTheta_o));
        float M_omega = 2 * Pi * (1 - b.cosTheta_o) +
                        Pi / 2 * (2 * theta_w * sinTheta_o - vspg_cosf(theta_o - 2 * theta_w) - 2 * theta_o * sinTheta_o + b.cosTheta_o);
        V3 d = bounds.Diagonal();
        float Kr = std::max({d.x, d.y, d.z}) / d[dim];
        return b.phi * M_omega * Kr * b.bounds.SurfaceArea();
    }
    // lights: the scene's lights; quads + quadArea: the emitting bilinear patches
    // tris / spheres + their areas: emitting triangles (shapes.cpp:290-314) and full spheres (shapes.cpp:33-36, shapes.h:134)
    void Build(const vspg_light *lights, int nLights, const vspg_quad *quads, const std::vector<float> &quadArea,
               const vspg_triangle *tris = nullptr, const std::vector<float> *triArea = nullptr,
               const vspg_sphere *spheres = nullptr, const std::vector<float> *sphereArea = nullptr) {
        nodes.clear(); infiniteLights.clear();
        lightToBitTrail.assign(nLights, 0u); hasKey.assign(nLights, 0);
        allLightBounds = LBox();
        std::vector<Item> bvhLights;
        for (int i = 0; i < nLights; ++i) {
            const vspg_light &l = lights[i];
            LightBounds lb;
            float maxL = std::max({l.L[0], l.L[1], l.L[2]});
            if (l.type == VSPG_LIGHT_POINT) {   // lights.cpp:200-205
                lb.bounds.pMin = lb.bounds.pMax = V3(l.pos);
                lb.w = V3(0, 0, 1); lb.phi = 4 * Pi * l.scale * maxL; lb.cosTheta_o = vspg_cosf(Pi); lb.cosTheta_e = vspg_cosf(Pi / 2); lb.twoSided = false;
            } else if (l.type == VSPG_LIGHT_SPOT) {   // lights.cpp:1457-1468 SpotLight::Bounds
                lb.bounds.pMin = lb.bounds.pMax = V3(l.pos);
                lb.w = V3(l.axis); lb.phi = l.scale * maxL * 4 * Pi; lb.cosTheta_o = l.cos_falloff_start;
                lb.cosTheta_e = vspg_cosf(vspg_acosf(l.cos_falloff_end) - vspg_acosf(l.cos_falloff_start));
                if (lb.cosTheta_e == 1 && l.cos_falloff_end != l.cos_falloff_start) lb.cosTheta_e = 0.999f;
                lb.twoSided = false;
            } else if (l.type == VSPG_LIGHT_DIFFUSE_AREA && l.shape_kind == 1) {   // Triangle::Bounds / NormalBounds
                const vspg_triangle &q = tris[l.shape_index];
                V3 p0(q.p0), p1(q.p1), p2(q.p2);
                lb.bounds = Union(Union(Union(LBox(), p0), p1), p2);
                V3 n = Normalize(Cross(p1 - p0, p2 - p0));
                if (q.flip_normal) n = -n;
                lb.w = Normalize(n); lb.phi = maxL; lb.phi *= l.scale * (*triArea)[l.shape_index] * Pi; lb.cosTheta_o = 1; lb.cosTheta_e = vspg_cosf(Pi / 2);
                lb.twoSided = l.two_sided != 0;
            } else if (l.type == VSPG_LIGHT_DIFFUSE_AREA && l.shape_kind == 2) {   // Sphere::Bounds, NormalBounds = EntireSphere
                const vspg_sphere &q = spheres[l.shape_index];
                V3 c(q.center);
                lb.bounds.pMin = V3(-q.radius + c.x, -q.radius + c.y, -q.radius + c.z);
                lb.bounds.pMax = V3(q.radius + c.x, q.radius + c.y, q.radius + c.z);
                lb.w = V3(0, 0, 1); lb.phi = maxL; lb.phi *= l.scale * (*sphereArea)[l.shape_index] * Pi; lb.cosTheta_o = -1; lb.cosTheta_e = vspg_cosf(Pi / 2);
                lb.twoSided = l.two_sided != 0;
            } else if (l.type == VSPG_LIGHT_DIFFUSE_AREA) {   // lights.cpp:845-864 + shapes.cpp:1070-1126
                const vspg_quad &q = quads[l.shape_index];
                V3 p00(q.p00), p10(q.p10), p01(q.p01), p11(q.p11);
                LBox b0, b1;
                b0 = Union(Union(LBox(), p00), p01); b1 = Union(Union(LBox(), p10), p11);
                lb.bounds = Union(b0, b1);
                V3 n; float cosTheta;
                if (p00 == p10 || p10 == p11 || p11 == p01 || p01 == p00) {
                    V3 dpdu = LerpV(0.5f, p10, p11) - LerpV(0.5f, p00, p01), dpdv = LerpV(0.5f, p01, p11) - LerpV(0.5f, p00, p10);
                    n = Normalize(Cross(dpdu, dpdv));
                    if (q.flip_normal) n = -n;
                    cosTheta = 1;
                } else {
                    V3 n00 = Normalize(Cross(p10 - p00, p01 - p00)), n10 = Normalize(Cross(p11 - p10, p00 - p10));
                    V3 n01 = Normalize(Cross(p00 - p01, p11 - p01)), n11 = Normalize(Cross(p01 - p11, p10 - p11));
                    if (q.flip_normal) { n00 = -n00; n10 = -n10; n01 = -n01; n11 = -n11; }
                    n = Normalize(n00 + n10 + n01 + n11);
                    cosTheta = std::min({Dot(n, n00), Dot(n, n01), Dot(n, n10), Dot(n, n11)});
                    cosTheta = std::min(1.f, std::max(-1.f, cosTheta));
                }
                lb.w = Normalize(n); lb.phi = maxL; lb.phi *= l.scale * quadArea[l.shape_index] * Pi; lb.cosTheta_o = cosTheta; lb.cosTheta_e = vspg_cosf(Pi / 2);
                lb.twoSided = l.two_sided != 0;
            } else { infiniteLights.push_back(i); continue; }
            if (lb.phi > 0) {
                bvhLights.push_back({i, lb});
                hasKey[i] = 1;
                allLightBounds = Union(allLightBounds, lb.bounds);
            }
        }
        if (!bvhLights.empty()) buildBVH(bvhLights, 0, (int)bvhLights.size(), 0, 0);
    }
    static int BucketOf(const LightBounds &lb, const LBox &cb, int dim) {
        float o = lb.Centroid()[dim] - cb.pMin[dim];
        if (cb.pMax[dim] > cb.pMin[dim]) o /= cb.pMax[dim] - cb.pMin[dim];   // LBox::Offset
        int b = (int)(12 * o);
        return b == 12 ? 11 : b;
    }
    std::pair<int, LightBounds> buildBVH(std::vector<Item> &bvhLights, int start, int end, uint32_t bitTrail, int depth) {
        if (end - start == 1) {
            int nodeIndex = (int)nodes.size();
            LightBVHNode nd; nd.lightBounds = CompactLightBounds(bvhLights[start].second, allLightBounds);
            nd.childOrLightIndex = (unsigned)bvhLights[start].first; nd.isLeaf = true;
            nodes.push_back(nd);
            lightToBitTrail[bvhLights[start].first] = bitTrail;
            return {nodeIndex, bvhLights[start].second};
        }
        LBox bounds, centroidBounds;
        for (int i = start; i < end; ++i) {
            bounds = Union(bounds, bvhLights[i].second.bounds);
            centroidBounds = Union(centroidBounds, bvhLights[i].second.Centroid());
        }
        float minCost = Infinity;
        int minCostSplitBucket = -1, minCostSplitDim = -1;
        constexpr int nBuckets = 12;
        for (int dim = 0; dim < 3; ++dim) {
            if (centroidBounds.pMax[dim] == centroidBounds.pMin[dim]) continue;
            LightBounds bucketLightBounds[nBuckets];
            for (int i = start; i < end; ++i) {
                int b = BucketOf(bvhLights[i].second, centroidBounds, dim);
                bucketLightBounds[b] = Union(bucketLightBounds[b], bvhLights[i].second);
            }
            float cost[nBuckets - 1];
            for (int i = 0; i < nBuckets - 1; ++i) {
                LightBounds b0, b1;
                for (int j = 0; j <= i; ++j) b0 = Union(b0, bucketLightBounds[j]);
                for (int j = i + 1; j < nBuckets; ++j) b1 = Union(b1, bucketLightBounds[j]);
                cost[i] = EvaluateCost(b0, bounds, dim) + EvaluateCost(b1, bounds, dim);
            }
            for (int i = 1; i < nBuckets - 1; ++i)
                if (cost[i] > 0 && cost[i] < minCost) { minCost = cost[i]; minCostSplitBucket = i; minCostSplitDim = dim; }
        }
        int mid;
        if (minCostSplitDim == -1) mid = (start + end) / 2;
        else {
            auto pmid = std::partition(bvhLights.begin() + start, bvhLights.begin() + end, [&](const Item &l) {
                return BucketOf(l.second, centroidBounds, minCostSplitDim) <= minCostSplitBucket;
            });
            mid = (int)(pmid - bvhLights.begin());
            if (mid == start || mid == end) mid = (start + end) / 2;
        }
        int nodeIndex = (int)nodes.size();
        nodes.push_back(LightBVHNode());
        auto child0 = buildBVH(bvhLights, start, mid, bitTrail, depth + 1);
        auto child1 = buildBVH(bvhLights, mid, end, bitTrail | (1u << depth), depth + 1);
        LightBounds lb = Union(child0.second, child1.second);
        LightBVHNode nd; nd.lightBounds = CompactLightBounds(lb, allLightBounds);
        nd.childOrLightIndex = (unsigned)child1.first; nd.isLeaf = false;
        nodes[nodeIndex] = nd;
        return {nodeIndex, lb};
    }
    // lightsamplers.h:266-321
    int Sample(V3 p, V3 n, float u, float *pmfOut) const {
        const int nInf = (int)infiniteLights.size();
        float pInfinite = float(nInf) / float(nInf + (nodes.empty() ? 0 : 1));
        if (u < pInfinite) {
            u /= pInfinite;
            int index = std::min<int>((int)(u * nInf), nInf - 1);
            *pmfOut = pInfinite / nInf;
            return infiniteLights[index];
        }
        if (nodes.empty()) return -1;
        u = std::min<float>((u - pInfinite) / (1 - pInfinite), OneMinusEpsilon);
        int nodeIndex = 0;
        float pmf = 1 - pInfinite;
        while (true) {
            LightBVHNode node = nodes[nodeIndex];
            if (!node.isLeaf) {
                const LightBVHNode *children[2] = {&nodes[nodeIndex + 1], &nodes[node.childOrLightIndex]};
                float ci[2] = {children[0]->lightBounds.Importance(p, n, allLightBounds), children[1]->lightBounds.Importance(p, n, allLightBounds)};
                if (ci[0] == 0 && ci[1] == 0) return -1;
                // SampleDiscrete(ci, u, &nodePMF, &u)   util/sampling.h:79-112
                float sumWeights = 0;
                for (float w : ci) sumWeights += w;
                float up = u * sumWeights;
                if (up == sumWeights) up = NextFloatDown(up);
                int offset = 0; float sum = 0;
                while (sum + ci[offset] <= up) sum += ci[offset++];
                float nodePMF = ci[offset] / sumWeights;
                u = std::min((up - sum) / ci[offset], OneMinusEpsilon);
                pmf *= nodePMF;
                nodeIndex = (offset == 0) ? (nodeIndex + 1) : (int)node.childOrLightIndex;
            } else {
                if (nodeIndex > 0 || node.lightBounds.Importance(p, n, allLightBounds) > 0) { *pmfOut = pmf; return (int)node.childOrLightIndex; }
                return -1;
            }
        }
    }
    // lightsamplers.h:324-359
    float PMF(V3 p, V3 n, int light) const {
        const int nInf = (int)infiniteLights.size();
        if (!hasKey[light]) return 1.f / (nInf + (nodes.empty() ? 0 : 1));
        uint32_t bitTrail = lightToBitTrail[light];
        float pInfinite = float(nInf) / float(nInf + (nodes.empty() ? 0 : 1));
        float pmf = 1 - pInfinite;
        int nodeIndex = 0;
        while (true) {
            const LightBVHNode *node = &nodes[nodeIndex];
            if (node->isLeaf) return pmf;
            const LightBVHNode *child0 = &nodes[nodeIndex + 1];
            const LightBVHNode *child1 = &nodes[node->childOrLightIndex];
            float ci[2] = {child0->lightBounds.Importance(p, n, allLightBounds), child1->lightBounds.Importance(p, n, allLightBounds)};
            pmf *= ci[bitTrail & 1] / (ci[0] + ci[1]);
            nodeIndex = (bitTrail & 1) ? (int)node->childOrLightIndex : (nodeIndex + 1);
            bitTrail >>= 1;
        }
    }
};

}  // namespace orc
