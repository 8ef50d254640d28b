// Light BVH of the reference's default light sampler ("bvh"), built on the host at set_scene time.
//   BVHLightSampler ctor / buildBVH / EvaluateCost      src/pbrt/lightsamplers.cpp:107-236, lightsamplers.h:383-396
//   LightBounds, Union                                   src/pbrt/lights.h:104-152
//   CompactLightBounds quantisation                      src/pbrt/lightsamplers.h:100-228
//   DirectionCone Union, OctahedralVector                src/pbrt/util/vecmath.cpp:56-83, util/vecmath.h:1733-1782
//   light Bounds(): DiffuseAreaLight / PointLight        src/pbrt/lights.cpp:845-864, 200-205 (distant / infinite: none)
//   BilinearPatch::Bounds / NormalBounds                 src/pbrt/shapes.cpp:1070-1126
// The nodes are stored DECODED (bounds, axis and cosines as the floats CompactLightBounds::Bounds / CosTheta_* /
// Vector3f(w) return): the decode is a fixed sequence of IEEE operations, so doing it once here gives the device the
// same numbers the reference recomputes at every Importance() call.
#pragma once
#include "../../include/vspg_libm.h"   // the path's libm (portable, identical on host and device)
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <utility>
#include <vector>

#include "host_util.h"

namespace vspg_host {

struct LBounds {
    float bmin[3] = {INFINITY, INFINITY, INFINITY}, bmax[3] = {-INFINITY, -INFINITY, -INFINITY};
    float phi = 0;
    H3 w;
    float cosTheta_o = 1, cosTheta_e = 1;
    bool twoSided = false;
    H3 centroid() const { return H3((bmin[0] + bmax[0]) / 2, (bmin[1] + bmax[1]) / 2, (bmin[2] + bmax[2]) / 2); }
};

struct HostLightNode {   // 64 bytes, mirrored by DLightNode on the device
    float bmin[3], bmax[3], w[3];
    float phi, cosTheta_o, cosTheta_e;
    int twoSided, childOrLight, isLeaf, pad;
};
struct HostLightBvh {
    std::vector<HostLightNode> nodes;
    std::vector<uint32_t> bitTrail;   // per light; meaningful for lights in the tree
    std::vector<int> inTree;          // per light: 1 if the light sits in the tree
    std::vector<int> infinite;        // lights without bounds (distant, infinite), in light order
};

namespace lbvh {
constexpr float kPi = 3.14159265358979323846f;
inline float safe_acos(float x) { return vspg_acosf(std::min(1.f, std::max(-1.f, x))); }
inline float safe_asin(float x) { return vspg_asinf(std::min(1.f, std::max(-1.f, x))); }
inline float safe_sqrt(float x) { return std::sqrt(std::max(0.f, x)); }
inline float angle_between(H3 a, H3 b) {   // util/vecmath.h:972-977
    if (hdot(a, b) < 0) return kPi - 2 * safe_asin(hlen(a + b) / 2);
    return 2 * safe_asin(hlen(b - a) / 2);
}
struct Cone { H3 w; float cosTheta = INFINITY; bool empty() const { return cosTheta == INFINITY; } };
inline Cone entire_sphere() { Cone c; c.w = H3(0, 0, 1); c.cosTheta = -1; return c; }
// util/transform.cpp Rotate(sinTheta, cosTheta, axis) applied to a vector
inline H3 rotate_about(float thetaRad, H3 axis, H3 v) {
    H3 a = hnormalize(axis);
    // Rotate(Degrees(theta_r), wr): the reference converts to degrees and back; sin/cos of the round-tripped angle
    float deg = (180 / kPi) * thetaRad;
    float th = (kPi / 180) * deg;
    float s = vspg_sinf(th), c = vspg_cosf(th);
    float m[3][3] = {{a.x * a.x + (1 - a.x * a.x) * c, a.x * a.y * (1 - c) - a.z * s, a.x * a.z * (1 - c) + a.y * s},
                     {a.x * a.y * (1 - c) + a.z * s, a.y * a.y + (1 - a.y * a.y) * c, a.y * a.z * (1 - c) - a.x * s},
                     {a.x * a.z * (1 - c) - a.y * s, a.y * a.z * (1 - c) + a.x * s, a.z * a.z + (1 - a.z * a.z) * c}};
    return H3(m[0][0] * v.x + m[0][1] * v.y + m[0][2] * v.z, m[1][0] * v.x + m[1][1] * v.y + m[1][2] * v.z,
              m[2][0] * v.x + m[2][1] * v.y + m[2][2] * v.z);
}
inline Cone cone_union(const Cone &a, const Cone &b) {   // util/vecmath.cpp:56-83
    if (a.empty()) return b;
    if (b.empty()) return a;
    float theta_a = safe_acos(a.cosTheta), theta_b = safe_acos(b.cosTheta);
    float theta_d = angle_between(a.w, b.w);
    if (std::min(theta_d + theta_b, kPi) <= theta_a) return a;
    if (std::min(theta_d + theta_a, kPi) <= theta_b) return b;
    float theta_o = (theta_a + theta_d + theta_b) / 2;
    if (theta_o >= kPi) return entire_sphere();
    float theta_r = theta_o - theta_a;
    H3 wr = hcross(a.w, b.w);
    if (hlen2(wr) == 0) return entire_sphere();
    Cone r; r.w = hnormalize(rotate_about(theta_r, wr, a.w)); r.cosTheta = vspg_cosf(theta_o);
    return r;
}
inline LBounds lb_union(const LBounds &a, const LBounds &b) {   // lights.h:137-152
    if (a.phi == 0) return b;
    if (b.phi == 0) return a;
    Cone ca; ca.w = hnormalize(a.w); ca.cosTheta = a.cosTheta_o;
    Cone cb; cb.w = hnormalize(b.w); cb.cosTheta = b.cosTheta_o;
    Cone c = cone_union(ca, cb);
    LBounds r;
    for (int k = 0; k < 3; k++) { r.bmin[k] = std::min(a.bmin[k], b.bmin[k]); r.bmax[k] = std::max(a.bmax[k], b.bmax[k]); }
    r.w = hnormalize(c.w); r.phi = a.phi + b.phi; r.cosTheta_o = c.cosTheta; r.cosTheta_e = std::min(a.cosTheta_e, b.cosTheta_e);
    r.twoSided = a.twoSided | b.twoSided;
    return r;
}
inline float evaluate_cost(const LBounds &b, const float *bmin, const float *bmax, int dim) {   // lightsamplers.h:383-396
    if (b.phi == 0) return 0;   // an empty side: the reference reads uninitialised cosines here; defined as zero cost
    float theta_o = vspg_acosf(b.cosTheta_o), theta_e = vspg_acosf(b.cosTheta_e);
    float theta_w = std::min(theta_o + theta_e, kPi);
    float sinTheta_o = safe_sqrt(1 - b.cosTheta_o * b.cosTheta_o);
    float M_omega = 2 * kPi * (1 - b.cosTheta_o) +
                    kPi / 2 * (2 * theta_w * sinTheta_o - vspg_cosf(theta_o - 2 * theta_w) - 2 * theta_o * sinTheta_o + b.cosTheta_o);
    float d[3] = {bmax[0] - bmin[0], bmax[1] - bmin[1], bmax[2] - bmin[2]};
    float Kr = std::max({d[0], d[1], d[2]}) / d[dim];
    float e[3] = {b.bmax[0] - b.bmin[0], b.bmax[1] - b.bmin[1], b.bmax[2] - b.bmin[2]};
    float area = 2 * (e[0] * e[1] + e[0] * e[2] + e[1] * e[2]);
    return b.phi * M_omega * Kr * area;
}
inline uint16_t oct_encode(float f) { return (uint16_t)std::round(std::min(1.f, std::max(0.f, (f + 1) / 2)) * 65535.f); }
inline float signf(float v) { return std::copysign(1.f, v); }
inline H3 oct_roundtrip(H3 v) {   // OctahedralVector(v) then Vector3f(oct)
    float s = std::abs(v.x) + std::abs(v.y) + std::abs(v.z);
    v = v / s;
    uint16_t x, y;
    if (v.z >= 0) { x = oct_encode(v.x); y = oct_encode(v.y); }
    else { x = oct_encode((1 - std::abs(v.y)) * signf(v.x)); y = oct_encode((1 - std::abs(v.x)) * signf(v.y)); }
    H3 r;
    r.x = -1 + 2 * (x / 65535.f);
    r.y = -1 + 2 * (y / 65535.f);
    r.z = 1 - (std::abs(r.x) + std::abs(r.y));
    if (r.z < 0) { float xo = r.x; r.x = (1 - std::abs(r.y)) * signf(xo); r.y = (1 - std::abs(xo)) * signf(r.y); }
    return hnormalize(r);
}
inline HostLightNode compact(const LBounds &lb, const float *amin, const float *amax) {   // CompactLightBounds ctor + accessors
    HostLightNode n{};
    H3 w = oct_roundtrip(hnormalize(lb.w));
    n.w[0] = w.x; n.w[1] = w.y; n.w[2] = w.z;
    n.phi = lb.phi;
    unsigned qo = (unsigned)std::floor(32767.f * ((lb.cosTheta_o + 1) / 2)), qe = (unsigned)std::floor(32767.f * ((lb.cosTheta_e + 1) / 2));
    n.cosTheta_o = 2 * (qo / 32767.f) - 1;
    n.cosTheta_e = 2 * (qe / 32767.f) - 1;
    n.twoSided = lb.twoSided;
    for (int c = 0; c < 3; c++) {
        auto quant = [&](float v) { return amin[c] == amax[c] ? 0.f : 65535.f * std::min(1.f, std::max(0.f, (v - amin[c]) / (amax[c] - amin[c]))); };
        uint16_t q0 = (uint16_t)std::floor(quant(lb.bmin[c])), q1 = (uint16_t)std::ceil(quant(lb.bmax[c]));
        float t0 = q0 / 65535.f, t1 = q1 / 65535.f;
        n.bmin[c] = (1 - t0) * amin[c] + t0 * amax[c];
        n.bmax[c] = (1 - t1) * amin[c] + t1 * amax[c];
    }
    return n;
}

struct Builder {
    HostLightBvh &out;
    float amin[3], amax[3];
    typedef std::pair<int, LBounds> Item;
    std::pair<int, LBounds> build(std::vector<Item> &L, int start, int end, uint32_t bitTrail, int depth) {
        if (end - start == 1) {
            int nodeIndex = (int)out.nodes.size();
            HostLightNode n = compact(L[start].second, amin, amax);
            n.childOrLight = L[start].first; n.isLeaf = 1;
            out.nodes.push_back(n);
            out.bitTrail[L[start].first] = bitTrail;
            return {nodeIndex, L[start].second};
        }
        float bmin[3] = {INFINITY, INFINITY, INFINITY}, bmax[3] = {-INFINITY, -INFINITY, -INFINITY};
        float cmin[3] = {INFINITY, INFINITY, INFINITY}, cmax[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int i = start; i < end; ++i) {
            const LBounds &lb = L[i].second;
            H3 c = lb.centroid();
            for (int k = 0; k < 3; k++) {
                bmin[k] = std::min(bmin[k], lb.bmin[k]); bmax[k] = std::max(bmax[k], lb.bmax[k]);
                cmin[k] = std::min(cmin[k], c[k]); cmax[k] = std::max(cmax[k], c[k]);
            }
        }
        auto bucket_of = [&](const LBounds &lb, int dim) {
            float o = lb.centroid()[dim] - cmin[dim];
            if (cmax[dim] > cmin[dim]) o /= cmax[dim] - cmin[dim];
            int b = (int)(12 * o);
            return b == 12 ? 11 : b;
        };
        float minCost = INFINITY;
        int minCostSplitBucket = -1, minCostSplitDim = -1;
        constexpr int nBuckets = 12;
        for (int dim = 0; dim < 3; ++dim) {
            if (cmax[dim] == cmin[dim]) continue;
            LBounds bucket[nBuckets];
            for (int i = start; i < end; ++i) { int b = bucket_of(L[i].second, dim); bucket[b] = lb_union(bucket[b], L[i].second); }
            float cost[nBuckets - 1];
            for (int i = 0; i < nBuckets - 1; ++i) {
                LBounds b0, b1;
                for (int j = 0; j <= i; ++j) b0 = lb_union(b0, bucket[j]);
                for (int j = i + 1; j < nBuckets; ++j) b1 = lb_union(b1, bucket[j]);
                cost[i] = evaluate_cost(b0, bmin, bmax, dim) + evaluate_cost(b1, bmin, bmax, dim);
            }
            for (int i = 1; i < nBuckets - 1; ++i)
                if (cost[i] > 0 && cost[i] < minCost) { minCost = cost[i]; minCostSplitBucket = i; minCostSplitDim = dim; }
        }
        int mid;
        if (minCostSplitDim == -1) mid = (start + end) / 2;
        else {
            auto pmid = std::partition(L.begin() + start, L.begin() + end,
                                       [&](const Item &l) { return bucket_of(l.second, minCostSplitDim) <= minCostSplitBucket; });
            mid = (int)(pmid - L.begin());
            if (mid == start || mid == end) mid = (start + end) / 2;
        }
        int nodeIndex = (int)out.nodes.size();
        out.nodes.push_back(HostLightNode{});
        auto c0 = build(L, start, mid, bitTrail, depth + 1);
        auto c1 = build(L, mid, end, bitTrail | (1u << depth), depth + 1);
        LBounds lb = lb_union(c0.second, c1.second);
        HostLightNode n = compact(lb, amin, amax);
        n.childOrLight = c1.first; n.isLeaf = 0;
        out.nodes[nodeIndex] = n;
        return {nodeIndex, lb};
    }
};
}  // namespace lbvh

// scene: lights + the shapes they emit from (render space); primArea[i]: area of the i-th primitive counted over
// quads, then triangles, then spheres (as set_scene computes them)
inline HostLightBvh build_light_bvh(const vspg_scene_desc *scene, const float *primArea) {
    using namespace lbvh;
    const vspg_light *lights = scene->lights; const int nLights = scene->n_lights;
    const vspg_quad *quads = scene->quads;
    const float *quadArea = primArea, *triArea = primArea + scene->n_quads, *sphereArea = primArea + scene->n_quads + scene->n_triangles;
    HostLightBvh out;
    out.bitTrail.assign(nLights, 0u); out.inTree.assign(nLights, 0);
    std::vector<Builder::Item> items;
    float amin[3] = {INFINITY, INFINITY, INFINITY}, amax[3] = {-INFINITY, -INFINITY, -INFINITY};
    const float cosHalfPi = vspg_cosf(kPi / 2);
    for (int i = 0; i < nLights; i++) {
        const vspg_light &l = lights[i];
        LBounds lb;
        float maxL = std::max({l.L[0], l.L[1], l.L[2]});
        if (l.type == VSPG_LIGHT_POINT) {   // lights.cpp:200-205
            for (int k = 0; k < 3; k++) lb.bmin[k] = lb.bmax[k] = l.pos[k];
            lb.w = H3(0, 0, 1); lb.phi = 4 * kPi * l.scale * maxL; lb.cosTheta_o = vspg_cosf(kPi); lb.cosTheta_e = cosHalfPi; lb.twoSided = false;
        } else if (l.type == VSPG_LIGHT_SPOT) {   // lights.cpp:1457-1468
            for (int k = 0; k < 3; k++) lb.bmin[k] = lb.bmax[k] = l.pos[k];
            lb.w = H3(l.axis); lb.phi = l.scale * maxL * 4 * kPi; lb.cosTheta_o = l.cos_falloff_start;
            lb.cosTheta_e = vspg_cosf(vspg_acosf(l.cos_falloff_end) - vspg_acosf(l.cos_falloff_start));
            if (lb.cosTheta_e == 1 && l.cos_falloff_end != l.cos_falloff_start) lb.cosTheta_e = 0.999f;
            lb.twoSided = false;
        } else if (l.type == VSPG_LIGHT_DIFFUSE_AREA && l.shape_kind == 1) {   // shapes.cpp:290-314 Triangle::Bounds / NormalBounds
            const vspg_triangle &q = scene->triangles[l.shape_index];
            H3 p0(q.p0), p1(q.p1), p2(q.p2);
            for (int k = 0; k < 3; k++) { lb.bmin[k] = std::min(std::min(p0[k], p1[k]), p2[k]); lb.bmax[k] = std::max(std::max(p0[k], p1[k]), p2[k]); }
            H3 n = hnormalize(hcross(p1 - p0, p2 - p0));
            if (q.flip_normal) n = -1.f * n;
            lb.w = hnormalize(n); lb.phi = maxL; lb.phi *= l.scale * triArea[l.shape_index] * kPi; lb.cosTheta_o = 1; lb.cosTheta_e = cosHalfPi;
            lb.twoSided = l.two_sided != 0;
        } else if (l.type == VSPG_LIGHT_DIFFUSE_AREA && l.shape_kind == 2) {   // shapes.cpp:33-36 Sphere::Bounds, shapes.h:134 EntireSphere
            const vspg_sphere &q = scene->spheres[l.shape_index];
            for (int k = 0; k < 3; k++) { lb.bmin[k] = -q.radius + q.center[k]; lb.bmax[k] = q.radius + q.center[k]; }
            lb.w = H3(0, 0, 1); lb.phi = maxL; lb.phi *= l.scale * sphereArea[l.shape_index] * kPi; lb.cosTheta_o = -1; lb.cosTheta_e = cosHalfPi;
            lb.twoSided = l.two_sided != 0;
        } else if (l.type == VSPG_LIGHT_DIFFUSE_AREA) {   // lights.cpp:845-864
            const vspg_quad &q = quads[l.shape_index];
            H3 p00(q.p00), p10(q.p10), p01(q.p01), p11(q.p11);
            for (int k = 0; k < 3; k++) {
                lb.bmin[k] = std::min(std::min(p00[k], p01[k]), std::min(p10[k], p11[k]));
                lb.bmax[k] = std::max(std::max(p00[k], p01[k]), std::max(p10[k], p11[k]));
            }
            H3 n; float cosTheta;
            if (p00 == p10 || p10 == p11 || p11 == p01 || p01 == p00) {   // shapes.cpp:1088-1099
                H3 dpdu = hlerp(0.5f, p10, p11) - hlerp(0.5f, p00, p01), dpdv = hlerp(0.5f, p01, p11) - hlerp(0.5f, p00, p10);
                n = hnormalize(hcross(dpdu, dpdv));
                if (q.flip_normal) n = -1.f * n;
                cosTheta = 1;
            } else {
                H3 n00 = hnormalize(hcross(p10 - p00, p01 - p00)), n10 = hnormalize(hcross(p11 - p10, p00 - p10));
                H3 n01 = hnormalize(hcross(p00 - p01, p11 - p01)), n11 = hnormalize(hcross(p01 - p11, p10 - p11));
                if (q.flip_normal) { n00 = -1.f * n00; n10 = -1.f * n10; n01 = -1.f * n01; n11 = -1.f * n11; }
                n = hnormalize(n00 + n10 + n01 + n11);
                cosTheta = std::min({hdot(n, n00), hdot(n, n01), hdot(n, n10), hdot(n, n11)});
                cosTheta = std::min(1.f, std::max(-1.f, cosTheta));
            }
            lb.w = hnormalize(n); lb.phi = maxL; lb.phi *= l.scale * quadArea[l.shape_index] * kPi; lb.cosTheta_o = cosTheta; lb.cosTheta_e = cosHalfPi;
            lb.twoSided = l.two_sided != 0;
        } else { out.infinite.push_back(i); continue; }
        if (lb.phi > 0) {
            items.push_back({i, lb});
            out.inTree[i] = 1;
            for (int k = 0; k < 3; k++) { amin[k] = std::min(amin[k], lb.bmin[k]); amax[k] = std::max(amax[k], lb.bmax[k]); }
        }
    }
    if (!items.empty()) {
        Builder B{out};
        for (int k = 0; k < 3; k++) { B.amin[k] = amin[k]; B.amax[k] = amax[k]; }
        B.build(items, 0, (int)items.size(), 0, 0);
    }
    return out;
}

}  // namespace vspg_host
