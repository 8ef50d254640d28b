// Host-side helpers shared by the C-ABI implementation files: float vector math with the reference's operation
// order (compiled with -ffp-contract=off), BilinearPatch::IsRectangle / Area, majorant-grid construction and the
// threaded-BVH builder.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../../include/vspg.h"

namespace vspg_host {

struct H3 {
    float x = 0, y = 0, z = 0;
    H3() = default;
    H3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit H3(const float *p) : x(p[0]), y(p[1]), z(p[2]) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline H3 operator+(H3 a, H3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline H3 operator-(H3 a, H3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline H3 operator*(float s, H3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline H3 operator/(H3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline bool operator==(H3 a, H3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline float hdop(float a, float b, float c, float d) {  // util/math.h:569 DifferenceOfProducts
    float cd = c * d;
    float dop = std::fma(a, b, -cd);
    float err = std::fma(-c, d, cd);
    return dop + err;
}
inline float hdot(H3 a, H3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float hdotn(H3 a, H3 b) {   // Dot with a Normal3 operand (util/vecmath.h:1056-1077): FMA + compensated SumOfProducts
    float cd = a.z * b.z;
    float sop = std::fma(a.y, b.y, cd);
    float err = std::fma(a.z, b.z, -cd);
    return std::fma(a.x, b.x, sop + err);
}
inline H3 hcross(H3 v, H3 w) { return {hdop(v.y, w.z, v.z, w.y), hdop(v.z, w.x, v.x, w.z), hdop(v.x, w.y, v.y, w.x)}; }
inline float hlen2(H3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
inline float hlen(H3 v) { return std::sqrt(hlen2(v)); }
inline H3 hnormalize(H3 v) { return v / hlen(v); }
inline H3 hlerp(float t, H3 a, H3 b) { return (1 - t) * a + t * b; }
inline float hdist2(H3 a, H3 b) { return hlen2(a - b); }
inline float hdist(H3 a, H3 b) { return hlen(a - b); }

// shapes.h BilinearPatch::IsRectangle
inline bool quad_is_rectangle(H3 p00, H3 p10, H3 p01, H3 p11) {
    if (p00 == p01 || p01 == p11 || p11 == p10 || p10 == p00) return false;
    H3 n = hnormalize(hcross(p10 - p00, p01 - p00));
    if (std::abs(hdotn(hnormalize(p11 - p00), n)) > 1e-5f) return false;
    H3 pCenter = (p00 + p01 + p10 + p11) / 4;
    float d2[4] = {hdist2(p00, pCenter), hdist2(p01, pCenter), hdist2(p10, pCenter), hdist2(p11, pCenter)};
    for (int i = 1; i < 4; ++i)
        if (std::abs(d2[i] - d2[0]) / d2[0] > 1e-4f) return false;
    return true;
}
// shapes.cpp:1038-1068 BilinearPatch area
inline float quad_area(H3 p00, H3 p10, H3 p01, H3 p11) {
    if (quad_is_rectangle(p00, p10, p01, p11)) return hdist(p00, p01) * hdist(p00, p10);
    constexpr int na = 3;
    H3 p[na + 1][na + 1];
    for (int i = 0; i <= na; ++i) {
        float u = float(i) / float(na);
        for (int j = 0; j <= na; ++j) {
            float v = float(j) / float(na);
            p[i][j] = hlerp(u, hlerp(v, p00, p01), hlerp(v, p10, p11));
        }
    }
    float area = 0;
    for (int i = 0; i < na; ++i)
        for (int j = 0; j < na; ++j) area += 0.5f * hlen(hcross(p[i + 1][j + 1] - p[i][j], p[i + 1][j] - p[i][j + 1]));
    return area;
}

// ------------------------------------------------------------------ majorant grid (media.cpp:256-269, 393-408)
struct GridView {
    const vspg_medium *m;
    float at(int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= m->nx || y >= m->ny || z >= m->nz) return 0.f;
        return m->density[(size_t)(z * m->ny + y) * m->nx + x];
    }
    float rgbMax(const float *g, int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= m->nx || y >= m->ny || z >= m->nz) return 0.f;
        const float *q = g + 3 * ((size_t)(z * m->ny + y) * m->nx + x);
        return std::max({q[0], q[1], q[2]});
    }
};
// util/containers.h:838-855 SampledGrid::MaxValue(bounds)
template <typename F>
inline float grid_max_value(const vspg_medium *m, const float *p0, const float *p1, F at) {
    float ps0[3] = {p0[0] * m->nx - .5f, p0[1] * m->ny - .5f, p0[2] * m->nz - .5f};
    float ps1[3] = {p1[0] * m->nx - .5f, p1[1] * m->ny - .5f, p1[2] * m->nz - .5f};
    int n[3] = {m->nx, m->ny, m->nz};
    int lo[3], hi[3];
    for (int a = 0; a < 3; a++) {
        lo[a] = std::max((int)std::floor(ps0[a]), 0);
        hi[a] = std::min((int)std::floor(ps1[a]) + 1, n[a] - 1);
    }
    float maxValue = at(lo[0], lo[1], lo[2]);
    for (int z = lo[2]; z <= hi[2]; ++z)
        for (int y = lo[1]; y <= hi[1]; ++y)
            for (int x = lo[0]; x <= hi[0]; ++x) maxValue = std::max(maxValue, at(x, y, z));
    return maxValue;
}
inline void build_majorant_grid(const vspg_medium *m, float *out) {
    int rx = m->maj_res[0], ry = m->maj_res[1], rz = m->maj_res[2];
    GridView gv{m};
    for (int z = 0; z < rz; ++z)
        for (int y = 0; y < ry; ++y)
            for (int x = 0; x < rx; ++x) {
                float p0[3] = {float(x) / rx, float(y) / ry, float(z) / rz};  // media.h:126-130 VoxelBounds
                float p1[3] = {float(x + 1) / rx, float(y + 1) / ry, float(z + 1) / rz};
                float v;
                if (m->type == VSPG_MEDIUM_GRID) v = grid_max_value(m, p0, p1, [&](int a, int b, int c) { return gv.at(a, b, c); });
                else {
                    float ma = m->sigma_a_rgb ? grid_max_value(m, p0, p1, [&](int a, int b, int c) { return gv.rgbMax(m->sigma_a_rgb, a, b, c); }) : 1.f;
                    float ms = m->sigma_s_rgb ? grid_max_value(m, p0, p1, [&](int a, int b, int c) { return gv.rgbMax(m->sigma_s_rgb, a, b, c); }) : 1.f;
                    v = m->sigma_scale * (ma + ms);
                }
                out[x + rx * (y + ry * z)] = v;
            }
}

// ------------------------------------------------------------------ PixelFilter "gaussian" sampling table
// filters.h:80-118 GaussianFilter, filters.h:26-45 + filters.cpp:133-147 FilterSampler, util/sampling.h:603-760
// PiecewiseConstant1D/2D construction.  Layout handed to the device: f[ny][nx], condCdf[ny][nx+1], condInt[ny]
// (= the marginal's func), margCdf[ny+1], margInt.
inline float host_fast_exp(float x) {   // util/math.h:450-475 (CPU branch)
    float xp = x * 1.442695041f;
    float fxp = std::floor(xp), f = xp - fxp;
    int i = (int)fxp;
    float twoToF = std::fma(f, std::fma(f, std::fma(f, 0.0781455737f, 0.226173572f), 0.695556856f), 1.f);
    uint32_t bits; memcpy(&bits, &twoToF, 4);
    int exponent = (int)(bits >> 23) - 127 + i;
    if (exponent < -126) return 0.f;
    if (exponent > 127) return std::numeric_limits<float>::infinity();
    bits &= 0x807fffffu;
    bits |= (uint32_t)(exponent + 127) << 23;
    float r; memcpy(&r, &bits, 4);
    return r;
}
struct HostFilterTable { int nx = 0, ny = 0; std::vector<float> f, condCdf, condInt, margCdf; float margInt = 0; };
inline void piecewise_cdf(const float *func, int n, float mn, float mx, float *cdf, float *funcInt) {
    cdf[0] = 0;
    for (int i = 1; i < n + 1; ++i) cdf[i] = cdf[i - 1] + std::abs(func[i - 1]) * (mx - mn) / n;
    *funcInt = cdf[n];
    if (*funcInt == 0) for (int i = 1; i < n + 1; ++i) cdf[i] = float(i) / float(n);
    else for (int i = 1; i < n + 1; ++i) cdf[i] /= *funcInt;
}
inline HostFilterTable build_gaussian_filter(float rx, float ry, float sigma) {
    const float Pi = 3.14159265358979323846f;
    auto gaussian = [&](float x) { return 1 / std::sqrt(2 * Pi * sigma * sigma) * host_fast_exp(-(x * x) / (2 * sigma * sigma)); };
    auto lerp = [](float t, float a, float b) { return (1 - t) * a + t * b; };
    HostFilterTable T;
    T.nx = int(32 * rx); T.ny = int(32 * ry);
    float expX = gaussian(rx), expY = gaussian(ry);
    T.f.resize((size_t)T.nx * T.ny); T.condCdf.resize((size_t)T.ny * (T.nx + 1)); T.condInt.resize(T.ny); T.margCdf.resize(T.ny + 1);
    for (int y = 0; y < T.ny; ++y)
        for (int x = 0; x < T.nx; ++x) {
            float px = lerp((x + 0.5f) / T.nx, -rx, rx), py = lerp((y + 0.5f) / T.ny, -ry, ry);
            T.f[(size_t)y * T.nx + x] = std::max(0.f, gaussian(px) - expX) * std::max(0.f, gaussian(py) - expY);
        }
    for (int v = 0; v < T.ny; ++v) piecewise_cdf(&T.f[(size_t)v * T.nx], T.nx, -rx, rx, &T.condCdf[(size_t)v * (T.nx + 1)], &T.condInt[v]);
    piecewise_cdf(T.condInt.data(), T.ny, -ry, ry, T.margCdf.data(), &T.margInt);
    return T;
}

// ------------------------------------------------------------------ threaded BVH over quads
struct HostBvhNode { float bmin[3]; int skip; float bmax[3]; int prim_begin; int prim_count; int pad[3]; };
struct HostBvh { std::vector<HostBvhNode> nodes; std::vector<int> prim_ids; };

inline void bvh_build_rec(HostBvh &out, std::vector<int> &ids, int begin, int end, const std::vector<float> &bmin,
                          const std::vector<float> &bmax, int leafSize) {
    HostBvhNode node;
    for (int a = 0; a < 3; a++) { node.bmin[a] = std::numeric_limits<float>::infinity(); node.bmax[a] = -std::numeric_limits<float>::infinity(); }
    float cmin[3] = {node.bmin[0], node.bmin[1], node.bmin[2]}, cmax[3] = {node.bmax[0], node.bmax[1], node.bmax[2]};
    for (int i = begin; i < end; i++)
        for (int a = 0; a < 3; a++) {
            node.bmin[a] = std::min(node.bmin[a], bmin[3 * ids[i] + a]);
            node.bmax[a] = std::max(node.bmax[a], bmax[3 * ids[i] + a]);
            float c = 0.5f * (bmin[3 * ids[i] + a] + bmax[3 * ids[i] + a]);
            cmin[a] = std::min(cmin[a], c); cmax[a] = std::max(cmax[a], c);
        }
    node.pad[0] = node.pad[1] = node.pad[2] = 0;
    int me = (int)out.nodes.size();
    out.nodes.push_back(node);
    int axis = 0;
    if (cmax[1] - cmin[1] > cmax[axis] - cmin[axis]) axis = 1;
    if (cmax[2] - cmin[2] > cmax[axis] - cmin[axis]) axis = 2;
    if (end - begin <= leafSize || !(cmax[axis] > cmin[axis])) {
        out.nodes[me].prim_begin = (int)out.prim_ids.size();
        out.nodes[me].prim_count = end - begin;
        std::sort(ids.begin() + begin, ids.begin() + end);
        for (int i = begin; i < end; i++) out.prim_ids.push_back(ids[i]);
    } else {
        int mid = (begin + end) / 2;
        std::nth_element(ids.begin() + begin, ids.begin() + mid, ids.begin() + end, [&](int a, int b) {
            float ca = bmin[3 * a + axis] + bmax[3 * a + axis], cb = bmin[3 * b + axis] + bmax[3 * b + axis];
            return ca < cb || (ca == cb && a < b);
        });
        out.nodes[me].prim_begin = -1; out.nodes[me].prim_count = 0;
        bvh_build_rec(out, ids, begin, mid, bmin, bmax, leafSize);
        bvh_build_rec(out, ids, mid, end, bmin, bmax, leafSize);
    }
    out.nodes[me].skip = (int)out.nodes.size();  // DFS order: everything after this subtree
}
inline HostBvh build_bvh(const vspg_quad *quads, int n) {
    HostBvh out;
    if (n == 0) return out;
    std::vector<float> bmin(3 * n), bmax(3 * n);
    std::vector<int> ids(n);
    for (int i = 0; i < n; i++) {
        ids[i] = i;
        const float *P[4] = {quads[i].p00, quads[i].p10, quads[i].p01, quads[i].p11};
        for (int a = 0; a < 3; a++) {
            bmin[3 * i + a] = std::min({P[0][a], P[1][a], P[2][a], P[3][a]});
            bmax[3 * i + a] = std::max({P[0][a], P[1][a], P[2][a], P[3][a]});
        }
    }
    bvh_build_rec(out, ids, 0, n, bmin, bmax, 2);
    return out;
}

}  // namespace vspg_host
