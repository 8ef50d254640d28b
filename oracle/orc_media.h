// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  (see orc_math.h header)
// CPU restatement of the reference's media: majorant grid construction, DDA majorant iterator,
// trilinear density lookup and the three SampleT_maj variants of the fork.
// Paths are relative to /root/reference/src/pbrt/.
#pragma once
#include <cstring>
#include <mutex>
#include <vector>

#include "../include/vspg.h"
#include "orc_math.h"

namespace orc {

struct Ray {
    V3 o, d;
    int medium = -1;
    V3 at(float t) const { return o + d * t; }  // ray.h operator()
};

struct Bounds3 {
    V3 pMin, pMax;
    V3 Offset(V3 p) const {  // util/vecmath.h Bounds3::Offset
        V3 o = p - pMin;
        if (pMax.x > pMin.x) o.x /= pMax.x - pMin.x;
        if (pMax.y > pMin.y) o.y /= pMax.y - pMin.y;
        if (pMax.z > pMin.z) o.z /= pMax.z - pMin.z;
        return o;
    }
    V3 Diagonal() const { return pMax - pMin; }
    // util/vecmath.h:1547-1571
    bool IntersectP(V3 o, V3 d, float tMax, float *hitt0, float *hitt1) const {
        float t0 = 0, t1 = tMax;
        for (int i = 0; i < 3; ++i) {
            float invRayDir = 1 / d[i];
            float tNear = (pMin[i] - o[i]) * invRayDir;
            float tFar = (pMax[i] - o[i]) * invRayDir;
            if (tNear > tFar) std::swap(tNear, tFar);
            tFar *= 1 + 2 * gamma_(3);
            t0 = tNear > t0 ? tNear : t0;
            t1 = tFar < t1 ? tFar : t1;
            if (t0 > t1) return false;
        }
        if (hitt0) *hitt0 = t0;
        if (hitt1) *hitt1 = t1;
        return true;
    }
};

// util/transform.cpp:263-303 + util/transform.h:416-429  Transform::ApplyInverse(const Ray&, Float *tMax)
inline void ApplyInverseRay(const vspg_transform &T, V3 ro, V3 rd, float *tMax, V3 *oOut, V3 *dOut) {
    const float(*mInv)[4] = T.mInv;
    float x = ro.x, y = ro.y, z = ro.z;
    float xp = (mInv[0][0] * x + mInv[0][1] * y) + (mInv[0][2] * z + mInv[0][3]);
    float yp = (mInv[1][0] * x + mInv[1][1] * y) + (mInv[1][2] * z + mInv[1][3]);
    float zp = (mInv[2][0] * x + mInv[2][1] * y) + (mInv[2][2] * z + mInv[2][3]);
    float wp = (mInv[3][0] * x + mInv[3][1] * y) + (mInv[3][2] * z + mInv[3][3]);
    V3 err;
    err.x = gamma_(3) * (std::abs(mInv[0][0] * x) + std::abs(mInv[0][1] * y) + std::abs(mInv[0][2] * z));
    err.y = gamma_(3) * (std::abs(mInv[1][0] * x) + std::abs(mInv[1][1] * y) + std::abs(mInv[1][2] * z));
    err.z = gamma_(3) * (std::abs(mInv[2][0] * x) + std::abs(mInv[2][1] * y) + std::abs(mInv[2][2] * z));
    // Point3fi(p, e) stores [SubRoundDown(p,e), AddRoundUp(p,e)]; Width()/2 of that interval (util/math.h Interval):
    auto halfWidth = [](float v, float e) {
        float lo, hi;
        if (e == 0) { lo = hi = v; }
        else { lo = NextFloatDown(v - e); hi = NextFloatUp(v + e); }
        return (hi - lo) / 2;
    };
    auto mid = [](float v, float e) {
        if (e == 0) return v;
        float lo = NextFloatDown(v - e), hi = NextFloatUp(v + e);
        return (lo + hi) / 2;
    };
    V3 o(xp, yp, zp);
    if (wp != 1) { /* affine transforms only (wp == 1); projective medium transforms are out of scope */ }
    // Vector: util/transform.h:401-406
    V3 d(mInv[0][0] * rd.x + mInv[0][1] * rd.y + mInv[0][2] * rd.z, mInv[1][0] * rd.x + mInv[1][1] * rd.y + mInv[1][2] * rd.z,
         mInv[2][0] * rd.x + mInv[2][1] * rd.y + mInv[2][2] * rd.z);
    float lengthSquared = LengthSquared(d);
    V3 om(mid(o.x, err.x), mid(o.y, err.y), mid(o.z, err.z));
    if (lengthSquared > 0) {
        V3 oError(halfWidth(o.x, err.x), halfWidth(o.y, err.y), halfWidth(o.z, err.z));
        float dt = Dot(Abs(d), oError) / lengthSquared;
        // o += d * dt on the interval, then Point3f(o) takes interval midpoints.  Interval + exact value keeps
        // the same half width up to rounding; we follow the midpoint arithmetic.
        V3 lo(NextFloatDown(o.x - err.x), NextFloatDown(o.y - err.y), NextFloatDown(o.z - err.z));
        V3 hi(NextFloatUp(o.x + err.x), NextFloatUp(o.y + err.y), NextFloatUp(o.z + err.z));
        if (err.x == 0) lo.x = hi.x = o.x;
        if (err.y == 0) lo.y = hi.y = o.y;
        if (err.z == 0) lo.z = hi.z = o.z;
        V3 a = d * dt;
        // Interval operator+(Interval, Interval{exact}) = {SubRoundDown? no: AddRoundDown(lo, a), AddRoundUp(hi, a)}
        for (int i = 0; i < 3; i++) {
            float l = lo[i] + a[i], h = hi[i] + a[i];
            l = NextFloatDown(l); h = NextFloatUp(h);
            om[i] = (l + h) / 2;
        }
        if (tMax) *tMax -= dt;
    }
    *oOut = om;
    *dOut = d;
}
inline V3 ApplyInversePoint(const vspg_transform &T, V3 p) {  // util/transform.h Transform::ApplyInverse(Point3<T>)
    const float(*mInv)[4] = T.mInv;
    float x = p.x, y = p.y, z = p.z;
    float xp = (mInv[0][0] * x + mInv[0][1] * y) + (mInv[0][2] * z + mInv[0][3]);
    float yp = (mInv[1][0] * x + mInv[1][1] * y) + (mInv[1][2] * z + mInv[1][3]);
    float zp = (mInv[2][0] * x + mInv[2][1] * y) + (mInv[2][2] * z + mInv[2][3]);
    float wp = (mInv[3][0] * x + mInv[3][1] * y) + (mInv[3][2] * z + mInv[3][3]);
    if (wp == 1) return {xp, yp, zp};
    return V3(xp, yp, zp) / wp;
}

struct MediumProps { S3 sigma_a, sigma_s; float g; S3 Le = S3(0.f); };
struct Segment { float tMin, tMax; S3 sigma_maj; };

struct Medium {
    const vspg_medium *m = nullptr;
    std::vector<float> ownMaj;
    const float *maj = nullptr;
    Bounds3 bounds;
    bool IsHomogeneous() const { return m->type == VSPG_MEDIUM_HOMOGENEOUS; }

    float gridAt(int x, int y, int z) const {  // util/containers.h:826-831 Lookup(Point3i)
        if (x < 0 || y < 0 || z < 0 || x >= m->nx || y >= m->ny || z >= m->nz) return 0.f;
        return m->density[(size_t)(z * m->ny + y) * m->nx + x];
    }
    S3 rgbAt(const float *g, int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= m->nx || y >= m->ny || z >= m->nz) return S3(0.f);
        return S3(g + 3 * ((size_t)(z * m->ny + y) * m->nx + x));
    }
    // util/containers.h:804-818 SampledGrid<Float>::Lookup(Point3f)
    float densityLookup(V3 p) const {
        V3 ps(p.x * m->nx - .5f, p.y * m->ny - .5f, p.z * m->nz - .5f);
        int ix = (int)std::floor(ps.x), iy = (int)std::floor(ps.y), iz = (int)std::floor(ps.z);
        V3 d(ps.x - (float)ix, ps.y - (float)iy, ps.z - (float)iz);
        float d00 = Lerp(d.x, gridAt(ix, iy, iz), gridAt(ix + 1, iy, iz));
        float d10 = Lerp(d.x, gridAt(ix, iy + 1, iz), gridAt(ix + 1, iy + 1, iz));
        float d01 = Lerp(d.x, gridAt(ix, iy, iz + 1), gridAt(ix + 1, iy, iz + 1));
        float d11 = Lerp(d.x, gridAt(ix, iy + 1, iz + 1), gridAt(ix + 1, iy + 1, iz + 1));
        return Lerp(d.z, Lerp(d.y, d00, d10), Lerp(d.y, d01, d11));
    }
    S3 rgbLookup(const float *g, V3 p) const {  // containers.h:785-801 with convert = Sample(lambda) = rgb
        V3 ps(p.x * m->nx - .5f, p.y * m->ny - .5f, p.z * m->nz - .5f);
        int ix = (int)std::floor(ps.x), iy = (int)std::floor(ps.y), iz = (int)std::floor(ps.z);
        V3 d(ps.x - (float)ix, ps.y - (float)iy, ps.z - (float)iz);
        auto L = [](float t, S3 a, S3 b) { return (1 - t) * a + t * b; };
        S3 d00 = L(d.x, rgbAt(g, ix, iy, iz), rgbAt(g, ix + 1, iy, iz));
        S3 d10 = L(d.x, rgbAt(g, ix, iy + 1, iz), rgbAt(g, ix + 1, iy + 1, iz));
        S3 d01 = L(d.x, rgbAt(g, ix, iy, iz + 1), rgbAt(g, ix + 1, iy, iz + 1));
        S3 d11 = L(d.x, rgbAt(g, ix, iy + 1, iz + 1), rgbAt(g, ix + 1, iy + 1, iz + 1));
        return L(d.z, L(d.y, d00, d10), L(d.y, d01, d11));
    }
    // ReadAccessor::getValue on the flattened grid: voxels outside the index bounding box read the background
    float nvdbAt(int i, int j, int k) const {
        int x = i - m->index_min[0], y = j - m->index_min[1], z = k - m->index_min[2];
        if (x < 0 || y < 0 || z < 0 || x >= m->nx || y >= m->ny || z >= m->nz) return m->background;
        return m->density[(size_t)(z * m->ny + y) * m->nx + x];
    }
    // media.h:699-703: worldToIndexF (NanoVDB.h:2028-2031 applyInverseMapF, :1408-1414 matMult with explicit fmaf) and
    // SampleFromVoxels<Tree, 1, false>::operator() (util/SampleFromVoxels.h:57-65 Floor, :421-463)
    float nvdbLookup(V3 p) const {
        const float *mt = m->index_from_medium;
        float x = p.x - m->index_origin[0], y = p.y - m->index_origin[1], z = p.z - m->index_origin[2];
        float xyz[3] = {std::fmaf(x, mt[0], std::fmaf(y, mt[1], z * mt[2])), std::fmaf(x, mt[3], std::fmaf(y, mt[4], z * mt[5])),
                        std::fmaf(x, mt[6], std::fmaf(y, mt[7], z * mt[8]))};
        const float fl[3] = {std::floor(xyz[0]), std::floor(xyz[1]), std::floor(xyz[2])};
        for (int a = 0; a < 3; a++) xyz[a] -= fl[a];
        int c[3] = {(int)fl[0], (int)fl[1], (int)fl[2]};
        auto lerp = [](float a, float b, float w) { return a + w * (b - a); };
        float vx, vx1, vy, vy1, vz, vz1;
        vz = nvdbAt(c[0], c[1], c[2]); c[2] += 1; vz1 = nvdbAt(c[0], c[1], c[2]); vy = lerp(vz, vz1, xyz[2]);
        c[1] += 1;
        vz1 = nvdbAt(c[0], c[1], c[2]); c[2] -= 1; vz = nvdbAt(c[0], c[1], c[2]); vy1 = lerp(vz, vz1, xyz[2]);
        vx = lerp(vy, vy1, xyz[1]);
        c[0] += 1;
        vz = nvdbAt(c[0], c[1], c[2]); c[2] += 1; vz1 = nvdbAt(c[0], c[1], c[2]); vy1 = lerp(vz, vz1, xyz[2]);
        c[1] -= 1;
        vz1 = nvdbAt(c[0], c[1], c[2]); c[2] -= 1; vz = nvdbAt(c[0], c[1], c[2]); vy = lerp(vz, vz1, xyz[2]);
        vx1 = lerp(vy, vy1, xyz[1]);
        return lerp(vx, vx1, xyz[0]);
    }
    bool IsEmissive() const { return std::max({m->Le[0], m->Le[1], m->Le[2]}) > 0; }   // media.h:250, media.cpp:261
    // util/containers.h:804-831 Lookup on the "Lescale" grid (a 1x1x1 grid of 1 when le_scale is NULL, media.cpp:319-320)
    float leScaleLookup(V3 p) const {
        static const float one = 1.f;
        const float *g = m->le_scale ? m->le_scale : &one;
        const int nx = m->le_scale ? m->le_nx : 1, ny = m->le_scale ? m->le_ny : 1, nz = m->le_scale ? m->le_nz : 1;
        auto at = [&](int x, int y, int z) {
            if (x < 0 || y < 0 || z < 0 || x >= nx || y >= ny || z >= nz) return 0.f;
            return g[(size_t)(z * ny + y) * nx + x];
        };
        V3 ps(p.x * nx - .5f, p.y * ny - .5f, p.z * nz - .5f);
        int ix = (int)std::floor(ps.x), iy = (int)std::floor(ps.y), iz = (int)std::floor(ps.z);
        V3 d(ps.x - (float)ix, ps.y - (float)iy, ps.z - (float)iz);
        float d00 = Lerp(d.x, at(ix, iy, iz), at(ix + 1, iy, iz));
        float d10 = Lerp(d.x, at(ix, iy + 1, iz), at(ix + 1, iy + 1, iz));
        float d01 = Lerp(d.x, at(ix, iy, iz + 1), at(ix + 1, iy, iz + 1));
        float d11 = Lerp(d.x, at(ix, iy + 1, iz + 1), at(ix + 1, iy + 1, iz + 1));
        return Lerp(d.z, Lerp(d.y, d00, d10), Lerp(d.y, d01, d11));
    }
    // media.h:256-261, 316-348, 417-441, 692-705 SamplePoint (emission: homogeneous and uniformgrid "Le" / "Lescale")
    MediumProps SamplePoint(V3 p) const {
        MediumProps mp; mp.g = m->g;
        if (m->type == VSPG_MEDIUM_HOMOGENEOUS) { mp.sigma_a = S3(m->sigma_a); mp.sigma_s = S3(m->sigma_s); mp.Le = S3(m->Le); return mp; }
        p = ApplyInversePoint(m->render_from_medium, p);
        if (m->type == VSPG_MEDIUM_NANOVDB) {
            float d = nvdbLookup(p);
            d += m->density_offset;
            mp.sigma_a = S3(m->sigma_a) * d; mp.sigma_s = S3(m->sigma_s) * d;
            return mp;
        }
        p = bounds.Offset(p);
        if (m->type == VSPG_MEDIUM_GRID) {
            float d = densityLookup(p);
            mp.sigma_a = S3(m->sigma_a) * d; mp.sigma_s = S3(m->sigma_s) * d;
            if (IsEmissive()) {   // media.h:329-345
                float scale = leScaleLookup(p);
                if (scale > 0) mp.Le = scale * S3(m->Le);
            }
        } else {
            mp.sigma_a = m->sigma_scale * (m->sigma_a_rgb ? rgbLookup(m->sigma_a_rgb, p) : S3(1.f));
            mp.sigma_s = m->sigma_scale * (m->sigma_s_rgb ? rgbLookup(m->sigma_s_rgb, p) : S3(1.f));
        }
        return mp;
    }
    // containers.h:838-855 SampledGrid::MaxValue(bounds, convert)
    template <typename F>
    float MaxValue(V3 bmin, V3 bmax, F at) const {
        V3 ps0(bmin.x * m->nx - .5f, bmin.y * m->ny - .5f, bmin.z * m->nz - .5f);
        V3 ps1(bmax.x * m->nx - .5f, bmax.y * m->ny - .5f, bmax.z * m->nz - .5f);
        int lo[3] = {std::max((int)std::floor(ps0.x), 0), std::max((int)std::floor(ps0.y), 0), std::max((int)std::floor(ps0.z), 0)};
        int hi[3] = {std::min((int)std::floor(ps1.x) + 1, m->nx - 1), std::min((int)std::floor(ps1.y) + 1, m->ny - 1),
                     std::min((int)std::floor(ps1.z) + 1, m->nz - 1)};
        float maxValue = at(lo[0], lo[1], lo[2]);
        for (int z = lo[2]; z <= hi[2]; ++z)
            for (int y = lo[1]; y <= hi[1]; ++y)
                for (int x = lo[0]; x <= hi[0]; ++x) maxValue = std::max(maxValue, at(x, y, z));
        return maxValue;
    }
    // media.cpp:256-269 (GridMedium) and media.cpp:393-408 (RGBGridMedium)
    // media.cpp:605-670 NanoVDBMedium's majorant grid: per cell, the maximum over the index-space box of the cell +- 1 voxel
    void BuildMajorantNanoVDB(float *out) const {
        const int rx = m->maj_res[0], ry = m->maj_res[1], rz = m->maj_res[2];
        const int bmax[3] = {m->index_min[0] + m->nx - 1, m->index_min[1] + m->ny - 1, m->index_min[2] + m->nz - 1};
        auto toIndex = [&](V3 w, double *o) {   // worldToIndexF(Vec3R): the float map applied to double coordinates
            double d[3] = {(double)w.x - (double)m->index_origin[0], (double)w.y - (double)m->index_origin[1], (double)w.z - (double)m->index_origin[2]};
            for (int r = 0; r < 3; r++) {
                const float *mt = m->index_from_medium + 3 * r;
                o[r] = (double)fmaf((float)d[0], mt[0], fmaf((float)d[1], mt[1], (float)(d[2] * (double)mt[2])));
            }
        };
        for (int z = 0; z < rz; ++z)
            for (int y = 0; y < ry; ++y)
                for (int x = 0; x < rx; ++x) {
                    V3 t0(float(x) / rx, float(y) / ry, float(z) / rz), t1(float(x + 1) / rx, float(y + 1) / ry, float(z + 1) / rz);
                    V3 w0(Lerp(t0.x, bounds.pMin.x, bounds.pMax.x), Lerp(t0.y, bounds.pMin.y, bounds.pMax.y), Lerp(t0.z, bounds.pMin.z, bounds.pMax.z));
                    V3 w1(Lerp(t1.x, bounds.pMin.x, bounds.pMax.x), Lerp(t1.y, bounds.pMin.y, bounds.pMax.y), Lerp(t1.z, bounds.pMin.z, bounds.pMax.z));
                    double i0[3], i1[3];
                    toIndex(w0, i0); toIndex(w1, i1);
                    float delta = 1.f;
                    int n0[3], n1[3];
                    for (int a = 0; a < 3; a++) { n0[a] = std::max(int(i0[a] - delta), m->index_min[a]); n1[a] = std::min(int(i1[a] + delta), bmax[a]); }
                    float maxValue = 0;
                    for (int nz = n0[2]; nz <= n1[2]; ++nz)
                        for (int ny = n0[1]; ny <= n1[1]; ++ny)
                            for (int nx = n0[0]; nx <= n1[0]; ++nx) maxValue = std::max(maxValue, nvdbAt(nx, ny, nz));
                    out[x + rx * (y + ry * z)] = (maxValue + m->density_offset) * m->majorant_scale;
                }
    }
    void BuildMajorant(float *out) const {
        if (m->type == VSPG_MEDIUM_NANOVDB) { BuildMajorantNanoVDB(out); return; }
        int rx = m->maj_res[0], ry = m->maj_res[1], rz = m->maj_res[2];
        for (int z = 0; z < rz; ++z)
            for (int y = 0; y < ry; ++y)
                for (int x = 0; x < rx; ++x) {
                    V3 p0(float(x) / rx, float(y) / ry, float(z) / rz);          // media.h:126-130 VoxelBounds
                    V3 p1(float(x + 1) / rx, float(y + 1) / ry, float(z + 1) / rz);
                    float v;
                    if (m->type == VSPG_MEDIUM_GRID) v = MaxValue(p0, p1, [&](int a, int b, int c) { return gridAt(a, b, c); });
                    else {
                        float ma = m->sigma_a_rgb ? MaxValue(p0, p1, [&](int a, int b, int c) { return rgbAt(m->sigma_a_rgb, a, b, c).MaxComponentValue(); }) : 1;
                        float ms = m->sigma_s_rgb ? MaxValue(p0, p1, [&](int a, int b, int c) { return rgbAt(m->sigma_s_rgb, a, b, c).MaxComponentValue(); }) : 1;
                        v = m->sigma_scale * (ma + ms);
                    }
                    out[x + rx * (y + ry * z)] = v;
                }
    }
    void init(const vspg_medium *mm, bool ownMajorant) {
        m = mm;
        bounds.pMin = V3(m->bounds_min); bounds.pMax = V3(m->bounds_max);
        maj = m->majorant;
        if (m->type != VSPG_MEDIUM_HOMOGENEOUS && (ownMajorant || !maj)) {
            const size_t n = (size_t)m->maj_res[0] * m->maj_res[1] * m->maj_res[2];
            if (m->type == VSPG_MEDIUM_NANOVDB) {
                // the per-sample entry points build a Scene per call: keep the last 64^3 grid (0.1 s to build) while the
                // medium it was built from is unchanged
                static std::mutex mu;
                static std::vector<float> cached;
                static vspg_medium key;
                std::lock_guard<std::mutex> lock(mu);
                vspg_medium k = *m; k.majorant = nullptr; k.g = 0; memset(k.sigma_a, 0, sizeof k.sigma_a); memset(k.sigma_s, 0, sizeof k.sigma_s);
                memset(&k.render_from_medium, 0, sizeof k.render_from_medium);
                if (cached.size() != n || memcmp(&k, &key, sizeof k) != 0) {
                    cached.assign(n, 0.f);
                    BuildMajorant(cached.data());
                    key = k;
                }
                ownMaj = cached;
            } else {
                ownMaj.resize(n);
                BuildMajorant(ownMaj.data());
            }
            maj = ownMaj.data();
        }
    }
};

// media.h:140-218 DDAMajorantIterator and media.h:76-107 HomogeneousMajorantIterator
struct MajorantIterator {
    bool homogeneous = false, called = false;
    Segment hseg;
    S3 sigma_t;
    float tMin = Infinity, tMax = -Infinity;
    const float *grid = nullptr;
    int res[3] = {0, 0, 0};
    float nextCrossingT[3], deltaT[3];
    int step[3], voxelLimit[3], voxel[3];
    long steps = 0;

    void initDDA(V3 ro, V3 rd, float tMin_, float tMax_, const Medium &med, S3 sigma_t_) {
        tMin = tMin_; tMax = tMax_; grid = med.maj; sigma_t = sigma_t_;
        for (int i = 0; i < 3; i++) res[i] = med.m->maj_res[i];
        V3 diag = med.bounds.Diagonal();
        V3 go = med.bounds.Offset(ro);
        V3 gd(rd.x / diag.x, rd.y / diag.y, rd.z / diag.z);
        V3 gridIntersect = go + gd * tMin;
        for (int axis = 0; axis < 3; ++axis) {
            voxel[axis] = Clamp((int)(gridIntersect[axis] * res[axis]), 0, res[axis] - 1);
            deltaT[axis] = 1 / (std::abs(gd[axis]) * res[axis]);
            if (gd[axis] == -0.f) gd[axis] = 0.f;
            if (gd[axis] >= 0) {
                float nextVoxelPos = float(voxel[axis] + 1) / res[axis];
                nextCrossingT[axis] = tMin + (nextVoxelPos - gridIntersect[axis]) / gd[axis];
                step[axis] = 1; voxelLimit[axis] = res[axis];
            } else {
                float nextVoxelPos = float(voxel[axis]) / res[axis];
                nextCrossingT[axis] = tMin + (nextVoxelPos - gridIntersect[axis]) / gd[axis];
                step[axis] = -1; voxelLimit[axis] = -1;
            }
        }
    }
    bool Next(Segment *seg) {
        steps++;
        if (homogeneous) {
            if (called) return false;
            called = true; *seg = hseg; return true;
        }
        if (tMin >= tMax) return false;
        int bits = ((nextCrossingT[0] < nextCrossingT[1]) << 2) + ((nextCrossingT[0] < nextCrossingT[2]) << 1) +
                   ((nextCrossingT[1] < nextCrossingT[2]));
        const int cmpToAxis[8] = {2, 1, 2, 1, 2, 2, 0, 0};
        int stepAxis = cmpToAxis[bits];
        float tVoxelExit = std::min(tMax, nextCrossingT[stepAxis]);
        S3 sigma_maj = sigma_t * grid[voxel[0] + res[0] * (voxel[1] + res[1] * voxel[2])];
        *seg = Segment{tMin, tVoxelExit, sigma_maj};
        tMin = tVoxelExit;
        if (nextCrossingT[stepAxis] > tMax) tMin = tMax;
        voxel[stepAxis] += step[stepAxis];
        if (voxel[stepAxis] == voxelLimit[stepAxis]) tMin = tMax;
        nextCrossingT[stepAxis] += deltaT[stepAxis];
        return true;
    }
};

// media.h:264-269, 351-366, 444-455  Medium::SampleRay
inline MajorantIterator SampleRay(const Medium &med, Ray ray, float raytMax) {
    MajorantIterator it;
    const vspg_medium *m = med.m;
    if (m->type == VSPG_MEDIUM_HOMOGENEOUS) {
        it.homogeneous = true;
        it.hseg = Segment{0.f, raytMax, S3(m->sigma_a) + S3(m->sigma_s)};
        return it;
    }
    V3 o, d;
    ApplyInverseRay(m->render_from_medium, ray.o, ray.d, &raytMax, &o, &d);
    float tMin, tMax;
    if (!med.bounds.IntersectP(o, d, raytMax, &tMin, &tMax)) return it;  // default iterator: tMin=inf >= tMax=-inf
    S3 sigma_t = (m->type == VSPG_MEDIUM_GRID || m->type == VSPG_MEDIUM_NANOVDB) ? (S3(m->sigma_a) + S3(m->sigma_s)) : S3(1.f);
    it.initDDA(o, d, tMin, tMax, med, sigma_t);
    return it;
}

struct TrackStats { long queries = 0, dda = 0; };

// media_sampleTMaj.h:49-117.  callback(p, mp, sigma_maj, T_maj) -> bool
template <typename F>
S3 SampleT_maj(const Medium &med, Ray ray, float tMax, float u, RNG &rng, int ch, TrackStats *st, F callback) {
    tMax *= Length(ray.d);
    ray.d = Normalize(ray.d);
    MajorantIterator iter = SampleRay(med, ray, tMax);
    S3 T_maj(1.f);
    bool done = false;
    Segment seg;
    while (!done) {
        if (!iter.Next(&seg)) { if (st) st->dda += iter.steps; return T_maj; }
        if (seg.sigma_maj[ch] == 0) {
            float dt = seg.tMax - seg.tMin;
            if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
            T_maj *= FastExp(-dt * seg.sigma_maj);
            continue;
        }
        float tMin = seg.tMin;
        while (true) {
            float t = tMin + SampleExponential(u, seg.sigma_maj[ch]);
            u = rng.Uniform();
            if (t < seg.tMax) {
                T_maj *= FastExp(-(t - tMin) * seg.sigma_maj);
                MediumProps mp = med.SamplePoint(ray.at(t));
                if (st) st->queries++;
                if (!callback(ray.at(t), mp, seg.sigma_maj, T_maj)) { done = true; break; }
                T_maj = S3(1.f);
                tMin = t;
            } else {
                float dt = seg.tMax - tMin;
                if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
                T_maj *= FastExp(-dt * seg.sigma_maj);
                break;
            }
        }
    }
    if (st) st->dda += iter.steps;
    return S3(1.f);
}

// media_sampleTMaj.h:136-248
template <typename F>
S3 SampleT_maj_Resampling(const Medium &med, Ray ray, float tMax, float u, RNG &rng, int ch, bool guideScatterDecision,
                          float volScatterProb, float &volumeRatioZeroCandidateCompensation, float &majorantScale,
                          TrackStats *st, F callback) {
    tMax *= Length(ray.d);
    ray.d = Normalize(ray.d);
    MajorantIterator iter = SampleRay(med, ray, tMax);
    MajorantIterator pre = iter;
    float totalLength = 0.f;
    Segment seg;
    while (true) {
        if (!pre.Next(&seg)) break;
        if (seg.sigma_maj[ch] == 0) continue;
        totalLength += seg.sigma_maj[ch] * (seg.tMax - seg.tMin);
    }
    if (st) st->dda += pre.steps;
    if (totalLength == 0.f) return S3(1.f);
    majorantScale = 1.0f;
    volumeRatioZeroCandidateCompensation = volScatterProb;
    if (guideScatterDecision) {
        float minTotalLength = -vspg_logf(1 - volScatterProb);
        if (minTotalLength > totalLength) {
            majorantScale = minTotalLength / totalLength;
            totalLength = minTotalLength;
        }
        float expNegTotalLength = FastExp(-totalLength);
        volumeRatioZeroCandidateCompensation = volScatterProb / (1 - expNegTotalLength);
    }
    S3 T_maj(1.f);
    bool done = false;
    int count = 0;
    iter.steps = 0;
    while (!done) {
        if (!iter.Next(&seg)) { if (st) st->dda += iter.steps; return T_maj; }
        seg.sigma_maj *= majorantScale;
        if (seg.sigma_maj[ch] == 0) {
            float dt = seg.tMax - seg.tMin;
            if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
            T_maj *= FastExp(-dt * seg.sigma_maj);
            continue;
        }
        float tMin = seg.tMin;
        while (true) {
            count++;
            float t = tMin + SampleExponential(u, seg.sigma_maj[ch]);
            u = rng.Uniform();
            if (t < seg.tMax) {
                if (count > 10000) break;
                T_maj *= FastExp(-(t - tMin) * seg.sigma_maj);
                MediumProps mp = med.SamplePoint(ray.at(t));
                if (st) st->queries++;
                if (!callback(ray.at(t), mp, seg.sigma_maj, T_maj)) { done = true; break; }
                T_maj = S3(1.f);
                tMin = t;
            } else {
                float dt = seg.tMax - tMin;
                if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
                T_maj *= FastExp(-dt * seg.sigma_maj);
                break;
            }
        }
    }
    if (st) st->dda += iter.steps;
    return S3(1.f);
}

// media_sampleTMaj.h:269-491.  callback(p, mp, sigma_maj, T_maj, activateNDS) -> bool
template <typename F>
S3 SampleT_maj_OpticalDepthSpace(const Medium &med, Ray ray, float tMax, float u, RNG &rng, int ch, bool guideScatterDecision,
                                 float vsp, float vspMISRatio, bool NDS, S3 &beta_factor, S3 &r_u_factor, TrackStats *st,
                                 F callback) {
    auto plain = [&](const V3 &p, const MediumProps &mp, const S3 &sm, const S3 &tm) { return callback(p, mp, sm, tm, false); };
    if (!guideScatterDecision || vspMISRatio == 0.f) return SampleT_maj(med, ray, tMax, u, rng, ch, st, plain);
    tMax *= Length(ray.d);
    ray.d = Normalize(ray.d);
    MajorantIterator iter = SampleRay(med, ray, tMax);
    MajorantIterator pre = iter;
    float t_v = 0.f;
    Segment seg;
    while (true) {
        if (!pre.Next(&seg)) break;
        if (std::isinf(seg.tMax)) { if (st) st->dda += pre.steps; return SampleT_maj(med, ray, tMax, u, rng, ch, st, plain); }
        if (seg.sigma_maj[ch] == 0) continue;
        t_v += seg.sigma_maj[ch] * (seg.tMax - seg.tMin);
    }
    if (st) st->dda += pre.steps;
    if (t_v == 0.f) return S3(1.f);
    float OneMinusENegTv = 1.f - FastExp(-t_v);
    float t_n = -1.f, t_n_current = -1.f;
    if (NDS) {
        if (vsp < 1 - FastExp(-t_v)) return SampleT_maj(med, ray, tMax, u, rng, ch, st, plain);  // (already-normalised ray, as in the reference)
        else {
            t_n = (float)(-vspg_log(1.0 - OneMinusENegTv / vsp));  // double arithmetic as in the reference (1.0 literal)
            t_n_current = t_n;
        }
    }
    S3 T_maj(1.f), tpScaleFactor(1.f);
    float t_v_current = t_v;
    float remainingDist = 0;
    bool deltaTracking = false;
    if (u > vspMISRatio) { deltaTracking = true; u = (u - vspMISRatio) / (1 - vspMISRatio); }
    else u /= vspMISRatio;
    bool done = false;
    int count = 0;
    bool overTheEnd = false;
    const float ScatterEpsilon = 1e-5;
    iter.steps = 0;
    while (!done) {
        if (!iter.Next(&seg)) { if (st) st->dda += iter.steps; return T_maj; }
        if (seg.sigma_maj[ch] == 0 || overTheEnd) {
            float dt = seg.tMax - seg.tMin;
            if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
            T_maj *= FastExp(-dt * seg.sigma_maj);
            continue;
        }
        float tMin = seg.tMin;
        S3 normalizedMaj = seg.sigma_maj / seg.sigma_maj[ch];
        if (remainingDist > 0) {
            tMin += remainingDist / seg.sigma_maj[ch];
            if (tMin > seg.tMax + ScatterEpsilon) {
                float dist = (seg.tMax - seg.tMin) * seg.sigma_maj[ch];
                t_v_current -= dist; t_n_current -= dist; remainingDist -= dist;
                T_maj *= FastExp(-(seg.tMax - seg.tMin) * seg.sigma_maj);
                continue;
            }
            t_v_current -= remainingDist; t_n_current -= remainingDist; remainingDist = 0;
            T_maj *= FastExp(-(tMin - seg.tMin) * seg.sigma_maj);
            MediumProps mp = med.SamplePoint(ray.at(tMin));
            if (st) st->queries++;
            r_u_factor = S3(vspMISRatio) / tpScaleFactor + S3(1 - vspMISRatio);
            if (!callback(ray.at(tMin), mp, seg.sigma_maj, T_maj, true)) break;
            T_maj = S3(1.f);
        }
        while (true) {
            count++;
            float dist = std::numeric_limits<float>::max();
            S3 tpSingle;
            if (NDS) {
                tpSingle = S3(1.0f) - FastExp(-t_n_current * normalizedMaj);
                if (!deltaTracking) dist = (float)(-vspg_log(1.0 - u * tpSingle[ch]));
            } else {
                tpSingle = (S3(1.0f) - FastExp(-t_v_current * normalizedMaj)) / vsp;
                if (!deltaTracking) {
                    if (u < vsp) dist = (float)(-vspg_log(1.0 - u * tpSingle[ch]));
                }
            }
            if (deltaTracking) dist = (float)(-vspg_log(1.0 - u));
            bool passThrough = t_v_current - dist < ScatterEpsilon || dist == 0;
            if (NDS || !passThrough) tpScaleFactor *= tpSingle;
            if (passThrough) {
                if (NDS) tpScaleFactor /= 1.0f - FastExp(-t_n + t_v);
                else tpScaleFactor *= FastExp(-t_v_current * normalizedMaj) / (1 - vsp);
                r_u_factor = S3(vspMISRatio) / tpScaleFactor + S3(1 - vspMISRatio);
                overTheEnd = true;
                T_maj *= FastExp(-(seg.tMax - tMin) * seg.sigma_maj);
                break;
            }
            float t = tMin + dist / seg.sigma_maj[ch];
            u = rng.Uniform();
            if (t <= seg.tMax + ScatterEpsilon) {
                if (count > 10000) break;
                t_v_current -= dist; t_n_current -= dist; remainingDist = 0;
                T_maj *= FastExp(-(t - tMin) * seg.sigma_maj);
                MediumProps mp = med.SamplePoint(ray.at(t));
                if (st) st->queries++;
                r_u_factor = S3(vspMISRatio) / tpScaleFactor + S3(1 - vspMISRatio);
                if (!callback(ray.at(t), mp, seg.sigma_maj, T_maj, true)) { done = true; break; }
                T_maj = S3(1.f);
                tMin = t;
            } else {
                float dt = seg.tMax - tMin;
                if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
                T_maj *= FastExp(-dt * seg.sigma_maj);
                float distWithinThisSeg = dt * seg.sigma_maj[ch];
                remainingDist = dist - distWithinThisSeg;
                t_v_current -= distWithinThisSeg; t_n_current -= distWithinThisSeg;
                break;
            }
        }
    }
    if (st) st->dda += iter.steps;
    return S3(1.f);
}

}  // namespace orc
