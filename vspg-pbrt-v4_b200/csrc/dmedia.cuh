// Participating media on the device: ray -> medium space, majorant-grid DDA, trilinear density fetch and the
// three majorant-space distance samplers of the fork.  Citations relative to /root/reference/src/pbrt.
#pragma once
#include "dscene.cuh"

namespace vspg {

struct Ray { V3 o, d; int medium; VD V3 at(float t) const { return o + d * t; } };
struct MediumProps { V3 sigma_a, sigma_s; float g; };
struct Segment { float tMin, tMax; V3 sigma_maj; };
struct TrackCounters { unsigned queries, dda; };

VD V3 ld3(const float *p) { return V3(p[0], p[1], p[2]); }

// util/vecmath.h Bounds3::Offset
VD V3 bounds_offset(const DMedium &m, V3 p) {
    V3 o(p.x - m.bmin[0], p.y - m.bmin[1], p.z - m.bmin[2]);
    if (m.extPow2 == 7) return V3(o.x * m.invExt[0], o.y * m.invExt[1], o.z * m.invExt[2]);   // extents 2^k: the product is the quotient, exactly
    if (m.bmax[0] > m.bmin[0]) o.x /= m.bmax[0] - m.bmin[0];
    if (m.bmax[1] > m.bmin[1]) o.y /= m.bmax[1] - m.bmin[1];
    if (m.bmax[2] > m.bmin[2]) o.z /= m.bmax[2] - m.bmin[2];
    return o;
}
// util/vecmath.h:1547-1571 Bounds3::IntersectP(o, d, tMax, &t0, &t1)
VD bool bounds_intersect(const DMedium &m, V3 o, V3 d, float tMax, float *hitt0, float *hitt1) {
    float t0 = 0, t1 = tMax;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float invRayDir = 1 / d[i];
        float tNear = (m.bmin[i] - o[i]) * invRayDir;
        float tFar = (m.bmax[i] - o[i]) * invRayDir;
        if (tNear > tFar) { float t = tNear; tNear = tFar; tFar = t; }
        tFar *= 1 + 2 * gammaN(3);
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return false;
    }
    *hitt0 = t0; *hitt1 = t1;
    return true;
}
// util/transform.cpp:263-303 + util/transform.h:416-429: Transform::ApplyInverse(const Ray&, Float *tMax).
// The origin is mapped as an interval (Point3fi), advanced along d by the error bound, and tMax shrinks by dt.
VD void apply_inverse_ray(const DMedium &m, V3 ro, V3 rd, float *tMax, V3 *oOut, V3 *dOut) {
    float o[3], err[3], lo[3], hi[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        float a = m.mInv[r][0] * ro.x, b = m.mInv[r][1] * ro.y, c = m.mInv[r][2] * ro.z;
        o[r] = (a + b) + (c + m.mInv[r][3]);
        err[r] = gammaN(3) * (fabsf(a) + fabsf(b) + fabsf(c));
        if (err[r] == 0) { lo[r] = hi[r] = o[r]; }
        else { lo[r] = next_down(o[r] - err[r]); hi[r] = next_up(o[r] + err[r]); }
    }
    V3 d(m.mInv[0][0] * rd.x + m.mInv[0][1] * rd.y + m.mInv[0][2] * rd.z,
         m.mInv[1][0] * rd.x + m.mInv[1][1] * rd.y + m.mInv[1][2] * rd.z,
         m.mInv[2][0] * rd.x + m.mInv[2][1] * rd.y + m.mInv[2][2] * rd.z);
    float lengthSquared = length_sq(d);
    V3 om((lo[0] + hi[0]) / 2, (lo[1] + hi[1]) / 2, (lo[2] + hi[2]) / 2);
    if (err[0] == 0) om.x = o[0];
    if (err[1] == 0) om.y = o[1];
    if (err[2] == 0) om.z = o[2];
    if (lengthSquared > 0) {
        V3 oError((hi[0] - lo[0]) / 2, (hi[1] - lo[1]) / 2, (hi[2] - lo[2]) / 2);
        float dt = dot(vabs(d), oError) / lengthSquared;
        V3 a = d * dt;
        float l, h;
        l = next_down(lo[0] + a.x); h = next_up(hi[0] + a.x); om.x = (l + h) / 2;
        l = next_down(lo[1] + a.y); h = next_up(hi[1] + a.y); om.y = (l + h) / 2;
        l = next_down(lo[2] + a.z); h = next_up(hi[2] + a.z); om.z = (l + h) / 2;
        *tMax -= dt;
    }
    *oOut = om; *dOut = d;
}
VD V3 apply_inverse_point(const DMedium &m, V3 p) {  // util/transform.h Transform::ApplyInverse(Point3)
    // pure translation: (1 * p.x + 0 * p.y) + (0 * p.z + t) is p.x + t in value (the zero products only ever decide the
    // sign of a zero result, which no consumer of the point distinguishes), and wp == 1
    if (m.xlateOnly) return V3(p.x + m.mInv[0][3], p.y + m.mInv[1][3], p.z + m.mInv[2][3]);
    float xp = (m.mInv[0][0] * p.x + m.mInv[0][1] * p.y) + (m.mInv[0][2] * p.z + m.mInv[0][3]);
    float yp = (m.mInv[1][0] * p.x + m.mInv[1][1] * p.y) + (m.mInv[1][2] * p.z + m.mInv[1][3]);
    float zp = (m.mInv[2][0] * p.x + m.mInv[2][1] * p.y) + (m.mInv[2][2] * p.z + m.mInv[2][3]);
    float wp = (m.mInv[3][0] * p.x + m.mInv[3][1] * p.y) + (m.mInv[3][2] * p.z + m.mInv[3][3]);
    if (wp == 1) return V3(xp, yp, zp);
    return V3(xp, yp, zp) / wp;
}

// ------------------------------------------------------------------ density fetch (util/containers.h:804-818)
VD float grid_at(const DMedium &m, int x, int y, int z) {
    if (x < 0 || y < 0 || z < 0 || x >= m.nx || y >= m.ny || z >= m.nz) return m.bg;
    return __ldg(&m.density[(size_t)(z * m.ny + y) * m.nx + x]);
}
VD float trilerp(float dx, float dy, float dz, float v000, float v100, float v010, float v110, float v001, float v101,
                 float v011, float v111) {
    float d00 = lerpf(dx, v000, v100);
    float d10 = lerpf(dx, v010, v110);
    float d01 = lerpf(dx, v001, v101);
    float d11 = lerpf(dx, v011, v111);
    return lerpf(dz, lerpf(dy, d00, d10), lerpf(dy, d01, d11));
}
// The 8 corners of interpolation cell (ix,iy,iz) (grid-local voxel indices): one aligned 32-byte record in the CELL8 layout.
struct DensityFetch { float4 a, b; float dx, dy, dz; };
VD void density_gather(const DMedium &m, int ix, int iy, int iz, float4 *a, float4 *b) {
    if (m.layout == GRID_CELL8) {
        // cells outside [-1,n-1] only see voxels outside the grid
        if (ix < -1 || iy < -1 || iz < -1 || ix >= m.nx || iy >= m.ny || iz >= m.nz) { *a = make_float4(m.bg, m.bg, m.bg, m.bg); *b = *a; return; }
        size_t cell = ((size_t)(iz + 1) * (m.ny + 1) + (iy + 1)) * (m.nx + 1) + (ix + 1);
        *a = __ldg(&m.cells[2 * cell]); *b = __ldg(&m.cells[2 * cell + 1]);
    } else {
        *a = make_float4(grid_at(m, ix, iy, iz), grid_at(m, ix + 1, iy, iz), grid_at(m, ix, iy + 1, iz), grid_at(m, ix + 1, iy + 1, iz));
        *b = make_float4(grid_at(m, ix, iy, iz + 1), grid_at(m, ix + 1, iy, iz + 1), grid_at(m, ix, iy + 1, iz + 1), grid_at(m, ix + 1, iy + 1, iz + 1));
    }
}
// util/containers.h:804-818 SampledGrid<Float>::Lookup(Point3f): p in [0,1]^3 medium-bounds coordinates
VD DensityFetch density_fetch_issue(const DMedium &m, V3 p) {
    DensityFetch f;
    float psx = p.x * m.nx - .5f, psy = p.y * m.ny - .5f, psz = p.z * m.nz - .5f;
    float fx = floorf(psx), fy = floorf(psy), fz = floorf(psz);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    f.dx = psx - (float)ix; f.dy = psy - (float)iy; f.dz = psz - (float)iz;
    density_gather(m, ix, iy, iz, &f.a, &f.b);
    return f;
}
VD float density_fetch_value(const DensityFetch &f) { return trilerp(f.dx, f.dy, f.dz, f.a.x, f.a.y, f.a.z, f.a.w, f.b.x, f.b.y, f.b.z, f.b.w); }
VD float density_lookup(const DMedium &m, V3 p) { return density_fetch_value(density_fetch_issue(m, p)); }

// NanoVDB grids (media.h:692-705): pIndex = Map::applyInverseMapF(p) (NanoVDB.h:2028-2031 with matMult, :1408-1414),
// Floor + fractional part (util/SampleFromVoxels.h:57-65), then the order-1 sampler (:421-463): lerps along z, then y,
// then x, each as a + w * (b - a).  p is in medium space.
VD DensityFetch nvdb_fetch_issue(const DMedium &m, V3 p) {
    DensityFetch f;
    float x = p.x - m.nvVec[0], y = p.y - m.nvVec[1], z = p.z - m.nvVec[2];
    float px = fmaf(x, m.nvInv[0], fmaf(y, m.nvInv[1], z * m.nvInv[2]));
    float py = fmaf(x, m.nvInv[3], fmaf(y, m.nvInv[4], z * m.nvInv[5]));
    float pz = fmaf(x, m.nvInv[6], fmaf(y, m.nvInv[7], z * m.nvInv[8]));
    float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
    f.dx = px - fx; f.dy = py - fy; f.dz = pz - fz;
    density_gather(m, (int)fx - m.imin[0], (int)fy - m.imin[1], (int)fz - m.imin[2], &f.a, &f.b);
    return f;
}
VD float nvdb_lerp(float a, float b, float w) { return a + w * (b - a); }
VD float nvdb_fetch_value(const DMedium &m, const DensityFetch &f) {
    // a = (v000, v100, v010, v110), b = the same at k + 1, indices in (i, j, k) order
    float vy = nvdb_lerp(f.a.x, f.b.x, f.dz), vy1 = nvdb_lerp(f.a.z, f.b.z, f.dz);
    float vx = nvdb_lerp(vy, vy1, f.dy);
    vy1 = nvdb_lerp(f.a.w, f.b.w, f.dz); vy = nvdb_lerp(f.a.y, f.b.y, f.dz);
    float vx1 = nvdb_lerp(vy, vy1, f.dy);
    return nvdb_lerp(vx, vx1, f.dx) + m.densityOffset;
}
// The density at render-space point p of a uniformgrid or nanovdb medium, split in two so that a caller can put
// independent work between the load and its first use (k_medium_rs issues the fetch for collision k, searches for
// collision k+1, then consumes k).
template <bool NVDB = true> VD DensityFetch medium_density_issue(const DMedium &m, V3 pRender) {
    V3 p = apply_inverse_point(m, pRender);
    return NVDB && m.type == VSPG_MEDIUM_NANOVDB ? nvdb_fetch_issue(m, p) : density_fetch_issue(m, bounds_offset(m, p));
}
template <bool NVDB = true> VD float medium_density_value(const DMedium &m, const DensityFetch &f) {
    return NVDB && m.type == VSPG_MEDIUM_NANOVDB ? nvdb_fetch_value(m, f) : density_fetch_value(f);
}

VD V3 rgb_at(const DMedium &m, const float *g, int x, int y, int z) {
    if (x < 0 || y < 0 || z < 0 || x >= m.nx || y >= m.ny || z >= m.nz) return V3();
    const float *q = g + 3 * ((size_t)(z * m.ny + y) * m.nx + x);
    return V3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
}
VD V3 rgb_lookup(const DMedium &m, const float *g, V3 p) {  // util/containers.h:785-801
    float psx = p.x * m.nx - .5f, psy = p.y * m.ny - .5f, psz = p.z * m.nz - .5f;
    int ix = (int)floorf(psx), iy = (int)floorf(psy), iz = (int)floorf(psz);
    float dx = psx - (float)ix, dy = psy - (float)iy, dz = psz - (float)iz;
    V3 d00 = lerpv(dx, rgb_at(m, g, ix, iy, iz), rgb_at(m, g, ix + 1, iy, iz));
    V3 d10 = lerpv(dx, rgb_at(m, g, ix, iy + 1, iz), rgb_at(m, g, ix + 1, iy + 1, iz));
    V3 d01 = lerpv(dx, rgb_at(m, g, ix, iy, iz + 1), rgb_at(m, g, ix + 1, iy, iz + 1));
    V3 d11 = lerpv(dx, rgb_at(m, g, ix, iy + 1, iz + 1), rgb_at(m, g, ix + 1, iy + 1, iz + 1));
    return lerpv(dz, lerpv(dy, d00, d10), lerpv(dy, d01, d11));
}
// One grid's half of an RGB cell record: 6 x float4 = corners c000, c100, c010, c110, c001, c101, c011, c111, rgb each;
// the interpolation is rgb_lookup's, term for term.
VD V3 rgb_cell_lerp(const float4 *c, bool inside, float dx, float dy, float dz) {
    float4 q0 = make_float4(0, 0, 0, 0), q1 = q0, q2 = q0, q3 = q0, q4 = q0, q5 = q0;
    if (inside) { q0 = __ldg(c); q1 = __ldg(c + 1); q2 = __ldg(c + 2); q3 = __ldg(c + 3); q4 = __ldg(c + 4); q5 = __ldg(c + 5); }
    V3 c000(q0.x, q0.y, q0.z), c100(q0.w, q1.x, q1.y), c010(q1.z, q1.w, q2.x), c110(q2.y, q2.z, q2.w);
    V3 c001(q3.x, q3.y, q3.z), c101(q3.w, q4.x, q4.y), c011(q4.z, q4.w, q5.x), c111(q5.y, q5.z, q5.w);
    V3 d00 = lerpv(dx, c000, c100), d10 = lerpv(dx, c010, c110), d01 = lerpv(dx, c001, c101), d11 = lerpv(dx, c011, c111);
    return lerpv(dz, lerpv(dy, d00, d10), lerpv(dy, d01, d11));
}
// media.h:256-261, 316-348, 417-441  Medium::SamplePoint (emission-free media)
VD MediumProps sample_point(const DMedium &m, V3 p) {
    MediumProps mp; mp.g = m.g;
    if (m.type == VSPG_MEDIUM_HOMOGENEOUS) { mp.sigma_a = ld3(m.sigma_a); mp.sigma_s = ld3(m.sigma_s); return mp; }
    if (m.type == VSPG_MEDIUM_NANOVDB) {
        float d = medium_density_value(m, medium_density_issue(m, p));
        mp.sigma_a = ld3(m.sigma_a) * d; mp.sigma_s = ld3(m.sigma_s) * d;
        return mp;
    }
    p = apply_inverse_point(m, p);
    p = bounds_offset(m, p);
    if (m.type == VSPG_MEDIUM_GRID) {
        float d = density_lookup(m, p);
        mp.sigma_a = ld3(m.sigma_a) * d; mp.sigma_s = ld3(m.sigma_s) * d;
    } else {
        if (m.cells_rgb) {   // both grids from one 192-byte cell record instead of 16 scattered 12-byte reads
            float psx = p.x * m.nx - .5f, psy = p.y * m.ny - .5f, psz = p.z * m.nz - .5f;
            int ix = (int)floorf(psx), iy = (int)floorf(psy), iz = (int)floorf(psz);
            float dx = psx - (float)ix, dy = psy - (float)iy, dz = psz - (float)iz;
            const bool inside = !(ix < -1 || iy < -1 || iz < -1 || ix >= m.nx || iy >= m.ny || iz >= m.nz);
            const float4 *c = m.cells_rgb + 12 * (inside ? ((size_t)(iz + 1) * (m.ny + 1) + (iy + 1)) * (m.nx + 1) + (ix + 1) : 0);
            mp.sigma_a = m.sigma_scale * rgb_cell_lerp(c, inside, dx, dy, dz);
            mp.sigma_s = m.sigma_scale * rgb_cell_lerp(c + 6, inside, dx, dy, dz);
            return mp;
        }
        mp.sigma_a = m.sigma_scale * (m.sa_rgb ? rgb_lookup(m, m.sa_rgb, p) : splat(1.f));
        mp.sigma_s = m.sigma_scale * (m.ss_rgb ? rgb_lookup(m, m.ss_rgb, p) : splat(1.f));
    }
    return mp;
}

// MediumProperties::Le (media.h:259, 329-345): homogeneous media emit Le; grid media s * Le where the trilinear lookup s
// of the "Lescale" grid (util/containers.h:804-831, zero outside the grid) is positive.  p in render space.
VD V3 medium_emission(const DMedium &m, V3 p) {
    if (m.type == VSPG_MEDIUM_HOMOGENEOUS) return ld3(m.Le);
    p = apply_inverse_point(m, p);
    p = bounds_offset(m, p);
    const int nx = m.leN[0], ny = m.leN[1], nz = m.leN[2];
    float psx = p.x * nx - .5f, psy = p.y * ny - .5f, psz = p.z * nz - .5f;
    int ix = (int)floorf(psx), iy = (int)floorf(psy), iz = (int)floorf(psz);
    float dx = psx - (float)ix, dy = psy - (float)iy, dz = psz - (float)iz;
    auto at = [&](int x, int y, int z) {
        if (x < 0 || y < 0 || z < 0 || x >= nx || y >= ny || z >= nz) return 0.f;
        return __ldg(&m.leScale[(size_t)(z * ny + y) * nx + x]);
    };
    float d00 = lerpf(dx, at(ix, iy, iz), at(ix + 1, iy, iz));
    float d10 = lerpf(dx, at(ix, iy + 1, iz), at(ix + 1, iy + 1, iz));
    float d01 = lerpf(dx, at(ix, iy, iz + 1), at(ix + 1, iy, iz + 1));
    float d11 = lerpf(dx, at(ix, iy + 1, iz + 1), at(ix + 1, iy + 1, iz + 1));
    float s = lerpf(dz, lerpf(dy, d00, d10), lerpf(dy, d01, d11));
    return s > 0 ? s * ld3(m.Le) : V3();
}

// ------------------------------------------------------------------ majorant iterators (media.h:76-107, 140-218)
struct MajorantIterator {
    bool homogeneous, called;
    Segment hseg;
    V3 sigma_t;
    float tMin, tMax;
    const float *grid;
    int rx, ry, rz;
    float ncx, ncy, ncz, dtx, dty, dtz;
    int vx, vy, vz, sx, sy, sz, lx, ly, lz;
    unsigned steps;

    VD void init_empty() { homogeneous = false; called = false; tMin = inff(); tMax = -inff(); steps = 0; grid = nullptr; }
    VD void init_dda(const DMedium &m, V3 ro, V3 rd, float tMin_, float tMax_, V3 sigma_t_) {
        tMin = tMin_; tMax = tMax_; grid = m.majorant; sigma_t = sigma_t_;
        rx = m.mres[0]; ry = m.mres[1]; rz = m.mres[2];
        V3 diag(m.bmax[0] - m.bmin[0], m.bmax[1] - m.bmin[1], m.bmax[2] - m.bmin[2]);
        V3 go = bounds_offset(m, ro);
        V3 gd(rd.x / diag.x, rd.y / diag.y, rd.z / diag.z);
        V3 gi = go + gd * tMin;
        axis_setup(gi.x, gd.x, rx, vx, dtx, ncx, sx, lx);
        axis_setup(gi.y, gd.y, ry, vy, dty, ncy, sy, ly);
        axis_setup(gi.z, gd.z, rz, vz, dtz, ncz, sz, lz);
    }
    VD void axis_setup(float gi, float gd, int res, int &voxel, float &deltaT, float &nextCrossingT, int &step, int &limit) {
        voxel = clampi((int)(gi * res), 0, res - 1);
        deltaT = 1 / (fabsf(gd) * res);
        if (gd == -0.f) gd = 0.f;
        if (gd >= 0) {
            float nextVoxelPos = float(voxel + 1) / res;
            nextCrossingT = tMin + (nextVoxelPos - gi) / gd;
            step = 1; limit = res;
        } else {
            float nextVoxelPos = float(voxel) / res;
            nextCrossingT = tMin + (nextVoxelPos - gi) / gd;
            step = -1; limit = -1;
        }
    }
    VD bool next(Segment *seg) {
        steps++;
        if (homogeneous) {
            if (called) return false;
            called = true; *seg = hseg; return true;
        }
        if (tMin >= tMax) return false;
        int bits = ((ncx < ncy) << 2) + ((ncx < ncz) << 1) + ((ncy < ncz));
        int axis = (0xA66 >> (2 * bits)) & 3;  // cmpToAxis = {2,1,2,1,2,2,0,0}  (media.h:186-190)
        float nc = axis == 0 ? ncx : (axis == 1 ? ncy : ncz);
        float tVoxelExit = fminf(tMax, nc);
        float gm = __ldg(&grid[vx + rx * (vy + ry * vz)]);
        seg->tMin = tMin; seg->tMax = tVoxelExit; seg->sigma_maj = sigma_t * gm;
        tMin = tVoxelExit;
        if (nc > tMax) tMin = tMax;
        if (axis == 0) { vx += sx; if (vx == lx) tMin = tMax; ncx += dtx; }
        else if (axis == 1) { vy += sy; if (vy == ly) tMin = tMax; ncy += dty; }
        else { vz += sz; if (vz == lz) tMin = tMax; ncz += dtz; }
        return true;
    }
};

// media.h:264-269, 351-366, 444-455  Medium::SampleRay
VD MajorantIterator sample_ray(const DMedium &m, const Ray &ray, float raytMax) {
    MajorantIterator it; it.init_empty();
    if (m.type == VSPG_MEDIUM_HOMOGENEOUS) {
        it.homogeneous = true;
        it.hseg.tMin = 0.f; it.hseg.tMax = raytMax; it.hseg.sigma_maj = ld3(m.sigma_a) + ld3(m.sigma_s);
        return it;
    }
    V3 o, d;
    apply_inverse_ray(m, ray.o, ray.d, &raytMax, &o, &d);
    float tMin, tMax;
    if (!bounds_intersect(m, o, d, raytMax, &tMin, &tMax)) return it;
    V3 sigma_t = (m.type == VSPG_MEDIUM_GRID || m.type == VSPG_MEDIUM_NANOVDB) ? (ld3(m.sigma_a) + ld3(m.sigma_s)) : splat(1.f);
    it.init_dda(m, o, d, tMin, tMax, sigma_t);
    return it;
}

VD float seg_dt(float a, float b) { float dt = a - b; if (isinf(dt)) dt = kFltMax; return dt; }

}  // namespace vspg
