// Path state in registers + SoA load/store, camera-ray generation and film accumulation.
#pragma once
#include "dshapes.cuh"

namespace vspg {

struct PathRegs {
    Ray ray;
    V3 beta, r_u, r_l, L;
    Rng rng;
    PathFlags fl;
    float rr_correction, vsp_next;
    V3 prevP, prevNs;
    int pixel, sampleIndex;
    float filterWeight;            // FilterSample::weight of the camera sample (1 for the box filter)
};

VD void path_load(const PathSoA &P, int s, PathRegs &r) {
    float4 a = P.o_dx[s], b = P.d_med[s], c = P.beta_ru[s], d = P.ru_rl[s], e = P.rl_L[s], f = P.misc[s], g = P.misc2[s];
    ulonglong2 q = P.rng[s];
    r.ray.o = V3(a.x, a.y, a.z); r.ray.d = V3(a.w, b.x, b.y); r.ray.medium = __float_as_int(b.z);
    r.fl = PathFlags::unpack(__float_as_int(b.w));
    r.beta = V3(c.x, c.y, c.z); r.r_u = V3(c.w, d.x, d.y); r.r_l = V3(d.z, d.w, e.x); r.L = V3(e.y, e.z, e.w);
    r.rng.state = q.x; r.rng.inc = q.y;
    r.rr_correction = f.x; r.vsp_next = f.y; r.prevP = V3(f.z, f.w, g.x);
    r.pixel = __float_as_int(g.y); r.sampleIndex = __float_as_int(g.z); r.filterWeight = g.w;
    float4 h = P.prev_ns[s]; r.prevNs = V3(h.x, h.y, h.z);
}
VD void path_store(const PathSoA &P, int s, const PathRegs &r) {
    P.o_dx[s] = make_float4(r.ray.o.x, r.ray.o.y, r.ray.o.z, r.ray.d.x);
    P.d_med[s] = make_float4(r.ray.d.y, r.ray.d.z, __int_as_float(r.ray.medium), __int_as_float(r.fl.pack()));
    P.beta_ru[s] = make_float4(r.beta.x, r.beta.y, r.beta.z, r.r_u.x);
    P.ru_rl[s] = make_float4(r.r_u.y, r.r_u.z, r.r_l.x, r.r_l.y);
    P.rl_L[s] = make_float4(r.r_l.z, r.L.x, r.L.y, r.L.z);
    P.rng[s] = make_ulonglong2(r.rng.state, r.rng.inc);
    P.misc[s] = make_float4(r.rr_correction, r.vsp_next, r.prevP.x, r.prevP.y);
    P.misc2[s] = make_float4(r.prevP.z, __int_as_float(r.pixel), __int_as_float(r.sampleIndex), r.filterWeight);
    P.prev_ns[s] = make_float4(r.prevNs.x, r.prevNs.y, r.prevNs.z, 0.f);
}

// util/sampling.h:657-675 PiecewiseConstant1D::Sample over cdf[n+1] with func value fo = func[offset] supplied by the caller
VD int find_interval_cdf(const float *cdf, int sz, float u) {   // util/math.h:508-519 with pred = cdf[i] <= u
    int size = sz - 2, first = 1;
    while (size > 0) {
        int half = size >> 1, middle = first + half;
        bool pr = __ldg(&cdf[middle]) <= u;
        first = pr ? middle + 1 : first;
        size = pr ? size - (half + 1) : half;
    }
    return clampi(first - 1, 0, sz - 2);
}
// filters.h:33-38 FilterSampler::Sample for the tabulated gaussian
VD void sample_gaussian_filter(const DScene &sc, float u0, float u1, float *px, float *py, float *weight) {
    const int nx = sc.filt_nx, ny = sc.filt_ny;
    const float rx = sc.cam.filter_radius[0], ry = sc.cam.filter_radius[1];
    int ov = find_interval_cdf(sc.filt_marg_cdf, ny + 1, u1);
    float c0 = __ldg(&sc.filt_marg_cdf[ov]), c1 = __ldg(&sc.filt_marg_cdf[ov + 1]);
    float du = u1 - c0;
    if (c1 - c0 > 0) du /= c1 - c0;
    float rowInt = __ldg(&sc.filt_cond_int[ov]);
    float pdf1 = sc.filt_marg_int > 0 ? rowInt / sc.filt_marg_int : 0.f;
    *py = lerpf((ov + du) / (float)ny, -ry, ry);
    const float *cdf = sc.filt_cond_cdf + (size_t)ov * (nx + 1);
    int ou = find_interval_cdf(cdf, nx + 1, u0);
    c0 = __ldg(&cdf[ou]); c1 = __ldg(&cdf[ou + 1]);
    du = u0 - c0;
    if (c1 - c0 > 0) du /= c1 - c0;
    float fv = __ldg(&sc.filt_f[(size_t)ov * nx + ou]);
    float pdf0 = rowInt > 0 ? fabsf(fv) / rowInt : 0.f;
    *px = lerpf((ou + du) / (float)nx, -rx, rx);
    *weight = fv / (pdf0 * pdf1);
}

// cpu/integrators.cpp:272-334 EvaluatePixelSample (front half) + samplers.h:442-476, 797-814 + cameras.cpp:410-437
VD void generate_camera_path(const DScene &sc, int seed, int px, int py, int sampleIndex, PathRegs &r) {
    const vspg_camera &c = sc.cam;
    r.rng.set_sequence(hash_3u32((uint32_t)px, (uint32_t)py, (uint32_t)seed));  // Hash(Point2i p, int seed)
    r.rng.advance((uint64_t)sampleIndex * 65536ull);
    float lu = r.rng.uniform();
    int ch = min((int)floorf(lu * 3), 2);  // util/spectrum.h:335-339 hero channel
    float uf0 = r.rng.uniform(), uf1 = r.rng.uniform();
    float fx = lerpf(uf0, -c.filter_radius[0], c.filter_radius[0]), fy = lerpf(uf1, -c.filter_radius[1], c.filter_radius[1]);   // BoxFilter::Sample
    r.filterWeight = 1.f;
    if (sc.filt_nx > 0) sample_gaussian_filter(sc, uf0, uf1, &fx, &fy, &r.filterWeight);
    float pFilmX = px + fx + 0.5f, pFilmY = py + fy + 0.5f;
    (void)r.rng.uniform();  // time
    float ul0 = r.rng.uniform(), ul1 = r.rng.uniform();
    const float(*m)[4] = c.camera_from_raster.m;
    float x = pFilmX, y = pFilmY, z = 0;
    float xp = m[0][0] * x + m[0][1] * y + m[0][2] * z + m[0][3];
    float yp = m[1][0] * x + m[1][1] * y + m[1][2] * z + m[1][3];
    float zp = m[2][0] * x + m[2][1] * y + m[2][2] * z + m[2][3];
    float wp = m[3][0] * x + m[3][1] * y + m[3][2] * z + m[3][3];
    V3 pCamera = (wp == 1) ? V3(xp, yp, zp) : V3(xp, yp, zp) / wp;
    V3 o(0, 0, 0), d = normalize(pCamera);
    if (c.lens_radius > 0) {
        float plx, ply; sample_uniform_disk_concentric(ul0, ul1, &plx, &ply);
        plx *= c.lens_radius; ply *= c.lens_radius;
        float ft = c.focal_distance / d.z;
        V3 pFocus = o + d * ft;
        o = V3(plx, ply, 0);
        d = normalize(pFocus - o);
    }
    const float(*rm)[4] = c.render_from_camera.m;
    r.ray.o = V3((rm[0][0] * o.x + rm[0][1] * o.y) + (rm[0][2] * o.z + rm[0][3]), (rm[1][0] * o.x + rm[1][1] * o.y) + (rm[1][2] * o.z + rm[1][3]),
                 (rm[2][0] * o.x + rm[2][1] * o.y) + (rm[2][2] * o.z + rm[2][3]));
    r.ray.d = V3(rm[0][0] * d.x + rm[0][1] * d.y + rm[0][2] * d.z, rm[1][0] * d.x + rm[1][1] * d.y + rm[1][2] * d.z,
                 rm[2][0] * d.x + rm[2][1] * d.y + rm[2][2] * d.z);
    r.ray.medium = c.medium;
    r.beta = r.r_u = r.r_l = splat(1.f); r.L = V3();
    r.fl.depth = 0; r.fl.channel = ch; r.fl.specular = false; r.fl.anyNonSpecular = false; r.fl.lastVertexVolume = false; r.fl.firstVolume = false;
    r.rr_correction = 1.f; r.vsp_next = -1.f; r.prevP = V3();
    r.pixel = py * c.width + px; r.sampleIndex = sampleIndex;
}

// film.h:251-267 RGBFilm::AddSample (fp64 accumulators), after the NaN/Inf guard of integrators.cpp:308-318
VD void film_add(const DScene &sc, double *film, int pixel, V3 L, float fw) {
    if (has_nan(L) || y_is_inf(L)) L = V3();
    const vspg_camera &c = sc.cam;
    float r = c.imaging_ratio * L.x, g = c.imaging_ratio * L.y, b = c.imaging_ratio * L.z;
    float m = fmaxf(r, fmaxf(g, b));
    if (m > c.max_component_value) { float s = c.max_component_value / m; r *= s; g *= s; b *= s; }
    double *pix = film + 4 * (size_t)pixel;
    atomicAdd(pix + 0, (double)(fw * r));
    atomicAdd(pix + 1, (double)(fw * g));
    atomicAdd(pix + 2, (double)(fw * b));
    atomicAdd(pix + 3, (double)fw);
}

}  // namespace vspg
