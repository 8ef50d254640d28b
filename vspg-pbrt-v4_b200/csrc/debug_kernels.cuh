// Layout-conversion kernel and the per-kernel parity hooks (include/vspg.h, "per-kernel parity hooks").
#pragma once
#include "kernels2.cuh"

namespace vspg {

// CELL8 layout: cell (cx,cy,cz), cx in [0,nx], holds the corners of interpolation cell ix = cx-1 .. (see dscene.cuh)
__global__ void k_build_cells(const float *density, int nx, int ny, int nz, float bg, float4 *cells) {
    size_t ncell = (size_t)(nx + 1) * (ny + 1) * (nz + 1);
    for (size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncell; c += (size_t)gridDim.x * blockDim.x) {
        int cx = (int)(c % (nx + 1)), cy = (int)((c / (nx + 1)) % (ny + 1)), cz = (int)(c / ((size_t)(nx + 1) * (ny + 1)));
        int ix = cx - 1, iy = cy - 1, iz = cz - 1;
        auto at = [&](int x, int y, int z) -> float {
            if (x < 0 || y < 0 || z < 0 || x >= nx || y >= ny || z >= nz) return bg;   // 0 for pbrt's grids, the tree's background for nanovdb
            return density[(size_t)(z * ny + y) * nx + x];
        };
        cells[2 * c] = make_float4(at(ix, iy, iz), at(ix + 1, iy, iz), at(ix, iy + 1, iz), at(ix + 1, iy + 1, iz));
        cells[2 * c + 1] = make_float4(at(ix, iy, iz + 1), at(ix + 1, iy, iz + 1), at(ix, iy + 1, iz + 1), at(ix + 1, iy + 1, iz + 1));
    }
}

// RGB cell records (dmedia.cuh rgb_cell_lerp): one thread per (cell, grid)
__global__ void k_build_cells_rgb(const float *sa, const float *ss, int nx, int ny, int nz, float4 *cells) {
    size_t ncell = (size_t)(nx + 1) * (ny + 1) * (nz + 1);
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < 2 * ncell; t += (size_t)gridDim.x * blockDim.x) {
        const size_t c = t >> 1;
        const float *g = (t & 1) ? ss : sa;
        int cx = (int)(c % (nx + 1)), cy = (int)((c / (nx + 1)) % (ny + 1)), cz = (int)(c / ((size_t)(nx + 1) * (ny + 1)));
        float v[24];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            int x = cx - 1 + (k & 1), y = cy - 1 + ((k >> 1) & 1), z = cz - 1 + (k >> 2);
            bool in = !(x < 0 || y < 0 || z < 0 || x >= nx || y >= ny || z >= nz);
            const float *q = g + 3 * ((size_t)(z * ny + y) * nx + x);
            v[3 * k] = in ? q[0] : 0.f; v[3 * k + 1] = in ? q[1] : 0.f; v[3 * k + 2] = in ? q[2] : 0.f;
        }
        float4 *o = cells + 12 * c + 6 * (t & 1);
#pragma unroll
        for (int k = 0; k < 6; k++) o[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
    }
}

// Majorant grid of a uniformgrid / rgbgrid medium (media.cpp:256-269, 393-408 with util/containers.h:838-855 MaxValue):
// one CTA per majorant cell takes the maximum over the voxel box MaxValue visits.  A maximum of floats does not depend
// on the order, so this equals the host loop (host_util.h build_majorant_grid) bit for bit.  comps = 1: density grid;
// comps = 3: per-voxel max of the three channels of `a` and of `b` (either may be null => 1), out = scale * (maxA + maxB).
__global__ void __launch_bounds__(256) k_build_majorant(const float *a, const float *b, int comps, float scale, int nx, int ny, int nz, int rx, int ry, int rz, float *out) {
    __shared__ float red[2][256];
    const int cell = blockIdx.x, x = cell % rx, y = (cell / rx) % ry, z = cell / (rx * ry);
    const float p0[3] = {float(x) / rx, float(y) / ry, float(z) / rz}, p1[3] = {float(x + 1) / rx, float(y + 1) / ry, float(z + 1) / rz};
    const int n[3] = {nx, ny, nz};
    int lo[3], hi[3];
    for (int k = 0; k < 3; k++) {
        lo[k] = max((int)floorf(p0[k] * n[k] - .5f), 0);
        hi[k] = min((int)floorf(p1[k] * n[k] - .5f) + 1, n[k] - 1);
    }
    auto at = [&](const float *g, int vx, int vy, int vz) -> float {
        if (vx < 0 || vy < 0 || vz < 0 || vx >= nx || vy >= ny || vz >= nz) return 0.f;   // (MaxValue starts from Lookup(lo), which may lie outside)
        const float *q = g + (size_t)comps * ((size_t)(vz * ny + vy) * nx + vx);
        return comps == 1 ? q[0] : fmaxf(q[0], fmaxf(q[1], q[2]));
    };
    float ma = a ? at(a, lo[0], lo[1], lo[2]) : 1.f, mb = b ? at(b, lo[0], lo[1], lo[2]) : 1.f;
    const int ex = hi[0] - lo[0] + 1, ey = hi[1] - lo[1] + 1, ez = hi[2] - lo[2] + 1;
    const long long total = (ex > 0 && ey > 0 && ez > 0) ? (long long)ex * ey * ez : 0;
    for (long long i = threadIdx.x; i < total; i += blockDim.x) {
        int vx = lo[0] + (int)(i % ex), vy = lo[1] + (int)((i / ex) % ey), vz = lo[2] + (int)(i / ((long long)ex * ey));
        if (a) ma = fmaxf(ma, at(a, vx, vy, vz));
        if (b) mb = fmaxf(mb, at(b, vx, vy, vz));
    }
    red[0][threadIdx.x] = ma; red[1][threadIdx.x] = mb;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) { red[0][threadIdx.x] = fmaxf(red[0][threadIdx.x], red[0][threadIdx.x + s]); red[1][threadIdx.x] = fmaxf(red[1][threadIdx.x], red[1][threadIdx.x + s]); }
        __syncthreads();
    }
    if (threadIdx.x == 0) out[cell] = comps == 1 ? red[0][0] : scale * (red[0][0] + red[1][0]);
}

// DDAMajorantIterator segments for explicit rays (already normalised d, tMax given)
__global__ void k_debug_segments(const DScene sc, int medium, int n, const float *rays7, int maxSeg, vspg_segment *out, int *outCount) {
    for (int i = gtid(); i < n; i += gstride()) {
        const float *r = rays7 + 7 * i;
        Ray ray; ray.o = V3(r[0], r[1], r[2]); ray.d = V3(r[3], r[4], r[5]); ray.medium = medium;
        MajorantIterator it = sample_ray(sc.media[medium], ray, r[6]);
        Segment s; int c = 0;
        while (it.next(&s)) {
            if (c < maxSeg) {
                vspg_segment &o = out[(size_t)i * maxSeg + c];
                o.t_min = s.tMin; o.t_max = s.tMax; o.sigma_maj[0] = s.sigma_maj.x; o.sigma_maj[1] = s.sigma_maj.y; o.sigma_maj[2] = s.sigma_maj.z;
            }
            c++;
        }
        outCount[i] = c;
    }
}

__global__ void k_debug_light_sample(const DScene sc, int n, const float *pos3, const float *ns3, const float *u, const int *lightOf,
                                     int *outLight, float *outPmf, float *outPmfOf) {
    for (int i = gtid(); i < n; i += gstride()) {
        V3 p(pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]), ns(ns3[3 * i], ns3[3 * i + 1], ns3[3 * i + 2]);
        float pmf = 0;
        int l = sample_light(sc, u[i], &pmf, p, ns);
        outLight[i] = l; outPmf[i] = l >= 0 ? pmf : 0.f;
        outPmfOf[i] = light_pmf(sc, lightOf[i], p, ns);
    }
}

__global__ void k_debug_intersect(const DScene sc, int n, const float *rays7, int *outPrim, float *out10) {
    for (int i = gtid(); i < n; i += gstride()) {
        const float *r = rays7 + 7 * i;
        V3 o(r[0], r[1], r[2]), d(r[3], r[4], r[5]);
        unsigned nodes = 0;
        Hit h = scene_intersect(sc, o, d, r[6], nodes);
        float *q = out10 + 10 * i;
        outPrim[i] = h.prim;
        for (int k = 0; k < 10; k++) q[k] = 0;
        if (h.prim < 0) continue;
        SurfaceIntr s = make_intr(sc, h, o, d, -1);
        q[0] = h.t; q[1] = s.p.x; q[2] = s.p.y; q[3] = s.p.z; q[4] = s.pError.x; q[5] = s.pError.y; q[6] = s.pError.z;
        q[7] = s.n.x; q[8] = s.n.y; q[9] = s.n.z;
    }
}

__global__ void k_debug_sample_li(const DScene sc, int n, const float *ctx9, const int *light, const float *u2, float *out12) {
    for (int i = gtid(); i < n; i += gstride()) {
        const float *c = ctx9 + 9 * i;
        V3 p(c[0], c[1], c[2]), nn(c[3], c[4], c[5]), ns(c[6], c[7], c[8]);
        LiSample ls = sample_li(sc, light[i], p, ns, u2[2 * i], u2[2 * i + 1]);
        float *o = out12 + 12 * i;
        for (int k = 0; k < 12; k++) o[k] = 0;
        o[11] = -1;
        if (!ls.valid) continue;
        o[0] = 1; o[1] = ls.pdf; o[2] = ls.wi.x; o[3] = ls.wi.y; o[4] = ls.wi.z; o[5] = ls.pLight.x; o[6] = ls.pLight.y; o[7] = ls.pLight.z;
        o[8] = ls.L.x; o[9] = ls.L.y; o[10] = ls.L.z;
        if (sc.lights[light[i]].type != VSPG_LIGHT_DIFFUSE_AREA) continue;
        V3 ro = offset_ray_origin(p, V3(), nn, ls.wi);   // ctx.SpawnRay(wi): pi is exact
        unsigned nodes = 0;
        Hit h = scene_intersect(sc, ro, ls.wi, inff(), nodes);
        if (h.prim < 0) continue;
        SurfaceIntr s = make_intr(sc, h, ro, ls.wi, -1);
        if (s.areaLight == light[i]) o[11] = area_pdf_li(sc, light[i], p, ns, ls.wi, ro);
    }
}

__global__ void k_debug_density(const DScene sc, int medium, int n, const float *pos3, float *out7) {
    for (int i = gtid(); i < n; i += gstride()) {
        MediumProps mp = sample_point(sc.media[medium], V3(pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]));
        float *o = out7 + 7 * i;
        o[0] = mp.sigma_a.x; o[1] = mp.sigma_a.y; o[2] = mp.sigma_a.z; o[3] = mp.sigma_s.x; o[4] = mp.sigma_s.y; o[5] = mp.sigma_s.z; o[6] = mp.g;
    }
}

__global__ void k_debug_sample_distance(const DScene sc, int n, const vspg_sd_input *in, vspg_sd_output *out) {
    for (int i = gtid(); i < n; i += gstride()) {
        const vspg_sd_input &I = in[i];
        vspg_sd_output &O = out[i];
        PathRegs ps;
        ps.ray.o = V3(I.o[0], I.o[1], I.o[2]); ps.ray.d = V3(I.d[0], I.d[1], I.d[2]); ps.ray.medium = I.medium;
        ps.beta = V3(I.beta[0], I.beta[1], I.beta[2]); ps.r_u = V3(I.r_u[0], I.r_u[1], I.r_u[2]); ps.r_l = V3(I.r_l[0], I.r_l[1], I.r_l[2]);
        ps.L = V3();
        ps.rng.state = I.rng_state; ps.rng.inc = I.rng_inc;
        ps.fl.depth = I.depth; ps.fl.channel = I.channel; ps.fl.specular = false; ps.fl.anyNonSpecular = false; ps.fl.lastVertexVolume = false;
        ps.rr_correction = 1.f; ps.vsp_next = I.vsp; ps.prevP = V3();
        ps.pixel = I.pixel_y * sc.cam.width + I.pixel_x; ps.sampleIndex = 0;
        TrackCounters tc; tc.queries = 0; tc.dda = 0;
        DistanceResult res; res.p = V3(); res.mp.sigma_a = V3(); res.mp.sigma_s = V3(); res.mp.g = 0;
        sample_distance_full(sc, ps, I.t_max, tc, res);
        O.scattered = res.scattered; O.terminated = res.terminated; O.depth = ps.fl.depth;
        V3 p = res.scattered ? res.p : V3();
        O.p[0] = p.x; O.p[1] = p.y; O.p[2] = p.z;
        V3 sa = res.scattered ? res.mp.sigma_a : V3(), ss = res.scattered ? res.mp.sigma_s : V3();
        O.sigma_a[0] = sa.x; O.sigma_a[1] = sa.y; O.sigma_a[2] = sa.z; O.sigma_s[0] = ss.x; O.sigma_s[1] = ss.y; O.sigma_s[2] = ss.z;
        O.beta[0] = ps.beta.x; O.beta[1] = ps.beta.y; O.beta[2] = ps.beta.z;
        O.r_u[0] = ps.r_u.x; O.r_u[1] = ps.r_u.y; O.r_u[2] = ps.r_u.z;
        O.r_l[0] = ps.r_l.x; O.r_l[1] = ps.r_l.y; O.r_l[2] = ps.r_l.z;
        O.rng_state = ps.rng.state; O.density_queries = (int)tc.queries; O.dda_steps = (int)tc.dda;
    }
}

__global__ void k_debug_cache_lookup(const DScene sc, int isSurface, int n, const float *pos3, const float *u, int *outLeaf, int *outRegion, float *outU) {
    for (int i = gtid(); i < n; i += gstride()) {
        float s = u[i]; int leaf;
        outRegion[i] = field_get_region(isSurface ? sc.surf : sc.vol, V3(pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]), &s, &leaf);
        outLeaf[i] = leaf; outU[i] = s;
    }
}

// GuidedBSDF / GuidedPhaseFunction on explicit inputs: one warp per item.
__global__ void __launch_bounds__(kBlock) k_debug_guided_sample(const DScene sc, int isSurface, int n, const vspg_gs_input *in, vspg_gs_output *out) {
    const vspg_integrator_params &ip = sc.ip;
    const bool leader = (threadIdx.x & 31) == 0;
    for (int i = gtid() >> 5; i < n; i += gstride() >> 5) {
        const vspg_gs_input &I = in[i];
        vspg_gs_output O; memset(&O, 0, sizeof O);
        Rng sampler; sampler.state = I.rng_state; sampler.inc = I.rng_inc;
        float v = I.v;
        V3 wo(I.wo[0], I.wo[1], I.wo[2]), p(I.p[0], I.p[1], I.p[2]), wq(I.wq[0], I.wq[1], I.wq[2]);
        unsigned lookups = 0;
        ScatterSample s;
        if (isSurface) {
            GuidedBSDF<32> gb;
            gb.enableGuiding = ip.surface_guiding; gb.enableScatterGuiding = ip.vsp_secondary_guiding; gb.ris = ip.surface_guiding_ris;
            V3 ns(I.n[0], I.n[1], I.n[2]);
            V3 dpdu, t2; coordinate_system(ns, &dpdu, &t2);
            gb.set_bsdf(ns, dpdu, V3(I.reflectance[0], I.reflectance[1], I.reflectance[2]));
            O.ok = gb.init(sc, p, -wo, ns, v, lookups);
            O.region = O.ok ? gb.region : -1;
            s = gb.sample_f(wo, I.u1, I.u2[0], I.u2[1], sampler);
            O.pdf_wq = gb.pdf(wo, wq); O.vsp_wq = gb.vsp(wq);
            if (s.valid) { O.f[0] = s.f.x; O.f[1] = s.f.y; O.f[2] = s.f.z; }
        } else {
            GuidedPhase<32> gp;
            gp.enableGuiding = ip.volume_guiding; gp.enableScatterGuiding = ip.vsp_secondary_guiding; gp.ris = ip.volume_guiding_ris;
            O.ok = gp.init(sc, I.g, p, -wo, v, lookups);
            O.region = O.ok ? gp.region : -1;
            s = gp.sample_p(wo, I.u2[0], I.u2[1], sampler);
            O.pdf_wq = gp.pdf(wo, wq); O.vsp_wq = gp.vsp(wq);
            if (s.valid) O.f[0] = s.f.x;
        }
        O.valid = s.valid;
        if (s.valid) { O.wi[0] = s.wi.x; O.wi[1] = s.wi.y; O.wi[2] = s.wi.z; O.pdf = s.pdf; O.scatter_pdf = s.scatPdf; O.mis_pdf = s.misPdf; }
        O.rng_state = sampler.state;
        if (leader) out[i] = O;
    }
}

// SampleLd at explicit vertices (vspg.cpp:1136-1252): one warp per item.  The same pieces the wavefront runs in
// k_lookup (nee_presample), k_shade / k_volscatter (f, guided pdf, shadow record) and k_shadow (shadow_trace).
template <bool GRAY>
__global__ void __launch_bounds__(kBlock) k_debug_sample_ld(const DScene sc, int n, const vspg_ld_input *in, vspg_ld_output *out) {
    const vspg_integrator_params &ip = sc.ip;
    const bool leader = (threadIdx.x & 31) == 0;
    for (int i = gtid() >> 5; i < n; i += gstride() >> 5) {
        const vspg_ld_input &I = in[i];
        vspg_ld_output O; O.status = 0; O.Ld[0] = O.Ld[1] = O.Ld[2] = 0;
        Rng rng;   // samplers.h:442-476 IndependentSampler::StartPixelSample
        rng.set_sequence(hash_3u32((uint32_t)I.px, (uint32_t)I.py, (uint32_t)I.seed));
        rng.advance((uint64_t)I.sample_index * 65536ull + (uint64_t)I.dim);
        V3 ro(I.ray_o[0], I.ray_o[1], I.ray_o[2]), rd(I.ray_d[0], I.ray_d[1], I.ray_d[2]);
        V3 r_p(I.r_p[0], I.r_p[1], I.r_p[2]);
        unsigned lookups = 0, nodes = 0;
        ShadowRec sr; bool emit = false;
        sr.beta = splat(1.f); sr.channel = I.channel;
        if (I.is_surface) {
            Hit h = scene_intersect(sc, ro, rd, inff(), nodes);
            if (h.prim < 0 || sc.quads[h.prim].material < 0) O.status = 2;
            else {
                SurfaceIntr isect = make_intr(sc, h, ro, rd, I.medium);
                GuidedBSDF<32> gb;
                gb.enableGuiding = ip.surface_guiding; gb.enableScatterGuiding = ip.vsp_secondary_guiding; gb.ris = ip.surface_guiding_ris;
                gb.set_bsdf(isect.ns, isect.dpdu, ld3(sc.materials[isect.material].reflectance));
                float v = rng.uniform();
                gb.init(sc, ro + h.t * rd, rd, isect.ns, v, lookups);
                if (gb.non_specular()) {
                    V3 ctxP = offset_ray_origin(isect.p, isect.pError, isect.n, isect.wo);
                    NeePre ls = nee_presample(sc, rng, ctxP, isect.ns);
                    if (ls.valid) {
                        V3 f_hat = gb.f(isect.wo, ls.wi) * absdotn(ls.wi, isect.ns);
                        float scatterPDF = 1.0f * gb.pdf(isect.wo, ls.wi);
                        if (nonzero(f_hat)) {
                            spawn_ray_to(isect.p, isect.pError, isect.n, ls.pLight, ls.pLightErr, ls.nLight, &sr.o, &sr.d);
                            sr.medium = isect.get_medium(sr.d);
                            sr.f_hat = f_hat; sr.L = ls.L; sr.rl = r_p * ls.p_l; sr.ru = r_p * scatterPDF;
                            sr.pLight = ls.pLight; sr.pLightErr = ls.pLightErr; sr.nLight = ls.nLight; sr.delta = ls.delta;
                            emit = true;
                        }
                    }
                }
            }
        } else if (I.medium < 0 || I.medium >= sc.n_media) O.status = 3;
        else {
            MediumProps mp = sample_point(sc.media[I.medium], ro);
            GuidedPhase<32> gp;
            gp.enableGuiding = ip.volume_guiding; gp.enableScatterGuiding = ip.vsp_secondary_guiding; gp.ris = ip.volume_guiding_ris;
            float v = rng.uniform();
            gp.init(sc, mp.g, ro, rd, v, lookups);
            NeePre ls = nee_presample(sc, rng, ro, V3());
            if (ls.valid) {
                float ph = gp.p(-rd, ls.wi);
                float scatterPDF = 1.0f * gp.pdf(-rd, ls.wi);
                if (ph != 0) {
                    spawn_ray_to(ro, V3(), V3(), ls.pLight, ls.pLightErr, ls.nLight, &sr.o, &sr.d);
                    sr.medium = I.medium;
                    sr.f_hat = splat(ph); sr.L = ls.L; sr.rl = r_p * ls.p_l; sr.ru = r_p * scatterPDF;
                    sr.pLight = ls.pLight; sr.pLightErr = ls.pLightErr; sr.nLight = ls.nLight; sr.delta = ls.delta;
                    emit = true;
                }
            }
        }
        if (emit) {
            TrackCounters tc; tc.queries = 0; tc.dda = 0;
            V3 Ld;
            if (shadow_trace<GRAY>(sc, sr, tc, nodes, &Ld)) { O.Ld[0] = Ld.x; O.Ld[1] = Ld.y; O.Ld[2] = Ld.z; }
        }
        if (leader) out[i] = O;
    }
}

// The path's libm on the device (include/vspg_libm.h, include/vspg_rsqrtss.h): out[i] = fn(a[i] [, b[i]]).
// fn: 0 logf, 1 expf, 2 powf(a, b), 3 sinf, 4 cosf, 5 tanf, 6 asinf, 7 acosf, 8 atanf, 9 atan2f(a, b), 10 embree rsqrt, 11 log (double, a widened)
__global__ void k_debug_libm(int fn, int n, const float *a, const float *b, float *out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float x = a[i], y = b ? b[i] : 0.f, r = 0.f;
        switch (fn) {
            case 0: r = vspg_logf(x); break;
            case 1: r = vspg_expf(x); break;
            case 2: r = vspg_powf(x, y); break;
            case 3: r = vspg_sinf(x); break;
            case 4: r = vspg_cosf(x); break;
            case 5: r = vspg_tanf(x); break;
            case 6: r = vspg_asinf(x); break;
            case 7: r = vspg_acosf(x); break;
            case 8: r = vspg_atanf(x); break;
            case 9: r = vspg_atan2f(x, y); break;
            case 10: r = vspg_embree_rsqrt(x); break;
            default: r = (float)vspg_log((double)x); break;
        }
        out[i] = r;
    }
}

}  // namespace vspg
