// Wavefront stage kernels, part 2: surface hit handling, BSDF shading, shadow rays, film.
#pragma once
#include "kernels.cuh"
#include "kmedium_rs.cuh"

namespace vspg {

// ------------------------------------------------------------------ escape / emission / interface (vspg.cpp:353-446)
__global__ void __launch_bounds__(kBlock, 5) k_surface(const DScene sc, const Wavefront W) {
    const int n = *W.surface.count;
    const vspg_integrator_params &ip = sc.ip;
    for (int i = gtid(); i < n; i += gstride()) {
        int slot = W.surface.items[i];
        PathRegs ps; path_load(W.P, slot, ps);
        float4 hr = W.P.hit[slot];
        Hit h; h.t = hr.x; h.u = hr.y; h.v = hr.z; h.prim = __float_as_int(hr.w);
        bool toDone = false, toNext = false, toShade = false;
        if (h.prim < 0) {
            // infiniteLights = LightType::Infinite and LightType::DeltaDirection lights in this fork (cpu/integrators.h:77-81);
            // DistantLight::Le = scale * Lemit (lights.h:290-293), zeroed beyond depth 0 (vspg.cpp:356-358);
            // PDF_Li(..., allowIncompletePDF = true) = 0 for both
            for (int k = 0; k < sc.n_inf_lights; k++) {
                int li = sc.inf_lights[k];
                V3 Le = ld3(sc.lights[li].L);
                if (sc.lights[li].type == VSPG_LIGHT_DISTANT && ps.fl.depth != 0) Le = V3();
                if (ps.fl.depth == 0 || ps.fl.specular) ps.L = ps.L + ps.beta * Le / avg3(ps.r_u);
                else {
                    float lightPDF = light_pmf(sc, li, ps.prevP, ps.prevNs) * 0.f;
                    ps.r_l = ps.r_l * lightPDF;
                    float w_b = ip.use_nee ? 1.0f / avg3(ps.r_u + ps.r_l) : 1.f;
                    ps.L = ps.L + ps.beta * w_b * Le;
                }
            }
            toDone = true;
        } else {
            SurfaceIntr isect = make_intr(sc, h, ps.ray.o, ps.ray.d, ps.ray.medium);
            if (isect.areaLight >= 0) {
                V3 Le = area_L(sc.lights[isect.areaLight], isect.n, -ps.ray.d);
                if (nonzero(Le)) {
                    if (ps.fl.depth == 0 || ps.fl.specular) ps.L = ps.L + ps.beta * Le / avg3(ps.r_u);
                    else {
                        V3 spawnO = ps.prevP;   // medium vertex: ctx.pi is exact and n = 0, SpawnRay starts at p
                        if (nonzero(ps.prevNs)) { float4 po = W.P.prev_o[slot]; spawnO = V3(po.x, po.y, po.z); }
                        float lightPDF = light_pmf(sc, isect.areaLight, ps.prevP, ps.prevNs) * area_pdf_li(sc, isect.areaLight, ps.prevP, ps.prevNs, ps.ray.d, spawnO);
                        ps.r_l = ps.r_l * lightPDF;
                        float w_l = ip.use_nee ? 1.0f / avg3(ps.r_u + ps.r_l) : 1.0f;
                        ps.L = ps.L + ps.beta * w_l * Le;
                    }
                }
            }
            if (isect.material < 0) {  // interface material: SkipIntersection (interaction.cpp:91-97)
                V3 d = ps.ray.d;
                ps.ray.o = offset_ray_origin(isect.p, isect.pError, isect.n, d);
                ps.ray.medium = isect.get_medium(d);
                toNext = true;
            } else if (ps.fl.depth++ >= ip.max_depth) toDone = true;
            else toShade = true;
        }
        path_store(W.P, slot, ps);
        queue_push(W.done, toDone, slot);
        queue_push(W.next, toNext, slot);
        queue_push(W.shade, toShade, slot);
    }
}

// ------------------------------------------------------------------ surface scatter vertex (vspg.cpp:448-607)
template <int G>
__global__ void __launch_bounds__(G > 0 ? 256 : kBlock, G > 0 ? 3 : 1) k_shade(const DScene sc, const Wavefront W, const Queue in) {
    const int n = *in.count;
    const GroupIter<G> gi;
    const bool leader = gi.leader;
    const vspg_integrator_params &ip = sc.ip;
    unsigned lookups = 0, shadows = 0;
    for (int base = gi.first; base < n; base += gi.step) {
        const int i0 = base + gi.warpOff;
        const bool live = i0 + gi.sub < n;
        if (G == 0 && !live) continue;             // one thread per vertex: no group shuffles, no barriers
        scatter_phase_sync<G>();
        int slot = in.items[live ? i0 + gi.sub : base];
        PathRegs ps; path_load(W.P, slot, ps);
        float4 hr = W.P.hit[slot];
        Hit h; h.t = hr.x; h.u = hr.y; h.v = hr.z; h.prim = __float_as_int(hr.w);
        SurfaceIntr isect = make_intr(sc, h, ps.ray.o, ps.ray.d, ps.ray.medium);
        GuidedBSDF<G> gbsdf;
        gbsdf.enableGuiding = ip.surface_guiding; gbsdf.enableScatterGuiding = ip.vsp_secondary_guiding; gbsdf.ris = ip.surface_guiding_ris;
        gbsdf.set_bsdf(isect.ns, isect.dpdu, ld3(sc.materials[isect.material].reflectance));
        if (G > 0) gbsdf.init_region(sc, ps.ray.o + h.t * ps.ray.d, ps.ray.d, isect.ns, W.P.lk[slot]);   // `v` was drawn by k_lookup
        else { float v = ps.rng.uniform(); gbsdf.init(sc, ps.ray.o + h.t * ps.ray.d, ps.ray.d, isect.ns, v, lookups); }
        scatter_phase_sync<G>();
        bool emitShadow = false;
        if (ip.use_nee && gbsdf.non_specular()) {
            V3 ctxP = offset_ray_origin(isect.p, isect.pError, isect.n, isect.wo);  // reflective-only BSDF (vspg.cpp:1147-1149)
            V3 wo = isect.wo;
            NeePre ls;
            if (G > 0) {   // drawn by k_lookup from these three sampler dimensions
                ps.rng.uniform(); ps.rng.uniform(); ps.rng.uniform();
                if (live) ls = nee_load(W.P, slot); else { ls.valid = false; ls.delta = false; ls.p_l = 0; }   // see k_volscatter
                __syncwarp(grp_mask<G>());   // per-vertex branch: only this group's lanes are known to be here
            } else ls = nee_presample(sc, ps.rng, ctxP, isect.ns), ps.rng.uniform(), ps.rng.uniform(), ps.rng.uniform();
            {
                if (ls.valid) {
                    float p_l = ls.p_l;
                    V3 f_hat = gbsdf.f(wo, ls.wi) * absdotn(ls.wi, isect.ns);
                    float scatterPDF = 1.0f * gbsdf.pdf(wo, ls.wi);
                    if (nonzero(f_hat)) {
                        ShadowRec sr;
                        spawn_ray_to(isect.p, isect.pError, isect.n, ls.pLight, ls.pLightErr, ls.nLight, &sr.o, &sr.d);
                        sr.medium = isect.get_medium(sr.d);
                        sr.f_hat = f_hat; sr.L = ls.L; sr.beta = ps.beta;
                        sr.rl = ps.r_u * p_l; sr.ru = ps.r_u * scatterPDF;
                        sr.pLight = ls.pLight; sr.pLightErr = ls.pLightErr; sr.nLight = ls.nLight;
                        sr.delta = ls.delta; sr.channel = ps.fl.channel;
                        if (leader && live) shadow_store(W.P, slot, sr);
                        emitShadow = true;
                    }
                }
            }
        }
        scatter_phase_sync<G>();
        ps.prevP = isect.p; ps.prevNs = isect.ns;
        V3 wo = -ps.ray.d;
        float u = ps.rng.uniform();
        float u0 = ps.rng.uniform(), u1 = ps.rng.uniform();
        ScatterSample bs = gbsdf.sample_f(wo, u, u0, u1, ps.rng);
        bool alive = bs.valid;
        if (alive) {
            ps.fl.lastVertexVolume = false;
            ps.rr_correction *= bs.pdf / bs.scatPdf;
            V3 bsdfWeight = bs.f * absdotn(bs.wi, isect.ns) / bs.pdf;
            ps.beta = ps.beta * bsdfWeight;
            ps.r_l = ps.r_u / bs.misPdf;
            ps.fl.specular = false; ps.fl.anyNonSpecular = true;
            ps.ray.o = offset_ray_origin(isect.p, isect.pError, isect.n, bs.wi);
            ps.ray.d = bs.wi;
            ps.ray.medium = isect.get_medium(bs.wi);
            if (leader && live) W.P.prev_o[slot] = make_float4(ps.ray.o.x, ps.ray.o.y, ps.ray.o.z, 0.f);
            ps.vsp_next = gbsdf.vsp(bs.wi);
            if (!nonzero(ps.beta)) alive = false;
        }
        if (alive) {
            float survivalProb = 1.f;
            if (ps.fl.depth > ip.min_rr_depth) {
                const float etaScale = 1.f;  // no transmissive BxDFs on this path
                V3 rrw = (ps.beta / avg3(ps.r_u)) * ps.rr_correction * etaScale;
                survivalProb = ps.fl.specular ? 0.95f : fmaxf(0.f, fminf(1.f, max_comp(rrw)));
            }
            if (survivalProb < 1 && ps.fl.depth > ip.min_rr_depth) {
                float q = fmaxf(0.f, 1 - survivalProb);
                if (ps.rng.uniform() < q) alive = false;
                else ps.beta = ps.beta / (1 - q);
            }
        }
        scatter_phase_sync<G>();
        const bool w = leader && live;
        if (w) {
            path_store(W.P, slot, ps);
            if (emitShadow) shadows++;
        }
        queue_push(W.shadow, w && emitShadow, slot);
        queue_push(W.next, w && alive, slot);
        queue_push(W.done, w && !alive, slot);
    }
    if (leader) { counter_add(W.counters, CNT_LOOKUPS, lookups); counter_add(W.counters, CNT_SHADOW, shadows); }
}

// ------------------------------------------------------------------ shadow rays (SampleLd back half, vspg.cpp:1193-1252)
// GRAY: all media gray (see dmath.cuh G1) -- T_ray, r_l, r_u of the ratio tracker are scalars until the final weighting.
template <bool GRAY>
__global__ void __launch_bounds__(kBlock, 8) k_shadow(const DScene sc, const Wavefront W) {
    typedef typename SpecT<GRAY>::type Sp;
    const int n = *W.shadow.count;
    TrackCounters tc; tc.queries = 0; tc.dda = 0;
    unsigned nodes = 0;
    for (int i = gtid(); i < n; i += gstride()) {
        int slot = W.shadow.items[i];
        ShadowRec sr; shadow_load(W.P, slot, sr);
        const int ch = sr.channel;
        Ray lightRay; lightRay.o = sr.o; lightRay.d = sr.d; lightRay.medium = sr.medium;
        Sp T_ray = to_spec<GRAY>(splat(1.f)), r_l = T_ray, r_u = T_ray;
        Rng rng; rng.set_sequence(hash_v3(lightRay.o), hash_v3(lightRay.d));
        bool blocked = false;
        while (!(lightRay.d == V3(0, 0, 0))) {
            Hit h = scene_intersect(sc, lightRay.o, lightRay.d, 1 - kShadowEpsilon, nodes);
            if (h.prim >= 0 && sc.quads[h.prim].material >= 0) { blocked = true; break; }
            if (lightRay.medium >= 0) {
                float tMax = h.prim >= 0 ? h.t : (1 - kShadowEpsilon);
                float uu = rng.uniform();
                Ray tr = lightRay;
                tMax *= length(tr.d);
                tr.d = normalize(tr.d);
                Sp T_maj = track_T_maj_t<GRAY>(sc.media[lightRay.medium], tr, tMax, uu, rng, ch, tc,
                                       [&](V3, const MediumProps &mp, Sp sigma_maj, Sp T_maj) {
                                           Sp sigma_n = clamp_zero(sigma_maj - to_spec<GRAY>(mp.sigma_a) - to_spec<GRAY>(mp.sigma_s));
                                           float pdf = T_maj[ch] * sigma_maj[ch];
                                           T_ray = T_ray * (T_maj * sigma_n / pdf);
                                           r_l = r_l * (T_maj * sigma_maj / pdf);
                                           r_u = r_u * (T_maj * sigma_n / pdf);
                                           Sp Tr = T_ray / avg3(r_l + r_u);
                                           if (max_comp(Tr) < 0.05f) {
                                               float q = 0.75f;
                                               if (rng.uniform() < q) T_ray = Sp();
                                               else T_ray = T_ray / (1 - q);
                                           }
                                           return nonzero(T_ray);
                                       });
                T_ray = T_ray * (T_maj / T_maj[ch]);
                r_l = r_l * (T_maj / T_maj[ch]);
                r_u = r_u * (T_maj / T_maj[ch]);
            }
            if (!nonzero(T_ray)) { blocked = true; break; }
            if (h.prim < 0) break;
            SurfaceIntr hs = make_intr(sc, h, lightRay.o, lightRay.d, lightRay.medium);
            spawn_ray_to(hs.p, hs.pError, hs.n, sr.pLight, sr.pLightErr, sr.nLight, &lightRay.o, &lightRay.d);
            lightRay.medium = hs.get_medium(lightRay.d);
        }
        if (!blocked) {
            V3 rl3 = to_v3(r_l) * sr.rl;
            V3 ru3 = to_v3(r_u) * sr.ru;
            // f_hat * T_ray * ls->L / (r_l + r_u).Average()  (vspg.cpp:1247-1250), then L += beta * Ld (:479, :833)
            V3 Ld = sr.delta ? sr.f_hat * to_v3(T_ray) * sr.L / avg3(rl3) : sr.f_hat * to_v3(T_ray) * sr.L / avg3(rl3 + ru3);
            Ld = sr.beta * Ld;
            float4 e = W.P.rl_L[slot];
            e.y += Ld.x; e.z += Ld.y; e.w += Ld.z;
            W.P.rl_L[slot] = e;
        }
    }
    counter_add(W.counters, CNT_QUERIES_SH, tc.queries);
    counter_add(W.counters, CNT_DDA_SH, tc.dda);
    counter_add(W.counters, CNT_BVH, nodes);
}

// ------------------------------------------------------------------ film + regeneration (integrators.cpp:308-334)
__global__ void __launch_bounds__(kBlock) k_finish(const DScene sc, const Wavefront W) {
    const int n = *W.done.count;
    unsigned paths = 0;
    for (int i = gtid(); i < n; i += gstride()) {
        int slot = W.done.items[i];
        float4 e = W.P.rl_L[slot];
        float4 m2 = W.P.misc2[slot];
        film_add(sc, W.film, __float_as_int(m2.y), V3(e.y, e.z, e.w), m2.w);
        paths++;
        PathRegs r;
        bool ok = regenerate(sc, W, slot, r);
        queue_push(W.next, ok, slot);
    }
    counter_add(W.counters, CNT_PATHS, paths);
}

__global__ void k_reset_counts(int *counts, int n, int keep) {
    int i = threadIdx.x;
    if (i < n && i != keep) counts[i] = 0;
}

}  // namespace vspg
