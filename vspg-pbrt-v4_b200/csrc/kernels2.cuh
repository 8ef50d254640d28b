// Wavefront stage kernels, part 2: surface hit handling, BSDF shading, shadow rays, film.
#pragma once
#include "kernels.cuh"
#include "kmedium_rs.cuh"

namespace vspg {

// ------------------------------------------------------------------ escape / emission / interface (vspg.cpp:353-446)
#ifndef VSPG_SURFACE_MINB
#define VSPG_SURFACE_MINB 8
#endif
// Streams are touched per branch (profiles/r02n: a whole-record load/store moved 2.9 GB per launch for a kernel that
// mostly re-spawns rays at interfaces): the hit, the ray and the flags always; throughput / MIS weights / radiance only
// where light is picked up.
template <int KINDS = kKindsAll>
__global__ void __launch_bounds__(kBlock, VSPG_SURFACE_MINB) k_surface(const DScene sc, const Wavefront W) {
    const int n = *W.surface.count;
    const vspg_integrator_params &ip = sc.ip;
    for (int i = gtid(); i < n; i += gstride()) {
        const int slot = W.surface.items[i];
        const float4 hr = W.P.hit[slot];
        float4 b = W.P.d_med[slot];
        Hit h; h.t = hr.x; h.u = hr.y; h.v = hr.z; h.prim = __float_as_int(hr.w);
        PathFlags fl = PathFlags::unpack(__float_as_int(b.w));
        bool toDone = false, toNext = false, toShade = false;
        // radiance pick-up: L += beta * Le / avg(r_u) on camera / specular vertices, else the MIS-weighted form
        auto add_emission = [&](V3 Le, auto lightPdfOf) {
            float4 c = W.P.beta_ru[slot], d = W.P.ru_rl[slot], e = W.P.rl_L[slot];
            V3 beta(c.x, c.y, c.z), r_u(c.w, d.x, d.y), r_l(d.z, d.w, e.x), L(e.y, e.z, e.w);
            if (fl.depth == 0 || fl.specular) L = L + beta * Le / avg3(r_u);
            else {
                float4 f = W.P.misc[slot], g = W.P.misc2[slot], pn = W.P.prev_ns[slot];
                float lightPDF = lightPdfOf(V3(f.z, f.w, g.x), V3(pn.x, pn.y, pn.z));
                r_l = r_l * lightPDF;
                float w_l = ip.use_nee ? 1.0f / avg3(r_u + r_l) : 1.0f;
                L = L + beta * w_l * Le;
            }
            W.P.ru_rl[slot] = make_float4(d.x, d.y, r_l.x, r_l.y);
            W.P.rl_L[slot] = make_float4(r_l.z, L.x, L.y, L.z);
        };
        if (h.prim < 0) {
            // infiniteLights = LightType::Infinite and LightType::DeltaDirection lights in this fork (cpu/integrators.h:77-81);
            // DistantLight::Le = scale * Lemit (lights.h:290-293), zeroed beyond depth 0 (vspg.cpp:356-358);
            // PDF_Li(..., allowIncompletePDF = true) = 0 for both
            for (int k = 0; k < sc.n_inf_lights; k++) {
                int li = sc.inf_lights[k];
                V3 Le = ld3(sc.lights[li].L);
                if (sc.lights[li].type == VSPG_LIGHT_DISTANT && fl.depth != 0) Le = V3();
                add_emission(Le, [&](V3 prevP, V3 prevNs) { return light_pmf(sc, li, prevP, prevNs) * 0.f; });
            }
            toDone = true;
        } else {
            const float4 a = W.P.o_dx[slot];
            const V3 ro(a.x, a.y, a.z), rd(a.w, b.x, b.y);
            SurfaceIntr isect = make_intr<KINDS>(sc, h, ro, rd, __float_as_int(b.z));
            if (isect.areaLight >= 0) {
                V3 Le = area_L(sc.lights[isect.areaLight], isect.n, -rd);
                if (nonzero(Le))
                    add_emission(Le, [&](V3 prevP, V3 prevNs) {
                        V3 spawnO = prevP;   // medium vertex: ctx.pi is exact and n = 0, SpawnRay starts at p
                        if (nonzero(prevNs)) { float4 po = W.P.prev_o[slot]; spawnO = V3(po.x, po.y, po.z); }
                        return light_pmf(sc, isect.areaLight, prevP, prevNs) * area_pdf_li<KINDS>(sc, isect.areaLight, prevP, prevNs, rd, spawnO);
                    });
            }
            if (isect.material < 0) {  // interface material: SkipIntersection (interaction.cpp:91-97)
                V3 o = offset_ray_origin(isect.p, isect.pError, isect.n, rd);
                W.P.o_dx[slot] = make_float4(o.x, o.y, o.z, a.w);
                b.z = __int_as_float(isect.get_medium(rd));
                W.P.d_med[slot] = b;
                toNext = true;
            } else if (fl.depth++ >= ip.max_depth) toDone = true;
            else {
                b.w = __int_as_float(fl.pack());
                W.P.d_med[slot] = b;
                toShade = true;
            }
        }
        queue_push(W.done, toDone, slot);
        queue_push(W.next, toNext, slot);
        queue_push(W.shade, toShade, slot);
    }
}

// ------------------------------------------------------------------ surface scatter vertex (vspg.cpp:448-607)
// Same register discipline as k_volscatter: the lanes carry the mixture, the shading frame and the sampler; what the
// scalar epilogue needs of the intersection (point interval, normals, media) waits in a small shared-memory record per
// vertex, and the path record is read and written by the group's leader only, straight from / to the SoA streams.
struct ShadeKeep { float p[3], pe[3], n[3], ns[3]; int mIn, mOut, rayMed, hasIf; };
#ifndef VSPG_SHADE_MINB
#define VSPG_SHADE_MINB 4   // r02: 4 CTAs of 256 threads (64 registers) beat 3 (80): C2 shade 13.5 -> 12.7 ms, C4 32.4 -> 29.7 ms
#endif
template <int G, int KINDS = kKindsAll>
__global__ void __launch_bounds__(G > 0 ? 256 : kBlock, G > 0 ? VSPG_SHADE_MINB : VSPG_SCATTER0_MINB) k_shade(const DScene sc, const Wavefront W, const Queue in) {
    __shared__ ShadeKeep keep_s[G > 0 ? 256 / G : kBlock];
    ShadeKeep &K = keep_s[G > 0 ? threadIdx.x / (G > 0 ? G : 1) : threadIdx.x];
    const int n = *in.count;
    const GroupIter<G> gi;
    const bool leader = gi.leader;
    const vspg_integrator_params &ip = sc.ip;
    unsigned lookups = 0, shadows = 0;
    for (int base = gi.first; base < n; base += gi.step) {
        const int i0 = base + gi.warpOff;
        const bool live = i0 + gi.sub < n;
        if (G == 0 && !live) continue;             // one thread per vertex: no group shuffles, no barriers
        if (G > 0) __syncthreads();                // phase boundary 0; also: the previous trip's epilogue is done with K (always a CTA barrier)
        const int slot = in.items[live ? i0 + gi.sub : base];
        Rng rng; { ulonglong2 q = W.P.rng[slot]; rng.state = q.x; rng.inc = q.y; }
        GuidedBSDF<G> gbsdf;
        gbsdf.enableGuiding = ip.surface_guiding; gbsdf.enableScatterGuiding = ip.vsp_secondary_guiding; gbsdf.ris = ip.surface_guiding_ris;
        V3 rd, woN;
        {
            float4 a = W.P.o_dx[slot], b = W.P.d_med[slot], hr = W.P.hit[slot];
            Hit h; h.t = hr.x; h.u = hr.y; h.v = hr.z; h.prim = __float_as_int(hr.w);
            V3 ro(a.x, a.y, a.z);
            rd = V3(a.w, b.x, b.y);
            SurfaceIntr isect = make_intr<KINDS>(sc, h, ro, rd, __float_as_int(b.z));
            woN = isect.wo;
            if (leader) {
                K.p[0] = isect.p.x; K.p[1] = isect.p.y; K.p[2] = isect.p.z;
                K.pe[0] = isect.pError.x; K.pe[1] = isect.pError.y; K.pe[2] = isect.pError.z;
                K.n[0] = isect.n.x; K.n[1] = isect.n.y; K.n[2] = isect.n.z;
                K.ns[0] = isect.ns.x; K.ns[1] = isect.ns.y; K.ns[2] = isect.ns.z;
                K.mIn = isect.mediumInside; K.mOut = isect.mediumOutside; K.rayMed = isect.rayMedium; K.hasIf = isect.hasInterface;
            }
            gbsdf.set_bsdf(isect.ns, isect.dpdu, ld3(sc.materials[isect.material].reflectance));
            if (G > 0) gbsdf.init_region(sc, ro + h.t * rd, rd, isect.ns, W.P.lk[slot]);   // `v` was drawn by k_lookup
            else { float v = rng.uniform(); gbsdf.init(sc, ro + h.t * rd, rd, isect.ns, v, lookups); }
        }
        scatter_phase_sync<G, 1, VSPG_SHADE_SYNC_SITES>();
        bool emitShadow = false;
        V3 nee_f; float nee_pdf = 0;
        if (ip.use_nee && gbsdf.non_specular()) {
            bool valid; V3 wi;
            if (G > 0) {   // drawn by k_lookup from these three sampler dimensions
                rng.uniform(); rng.uniform(); rng.uniform();
                float4 r = live ? W.P.sh_rl[slot] : make_float4(0, 0, 0, 0);   // see k_volscatter
                valid = __float_as_int(r.w) & 1; wi = V3(r.x, r.y, r.z);
            } else {
                V3 ctxP = offset_ray_origin(ld3(K.p), ld3(K.pe), ld3(K.n), woN);  // reflective-only BSDF (vspg.cpp:1147-1149)
                NeePre ls = nee_presample<KINDS>(sc, rng, ctxP, ld3(K.ns));
                rng.uniform(); rng.uniform(); rng.uniform();
                nee_store(W.P, slot, ls);
                valid = ls.valid; wi = ls.wi;
            }
            if (valid) {
                nee_f = gbsdf.f(woN, wi) * absdotn(wi, gbsdf.frame.z);
                nee_pdf = 1.0f * gbsdf.pdf(woN, wi);
                emitShadow = nonzero(nee_f);
            }
        }
        scatter_phase_sync<G, 2, VSPG_SHADE_SYNC_SITES>();
        float vspNext = 0;
        ScatterSample bs;
        {
            float u = rng.uniform();
            float u0 = rng.uniform(), u1 = rng.uniform();
            bs = gbsdf.sample_f(-rd, u, u0, u1, rng);
            if (bs.valid) vspNext = gbsdf.vsp(bs.wi);
        }
        scatter_phase_sync<G, 3, VSPG_SHADE_SYNC_SITES>();
        const bool w = leader && live;
        bool alive = false;
        if (w) {
            const V3 ip_ = ld3(K.p), ipe = ld3(K.pe), in_ = ld3(K.n), ins = ld3(K.ns);
            float4 c = W.P.beta_ru[slot], d = W.P.ru_rl[slot];
            V3 beta(c.x, c.y, c.z), r_u(c.w, d.x, d.y);
            const int flw = __float_as_int(W.P.d_med[slot].w);
            if (emitShadow) {
                // sh_pl / sh_ple / sh_ru.w and sh_c.xyz already hold the shadow record's values (see k_volscatter)
                float4 lc = W.P.sh_c[slot], pl = W.P.sh_pl[slot], pe = W.P.sh_ple[slot], nr = W.P.sh_rl[slot];
                const float nz = W.P.sh_ru[slot].w;
                V3 so, sd;
                spawn_ray_to(ip_, ipe, in_, V3(pl.x, pl.y, pl.z), V3(pe.x, pe.y, pe.z), V3(pl.w, pe.w, nz), &so, &sd);
                const int smed = K.hasIf ? (dotn(sd, in_) > 0 ? K.mOut : K.mIn) : K.rayMed;   // isect.get_medium(sr.d)
                const int delta = (__float_as_int(nr.w) >> 1) & 1, channel = (flw >> 8) & 3;
                V3 rl = r_u * lc.w, ru = r_u * nee_pdf;
                W.P.sh_o[slot] = make_float4(so.x, so.y, so.z, sd.x);
                W.P.sh_d[slot] = make_float4(sd.y, sd.z, __int_as_float(smed), __int_as_float(delta | (channel << 8)));
                W.P.sh_c[slot] = make_float4(lc.x, lc.y, lc.z, beta.x);
                W.P.sh_rl[slot] = make_float4(rl.x, rl.y, rl.z, beta.y);
                W.P.sh_f[slot] = make_float4(nee_f.x, nee_f.y, nee_f.z, beta.z);
                W.P.sh_ru[slot] = make_float4(ru.x, ru.y, ru.z, nz);
                shadows++;
            }
            alive = bs.valid;
            if (alive) {
                PathFlags fl = PathFlags::unpack(flw);
                float4 f = W.P.misc[slot];
                float rr_correction = f.x * (bs.pdf / bs.scatPdf);
                V3 bsdfWeight = bs.f * absdotn(bs.wi, ins) / bs.pdf;
                beta = beta * bsdfWeight;
                V3 r_l = r_u / bs.misPdf;
                fl.lastVertexVolume = false; fl.specular = false; fl.anyNonSpecular = true;
                V3 no = offset_ray_origin(ip_, ipe, in_, bs.wi);
                const int nmed = K.hasIf ? (dotn(bs.wi, in_) > 0 ? K.mOut : K.mIn) : K.rayMed;   // isect.get_medium(bs.wi)
                if (!nonzero(beta)) alive = false;
                if (alive) {
                    float survivalProb = 1.f;
                    if (fl.depth > ip.min_rr_depth) {
                        const float etaScale = 1.f;  // no transmissive BxDFs on this path
                        V3 rrw = (beta / avg3(r_u)) * rr_correction * etaScale;
                        survivalProb = fl.specular ? 0.95f : fmaxf(0.f, fminf(1.f, max_comp(rrw)));
                    }
                    if (survivalProb < 1 && fl.depth > ip.min_rr_depth) {
                        float q = fmaxf(0.f, 1 - survivalProb);
                        if (rng.uniform() < q) alive = false;
                        else beta = beta / (1 - q);
                    }
                }
                if (alive) {
                    float4 e = W.P.rl_L[slot], g2 = W.P.misc2[slot];
                    W.P.o_dx[slot] = make_float4(no.x, no.y, no.z, bs.wi.x);
                    W.P.d_med[slot] = make_float4(bs.wi.y, bs.wi.z, __int_as_float(nmed), __int_as_float(fl.pack()));
                    W.P.beta_ru[slot] = make_float4(beta.x, beta.y, beta.z, r_u.x);
                    W.P.ru_rl[slot] = make_float4(r_u.y, r_u.z, r_l.x, r_l.y);
                    W.P.rl_L[slot] = make_float4(r_l.z, e.y, e.z, e.w);
                    W.P.rng[slot] = make_ulonglong2(rng.state, rng.inc);
                    W.P.misc[slot] = make_float4(rr_correction, vspNext, ip_.x, ip_.y);
                    W.P.misc2[slot] = make_float4(ip_.z, g2.y, g2.z, g2.w);
                    W.P.prev_ns[slot] = make_float4(ins.x, ins.y, ins.z, 0.f);
                    W.P.prev_o[slot] = make_float4(no.x, no.y, no.z, 0.f);
                }
            }
        }
        queue_push(W.shadow, w && emitShadow, slot);
        queue_push(W.next, w && alive, slot);
        queue_push(W.done, w && !alive, slot);
    }
    if (leader) { counter_add(W.counters, CNT_LOOKUPS, lookups); counter_add(W.counters, CNT_SHADOW, shadows); }
}

// ------------------------------------------------------------------ shadow rays (SampleLd back half, vspg.cpp:1193-1252)
// GRAY: all media gray (see dmath.cuh G1) -- T_ray, r_l, r_u of the ratio tracker are scalars until the final weighting.
#ifndef VSPG_SHADOW_MINB
#define VSPG_SHADOW_MINB 8
#endif
// One shadow ray: the chain of segments through interface surfaces towards the light point with ratio tracking + Russian
// roulette in every medium segment.  Returns false when the ray is blocked / its transmittance was terminated; else *Ld is
// beta * (f_hat * T_ray * L / avg(MIS weights)).
template <bool GRAY, int KINDS = kKindsAll>
VD bool shadow_trace(const DScene &sc, const ShadowRec &sr, TrackCounters &tc, unsigned &nodes, V3 *Ld) {
    typedef typename SpecT<GRAY>::type Sp;
    const int ch = sr.channel;
    Ray lightRay; lightRay.o = sr.o; lightRay.d = sr.d; lightRay.medium = sr.medium;
    Sp T_ray = to_spec<GRAY>(splat(1.f)), r_l = T_ray, r_u = T_ray;
    Rng rng; rng.set_sequence(hash_v3(lightRay.o), hash_v3(lightRay.d));
    while (!(lightRay.d == V3(0, 0, 0))) {
        Hit h = scene_intersect<KINDS>(sc, lightRay.o, lightRay.d, 1 - kShadowEpsilon, nodes);
        if (h.prim >= 0 && sc.quads[h.prim].material >= 0) return false;
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
        if (!nonzero(T_ray)) return false;
        if (h.prim < 0) break;
        SurfaceIntr hs = make_intr<KINDS>(sc, h, lightRay.o, lightRay.d, lightRay.medium);
        spawn_ray_to(hs.p, hs.pError, hs.n, sr.pLight, sr.pLightErr, sr.nLight, &lightRay.o, &lightRay.d);
        lightRay.medium = hs.get_medium(lightRay.d);
    }
    V3 rl3 = to_v3(r_l) * sr.rl;
    V3 ru3 = to_v3(r_u) * sr.ru;
    // f_hat * T_ray * ls->L / (r_l + r_u).Average()  (vspg.cpp:1247-1250), then L += beta * Ld (:479, :833)
    V3 v = sr.delta ? sr.f_hat * to_v3(T_ray) * sr.L / avg3(rl3) : sr.f_hat * to_v3(T_ray) * sr.L / avg3(rl3 + ru3);
    *Ld = sr.beta * v;
    return true;
}

template <bool GRAY, int KINDS = kKindsAll>
__global__ void __launch_bounds__(kBlock, VSPG_SHADOW_MINB) k_shadow(const DScene sc, const Wavefront W) {
    const int n = *W.shadow.count;
    TrackCounters tc; tc.queries = 0; tc.dda = 0;
    unsigned nodes = 0;
    for (int i = gtid(); i < n; i += gstride()) {
        int slot = W.shadow.items[i];
        ShadowRec sr; shadow_load(W.P, slot, sr);
        V3 Ld;
        if (shadow_trace<GRAY, KINDS>(sc, sr, tc, nodes, &Ld)) {
            float4 e = W.P.rl_L[slot];
            e.y += Ld.x; e.z += Ld.y; e.w += Ld.z;
            W.P.rl_L[slot] = e;
        }
    }
    counter_add(W.counters, CNT_QUERIES_SH, tc.queries);
    counter_add(W.counters, CNT_DDA_SH, tc.dda);
    counter_add(W.counters, CNT_BVH, nodes);
}

// ------------------------------------------------------------------ storeISGBuffer (vspg.cpp:613-622)
// openpgl/imagespace/ImageSpaceGuidingBuffer.h:744-797 addSample with VSP_USE_PVOL_EST: the running mean of the path
// contributions (squared for the variance criterion) per first-event type.  The reference adds one sample per pixel
// and wave, in wave order; vspg_render enforces the same (one sample index per render while the store is active), so the
// record is owned by one thread and the float arithmetic is the reference's.
VD void isg_add_sample(const DScene &sc, int pixel, V3 L, bool volumeEvent) {
    float *r = sc.isg_rec + 8 * (size_t)pixel;
    float *spp = r + (volumeEvent ? 1 : 0), *c = r + (volumeEvent ? 5 : 2);
    *spp += 1;
    float alpha = 1.f / *spp;
    V3 q = sc.isg_variance ? L * L : L;
    c[0] = (1.f - alpha) * c[0] + alpha * q.x;
    c[1] = (1.f - alpha) * c[1] + alpha * q.y;
    c[2] = (1.f - alpha) * c[2] + alpha * q.z;
}
// :632-742 update() with the denoiser taken as the identity (un-denoised estimate; the reference filters pVol and the two
// contribution buffers with OIDN first): VSP = max(volume term) / max(surface term + volume term), -1 where both vanish
__global__ void k_isg_update(const float *rec, int npix, int variance, float *vsp) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const float *r = rec + 8 * (size_t)i;
        float pVol = r[1] / (r[0] + r[1]);
        pVol = pVol < 1.f ? pVol : 1.f;       // std::min(1.f, pVol): NaN -> 1
        pVol = 0.f < pVol ? pVol : 0.f;       // std::max(0.f, .)
        float sv[3], vv[3];
        for (int k = 0; k < 3; k++) {
            float s = r[2 + k], v = r[5 + k];
            if (variance) { sv[k] = s > 0.f ? (1.f - pVol) * sqrtf(s) : 0.f; vv[k] = v > 0.f ? pVol * sqrtf(v) : 0.f; }
            else { sv[k] = (1.f - pVol) * s; vv[k] = pVol * v; }
        }
        auto mx = [](float a, float b) { return a < b ? b : a; };   // std::max
        float c0 = sv[0] + vv[0], c1 = sv[1] + vv[1], c2 = sv[2] + vv[2];
        float cs = mx(c0, mx(c1, c2)), vs = mx(vv[0], mx(vv[1], vv[2]));
        vsp[i] = cs > 0.f ? vs / cs : -1.f;
    }
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
        if (sc.isg_rec) isg_add_sample(sc, __float_as_int(m2.y), V3(e.y, e.z, e.w), PathFlags::unpack(__float_as_int(W.P.d_med[slot].w)).firstVolume);
        paths++;
        PathRegs r;
        bool ok = regenerate(sc, W, slot, r);
        queue_push(W.next, ok, slot);
    }
    counter_add(W.counters, CNT_PATHS, paths);
}

// Last node of the wave graph's loop body: counts the trip and keeps the WHILE node going while paths are in flight.
__global__ void k_wave_condition(cudaGraphConditionalHandle handle, const int *activeCount, int *trips) {
    *trips += 1;
    cudaGraphSetConditional(handle, *activeCount > 0 ? 1u : 0u);
}

__global__ void k_reset_counts(int *counts, int n, int keep) {
    int i = threadIdx.x;
    if (i < n && i != keep) counts[i] = 0;
}

}  // namespace vspg
