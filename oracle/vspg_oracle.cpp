// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
// C entry points of the CPU oracle (restatement of the reference's algorithm for the render pass of
// GuidedVolPathVSPGIntegrator).  Loaded through ctypes by tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs ONLY.  The product library (libvspg.so) never
// links or calls this.  Build: oracle/Makefile (g++ -O2 -ffp-contract=off).
//
// Parity status: the OpenPGL query math here is pinned against the reference's own library built
// into oracle/_ref (tests/test_oracle_vs_ref.py, tests/golden/pgl_*.npz).  The pbrt-side pieces are
// pinned by the reference's unit-test vectors where they exist (PCG32, hash, FastExp tolerance, HG);
// SampleDistance / SampleT_maj_* have no reference fixtures ("parity unpinned" by reference tests,
// SURVEY.md 8c) and follow the cited source lines.
#include <cstdio>
#include <cstring>
#include <mutex>

#include "orc_integrator.h"

namespace orc {
HGRep g_hg_table[128];
bool g_hg_table_set = false;
}  // namespace orc

using namespace orc;

extern "C" {

// 128 rows x 18 floats: K, g, weights[4], meanCosines[4], kappas[4], normalizations[4]
int orc_set_hg_table(const float *rows, int n_rows) {
    if (n_rows != 128) return -1;
    for (int i = 0; i < 128; i++) {
        const float *r = rows + 18 * i;
        g_hg_table[i].K = (int)r[0]; g_hg_table[i].g = r[1];
        for (int j = 0; j < 4; j++) {
            g_hg_table[i].weights[j] = r[2 + j]; g_hg_table[i].meanCosines[j] = r[6 + j];
            g_hg_table[i].kappas[j] = r[10 + j]; g_hg_table[i].normalizations[j] = r[14 + j];
        }
    }
    g_hg_table_set = true;
    return 0;
}

// ------------------------------------------------------------- known-answer probes
void orc_pcg32_sequence(uint64_t seq, uint64_t off, int has_off, int n, uint32_t *out) {
    RNG r; if (has_off) r.SetSequence(seq, off); else r.SetSequence(seq);
    for (int i = 0; i < n; i++) out[i] = r.U32();
}
void orc_pcg32_floats(uint64_t seq, uint64_t off, int64_t advance, int n, float *out) {
    RNG r(seq, off); r.Advance(advance);
    for (int i = 0; i < n; i++) out[i] = r.Uniform();
}
uint64_t orc_murmur(const unsigned char *key, size_t len, uint64_t seed) { return MurmurHash64A(key, len, seed); }
uint64_t orc_mixbits(uint64_t v) { return MixBits(v); }
float orc_fastexp(float x) { return FastExp(x); }
float orc_pgl_exp(float x) { return pgl_exp(x); }
float orc_pgl_log(float x) { return pgl_log(x); }
float orc_hg(float cosTheta, float g) { return HenyeyGreenstein(cosTheta, g); }
void orc_sample_hg(const float *wo, float g, const float *u, float *wi, float *pdf) {
    V3 w = SampleHenyeyGreenstein(V3(wo), g, u, pdf);
    wi[0] = w.x; wi[1] = w.y; wi[2] = w.z;
}
void orc_sampler_stream(int px, int py, int seed, int sampleIndex, int n, float *out) {
    Sampler s; s.StartPixelSample(px, py, seed, sampleIndex);
    for (int i = 0; i < n; i++) out[i] = s.Get1D();
}

// ------------------------------------------------------------- media
int orc_build_majorant_grid(const vspg_medium *m, float *out) {
    Medium med; med.m = m; med.bounds.pMin = V3(m->bounds_min); med.bounds.pMax = V3(m->bounds_max);
    med.BuildMajorant(out);
    return 0;
}
float orc_density_lookup(const vspg_medium *m, const float *p) {
    Medium med; med.m = m; med.bounds.pMin = V3(m->bounds_min); med.bounds.pMax = V3(m->bounds_max);   // (no majorant grid needed)
    MediumProps mp = med.SamplePoint(V3(p));
    return mp.sigma_a[0] + mp.sigma_s[0];
}
int orc_majorant_segments(const vspg_medium *m, int n, const float *rays7, int max_segments, vspg_segment *out, int *out_count) {
    Medium med; med.init(m, true);
    for (int i = 0; i < n; i++) {
        const float *r = rays7 + 7 * i;
        Ray ray; ray.o = V3(r); ray.d = V3(r + 3);
        MajorantIterator it = SampleRay(med, ray, r[6]);
        Segment s; int c = 0;
        while (it.Next(&s)) {
            if (c < max_segments) {
                vspg_segment &o = out[(size_t)i * max_segments + c];
                o.t_min = s.tMin; o.t_max = s.tMax;
                for (int k = 0; k < 3; k++) o.sigma_maj[k] = s.sigma_maj[k];
            }
            c++;
        }
        out_count[i] = c;
    }
    return 0;
}

// LightSampler::Sample / PMF at explicit contexts (lightsamplers.h:33-38, 266-359)
int orc_light_sample(const vspg_scene_desc *d, const vspg_integrator_params *ip, int n, const float *pos3, const float *ns3, const float *u,
                     const int *light_of, int *out_light, float *out_pmf, float *out_pmf_of) {
    Scene sc; sc.init(d, nullptr, ip, true);
    Integrator in; in.sc = &sc;
    for (int i = 0; i < n; i++) {
        LightSampleCtx ctx; ctx.p = V3(pos3 + 3 * i); ctx.ns = V3(ns3 + 3 * i);
        float pmf = 0;
        int l = in.SampleLight(u[i], &pmf, ctx.p, ctx.ns);
        out_light[i] = l; out_pmf[i] = l >= 0 ? pmf : 0.f;
        out_pmf_of[i] = in.LightPMF(light_of[i], ctx);
    }
    return 0;
}

int orc_intersect(const vspg_scene_desc *d, const vspg_integrator_params *ip, int n, const float *rays7, int *out_prim, float *out10) {
    Scene sc; sc.init(d, nullptr, ip, true);
    for (int i = 0; i < n; i++) {
        const float *r = rays7 + 7 * i;
        V3 o(r), dir(r + 3);
        Hit h = sc.Intersect(o, dir, r[6]);
        float *q = out10 + 10 * i;
        for (int k = 0; k < 10; k++) q[k] = 0;
        out_prim[i] = -1;
        if (!h.valid) continue;
        out_prim[i] = h.kind == 0 ? h.index : h.kind == 1 ? d->n_quads + h.index : d->n_quads + d->n_triangles + h.index;
        SurfaceIntr s = sc.MakeIntr(h, dir, -1);
        q[0] = h.t;
        for (int k = 0; k < 3; k++) { q[1 + k] = s.p[k]; q[4 + k] = s.pError[k]; q[7 + k] = s.n[k]; }
    }
    return 0;
}

int orc_sample_li(const vspg_scene_desc *d, const vspg_integrator_params *ip, int n, const float *ctx9, const int *light, const float *u2, float *out12) {
    Scene sc; sc.init(d, nullptr, ip, true);
    Integrator in; in.sc = &sc;
    for (int i = 0; i < n; i++) {
        const float *c = ctx9 + 9 * i;
        LightSampleCtx ctx; ctx.p = V3(c); ctx.n = V3(c + 3); ctx.ns = V3(c + 6);
        Integrator::LiSample ls = in.SampleLi(light[i], ctx, u2 + 2 * i);
        float *o = out12 + 12 * i;
        for (int k = 0; k < 12; k++) o[k] = 0;
        o[11] = -1;
        if (!ls.valid) continue;
        o[0] = 1; o[1] = ls.pdf;
        for (int k = 0; k < 3; k++) { o[2 + k] = ls.wi[k]; o[5 + k] = ls.pLight[k]; o[8 + k] = ls.L[k]; }
        if (d->lights[light[i]].type != VSPG_LIGHT_DIFFUSE_AREA) continue;
        V3 ro = OffsetRayOrigin(ctx.p, V3(), ctx.n, ls.wi);
        Hit h = sc.Intersect(ro, ls.wi, Infinity);
        if (!h.valid) continue;
        SurfaceIntr s = sc.MakeIntr(h, ls.wi, -1);
        if (s.areaLight == light[i]) o[11] = in.AreaPDF_Li(light[i], ctx, ls.wi, s);
    }
    return 0;
}

// SampleT_maj on medium `medium` of the scene: same contract as oracle/ref_harness/pbrt_harness.cpp refp_sample_tmaj.
int orc_sample_tmaj(const vspg_scene_desc *d, int medium, const float *ray6, float tMax, float u, uint64_t seq, uint64_t off, int ch,
                    int stop_after, int max_records, float *records15, float *out_final) {
    if (medium < 0 || medium >= d->n_media) return -1;
    Medium med; med.init(&d->media[medium], true);
    Ray ray; ray.o = V3(ray6); ray.d = V3(ray6 + 3); ray.medium = medium;
    RNG rng(seq, off);
    int n = 0;
    S3 T = SampleT_maj(med, ray, tMax, u, rng, ch, nullptr, [&](V3 p, const MediumProps &mp, S3 sigma_maj, S3 T_maj) {
        if (n < max_records) {
            float *r = records15 + 15 * n;
            r[0] = p.x; r[1] = p.y; r[2] = p.z;
            for (int k = 0; k < 3; ++k) { r[3 + k] = mp.sigma_a[k]; r[6 + k] = mp.sigma_s[k]; r[9 + k] = sigma_maj[k]; r[12 + k] = T_maj[k]; }
        }
        ++n;
        return stop_after < 0 || n < stop_after;
    });
    out_final[0] = T[0]; out_final[1] = T[1]; out_final[2] = T[2];
    out_final[3] = rng.Uniform();
    return n;
}

// SampleLd at one vertex: same contract as oracle/ref_harness/pbrt_harness.cpp refp_sample_ld (medium = index, -1 none).
int orc_sample_ld(const vspg_scene_desc *d, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int is_surface, const float *ray6,
                  int medium, int px, int py, int sampleIndex, int dim, int seed, const float *r_p3, int ch, float *out) {
    Scene sc; sc.init(d, g, ip, true); sc.seed = seed;
    Integrator in; in.sc = &sc;
    Sampler sampler; sampler.StartPixelSample(px, py, seed, sampleIndex, dim);
    GuidedBSDF gbsdf; gbsdf.sc = &sc; gbsdf.sampler = &sampler; gbsdf.enableGuiding = ip->surface_guiding;
    gbsdf.enableScatterGuiding = ip->vsp_secondary_guiding; gbsdf.ris = ip->surface_guiding_ris;
    GuidedPhase gphase; gphase.sc = &sc; gphase.sampler = &sampler; gphase.enableGuiding = ip->volume_guiding;
    gphase.enableScatterGuiding = ip->vsp_secondary_guiding; gphase.ris = ip->volume_guiding_ris;
    S3 r_p(r_p3[0], r_p3[1], r_p3[2]);
    S3 Ld(0.f);
    V3 ro(ray6), rd(ray6 + 3);
    if (is_surface) {
        Hit h = sc.Intersect(ro, rd, Infinity);
        if (!h.valid) return 2;
        SurfaceIntr isect = sc.MakeIntr(h, rd, medium);
        if (isect.material < 0) return 2;
        const vspg_material &mat = d->materials[isect.material];
        gbsdf.setBSDF(isect.ns, isect.dpdu, S3(mat.reflectance));
        float v = sampler.Get1D();
        gbsdf.init(ro + h.t * rd, rd, isect.ns, v);
        Ld = in.SampleLd(true, &isect, V3(), V3(), medium, &gbsdf, nullptr, sampler, r_p, ch);
    } else {
        if (medium < 0 || medium >= d->n_media) return 3;
        MediumProps mp = sc.media[medium].SamplePoint(ro);
        float v = sampler.Get1D();
        gphase.init(mp.g, ro, rd, v);
        Ld = in.SampleLd(false, nullptr, ro, -rd, medium, nullptr, &gphase, sampler, r_p, ch);
    }
    out[0] = Ld[0]; out[1] = Ld[1]; out[2] = Ld[2];
    return 0;
}

// ------------------------------------------------------------- guiding
int orc_cache_lookup(const vspg_field *f, int n, const float *pos3, const float *u, int *out_leaf, int *out_region, float *out_u) {
    for (int i = 0; i < n; i++) {
        float s = u[i]; int leaf;
        out_region[i] = field_get_region(f, V3(pos3 + 3 * i), &s, &leaf);
        out_leaf[i] = leaf; out_u[i] = s;
    }
    return 0;
}

// Raw mixture probes (for pinning against the reference library): mode 0 surface (cosine product with n),
// mode 1 volume (HG product with wo,g), mode 2 no product.
int orc_distribution_probe(const vspg_field *f, int mode, int n, const float *pos3, const float *u, const float *aux3, const float *g,
                           const float *s2, const float *wq3, int *region, float *pdf_wq, float *vsp_wq, float *li_pdf_wq,
                           float *sampled3, float *sampled_pdf) {
    if (mode == 1 && !g_hg_table_set) return -2;
    for (int i = 0; i < n; i++) {
        float s = u[i]; int leaf;
        V3 p(pos3 + 3 * i);
        int r = field_get_region(f, p, &s, &leaf);
        region[i] = r;
        if (r < 0 || !f->regions[r].valid) { region[i] = -1; continue; }
        SamplingDistribution d; d.init(&f->regions[r], p);
        if (mode == 0) d.applyCosineProduct(V3(aux3 + 3 * i));
        else if (mode == 1) d.applyHGProduct(V3(aux3 + 3 * i), g[i]);
        V3 wq(wq3 + 3 * i);
        pdf_wq[i] = d.pdf(wq); vsp_wq[i] = d.vsp(wq); li_pdf_wq[i] = d.pdfLi(wq);
        V3 dir; sampled_pdf[i] = d.samplePdf(s2[2 * i], s2[2 * i + 1], &dir);
        sampled3[3 * i] = dir.x; sampled3[3 * i + 1] = dir.y; sampled3[3 * i + 2] = dir.z;
    }
    return 0;
}

int orc_guided_sample(const vspg_scene_desc *scene, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int is_surface, int n,
                      const vspg_gs_input *in, vspg_gs_output *out) {
    Scene sc; sc.init(scene, g, ip, false);
    for (int i = 0; i < n; i++) {
        const vspg_gs_input &I = in[i]; vspg_gs_output &O = out[i];
        std::memset(&O, 0, sizeof(O));
        Sampler sampler; sampler.rng.state = I.rng_state; sampler.rng.inc = I.rng_inc;
        float v = I.v;
        V3 wo(I.wo), p(I.p), wq(I.wq);
        if (is_surface) {
            GuidedBSDF gb; gb.sc = &sc; gb.sampler = &sampler; gb.enableGuiding = ip->surface_guiding;
            gb.enableScatterGuiding = ip->vsp_secondary_guiding; gb.ris = ip->surface_guiding_ris;
            V3 ns(I.n);
            V3 dpdu, t2; CoordinateSystem(ns, &dpdu, &t2);
            gb.setBSDF(ns, dpdu, S3(I.reflectance));
            O.ok = gb.init(p, -wo, ns, v);
            O.region = O.ok ? gb.dist.region : -1;
            GuidedBSDF::Sample s = gb.Sample_f(wo, I.u1, I.u2);
            O.valid = s.valid;
            O.wi[0] = s.wi.x; O.wi[1] = s.wi.y; O.wi[2] = s.wi.z;
            for (int k = 0; k < 3; k++) O.f[k] = s.f[k];
            O.pdf = s.pdf; O.scatter_pdf = s.bsdfPdf; O.mis_pdf = s.misPdf;
            O.pdf_wq = gb.PDF(wo, wq); O.vsp_wq = gb.VolumeScatterProbability(wq);
        } else {
            GuidedPhase gp; gp.sc = &sc; gp.sampler = &sampler; gp.enableGuiding = ip->volume_guiding;
            gp.enableScatterGuiding = ip->vsp_secondary_guiding; gp.ris = ip->volume_guiding_ris;
            O.ok = gp.init(I.g, p, -wo, v);
            O.region = O.ok ? gp.dist.region : -1;
            GuidedPhase::Sample s = gp.Sample_p(wo, I.u2);
            O.valid = s.valid;
            O.wi[0] = s.wi.x; O.wi[1] = s.wi.y; O.wi[2] = s.wi.z;
            O.f[0] = s.p; O.pdf = s.pdf; O.scatter_pdf = s.phasePdf; O.mis_pdf = s.misPdf;
            O.pdf_wq = gp.PDF(wo, wq); O.vsp_wq = gp.VolumeScatterProbability(wq);
        }
        O.rng_state = sampler.rng.state;
    }
    return 0;
}

// ------------------------------------------------------------- SampleDistance probe
int orc_sample_distance(const vspg_scene_desc *scene, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int n,
                        const vspg_sd_input *in, vspg_sd_output *out) {
    Scene sc; sc.init(scene, g, ip, true);
    Integrator integ; integ.sc = &sc;
    for (int i = 0; i < n; i++) {
        const vspg_sd_input &I = in[i]; vspg_sd_output &O = out[i];
        std::memset(&O, 0, sizeof(O));
        Sampler sampler; sampler.rng.state = I.rng_state; sampler.rng.inc = I.rng_inc;
        Integrator::PathState ps;
        ps.ray.o = V3(I.o); ps.ray.d = V3(I.d); ps.ray.medium = I.medium;
        ps.depth = I.depth; ps.beta = S3(I.beta); ps.r_u = S3(I.r_u); ps.r_l = S3(I.r_l);
        GuidedBSDF gb; gb.sc = &sc; gb.sampler = &sampler; gb.enableGuiding = false; gb.enableScatterGuiding = false; gb.ris = false;
        GuidedPhase gp; gp.sc = &sc; gp.sampler = &sampler; gp.enableGuiding = false; gp.enableScatterGuiding = false; gp.ris = false;
        // vspg.cpp:323-325
        uint64_t h0 = HashFloat1(sampler.Get1D());
        uint64_t h1 = HashFloat1(sampler.Get1D());
        RNG rng(h0, h1);
        bool scattered = false, terminated = false;
        Integrator::DistanceEvent ev;
        float vsp = I.vsp;
        integ.SampleDistance(ps, I.pixel_x, I.pixel_y, I.t_max, sampler, rng, I.channel, gb, gp, scattered, terminated, true, &vsp, &ev);
        O.scattered = ev.scattered; O.terminated = terminated; O.depth = ps.depth;
        O.p[0] = ev.p.x; O.p[1] = ev.p.y; O.p[2] = ev.p.z;
        for (int k = 0; k < 3; k++) {
            O.sigma_a[k] = ev.mp.sigma_a[k]; O.sigma_s[k] = ev.mp.sigma_s[k];
            O.beta[k] = ps.beta[k]; O.r_u[k] = ps.r_u[k]; O.r_l[k] = ps.r_l[k];
        }
        O.rng_state = sampler.rng.state; O.density_queries = ev.queries; O.dda_steps = ev.dda;
    }
    return 0;
}

// ------------------------------------------------------------- full render (ImageTileIntegrator::Render restated)
// film_rgbw: H*W*4 doubles, accumulated into.  counters5: paths, queries, dda, lookups, shadow.
int orc_render_ex(const vspg_scene_desc *scene, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int spp_begin, int spp_end,
                  int spp_stride, int x0, int y0, int x1, int y1, int seed, int nthreads, double *film_rgbw, long long *counters5,
                  float *tr_rgb, int *tr_spp);
int orc_render(const vspg_scene_desc *scene, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int spp_begin, int spp_end,
               int spp_stride, int x0, int y0, int x1, int y1, int seed, int nthreads, double *film_rgbw, long long *counters5) {
    return orc_render_ex(scene, g, ip, spp_begin, spp_end, spp_stride, x0, y0, x1, y1, seed, nthreads, film_rgbw, counters5, nullptr, nullptr);
}
// tr_rgb [h*w*3] / tr_spp [h*w] (zero-initialised by the caller, or carrying earlier passes): the storeTrBuffer pre-pass
int orc_render_ex(const vspg_scene_desc *scene, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int spp_begin, int spp_end,
                  int spp_stride, int x0, int y0, int x1, int y1, int seed, int nthreads, double *film_rgbw, long long *counters5,
                  float *tr_rgb, int *tr_spp) {
    if ((ip->volume_guiding || ip->vsp_secondary_guiding) && g && g->volume && !g_hg_table_set) return -2;
    Scene sc; sc.init(scene, g, ip, true); sc.seed = seed;
    if (nthreads < 1) nthreads = 1;
    if (spp_stride < 1) spp_stride = 1;
    std::atomic<int> nextRow(y0);
    std::mutex mu;
    long long tot[5] = {0, 0, 0, 0, 0};
    auto worker = [&]() {
        Integrator integ; integ.sc = &sc;
        if (tr_rgb && tr_spp) { integ.trBuf = tr_rgb; integ.trSpp = tr_spp; }
        while (true) {
            int y = nextRow.fetch_add(1);
            if (y >= y1) break;
            for (int x = x0; x < x1; x++)
                for (int s = spp_begin; s < spp_end; s += spp_stride) integ.RenderPixelSample(x, y, s, film_rgbw);
        }
        std::lock_guard<std::mutex> lk(mu);
        tot[0] += integ.cnt.paths; tot[1] += integ.cnt.queries; tot[2] += integ.cnt.dda; tot[3] += integ.cnt.lookups; tot[4] += integ.cnt.shadow;
    };
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; i++) th.emplace_back(worker);
    for (auto &t : th) t.join();
    if (counters5) for (int i = 0; i < 5; i++) counters5[i] = tot[i];
    return 0;
}

// The reference's wave loop with the ISG buffer trained on the fly (cpu/integrators.cpp:123-239 with nextWaveSize = 1;
// vspg.cpp:230-258 PostProcessWave: Update() after waves 1, 2, 4, 8, ...).  out_vsp [h*w] = the estimate of the last
// update, out_rec8 [h*w*8] the records after the last wave.
int orc_render_isg(const vspg_scene_desc *scene, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int spp, int seed, int nthreads,
                   double *film_rgbw, float *out_vsp, float *out_rec8) {
    if ((ip->volume_guiding || ip->vsp_secondary_guiding) && g && g->volume && !g_hg_table_set) return -2;
    Scene sc; sc.init(scene, g, ip, true); sc.seed = seed;
    if (nthreads < 1) nthreads = 1;
    const int w = scene->camera.width, h = scene->camera.height;
    std::vector<float> rec((size_t)w * h * 8, 0.f), trained((size_t)w * h, 0.f);
    bool ready = false;
    int nextUpdate = 1;
    for (int wave = 1; wave <= spp; wave++) {
        std::atomic<int> nextRow(0);
        auto worker = [&]() {
            Integrator integ; integ.sc = &sc;
            integ.isgRec = rec.data(); integ.isgTrained = ready ? trained.data() : nullptr; integ.isgVariance = ip->vsp_criterion_variance != 0;
            while (true) {
                int y = nextRow.fetch_add(1);
                if (y >= h) break;
                for (int x = 0; x < w; x++) integ.RenderPixelSample(x, y, wave - 1, film_rgbw);
            }
        };
        std::vector<std::thread> th;
        for (int i = 0; i < nthreads; i++) th.emplace_back(worker);
        for (auto &t : th) t.join();
        if (wave == nextUpdate) {
            Integrator::IsgUpdate(rec.data(), (size_t)w * h, ip->vsp_criterion_variance != 0, trained.data());
            ready = true;
            nextUpdate *= 2;
        }
    }
    if (out_vsp) std::memcpy(out_vsp, trained.data(), trained.size() * sizeof(float));
    if (out_rec8) std::memcpy(out_rec8, rec.data(), rec.size() * sizeof(float));
    return 0;
}

// One path with a trace of the radiance, for debugging single pixels
int orc_render_pixel_sample(const vspg_scene_desc *scene, const vspg_guiding_desc *g, const vspg_integrator_params *ip, int px, int py,
                            int sampleIndex, int seed, float *L3) {
    Scene sc; sc.init(scene, g, ip, true); sc.seed = seed;
    Integrator integ; integ.sc = &sc;
    Sampler sampler; sampler.StartPixelSample(px, py, seed, sampleIndex);
    Ray ray; int ch; float fw;
    integ.CameraRay(px, py, sampler, &ray, &ch, &fw);
    S3 L = integ.Li(px, py, ray, sampler, ch);
    for (int k = 0; k < 3; k++) L3[k] = L[k];
    return 0;
}

int orc_camera_ray(const vspg_scene_desc *scene, int px, int py, int sampleIndex, int seed, float *o3, float *d3, int *ch) {
    vspg_integrator_params ip; std::memset(&ip, 0, sizeof(ip));
    Scene sc; sc.init(scene, nullptr, &ip, true); sc.seed = seed;   // (the tabulated gaussian filter is part of the scene set-up)
    Integrator integ; integ.sc = &sc;
    Sampler sampler; sampler.StartPixelSample(px, py, seed, sampleIndex);
    Ray ray; float fw;
    integ.CameraRay(px, py, sampler, &ray, ch, &fw);
    o3[0] = ray.o.x; o3[1] = ray.o.y; o3[2] = ray.o.z; d3[0] = ray.d.x; d3[1] = ray.d.y; d3[2] = ray.d.z;
    return 0;
}

}  // extern "C"
