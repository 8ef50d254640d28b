// Test infrastructure (NOT product code).
// C-callable probe harness over the reference renderer's OWN code: it is compiled against the headers under
// /root/reference/src and linked with oracle/_ref/libpbrt_ref.so (oracle/build_ref_pbrt.sh).  Every function below
// calls the reference's functions and returns what they return; nothing is re-implemented here.  It exists so that
//   * tests/golden/make_golden_pbrt.py can emit golden vectors (camera rays, hits, majorant segments, tracking
//     callbacks, SampleDistance / Li results, light picks) for the oracle restatement to be pinned against, and
//   * bench.py --impl reference can time the reference's own Render() on the host cores.
// `#define private public` opens GuidedVolPathVSPGIntegrator::SampleDistance / SampleLd and TrBuffer to the harness;
// the reference's objects are not modified.
#include <sstream>   // (standard headers first: they must not see the access hack)
#include <string>
#include <vector>
#include <map>
#include <memory>
#include <mutex>
#include <functional>
#include <iostream>
#define private public
#define protected public
#include <pbrt/cpu/integrators.h>
#undef private
#undef protected
#include <pbrt/pbrt.h>
#include <pbrt/cameras.h>
#include <pbrt/cpu/aggregates.h>
#include <pbrt/film.h>
#include <pbrt/filters.h>
#include <pbrt/lights.h>
#include <pbrt/lightsamplers.h>
#include <pbrt/materials.h>
#include <pbrt/media.h>
#include <pbrt/media_sampleTMaj.h>
#include <pbrt/options.h>
#include <pbrt/parser.h>
#include <pbrt/samplers.h>
#include <pbrt/scene.h>
#include <pbrt/shapes.h>
#include <pbrt/textures.h>
#include <pbrt/util/colorspace.h>
#include <pbrt/util/memory.h>
#include <pbrt/util/parallel.h>
#include <pbrt/util/rng.h>
#include <pbrt/util/sampling.h>

#include <chrono>
#include <cstring>
#include <csignal>
#include <execinfo.h>
#include <unistd.h>

using namespace pbrt;

namespace {
struct RefScene {
    std::unique_ptr<BasicScene> scene;
    std::unique_ptr<BasicSceneBuilder> builder;   // owns the transform cache the shapes point into: lives as long as the scene
    std::map<std::string, Medium> media;
    NamedTextures textures;
    std::map<int, pstd::vector<Light> *> shapeIndexToAreaLights;
    std::vector<Light> lights;
    std::map<std::string, pbrt::Material> namedMaterials;
    std::vector<pbrt::Material> materials;
    Primitive accel;
    Camera camera;
    Sampler sampler;
    std::unique_ptr<Integrator> integrator;
    ScratchBuffer scratch{65536};
};
bool g_inited = false;
void segv_backtrace(int sig) {   // debugging aid for the harness: where did the reference crash
    void *frames[64];
    int n = backtrace(frames, 64);
    backtrace_symbols_fd(frames, n, 2);
    _exit(128 + sig);
}
}  // namespace

extern "C" {

// InitPBRT (pbrt.cpp:57-140) with the options a `pbrt --nthreads N --render-coord-sys world --quiet` run sets.
int refp_init(int nthreads, int render_space_world) {
    if (g_inited) return 0;
    PBRTOptions opt;
    opt.nThreads = nthreads;
    opt.quiet = true;
    opt.renderingSpace = render_space_world ? RenderingCoordinateSystem::World : RenderingCoordinateSystem::CameraWorld;
    InitPBRT(opt);
    if (getenv("REFP_BACKTRACE")) std::signal(SIGSEGV, segv_backtrace);
    g_inited = true;
    return 0;
}

// The set-up half of RenderCPU (cpu/render.cpp:24-58): parse, create media / textures / lights / materials /
// aggregate / camera / sampler / integrator with the reference's own factories.
void *refp_open(const char *scene_file) {
    if (!g_inited) return nullptr;
    auto *rs = new RefScene;
    rs->scene = std::make_unique<BasicScene>();
    rs->builder = std::make_unique<BasicSceneBuilder>(rs->scene.get());
    std::vector<std::string> files{std::string(scene_file)};
    ParseFiles(rs->builder.get(), files);
    BasicScene &ps = *rs->scene;
    rs->media = ps.CreateMedia();
    rs->textures = ps.CreateTextures();
    rs->lights = ps.CreateLights(rs->textures, &rs->shapeIndexToAreaLights);
    ps.CreateMaterials(rs->textures, &rs->namedMaterials, &rs->materials);
    rs->accel = ps.CreateAggregate(rs->textures, rs->shapeIndexToAreaLights, rs->media, rs->namedMaterials, rs->materials);
    rs->camera = ps.GetCamera();
    rs->sampler = ps.GetSampler();
    rs->integrator = ps.CreateIntegrator(rs->camera, rs->sampler, rs->accel, rs->lights);
    return rs;
}

void refp_close(void *h) { delete (RefScene *)h; }

// integrator->Render() (cpu/integrators.cpp:75-269) timed around the call; the film is then read back with
// RGBFilm::GetPixelRGB (film.h:269).  out_rgb = [H*W*3] floats or null.
double refp_render(void *h, float *out_rgb) {
    RefScene *rs = (RefScene *)h;
    auto t0 = std::chrono::steady_clock::now();
    rs->integrator->Render();
    double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (out_rgb) {
        Film film = rs->camera.GetFilm();
        Bounds2i pb = film.PixelBounds();
        int w = pb.pMax.x - pb.pMin.x;
        for (int y = pb.pMin.y; y < pb.pMax.y; ++y)
            for (int x = pb.pMin.x; x < pb.pMax.x; ++x) {
                RGB c = film.GetPixelRGB({x, y});
                float *o = out_rgb + 3 * ((size_t)(y - pb.pMin.y) * w + (x - pb.pMin.x));
                o[0] = c.r; o[1] = c.g; o[2] = c.b;
            }
    }
    return s;
}

// The front half of RayIntegrator::EvaluatePixelSample (cpu/integrators.cpp:272-287) for one (pixel, sampleIndex):
// out = {o.xyz, d.xyz, time, weight(rgb), filterWeight, channelIdx, pFilm.xy}
int refp_camera_ray(void *h, int px, int py, int sampleIndex, float *out) {
    RefScene *rs = (RefScene *)h;
    Sampler sampler = rs->sampler.Clone(Allocator());
    sampler.StartPixelSample({px, py}, sampleIndex);
    Float lu = sampler.Get1D();
    SampledWavelengths lambda = rs->camera.GetFilm().SampleWavelengths(lu);
    Filter filter = rs->camera.GetFilm().GetFilter();
    CameraSample cs = GetCameraSample(sampler, {px, py}, filter);
    pstd::optional<CameraRayDifferential> cr = rs->camera.GenerateRayDifferential(cs, lambda);
    if (!cr) return 1;
    out[0] = cr->ray.o.x; out[1] = cr->ray.o.y; out[2] = cr->ray.o.z;
    out[3] = cr->ray.d.x; out[4] = cr->ray.d.y; out[5] = cr->ray.d.z;
    out[6] = cr->ray.time;
    out[7] = cr->weight[0]; out[8] = cr->weight[1]; out[9] = cr->weight[2];
    out[10] = cs.filterWeight; out[11] = (float)lambda.ChannelIdx();
    out[12] = cs.pFilm.x; out[13] = cs.pFilm.y;
    return 0;
}

// RayIntegrator::EvaluatePixelSample without the film add: out = {L.rgb (after the NaN/Inf guard), filterWeight}.
int refp_pixel_sample(void *h, int px, int py, int sampleIndex, float *out) {
    RefScene *rs = (RefScene *)h;
    RayIntegrator *ri = dynamic_cast<RayIntegrator *>(rs->integrator.get());
    if (!ri) return 2;
    Sampler sampler = rs->sampler.Clone(Allocator());
    sampler.StartPixelSample({px, py}, sampleIndex);
    Float lu = sampler.Get1D();
    SampledWavelengths lambda = rs->camera.GetFilm().SampleWavelengths(lu);
    Filter filter = rs->camera.GetFilm().GetFilter();
    CameraSample cs = GetCameraSample(sampler, {px, py}, filter);
    pstd::optional<CameraRayDifferential> cr = rs->camera.GenerateRayDifferential(cs, lambda);
    SampledSpectrum L(0.f);
    if (cr) {
        Float rayDiffScale = std::max<Float>(.125f, 1 / std::sqrt((Float)sampler.SamplesPerPixel()));
        cr->ray.ScaleDifferentials(rayDiffScale);
        L = cr->weight * ri->Li({px, py}, cr->ray, lambda, sampler, rs->scratch, nullptr);
        out[4] = L[0]; out[5] = L[1]; out[6] = L[2]; out[7] = L.y(lambda);
        if (L.HasNaNs() || IsInf(L.y(lambda))) L = SampledSpectrum(0.f);
        rs->scratch.Reset();
    }
    out[0] = L[0]; out[1] = L[1]; out[2] = L[2]; out[3] = cs.filterWeight;
    return 0;
}

// Integrator::Intersect (cpu/integrators.cpp:341-349): rays7 = {o, d, tMax}; out10 = {tHit, p.xyz, n.xyz, ns.xyz}
// hit[i] = 1 / 0.
int refp_intersect(void *h, int n, const float *rays7, int *hit, float *out10) {
    RefScene *rs = (RefScene *)h;
    for (int i = 0; i < n; ++i) {
        const float *r = rays7 + 7 * i;
        Ray ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]));
        pstd::optional<ShapeIntersection> si = rs->integrator->Intersect(ray, r[6]);
        hit[i] = si ? 1 : 0;
        float *o = out10 + 10 * i;
        std::memset(o, 0, 40);
        if (si) {
            o[0] = si->tHit;
            Point3f p = si->intr.p();
            o[1] = p.x; o[2] = p.y; o[3] = p.z;
            o[4] = si->intr.n.x; o[5] = si->intr.n.y; o[6] = si->intr.n.z;
            o[7] = si->intr.shading.n.x; o[8] = si->intr.shading.n.y; o[9] = si->intr.shading.n.z;
        }
    }
    return 0;
}

// Light::SampleLi(ctx, u, lambda, true) (lights.cpp:796-818) then Light::PDF_Li(ctx, wi, true) (:820-824), same layout as
// vspg_debug_sample_li: ctx9 = p, n, ns; out12 = valid, pdf, wi, pLight, L, PDF_Li(ctx, wi).
int refp_sample_li(void *h, int n, const float *ctx9, const int *light, const float *u2, float *out12) {
    RefScene *rs = (RefScene *)h;
    SampledWavelengths lambda = rs->camera.GetFilm().SampleWavelengths(0.1f);
    for (int i = 0; i < n; ++i) {
        const float *c = ctx9 + 9 * i;
        LightSampleContext ctx(Point3fi(Point3f(c[0], c[1], c[2])), Normal3f(c[3], c[4], c[5]), Normal3f(c[6], c[7], c[8]));
        float *o = out12 + 12 * i;
        for (int k = 0; k < 12; ++k) o[k] = 0;
        o[11] = -1;
        if (light[i] < 0 || light[i] >= (int)rs->lights.size()) return 1;
        Light l = rs->lights[light[i]];
        pstd::optional<LightLiSample> ls = l.SampleLi(ctx, Point2f(u2[2 * i], u2[2 * i + 1]), lambda, true);
        if (!ls || !ls->L || ls->pdf == 0) continue;
        o[0] = 1; o[1] = ls->pdf;
        o[2] = ls->wi.x; o[3] = ls->wi.y; o[4] = ls->wi.z;
        Point3f pl = ls->pLight.p();
        o[5] = pl.x; o[6] = pl.y; o[7] = pl.z;
        o[8] = ls->L[0]; o[9] = ls->L[1]; o[10] = ls->L[2];
        o[11] = l.PDF_Li(ctx, ls->wi, true);
    }
    return 0;
}

// SampleT_maj (media_sampleTMaj.h:39-117) on medium `name` for the ray {o, d} up to tMax (in units of |d|), first sample u,
// tracking RNG(seq, off), hero channel `ch`.  The callback records {p.xyz, sigma_a.rgb, sigma_s.rgb, sigma_maj.rgb,
// T_maj.rgb} (15 floats) per tentative collision and returns true until `stop_after` collisions were seen (< 0: never
// stops).  Returns the number of callbacks; out_final = returned T_maj.rgb + the next rng.Uniform<Float>().
int refp_sample_tmaj(void *h, const char *name, const float *ray6, float tMax, float u, uint64_t seq, uint64_t off, int ch,
                     int stop_after, int max_records, float *records15, float *out_final) {
    RefScene *rs = (RefScene *)h;
    auto it = rs->media.find(name);
    if (it == rs->media.end()) return -1;
    Ray ray(Point3f(ray6[0], ray6[1], ray6[2]), Vector3f(ray6[3], ray6[4], ray6[5]), 0.f, it->second);
    SampledWavelengths lambda = rs->camera.GetFilm().SampleWavelengths((ch + 0.5f) / 3.f);
    RNG rng(seq, off);
    int n = 0;
    SampledSpectrum T = SampleT_maj(ray, tMax, u, rng, lambda,
                                    [&](Point3f p, MediumProperties mp, SampledSpectrum sigma_maj, SampledSpectrum T_maj) {
                                        if (n < max_records) {
                                            float *r = records15 + 15 * n;
                                            r[0] = p.x; r[1] = p.y; r[2] = p.z;
                                            for (int k = 0; k < 3; ++k) { r[3 + k] = mp.sigma_a[k]; r[6 + k] = mp.sigma_s[k]; r[9 + k] = sigma_maj[k]; r[12 + k] = T_maj[k]; }
                                        }
                                        ++n;
                                        return stop_after < 0 || n < stop_after;
                                    });
    out_final[0] = T[0]; out_final[1] = T[1]; out_final[2] = T[2];
    out_final[3] = rng.Uniform<Float>();
    return n;
}

// GuidedVolPathVSPGIntegrator::SampleLd (guidedvolpathvspgintegrator.cpp:1136-1252) at one vertex, set up the way Li does
// (:398-480 surface, :805-835 volume).  The sampler is IndependentSampler positioned at (pixel, sampleIndex, dim): the
// guiding-lookup draw v, then SampleLd's own u, uLight come from it, so a restatement given the same position draws
// the same numbers.
//   surface vertex (is_surface = 1): ray6 = incoming ray {o, d}; the vertex is Intersect(ray).  Returns 2 when the ray
//       misses or hits an interface (no BSDF).
//   volume vertex  (is_surface = 0): ray6 = {p, d} with p the scattering point and d the incoming ray direction; the
//       medium is `medium_name`, the phase function its HG(g) as SamplePoint returns it.
// r_p3 = the path's r_u at the vertex, ch = hero channel.  out = Ld.rgb.
int refp_sample_ld(void *h, int is_surface, const float *ray6, const char *medium_name, int px, int py, int sampleIndex, int dim,
                   const float *r_p3, int ch, float *out) {
    RefScene *rs = (RefScene *)h;
    auto *integ = dynamic_cast<GuidedVolPathVSPGIntegrator *>(rs->integrator.get());
    if (!integ) return 1;
    Sampler sampler = rs->sampler.Clone(Allocator());
    sampler.StartPixelSample({px, py}, sampleIndex, dim);
    SampledWavelengths lambda = rs->camera.GetFilm().SampleWavelengths((ch + 0.5f) / 3.f);
    SampledSpectrum r_p; r_p[0] = r_p3[0]; r_p[1] = r_p3[1]; r_p[2] = r_p3[2];
    const auto &gs = integ->guideSettings;
    auto *sdist = integ->guiding_threadSurfaceSamplingDistribution->Get();
    auto *vdist = integ->guiding_threadVolumeSamplingDistribution->Get();
    GuidedBSDF gbsdf(&sampler, integ->guiding_field, sdist, gs.guideSurface, gs.guideSecondaryVSP, gs.surfaceGuidingType);
    GuidedPhaseFunction gphase(&sampler, integ->guiding_field, vdist, gs.guideVolume, gs.guideSecondaryVSP, gs.volumeGuidingType);
    SampledSpectrum Ld(0.f);
    if (is_surface) {
        RayDifferential ray(Point3f(ray6[0], ray6[1], ray6[2]), Vector3f(ray6[3], ray6[4], ray6[5]));
        if (medium_name && *medium_name) { auto it = rs->media.find(medium_name); if (it == rs->media.end()) return 3; ray.medium = it->second; }
        pstd::optional<ShapeIntersection> si = integ->Intersect(ray);
        if (!si) return 2;
        SurfaceInteraction &isect = si->intr;
        BSDF bsdf = isect.GetBSDF(ray, lambda, rs->camera, rs->scratch, sampler);
        if (!bsdf) return 2;
        Float v = sampler.Get1D();
        gbsdf.init(&bsdf, ray, si, v);
        Ld = integ->SampleLd(isect, &gbsdf, nullptr, 1.0f, lambda, sampler, r_p);
        rs->scratch.Reset();
    } else {
        auto it = rs->media.find(medium_name ? medium_name : "");
        if (it == rs->media.end()) return 3;
        Point3f p(ray6[0], ray6[1], ray6[2]);
        Vector3f d(ray6[3], ray6[4], ray6[5]);
        MediumProperties mp = it->second.SamplePoint(p, lambda);
        MediumInteraction intr(p, -d, 0.f, it->second, mp.phase);
        Float v = sampler.Get1D();
        gphase.init(&intr.phase, p, d, v);
        Ld = integ->SampleLd(intr, nullptr, &gphase, 1.0f, lambda, sampler, r_p);
    }
    out[0] = Ld[0]; out[1] = Ld[1]; out[2] = Ld[2];
    return 0;
}

// The geometric front half of SampleLd at a surface vertex, piece by piece, with the reference's own calls: the hit's
// pi (point + error), n, shading.n, wo; ctx.pi = intr.OffsetRayOrigin(intr.wo); light.SampleLi(ctx, uLight); the
// shadow ray intr.SpawnRayTo(ls->pLight) and the two hashes that seed its tracker.
// out = p(3) pError(3) n(3) ns(3) wo(3) ctxp(3) pLight(3) nLight(3) ray.o(3) ray.d(3)  [30], then hash(o), hash(d) in out_hash.
int refp_surface_nee_geometry(void *h, const float *ray6, int light, const float *uLight, float *out, uint64_t *out_hash) {
    RefScene *rs = (RefScene *)h;
    RayDifferential ray(Point3f(ray6[0], ray6[1], ray6[2]), Vector3f(ray6[3], ray6[4], ray6[5]));
    pstd::optional<ShapeIntersection> si = rs->integrator->Intersect(ray);
    if (!si) return 2;
    SurfaceInteraction &isect = si->intr;
    auto put = [&](int k, Float x, Float y, Float z) { out[3 * k] = x; out[3 * k + 1] = y; out[3 * k + 2] = z; };
    Point3f p = isect.p(); Vector3f e = isect.pi.Error();
    put(0, p.x, p.y, p.z); put(1, e.x, e.y, e.z); put(2, isect.n.x, isect.n.y, isect.n.z);
    put(3, isect.shading.n.x, isect.shading.n.y, isect.shading.n.z); put(4, isect.wo.x, isect.wo.y, isect.wo.z);
    LightSampleContext ctx(isect);
    ctx.pi = isect.OffsetRayOrigin(isect.wo);
    put(5, ctx.p().x, ctx.p().y, ctx.p().z);
    SampledWavelengths lambda = rs->camera.GetFilm().SampleWavelengths(0.1f);
    pstd::optional<LightLiSample> ls = rs->lights[light].SampleLi(ctx, Point2f(uLight[0], uLight[1]), lambda, true);
    if (!ls) return 3;
    Point3f pl = ls->pLight.p();
    put(6, pl.x, pl.y, pl.z); put(7, ls->pLight.n.x, ls->pLight.n.y, ls->pLight.n.z);
    Ray lr = isect.SpawnRayTo(ls->pLight);
    put(8, lr.o.x, lr.o.y, lr.o.z); put(9, lr.d.x, lr.d.y, lr.d.z);
    out_hash[0] = Hash(lr.o); out_hash[1] = Hash(lr.d);
    return 0;
}

// The integrator's own light sampler: LightSampler::Sample(ctx, u) (lightsamplers.h:60-360) and PMF(ctx, light_of).
// ctx9 = p, n, ns.  out_light = index into the scene's light array (-1 none).
int refp_light_sample(void *h, int n, const float *ctx9, const float *u, const int *light_of, int *out_light, float *out_pmf, float *out_pmf_of) {
    RefScene *rs = (RefScene *)h;
    auto *integ = dynamic_cast<GuidedVolPathVSPGIntegrator *>(rs->integrator.get());
    if (!integ) return 1;
    for (int i = 0; i < n; ++i) {
        const float *c = ctx9 + 9 * i;
        LightSampleContext ctx(Point3fi(Point3f(c[0], c[1], c[2])), Normal3f(c[3], c[4], c[5]), Normal3f(c[6], c[7], c[8]));
        pstd::optional<SampledLight> sl = integ->lightSampler.Sample(ctx, u[i]);
        out_light[i] = -1; out_pmf[i] = 0;
        if (sl) {
            for (size_t k = 0; k < rs->lights.size(); ++k) if (rs->lights[k] == sl->light) out_light[i] = (int)k;
            out_pmf[i] = sl->p;
        }
        out_pmf_of[i] = (light_of[i] >= 0 && light_of[i] < (int)rs->lights.size()) ? integ->lightSampler.PMF(ctx, rs->lights[light_of[i]]) : 0.f;
    }
    return 0;
}

// Light::Phi(SampledWavelengths::SampleVisible(0.5f)) and that lambda's PDF, per light: out = phi.rgb, pdf.rgb (6 floats each)
int refp_light_phi(void *h, float *out6) {
    RefScene *rs = (RefScene *)h;
    SampledWavelengths lambda = SampledWavelengths::SampleVisible(0.5f);
    for (size_t i = 0; i < rs->lights.size(); ++i) {
        SampledSpectrum phi = rs->lights[i].Phi(lambda), pdf = lambda.PDF();
        for (int c = 0; c < 3; ++c) { out6[6 * i + c] = phi[c]; out6[6 * i + 3 + c] = pdf[c]; }
    }
    return (int)rs->lights.size();
}

// ProjectiveCamera's matrices (cameras.h:198-236): out = cameraFromRaster[16], screenFromCamera[16],
// renderFromCamera[16] (start transform), worldFromRender[16]; returns lensRadius / focalDistance in out[64..65].
int refp_camera_matrices(void *h, float *out) {
    RefScene *rs = (RefScene *)h;
    if (!rs->camera.Is<PerspectiveCamera>()) return 1;
    PerspectiveCamera *pc = rs->camera.Cast<PerspectiveCamera>();
    auto put = [&](const Transform &t, float *o) {
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) o[4 * i + j] = t.GetMatrix()[i][j];
    };
    put(pc->cameraFromRaster, out);
    put(pc->screenFromCamera, out + 16);
    put(pc->cameraTransform.RenderFromCamera().startTransform, out + 32);
    put(pc->cameraTransform.WorldFromRender(), out + 48);
    out[64] = pc->lensRadius; out[65] = pc->focalDistance;
    return 0;
}

}  // extern "C"
