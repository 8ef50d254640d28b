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

using namespace pbrt;

namespace {
struct RefScene {
    std::unique_ptr<BasicScene> scene;
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
    g_inited = true;
    return 0;
}

// The set-up half of RenderCPU (cpu/render.cpp:24-58): parse, create media / textures / lights / materials /
// aggregate / camera / sampler / integrator with the reference's own factories.
void *refp_open(const char *scene_file) {
    if (!g_inited) return nullptr;
    auto *rs = new RefScene;
    rs->scene = std::make_unique<BasicScene>();
    {
        BasicSceneBuilder builder(rs->scene.get());
        std::vector<std::string> files{std::string(scene_file)};
        ParseFiles(&builder, files);
    }
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
