// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
// C-callable probes into the REFERENCE's own OpenPGL library (built from /root/reference into
// oracle/_ref/libopenpgl_ref.so by oracle/build_ref.sh).  Used by tests to pin the oracle's (and the
// product loader's) guiding-cache arithmetic against the real thing, and by
// tests/golden/make_golden.py to generate the committed golden vectors.
#include <openpgl/cpp/OpenPGL.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <random>
#include <string>
#include <vector>

namespace hx {
namespace pg = openpgl::cpp;
using pg::Device; using pg::Field; using pg::FieldConfig; using pg::SampleStorage; using pg::SampleData;
using pg::SurfaceSamplingDistribution; using pg::VolumeSamplingDistribution;

extern "C" {

struct RefField {
    Device *dev;
    Field *field;
    SurfaceSamplingDistribution *ssd;
    VolumeSamplingDistribution *vsd;
};

void *ref_field_load(const char *path) {
    RefField *f = new RefField();
    f->dev = new Device(PGL_DEVICE_TYPE_CPU_4);
    f->field = new Field(f->dev, std::string(path));
    f->ssd = new SurfaceSamplingDistribution(f->field);
    f->vsd = new VolumeSamplingDistribution(f->field);
    return f;
}
void ref_field_free(void *h) {
    RefField *f = (RefField *)h;
    delete f->ssd; delete f->vsd; delete f->field; delete f->dev; delete f;
}
int ref_field_iteration(void *h) { return (int)((RefField *)h)->field->GetIteration(); }
int ref_field_validate(void *h) { return ((RefField *)h)->field->Validate() ? 1 : 0; }

// Trains a real field from deterministic synthetic samples through the public API (the same calls the
// integrator's training pass makes: SampleStorage::AddSample + Field::Update) and stores it.
// Samples: positions uniform in [bmin,bmax]; directions drawn around `n_lights` pseudo light positions with
// distance = |light - p|; a fraction are volume samples (EInsideVolume).
int ref_field_train(const char *out_path, const float *bmin, const float *bmax, int n_surface, int n_volume, int iterations, int seed,
                    int variance_vsp, int max_samples_per_leaf) {
    Device dev(PGL_DEVICE_TYPE_CPU_4, 1);
    FieldConfig cfg;
    cfg.Init(PGL_SPATIAL_STRUCTURE_KDTREE, PGL_DIRECTIONAL_DISTRIBUTION_PARALLAX_AWARE_VMM, true, (size_t)max_samples_per_leaf);
    cfg.SetVarianceBasedVSP(variance_vsp != 0);
    Field field(&dev, cfg);
    pgl_box3f bounds;
    bounds.lower = {bmin[0], bmin[1], bmin[2]};
    bounds.upper = {bmax[0], bmax[1], bmax[2]};
    field.SetSceneBounds(bounds);
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    const int nl = 3;
    float lights[nl][3];
    for (int i = 0; i < nl; i++)
        for (int k = 0; k < 3; k++) lights[i][k] = bmin[k] + (bmax[k] - bmin[k]) * (U(rng) * 2.f - 0.5f);
    SampleStorage storage;
    for (int it = 0; it < iterations; it++) {
        storage.Clear();
        for (int s = 0; s < n_surface + n_volume; s++) {
            bool vol = s >= n_surface;
            SampleData sd;
            std::memset(&sd, 0, sizeof(sd));
            float p[3];
            for (int k = 0; k < 3; k++) p[k] = bmin[k] + (bmax[k] - bmin[k]) * U(rng);
            if (!vol) p[1] = bmin[1] + 0.02f * (bmax[1] - bmin[1]) * U(rng);  // surface samples near the floor
            int li = std::min(nl - 1, (int)(U(rng) * nl));
            float d[3], len = 0;
            for (int k = 0; k < 3; k++) { d[k] = lights[li][k] - p[k] + 0.3f * (U(rng) - 0.5f); len += d[k] * d[k]; }
            len = std::sqrt(len);
            for (int k = 0; k < 3; k++) d[k] /= len;
            sd.position = {p[0], p[1], p[2]};
            sd.direction = {d[0], d[1], d[2]};
            sd.weight = 0.2f + 2.f * U(rng) * (1 + li);
            sd.pdf = 0.05f + U(rng);
            sd.distance = len;
            sd.flags = 0;
            if (vol) sd.flags |= PGLSampleData::EInsideVolume;
            if (U(rng) < 0.3f) sd.flags |= PGLSampleData::EDirectLight;
            if (U(rng) < 0.45f) sd.flags |= PGLSampleData::ENextEventVolume;
            storage.AddSample(sd);
        }
        field.Update(storage);
    }
    if (!field.Store(std::string(out_path))) return -1;
    return (int)field.GetIteration();
}

// mode 0: surface distribution + cosine product with aux (normal); mode 1: volume distribution + HG product with
// aux (dir) and g; mode 2: surface without product; mode 3: volume without product.
int ref_distribution_probe(void *h, int mode, int n, const float *pos3, const float *u, const float *aux3, const float *g, const float *s2,
                           const float *wq3, int *id, float *u_after, float *pdf_wq, float *vsp_wq, float *li_pdf_wq, float *sampled3,
                           float *sampled_pdf) {
    RefField *f = (RefField *)h;
    for (int i = 0; i < n; i++) {
        pgl_point3f p = {pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]};
        pgl_vec3f aux = {aux3[3 * i], aux3[3 * i + 1], aux3[3 * i + 2]};
        pgl_vec3f wq = {wq3[3 * i], wq3[3 * i + 1], wq3[3 * i + 2]};
        pgl_point2f s = {s2[2 * i], s2[2 * i + 1]};
        float uu = u[i];
        bool ok;
        pgl_vec3f dir = {0, 0, 0};
        if (mode == 0 || mode == 2) {
            ok = f->ssd->Init(f->field, p, uu);
            if (ok) {
                if (mode == 0) f->ssd->ApplyCosineProduct(aux);
                id[i] = (int)f->ssd->GetId();
                pdf_wq[i] = f->ssd->PDF(wq); vsp_wq[i] = f->ssd->VolumeScatterProbability(wq);
                li_pdf_wq[i] = f->ssd->IncomingRadiancePDF(wq);
                sampled_pdf[i] = f->ssd->SamplePDF(s, dir);
            }
        } else {
            ok = f->vsd->Init(f->field, p, uu);
            if (ok) {
                if (mode == 1) f->vsd->ApplySingleLobeHenyeyGreensteinProduct(aux, g[i]);
                id[i] = (int)f->vsd->GetId();
                pdf_wq[i] = f->vsd->PDF(wq); vsp_wq[i] = f->vsd->VolumeScatterProbability(wq);
                li_pdf_wq[i] = f->vsd->IncomingRadiancePDF(wq);
                sampled_pdf[i] = f->vsd->SamplePDF(s, dir);
            }
        }
        u_after[i] = uu;
        if (!ok) { id[i] = -1; continue; }
        sampled3[3 * i] = dir.x; sampled3[3 * i + 1] = dir.y; sampled3[3 * i + 2] = dir.z;
    }
    return 0;
}

// Image-space guiding buffer through the reference's public API: accumulate deterministic pixel samples, Update()
// (OIDN is the identity stub here, so the stored estimate is the un-denoised one), Store() -> multi-layer EXR written
// by the reference's tinyexr path (imagespace/ImageSpaceGuidingBuffer.h:360-630).  (Loading the file back with the
// reference's own loader, :120-345, crashes in this trimmed build, so the in-memory estimate is what gets pinned.)
int ref_isgb_make(const char *path, int w, int h, int spp, int seed, float *vsp_out) {
    using pg::util::ImageSpaceGuidingBuffer;
    ImageSpaceGuidingBuffer::Config cfg(pgl_point2i{w, h});
    cfg.EnableContributionEstimate(true);
    cfg.EnableVolumeScatterProbabilityEstimate(true);
    ImageSpaceGuidingBuffer buf(cfg);
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    for (int s = 0; s < spp; s++)
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                ImageSpaceGuidingBuffer::Sample smp;
                float pv = 0.5f + 0.45f * std::sin(0.37f * x) * std::cos(0.23f * y);
                bool vol = U(rng) < pv;
                float c = (vol ? 0.8f : 0.3f) * (0.5f + U(rng));
                smp.contribution = {c, 0.9f * c, 0.7f * c};
                smp.albedo = {0.5f, 0.5f, 0.5f};
                smp.normal = {0.f, 0.f, -1.f};
                smp.SetSurfaceEvent(!vol);
                buf.AddSample(pgl_point2i{x, y}, smp);
            }
    buf.Update();
    buf.Store(std::string(path));
    // the estimate the integrator reads at depth 0 (vspg.cpp:1098-1112 -> ImageSpaceGuidingBuffer.h:802-809), taken from
    // the very buffer Store() has just written as layer "VSP"
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) vsp_out[y * w + x] = buf.GetVolumeScatterProbabilityEstimate(pgl_point2i{x, y});
    return buf.IsReady() ? 1 : 0;
}
float ref_standard_rr(float x, float y, float z) {
    pgl_vec3f t = {x, y, z};
    return openpgl::cpp::util::StandardThroughputBasedRussianRoulette(t);
}

}  // extern "C"
}  // namespace hx
