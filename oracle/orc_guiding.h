// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  (see orc_math.h header)
// CPU restatement of the OpenPGL 0.7.0 (VSP branch) QUERY side used by the render pass:
// KD-tree lookup, stochastic nearest-neighbour region pick, parallax-aware vMF mixture ops,
// surface / volume sampling distributions and the pbrt GuidedBSDF / GuidedPhaseFunction glue.
// Paths cited are relative to /root/reference/src/ext/openpgl/openpgl/ unless they start with pbrt/.
//
// The reference evaluates mixtures as 8 SSE vectors of 4 lanes (CPU_4 device).  To reproduce its
// float results the restatement keeps the same association: per-lane accumulators over vectors
// k = 0..cnt-1, then reduce_add = (l0+l1)+(l2+l3)  (embree vfloat4 vreduce_add: shuffle<1,0,3,2>+v,
// then shuffle<2,3,0,1>+h).
#pragma once
#include "../include/vspg_rsqrtss.h"
#include "../include/vspg.h"
#include "orc_math.h"
#if defined(__SSE__)
#include <xmmintrin.h>
#endif

namespace orc {

static constexpr float M_PI_F = 3.14159265358979323846f;
static constexpr float ONE_OVER_FOUR_PI = 0.07957747154594767f;  // openpgl_common.h
static constexpr float FLT_EPS = 1.1920928955078125e-7f;

// third-party/embreeSrc/common/math/transcendental.h:333-384
inline float pgl_exp(float v) {
    const float ln2Part1 = 0.6931457519f, ln2Part2 = 1.4286067653e-6f, oneOverLn2 = 1.44269502162933349609375f;
    float scaled = v * oneOverLn2;
    float kReal = std::floor(scaled);
    int k = (int)kReal;
    float x = v - kReal * ln2Part1;
    x -= kReal * ln2Part2;
    const float one = 1.f, c2 = 0.4999999105930328369140625f, c3 = 0.166668415069580078125f,
                c4 = 4.16539050638675689697265625e-2f, c5 = 8.378830738365650177001953125e-3f,
                c6 = 1.304379315115511417388916015625e-3f, c7 = 2.7555381529964506626129150390625e-4f;
    float r = x * c7 + c6;
    r = x * r + c5; r = x * r + c4; r = x * r + c3; r = x * r + c2; r = x * r + one; r = x * r + one;
    const int fpbias = 127;
    int biasedN = k + fpbias;
    bool overflow = kReal > fpbias;
    bool underflow = kReal <= -fpbias;
    biasedN = (int)((uint32_t)biasedN << 23);
    float twoToTheN = BitsToFloat((uint32_t)biasedN);
    r *= twoToTheN;
    if (overflow) r = Infinity;
    if (underflow) r = 0.f;
    return r;
}
// transcendental.h:388-472
inline float pgl_log(float v) {
    bool useNan = v < 0.f, useInf = v == 0.f;
    bool exceptional = useNan || useInf;
    float patched = exceptional ? 1.f : v;
    int iv = (int)FloatToBits(patched);
    int biasedExponent = iv >> 23;
    int exponent = biasedExponent + 1 - 127;
    int blended = (iv & (int)0x807FFFFF) | (126 << 23);
    float reduced = BitsToFloat((uint32_t)blended);
    const float ln2 = 0.693147182464599609375f;
    float x1 = 1.f - reduced;
    const float c1 = 0.50000095367431640625f, c2 = 0.33326041698455810546875f, c3 = 0.2519190013408660888671875f,
                c4 = 0.17541764676570892333984375f, c5 = 0.3424419462680816650390625f,
                c6 = -0.599632322788238525390625f, c7 = 1.98442304134368896484375f,
                c8 = -2.4899270534515380859375f, c9 = 1.7491014003753662109375f;
    float r = x1 * c9 + c8;
    r = x1 * r + c7; r = x1 * r + c6; r = x1 * r + c5; r = x1 * r + c4; r = x1 * r + c3; r = x1 * r + c2;
    r = x1 * r + c1; r = x1 * r + 1.f;
    r *= -x1;
    r += (float)exponent * ln2;
    if (exceptional) return useNan ? std::numeric_limits<float>::quiet_NaN() : -Infinity;
    return r;
}

// embree::rsqrt (emath.h:129-149): RSQRTSS estimate + one Newton step.  The estimate is taken from the table of
// include/vspg_rsqrtss.h (what Intel processors return, verified exhaustively), not from this machine's instruction,
// so the oracle gives the same bits on every host.
inline float embree_rsqrt(float x) { return vspg_embree_rsqrt(x); }

inline float reduce4(const float *l) { return (l[0] + l[1]) + (l[2] + l[3]); }

// ---------------------------------------------------------------- spatial lookup
// spatial/kdtree/KDTree.h:402-453 (the treelet walk visits the same nodes as the plain walk)
inline uint32_t kd_lookup(const vspg_field *f, V3 p) {
    uint32_t idx = 0;
    while ((f->nodes[idx].dim_idx >> 30) != 3) {
        uint32_t dim = f->nodes[idx].dim_idx >> 30;
        float pivot = f->nodes[idx].split;
        idx = (f->nodes[idx].dim_idx << 2) >> 2;
        idx += p[(int)dim] >= pivot ? 1 : 0;
    }
    return (f->nodes[idx].dim_idx << 2) >> 2;
}
// spatial/KNN.h:25-32
inline uint32_t knn_draw(float *sample, uint32_t size) {
    size = std::min<uint32_t>(4, size);
    uint32_t selected = (uint32_t)(*sample * size);
    *sample = (*sample - float(selected) / size) * size;
    return std::min(selected, size - 1);
}
// spatial/KNN.h:88-124  RegionNeighbours<4>::prepare + sampleApproximateClosestRegionIdx
inline uint32_t knn_pick(const vspg_neighbours *nh, V3 p, float *sample) {
    uint32_t selected = knn_draw(sample, nh->size);
    float d[2][4];
    for (int i = 0; i < 2; i++) {
        for (int j = 0; j < 4; j++) {
            int s = 4 * i + j;
            float dx = nh->px[s] - p.x, dy = nh->py[s] - p.y, dz = nh->pz[s] - p.z;
            // embree::dot(Vec3<vfloat4>) = madd(a.x,b.x, madd(a.y,b.y, a.z*b.z)) without FMA on SSE4.2
            float dist = dx * dx + (dy * dy + dz * dz);
            uint32_t bits = (FloatToBits(dist) & ~7u) | (uint32_t)s;
            dist = BitsToFloat(bits);
            if (nh->ids[s] == ~0u) dist = Infinity;
            d[i][j] = dist;
        }
        std::sort(d[i], d[i] + 4);  // sort_ascending(vfloat4)
    }
    uint32_t i0 = 0, i1 = 0;
    for (uint32_t i = 0; i < selected; i++) {
        if (d[0][i0] < d[1][i1]) i0++; else i1++;
    }
    if (d[0][i0] < d[1][i1]) return nh->ids[0 + (FloatToBits(d[0][i0]) & 3)];
    else return nh->ids[4 + (FloatToBits(d[1][i1]) & 3)];
}
// field/Field.h:114-150 Field::getRegion.  Returns region index or -1; *leaf gets the KD leaf id.
inline int field_get_region(const vspg_field *f, V3 p, float *sample1D, int *leaf) {
    *leaf = -1;
    if (!f || !f->usable) return -1;
    // embree::inside(BBox, p): lower <= p <= upper on all axes
    if (!(p.x >= f->bounds_min[0] && p.y >= f->bounds_min[1] && p.z >= f->bounds_min[2] && p.x <= f->bounds_max[0] &&
          p.y <= f->bounds_max[1] && p.z <= f->bounds_max[2]))
        return -1;
    uint32_t dataIdx = kd_lookup(f, p);
    *leaf = (int)dataIdx;
    if (f->stochastic_nn && *sample1D >= 0.f && f->neighbours) return (int)knn_pick(&f->neighbours[dataIdx], p, sample1D);
    return (int)dataIdx;
}

// ---------------------------------------------------------------- vMF mixture
struct VMM {
    float w[32], k[32], mx[32], my[32], mz[32], nrm[32], e2k[32], mc[32], dist[32], vspw[32];
    int n = 0;
    V3 pivot;
    void load(const vspg_mixture *m) {
        std::memcpy(w, m->weight, 128); std::memcpy(k, m->kappa, 128);
        std::memcpy(mx, m->mu_x, 128); std::memcpy(my, m->mu_y, 128); std::memcpy(mz, m->mu_z, 128);
        std::memcpy(nrm, m->norm, 128); std::memcpy(e2k, m->e_minus_2k, 128); std::memcpy(mc, m->mean_cos, 128);
        std::memcpy(dist, m->distance, 128); std::memcpy(vspw, m->vsp_weight, 128);
        n = m->num_components; pivot = V3(m->pivot);
    }
    int cnt() const { return (n + 3) / 4; }

    // directional/vmm/ParallaxAwareVonMisesFisherMixture.h:1132-1155
    void parallaxShift(V3 s) {
        for (int i = 0; i < cnt() * 4; i++) {
            float px = mx[i] * dist[i] + s.x, py = my[i] * dist[i] + s.y, pz = mz[i] * dist[i] + s.z;
            // embree::length = sqrt(dot): dot(Vec3<vfloat4>) = x*x + (y*y + z*z)
            float len = std::sqrt(px * px + (py * py + pz * pz));
            px /= len; py /= len; pz /= len;
            bool ok = (dist[i] > 0.0f) && std::isfinite(dist[i]);
            if (ok) { mx[i] = px; my[i] = py; mz[i] = pz; }
            dist[i] = dist[i] > 0.0f ? len : 0.f;
        }
        pivot = pivot - s;
    }
    // :861-880
    float pdf(V3 d) const {
        float acc[4] = {0, 0, 0, 0};
        for (int v = 0; v < cnt(); v++)
            for (int l = 0; l < 4; l++) {
                int i = 4 * v + l;
                float cosTheta = d.x * mx[i] + (d.y * my[i] + d.z * mz[i]);
                float cm1 = std::min(cosTheta - 1.f, 0.f);
                float eval = nrm[i] * pgl_exp(k[i] * cm1);
                acc[l] += w[i] * eval;
            }
        return reduce4(acc);
    }
    // :1237-1258
    float vsp(V3 d) const {
        float av[4] = {0, 0, 0, 0}, ap[4] = {0, 0, 0, 0};
        for (int v = 0; v < cnt(); v++)
            for (int l = 0; l < 4; l++) {
                int i = 4 * v + l;
                float cosTheta = d.x * mx[i] + (d.y * my[i] + d.z * mz[i]);
                float cm1 = std::min(cosTheta - 1.f, 0.f);
                float eval = w[i] * nrm[i] * pgl_exp(k[i] * cm1);
                ap[l] += eval;
                av[l] += vspw[i] * eval;
            }
        return reduce4(av) / reduce4(ap);
    }
    // :776-858  product(weight, meanDirection, kappa, normalization)
    float product(float _w, V3 md, float _k, float _n) {
        float acc[4] = {0, 0, 0, 0};
        const float twoPi = 2.0f * M_PI_F;
        int c = cnt();
        for (int v = 0; v < c; v++)
            for (int l = 0; l < 4; l++) {
                int i = 4 * v + l;
                float nx = k[i] * mx[i] + _k * md.x, ny = k[i] * my[i] + _k * md.y, nz = k[i] * mz[i] + _k * md.z;
                float nk = std::sqrt(nx * nx + (ny * ny + nz * nz));
                bool chk = nk > 1e-3f;
                nk = chk ? nk : 0.f;
                nx = chk ? nx / nk : mx[i]; ny = chk ? ny / nk : my[i]; nz = chk ? nz / nk : mz[i];
                float ne2k = pgl_exp(-2.0f * nk);
                float nn = nk / (twoPi * (1.f - ne2k));
                nn = chk ? nn : ONE_OVER_FOUR_PI;
                float scale = (nrm[i] * _n) / nn;
                float cos0 = mx[i] * nx + (my[i] * ny + mz[i] * nz);
                float cos1 = md.x * nx + (md.y * ny + md.z * nz);
                float eval = pgl_exp(k[i] * (cos0 - 1.f) + _k * (cos1 - 1.f));
                scale *= eval;
                scale *= w[i] * _w;
                w[i] = scale; k[i] = nk; mx[i] = nx; my[i] = ny; mz[i] = nz; nrm[i] = nn; e2k[i] = ne2k;
                acc[l] += w[i];
            }
        float integral = reduce4(acc);
        for (int i = 0; i < c * 4; i++) w[i] /= integral;
        int rem = n % 4;
        if (rem > 0)
            for (int l = rem; l < 4; l++) {
                int i = 4 * (c - 1) + l;
                mx[i] = 0; my[i] = 0; mz[i] = 1; mc[i] = 0; k[i] = 0; nrm[i] = ONE_OVER_FOUR_PI; e2k[i] = 1; dist[i] = 0;
            }
        return integral;
    }
    // :970-1031 selectComponentSIMD ; prefix sum (openpgl_common.h:164-170): (v0, v0+v1, (v1+v2)+v0, (v2+v3)+(v0+v1))
    void selectComponent(int *selVec, int *selComp, float *s1) const {
        const float searched = *s1;
        float sumWeights = 0.f, cdf = 0.f;
        int quot = (n - 1) / 4, rem = (n - 1) % 4;
        float cd[4] = {0, 0, 0, 0};
        int sv = 0;
        for (sv = 0; sv < 8; sv++) {
            const float *v = &w[4 * sv];
            cd[0] = v[0]; cd[1] = v[0] + v[1]; cd[2] = (v[1] + v[2]) + v[0]; cd[3] = (v[2] + v[3]) + (v[0] + v[1]);
            cdf = cd[3];
            if (sumWeights + cdf >= searched || sv >= quot) break;
            else sumWeights += cdf;
        }
        const int maxSel = sv == quot ? rem + 1 : 4;
        for (int l = 0; l < 4; l++) cd[l] += sumWeights;
        // embree::select_min(valid, v): index of the minimum among valid lanes; if none valid -> bsf(movemask(valid)) of 0
        int sc = -1; float best = Infinity; bool any = false;
        for (int l = 0; l < 4; l++) if (cd[l] >= searched) { any = true; if (cd[l] < best) { best = cd[l]; sc = l; } }
        if (!any) sc = 32;  // bsf(0): undefined in the reference; clamped by the min() below
        if (any && sc < 0) sc = 0;
        sc = std::min(sc, maxSel - 1);
        sumWeights = sc > 0 ? cd[sc - 1] : sumWeights;
        cdf = w[4 * sv + sc];
        cdf = std::max(FLT_EPS, cdf);
        *s1 = std::min(1 - FLT_EPS, (searched - sumWeights) / cdf);
        *selVec = sv; *selComp = sc;
    }
    // :1086-1129 sample
    V3 sample(float s0, float s1) const {
        int sv, sc;
        selectComponent(&sv, &sc, &s1);
        int i = 4 * sv + sc;
        float sK = k[i], sE = e2k[i];
        V3 mu(mx[i], my[i], mz[i]);
        V3 dir(0, 0, 1);
        if (sK == 0.0f) {  // openpgl_common.h:330 squareToUniformSphere
            float z = 1.0f - 2.0f * s1;
            float r = std::sqrt(std::max(0.f, (1.0f - z * z)));
            float sp, cp; vspg_sincosf(2.0f * M_PI_F * s0, &sp, &cp);
            dir = V3(r * cp, r * sp, z);
        } else {
            float cosTheta = 1.f + (pgl_log(1.0f + ((sE - 1.f) * s0))) / sK;
            cosTheta = std::min(1.0f, std::max(cosTheta, -1.f));
            float sinTheta = std::sqrt(1.f - cosTheta * cosTheta);
            float phi = 2.f * M_PI_F * s1;
            float sp, cp; vspg_sincosf(phi, &sp, &cp);
            dir = V3(sinTheta * cp, sinTheta * sp, cosTheta);
        }
        // embree::frame(N) linearspace3.h:117-124; normalize(a) = a*rsqrt(dot(a,a)) with the scalar
        // rsqrt of math/emath.h:129-149 (rsqrtss + one Newton step), with the Intel RSQRTSS table of include/vspg_rsqrtss.h.
        V3 dx0(0, mu.z, -mu.y), dx1(-mu.z, 0, mu.x);
        auto dot3 = [](V3 a, V3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); };
        V3 dxs = dot3(dx0, dx0) > dot3(dx1, dx1) ? dx0 : dx1;
        V3 dx = dxs * embree_rsqrt(dot3(dxs, dxs));
        V3 cr(mu.y * dx.z - mu.z * dx.y, mu.z * dx.x - mu.x * dx.z, mu.x * dx.y - mu.y * dx.x);
        V3 dy = cr * embree_rsqrt(dot3(cr, cr));
        // LinearSpace3 * v = v.x*vx + v.y*vy + v.z*vz  (madd chain: x*vx + (y*vy + z*vz))
        return V3(dir.x * dx.x + (dir.y * dy.x + dir.z * mu.x), dir.x * dx.y + (dir.y * dy.y + dir.z * mu.y),
                  dir.x * dx.z + (dir.y * dy.z + dir.z * mu.z));
    }
};

// directional/vmm/VMMPhaseFunctions.h:20-58; the 128-row table itself is DATA dumped from the reference
// library by oracle/ref_harness (tests/golden/hg_vmm_table.bin) and handed in through orc_set_hg_table.
struct HGRep { int K; float g; float weights[4], meanCosines[4], kappas[4], normalizations[4]; };
extern HGRep g_hg_table[128];
extern bool g_hg_table_set;
inline const HGRep &hg_representation(float meanCosine) {
    const float absMeanCosine = std::fabs(meanCosine);
    const float stepSize = (0.99f - 0.f) / float(128);
    int idx = (int)std::floor((absMeanCosine - 0.f) / stepSize);
    idx = std::min(idx, 127);
    return g_hg_table[idx];
}

// VMMSurfaceSamplingDistribution.h:37-101 / VMMVolumeSamplingDistribution.h:35-117,199-204
struct SamplingDistribution {
    VMM li;
    VMM dists[4];
    float weights[4];
    int numDist = 0;
    float productIntegral = 0;
    int region = -1;
    void init(const vspg_mixture *m, V3 samplePos) {
        li.load(m);
        li.parallaxShift(li.pivot - samplePos);
        dists[0] = li; weights[0] = 1.f; numDist = 1; productIntegral = 1.f;
    }
    void applyCosineProduct(V3 normal) {
        const float cosine_kappa = 2.18853f;
        const float cosine_normalization = 2.18853f / (2.0f * M_PI_F * (1.0f - vspg_expf(-2.0f * 2.18853f)));
        if (numDist > 0) productIntegral = dists[0].product(1.0f, normal, cosine_kappa, cosine_normalization);
    }
    void applyHGProduct(V3 dir, float meanCosine) {
        float sumWeights = 0.f;
        const HGRep &pf = hg_representation(meanCosine);
        for (int i = 0; i < pf.K; i++) {
            dists[i] = li;
            V3 outDir = meanCosine * pf.meanCosines[i] > 0.f ? dir : -dir;
            weights[i] = dists[i].product(pf.weights[i], outDir, pf.kappas[i], pf.normalizations[i]);
            sumWeights += weights[i];
        }
        for (int i = 0; i < pf.K; i++) weights[i] /= sumWeights;
        numDist = pf.K; productIntegral = sumWeights;
    }
    V3 sample(float sx, float sy) const {
        float weight = 0.f; int i = 0;
        for (; i < numDist - 1; ++i) {
            float nextWeight = weight + weights[i];
            if (nextWeight > sx) break;
            weight = nextWeight;
        }
        return dists[i].sample((sx - weight) / weights[i], sy);
    }
    float pdf(V3 d) const { float p = 0.f; for (int i = 0; i < numDist; i++) p += weights[i] * dists[i].pdf(d); return p; }
    float samplePdf(float sx, float sy, V3 *dir) const { *dir = sample(sx, sy); return pdf(*dir); }
    float pdfLi(V3 d) const { return li.pdf(d); }
    float vsp(V3 d) const { return li.vsp(d); }
};

}  // namespace orc
