// OpenPGL 0.7.0 (VSP branch) query side on the device: KD-tree walk, stochastic nearest-region pick and the
// parallax-aware vMF mixture operations.  A mixture (<= 32 lobes) is spread over the G lanes of a lane GROUP
// (G = 8, 16 or 32, chosen per vertex from the region's lobe count so that 4*ceil(n/4) <= G; a warp serves 32/G
// vertices at once): lane i of the group holds lobe i (the reference's SSE vector i/4, lane i%4), so every mixture
// op is one pass of lane-parallel math plus a shuffle reduction that keeps the reference's summation order
// (per-SSE-lane accumulators over vectors, then (l0+l1)+(l2+l3)).  The functions below must be called by all G
// lanes of a group together; scalar arguments and results are group-uniform.  Citations relative to /root/reference/src/ext/openpgl/openpgl unless noted.
#pragma once
#include "dscene.cuh"

namespace vspg {

constexpr unsigned kFull = 0xffffffffu;
constexpr float kFltEps = 1.1920928955078125e-7f;

template <int G> VD unsigned grp_mask() {
    if (G == 32) return kFull;
    return ((1u << (G & 31)) - 1u) << ((threadIdx.x & 31) & ~(G - 1));
}
template <int G> VD int grp_lane() { return threadIdx.x & (G - 1); }
template <int G> VD float gshfl(float v, int src) { return __shfl_sync(grp_mask<G>(), v, src, G); }
template <int G> VD float gshfl_xor(float v, int m) { return __shfl_xor_sync(grp_mask<G>(), v, m, G); }

// third-party/embreeSrc/common/math/transcendental.h:333-384
VD float pgl_exp(float v) {
    const float ln2Part1 = 0.6931457519f, ln2Part2 = 1.4286067653e-6f, oneOverLn2 = 1.44269502162933349609375f;
    float scaled = v * oneOverLn2;
    float kReal = floorf(scaled);
    int k = (int)kReal;
    float x = v - kReal * ln2Part1;
    x -= kReal * ln2Part2;
    float r = x * 2.7555381529964506626129150390625e-4f + 1.304379315115511417388916015625e-3f;
    r = x * r + 8.378830738365650177001953125e-3f;
    r = x * r + 4.16539050638675689697265625e-2f;
    r = x * r + 0.166668415069580078125f;
    r = x * r + 0.4999999105930328369140625f;
    r = x * r + 1.f;
    r = x * r + 1.f;
    bool overflow = kReal > 127.f, underflow = kReal <= -127.f;
    float twoToTheN = __int_as_float((int)((uint32_t)(k + 127) << 23));
    r *= twoToTheN;
    if (overflow) r = inff();
    if (underflow) r = 0.f;
    return r;
}
// transcendental.h:388-472
VD float pgl_log(float v) {
    bool useNan = v < 0.f, useInf = v == 0.f;
    bool exceptional = useNan || useInf;
    float patched = exceptional ? 1.f : v;
    int iv = __float_as_int(patched);
    int exponent = (iv >> 23) + 1 - 127;
    float reduced = __int_as_float((iv & (int)0x807FFFFF) | (126 << 23));
    float x1 = 1.f - reduced;
    float r = x1 * 1.7491014003753662109375f + -2.4899270534515380859375f;
    r = x1 * r + 1.98442304134368896484375f;
    r = x1 * r + -0.599632322788238525390625f;
    r = x1 * r + 0.3424419462680816650390625f;
    r = x1 * r + 0.17541764676570892333984375f;
    r = x1 * r + 0.2519190013408660888671875f;
    r = x1 * r + 0.33326041698455810546875f;
    r = x1 * r + 0.50000095367431640625f;
    r = x1 * r + 1.f;
    r *= -x1;
    r += (float)exponent * 0.693147182464599609375f;
    if (exceptional) return useNan ? __int_as_float(0x7fc00000) : -inff();
    return r;
}

// ------------------------------------------------------------------ spatial lookup (per thread / warp-uniform)
// spatial/kdtree/KDTree.h:402-453
VD uint32_t kd_lookup(const DField &f, V3 p) {
    uint32_t idx = 0;
    while (true) {
        uint2 nd = __ldg(reinterpret_cast<const uint2 *>(&f.nodes[idx]));
        uint32_t dim = nd.y >> 30;
        if (dim == 3) return (nd.y << 2) >> 2;
        float pivot = __uint_as_float(nd.x);
        idx = (nd.y << 2) >> 2;
        idx += p[(int)dim] >= pivot ? 1 : 0;
    }
}
VD void cswap(float &a, float &b) { float lo = fminf(a, b), hi = fmaxf(a, b); a = lo; b = hi; }
VD void sort4(float &a, float &b, float &c, float &d) { cswap(a, b); cswap(c, d); cswap(a, c); cswap(b, d); cswap(b, c); }
// spatial/KNN.h:25-32, 88-124  RegionNeighbours<4>::sampleApproximateClosestRegionIdx
VD uint32_t knn_pick(const vspg_neighbours *nh, V3 p, float *sample) {
    uint32_t size = min(4u, __ldg(&nh->size));
    uint32_t selected = (uint32_t)(*sample * size);
    *sample = (*sample - float(selected) / size) * size;
    selected = min(selected, size - 1);
    float d[8];
#pragma unroll
    for (int s = 0; s < 8; s++) {
        float dx = __ldg(&nh->px[s]) - p.x, dy = __ldg(&nh->py[s]) - p.y, dz = __ldg(&nh->pz[s]) - p.z;
        float dist = dx * dx + (dy * dy + dz * dz);
        dist = __uint_as_float((__float_as_uint(dist) & ~7u) | (uint32_t)s);
        if (__ldg(&nh->ids[s]) == ~0u) dist = inff();
        d[s] = dist;
    }
    sort4(d[0], d[1], d[2], d[3]);
    sort4(d[4], d[5], d[6], d[7]);
    // walk `selected` steps of the two-way merge
    float a0 = d[0], a1 = d[1], a2 = d[2], a3 = d[3], b0 = d[4], b1 = d[5], b2 = d[6], b3 = d[7];
    for (uint32_t i = 0; i < selected; i++) {
        if (a0 < b0) { a0 = a1; a1 = a2; a2 = a3; a3 = inff(); }
        else { b0 = b1; b1 = b2; b2 = b3; b3 = inff(); }
    }
    if (a0 < b0) return __ldg(&nh->ids[0 + (__float_as_uint(a0) & 3)]);
    return __ldg(&nh->ids[4 + (__float_as_uint(b0) & 3)]);
}
// field/Field.h:114-150 Field::getRegion.  Returns region index or -1; *leaf gets the KD leaf data index.
VD int field_get_region(const DField &f, V3 p, float *sample1D, int *leaf) {
    *leaf = -1;
    if (!f.present || !f.usable) return -1;
    if (!(p.x >= f.bmin[0] && p.y >= f.bmin[1] && p.z >= f.bmin[2] && p.x <= f.bmax[0] && p.y <= f.bmax[1] && p.z <= f.bmax[2]))
        return -1;
    uint32_t dataIdx = kd_lookup(f, p);
    *leaf = (int)dataIdx;
    if (f.stochastic_nn && *sample1D >= 0.f && f.neigh) return (int)knn_pick(&f.neigh[dataIdx], p, sample1D);
    return (int)dataIdx;
}

// ------------------------------------------------------------------ lane-distributed vMF mixtures
struct Lobe { float w, k, mx, my, mz, nrm, e2k; };
struct LiLobe : Lobe { float mc, dist, vspw; };

// Sum over lobes in the reference's association.  `val` must be 0 on lanes >= 4*cnt.
// Fully unrolled over the G/4 vectors of the group: vectors >= cnt hold exact +0 (every term is a product of
// non-negative factors, never -0), and x + (+0) == x bit for bit, so this equals the reference's loop over cnt vectors
// without the loop-carried branch (the counted loop was 22 % of the scatter kernels' instructions, profiles/r02n).
template <int G> VD float mix_reduce(float val, int cnt) {
    const unsigned mask = grp_mask<G>();
    const int l = threadIdx.x & 3;
    float acc = __shfl_sync(mask, val, l, G);
#pragma unroll
    for (int v = 1; v < G / 4; v++) acc += __shfl_sync(mask, val, 4 * v + l, G);
    float s = acc + __shfl_xor_sync(mask, acc, 1, G);
    return s + __shfl_xor_sync(mask, s, 2, G);
}
VD float dot_e(float ax, float ay, float az, float bx, float by, float bz) { return ax * bx + (ay * by + az * bz); }  // embree madd chain

template <int G> VD void mix_load(const vspg_mixture *m, LiLobe &L, int &n, V3 &pivot) {
    int i = grp_lane<G>();
    L.w = __ldg(&m->weight[i]); L.k = __ldg(&m->kappa[i]);
    L.mx = __ldg(&m->mu_x[i]); L.my = __ldg(&m->mu_y[i]); L.mz = __ldg(&m->mu_z[i]);
    L.nrm = __ldg(&m->norm[i]); L.e2k = __ldg(&m->e_minus_2k[i]); L.mc = __ldg(&m->mean_cos[i]);
    L.dist = __ldg(&m->distance[i]); L.vspw = __ldg(&m->vsp_weight[i]);
    n = __ldg(&m->num_components);
    pivot = V3(__ldg(&m->pivot[0]), __ldg(&m->pivot[1]), __ldg(&m->pivot[2]));
}
// directional/vmm/ParallaxAwareVonMisesFisherMixture.h:1132-1155
template <int G> VD void mix_parallax_shift(LiLobe &L, int cnt, V3 s) {
    int i = grp_lane<G>();
    if (i < 4 * cnt) {
        float px = L.mx * L.dist + s.x, py = L.my * L.dist + s.y, pz = L.mz * L.dist + s.z;
        float len = sqrtf(dot_e(px, py, pz, px, py, pz));
        px /= len; py /= len; pz /= len;
        bool ok = (L.dist > 0.0f) && isfinite(L.dist);
        if (ok) { L.mx = px; L.my = py; L.mz = pz; }
        L.dist = L.dist > 0.0f ? len : 0.f;
    }
}
// :861-880
#ifndef VSPG_MIX_ATTR
#define VSPG_MIX_ATTR __forceinline__   // r02: once the kernels had shrunk (KINDS, unrolled reductions) inlining won 2 % (was a loss in round 1)
#endif
template <int G> __device__ VSPG_MIX_ATTR float mix_pdf(Lobe L, int cnt, V3 d) {
    int i = grp_lane<G>();
    float val = 0.f;
    if (i < 4 * cnt) {
        float cosTheta = dot_e(d.x, d.y, d.z, L.mx, L.my, L.mz);
        float cm1 = fminf(cosTheta - 1.f, 0.f);
        float eval = L.nrm * pgl_exp(L.k * cm1);
        val = L.w * eval;
    }
    return mix_reduce<G>(val, cnt);
}
// :1237-1258
template <int G> __device__ VSPG_MIX_ATTR float mix_vsp(LiLobe L, int cnt, V3 d) {
    int i = grp_lane<G>();
    float e = 0.f, ev = 0.f;
    if (i < 4 * cnt) {
        float cosTheta = dot_e(d.x, d.y, d.z, L.mx, L.my, L.mz);
        float cm1 = fminf(cosTheta - 1.f, 0.f);
        e = L.w * L.nrm * pgl_exp(L.k * cm1);
        ev = L.vspw * e;
    }
    float sv = mix_reduce<G>(ev, cnt), sp = mix_reduce<G>(e, cnt);
    return sv / sp;
}
// :776-858 product(weight, meanDirection, kappa, normalization); returns the product integral
template <int G> VD float mix_product(Lobe &L, int n, int cnt, float _w, V3 md, float _k, float _n) {
    int i = grp_lane<G>();
    const float twoPi = 2.0f * kPi;
    float wv = 0.f;
    if (i < 4 * cnt) {
        float nx = L.k * L.mx + _k * md.x, ny = L.k * L.my + _k * md.y, nz = L.k * L.mz + _k * md.z;
        float nk = sqrtf(dot_e(nx, ny, nz, nx, ny, nz));
        bool chk = nk > 1e-3f;
        nk = chk ? nk : 0.f;
        nx = chk ? nx / nk : L.mx; ny = chk ? ny / nk : L.my; nz = chk ? nz / nk : L.mz;
        float ne2k = pgl_exp(-2.0f * nk);
        float nn = nk / (twoPi * (1.f - ne2k));
        nn = chk ? nn : kInv4Pi;
        float scale = (L.nrm * _n) / nn;
        float cos0 = dot_e(L.mx, L.my, L.mz, nx, ny, nz);
        float cos1 = dot_e(md.x, md.y, md.z, nx, ny, nz);
        float eval = pgl_exp(L.k * (cos0 - 1.f) + _k * (cos1 - 1.f));
        scale *= eval;
        scale *= L.w * _w;
        L.w = scale; L.k = nk; L.mx = nx; L.my = ny; L.mz = nz; L.nrm = nn; L.e2k = ne2k;
        wv = L.w;
    }
    float integral = mix_reduce<G>(wv, cnt);
    if (i < 4 * cnt) L.w /= integral;
    int rem = n % 4;
    if (rem > 0 && i >= 4 * (cnt - 1) + rem && i < 4 * cnt) { L.mx = 0; L.my = 0; L.mz = 1; L.k = 0; L.nrm = kInv4Pi; L.e2k = 1; }
    return integral;
}
// :970-1031 selectComponentSIMD + :1086-1129 sample.  s0 drives cos(theta), s1 picks the lobe and becomes phi.
#ifndef VSPG_MIXS_ATTR
#define VSPG_MIXS_ATTR __noinline__
#endif
template <int G> __device__ VSPG_MIXS_ATTR V3 mix_sample(Lobe L, int n, float s0, float s1) {
    int lane = grp_lane<G>(), l = lane & 3, base = lane & ~3;
    float v0 = gshfl<G>(L.w, base), v1 = gshfl<G>(L.w, base + 1);
    float v2 = gshfl<G>(L.w, base + 2), v3 = gshfl<G>(L.w, base + 3);
    // openpgl_common.h:164-170 prefix sum inside one SSE vector
    float cdl = l == 0 ? v0 : (l == 1 ? v0 + v1 : (l == 2 ? (v1 + v2) + v0 : (v2 + v3) + (v0 + v1)));
    const float searched = s1;
    float sumWeights = 0.f;
    int quot = (n - 1) / 4, rem = (n - 1) % 4;
    int sv = 0;
    for (sv = 0; sv < 8; sv++) {
        float cdf = gshfl<G>(cdl, 4 * sv + 3);
        if (sumWeights + cdf >= searched || sv >= quot) break;
        sumWeights += cdf;
    }
    if (sv > 7) sv = 7;
    const int maxSel = sv == quot ? rem + 1 : 4;
    float cd0 = gshfl<G>(cdl, 4 * sv) + sumWeights, cd1 = gshfl<G>(cdl, 4 * sv + 1) + sumWeights;
    float cd2 = gshfl<G>(cdl, 4 * sv + 2) + sumWeights, cd3 = gshfl<G>(cdl, 4 * sv + 3) + sumWeights;
    // embree select_min(valid, v): lowest lane holding the minimum among valid lanes
    int sc = -1; float best = inff();
    if (cd0 >= searched && cd0 < best) { best = cd0; sc = 0; }
    if (cd1 >= searched && cd1 < best) { best = cd1; sc = 1; }
    if (cd2 >= searched && cd2 < best) { best = cd2; sc = 2; }
    if (cd3 >= searched && cd3 < best) { best = cd3; sc = 3; }
    bool any = cd0 >= searched || cd1 >= searched || cd2 >= searched || cd3 >= searched;
    if (!any) sc = 32;
    if (any && sc < 0) sc = 0;
    sc = min(sc, maxSel - 1);
    float prev = sc == 1 ? cd0 : (sc == 2 ? cd1 : cd2);
    sumWeights = sc > 0 ? prev : sumWeights;
    int sel = 4 * sv + sc;
    float cdfw = fmaxf(kFltEps, gshfl<G>(L.w, sel));
    s1 = fminf(1 - kFltEps, (searched - sumWeights) / cdfw);
    float sK = gshfl<G>(L.k, sel), sE = gshfl<G>(L.e2k, sel);
    V3 mu(gshfl<G>(L.mx, sel), gshfl<G>(L.my, sel), gshfl<G>(L.mz, sel));
    V3 dir(0, 0, 1);
    if (sK == 0.0f) {  // openpgl_common.h:330 squareToUniformSphere
        float z = 1.0f - 2.0f * s1;
        float r = sqrtf(fmaxf(0.f, (1.0f - z * z)));
        float sp, cp; vspg_sincosf(2.0f * kPi * s0, &sp, &cp);   // sincosf(2.0f * M_PI_F * s0)
        dir = V3(r * cp, r * sp, z);
    } else {
        float cosTheta = 1.f + (pgl_log(1.0f + ((sE - 1.f) * s0))) / sK;
        cosTheta = fminf(1.0f, fmaxf(cosTheta, -1.f));
        float sinTheta = sqrtf(1.f - cosTheta * cosTheta);
        float phi = 2.f * kPi * s1;
        float sp, cp; vspg_sincosf(phi, &sp, &cp);
        dir = V3(sinTheta * cp, sinTheta * sp, cosTheta);
    }
    // embree::frame(N), linearspace3.h:117-124; normalize(a) = a * rsqrt(dot(a, a)) with embree::rsqrt = RSQRTSS estimate
    // (Intel's lookup, include/vspg_rsqrtss.h) + one Newton step
    V3 dx0(0, mu.z, -mu.y), dx1(-mu.z, 0, mu.x);
    float l0 = dot_e(dx0.x, dx0.y, dx0.z, dx0.x, dx0.y, dx0.z), l1 = dot_e(dx1.x, dx1.y, dx1.z, dx1.x, dx1.y, dx1.z);
    V3 dxs = l0 > l1 ? dx0 : dx1;
    V3 dx = dxs * vspg_embree_rsqrt(l0 > l1 ? l0 : l1);
    V3 cr(mu.y * dx.z - mu.z * dx.y, mu.z * dx.x - mu.x * dx.z, mu.x * dx.y - mu.y * dx.x);
    V3 dy = cr * vspg_embree_rsqrt(dot_e(cr.x, cr.y, cr.z, cr.x, cr.y, cr.z));
    return V3(dir.x * dx.x + (dir.y * dy.x + dir.z * mu.x), dir.x * dx.y + (dir.y * dy.y + dir.z * mu.y),
              dir.x * dx.z + (dir.y * dy.z + dir.z * mu.z));
}

// VMMSurfaceSamplingDistribution.h:37-101 / VMMVolumeSamplingDistribution.h:35-117,199-204
template <int G>
struct SamplingDist {
    LiLobe li;
    Lobe d0, d1, d2;
    float w0, w1, w2;
    int numDist, n, cnt;

    VD void init(const vspg_mixture *m, V3 samplePos) {
        V3 pivot; mix_load<G>(m, li, n, pivot);
        cnt = (n + 3) / 4;
        mix_parallax_shift<G>(li, cnt, pivot - samplePos);
        d0 = li; w0 = 1.f; w1 = w2 = 0.f; numDist = 1;
    }
    VD void apply_cosine_product(V3 normal) {
        const float cosine_kappa = 2.18853f;
        // VMMSurfaceSamplingDistribution.h:55: 2.18853f/(2*pi*(1-exp(-2*2.18853f))) as the CPU evaluates it in float
        const float cosine_normalization = 0x1.69366ap-2f;
        mix_product<G>(d0, n, cnt, 1.0f, normal, cosine_kappa, cosine_normalization);
    }
    VD void apply_hg_product(const DHGRow *table, V3 dir, float meanCosine) {
        const float stepSize = (0.99f - 0.f) / float(128);
        int idx = min((int)floorf((fabsf(meanCosine) - 0.f) / stepSize), 127);
        const DHGRow &pf = table[idx];
        float sumWeights = 0.f;
        int K = pf.K;
        if (K > 0) { d0 = li; V3 od = meanCosine * pf.mc[0] > 0.f ? dir : -dir; w0 = mix_product<G>(d0, n, cnt, pf.w[0], od, pf.kappa[0], pf.norm[0]); sumWeights += w0; }
        if (K > 1) { d1 = li; V3 od = meanCosine * pf.mc[1] > 0.f ? dir : -dir; w1 = mix_product<G>(d1, n, cnt, pf.w[1], od, pf.kappa[1], pf.norm[1]); sumWeights += w1; }
        if (K > 2) { d2 = li; V3 od = meanCosine * pf.mc[2] > 0.f ? dir : -dir; w2 = mix_product<G>(d2, n, cnt, pf.w[2], od, pf.kappa[2], pf.norm[2]); sumWeights += w2; }
        if (K > 0) w0 /= sumWeights;
        if (K > 1) w1 /= sumWeights;
        if (K > 2) w2 /= sumWeights;
        numDist = K;
    }
    VD float pdf(V3 d) const {
        float p = 0.f;
        if (numDist > 0) p += w0 * mix_pdf<G>(d0, cnt, d);
        if (numDist > 1) p += w1 * mix_pdf<G>(d1, cnt, d);
        if (numDist > 2) p += w2 * mix_pdf<G>(d2, cnt, d);
        return p;
    }
    VD V3 sample(float sx, float sy) const {
        float weight = 0.f; int i = 0;
        for (; i < numDist - 1; ++i) {
            float wi = i == 0 ? w0 : w1;
            float nextWeight = weight + wi;
            if (nextWeight > sx) break;
            weight = nextWeight;
        }
        float wsel = i == 0 ? w0 : (i == 1 ? w1 : w2);
        float s0 = (sx - weight) / wsel;
        Lobe L = i == 0 ? d0 : (i == 1 ? d1 : d2);
        return mix_sample<G>(L, n, s0, sy);
    }
    VD float sample_pdf(float sx, float sy, V3 *dir) const { *dir = sample(sx, sy); return pdf(*dir); }
    VD float pdf_li(V3 d) const { return mix_pdf<G>(li, cnt, d); }
    VD float vsp(V3 d) const { return mix_vsp<G>(li, cnt, d); }
};

}  // namespace vspg
