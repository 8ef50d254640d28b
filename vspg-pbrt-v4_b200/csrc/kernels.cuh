// Wavefront stage kernels of the GuidedVolPathVSPGIntegrator render pass (Li, vspg.cpp:262-635, re-expressed as
// queues of path slots).  One iteration = intersect -> medium -> volume-scatter | surface -> shade -> shadow -> finish.
#pragma once
#include "dsampledistance.cuh"
#include "dvertex.cuh"

namespace vspg {

constexpr int kBlock = 128;
// resident CTAs per SM the thread-per-item kernels are compiled for (A/B knobs; 1 = let the compiler choose)
#ifndef VSPG_LOOKUP_MINB
#define VSPG_LOOKUP_MINB 8   // latency-bound KD walk: 64 registers, 8 CTAs per SM (C2 -0.7 ms)
#endif
#ifndef VSPG_INTERSECT_MINB
#define VSPG_INTERSECT_MINB 1
#endif
#ifndef VSPG_SCATTER0_MINB
#define VSPG_SCATTER0_MINB 4   // CTAs / SM of the one-thread-per-vertex scatter kernels (k_volscatter<0>, k_shade<0>: unguided configurations; C1 shade 1.77 -> 1.67 ms, 6 / 8 the same, 16 none)
#endif

VD int gtid() { return blockIdx.x * blockDim.x + threadIdx.x; }
VD int gstride() { return gridDim.x * blockDim.x; }

// ------------------------------------------------------------------ regenerate: new camera samples into free slots
VD bool regenerate(const DScene &sc, const Wavefront &W, int slot, PathRegs &r) {
    // warp-aggregated fetch of the next sample id
    unsigned mask = __activemask();
    int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(W.next_sample, (unsigned long long)__popc(mask));
    base = __shfl_sync(mask, base, leader);
    unsigned long long id = base + __popc(mask & ((1u << lane) - 1));
    const RenderVars rv = *W.rv;
    if (id >= rv.total_samples) return false;
    const int w = rv.x1 - rv.x0, npix = w * (rv.y1 - rv.y0);
    int s, pix;
    if (rv.band_rows > 0) {
        // band-major order: all sample passes of a band of rows before the next band, so that the paths in flight stay in
        // one slab of the scene (a DRAM-resident grid is then read through L2; measured on C4, profiles/r02_summary.md)
        const unsigned long long perBand = (unsigned long long)w * rv.band_rows * rv.n_spp;
        const int band = (int)(id / perBand);
        const unsigned long long r = id % perBand;
        const int yb = band * rv.band_rows, rows = min(rv.band_rows, (rv.y1 - rv.y0) - yb), bandPix = w * rows;
        s = (int)(r / (unsigned long long)bandPix);
        pix = yb * w + (int)(r % (unsigned long long)bandPix);
    } else { s = (int)(id / (unsigned long long)npix); pix = (int)(id % (unsigned long long)npix); }
    int px = rv.x0 + pix % w, py = rv.y0 + pix / w;
    generate_camera_path(sc, rv.seed, px, py, rv.spp_begin + s * rv.spp_stride, r);
    path_store(W.P, slot, r);
    return true;
}

__global__ void __launch_bounds__(kBlock) k_generate(const DScene sc, const Wavefront W, int nslots) {
    for (int slot = gtid(); slot < nslots; slot += gstride()) {
        PathRegs r;
        bool ok = regenerate(sc, W, slot, r);
        queue_push(W.active, ok, slot);
    }
}

// ------------------------------------------------------------------ closest hit (vspg.cpp:314)
template <int KINDS = kKindsAll>
__global__ void __launch_bounds__(kBlock, VSPG_INTERSECT_MINB) k_intersect(const DScene sc, const Wavefront W) {
    const int n = *W.active.count;
    unsigned nodes = 0, preSteps = 0;
    for (int i = gtid(); i < n; i += gstride()) {
        int slot = W.active.items[i];
        float4 a = W.P.o_dx[slot], b = W.P.d_med[slot];
        V3 o(a.x, a.y, a.z), d(a.w, b.x, b.y);
        int medium = __float_as_int(b.z);
        Hit h = scene_intersect<KINDS>(sc, o, d, inff(), nodes);
        W.P.hit[slot] = make_float4(h.t, h.u, h.v, __int_as_float(h.prim));
        bool toMedium = medium >= 0 && h.prim >= 0;  // vspg.cpp:318: rays escaping to infinity skip the medium
        // vspg.cpp:679 use_resampling: the persistent-lane kernel takes those, the generic kernel the rest
        bool rs = toMedium && !sc.ip.vsp_sampling_nds && sc.media[medium].type != VSPG_MEDIUM_HOMOGENEOUS;
        if (rs) {
            // pre-pass of SampleT_maj_Resampling (media_sampleTMaj.h:153-164): total majorant optical depth of the
            // hero channel along [0, tHit]; done here, one thread per ray, so the persistent walk kernel starts sampling
            const int ch = PathFlags::unpack(__float_as_int(b.w)).channel;
            Ray ray; ray.o = o; ray.medium = medium;
            float tMax = h.t * length(d);
            ray.d = normalize(d);
            MajorantIterator pre = sample_ray(sc.media[medium], ray, tMax);
            float totalLength = 0.f;
            Segment seg;
            while (pre.next(&seg)) {
                if (seg.sigma_maj[ch] == 0) continue;
                totalLength += seg.sigma_maj[ch] * (seg.tMax - seg.tMin);
            }
            W.P.rs_pre[slot] = make_float2(totalLength, __int_as_float((int)pre.steps));
            preSteps += pre.steps;
        }
        queue_push(W.medium_rs, rs, slot);
        queue_push(W.medium, toMedium && !rs, slot);
        queue_push(W.surface, !toMedium, slot);
    }
    counter_add(W.counters, CNT_BVH, nodes);
    counter_add(W.counters, CNT_DDA_PRE, preSteps);
}

// ------------------------------------------------------------------ SampleDistance (vspg.cpp:318-351, 637-1096)
// One thread per ray, start to end of the delta-tracking walk.  The walk is a chain of dependent gathers (majorant cell,
// density cell) and the kernel is bound by their latency, not by registers: on C3 (NDS sampling over a 128^3 grid) the medium
// stage went 63.6 -> 59.9 -> 55.3 -> 50.6 -> 47.1 -> 43.5 ms at 3 (ptxas' own choice: 168 registers) / 4 / 6 / 8 / 12 / 16
// CTAs per SM -- 32 registers and 4.6 KB of spill traffic per thread at full occupancy still win (profiles/r02_summary.md
// section 7).  Scenes whose media are all homogeneous have no gathers to hide and keep the unconstrained build (C1: 0.36 vs
// 0.47 ms).  Three plain kernels from one macro (emissive scenes always take the
// full-occupancy one): as templates the same source compiled to
// a 1392-byte stack frame (dsampledistance_body.inl).
#ifndef VSPG_MEDIUM_MINB
#define VSPG_MEDIUM_MINB 16
#endif
#define VSPG_DEFINE_K_MEDIUM(NAME, BOUNDS, SAMPLE_DISTANCE)                                               \
    __global__ void BOUNDS NAME(const DScene sc, const Wavefront W) {                                     \
        const int n = *W.medium.count;                                                                    \
        TrackCounters tc; tc.queries = 0; tc.dda = 0;                                                     \
        for (int i = gtid(); i < n; i += gstride()) {                                                     \
            int slot = W.medium.items[i];                                                                 \
            PathRegs ps; path_load(W.P, slot, ps);                                                        \
            float tHit = W.P.hit[slot].x;                                                                 \
            DistanceResult res;                                                                           \
            SAMPLE_DISTANCE(sc, ps, tHit, tc, res);                                                       \
            path_store(W.P, slot, ps);                                                                    \
            bool dead = res.terminated || !nonzero(ps.beta) || !nonzero(ps.r_u);                          \
            bool scat = !dead && res.scattered;                                                           \
            if (scat) W.P.ev[slot] = make_float4(res.p.x, res.p.y, res.p.z, res.mp.g);                    \
            queue_push(W.done, dead, slot);                                                               \
            queue_push(W.volscatter, scat, slot);                                                         \
            queue_push(W.surface, !dead && !scat, slot);                                                  \
        }                                                                                                 \
        counter_add(W.counters, CNT_QUERIES, tc.queries);                                                 \
        counter_add(W.counters, CNT_DDA, tc.dda);                                                         \
    }
VSPG_DEFINE_K_MEDIUM(k_medium, __launch_bounds__(kBlock, VSPG_MEDIUM_MINB), sample_distance)             // some grid medium in the scene
VSPG_DEFINE_K_MEDIUM(k_medium_emit, __launch_bounds__(kBlock, VSPG_MEDIUM_MINB), sample_distance_emit)   // some medium emits (homogeneous or grid)
VSPG_DEFINE_K_MEDIUM(k_medium_hom, __launch_bounds__(kBlock), sample_distance)                           // homogeneous media only

// ------------------------------------------------------------------ shadow-ray record (SampleLd front half, vspg.cpp:1136-1192)
// f_hat, L and beta travel apart: the reference forms ((f_hat * T_ray) * L) / w first and multiplies by beta last
// (vspg.cpp:1247-1250, :479 / :833), and float products do not re-associate
struct ShadowRec { V3 o, d; int medium; V3 f_hat, L, beta, rl, ru; V3 pLight, pLightErr, nLight; bool delta; int channel; };
VD void shadow_store(const PathSoA &P, int s, const ShadowRec &r) {
    P.sh_o[s] = make_float4(r.o.x, r.o.y, r.o.z, r.d.x);
    P.sh_d[s] = make_float4(r.d.y, r.d.z, __int_as_float(r.medium), __int_as_float((r.delta ? 1 : 0) | (r.channel << 8)));
    P.sh_c[s] = make_float4(r.L.x, r.L.y, r.L.z, r.beta.x);
    P.sh_rl[s] = make_float4(r.rl.x, r.rl.y, r.rl.z, r.beta.y);
    P.sh_f[s] = make_float4(r.f_hat.x, r.f_hat.y, r.f_hat.z, r.beta.z);
    P.sh_ru[s] = make_float4(r.ru.x, r.ru.y, r.ru.z, r.nLight.z);
    P.sh_pl[s] = make_float4(r.pLight.x, r.pLight.y, r.pLight.z, r.nLight.x);
    P.sh_ple[s] = make_float4(r.pLightErr.x, r.pLightErr.y, r.pLightErr.z, r.nLight.y);
}

// Next-event light sample drawn ahead of the lane-group scatter kernels.  The light pick and Light::SampleLi are
// scalar work (one result per vertex), so for guided configurations k_lookup -- one THREAD per vertex -- computes
// them from a copy of the vertex's sampler stream and parks the result in the slot's (still unused) shadow-record
// streams; the G-lane kernels then only skip the three sampler dimensions and read the record back.  Same
// arithmetic, same stream positions: bit-identical to sampling inside the scatter kernel.
struct NeePre { bool valid, delta; V3 L, wi, pLight, pLightErr, nLight; float p_l; };
// second half of the pre-sample: Light::SampleLi for the picked light
template <int KINDS = kKindsAll>
VD NeePre nee_finish(const DScene &sc, int light, float pmf, V3 ctxP, V3 ctxNs, float ul0, float ul1) {
    NeePre np; np.valid = false; np.delta = false; np.p_l = 0;
    if (light < 0) return np;
    LiSample ls = sample_li<KINDS>(sc, light, ctxP, ctxNs, ul0, ul1);
    if (!(ls.valid && nonzero(ls.L) && ls.pdf != 0)) return np;
    np.valid = true; np.delta = ls.delta; np.L = ls.L; np.wi = ls.wi; np.pLight = ls.pLight; np.pLightErr = ls.pLightErr; np.nLight = ls.nLight;
    np.p_l = pmf * ls.pdf;
    return np;
}
template <int KINDS = kKindsAll>
VD NeePre nee_presample(const DScene &sc, Rng rng, V3 ctxP, V3 ctxNs) {
    float u = rng.uniform();
    float pmf = 0;
    int light = sample_light(sc, u, &pmf, ctxP, ctxNs);
    float ul0 = rng.uniform(), ul1 = rng.uniform();
    return nee_finish<KINDS>(sc, light, pmf, ctxP, ctxNs, ul0, ul1);
}
VD void nee_store(const PathSoA &P, int s, const NeePre &n) {
    P.sh_rl[s] = make_float4(n.wi.x, n.wi.y, n.wi.z, __int_as_float((n.valid ? 1 : 0) | (n.delta ? 2 : 0)));
    if (!n.valid) return;
    P.sh_c[s] = make_float4(n.L.x, n.L.y, n.L.z, n.p_l);
    P.sh_pl[s] = make_float4(n.pLight.x, n.pLight.y, n.pLight.z, n.nLight.x);
    P.sh_ple[s] = make_float4(n.pLightErr.x, n.pLightErr.y, n.pLightErr.z, n.nLight.y);
    P.sh_ru[s] = make_float4(0.f, 0.f, 0.f, n.nLight.z);   // same 32-byte record as sh_pl
}
VD NeePre nee_load(const PathSoA &P, int s) {
    NeePre n;
    float4 a = P.sh_rl[s];
    int fl = __float_as_int(a.w);
    n.valid = fl & 1; n.delta = (fl >> 1) & 1; n.wi = V3(a.x, a.y, a.z); n.p_l = 0;
    if (n.valid) {
        float4 c = P.sh_c[s], pl = P.sh_pl[s], pe = P.sh_ple[s];
        n.L = V3(c.x, c.y, c.z); n.p_l = c.w;
        n.pLight = V3(pl.x, pl.y, pl.z); n.pLightErr = V3(pe.x, pe.y, pe.z); n.nLight = V3(pl.w, pe.w, P.sh_ru[s].w);
    }
    return n;
}
VD void shadow_load(const PathSoA &P, int s, ShadowRec &r) {
    float4 a = P.sh_o[s], b = P.sh_d[s], c = P.sh_c[s], d = P.sh_rl[s], e = P.sh_ru[s], f = P.sh_pl[s], g = P.sh_ple[s];
    r.o = V3(a.x, a.y, a.z); r.d = V3(a.w, b.x, b.y); r.medium = __float_as_int(b.z);
    int fl = __float_as_int(b.w); r.delta = fl & 1; r.channel = (fl >> 8) & 3;
    float4 ff = P.sh_f[s];
    r.L = V3(c.x, c.y, c.z); r.beta = V3(c.w, d.w, ff.w); r.f_hat = V3(ff.x, ff.y, ff.z); r.rl = V3(d.x, d.y, d.z); r.ru = V3(e.x, e.y, e.z);
    r.pLight = V3(f.x, f.y, f.z); r.pLightErr = V3(g.x, g.y, g.z); r.nLight = V3(f.w, g.w, e.w);
}
// interaction.h:105-109 SpawnRayTo(pFrom, pTo): both ends offset (ray.h:104-109)
VD void spawn_ray_to(V3 pf_, V3 pfErr, V3 nf, V3 pLight, V3 pLightErr, V3 nLight, V3 *o, V3 *d) {
    V3 pf = offset_ray_origin(pf_, pfErr, nf, pLight - pf_);
    V3 pt = offset_ray_origin(pLight, pLightErr, nLight, pf - pLight);
    *o = pf; *d = pt - pf;
}

// ------------------------------------------------------------------ volume scatter vertex (vspg.cpp:804-875 / 988-1058)
// Lane-group scheduling shared by the two scatter-vertex kernels: G = 0 -> one thread per vertex; G = 8/16/32 ->
// one G-lane group per vertex (32/G vertices per warp).  The warp iterates uniformly; a group whose item index is
// past the end re-runs the warp's first item with every side effect suppressed (`live` false) so that the
// group-mask shuffles of the mixture code stay well-formed.
//
// The scatter kernels are ~90 KB of mostly straight-line SASS; with the warps of an SM scattered over that much code
// the top stall was instruction fetch (ncu r01c: no_instruction 3.8-5.9 per issue).  Trip counts are therefore
// uniform over the whole CTA (`base` is a CTA quantity) and the loop body carries __syncthreads() between its
// phases (scatter_phase_sync), so that the warps of a CTA walk the code together and share instruction-cache lines.
#ifndef VSPG_VOL_MINB
#define VSPG_VOL_MINB 3                   // resident CTAs/SM the volume scatter kernel is compiled for (A/B knob, see profiles/)
#endif
constexpr int kScatterBlock = 256;        // launch bound and default block of the lane-group scatter kernels (r01e A/B: 512 -> 116.5, 256 -> 118.0, 128 -> 109.4 Mpaths/s)
template <int G> struct GroupIter {
    static constexpr int W = G > 0 ? G : 1;        // lanes per vertex
    int first, step, warpOff, sub; bool leader;
    VD GroupIter() {
        if (G > 0) {
            const int perCta = (blockDim.x >> 5) * (32 / W);
            first = blockIdx.x * perCta; step = gridDim.x * perCta;
            warpOff = (threadIdx.x >> 5) * (32 / W); sub = (threadIdx.x & 31) / W; leader = (threadIdx.x & (W - 1)) == 0;
        } else { first = blockIdx.x * blockDim.x; step = gstride(); warpOff = threadIdx.x; sub = 0; leader = true; }
    }
};
#ifndef VSPG_SCATTER_SYNC
#define VSPG_SCATTER_SYNC 1
#endif
// bit i = phase boundary i keeps its CTA barrier.  Measured (profiles/r02_summary.md): the volume kernel is best with the loop-top
// barrier only (C2 volscatter 29.1 -> 28.0 ms), the surface kernel with all four
#ifndef VSPG_VOL_SYNC_SITES
#define VSPG_VOL_SYNC_SITES 1
#endif
#ifndef VSPG_SHADE_SYNC_SITES
#define VSPG_SHADE_SYNC_SITES 15
#endif
template <int G, int SITE = 0, int SITES = 15> VD void scatter_phase_sync() {
    if (G > 0) { if (VSPG_SCATTER_SYNC && ((SITES >> SITE) & 1)) __syncthreads(); else __syncwarp(); }
}

// ------------------------------------------------------------------ Field::getRegion for a whole stage queue
// (field/Field.h:114-150; guiding.h:83-109 / 383-398 front half).  One THREAD per vertex: draws the sampler's `v`,
// walks the KD tree, picks the stochastic neighbour, and sorts the vertex into the queue of the lane-group width its
// mixture needs (<= 8 lobes -> G8, <= 16 -> G16, else G32; vertices without a usable region go to G8).
// Area-light samples (spherical rectangle / triangle / cone sampling: ~10x the work of a point or distant light) are not
// taken where they are drawn: the CTA collects them in shared memory and then works the list off with consecutive
// threads, so that the heavy code runs on full warps instead of on the few lanes per warp that picked such a light
// (profiles/r02r: 8.5 of 32 lanes).
struct HeavyNee { int slot, light; float pmf, u0, u1, p[3], ns[3]; };
template <bool SURFACE, int KINDS = kKindsAll>
__global__ void __launch_bounds__(kBlock, VSPG_LOOKUP_MINB) k_lookup(const DScene sc, const Wavefront W, const Queue in, const Queue q8, const Queue q16, const Queue q32) {
    __shared__ HeavyNee heavy[kBlock];
    __shared__ int heavyCount;
    const int n = *in.count;
    unsigned lookups = 0;
    for (int base = blockIdx.x * blockDim.x; base < n; base += gstride()) {
        if (threadIdx.x == 0) heavyCount = 0;
        __syncthreads();
        const int i = base + threadIdx.x;
        if (i < n) {
            int slot = in.items[i];
            ulonglong2 q = W.P.rng[slot];
            Rng rng; rng.state = q.x; rng.inc = q.y;
            float v = rng.uniform();
            W.P.rng[slot] = make_ulonglong2(rng.state, rng.inc);
            V3 p; bool want = true;
            if (SURFACE) {
                float4 a = W.P.o_dx[slot], b = W.P.d_med[slot], hr = W.P.hit[slot];
                p = V3(a.x, a.y, a.z) + hr.x * V3(a.w, b.x, b.y);
                const DQuad &Q = sc.quads[__float_as_int(hr.w)];
                want = nonzero(ld3(sc.materials[Q.material].reflectance));   // GuidedBSDF::init only looks up for non-specular BSDFs
            } else { float4 e = W.P.ev[slot]; p = V3(e.x, e.y, e.z); }
            const DField &f = SURFACE ? sc.surf : sc.vol;
            int r = -1, nl = 0;
            if (want) {
                int leaf;
                r = field_get_region(f, p, &v, &leaf);
                if (leaf >= 0) lookups++;
                if (r >= 0) { if (__ldg(&f.regions[r].valid)) nl = __ldg(&f.regions[r].num_components); else r = -1; }
            }
            W.P.lk[slot] = r;
            if (sc.ip.use_nee && want) {
                V3 ctxP = p, ctxNs;
                if (SURFACE) {   // GuidedBSDF::non_specular(): reflective-only BSDF, ctx.pi = intr.OffsetRayOrigin(intr.wo) (vspg.cpp:1147-1149)
                    float4 a = W.P.o_dx[slot], b = W.P.d_med[slot], hr = W.P.hit[slot];
                    Hit h; h.t = hr.x; h.u = hr.y; h.v = hr.z; h.prim = __float_as_int(hr.w);
                    V3 ro(a.x, a.y, a.z), rd(a.w, b.x, b.y);
                    SurfaceIntr isect = make_intr<KINDS>(sc, h, ro, rd, __float_as_int(b.z));
                    ctxP = offset_ray_origin(isect.p, isect.pError, isect.n, isect.wo); ctxNs = isect.ns;
                }
                // nee_presample, with SampleLi of area lights deferred (rng is a copy: the path's stream is not advanced here)
                float u = rng.uniform();
                float pmf = 0;
                int light = sample_light(sc, u, &pmf, ctxP, ctxNs);
                float ul0 = rng.uniform(), ul1 = rng.uniform();
                if (light >= 0 && sc.lights[light].type == VSPG_LIGHT_DIFFUSE_AREA) {
                    HeavyNee &hn = heavy[atomicAdd(&heavyCount, 1)];
                    hn.slot = slot; hn.light = light; hn.pmf = pmf; hn.u0 = ul0; hn.u1 = ul1;
                    hn.p[0] = ctxP.x; hn.p[1] = ctxP.y; hn.p[2] = ctxP.z; hn.ns[0] = ctxNs.x; hn.ns[1] = ctxNs.y; hn.ns[2] = ctxNs.z;
                } else nee_store(W.P, slot, nee_finish<KINDS>(sc, light, pmf, ctxP, ctxNs, ul0, ul1));
            }
            queue_push(q8, nl <= 8, slot);
            queue_push(q16, nl > 8 && nl <= 16, slot);
            queue_push(q32, nl > 16, slot);
        }
        __syncthreads();
        for (int k = threadIdx.x; k < heavyCount; k += blockDim.x) {
            const HeavyNee hn = heavy[k];
            nee_store(W.P, hn.slot, nee_finish<KINDS>(sc, hn.light, hn.pmf, ld3(hn.p), ld3(hn.ns), hn.u0, hn.u1));
        }
        __syncthreads();
    }
    counter_add(W.counters, CNT_LOOKUPS, lookups);
}

// ------------------------------------------------------------------ volume scatter vertex (vspg.cpp:804-875 / 988-1058)
// Register discipline (profiles/r02n: the 64-register build spilled 540 B per lane and its local-memory sectors were 5x
// its global ones): the G lanes of a group only carry what the mixture operations need -- the sampling distribution,
// the sampler, p and the incoming direction.  The path record is never held as a whole: the Russian-roulette factor
// is formed from a transient read of beta / r_u, NEE only evaluates the phase function and the guided pdf towards the
// pre-drawn light sample, and the group's leader assembles the shadow-ray record and the outgoing path record in a
// scalar epilogue, straight from the SoA streams, once the mixture is dead.
template <int G>
__global__ void __launch_bounds__(G > 0 ? 256 : kBlock, G > 0 ? VSPG_VOL_MINB : VSPG_SCATTER0_MINB) k_volscatter(const DScene sc, const Wavefront W, const Queue in) {
    const int n = *in.count;
    const GroupIter<G> gi;
    const bool leader = gi.leader;
    const vspg_integrator_params &ip = sc.ip;
    unsigned lookups = 0, shadows = 0;
    for (int base = gi.first; base < n; base += gi.step) {
        const int i0 = base + gi.warpOff;
        const bool live = i0 + gi.sub < n;
        if (G == 0 && !live) continue;             // one thread per vertex: no group shuffles, no barriers
        scatter_phase_sync<G, 0, VSPG_VOL_SYNC_SITES>();
        const int slot = in.items[live ? i0 + gi.sub : base];
        V3 p, rd; float g_; int flw;
        {
            float4 e = W.P.ev[slot], a = W.P.o_dx[slot], b = W.P.d_med[slot];
            p = V3(e.x, e.y, e.z); g_ = e.w; rd = V3(a.w, b.x, b.y); flw = __float_as_int(b.w);
        }
        const int depth = flw & 0xff;
        Rng rng; { ulonglong2 q = W.P.rng[slot]; rng.state = q.x; rng.inc = q.y; }
        GuidedPhase<G> gphase;
        gphase.enableGuiding = ip.volume_guiding; gphase.enableScatterGuiding = ip.vsp_secondary_guiding; gphase.ris = ip.volume_guiding_ris;
        if (G > 0) gphase.init_region(sc, g_, p, rd, W.P.lk[slot]);   // `v` was drawn by k_lookup
        else { float v = rng.uniform(); gphase.init(sc, g_, p, rd, v, lookups); }
        scatter_phase_sync<G, 1, VSPG_VOL_SYNC_SITES>();
        float survivalProb = 1.0f;
        if (depth > ip.min_rr_depth) {
            float4 c = W.P.beta_ru[slot], d = W.P.ru_rl[slot];
            V3 rrw = (V3(c.x, c.y, c.z) / avg3(V3(c.w, d.x, d.y))) * W.P.misc[slot].x;   // openpgl/cpp/RussianRoulette.h:32-35
            survivalProb = ((flw >> 10) & 1) ? 0.95f : fmaxf(0.f, fminf(1.f, max_comp(rrw)));
        }
        bool emitShadow = false;
        float nee_ph = 0, nee_pdf = 0;
        if (ip.use_nee) {
            bool valid; V3 wi;
            if (G > 0) {   // drawn by k_lookup from these three sampler dimensions
                rng.uniform(); rng.uniform(); rng.uniform();
                // A lane group past the end of the queue re-runs the CTA's first item for the barriers only: it must not read
                // that slot's parked record, which the owning group's leader (possibly in another warp) overwrites below.
                float4 r = live ? W.P.sh_rl[slot] : make_float4(0, 0, 0, 0);
                valid = __float_as_int(r.w) & 1; wi = V3(r.x, r.y, r.z);
            } else {
                NeePre ls = nee_presample(sc, rng, p, V3());
                rng.uniform(); rng.uniform(); rng.uniform();
                nee_store(W.P, slot, ls);      // the epilogue reads the record back like the lane-group builds do
                valid = ls.valid; wi = ls.wi;
            }
            if (valid) {
                nee_ph = gphase.p(-rd, wi);
                nee_pdf = 1.0f * gphase.pdf(-rd, wi);
                emitShadow = nee_ph != 0;
            }
        }
        scatter_phase_sync<G, 2, VSPG_VOL_SYNC_SITES>();
        bool alive = true, rr = false;
        if (survivalProb < 1 && depth > ip.min_rr_depth) {
            float q = fmaxf(0.f, 1 - survivalProb);
            if (rng.uniform() < q) alive = false;
            else rr = true;
        }
        V3 wiNext; float pfw = 0, sPdf = 0, vspNext = 0;
        if (alive) {
            float u0 = rng.uniform(), u1 = rng.uniform();
            ScatterSample s = gphase.sample_p(-rd, u0, u1, rng);
            if (!s.valid || s.pdf == 0) alive = false;
            else {
                pfw = s.f.x / s.pdf; sPdf = s.pdf; wiNext = s.wi;
                vspNext = gphase.vsp(s.wi);
            }
        }
        scatter_phase_sync<G, 3, VSPG_VOL_SYNC_SITES>();
        const bool w = leader && live;
        if (w && (emitShadow || alive)) {
            float4 c = W.P.beta_ru[slot], d = W.P.ru_rl[slot];
            V3 beta(c.x, c.y, c.z), r_u(c.w, d.x, d.y);
            if (emitShadow) {
                // sh_pl / sh_ple / sh_ru.w (light point, its error bounds, its normal) and sh_c.xyz (L) already hold the
                // shadow record's values: the parked NEE record shares their layout
                float4 lc = W.P.sh_c[slot], pl = W.P.sh_pl[slot], pe = W.P.sh_ple[slot], nr = W.P.sh_rl[slot];
                const float nz = W.P.sh_ru[slot].w;
                V3 so, sd;
                spawn_ray_to(p, V3(), V3(), V3(pl.x, pl.y, pl.z), V3(pe.x, pe.y, pe.z), V3(pl.w, pe.w, nz), &so, &sd);
                const int delta = (__float_as_int(nr.w) >> 1) & 1, channel = (flw >> 8) & 3;
                V3 rl = r_u * lc.w, ru = r_u * nee_pdf;
                W.P.sh_o[slot] = make_float4(so.x, so.y, so.z, sd.x);
                W.P.sh_d[slot] = make_float4(sd.y, sd.z, W.P.d_med[slot].z, __int_as_float(delta | (channel << 8)));
                W.P.sh_c[slot] = make_float4(lc.x, lc.y, lc.z, beta.x);
                W.P.sh_rl[slot] = make_float4(rl.x, rl.y, rl.z, beta.y);
                W.P.sh_f[slot] = make_float4(nee_ph, nee_ph, nee_ph, beta.z);
                W.P.sh_ru[slot] = make_float4(ru.x, ru.y, ru.z, nz);
                shadows++;
            }
            if (alive) {
                if (rr) beta = beta / (1 - fmaxf(0.f, 1 - survivalProb));
                beta = beta * pfw;
                V3 r_l = r_u / sPdf;
                PathFlags fl = PathFlags::unpack(flw);
                fl.specular = false; fl.anyNonSpecular = true; fl.lastVertexVolume = true;
                float4 b = W.P.d_med[slot], e = W.P.rl_L[slot], f = W.P.misc[slot], g2 = W.P.misc2[slot];
                W.P.o_dx[slot] = make_float4(p.x, p.y, p.z, wiNext.x);
                W.P.d_med[slot] = make_float4(wiNext.y, wiNext.z, b.z, __int_as_float(fl.pack()));
                W.P.beta_ru[slot] = make_float4(beta.x, beta.y, beta.z, r_u.x);
                W.P.ru_rl[slot] = make_float4(r_u.y, r_u.z, r_l.x, r_l.y);
                W.P.rl_L[slot] = make_float4(r_l.z, e.y, e.z, e.w);
                W.P.rng[slot] = make_ulonglong2(rng.state, rng.inc);
                W.P.misc[slot] = make_float4(f.x, vspNext, p.x, p.y);
                W.P.misc2[slot] = make_float4(p.z, g2.y, g2.z, g2.w);
                W.P.prev_ns[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        queue_push(W.shadow, w && emitShadow, slot);
        queue_push(W.next, w && alive, slot);
        queue_push(W.done, w && !alive, slot);
    }
    if (leader) { counter_add(W.counters, CNT_LOOKUPS, lookups); counter_add(W.counters, CNT_SHADOW, shadows); }
}

}  // namespace vspg
