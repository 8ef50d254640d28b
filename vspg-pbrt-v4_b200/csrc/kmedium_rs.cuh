// SampleDistance, resampling routine (vspg.cpp:682-877 + media_sampleTMaj.h:136-248) as a persistent-lane kernel.
//
// One ray per LANE, but lanes do not wait for each other: a lane whose walk has ended finalises its path and a
// fresh ray is fetched from the stage queue as soon as enough lanes of the warp are idle.  Every trip of the main
// loop is (1) a light, divergent search for the lane's next tentative collision (DDA voxel steps + exponential
// steps) and (2) the heavy part -- trilinear density fetch + reservoir update -- executed by all busy lanes
// together and branch-free, so the 100x spread in walk length between rays (SURVEY 7 "divergence") costs idle
// lanes only at the very end of the queue.  The pre-pass (total majorant optical depth) was done by k_intersect.
// The arithmetic per ray is exactly the sequence of track_resampling() + the resampling branch of sample_distance().
#pragma once
#include "dsampledistance.cuh"

namespace vspg {

enum { RS_NEED_SEG = 1, RS_IN_SEG = 2, RS_DONE = 3 };

template <bool GRAY>
struct RsLane {
    typedef typename SpecT<GRAY>::type Sp;
    int slot, phase;
    Ray ray;                             // normalised ray (media_sampleTMaj.h:55-56)
    MajorantIterator it;
    float majorantScale, volumeRatioComp;
    bool guide;
    Sp segSigma; float segTMax, tMin;    // current majorant segment (sigma_maj already times majorantScale)
    Sp T_maj; float u; Rng rng, prng;    // tracking RNG, path sampler
    int count, ch, medium;
    // reservoir (vspg.cpp:690-731)
    float weightSum, selWi, selG;
    Sp trRatioEst, beta_rs, r_u_rs, selNum, selDen;
    V3 selP;
};

template <bool GRAY>
VD void rs_fetch(const DScene &sc, const Wavefront &W, int slot, RsLane<GRAY> &s, unsigned &ddaSteps) {
    typedef typename SpecT<GRAY>::type Sp;
    const vspg_integrator_params &ip = sc.ip;
    float4 a = W.P.o_dx[slot], b = W.P.d_med[slot];
    ulonglong2 q = W.P.rng[slot];
    float2 pre = W.P.rs_pre[slot];
    s.slot = slot;
    s.ray.o = V3(a.x, a.y, a.z); s.ray.d = V3(a.w, b.x, b.y); s.ray.medium = __float_as_int(b.z);
    s.medium = s.ray.medium;
    PathFlags fl = PathFlags::unpack(__float_as_int(b.w));
    s.ch = fl.channel;
    s.prng.state = q.x; s.prng.inc = q.y;
    uint64_t h0 = hash_float(s.prng.uniform());
    uint64_t h1 = hash_float(s.prng.uniform());
    s.rng.set_sequence(h0, h1);
    s.guide = false;
    float vsp = -1.f;
    if (fl.depth == 0) {
        if (ip.vsp_guiding && ip.vsp_primary_guiding) {
            int pixel = __float_as_int(W.P.misc2[slot].y);
            const int px = pixel % sc.cam.width, py = pixel / sc.cam.width;
            vsp = sc.isg_vsp ? __ldg(&sc.isg_vsp[(size_t)py * sc.isg_w + px]) : 0.5f;
            s.guide = !(isnan(vsp) || vsp < 0.f || vsp > 1.f);
        }
    } else if (ip.vsp_guiding && ip.vsp_secondary_guiding) {
        vsp = W.P.misc[slot].y;
        s.guide = !(isnan(vsp) || vsp < 0.f || vsp > 1.f);
    }
    if (s.guide) vsp = fmaxf(fminf(vsp, 0.999f), 0.001f);
    float tHit = W.P.hit[slot].x;
    float tMax = tHit * length(s.ray.d);
    s.ray.d = normalize(s.ray.d);
    s.u = s.prng.uniform();
    s.weightSum = 0; s.selWi = 0; s.selG = 0;
    s.trRatioEst = s.beta_rs = s.r_u_rs = to_spec<GRAY>(splat(1.f));
    s.selP = V3(); s.selNum = Sp(); s.selDen = Sp();
    s.T_maj = to_spec<GRAY>(splat(1.f));
    s.count = 0;
    // media_sampleTMaj.h:166-181 with the pre-pass sum from k_intersect
    float totalLength = pre.x;
    s.majorantScale = 1.0f;
    s.volumeRatioComp = vsp;
    if (totalLength == 0.f) { s.phase = RS_DONE; s.it.init_empty(); return; }   // T_maj = (1,1,1)
    if (s.guide) {
        float minTotalLength = -vspg_logf(1 - vsp);
        if (minTotalLength > totalLength) {
            s.majorantScale = minTotalLength / totalLength;
            totalLength = minTotalLength;
        }
        float expNegTotalLength = fast_exp(-totalLength);
        s.volumeRatioComp = vsp / (1 - expNegTotalLength);
    }
    s.it = sample_ray(sc.media[s.medium], s.ray, tMax);
    s.phase = RS_NEED_SEG;
}

// vspg.cpp:733-800: surface candidate, final pick, beta / r_u update; then hand-over to the next stage
template <bool GRAY>
VD void rs_finalize(const DScene &sc, const Wavefront &W, RsLane<GRAY> &s) {
    typedef typename SpecT<GRAY>::type Sp;
    const vspg_integrator_params &ip = sc.ip;
    const int ch = s.ch, slot = s.slot;
    Sp T_maj = s.T_maj;
    s.beta_rs = s.beta_rs * (T_maj / T_maj[ch]);
    s.r_u_rs = s.r_u_rs * (T_maj / T_maj[ch]);
    if (sc.tr_acc) {   // vspg.cpp:727-728: depth == 0 && calculateTrBuffer -> TrBuffer::AddSample(pPixel, trRatioEst RGB)
        PathFlags fl0 = PathFlags::unpack(__float_as_int(W.P.d_med[slot].w));
        if (fl0.depth == 0) {
            V3 tr = to_v3(s.trRatioEst);
            double *acc = sc.tr_acc + 4 * (size_t)__float_as_int(W.P.misc2[slot].y);
            atomicAdd(acc, (double)tr.x); atomicAdd(acc + 1, (double)tr.y); atomicAdd(acc + 2, (double)tr.z); atomicAdd(acc + 3, 1.0);
        }
    }
    float trS = s.trRatioEst[ch];
    float surfWi = trS;
    if (s.guide && trS < 1 && trS > 0 && s.weightSum > 0) {
        float volRatio = s.volumeRatioComp * ip.vsp_mis_ratio + (1 - trS) * (1 - ip.vsp_mis_ratio);
        float surfRatio = 1 - volRatio;
        surfWi = surfRatio / volRatio * s.weightSum;
    }
    s.weightSum += surfWi;
    float4 c = W.P.beta_ru[slot], d = W.P.ru_rl[slot];
    V3 beta(c.x, c.y, c.z), r_u(c.w, d.x, d.y);
    bool terminated = false, scattered = false;
    if (s.weightSum != 0) {
        bool selectSurface = false;
        float selSig = s.selWi;
        if (s.prng.uniform() < surfWi / s.weightSum) { selectSurface = true; s.selWi = surfWi; selSig = trS; s.selNum = s.beta_rs; s.selDen = s.r_u_rs; }
        float rf = s.weightSum * selSig / s.selWi;
        if (!selectSurface) {
            float4 b = W.P.d_med[slot];
            PathFlags fl = PathFlags::unpack(__float_as_int(b.w));
            if (fl.depth == 0) fl.firstVolume = true;
            if (fl.depth++ >= ip.max_depth) terminated = true;
            else { b.w = __int_as_float(fl.pack()); W.P.d_med[slot] = b; }
        }
        if (!terminated) {
            beta = beta * to_v3(s.selNum * rf);
            r_u = r_u * to_v3(s.selDen);
            if (has_nan(beta) || has_nan(r_u) || y_is_inf(beta) || y_is_inf(r_u)) terminated = true;
            else scattered = !selectSurface;
        }
    }
    W.P.beta_ru[slot] = make_float4(beta.x, beta.y, beta.z, r_u.x);
    W.P.ru_rl[slot] = make_float4(r_u.y, r_u.z, d.z, d.w);
    W.P.rng[slot] = make_ulonglong2(s.prng.state, s.prng.inc);
    bool dead = terminated || !nonzero(beta) || !nonzero(r_u);
    bool scat = !dead && scattered;
    if (scat) W.P.ev[slot] = make_float4(s.selP.x, s.selP.y, s.selP.z, s.selG);
    queue_push(W.done, dead, slot);
    queue_push(W.volscatter, scat, slot);
    queue_push(W.surface, !dead && !scat, slot);
}

#ifndef VSPG_RS_MINB
#define VSPG_RS_MINB 5                    // resident CTAs/SM of the gray resampling kernel's default build (A/B knob)
#endif
constexpr int kRsRefillIdleDefault = 16;    // fetch new rays once this many lanes of the warp are idle (kernel argument: VSPG_RS_REFILL for A/B runs)
#ifndef VSPG_RS_TRIPS
#define VSPG_RS_TRIPS 4
#endif
#ifndef VSPG_RS_SEARCH
#define VSPG_RS_SEARCH 4
#endif
constexpr int kRsTripsPerCheck = VSPG_RS_TRIPS; // collisions between two refill checks
constexpr int kRsSearchMax = VSPG_RS_SEARCH;     // search steps before the collision body runs for the lanes that are ready

// One search step of media_sampleTMaj.h:183-246 for one lane, straight-line: (A) if the lane needs a majorant segment,
// pull the next one from the DDA; (B) if it is inside a segment, draw the next exponential step.  Returns with `hit`
// set when a tentative collision at `t` lies inside the segment.
template <bool GRAY>
VD void rs_search_step(RsLane<GRAY> &s, float &t, bool &hit) {
    const int ch = s.ch;
    if (s.phase == RS_NEED_SEG) {
        Segment seg;
        if (!s.it.next(&seg)) s.phase = RS_DONE;                        // T_maj = residual transmittance
        else {
            typename SpecT<GRAY>::type sm = to_spec<GRAY>(seg.sigma_maj) * s.majorantScale;
            if (sm[ch] == 0) {
                float dt = seg_dt(seg.tMax, seg.tMin);
                s.T_maj = s.T_maj * fast_exp3(-dt * sm);
            } else { s.segSigma = sm; s.segTMax = seg.tMax; s.tMin = seg.tMin; s.phase = RS_IN_SEG; }
        }
    }
    if (s.phase == RS_IN_SEG) {
        s.count++;
        t = s.tMin + sample_exponential(s.u, s.segSigma[ch]);
        s.u = s.rng.uniform();
        if (t < s.segTMax) {
            if (s.count > 10000) s.phase = RS_NEED_SEG;                 // media_sampleTMaj.h:216-219 leaves the segment
            else hit = true;
        } else {
            float dt = seg_dt(s.segTMax, s.tMin);
            s.T_maj = s.T_maj * fast_exp3(-dt * s.segSigma);
            s.phase = RS_NEED_SEG;
        }
    }
}

// The whole warp runs every loop below (idle lanes carry phase == RS_DONE and do nothing), so the votes are full-warp.
// Per trip: up to kRsSearchMax search steps, leaving early as soon as at least half of the lanes still walking have a
// collision ready; then the collision body, predicated on `hit`.  Bounding the search keeps both halves of the trip
// above ~60 % lane utilisation (an unbounded search ran at 7 lanes of 32, profiles/r01c).
// NVDB: some medium of the scene is a NanoVDB grid (its sampler is compiled in only then)
template <bool GRAY, int MINB, bool NVDB = false>
__global__ void __launch_bounds__(kBlock, MINB) k_medium_rs(const DScene sc, const Wavefront W, int *cursor, const int kRsRefillIdle) {
    typedef typename SpecT<GRAY>::type Sp;
    const int n = *W.medium_rs.count;
    const int lane = threadIdx.x & 31;
    TrackCounters tc; tc.queries = 0; tc.dda = 0;
    RsLane<GRAY> s;
    s.phase = RS_DONE; s.it.init_empty(); s.medium = 0; s.ch = 0;
    bool busy = false, drained = false;
    // pending collision of the software pipeline (GRAY only)
    bool pend = false; Sp pT_maj, pSigma; V3 pP; DensityFetch pf;
    pf.a = pf.b = make_float4(0, 0, 0, 0); pf.dx = pf.dy = pf.dz = 0;
    while (true) {
        __syncwarp();
        unsigned idle = __ballot_sync(kFull, !busy);
        if (!drained && __popc(idle) >= kRsRefillIdle) {
            int nIdle = __popc(idle), base = 0;
            if (lane == 0) base = atomicAdd(cursor, nIdle);
            base = __shfl_sync(kFull, base, 0);
            if (base + nIdle >= n) drained = true;
            if (!busy) {
                int idx = base + __popc(idle & ((1u << lane) - 1));
                if (idx < n) { rs_fetch<GRAY>(sc, W, W.medium_rs.items[idx], s, tc.dda); busy = true; }
            }
            __syncwarp();
            idle = __ballot_sync(kFull, !busy);
        }
        if (idle == kFull) { if (drained) break; else continue; }
        const int ch = s.ch;
        bool flush = false;
#pragma unroll 1
        for (int trip = 0; trip < kRsTripsPerCheck; trip++) {
            // ---- (1) advance towards the next tentative collision inside the medium (or the end of the ray)
            float t = 0;
            bool hit = false;
#pragma unroll 1
            for (int j = 0; j < kRsSearchMax; j++) {
                if (s.phase != RS_DONE && !hit) rs_search_step<GRAY>(s, t, hit);
                unsigned hits = __ballot_sync(kFull, hit);
                unsigned walking = __ballot_sync(kFull, s.phase != RS_DONE && !hit);
                if (2 * __popc(hits) >= __popc(hits | walking)) break;
            }
            // ---- (2) reservoir callback (vspg.cpp:690-731).  GRAY (uniformgrid media): software-pipelined -- the
            // density cell of collision k was requested at the end of the previous trip and is consumed here, after the
            // search for collision k+1 has run, so the DRAM/L2 latency of the gather overlaps the search (the kernel was
            // long_scoreboard-bound, profiles/r01j).  The search never reads the density: it only needs tMin / T_maj,
            // which are committed when the fetch is issued.  Order of the path sampler's draws is unchanged (the
            // search uses the tracking RNG only).
            if (GRAY ? pend : hit) {
                Sp T_maj, sigma_maj; V3 p; MediumProps mp;
                if (GRAY) {
                    const DMedium &med = sc.media[s.medium];
                    T_maj = pT_maj; sigma_maj = pSigma; p = pP;
                    float d = medium_density_value<NVDB>(med, pf);
                    mp.g = med.g; mp.sigma_a = ld3(med.sigma_a) * d; mp.sigma_s = ld3(med.sigma_s) * d;   // sample_point, GRID branch
                    pend = false;
                } else {
                    T_maj = s.T_maj * fast_exp3(-(t - s.tMin) * s.segSigma);
                    p = s.ray.at(t);
                    mp = sample_point(sc.media[s.medium], p);
                    sigma_maj = s.segSigma;
                    s.T_maj = to_spec<GRAY>(splat(1.f));
                    s.tMin = t;
                }
                tc.queries++;
                Sp sigma_s = to_spec<GRAY>(mp.sigma_s);
                Sp sigma_t = sigma_s + to_spec<GRAY>(mp.sigma_a);
                Sp sigma_n = clamp_zero(sigma_maj - sigma_t);
                float wi = (sigma_t / sigma_maj * s.trRatioEst)[ch];
                const bool cand = wi > 0;
                Rng pr = s.prng;                        // the reservoir draw happens only for wi > 0 (:699-702)
                float ur = pr.uniform();
                float ws = s.weightSum + wi;
                if (cand) { s.prng = pr; s.weightSum = ws; }
                const bool take = cand && (ur < wi / ws);
                float pdfT = T_maj[ch] * sigma_t[ch];
                Sp num = s.beta_rs * T_maj * sigma_s / pdfT;
                Sp den = s.r_u_rs * T_maj * sigma_t / pdfT;
                if (take) { s.selP = p; s.selG = mp.g; s.selWi = wi; s.selNum = num; s.selDen = den; }
                float pdf = T_maj[ch] * sigma_n[ch];
                Sp f = T_maj * sigma_n / pdf;
                s.beta_rs = s.beta_rs * f;
                s.r_u_rs = s.r_u_rs * f;
                s.trRatioEst = s.trRatioEst * (sigma_n / sigma_maj);
            }
            // ---- (3) GRAY: request the density cell of the collision just found and commit the walk state
            if (GRAY && hit) {
                const DMedium &med = sc.media[s.medium];
                pT_maj = s.T_maj * fast_exp3(-(t - s.tMin) * s.segSigma);
                pP = s.ray.at(t);
                pSigma = s.segSigma;
                pf = medium_density_issue<NVDB>(med, pP);
                s.T_maj = to_spec<GRAY>(splat(1.f));
                s.tMin = t;
                pend = true;
            }
            // leave the trip loop once enough lanes have finished their walk (or are idle) to finalise / refill as a batch:
            // finalising one lane at a time cost 16 % of the kernel at 1.6 active lanes (profiles/r01j source page)
            // While the queue still has rays: flush when finished + idle lanes make up a refill batch.  Once it is
            // drained nothing refills, so wait for a full batch of finished lanes or for the last walker (r01u: the
            // unconditional flush after kRsTripsPerCheck trips still finalised 2 lanes at a time, 12 % of the samples).
            const unsigned doneMask = __ballot_sync(kFull, busy && s.phase == RS_DONE && !pend);
            const unsigned busyMask = __ballot_sync(kFull, busy);
            flush = drained ? (__popc(doneMask) >= kRsRefillIdle || doneMask == busyMask)
                            : (__popc(doneMask) + 32 - __popc(busyMask) >= kRsRefillIdle);
            if (flush) break;
        }
        if (flush && busy && s.phase == RS_DONE && !pend) {
            tc.dda += s.it.steps;
            rs_finalize<GRAY>(sc, W, s);
            busy = false;
        }
    }
    counter_add(W.counters, CNT_QUERIES, tc.queries);
    counter_add(W.counters, CNT_DDA, tc.dda);
}

}  // namespace vspg
