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

struct RsLane {
    int slot, phase;
    Ray ray;                             // normalised ray (media_sampleTMaj.h:55-56)
    MajorantIterator it;
    float majorantScale, volumeRatioComp;
    bool guide;
    Segment seg; float tMin;
    V3 T_maj; float u; Rng rng, prng;    // tracking RNG, path sampler
    int count, ch, medium;
    // reservoir (vspg.cpp:690-731)
    float weightSum, selWi, selG;
    V3 trRatioEst, beta_rs, r_u_rs, selP, selNum, selDen;
};

VD void rs_fetch(const DScene &sc, const Wavefront &W, int slot, RsLane &s, unsigned &ddaSteps) {
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
    s.trRatioEst = s.beta_rs = s.r_u_rs = splat(1.f);
    s.selP = V3(); s.selNum = V3(); s.selDen = V3();
    s.T_maj = splat(1.f);
    s.count = 0;
    ddaSteps += (unsigned)__float_as_int(pre.y);
    // media_sampleTMaj.h:166-181 with the pre-pass sum from k_intersect
    float totalLength = pre.x;
    s.majorantScale = 1.0f;
    s.volumeRatioComp = vsp;
    if (totalLength == 0.f) { s.phase = RS_DONE; s.it.init_empty(); return; }   // T_maj = (1,1,1)
    if (s.guide) {
        float minTotalLength = -logf(1 - vsp);
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
VD void rs_finalize(const DScene &sc, const Wavefront &W, RsLane &s) {
    const vspg_integrator_params &ip = sc.ip;
    const int ch = s.ch, slot = s.slot;
    V3 T_maj = s.T_maj;
    s.beta_rs = s.beta_rs * (T_maj / T_maj[ch]);
    s.r_u_rs = s.r_u_rs * (T_maj / T_maj[ch]);
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
            if (fl.depth++ >= ip.max_depth) terminated = true;
            else { b.w = __int_as_float(fl.pack()); W.P.d_med[slot] = b; }
        }
        if (!terminated) {
            beta = beta * (s.selNum * rf);
            r_u = r_u * s.selDen;
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

constexpr int kRsRefillIdle = 8;    // fetch new rays once this many lanes of the warp are idle
constexpr int kRsTripsPerCheck = 4; // collisions between two refill checks

__global__ void __launch_bounds__(kBlock) k_medium_rs(const DScene sc, const Wavefront W, int *cursor) {
    const int n = *W.medium_rs.count;
    const int lane = threadIdx.x & 31;
    TrackCounters tc; tc.queries = 0; tc.dda = 0;
    RsLane s;
    s.phase = RS_DONE; s.it.init_empty();
    bool busy = false, drained = false;
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
                if (idx < n) { rs_fetch(sc, W, W.medium_rs.items[idx], s, tc.dda); busy = true; }
            }
            __syncwarp();
            idle = __ballot_sync(kFull, !busy);
        }
        if (idle == kFull) { if (drained) break; else continue; }
        if (busy) {
            const DMedium &med = sc.media[s.medium];
            const int ch = s.ch;
#pragma unroll 1
            for (int trip = 0; trip < kRsTripsPerCheck; trip++) {
                // ---- (1) light: advance to the next tentative collision inside the medium (or the end of the ray)
                float t = 0;
                bool hit = false;
#pragma unroll 1
                while (s.phase != RS_DONE && !hit) {
                    if (s.phase == RS_NEED_SEG) {
                        Segment seg;
                        if (!s.it.next(&seg)) { s.phase = RS_DONE; break; }     // T_maj = residual transmittance
                        seg.sigma_maj = seg.sigma_maj * s.majorantScale;
                        if (seg.sigma_maj[ch] == 0) {
                            float dt = seg_dt(seg.tMax, seg.tMin);
                            s.T_maj = s.T_maj * fast_exp3(-dt * seg.sigma_maj);
                            continue;
                        }
                        s.seg = seg; s.tMin = seg.tMin; s.phase = RS_IN_SEG;
                    }
                    s.count++;
                    t = s.tMin + sample_exponential(s.u, s.seg.sigma_maj[ch]);
                    s.u = s.rng.uniform();
                    if (t < s.seg.tMax) {
                        if (s.count > 10000) s.phase = RS_NEED_SEG;   // media_sampleTMaj.h:216-219 leaves the segment
                        else hit = true;
                    } else {
                        float dt = seg_dt(s.seg.tMax, s.tMin);
                        s.T_maj = s.T_maj * fast_exp3(-dt * s.seg.sigma_maj);
                        s.phase = RS_NEED_SEG;
                    }
                }
                // ---- (2) heavy, branch-free: density fetch + reservoir callback (vspg.cpp:690-731)
                if (hit) {
                    V3 T_maj = s.T_maj * fast_exp3(-(t - s.tMin) * s.seg.sigma_maj);
                    V3 p = s.ray.at(t);
                    MediumProps mp = sample_point(med, p);
                    tc.queries++;
                    V3 sigma_maj = s.seg.sigma_maj;
                    V3 sigma_t = mp.sigma_s + mp.sigma_a;
                    V3 sigma_n = clamp_zero(sigma_maj - sigma_t);
                    float wi = (sigma_t / sigma_maj * s.trRatioEst)[ch];
                    const bool cand = wi > 0;
                    Rng pr = s.prng;                        // the reservoir draw happens only for wi > 0 (:699-702)
                    float ur = pr.uniform();
                    float ws = s.weightSum + wi;
                    if (cand) { s.prng = pr; s.weightSum = ws; }
                    const bool take = cand && (ur < wi / ws);
                    float pdfT = T_maj[ch] * sigma_t[ch];
                    V3 num = s.beta_rs * T_maj * mp.sigma_s / pdfT;
                    V3 den = s.r_u_rs * T_maj * sigma_t / pdfT;
                    if (take) { s.selP = p; s.selG = mp.g; s.selWi = wi; s.selNum = num; s.selDen = den; }
                    float pdf = T_maj[ch] * sigma_n[ch];
                    V3 f = T_maj * sigma_n / pdf;
                    s.beta_rs = s.beta_rs * f;
                    s.r_u_rs = s.r_u_rs * f;
                    s.trRatioEst = s.trRatioEst * (sigma_n / sigma_maj);
                    s.T_maj = splat(1.f);
                    s.tMin = t;
                }
                if (s.phase == RS_DONE) break;
            }
            if (s.phase == RS_DONE) {
                tc.dda += s.it.steps;
                rs_finalize(sc, W, s);
                busy = false;
            }
        }
    }
    counter_add(W.counters, CNT_QUERIES, tc.queries);
    counter_add(W.counters, CNT_DDA, tc.dda);
}

}  // namespace vspg
