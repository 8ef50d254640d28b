// The three majorant-space distance samplers of the fork (media_sampleTMaj.h), one thread per ray.
// Callbacks are device lambdas: cb(p, mp, sigma_maj, T_maj[, activateNDS]) -> bool (false = stop).
#pragma once
#include "dmedia.cuh"

namespace vspg {

// media_sampleTMaj.h:49-117  SampleT_maj (delta / ratio tracking driver).  `ray.d` is already normalised and
// tMax already scaled by |d| (the callers do lines :55-56).
// GRAY: sigma_maj / T_maj carried as one float (dmath.cuh G1) -- valid when the medium's sigma_a, sigma_s are equal across RGB.
template <bool GRAY, typename F>
VD typename SpecT<GRAY>::type track_T_maj_t(const DMedium &med, const Ray &ray, float tMax, float u, Rng &rng, int ch, TrackCounters &tc, F cb) {
    typedef typename SpecT<GRAY>::type Sp;
    MajorantIterator iter = sample_ray(med, ray, tMax);
    Sp T_maj = to_spec<GRAY>(splat(1.f));
    bool done = false;
    Segment seg;
    while (!done) {
        if (!iter.next(&seg)) { tc.dda += iter.steps; return T_maj; }
        const Sp sm = to_spec<GRAY>(seg.sigma_maj);
        float smc = sm[ch];
        if (smc == 0) {
            float dt = seg_dt(seg.tMax, seg.tMin);
            T_maj = T_maj * fast_exp3(-dt * sm);
            continue;
        }
        float tMin = seg.tMin;
        while (true) {
            float t = tMin + sample_exponential(u, smc);
            u = rng.uniform();
            if (t < seg.tMax) {
                T_maj = T_maj * fast_exp3(-(t - tMin) * sm);
                V3 p = ray.at(t);
                MediumProps mp = sample_point(med, p);
                tc.queries++;
                if (!cb(p, mp, sm, T_maj)) { done = true; break; }
                T_maj = to_spec<GRAY>(splat(1.f));
                tMin = t;
            } else {
                float dt = seg_dt(seg.tMax, tMin);
                T_maj = T_maj * fast_exp3(-dt * sm);
                break;
            }
        }
    }
    tc.dda += iter.steps;
    return to_spec<GRAY>(splat(1.f));
}
template <typename F>
VD V3 track_T_maj(const DMedium &med, const Ray &ray, float tMax, float u, Rng &rng, int ch, TrackCounters &tc, F cb) {
    return track_T_maj_t<false>(med, ray, tMax, u, rng, ch, tc, cb);
}

// media_sampleTMaj.h:136-248  SampleT_maj_Resampling
template <typename F>
VD V3 track_resampling(const DMedium &med, const Ray &ray, float tMax, float u, Rng &rng, int ch, bool guideScatterDecision,
                       float volScatterProb, float &volumeRatioComp, float &majorantScale, TrackCounters &tc, F cb) {
    MajorantIterator iter = sample_ray(med, ray, tMax);
    float totalLength = 0.f;
    Segment seg;
    {
        MajorantIterator pre = iter;  // pre-pass: total majorant optical depth along the segment
        while (pre.next(&seg)) {
            if (seg.sigma_maj[ch] == 0) continue;
            totalLength += seg.sigma_maj[ch] * (seg.tMax - seg.tMin);
        }
        tc.dda += pre.steps;
    }
    if (totalLength == 0.f) return splat(1.f);
    majorantScale = 1.0f;
    volumeRatioComp = volScatterProb;
    if (guideScatterDecision) {
        float minTotalLength = -vspg_logf(1 - volScatterProb);
        if (minTotalLength > totalLength) {
            majorantScale = minTotalLength / totalLength;
            totalLength = minTotalLength;
        }
        float expNegTotalLength = fast_exp(-totalLength);
        volumeRatioComp = volScatterProb / (1 - expNegTotalLength);
    }
    V3 T_maj = splat(1.f);
    bool done = false;
    int count = 0;
    while (!done) {
        if (!iter.next(&seg)) { tc.dda += iter.steps; return T_maj; }
        seg.sigma_maj = seg.sigma_maj * majorantScale;
        float smc = seg.sigma_maj[ch];
        if (smc == 0) {
            float dt = seg_dt(seg.tMax, seg.tMin);
            T_maj = T_maj * fast_exp3(-dt * seg.sigma_maj);
            continue;
        }
        float tMin = seg.tMin;
        while (true) {
            count++;
            float t = tMin + sample_exponential(u, smc);
            u = rng.uniform();
            if (t < seg.tMax) {
                if (count > 10000) break;
                T_maj = T_maj * fast_exp3(-(t - tMin) * seg.sigma_maj);
                V3 p = ray.at(t);
                MediumProps mp = sample_point(med, p);
                tc.queries++;
                if (!cb(p, mp, seg.sigma_maj, T_maj)) { done = true; break; }
                T_maj = splat(1.f);
                tMin = t;
            } else {
                float dt = seg_dt(seg.tMax, tMin);
                T_maj = T_maj * fast_exp3(-dt * seg.sigma_maj);
                break;
            }
        }
    }
    tc.dda += iter.steps;
    return splat(1.f);
}

// media_sampleTMaj.h:269-491  SampleT_maj_OpticalDepthSpace (analytic VSP / NDS / NDS+, MIS with delta tracking)
template <typename F>
VD V3 track_optical_depth(const DMedium &med, const Ray &ray, float tMax, float u, Rng &rng, int ch, bool guideScatterDecision,
                          float vsp, float vspMISRatio, bool NDS, V3 &beta_factor, V3 &r_u_factor, TrackCounters &tc, F cb) {
    auto plain = [&](V3 p, const MediumProps &mp, V3 sm, V3 tm) { return cb(p, mp, sm, tm, false); };
    if (!guideScatterDecision || vspMISRatio == 0.f) return track_T_maj(med, ray, tMax, u, rng, ch, tc, plain);
    // The two later fall-backs (:305-307 infinite segment, :330-332 NDS with vsp below the segment's scatter probability)
    // call SampleT_maj with the ALREADY normalised ray and scaled tMax, and SampleT_maj normalises again (:55-56):
    // |normalize(d)| is not exactly 1, so tMax and d move by an ulp -- and so do the collision points.
    auto plain_renormalised = [&]() {
        Ray r2 = ray;
        float tMax2 = tMax * length(r2.d);
        r2.d = normalize(r2.d);
        return track_T_maj(med, r2, tMax2, u, rng, ch, tc, plain);
    };
    MajorantIterator iter = sample_ray(med, ray, tMax);
    float t_v = 0.f;
    Segment seg;
    {
        MajorantIterator pre = iter;
        bool infinite = false;
        while (pre.next(&seg)) {
            if (isinf(seg.tMax)) { infinite = true; break; }
            if (seg.sigma_maj[ch] == 0) continue;
            t_v += seg.sigma_maj[ch] * (seg.tMax - seg.tMin);
        }
        tc.dda += pre.steps;
        if (infinite) return plain_renormalised();
    }
    if (t_v == 0.f) return splat(1.f);
    float OneMinusENegTv = 1.f - fast_exp(-t_v);
    float t_n = -1.f, t_n_current = -1.f;
    if (NDS) {
        if (vsp < 1 - fast_exp(-t_v)) return plain_renormalised();
        t_n = (float)(-vspg_log(1.0 - (double)(OneMinusENegTv / vsp)));  // the reference's 1.0 literal promotes to double
        t_n_current = t_n;
    }
    V3 T_maj = splat(1.f), tpScaleFactor = splat(1.f);
    float t_v_current = t_v;
    float remainingDist = 0;
    bool deltaTracking = false;
    if (u > vspMISRatio) { deltaTracking = true; u = (u - vspMISRatio) / (1 - vspMISRatio); }
    else u /= vspMISRatio;
    bool done = false;
    int count = 0;
    bool overTheEnd = false;
    const float ScatterEpsilon = 1e-5f;
    while (!done) {
        if (!iter.next(&seg)) { tc.dda += iter.steps; return T_maj; }
        float smc = seg.sigma_maj[ch];
        if (smc == 0 || overTheEnd) {
            float dt = seg_dt(seg.tMax, seg.tMin);
            T_maj = T_maj * fast_exp3(-dt * seg.sigma_maj);
            continue;
        }
        float tMin = seg.tMin;
        V3 normalizedMaj = seg.sigma_maj / smc;
        if (remainingDist > 0) {
            tMin += remainingDist / smc;
            if (tMin > seg.tMax + ScatterEpsilon) {
                float dist = (seg.tMax - seg.tMin) * smc;
                t_v_current -= dist; t_n_current -= dist; remainingDist -= dist;
                T_maj = T_maj * fast_exp3(-(seg.tMax - seg.tMin) * seg.sigma_maj);
                continue;
            }
            t_v_current -= remainingDist; t_n_current -= remainingDist; remainingDist = 0;
            T_maj = T_maj * fast_exp3(-(tMin - seg.tMin) * seg.sigma_maj);
            V3 p = ray.at(tMin);
            MediumProps mp = sample_point(med, p);
            tc.queries++;
            r_u_factor = splat(vspMISRatio) / tpScaleFactor + splat(1 - vspMISRatio);
            if (!cb(p, mp, seg.sigma_maj, T_maj, true)) break;
            T_maj = splat(1.f);
        }
        while (true) {
            count++;
            float dist = kFltMax;
            V3 tpSingle;
            if (NDS) {
                tpSingle = splat(1.0f) - fast_exp3(-t_n_current * normalizedMaj);
                if (!deltaTracking) dist = (float)(-vspg_log(1.0 - (double)(u * tpSingle[ch])));
            } else {
                tpSingle = (splat(1.0f) - fast_exp3(-t_v_current * normalizedMaj)) / vsp;
                if (!deltaTracking) {
                    if (u < vsp) dist = (float)(-vspg_log(1.0 - (double)(u * tpSingle[ch])));
                }
            }
            if (deltaTracking) dist = (float)(-vspg_log(1.0 - (double)u));
            bool passThrough = t_v_current - dist < ScatterEpsilon || dist == 0;
            if (NDS || !passThrough) tpScaleFactor = tpScaleFactor * tpSingle;
            if (passThrough) {
                if (NDS) tpScaleFactor = tpScaleFactor / (1.0f - fast_exp(-t_n + t_v));
                else tpScaleFactor = tpScaleFactor * fast_exp3(-t_v_current * normalizedMaj) / (1 - vsp);
                r_u_factor = splat(vspMISRatio) / tpScaleFactor + splat(1 - vspMISRatio);
                overTheEnd = true;
                T_maj = T_maj * fast_exp3(-(seg.tMax - tMin) * seg.sigma_maj);
                break;
            }
            float t = tMin + dist / smc;
            u = rng.uniform();
            if (t <= seg.tMax + ScatterEpsilon) {
                if (count > 10000) break;
                t_v_current -= dist; t_n_current -= dist; remainingDist = 0;
                T_maj = T_maj * fast_exp3(-(t - tMin) * seg.sigma_maj);
                V3 p = ray.at(t);
                MediumProps mp = sample_point(med, p);
                tc.queries++;
                r_u_factor = splat(vspMISRatio) / tpScaleFactor + splat(1 - vspMISRatio);
                if (!cb(p, mp, seg.sigma_maj, T_maj, true)) { done = true; break; }
                T_maj = splat(1.f);
                tMin = t;
            } else {
                float dt = seg_dt(seg.tMax, tMin);
                T_maj = T_maj * fast_exp3(-dt * seg.sigma_maj);
                float distWithinThisSeg = dt * smc;
                remainingDist = dist - distWithinThisSeg;
                t_v_current -= distWithinThisSeg; t_n_current -= distWithinThisSeg;
                break;
            }
        }
    }
    tc.dda += iter.steps;
    return splat(1.f);
}

}  // namespace vspg
