// Body of sample_distance() / sample_distance_emit() / sample_distance_full(): included by dsampledistance.cuh with VSPG_SD_NAME,
// VSPG_SD_EMIT and VSPG_SD_RESAMPLING set.
// Two plain functions instead of one template: as a template <bool EMIT> the same source compiled to a k_medium with a
// 1392-byte stack frame instead of 344 bytes (C3 medium stage 62 -> 71 ms), also for EMIT = false.
VD void VSPG_SD_NAME(const DScene &sc, PathRegs &ps, float tHit, TrackCounters &tc, DistanceResult &res) {
    const vspg_integrator_params &ip = sc.ip;
    const DMedium &med = sc.media[ps.ray.medium];
    const int ch = ps.fl.channel;
    res.scattered = false; res.terminated = false;
    uint64_t h0 = hash_float(ps.rng.uniform());
    uint64_t h1 = hash_float(ps.rng.uniform());
    Rng rng; rng.set_sequence(h0, h1);

    // VSP target (:652-671, 1098-1134)
    bool guideScatterDecision = false;
    float vsp = -1.f;
    const int px = ps.pixel % sc.cam.width, py = ps.pixel / sc.cam.width;
    if (ps.fl.depth == 0) {
        if (ip.vsp_guiding && ip.vsp_primary_guiding) {
            vsp = sc.isg_vsp ? __ldg(&sc.isg_vsp[(size_t)py * sc.isg_w + px]) : 0.5f;
            guideScatterDecision = !(isnan(vsp) || vsp < 0.f || vsp > 1.f);
        }
    } else if (ip.vsp_guiding && ip.vsp_secondary_guiding) {
        vsp = ps.vsp_next;
        guideScatterDecision = !(isnan(vsp) || vsp < 0.f || vsp > 1.f);
    }
    if (guideScatterDecision) vsp = fmaxf(fminf(vsp, 0.999f), 0.001f);

    // media_sampleTMaj.h:55-56 (all three variants): tMax *= |d|, d = normalize(d)
    Ray ray = ps.ray;
    float tMax = tHit * length(ray.d);
    ray.d = normalize(ray.d);
    float uT = ps.rng.uniform();

    const bool use_resampling = VSPG_SD_RESAMPLING && !ip.vsp_sampling_nds && med.type != VSPG_MEDIUM_HOMOGENEOUS;
    if (use_resampling) {
        float volumeRatioCompensated = 0, majorantScale = 1;
        float weightSum = 0;
        V3 trRatioEst = splat(1.f), beta_rs = splat(1.f), r_u_rs = splat(1.f);
        V3 selP; MediumProps selMp; float selWi = 0, selSig = 0; V3 selNum, selDen;
        selMp.g = 0;
        V3 T_maj = track_resampling(
            med, ray, tMax, uT, rng, ch, guideScatterDecision, vsp, volumeRatioCompensated, majorantScale, tc,
            [&](V3 p, const MediumProps &mp, V3 sigma_maj, V3 T_maj) {
                V3 sigma_t = mp.sigma_s + mp.sigma_a;
                V3 sigma_n = clamp_zero(sigma_maj - sigma_t);
                float wi = (sigma_t / sigma_maj * trRatioEst)[ch];
                if (wi > 0) {
                    weightSum += wi;
                    if (ps.rng.uniform() < wi / weightSum) {
                        float pdf = T_maj[ch] * sigma_t[ch];
                        selP = p; selMp = mp; selWi = wi; selSig = wi;
                        selNum = beta_rs * T_maj * mp.sigma_s / pdf;
                        selDen = r_u_rs * T_maj * sigma_t / pdf;
                    }
                }
                float pdf = T_maj[ch] * sigma_n[ch];
                beta_rs = beta_rs * (T_maj * sigma_n / pdf);
                r_u_rs = r_u_rs * (T_maj * sigma_n / pdf);
                trRatioEst = trRatioEst * (sigma_n / sigma_maj);
                return true;
            });
        beta_rs = beta_rs * (T_maj / T_maj[ch]);
        r_u_rs = r_u_rs * (T_maj / T_maj[ch]);
        float trS = trRatioEst[ch];
        float surfWi = trS;
        if (guideScatterDecision && trS < 1 && trS > 0 && weightSum > 0) {
            float volRatio = volumeRatioCompensated * ip.vsp_mis_ratio + (1 - trS) * (1 - ip.vsp_mis_ratio);
            float surfRatio = 1 - volRatio;
            surfWi = surfRatio / volRatio * weightSum;
        }
        weightSum += surfWi;
        bool selectSurface = false;
        if (weightSum == 0) return;
        else if (ps.rng.uniform() < surfWi / weightSum) { selectSurface = true; selWi = surfWi; selSig = trS; selNum = beta_rs; selDen = r_u_rs; }
        float rf = weightSum * selSig / selWi;
        if (selectSurface) {
            ps.beta = ps.beta * (selNum * rf);
            ps.r_u = ps.r_u * selDen;
            if (has_nan(ps.beta) || has_nan(ps.r_u) || y_is_inf(ps.beta) || y_is_inf(ps.r_u)) { res.terminated = true; return; }
        } else {
            if (ps.fl.depth == 0) ps.fl.firstVolume = true;
            if (ps.fl.depth++ >= ip.max_depth) { res.terminated = true; return; }
            ps.beta = ps.beta * (selNum * rf);
            ps.r_u = ps.r_u * selDen;
            if (has_nan(ps.beta) || has_nan(ps.r_u) || y_is_inf(ps.beta) || y_is_inf(ps.r_u)) { res.terminated = true; return; }
            res.scattered = true; res.p = selP; res.mp = selMp;
        }
    } else {
        V3 beta_factor = splat(1.f), r_u_factor = splat(1.f);
        const bool trLoaded = sc.tr_rgb != nullptr;
        V3 T_maj = track_optical_depth(
            med, ray, tMax, uT, rng, ch, guideScatterDecision, vsp, ip.vsp_mis_ratio, ip.vsp_sampling_nds != 0, beta_factor, r_u_factor, tc,
            [&](V3 p, const MediumProps &mp, V3 sigma_maj, V3 T_maj, bool activateNDS) {
                if (!nonzero(ps.beta)) { res.terminated = true; return false; }
                if (VSPG_SD_EMIT && med.emissive && ps.fl.depth < ip.max_depth) {   // :895-906 emission at the tentative collision
                    V3 Le = medium_emission(med, p);
                    if (nonzero(Le)) {
                        float pdf = sigma_maj[ch] * T_maj[ch];
                        V3 betap = ps.beta * T_maj / pdf;
                        V3 r_e = ps.r_u * sigma_maj * T_maj / pdf;
                        if (nonzero(r_e)) ps.L = ps.L + betap * mp.sigma_a * Le / avg3(r_e);
                    }
                }
                V3 sigma_t = mp.sigma_s + mp.sigma_a;
                float pScatter = sigma_t[ch] / sigma_maj[ch];
                bool NDS_plus = false;
                if (ps.fl.depth == 0 && ip.vsp_sampling_nds && ip.collision_probability_bias && trLoaded && activateNDS) {
                    NDS_plus = true;
                    float trEstCache = __ldg(&sc.tr_rgb[3 * ((size_t)py * sc.tr_w + px) + ch]);
                    float gamma = 1 + trEstCache;
                    pScatter = vspg_powf(pScatter, 1 / gamma);
                }
                float pNull = fmaxf(0.f, 1 - pScatter);
                float um = rng.uniform();
                // SampleDiscrete({pScatter, pNull}, um)  util/sampling.h:79-112
                int mode;
                {
                    float sumW = pScatter + pNull;
                    float up = um * sumW;
                    if (up == sumW) up = next_down(up);
                    mode = (0.f + pScatter <= up) ? 1 : 0;
                }
                if (mode == 0) {
                    if (ps.fl.depth == 0) ps.fl.firstVolume = true;
                    if (ps.fl.depth++ >= ip.max_depth) { res.terminated = true; return false; }
                    float pdf = T_maj[ch] * sigma_t[ch];
                    ps.beta = ps.beta * (T_maj * mp.sigma_s / pdf);
                    ps.r_u = ps.r_u * (T_maj * sigma_t / pdf);
                    if (NDS_plus) ps.r_u = ps.r_u * (sigma_maj * pScatter / sigma_t);
                    ps.beta = ps.beta * beta_factor;
                    ps.r_u = ps.r_u * r_u_factor;
                    res.scattered = true; res.p = p; res.mp = mp;
                    return false;
                } else {
                    V3 sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                    float pdf = T_maj[ch] * sigma_n[ch];
                    ps.beta = ps.beta * (T_maj * sigma_n / pdf);
                    if (pdf == 0) ps.beta = V3();
                    ps.r_u = ps.r_u * (T_maj * sigma_n / pdf);
                    if (NDS_plus) ps.r_u = ps.r_u * (sigma_maj * (1 - pScatter) / sigma_n);
                    ps.r_l = ps.r_l * (T_maj * sigma_maj / pdf);
                    return nonzero(ps.beta) && nonzero(ps.r_u);
                }
            });
        if (!(res.scattered || res.terminated || !nonzero(ps.beta) || !nonzero(ps.r_u))) {
            ps.beta = ps.beta * (T_maj / T_maj[ch]);
            ps.r_u = ps.r_u * (T_maj / T_maj[ch]);
            ps.r_l = ps.r_l * (T_maj / T_maj[ch]);
            ps.beta = ps.beta * beta_factor;
            ps.r_u = ps.r_u * r_u_factor;
            ps.r_l = ps.r_l * r_u_factor;
        }
    }
}

