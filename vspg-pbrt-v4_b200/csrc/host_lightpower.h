// PowerLightSampler set-up on the host (lightsamplers.cpp:76-96): each light's power is Phi(lambda) / lambda.PDF()
// averaged over the three channels, with lambda = SampledWavelengths::SampleVisible(0.5) -- in the RGB build the three
// "wavelengths" still carry the visible-wavelength pdf (util/spectrum.h:369-386), so a coloured light's channels are
// weighted unevenly -- and the pick goes through an AliasTable (util/sampling.cpp:563-645), not a CDF walk.
#pragma once
#include <cmath>
#include <vector>
#include "../../include/vspg.h"
#include "../../include/vspg_libm.h"

struct HostPowerSampler {
    std::vector<float> p, q;      // bins[i].p (the pmf), bins[i].q
    std::vector<int> alias;
};

// area[i] = Shape::Area() of light i's shape (area lights only)
inline HostPowerSampler build_power_sampler(const vspg_light *lights, int n, const float *area, float sceneRadius) {
    const float Pi = 3.14159265358979323846f;
    float pdf[3];
    for (int i = 0; i < 3; i++) {
        float up = 0.5f + float(i) / 3;
        if (up > 1) up -= 1;
        float lambda = 538 - 138.888889f * vspg_atanhf(0.85691062f - 1.82750197f * up);   // SampleVisibleWavelengths (util/sampling.h:169-171)
        float c = vspg_coshf(0.0072f * (lambda - 538));
        pdf[i] = (lambda < 360 || lambda > 830) ? 0.f : 0.0039398042f / (c * c);        // VisibleWavelengthsPDF (:163-167)
    }
    std::vector<float> power(n);
    for (int i = 0; i < n; i++) {
        const vspg_light &l = lights[i];
        float phi[3];
        for (int c = 0; c < 3; c++) {
            switch (l.type) {   // lights.cpp:196-198, 251-253, 826-842, 1038-1040 with the products in the reference's order
            case VSPG_LIGHT_DIFFUSE_AREA: phi[c] = Pi * (l.two_sided ? 2 : 1) * area[i] * (l.L[c] * l.scale); break;
            case VSPG_LIGHT_POINT: phi[c] = 4 * Pi * l.scale * l.L[c]; break;
            case VSPG_LIGHT_SPOT:    // lights.cpp:1452-1455
                phi[c] = l.scale * l.L[c] * 2 * Pi * ((1 - l.cos_falloff_start) + (l.cos_falloff_start - l.cos_falloff_end) / 2); break;
            case VSPG_LIGHT_DISTANT: phi[c] = l.scale * l.L[c] * Pi * (sceneRadius * sceneRadius); break;
            default: phi[c] = 4 * Pi * Pi * (sceneRadius * sceneRadius) * l.scale * l.L[c]; break;
            }
            phi[c] = pdf[c] != 0 ? phi[c] / pdf[c] : 0.f;   // SafeDiv
        }
        float s = phi[0]; s += phi[1]; s += phi[2];
        power[i] = s / 3;
    }
    float acc = 0;
    for (float v : power) acc += v;                        // std::accumulate(..., 0.f)
    if (acc == 0.f) for (float &v : power) v = 1.f;
    HostPowerSampler ps;
    ps.p.resize(n); ps.q.assign(n, 1.f); ps.alias.assign(n, -1);
    double dsum = 0;
    for (float v : power) dsum += v;                       // AliasTable: Float sum = std::accumulate(..., 0.): double accumulation, float result
    const float sum = (float)dsum;
    for (int i = 0; i < n; i++) ps.p[i] = power[i] / sum;
    struct Outcome { float pHat; int index; };
    std::vector<Outcome> under, over;
    for (int i = 0; i < n; i++) {
        float pHat = ps.p[i] * n;
        (pHat < 1 ? under : over).push_back(Outcome{pHat, i});
    }
    while (!under.empty() && !over.empty()) {
        Outcome un = under.back(), ov = over.back();
        under.pop_back(); over.pop_back();
        ps.q[un.index] = un.pHat; ps.alias[un.index] = ov.index;
        float pExcess = un.pHat + ov.pHat - 1;
        (pExcess < 1 ? under : over).push_back(Outcome{pExcess, ov.index});
    }
    while (!over.empty()) { ps.q[over.back().index] = 1; ps.alias[over.back().index] = -1; over.pop_back(); }
    while (!under.empty()) { ps.q[under.back().index] = 1; ps.alias[under.back().index] = -1; under.pop_back(); }
    return ps;
}
