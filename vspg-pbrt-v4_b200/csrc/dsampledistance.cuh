// GuidedVolPathVSPGIntegrator::SampleDistance (cpu/guidedvolpathvspgintegrator.cpp:637-1096) up to and including
// the chosen event; the scatter-vertex work (:804-875 / :988-1058) runs in the volume-scatter stage.
#pragma once
#include "dpath.cuh"
#include "dtrack.cuh"

namespace vspg {

struct DistanceResult {
    bool scattered;    // a real-scatter vertex was chosen at p
    bool terminated;
    V3 p; MediumProps mp;
};

// sample_distance(): media without emission; sample_distance_emit(): the same with the emission term of :895-906 compiled in.
// Inputs: path registers (ray, beta, r_u, r_l, depth, channel, vsp_next), tHit, pixel.  The two sampler draws that
// seed the tracking RNG (:323-325) are taken here.
#define VSPG_SD_NAME sample_distance
#define VSPG_SD_EMIT 0
#include "dsampledistance_body.inl"
#undef VSPG_SD_NAME
#undef VSPG_SD_EMIT
#define VSPG_SD_NAME sample_distance_emit
#define VSPG_SD_EMIT 1
#include "dsampledistance_body.inl"
#undef VSPG_SD_NAME
#undef VSPG_SD_EMIT

}  // namespace vspg
