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
// sample_distance / sample_distance_emit are what k_medium / k_medium_emit run: k_intersect routes every ray that takes the
// resampling branch (:679) to k_medium_rs, so these two leave that branch out (VSPG_SD_RESAMPLING 0);
// sample_distance_full is the whole function, for the per-kernel parity hook vspg_debug_sample_distance.
#ifndef VSPG_SD_KEEP_RS
#define VSPG_SD_KEEP_RS 0   // A/B switch: 1 compiles the (never taken) resampling branch into k_medium as well
#endif
#define VSPG_SD_NAME sample_distance
#define VSPG_SD_EMIT 0
#define VSPG_SD_RESAMPLING VSPG_SD_KEEP_RS
#include "dsampledistance_body.inl"
#undef VSPG_SD_NAME
#undef VSPG_SD_EMIT
#define VSPG_SD_NAME sample_distance_emit
#define VSPG_SD_EMIT 1
#include "dsampledistance_body.inl"
#undef VSPG_SD_NAME
#undef VSPG_SD_RESAMPLING
#define VSPG_SD_NAME sample_distance_full
#define VSPG_SD_RESAMPLING 1
#include "dsampledistance_body.inl"
#undef VSPG_SD_NAME
#undef VSPG_SD_EMIT
#undef VSPG_SD_RESAMPLING

}  // namespace vspg
