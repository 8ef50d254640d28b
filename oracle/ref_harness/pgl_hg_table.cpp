// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
// Reads the 128-row HG->vMF table (directional/vmm/VMMPhaseFunctions.cpp) back from the REFERENCE's own
// OpenPGL library (oracle/_ref/libopenpgl_ref.so).  Separate translation unit: the internal header
// clashes with the public cpp API used by pgl_harness.cpp.
#include "directional/vmm/VMMPhaseFunctions.h"

extern "C" {
// 128 rows x 18 floats: K, g, weights[4], meanCosines[4], kappas[4], normalizations[4]
// (directional/vmm/VMMPhaseFunctions.cpp table, read back from the library's static storage).
int ref_hg_table(float *out) {
    openpgl::VMMSingleLobeHenyeyGreensteinOracle::init();
    for (int i = 0; i < OPENPGL_VMM_NUM_PHASE_REP; i++) {
        const openpgl::VMMPhaseFunctionRepresentation &r = openpgl::VMMSingleLobeHenyeyGreensteinOracle::representations[i];
        float *o = out + 18 * i;
        o[0] = (float)r.K; o[1] = r.g;
        for (int j = 0; j < 4; j++) { o[2 + j] = r.weights[j]; o[6 + j] = r.meanCosines[j]; o[10 + j] = r.kappas[j]; o[14 + j] = r.normalizations[j]; }
    }
    return OPENPGL_VMM_NUM_PHASE_REP;
}

}
