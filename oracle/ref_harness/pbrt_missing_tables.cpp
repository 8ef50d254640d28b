// Test-infrastructure stub (NOT product code), linked into oracle/_ref/libpbrt_ref.so by build_ref_pbrt.sh.
// Stand-ins for what the trimmed build of the reference lacks:
//  * util/bluenoise.cpp and util/pmj02tables.cpp are missing from the reference checkout (.MISSING_LARGE_BLOBS).
//    Their tables are read by PNG dithering (util/image.cpp) and the "pmj02bn" sampler only; neither is on the
//    guidedvolpath / independent-sampler path.  Zero tables keep the link closed.
//  * RenderWavefront (wavefront/ and gpu/ are dropped: the CUDA/OptiX path) aborts if ever reached.
#include <pbrt/pbrt.h>
#include <pbrt/scene.h>
#include <pbrt/util/bluenoise.h>
#include <pbrt/util/pmj02tables.h>
#include <cstdio>
#include <cstdlib>

namespace pbrt {
PBRT_CONST uint16_t BlueNoiseTextures[NumBlueNoiseTextures][BlueNoiseResolution][BlueNoiseResolution] = {};
PBRT_CONST uint32_t pmj02bnSamples[nPMJ02bnSets][nPMJ02bnSamples][2] = {};

void RenderWavefront(BasicScene &) {
    std::fprintf(stderr, "oracle/_ref: the wavefront/GPU path is not part of the trimmed reference build\n");
    std::abort();
}
}  // namespace pbrt
