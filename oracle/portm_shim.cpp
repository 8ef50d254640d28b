// Test infrastructure (NOT product code).
// include/vspg_libm.h under the C library's names: LD_PRELOAD=oracle/libportm.so makes a process -- the reference
// renderer built in oracle/_ref, or any other -- compute log / exp / pow / sin / cos / tan / asin / acos / atan / atan2 /
// atanh / cosh / sinh / tanh with the same portable code the CUDA kernels run, instead of whatever variant the host's
// libm selects for this CPU.  tests/golden/make_golden_pbrt.py produces the golden vectors under this shim.
// Built by oracle/Makefile with -fno-builtin -ffp-contract=off.
#include "../include/vspg_libm.h"
#include "../include/vspg_rsqrtss.h"

extern "C" {
float logf(float x) { return vspg_logf(x); }
double log(double x) { return vspg_log(x); }
float expf(float x) { return vspg_expf(x); }
float powf(float x, float y) { return vspg_powf(x, y); }
float sinf(float x) { return vspg_sinf(x); }
float cosf(float x) { return vspg_cosf(x); }
void sincosf(float x, float *s, float *c) { vspg_sincosf(x, s, c); }
float tanf(float x) { return vspg_tanf(x); }
float asinf(float x) { return vspg_asinf(x); }
float acosf(float x) { return vspg_acosf(x); }
float atanf(float x) { return vspg_atanf(x); }
float atan2f(float y, float x) { return vspg_atan2f(y, x); }
float atanhf(float x) { return vspg_atanhf(x); }
float coshf(float x) { return vspg_coshf(x); }
float sinhf(float x) { return vspg_sinhf(x); }
float tanhf(float x) { return vspg_tanhf(x); }
int vspg_portm_loaded(void) { return 1; }
float vspg_portm_embree_rsqrt(float x) { return vspg_embree_rsqrt(x); }   // (host value of include/vspg_rsqrtss.h, for tests/test_libm_device.py)
}
