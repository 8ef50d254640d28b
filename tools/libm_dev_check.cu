// Diagnostic: include/vspg_libm.h evaluated on the device vs on the host, bit for bit.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include "../include/vspg_libm.h"
#include "../include/vspg_rsqrtss.h"
__global__ void k(const float *x, const float *y, int n, float *o, double *od) {
    int i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    float a = x[i], b = y[i]; float s, c; vspg_sincosf(a * 7 - 3, &s, &c);
    o[i * 12 + 0] = vspg_logf(a); o[i * 12 + 1] = s; o[i * 12 + 2] = c; o[i * 12 + 3] = vspg_asinf(2 * a - 1); o[i * 12 + 4] = vspg_acosf(2 * a - 1);
    o[i * 12 + 5] = vspg_atan2f(2 * a - 1, 2 * b - 1); o[i * 12 + 6] = vspg_powf(a, 0.5f + 0.5f * b); o[i * 12 + 7] = vspg_expf(a * 20 - 10);
    o[i * 12 + 8] = vspg_tanf(a * 3 - 1.5f); o[i * 12 + 9] = vspg_embree_rsqrt(a + 0.01f); o[i * 12 + 10] = (float)(-vspg_log(1.0 - (double)(a * b))); o[i * 12 + 11] = vspg_atanhf(2 * a - 1);
    od[i] = vspg_log(1.0 - (double)(a * b));
}
int main() {
    int n = 1 << 22; std::vector<float> x(n), y(n); srand48(3);
    for (int i = 0; i < n; i++) { x[i] = (float)drand48(); y[i] = (float)drand48(); }
    float *dx, *dy, *dout; double *dd; cudaMalloc(&dx, n * 4); cudaMalloc(&dy, n * 4); cudaMalloc(&dout, (size_t)n * 48); cudaMalloc(&dd, (size_t)n * 8);
    cudaMemcpy(dx, x.data(), n * 4, cudaMemcpyHostToDevice); cudaMemcpy(dy, y.data(), n * 4, cudaMemcpyHostToDevice);
    k<<<(n + 255) / 256, 256>>>(dx, dy, n, dout, dd);
    std::vector<float> o((size_t)n * 12); std::vector<double> od(n);
    cudaMemcpy(o.data(), dout, (size_t)n * 48, cudaMemcpyDeviceToHost); cudaMemcpy(od.data(), dd, (size_t)n * 8, cudaMemcpyDeviceToHost);
    long bad[13] = {0};
    for (int i = 0; i < n; i++) {
        float a = x[i], b = y[i], h[12]; float s, c; vspg_sincosf(a * 7 - 3, &s, &c);
        h[0] = vspg_logf(a); h[1] = s; h[2] = c; h[3] = vspg_asinf(2 * a - 1); h[4] = vspg_acosf(2 * a - 1); h[5] = vspg_atan2f(2 * a - 1, 2 * b - 1);
        h[6] = vspg_powf(a, 0.5f + 0.5f * b); h[7] = vspg_expf(a * 20 - 10); h[8] = vspg_tanf(a * 3 - 1.5f); h[9] = vspg_embree_rsqrt(a + 0.01f);
        h[10] = (float)(-vspg_log(1.0 - (double)(a * b))); h[11] = vspg_atanhf(2 * a - 1);
        for (int j = 0; j < 12; j++) if (memcmp(&h[j], &o[(size_t)i * 12 + j], 4)) { if (bad[j] < 2) printf("fn %d x=%.9g y=%.9g host %.9g dev %.9g\n", j, a, b, h[j], o[(size_t)i * 12 + j]); bad[j]++; }
        double hd = vspg_log(1.0 - (double)(a * b)); if (memcmp(&hd, &od[i], 8)) { if (bad[12] < 3) printf("logd x=%.17g host %.17g dev %.17g\n", 1.0 - (double)(a * b), hd, od[i]); bad[12]++; }
    }
    const char *nm[13] = {"logf", "sin", "cos", "asinf", "acosf", "atan2f", "powf", "expf", "tanf", "embree_rsqrt", "-log(1-ab)->f", "atanhf", "log(double)"};
    for (int j = 0; j < 13; j++) printf("%-14s device/host mismatches: %ld of %d\n", nm[j], bad[j], n);
    return 0;
}
