// Coherence pass for the two tracking stages: a counting sort of a stage queue by (coarse cell of the ray origin, octant
// of the ray direction), so that the rays a warp of k_medium_rs / k_shadow picks up together start in the same part of
// the grid and walk it in the same direction (their DDA steps, majorant reads and density cell records then share
// cache lines, and their walks have similar lengths).  Only the ORDER of queue items changes: every path keeps its own
// slot, sampler stream and arithmetic, so results are unchanged (film sums already are order-free fp64 atomics).
//
//   k_bin_count   : key per queue item + histogram (warp-aggregated atomics)
//   k_bin_scan    : exclusive scan of the kBins counters (one CTA), histogram cleared for the next use
//   k_bin_scatter : items written to their bin's range
#pragma once
#include "kernels.cuh"

namespace vspg {

constexpr int kBinCells = 8;                                   // cells per axis of the binning box
constexpr int kBins = kBinCells * kBinCells * kBinCells * 8;   // x 8 direction octants = 4096

// which = 0: the path ray (o_dx / d_med), 1: the parked shadow ray (sh_o / sh_d)
__global__ void __launch_bounds__(kBlock) k_bin_count(const DScene sc, const Wavefront W, const Queue in, int which, int *keys, int *hist) {
    const int n = *in.count;
    for (int i = gtid(); i < n; i += gstride()) {
        const int slot = in.items[i];
        float4 a, b;
        if (which == 0) { a = W.P.o_dx[slot]; b = W.P.d_med[slot]; } else { a = W.P.sh_o[slot]; b = W.P.sh_d[slot]; }
        int cx = clampi((int)((a.x - sc.bin_lo[0]) * sc.bin_scale[0]), 0, kBinCells - 1);
        int cy = clampi((int)((a.y - sc.bin_lo[1]) * sc.bin_scale[1]), 0, kBinCells - 1);
        int cz = clampi((int)((a.z - sc.bin_lo[2]) * sc.bin_scale[2]), 0, kBinCells - 1);
        int oct = (a.w < 0 ? 1 : 0) | (b.x < 0 ? 2 : 0) | (b.y < 0 ? 4 : 0);
        int key = (((cz * kBinCells + cy) * kBinCells + cx) << 3) | oct;
        keys[i] = key;
        // one atomic per distinct key of the warp
        unsigned peers = __match_any_sync(__activemask(), key);
        if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&hist[key], __popc(peers));
    }
}

__global__ void __launch_bounds__(1024) k_bin_scan(int *hist, int *offsets) {
    __shared__ int part[1024];
    constexpr int per = kBins / 1024;
    int v[per], sum = 0;
#pragma unroll
    for (int k = 0; k < per; k++) { v[k] = hist[threadIdx.x * per + k]; sum += v[k]; hist[threadIdx.x * per + k] = 0; }
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        int t = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += t;
        __syncthreads();
    }
    int base = part[threadIdx.x] - sum;
#pragma unroll
    for (int k = 0; k < per; k++) { offsets[threadIdx.x * per + k] = base; base += v[k]; }
}

__global__ void __launch_bounds__(kBlock) k_bin_scatter(const Queue in, const int *keys, int *offsets, int *outItems) {
    const int n = *in.count;
    for (int i = gtid(); i < n; i += gstride()) {
        const int slot = in.items[i], key = keys[i];
        unsigned peers = __match_any_sync(__activemask(), key);
        const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&offsets[key], __popc(peers));
        base = __shfl_sync(peers, base, leader);
        outItems[base + __popc(peers & ((1u << lane) - 1))] = slot;
    }
}

}  // namespace vspg
