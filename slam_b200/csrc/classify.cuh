// classify.cuh -- GA / NGA classification of segmented obstacle points on the GPU
// (CCICP::classifyPoints, ccicp2d/src/icpTools.cpp:36-103; constants icpTools.h:24-26): SURVEY.md 8(f) rank 2.
//
//   classify_bin_kernel   one thread per point: bin index with the reference's double arithmetic, occupancy bit
//                         (per-cloud bitmap, 1200 x 1200 bits = 180 KB), sort key (cloud << 21 | bin); points the
//                         reference drops (outside the grid, NaN, outermost ring of bins) get the largest bin value
//                         so they end up behind the kept points of their cloud
//   rsort::sort_pairs     stable LSD radix sort on (cloud, bin): bin-major output, input order inside a bin
//   classify_emit_kernel  one thread per sorted position: gathers the point, counts empty bins among the 8
//                         neighbours in the bitmap, writes xyz / ground_adj / source index and the per-cloud count
// All stages are HBM-streaming: per point 12 B in, 12+1(+4) B out, plus 8 B x 2 of key/value traffic and 20 B per
// radix pass.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "radix_sort.cuh"

namespace classify {

constexpr int kBins = 1200;                 // NUMBINSGA   (icpTools.h:24)
constexpr double kRes = 0.5;                // RESOLUTION  (icpTools.h:25)
constexpr int kAdjThresh = 2;               // GRD_ADJ_THRESH (icpTools.h:26)
constexpr uint32_t kDropped = 0x1FFFFFu;    // > any real bin index (1200*1200 = 1,440,000 < 2^21)
constexpr int kBinBits = 21;
constexpr int kWordsPerCloud = (kBins * kBins + 31) / 32;
constexpr int kMaxClouds = 2048;            // 11 bits of cloud id above the 21 bin bits

__device__ __forceinline__ int cloud_of(const int32_t* __restrict__ off, int n_clouds, int64_t i) {
  int lo = 0, hi = n_clouds;                // largest c with off[c] <= i
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((int64_t)off[mid] <= i) lo = mid; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(256) classify_bin_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ off,
                                                           int n_clouds, int64_t n, uint32_t* __restrict__ keys,
                                                           uint32_t* __restrict__ vals, uint32_t* __restrict__ bitmap) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = cloud_of(off, n_clouds, i);
  const float x = xyz[3 * i], y = xyz[3 * i + 1];
  // icpTools.cpp:44-52: offset = NUMBINSGA*RESOLUTION/2 (double); floor((float + double) / double)
  const double offset = (double)kBins * kRes / 2.0;
  const double fx = floor(((double)x + offset) / kRes), fy = floor(((double)y + offset) / kRes);
  uint32_t bin = kDropped;
  if (fx >= 0.0 && fx < (double)kBins && fy >= 0.0 && fy < (double)kBins) {       // false for NaN
    const int bx = (int)fx, by = (int)fy;
    const int h = bx * kBins + by;                                                  // :55
    atomicOr(&bitmap[(size_t)c * kWordsPerCloud + (h >> 5)], 1u << (h & 31));
    if (bx != 0 && bx != kBins - 1 && by != 0 && by != kBins - 1) bin = (uint32_t)h;   // :72-77 border bins are not emitted
  }
  keys[i] = ((uint32_t)c << kBinBits) | bin;
  vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) classify_emit_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ off,
                                                            int64_t n, const uint32_t* __restrict__ keys,
                                                            const uint32_t* __restrict__ vals, const uint32_t* __restrict__ bitmap,
                                                            float* __restrict__ out_xyz, uint8_t* __restrict__ out_ga,
                                                            int32_t* __restrict__ out_src, int32_t* __restrict__ out_count) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t key = keys[j];
  const uint32_t bin = key & kDropped;
  if (bin == kDropped) return;
  const int c = (int)(key >> kBinBits);
  const uint32_t src = vals[j];
  const int bx = (int)(bin / kBins), by = (int)(bin % kBins);
  const uint32_t* bm = bitmap + (size_t)c * kWordsPerCloud;
  int empty = 0;                                                                    // :79-90
#pragma unroll
  for (int dq = -1; dq <= 1; dq++) {
#pragma unroll
    for (int dr = -1; dr <= 1; dr++) {
      if (dq == 0 && dr == 0) continue;
      const int h = (bx + dq) * kBins + (by + dr);
      empty += ((bm[h >> 5] >> (h & 31)) & 1u) ? 0 : 1;
    }
  }
  out_xyz[3 * j] = xyz[3 * (size_t)src]; out_xyz[3 * j + 1] = xyz[3 * (size_t)src + 1]; out_xyz[3 * j + 2] = xyz[3 * (size_t)src + 2];
  out_ga[j] = empty >= kAdjThresh ? 1 : 0;                                          // :98
  if (out_src) out_src[j] = (int32_t)((int64_t)src - off[c]);
  // last kept element of the cloud -> count (clouds with nothing kept stay at the memset 0)
  const bool last = (j + 1 == (int64_t)off[c + 1]) || ((keys[j + 1] & kDropped) == kDropped) || ((int)(keys[j + 1] >> kBinBits) != c);
  if (last) out_count[c] = (int32_t)(j + 1 - off[c]);
}

struct Scratch {          // device pointers, sized by the caller
  uint32_t *keys, *vals, *keys_tmp, *vals_tmp, *hist, *bitmap;
};

inline size_t hist_words(int64_t n) { return rsort::hist_words(n); }

// Enqueues the whole classification; returns the number of kernels launched.
inline int run(const float* d_xyz, const int32_t* d_off, int n_clouds, int64_t n, const Scratch& s, float* d_out_xyz,
               uint8_t* d_out_ga, int32_t* d_out_src, int32_t* d_out_count, cudaStream_t st) {
  int launches = 0;
  cudaMemsetAsync(s.bitmap, 0, (size_t)n_clouds * kWordsPerCloud * 4, st);
  cudaMemsetAsync(d_out_count, 0, (size_t)n_clouds * 4, st);
  if (n <= 0) return 0;
  const int blocks = (int)((n + 255) / 256);
  classify_bin_kernel<<<blocks, 256, 0, st>>>(d_xyz, d_off, n_clouds, n, s.keys, s.vals, s.bitmap);
  launches++;
  int bits = kBinBits;
  for (int c = n_clouds - 1; c > 0; c >>= 1) bits++;
  const int side = rsort::sort_pairs(s.keys, s.vals, s.keys_tmp, s.vals_tmp, n, bits, s.hist, st, &launches);
  const uint32_t* k = side ? s.keys_tmp : s.keys;
  const uint32_t* v = side ? s.vals_tmp : s.vals;
  classify_emit_kernel<<<blocks, 256, 0, st>>>(d_xyz, d_off, n, k, v, s.bitmap, d_out_xyz, d_out_ga, d_out_src, d_out_count);
  launches++;
  return launches;
}

}  // namespace classify
