// classify.cuh -- GA / NGA classification of segmented obstacle points on the GPU
// (CCICP::classifyPoints, ccicp2d/src/icpTools.cpp:36-103; constants icpTools.h:24-26): SURVEY.md 8(f) rank 2.
//
//   classify_key_kernel   one CTA per tile of 4096 points of one cloud: bin index with the reference's double
//                         arithmetic, occupancy bit (per-cloud bitmap, 1200 x 1200 bits = 180 KB), sort key = bin;
//                         points the reference drops (outside the grid, NaN, outermost ring of bins) get the largest
//                         key so they end up behind the kept points of their cloud.  Also produces the digit
//                         histogram of the first radix pass and the per-cloud count of kept points.
//   ssort::*              stable segmented LSD radix sort on the 21-bit bin (3 passes of 7 bits): bin-major output,
//                         input order inside a bin
//   EmitPoint             fused into the last scatter pass: gathers the point, counts empty bins among the 8
//                         neighbours in the bitmap, writes xyz / ground_adj / source index at the sorted position
// All stages are HBM-streaming; per point: 12 B in + 4 B key, 4+8 B (pass 0), 4+8+8 B (pass 1), 4+8 B + the 12 B
// gather + 12+1(+4) B out (pass 2).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "seg_sort.cuh"

namespace classify {

constexpr int kBins = 1200;                 // NUMBINSGA   (icpTools.h:24)
constexpr double kRes = 0.5;                // RESOLUTION  (icpTools.h:25)
constexpr int kAdjThresh = 2;               // GRD_ADJ_THRESH (icpTools.h:26)
constexpr uint32_t kDropped = 0x1FFFFFu;    // > any real bin index (1200*1200 = 1,440,000 < 2^21)
static_assert(3 * ssort::kBits == 21, "three radix passes cover the 21-bit bin index");
constexpr int kWordsPerCloud = (kBins * kBins + 31) / 32;
constexpr int kMaxClouds = 2048;            // bitmap memory: 180 KB per cloud

__global__ void __launch_bounds__(ssort::kThreads) classify_key_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ off,
                                                                      ssort::Plan plan, uint32_t* __restrict__ keys,
                                                                      uint32_t* __restrict__ bitmap, uint32_t* __restrict__ hist,
                                                                      int32_t* __restrict__ out_count) {
  __shared__ uint32_t h[ssort::kDigits];
  __shared__ int kept_s;
  ssort::TileRef r;
  if (!ssort::locate_tile(plan, off, r)) return;
  const int tid = threadIdx.x;
  if (tid < ssort::kDigits) h[tid] = 0;
  if (tid == 0) kept_s = 0;
  __syncthreads();
  float x[ssort::kItems], y[ssort::kItems];
#pragma unroll
  for (int k = 0; k < ssort::kItems; k++) {
    const int e = k * ssort::kThreads + tid;
    const size_t i = (size_t)(r.start + e);
    x[k] = e < r.count ? xyz[3 * i] : 0.0f;
    y[k] = e < r.count ? xyz[3 * i + 1] : 0.0f;
  }
  uint32_t* bm = bitmap + (size_t)r.seg * kWordsPerCloud;
  int kept = 0;
#pragma unroll
  for (int k = 0; k < ssort::kItems; k++) {
    const int e = k * ssort::kThreads + tid;
    if (e >= r.count) continue;
    // icpTools.cpp:44-52: offset = NUMBINSGA*RESOLUTION/2 (double); floor((float + double) / double)
    const double offset = (double)kBins * kRes / 2.0;
    const double fx = floor(((double)x[k] + offset) / kRes), fy = floor(((double)y[k] + offset) / kRes);
    uint32_t bin = kDropped;
    if (fx >= 0.0 && fx < (double)kBins && fy >= 0.0 && fy < (double)kBins) {       // false for NaN
      const int bx = (int)fx, by = (int)fy;
      const int hh = bx * kBins + by;                                               // :55
      atomicOr(&bm[hh >> 5], 1u << (hh & 31));
      if (bx != 0 && bx != kBins - 1 && by != 0 && by != kBins - 1) { bin = (uint32_t)hh; kept++; }   // :72-77 border bins are not emitted
    }
    keys[r.start + e] = bin;
    atomicAdd(&h[bin & (ssort::kDigits - 1)], 1u);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) kept += __shfl_xor_sync(0xffffffffu, kept, o);
  if ((tid & 31) == 0 && kept) atomicAdd(&kept_s, kept);
  __syncthreads();
  if (tid < ssort::kDigits) hist[ssort::hist_index(r.tb, r.tc, r.t, tid)] = h[tid];
  if (tid == 0 && kept_s) atomicAdd(&out_count[r.seg], kept_s);
}

// Consumer of the last radix pass: element `src` (global input index) of cloud `c` lands at sorted position `pos`.
struct EmitPoint {
  const float* __restrict__ xyz; const int32_t* __restrict__ off; const uint32_t* __restrict__ bitmap;
  float* __restrict__ out_xyz; uint8_t* __restrict__ out_ga; int32_t* __restrict__ out_src;
  __device__ __forceinline__ void operator()(uint32_t pos, uint32_t bin, uint32_t src, int c) const {
    if (bin == kDropped) return;
    const int bx = (int)(bin / kBins), by = (int)(bin % kBins);
    const uint32_t* bm = bitmap + (size_t)c * kWordsPerCloud;
    int empty = 0;                                                                  // :79-90
#pragma unroll
    for (int dq = -1; dq <= 1; dq++) {
#pragma unroll
      for (int dr = -1; dr <= 1; dr++) {
        if (dq == 0 && dr == 0) continue;
        const int hh = (bx + dq) * kBins + (by + dr);
        empty += ((bm[hh >> 5] >> (hh & 31)) & 1u) ? 0 : 1;
      }
    }
    const size_t s3 = 3 * (size_t)src, p3 = 3 * (size_t)pos;
    const float px = xyz[s3], py = xyz[s3 + 1], pz = xyz[s3 + 2];
    out_xyz[p3] = px; out_xyz[p3 + 1] = py; out_xyz[p3 + 2] = pz;
    out_ga[pos] = empty >= kAdjThresh ? 1 : 0;                                      // :98
    if (out_src) out_src[pos] = (int32_t)((int64_t)src - off[c]);
  }
};

struct Scratch {          // device pointers, sized by the caller
  uint32_t *keys, *vals, *keys_tmp, *vals_tmp, *hist, *bitmap;
  int32_t *tile_base, *tile_seg;              // [n_clouds + 1], [ssort::max_tiles(n, n_clouds)]
};

inline size_t hist_words(int64_t n, int n_clouds) { return ssort::hist_words(n, n_clouds); }

// Enqueues the whole classification; returns the number of kernels launched.
inline int run(const float* d_xyz, const int32_t* d_off, int n_clouds, int64_t n, const Scratch& s, float* d_out_xyz,
               uint8_t* d_out_ga, int32_t* d_out_src, int32_t* d_out_count, cudaStream_t st) {
  int launches = 0;
  cudaMemsetAsync(s.bitmap, 0, (size_t)n_clouds * kWordsPerCloud * 4, st);
  cudaMemsetAsync(d_out_count, 0, (size_t)n_clouds * 4, st);
  if (n <= 0) return 0;
  const ssort::Plan plan{s.tile_base, s.tile_seg, n_clouds};
  const int tiles = ssort::max_tiles(n, n_clouds);
  ssort::plan_kernel<<<1, 1024, 0, st>>>(d_off, plan);
  ssort::plan_fill_kernel<<<n_clouds, 256, 0, st>>>(plan);
  classify_key_kernel<<<tiles, ssort::kThreads, 0, st>>>(d_xyz, d_off, plan, s.keys, s.bitmap, s.hist, d_out_count);
  launches += 3;
  // pass 0 (bits 0-6): values implicit; pass 1 (7-13); pass 2 (14-20) emits
  ssort::scan_table(s.hist, plan, n, st, &launches);
  ssort::seg_scatter<true, ssort::StorePairs><<<tiles, ssort::kThreads, 0, st>>>(s.keys, nullptr, d_off, plan, 0, s.hist,
                                                                                 ssort::StorePairs{s.keys_tmp, s.vals_tmp});
  ssort::seg_hist<<<tiles, ssort::kThreads, 0, st>>>(s.keys_tmp, d_off, plan, ssort::kBits, s.hist);
  ssort::scan_table(s.hist, plan, n, st, &launches);
  ssort::seg_scatter<false, ssort::StorePairs><<<tiles, ssort::kThreads, 0, st>>>(s.keys_tmp, s.vals_tmp, d_off, plan, ssort::kBits,
                                                                                  s.hist, ssort::StorePairs{s.keys, s.vals});
  ssort::seg_hist<<<tiles, ssort::kThreads, 0, st>>>(s.keys, d_off, plan, 2 * ssort::kBits, s.hist);
  ssort::scan_table(s.hist, plan, n, st, &launches);
  ssort::seg_scatter<false, EmitPoint><<<tiles, ssort::kThreads, 0, st>>>(
      s.keys, s.vals, d_off, plan, 2 * ssort::kBits, s.hist, EmitPoint{d_xyz, d_off, s.bitmap, d_out_xyz, d_out_ga, d_out_src});
  launches += 5;
  return launches;
}

}  // namespace classify
