// classify.cuh -- GA / NGA classification of segmented obstacle points on the GPU
// (CCICP::classifyPoints, ccicp2d/src/icpTools.cpp:36-103; constants icpTools.h:24-26): SURVEY.md 8(f) rank 2.
//
//   classify_key_kernel   one CTA per <= 32 tiles of 4096 points of one cloud: bin index with the reference's double
//                         arithmetic, occupancy bit (per-cloud bitmap, 1200 x 1200 bits = 180 KB, gathered in shared
//                         memory first), sort key = bin;
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

// One CTA per group of up to 32 tiles of one cloud.  The cloud's occupancy bitmap is accumulated in SHARED memory
// (180 KB) and only its non-zero words go to the global bitmap: one global atomic per point -- 17.9 M scattered REDs --
// cost 100 us of a 155 us kernel, the points of a cloud touch ~7x fewer distinct words.
constexpr int kKeyThreads = 1024;
constexpr int kKeyItems = ssort::kTile / kKeyThreads;          // points per thread per tile
constexpr size_t kKeySmemBytes = (size_t)kWordsPerCloud * 4 + 4 * ssort::kDigits * 4;

__global__ void __launch_bounds__(kKeyThreads, 1) classify_key_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ off,
                                                                      ssort::Plan plan, uint32_t* __restrict__ keys,
                                                                      uint32_t* __restrict__ bitmap, uint32_t* __restrict__ hist,
                                                                      uint32_t* __restrict__ part, int32_t* __restrict__ out_count) {
  extern __shared__ __align__(16) uint32_t key_smem[];
  uint32_t* sbm = key_smem;                                    // [kWordsPerCloud]
  uint32_t* h = key_smem + kWordsPerCloud;                     // [4][kDigits]: two tiles in flight, double buffered
  __shared__ int kept_s;
  const int group = blockIdx.x;
  if (group >= plan.group_base[plan.n_seg]) return;
  const int seg = ssort::seg_of(plan.group_base, plan.n_seg, group);
  const int gb = plan.group_base[seg], ng = plan.group_base[seg + 1] - gb;
  const int tb = plan.tile_base[seg], tc = plan.tile_base[seg + 1] - tb;
  const int t0 = (int)((int64_t)(group - gb) * tc / ng), t1 = (int)((int64_t)(group - gb + 1) * tc / ng);
  const int64_t seg_start = off[seg], seg_end = off[seg + 1];
  const int tid = threadIdx.x;
  for (int k = tid; k < kWordsPerCloud / 4; k += kKeyThreads) reinterpret_cast<uint4*>(sbm)[k] = make_uint4(0u, 0u, 0u, 0u);
  if (tid < 4 * ssort::kDigits) h[tid] = 0;
  if (tid == 0) kept_s = 0;
  __syncthreads();
  int kept = 0;
  for (int t = t0, pair = 0; t < t1; t += 2, pair++) {
    uint32_t* h2 = h + (pair & 1) * 2 * ssort::kDigits;         // histograms of tiles t and t + 1
    float x[2 * kKeyItems], y[2 * kKeyItems];
#pragma unroll
    for (int k = 0; k < 2 * kKeyItems; k++) {
      const int64_t i = seg_start + (int64_t)(t + k / kKeyItems) * ssort::kTile + (k % kKeyItems) * kKeyThreads + tid;
      const bool in = i < seg_end && t + k / kKeyItems < t1;
      x[k] = in ? xyz[3 * (size_t)i] : 0.0f;
      y[k] = in ? xyz[3 * (size_t)i + 1] : 0.0f;
    }
#pragma unroll
    for (int k = 0; k < 2 * kKeyItems; k++) {
      const int64_t i = seg_start + (int64_t)(t + k / kKeyItems) * ssort::kTile + (k % kKeyItems) * kKeyThreads + tid;
      if (!(i < seg_end && t + k / kKeyItems < t1)) continue;
      // icpTools.cpp:44-52: offset = NUMBINSGA*RESOLUTION/2 (double); floor((float + double) / double)
      const double offset = (double)kBins * kRes / 2.0;
      const double fx = floor(((double)x[k] + offset) / kRes), fy = floor(((double)y[k] + offset) / kRes);
      uint32_t bin = kDropped;
      if (fx >= 0.0 && fx < (double)kBins && fy >= 0.0 && fy < (double)kBins) {     // false for NaN
        const int bx = (int)fx, by = (int)fy;
        const int hh = bx * kBins + by;                                             // :55
        atomicOr(&sbm[hh >> 5], 1u << (hh & 31));
        if (bx != 0 && bx != kBins - 1 && by != 0 && by != kBins - 1) { bin = (uint32_t)hh; kept++; }   // :72-77 border bins are not emitted
      }
      keys[i] = bin;
      atomicAdd(&h2[(k / kKeyItems) * ssort::kDigits + (bin & (ssort::kDigits - 1))], 1u);
    }
    __syncthreads();
    if (tid < 2 * ssort::kDigits) {              // write the pair's histograms out, clear them for the pair after next
      const int tt = t + tid / ssort::kDigits;        // warp-uniform
      if (tt < t1) ssort::publish_hist(hist, part, ssort::hist_index(tb, tc, tt, tid % ssort::kDigits), h2[tid]);
      h2[tid] = 0;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) kept += __shfl_xor_sync(0xffffffffu, kept, o);
  if ((tid & 31) == 0 && kept) atomicAdd(&kept_s, kept);
  __syncthreads();
  uint32_t* bm = bitmap + (size_t)seg * kWordsPerCloud;
  for (int k = tid; k < kWordsPerCloud; k += kKeyThreads) {
    const uint32_t v = sbm[k];
    if (v) { if (ng == 1) bm[k] = v; else atomicOr(&bm[k], v); }
  }
  if (tid == 0 && kept_s) atomicAdd(&out_count[seg], kept_s);
}

// Consumer of the last radix pass: element `src` (global input index) of cloud `c` lands at sorted position `pos`.
struct EmitPoint {
  const float* __restrict__ xyz; const int32_t* __restrict__ off; const uint32_t* __restrict__ bitmap;
  float* __restrict__ out_xyz; uint8_t* __restrict__ out_ga; int32_t* __restrict__ out_src;
  struct Item { float x, y, z; uint32_t rows[3]; };
  // rows[q]: bits (by-1, by, by+1) of bitmap row bx-1+q in its low three bits
  __device__ __forceinline__ Item fetch(uint32_t, uint32_t bin, uint32_t src, int c) const {
    Item it{0.0f, 0.0f, 0.0f, {0u, 0u, 0u}};
    if (bin == kDropped) return it;
    const int bx = (int)(bin / kBins), by = (int)(bin % kBins);
    const uint32_t* bm = bitmap + (size_t)c * kWordsPerCloud;
#pragma unroll
    for (int q = 0; q < 3; q++) {
      const int h0 = (bx - 1 + q) * kBins + (by - 1);            // >= 0: border bins are never emitted
      const uint32_t lo = bm[h0 >> 5], hi = bm[(h0 + 2) >> 5];   // same word unless the window straddles
      it.rows[q] = __funnelshift_r(lo, ((h0 + 2) >> 5) != (h0 >> 5) ? hi : 0u, h0 & 31) & 7u;
    }
    const size_t s3 = 3 * (size_t)src;
    it.x = xyz[s3]; it.y = xyz[s3 + 1]; it.z = xyz[s3 + 2];
    return it;
  }
  __device__ __forceinline__ void commit(uint32_t pos, uint32_t bin, uint32_t src, int c, const Item& it) const {
    if (bin == kDropped) return;
    // icpTools.cpp:79-90: empty bins among the 8 neighbours
    const int occupied = __popc(it.rows[0]) + __popc(it.rows[1] & 5u) + __popc(it.rows[2]);
    const size_t p3 = 3 * (size_t)pos;
    out_xyz[p3] = it.x; out_xyz[p3 + 1] = it.y; out_xyz[p3 + 2] = it.z;
    out_ga[pos] = (8 - occupied) >= kAdjThresh ? 1 : 0;          // :98
    if (out_src) out_src[pos] = (int32_t)((int64_t)src - off[c]);
  }
};

struct Scratch {          // device pointers, sized by the caller
  uint32_t *keys, *vals, *keys_tmp, *vals_tmp, *hist, *bitmap;
  int32_t *tile_base, *tile_seg, *group_base; // [n_clouds + 1], [ssort::max_tiles(n, n_clouds)], [n_clouds + 1]
};

inline size_t hist_words(int64_t n, int n_clouds) { return ssort::hist_words(n, n_clouds); }

// Enqueues the whole classification; returns the number of kernels launched.
inline int run(const float* d_xyz, const int32_t* d_off, int n_clouds, int64_t n, const Scratch& s, float* d_out_xyz,
               uint8_t* d_out_ga, int32_t* d_out_src, int32_t* d_out_count, cudaStream_t st) {
  int launches = 0;
  cudaMemsetAsync(s.bitmap, 0, (size_t)n_clouds * kWordsPerCloud * 4, st);
  cudaMemsetAsync(d_out_count, 0, (size_t)n_clouds * 4, st);
  if (n <= 0) return 0;
  const ssort::Plan plan{s.tile_base, s.tile_seg, s.group_base, n_clouds};
  const int tiles = ssort::max_tiles(n, n_clouds);
  uint32_t* part[3];
  for (int k = 0; k < 3; k++) part[k] = ssort::part_of(s.hist, n, n_clouds, k);
  cudaMemsetAsync(part[0], 0, 3 * ssort::part_words(n, n_clouds) * 4, st);
  ssort::plan_kernel<<<1, 1024, 0, st>>>(d_off, plan);
  cudaFuncSetAttribute(classify_key_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kKeySmemBytes);   // per device
  classify_key_kernel<<<ssort::max_groups(n, n_clouds), kKeyThreads, kKeySmemBytes, st>>>(d_xyz, d_off, plan, s.keys, s.bitmap, s.hist,
                                                                                         part[0], d_out_count);
  launches += 2;
  // pass 0 (bits 0-6): values implicit; pass 1 (7-13); pass 2 (14-20) emits
  ssort::scan_table(s.hist, part[0], plan, n, st, &launches);
  ssort::seg_scatter<true, ssort::StorePairs><<<tiles, ssort::kThreads, 0, st>>>(s.keys, nullptr, d_off, plan, 0, s.hist,
                                                                                 ssort::StorePairs{s.keys_tmp, s.vals_tmp});
  ssort::seg_hist<<<tiles, ssort::kThreads, 0, st>>>(s.keys_tmp, d_off, plan, ssort::kBits, s.hist, part[1]);
  ssort::scan_table(s.hist, part[1], plan, n, st, &launches);
  ssort::seg_scatter<false, ssort::StorePairs><<<tiles, ssort::kThreads, 0, st>>>(s.keys_tmp, s.vals_tmp, d_off, plan, ssort::kBits,
                                                                                  s.hist, ssort::StorePairs{s.keys, s.vals});
  ssort::seg_hist<<<tiles, ssort::kThreads, 0, st>>>(s.keys, d_off, plan, 2 * ssort::kBits, s.hist, part[2]);
  ssort::scan_table(s.hist, part[2], plan, n, st, &launches);
  ssort::seg_scatter<false, EmitPoint><<<tiles, ssort::kThreads, 0, st>>>(
      s.keys, s.vals, d_off, plan, 2 * ssort::kBits, s.hist, EmitPoint{d_xyz, d_off, s.bitmap, d_out_xyz, d_out_ga, d_out_src});
  launches += 5;
  return launches;
}

}  // namespace classify
