// radix_sort.cuh -- hand-written stable LSD radix sort of (uint32 key, uint32 value) pairs for sm_100a.
//
// Used by the GA / NGA classification (classify_api.cu): points are ordered by (cloud, bin) while points of one
// bin keep their input order, which is exactly the order the reference builds with one std::vector per bin
// (ccicp2d/src/icpTools.cpp:39-60).  8 bits per pass, three kernels per pass:
//   rs_hist    per-tile digit histogram (shared-memory atomics)            -> hist[digit][tile]
//   rs_scan    exclusive scan of hist in digit-major order (one CTA; the array is 256 x tiles ints)
//   rs_scatter re-reads the tile and scatters; the position of an element is
//              hist[digit][tile] + (elements of the same digit earlier in the tile), the latter from
//              __match_any_sync peer masks (rank inside the warp) + a per-round scan over the warps.
// HBM traffic per pass: read keys (hist) + read keys, values + write keys, values (scatter) = 20 B / element.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rsort {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 16;                      // elements per thread per tile
constexpr int kTile = kThreads * kItems;        // 4096 elements per CTA

__host__ __device__ inline int num_tiles(int64_t n) { return (int)((n + kTile - 1) / kTile); }

__global__ void __launch_bounds__(kThreads) rs_hist(const uint32_t* __restrict__ keys, int64_t n, int shift,
                                                    uint32_t* __restrict__ hist, int tiles) {
  __shared__ uint32_t h[256];
  const int tid = threadIdx.x;
  h[tid] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kTile;
#pragma unroll
  for (int r = 0; r < kItems; r++) {
    const int64_t i = base + r * kThreads + tid;
    if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
  }
  __syncthreads();
  hist[(size_t)tid * tiles + blockIdx.x] = h[tid];
}

// one CTA of 1024 threads: exclusive scan of `m` uint32 in place
__global__ void __launch_bounds__(1024) rs_scan(uint32_t* __restrict__ a, int64_t m) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  constexpr int kPer = 8;
  for (int64_t base = 0; base < m; base += 1024 * kPer) {
    uint32_t v[kPer], sum = 0;
    const int64_t i0 = base + (int64_t)tid * kPer;
#pragma unroll
    for (int k = 0; k < kPer; k++) { v[k] = (i0 + k < m) ? a[i0 + k] : 0u; sum += v[k]; }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (w == 0) {
      uint32_t x = wsum[lane], xi = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, xi, o); if (lane >= o) xi += t; }
      wsum[lane] = xi - x;                                   // exclusive offset of warp `lane`
    }
    __syncthreads();
    uint32_t run = carry_s + wsum[w] + incl - sum;
#pragma unroll
    for (int k = 0; k < kPer; k++) { if (i0 + k < m) a[i0 + k] = run; run += v[k]; }
    __syncthreads();
    if (tid == 1023) carry_s = run;                          // run of the last thread = carry + chunk total
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kThreads) rs_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                       uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                       int64_t n, int shift, const uint32_t* __restrict__ hist, int tiles) {
  __shared__ uint32_t base_s[256];            // global position of the next element of each digit from this tile
  __shared__ uint32_t wcnt[kWarps][256];      // per round: elements of each digit in each warp
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  base_s[tid] = hist[(size_t)tid * tiles + blockIdx.x];
  const int64_t tbase = (int64_t)blockIdx.x * kTile;
  for (int r = 0; r < kItems; r++) {
#pragma unroll
    for (int k = 0; k < kWarps; k++) wcnt[k][tid] = 0;
    __syncthreads();
    const int64_t i = tbase + r * kThreads + tid;
    const bool valid = i < n;
    uint32_t key = 0, val = 0, d = 256;        // 256 = "no element" (never matches a real digit)
    if (valid) { key = keys_in[i]; val = vals_in[i]; d = (key >> shift) & 255u; }
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const uint32_t before = __popc(peers & ((1u << lane) - 1u));
    if (valid && before == 0) wcnt[w][d] = __popc(peers);    // one writer per (warp, digit)
    __syncthreads();
    // digit `tid`: exclusive scan over the warps, then advance the tile's running base
    uint32_t run = base_s[tid];
#pragma unroll
    for (int k = 0; k < kWarps; k++) { const uint32_t c = wcnt[k][tid]; wcnt[k][tid] = run; run += c; }
    base_s[tid] = run;
    __syncthreads();
    if (valid) {
      const uint32_t pos = wcnt[w][d] + before;
      keys_out[pos] = key; vals_out[pos] = val;
    }
    __syncthreads();
  }
}

// Sorts n pairs by the low `bits` bits of the key.  keys/vals are ping-ponged with keys_tmp/vals_tmp; returns 0 if
// the result is in (keys, vals), 1 if it is in (keys_tmp, vals_tmp).  hist needs 256 * num_tiles(n) uint32.
inline int sort_pairs(uint32_t* keys, uint32_t* vals, uint32_t* keys_tmp, uint32_t* vals_tmp, int64_t n, int bits,
                      uint32_t* hist, cudaStream_t st, int* launches) {
  if (n <= 0) return 0;
  const int tiles = num_tiles(n);
  int side = 0;
  for (int shift = 0; shift < bits; shift += 8) {
    uint32_t* ki = side ? keys_tmp : keys; uint32_t* vi = side ? vals_tmp : vals;
    uint32_t* ko = side ? keys : keys_tmp; uint32_t* vo = side ? vals : vals_tmp;
    rs_hist<<<tiles, kThreads, 0, st>>>(ki, n, shift, hist, tiles);
    rs_scan<<<1, 1024, 0, st>>>(hist, (int64_t)256 * tiles);
    rs_scatter<<<tiles, kThreads, 0, st>>>(ki, vi, ko, vo, n, shift, hist, tiles);
    if (launches) *launches += 3;
    side ^= 1;
  }
  return side;
}

}  // namespace rsort
