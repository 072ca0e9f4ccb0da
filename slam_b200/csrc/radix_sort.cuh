// radix_sort.cuh -- hand-written stable LSD radix sort of (uint32 key, uint32 value) pairs for sm_100a.
//
// Used by the GA / NGA classification (classify_api.cu): points are ordered by (cloud, bin) while points of one
// bin keep their input order, which is exactly the order the reference builds with one std::vector per bin
// (ccicp2d/src/icpTools.cpp:39-60).  8 bits per pass, three kernels per pass:
//   rs_hist    per-tile digit histogram (shared-memory atomics)            -> hist[digit][tile]
//   rs_scan_*  exclusive scan of hist in digit-major order (chunk totals, scan of totals, per-chunk scan)
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
// digit-major histogram table + one total per scan chunk
__host__ __device__ inline size_t hist_words(int64_t n) {
  const size_t m = (size_t)256 * num_tiles(n);
  return m + (m + 8191) / 8192 + 16;
}

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

// Exclusive scan of `m` uint32 in place, three small kernels: per-chunk totals, scan of the totals (one CTA),
// per-chunk scan seeded with its offset.  (A single-CTA scan of the 256 x tiles table took 0.8 ms per pass for
// 18 M keys -- 4x the scatter -- which is why this is spread over the GPU.)
constexpr int kScanChunk = 1024 * 8;            // entries per CTA: 1024 threads x 8

__device__ __forceinline__ uint32_t block_excl_scan_1024(uint32_t sum, uint32_t* wsum, uint32_t& total) {
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  uint32_t incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  __syncthreads();
  if (lane == 31) wsum[w] = incl;
  __syncthreads();
  if (w == 0) {
    uint32_t x = wsum[lane], xi = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, xi, o); if (lane >= o) xi += t; }
    wsum[lane] = xi - x;                                   // exclusive offset of warp `lane`
    if (lane == 31) wsum[32] = xi;                         // block total
  }
  __syncthreads();
  total = wsum[32];
  return wsum[w] + incl - sum;
}

__global__ void __launch_bounds__(1024) rs_scan_reduce(const uint32_t* __restrict__ a, int64_t m, uint32_t* __restrict__ part) {
  __shared__ uint32_t wsum[33];
  const int64_t i0 = (int64_t)blockIdx.x * kScanChunk + (int64_t)threadIdx.x * 8;
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) sum += (i0 + k < m) ? a[i0 + k] : 0u;
  uint32_t total;
  block_excl_scan_1024(sum, wsum, total);
  if (threadIdx.x == 0) part[blockIdx.x] = total;
}

// one CTA: exclusive scan of the chunk totals (any count, 1024 x 8 at a time)
__global__ void __launch_bounds__(1024) rs_scan_top(uint32_t* __restrict__ part, int64_t np) {
  __shared__ uint32_t wsum[33];
  uint32_t carry = 0;
  for (int64_t base = 0; base < np; base += kScanChunk) {
    uint32_t v[8], sum = 0;
    const int64_t i0 = base + (int64_t)threadIdx.x * 8;
#pragma unroll
    for (int k = 0; k < 8; k++) { v[k] = (i0 + k < np) ? part[i0 + k] : 0u; sum += v[k]; }
    uint32_t total;
    uint32_t run = carry + block_excl_scan_1024(sum, wsum, total);
#pragma unroll
    for (int k = 0; k < 8; k++) { if (i0 + k < np) part[i0 + k] = run; run += v[k]; }
    carry += total;
  }
}

__global__ void __launch_bounds__(1024) rs_scan_apply(uint32_t* __restrict__ a, int64_t m, const uint32_t* __restrict__ part) {
  __shared__ uint32_t wsum[33];
  const int64_t i0 = (int64_t)blockIdx.x * kScanChunk + (int64_t)threadIdx.x * 8;
  uint32_t v[8], sum = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) { v[k] = (i0 + k < m) ? a[i0 + k] : 0u; sum += v[k]; }
  uint32_t total;
  uint32_t run = part[blockIdx.x] + block_excl_scan_1024(sum, wsum, total);
#pragma unroll
  for (int k = 0; k < 8; k++) { if (i0 + k < m) a[i0 + k] = run; run += v[k]; }
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
// the result is in (keys, vals), 1 if it is in (keys_tmp, vals_tmp).  hist needs hist_words(n) uint32.
inline int sort_pairs(uint32_t* keys, uint32_t* vals, uint32_t* keys_tmp, uint32_t* vals_tmp, int64_t n, int bits,
                      uint32_t* hist, cudaStream_t st, int* launches) {
  if (n <= 0) return 0;
  const int tiles = num_tiles(n);
  int side = 0;
  for (int shift = 0; shift < bits; shift += 8) {
    uint32_t* ki = side ? keys_tmp : keys; uint32_t* vi = side ? vals_tmp : vals;
    uint32_t* ko = side ? keys : keys_tmp; uint32_t* vo = side ? vals : vals_tmp;
    rs_hist<<<tiles, kThreads, 0, st>>>(ki, n, shift, hist, tiles);
    const int64_t m = (int64_t)256 * tiles;
    const int chunks = (int)((m + kScanChunk - 1) / kScanChunk);
    uint32_t* part = hist + m;                   // chunk totals live behind the table
    rs_scan_reduce<<<chunks, 1024, 0, st>>>(hist, m, part);
    rs_scan_top<<<1, 1024, 0, st>>>(part, chunks);
    rs_scan_apply<<<chunks, 1024, 0, st>>>(hist, m, part);
    rs_scatter<<<tiles, kThreads, 0, st>>>(ki, vi, ko, vo, n, shift, hist, tiles);
    if (launches) *launches += 5;
    side ^= 1;
  }
  return side;
}

}  // namespace rsort
