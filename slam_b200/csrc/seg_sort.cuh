// seg_sort.cuh -- hand-written stable SEGMENTED LSD radix sort of (uint32 key, uint32 value) pairs for sm_100a.
//
// Used by the GA / NGA classification (classify.cuh): the points of every cloud are ordered by bin while points of
// one bin keep their input order -- exactly the order the reference builds with one std::vector per bin
// (ccicp2d/src/icpTools.cpp:39-60).  The clouds are already contiguous in the input, so the cloud id is not part
// of the key: every tile of 4096 elements belongs to one segment, and the digit histogram table is laid out
// (segment, digit, tile-in-segment) so that ONE exclusive scan over the whole table yields global positions.
// 7 bits per pass (three passes for the 21-bit bin index).  Per pass:
//   seg_hist     per-tile digit histogram                       (pass 0: produced by the caller's key kernel)
//   scan_table   exclusive scan of the table, one kernel (the producers publish per-chunk totals with their counts)
//   seg_scatter  ranks the tile's elements (ballot peer masks inside a warp + per-warp digit counters, warp w owns the
//                w-th contiguous 512 elements so the counters need no block barrier between rounds), sorts the tile
//                in shared memory and writes runs of equal digits to consecutive addresses.  The last pass hands
//                (position, key, value) to a functor instead of storing the pair: the consumer is fused in.
// HBM traffic per element: pass 0: 4 B in (values are implicit) + 8 B out; later passes 4 B (hist) + 8 B + 8 B.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ssort {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 16;                      // elements per thread per tile
constexpr int kTile = kThreads * kItems;        // 4096 elements per CTA
constexpr int kWarpSpan = kTile / kWarps;       // 512 consecutive elements per warp
constexpr int kBits = 7;
constexpr int kDigits = 1 << kBits;

struct Plan {                // device arrays, filled by plan_kernel
  int32_t* tile_base;        // [n_seg + 1] first tile of each segment; tile_base[n_seg] = number of tiles
  int32_t* tile_seg;         // [max_tiles] segment of each tile
  int32_t* group_base;       // [n_seg + 1] or NULL: first tile GROUP of each segment (a segment's tiles split evenly into
                             // ceil(tiles / kGroupTiles) groups, for kernels that want one CTA per many tiles)
  int32_t n_seg;
};
constexpr int kGroupTiles = 32;

// sum over segments of ceil(n_s / kTile) <= n / kTile + n_seg
__host__ __device__ inline int max_tiles(int64_t n, int n_seg) { return (int)(n / kTile) + n_seg; }
__host__ __device__ inline int max_groups(int64_t n, int n_seg) { return max_tiles(n, n_seg) / kGroupTiles + n_seg; }
constexpr int kScanChunk = 1024 * 8;            // table entries per scan CTA: 1024 threads x 8
constexpr int kMaxPasses = 4;                   // chunk-total arrays behind the table, one per pass of a sort
__host__ __device__ inline size_t table_words(int64_t n, int n_seg) { return (size_t)kDigits * max_tiles(n, n_seg); }
__host__ __device__ inline size_t part_words(int64_t n, int n_seg) { return (table_words(n, n_seg) + kScanChunk - 1) / kScanChunk + 4; }
__host__ __device__ inline size_t hist_words(int64_t n, int n_seg) { return table_words(n, n_seg) + kMaxPasses * part_words(n, n_seg); }
// chunk totals of pass `pass` (zeroed by the caller before the first producer runs)
inline uint32_t* part_of(uint32_t* hist, int64_t n, int n_seg, int pass) { return hist + table_words(n, n_seg) + (size_t)pass * part_words(n, n_seg); }
__device__ __forceinline__ size_t hist_index(int tb, int tc, int t, int d) {
  return (size_t)kDigits * tb + (size_t)d * tc + t;
}
// A producer's warp publishes 32 table entries and adds them to the totals of their scan chunks (one atomic per
// distinct chunk of the warp), so that the scan needs no reduce pass.  All 32 lanes must call.
__device__ __forceinline__ void publish_hist(uint32_t* __restrict__ hist, uint32_t* __restrict__ part, size_t idx, uint32_t count) {
  hist[idx] = count;
  const uint32_t chunk = (uint32_t)(idx / kScanChunk);
  const uint32_t peers = __match_any_sync(0xffffffffu, chunk);
  const uint32_t sum = __reduce_add_sync(peers, count);
  if ((int)(threadIdx.x & 31) == __ffs(peers) - 1 && sum) atomicAdd(&part[chunk], sum);
}

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

// ---- tiling plan (one CTA): scan of the per-segment tile (and group) counts, then the tile -> segment labels
__device__ __forceinline__ int seg_of(const int32_t* base, int n_seg, int i) {
  int lo = 0, hi = n_seg;                    // largest s with base[s] <= i (empty segments share a base: take the last)
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (base[mid] <= i) lo = mid; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(1024) plan_kernel(const int32_t* __restrict__ off, Plan p) {
  __shared__ uint32_t wsum[33];
  uint32_t carry = 0, gcarry = 0;
  for (int base = 0; base < p.n_seg; base += 1024) {
    const int s = base + threadIdx.x;
    const uint32_t t = s < p.n_seg ? (uint32_t)((off[s + 1] - off[s] + kTile - 1) / kTile) : 0u;
    uint32_t total;
    const uint32_t ex = block_excl_scan_1024(t, wsum, total);
    if (s < p.n_seg) p.tile_base[s] = (int32_t)(carry + ex);
    carry += total;
    if (p.group_base) {
      const uint32_t g = (t + kGroupTiles - 1) / kGroupTiles;
      const uint32_t gex = block_excl_scan_1024(g, wsum, total);
      if (s < p.n_seg) p.group_base[s] = (int32_t)(gcarry + gex);
      gcarry += total;
    }
  }
  if (threadIdx.x == 0) {
    p.tile_base[p.n_seg] = (int32_t)carry;
    if (p.group_base) p.group_base[p.n_seg] = (int32_t)gcarry;
  }
  __syncthreads();                             // this CTA's global writes above are visible to all its threads
  for (int t = threadIdx.x; t < (int)carry; t += 1024) p.tile_seg[t] = seg_of(p.tile_base, p.n_seg, t);
}

// The tile a CTA works on.  false = this CTA is beyond the last tile (the grid is sized by max_tiles()).
struct TileRef { int seg, tb, tc, t, count; int64_t start; };
__device__ __forceinline__ bool locate_tile(const Plan& p, const int32_t* __restrict__ off, TileRef& r) {
  const int tile = blockIdx.x;
  if (tile >= p.tile_base[p.n_seg]) return false;
  r.seg = p.tile_seg[tile];
  r.tb = p.tile_base[r.seg];
  r.tc = p.tile_base[r.seg + 1] - r.tb;
  r.t = tile - r.tb;
  r.start = (int64_t)off[r.seg] + (int64_t)r.t * kTile;
  const int64_t left = (int64_t)off[r.seg + 1] - r.start;
  r.count = left < kTile ? (int)left : kTile;
  return true;
}

__global__ void __launch_bounds__(kThreads) seg_hist(const uint32_t* __restrict__ keys, const int32_t* __restrict__ off,
                                                     Plan p, int shift, uint32_t* __restrict__ hist, uint32_t* __restrict__ part) {
  __shared__ uint32_t h[kDigits];
  TileRef r;
  if (!locate_tile(p, off, r)) return;
  const int tid = threadIdx.x;
  if (tid < kDigits) h[tid] = 0;
  __syncthreads();
  uint32_t k[kItems];
#pragma unroll
  for (int i = 0; i < kItems; i++) {
    const int e = i * kThreads + tid;
    k[i] = e < r.count ? keys[r.start + e] : 0u;
  }
#pragma unroll
  for (int i = 0; i < kItems; i++)
    if (i * kThreads + tid < r.count) atomicAdd(&h[(k[i] >> shift) & (kDigits - 1)], 1u);
  __syncthreads();
  if (tid < kDigits) publish_hist(hist, part, hist_index(r.tb, r.tc, r.t, tid), h[tid]);
}

// ---- exclusive scan of the first kDigits * (number of tiles) table entries, in place: one kernel, every CTA adds up the
// totals of the chunks before its own (published by the producers) and scans its chunk of 8192 entries
__device__ __forceinline__ int64_t table_len(const Plan& p) { return (int64_t)kDigits * p.tile_base[p.n_seg]; }

__global__ void __launch_bounds__(1024) scan_table_kernel(uint32_t* __restrict__ a, Plan p, const uint32_t* __restrict__ part) {
  __shared__ uint32_t wsum[33];
  const int64_t m = table_len(p);
  if ((int64_t)blockIdx.x * kScanChunk >= m) return;
  uint32_t pre = 0, before;
  for (int c = threadIdx.x; c < (int)blockIdx.x; c += 1024) pre += part[c];
  block_excl_scan_1024(pre, wsum, before);
  const int64_t i0 = (int64_t)blockIdx.x * kScanChunk + (int64_t)threadIdx.x * 8;
  uint32_t v[8], sum = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) { v[k] = (i0 + k < m) ? a[i0 + k] : 0u; sum += v[k]; }
  uint32_t total;
  uint32_t run = before + block_excl_scan_1024(sum, wsum, total);
#pragma unroll
  for (int k = 0; k < 8; k++) { if (i0 + k < m) a[i0 + k] = run; run += v[k]; }
}

inline void scan_table(uint32_t* hist, const uint32_t* part, const Plan& p, int64_t n, cudaStream_t st, int* launches) {
  const int64_t m = (int64_t)table_words(n, p.n_seg);
  scan_table_kernel<<<(int)((m + kScanChunk - 1) / kScanChunk), 1024, 0, st>>>(hist, p, part);
  if (launches) *launches += 1;
}

// Consumers of a scatter pass are split in fetch (loads only) and commit (stores only) so that the loads of a batch of
// elements are all in flight before the first store: nothing here may alias, but the compiler cannot know that.
struct StorePairs {          // the ordinary pass: write the pair to its sorted position
  uint32_t* keys_out; uint32_t* vals_out;
  struct Item {};
  __device__ __forceinline__ Item fetch(uint32_t, uint32_t, uint32_t, int) const { return Item{}; }
  __device__ __forceinline__ void commit(uint32_t pos, uint32_t key, uint32_t val, int, const Item&) const {
    keys_out[pos] = key; vals_out[pos] = val;
  }
};
constexpr int kOutBatch = 4;

// One tile: rank, sort in shared memory, write out.  kFirst: the value of element i is i itself (nothing to read).
template <bool kFirst, class Out>
__global__ void __launch_bounds__(kThreads, 4) seg_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                        const int32_t* __restrict__ off, Plan p, int shift,
                                                        const uint32_t* __restrict__ hist, Out out) {
  __shared__ uint32_t skey[kTile], sval[kTile];
  __shared__ uint16_t wcnt[kWarps][kDigits];     // per warp: count, then exclusive offset inside the digit's run
  __shared__ uint32_t dstart[kDigits];           // first slot of each digit in the sorted tile
  __shared__ uint32_t gbase[kDigits];            // global position of that slot
  __shared__ uint32_t wtot[kDigits / 32];
  TileRef r;
  if (!locate_tile(p, off, r)) return;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (int k = tid; k < kWarps * kDigits / 2; k += kThreads) reinterpret_cast<uint32_t*>(&wcnt[0][0])[k] = 0u;
  if (tid < kDigits) gbase[tid] = hist[hist_index(r.tb, r.tc, r.t, tid)];
  uint32_t key[kItems];
#pragma unroll
  for (int i = 0; i < kItems; i++) {
    const int e = w * kWarpSpan + i * 32 + lane;
    key[i] = e < r.count ? keys_in[r.start + e] : 0u;
  }
  __syncthreads();
  // Rank inside the warp's span.  The peer mask (lanes of this round with the same digit) is built from one ballot per
  // digit bit -- MATCH.ANY has a latency of several hundred cycles here and the ballots of eight rounds pipeline
  // (eight pipelined MATCH.ANY per half were measured too: 0.73 ms per classification instead of 0.69).
  uint32_t rank2[kItems / 2];                    // two 16-bit ranks per register
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int half = 0; half < 2; half++) {
    uint32_t pm[kItems / 2];
#pragma unroll
    for (int j = 0; j < kItems / 2; j++) {
      const int i = half * (kItems / 2) + j;
      const bool valid = w * kWarpSpan + i * 32 + lane < r.count;
      const uint32_t d = (key[i] >> shift) & (kDigits - 1);
      uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
      for (int b = 0; b < kBits; b++) {
        const bool bit = (d >> b) & 1u;
        const uint32_t m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
      }
      pm[j] = peers;
    }
#pragma unroll
    for (int j = 0; j < kItems / 2; j++) {
      const int i = half * (kItems / 2) + j;
      const bool valid = w * kWarpSpan + i * 32 + lane < r.count;
      const uint32_t d = (key[i] >> shift) & (kDigits - 1);
      const uint32_t before = __popc(pm[j] & lt);
      const uint32_t old = valid ? (uint32_t)wcnt[w][d] : 0u;
      __syncwarp();
      if (valid && before == 0) wcnt[w][d] = (uint16_t)(old + __popc(pm[j]));   // lowest peer lane updates the counter
      __syncwarp();
      if (i & 1) rank2[i >> 1] |= (old + before) << 16; else rank2[i >> 1] = old + before;
    }
  }
  __syncthreads();
  uint32_t tot = 0, incl = 0;
  if (tid < kDigits) {
#pragma unroll
    for (int k = 0; k < kWarps; k++) { const uint32_t c = wcnt[k][tid]; wcnt[k][tid] = (uint16_t)tot; tot += c; }
    incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) wtot[w] = incl;
  }
  __syncthreads();
  if (tid < kDigits) {
    uint32_t base = 0;
#pragma unroll
    for (int k = 0; k < kDigits / 32; k++) base += k < w ? wtot[k] : 0u;
    dstart[tid] = base + incl - tot;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kItems; i++) {
    const int e = w * kWarpSpan + i * 32 + lane;
    if (e < r.count) {
      const uint32_t d = (key[i] >> shift) & (kDigits - 1);
      const uint32_t rk = (i & 1) ? (rank2[i >> 1] >> 16) : (rank2[i >> 1] & 0xFFFFu);
      const uint32_t slot = dstart[d] + wcnt[w][d] + rk;
      skey[slot] = key[i];
      sval[slot] = kFirst ? (uint32_t)(r.start + e) : vals_in[r.start + e];
    }
  }
  __syncthreads();
  for (int i0 = 0; i0 < kItems; i0 += kOutBatch) {
    uint32_t k[kOutBatch], v[kOutBatch], pos[kOutBatch];
    typename Out::Item item[kOutBatch];
#pragma unroll
    for (int j = 0; j < kOutBatch; j++) {
      const int slot = (i0 + j) * kThreads + tid;
      if (slot < r.count) {
        k[j] = skey[slot]; v[j] = sval[slot];
        const uint32_t d = (k[j] >> shift) & (kDigits - 1);
        pos[j] = gbase[d] + ((uint32_t)slot - dstart[d]);
        item[j] = out.fetch(pos[j], k[j], v[j], r.seg);
      }
    }
#pragma unroll
    for (int j = 0; j < kOutBatch; j++)
      if ((i0 + j) * kThreads + tid < r.count) out.commit(pos[j], k[j], v[j], r.seg, item[j]);
  }
}

}  // namespace ssort
