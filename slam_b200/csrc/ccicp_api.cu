// ccicp_api.cu -- host side of libccicp2d_b200.so: the C ABI declared in
// include/ccicp2d_b200.h.  CUDA runtime only (no torch, no CPU compute path).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <algorithm>
#include <cmath>
#include <vector>

#include <nvtx3/nvToolsExt.h>      // header-only; ranges show up when a profiler injects itself, no-ops otherwise

#include "ccicp_pair_main.cuh"
#include "ccicp_cluster.cuh"
#include "pose_math.h"
#include "classify.cuh"
#include "lum.cuh"
#include "lum_grid.cuh"

using namespace ccicp;

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

struct ccicp_ctx {
  DevBuf h_init, h_out;       // pinned staging for the small per-pair arrays when the caller's are pageable
  int device = 0;
  int sm_count = 0;
  int max_smem_optin = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr, kstream[4] = {nullptr, nullptr, nullptr, nullptr};   // chunked host path: H2D overlaps compute
  cudaEvent_t chunk_ev[16] = {};
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // h2d0,h2d1,k0,k1,d2h0,d2h1
  bool timed = false, timed_copies = false;
  bool chunk_ramp = true;     // CCICP_CHUNK_RAMP=0: equal chunks
  int chunk_pairs = 256;      // CCICP_CHUNK_PAIRS: 1081-beam pairs per H2D/compute chunk of the host-buffer call (C2 end to end: 128: 3.51,
                              // 192: 3.48, 256: 3.49, 384: 3.53, 512: 3.56, 768: 3.66 ms)
  int stage_kb = 48;          // CCICP_STAGE_KB: shared-memory staging of the large-scan tie replay; the rest of the SM's 256 KB is L1
                              // for the workspace (one C4 wave: 8..64 KB 10.0-10.1 ms, 128 KB 10.3, 200 KB 11.0)
  bool force_l2 = false;      // diagnostics: CCICP_FORCE_L2=1 routes every batch through the large-scan variants
  bool cluster = false;       // CCICP_CLUSTER=1: large scans run one thread-block cluster per pair (measured slower than the L2 variant on C4)
  void* timed_stream = nullptr;
  DevBuf counter, scratch, all_results, workspace, diag, retry_list;
  DevBuf d_model, d_moff, d_scene, d_soff, d_init, d_out, d_corr;
  DevBuf f_xy, f_off, f_nrm, f_curv;       // ccicp_scan_features
  DevBuf t_xy, t_off, t_mid, t_sid;        // scan table (ccicp_scans_upload) and the id lists of the last scan-pair call
  int32_t t_scans = 0, t_max_pts = 0, t_max_cells = 0;
  std::vector<int32_t> t_size;             // points per scan (host copy, argument checks)
  DevBuf c_keys, c_vals, c_keys_tmp, c_vals_tmp, c_hist, c_bitmap, c_tile_base, c_tile_seg, c_group_base;            // classification scratch
  DevBuf l_src, l_soff, l_tgt, l_toff, l_dist, l_idx, l_info, l_num, l_status, l_head, l_next, l_tabo;   // Lu-Milios edge information
  DevBuf c_xyz, c_off, c_out_xyz, c_out_ga, c_out_src, c_out_cnt;             // classification, host-buffer path
  ccicp_launch_info info{};
  std::string last_err;
  // chunked host path: the exact-order re-run is issued once per call, after the last chunk kernel (launch_retry)
  struct { IcpArgs a; size_t smem; int threads, occ, n_clusters; bool use_l2, use_cluster, valid; ccicp_result* out; } deferred{};
};

#ifndef CCICP_ITEMS_PER_CTA
#define CCICP_ITEMS_PER_CTA 32
#endif
namespace {

struct NvtxRange {       // one named range per C-ABI entry point (tracing hooks)
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

int fail_cuda(ccicp_ctx* ctx, cudaError_t e, const char* what) {
  if (ctx) ctx->last_err = std::string(what) + ": " + cudaGetErrorString(e);
  cudaGetLastError();
  return e == cudaErrorMemoryAllocation ? CCICP_E_NOMEM : CCICP_E_CUDA;
}
#define CK(call)                                              \
  do {                                                        \
    cudaError_t e_ = (call);                                  \
    if (e_ != cudaSuccess) return fail_cuda(ctx, e_, #call);  \
  } while (0)

int reserve(ccicp_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return CCICP_OK;
  if (b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; }
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) { b.p = nullptr; return fail_cuda(ctx, e, "cudaMalloc"); }
  b.cap = want;
  return CCICP_OK;
}

int reserve_pinned(ccicp_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return CCICP_OK;
  if (b.p) { cudaFreeHost(b.p); b.p = nullptr; b.cap = 0; }
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMallocHost(&b.p, want);
  if (e != cudaSuccess) { b.p = nullptr; return fail_cuda(ctx, e, "cudaMallocHost"); }
  b.cap = want;
  return CCICP_OK;
}

// true when an asynchronous copy from / to this host pointer does not block the calling thread
bool is_pinned(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

int round_up(int x, int m) { return (x + m - 1) / m * m; }

// The dynamic shared-memory limit is per-FUNCTION state of the device: it is raised to the opt-in maximum once per
// context creation, never per launch -- two host threads (one ctx each) launching pairs of different sizes would
// otherwise race on it (one thread lowering the limit under the other's launch: "invalid argument").
template <class K>
bool allow_max_smem(K kernel, int max_optin) {
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, kernel) != cudaSuccess) return false;
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin - (int)fa.sharedSizeBytes) == cudaSuccess;
}

// worst-case number of grid cells for a pair with at most max_model model points
int max_cells_for(int max_model) {
  // one candidate per step of grid_side(): the smaller class just above each threshold, the rest in the other class
  static const int lows[] = {0, 4, 33, 129, 513, 2049, 8193};
  int best = 0;
  for (int a : lows) {
    if (a > max_model) break;
    int b = max_model - a;
    int ga = grid_side(a), gb = grid_side(b);
    int c = ga * ga + gb * gb;
    if (c > best) best = c;
  }
  return best;
}

// Exact-order re-run of the problems [first, first + n_prob) whose default-mode record is flagged CCICP_ST_DEGENERATE:
// collect_retry_kernel lists them on the device, the exact instantiation pulls them from the list (no host round trip).
int launch_retry(ccicp_ctx* ctx, const IcpArgs& a, ccicp_result* kernel_out, size_t first, size_t n_prob, int32_t* cnt, bool may_reserve,
                 bool use_l2, bool use_cluster, int threads, size_t smem, int occ, int n_clusters, cudaStream_t st) {
  if (ctx->retry_list.cap < (first + n_prob) * 4) {
    if (!may_reserve) return CCICP_E_INVALID;      // the host-buffer call reserves the whole batch up front
    int rc = reserve(ctx, ctx->retry_list, (first + n_prob) * 4);
    if (rc) return rc;
  }
  int32_t* list = (int32_t*)ctx->retry_list.p + first;
  collect_retry_kernel<<<(unsigned)((n_prob + 255) / 256), 256, 0, st>>>(kernel_out, (int)first, (int)n_prob, list, cnt + 2);
  CK(cudaGetLastError());
  IcpArgs a2 = a;
  a2.item_list = list; a2.item_count = cnt + 2; a2.work_counter = cnt + 1;
  auto k2_on = icp_pair_exact_kernel;
  auto k2_l2 = icp_pair_l2_exact_kernel;
  int occ2 = 0;
  if (use_l2) {
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, k2_l2, threads, smem));
  } else {
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, k2_on, threads, smem));
  }
  if (occ2 < 1) return CCICP_E_TOO_LARGE;
  if (occ2 > occ) occ2 = occ;                   // the scratch and the L2 workspace are sized for `resident` CTAs
  long long res2 = (long long)occ2 * ctx->sm_count;
  if (use_cluster && res2 > n_clusters) res2 = n_clusters;     // the workspace holds one working set per CLUSTER of the main launch
  const int grid2 = (int)((long long)n_prob < res2 ? (long long)n_prob : res2);
  if (use_l2) k2_l2<<<grid2, threads, smem, st>>>(a2);
  else k2_on<<<grid2, threads, smem, st>>>(a2);
  CK(cudaGetLastError());
  ctx->info.n_launches += 2;
  return CCICP_OK;
}

int launch_icp(ccicp_ctx* ctx, int32_t pair_base, int32_t slot, int32_t n_pairs, int32_t n_seeds, const double* d_model_xy, const int32_t* d_model_off,
               const double* d_scene_xy, const int32_t* d_scene_off, const double* d_init_Rt,
               const ccicp_params* prm, ccicp_result* d_out, int32_t* d_corr, int32_t max_model, int32_t max_scene,
               int32_t max_cells, int64_t total_scene, cudaStream_t st, int n_slots = 1,
               const int32_t* d_model_id = nullptr, const int32_t* d_scene_id = nullptr, bool defer_retry = false) {
  // n_slots = launches of this call that may be in flight together (chunked host path: up to 4); each gets its own
  // tie-replay scratch and L2 workspace, sized by the first launch of the call
  if (slot == 0) ctx->info.n_launches = 0;
  if (n_pairs == 0) return CCICP_OK;
  if (max_model > CCICP_MAX_POINTS - 1 || max_scene > CCICP_MAX_POINTS) return CCICP_E_TOO_LARGE;
  IcpArgs a{};
  a.mcap = round_up(max_model < 32 ? 32 : max_model, 32);
  a.scap = round_up(max_scene < 32 ? 32 : max_scene, 32);
  a.ccap = max_cells > 0 ? max_cells : max_cells_for(max_model);
  if (a.ccap < 64) a.ccap = 64;
  // on-chip variant when the pair's working set fits the SM's shared memory, else the L2 variant
  const bool exact = (prm->flags & CCICP_F_EXACT_SUMS) != 0;
  const bool trim = (prm->flags & CCICP_F_TRIM) != 0;
  const bool diag = (prm->flags & CCICP_F_DIAG_COUNTS) != 0;
  const bool curv = (prm->flags & CCICP_F_CURV_GATE) != 0;
  if (curv && (exact || trim || diag || !(prm->curv_tol >= 0.0) || !(prm->feat_radius > 0.0) || prm->feat_k < 1)) return CCICP_E_INVALID;
  if (trim && (exact || !(prm->trim_frac > 0.0) || prm->trim_frac > 1.0)) return CCICP_E_INVALID;
  if (diag && (exact || trim)) return CCICP_E_INVALID;
  auto k_on = exact ? icp_pair_exact_kernel : trim ? icp_pair_trim_kernel : diag ? icp_pair_diag_kernel : curv ? icp_pair_curv_kernel : n_seeds > 1 ? icp_pair_seeds_kernel : icp_pair_kernel;
  auto k_l2 = exact ? icp_pair_l2_exact_kernel : trim ? icp_pair_l2_trim_kernel : diag ? icp_pair_l2_diag_kernel : curv ? icp_pair_l2_curv_kernel : icp_pair_l2_kernel;
  cudaFuncAttributes fa_on;
  CK(cudaFuncGetAttributes(&fa_on, k_on));
  // problems whose default-mode solve meets a (nearly) rank-deficient H are re-run by the exact-order instantiation
  const bool retry = !exact && !trim && !diag && !curv && !(prm->flags & CCICP_F_NO_EXACT_RETRY);
  if (retry) {
    cudaFuncAttributes fa_ex;
    CK(cudaFuncGetAttributes(&fa_ex, icp_pair_exact_kernel));
    if (fa_ex.sharedSizeBytes > fa_on.sharedSizeBytes) fa_on.sharedSizeBytes = fa_ex.sharedSizeBytes;
  }
  // CCICP_F_TRIM: one u32 per scene point behind the carved working set (this iteration's squared distances, the keys
  // of the radix select)
  // CCICP_F_CURV_GATE: one float per model slot and per scene point (their curvatures) in the same place
  const size_t warm_bytes = trim ? align_up((size_t)a.scap * 4, 128) : curv ? align_up(((size_t)a.mcap + a.scap) * 4, 128) : 0;
  size_t smem = workset_bytes(a.mcap, a.scap, a.ccap, false) + warm_bytes;
  const bool use_l2 = smem + fa_on.sharedSizeBytes > (size_t)ctx->max_smem_optin || ctx->force_l2;
  const int threads = use_l2 ? kThreadsL2 : kThreadsOnChip;
  // large scans: the one-CTA-per-pair L2 variant by default; with CCICP_CLUSTER=1 one thread-block cluster per pair
  // (working sets of ~148 / kCluster pairs stay L2-resident: 0.23 GB instead of 4.6 GB of DRAM reads per wave, but the
  // cluster barriers cost more than the DRAM latency they remove: 7.1 k against 10.5 k pairs/s on C4)
  const bool use_cluster = use_l2 && !exact && !trim && !diag && !curv && ctx->cluster;
  int occ = 0, n_clusters = 0;
  cudaLaunchConfig_t ccfg{};
  cudaLaunchAttribute cattr[1];
  if (use_cluster) {
    a.stage_cap = (ctx->stage_kb * 1024 / 6) & ~31;
    smem = (size_t)a.stage_cap * 6;
    cattr[0].id = cudaLaunchAttributeClusterDimension;
    cattr[0].val.clusterDim.x = kCluster; cattr[0].val.clusterDim.y = 1; cattr[0].val.clusterDim.z = 1;
    ccfg.gridDim = dim3((unsigned)(kCluster * ctx->sm_count), 1, 1); ccfg.blockDim = dim3(kThreadsCl, 1, 1);
    ccfg.dynamicSmemBytes = smem; ccfg.stream = st; ccfg.attrs = cattr; ccfg.numAttrs = 1;
    CK(cudaOccupancyMaxActiveClusters(&n_clusters, icp_pair_cluster_kernel, &ccfg));
    if (n_clusters < 1) return CCICP_E_TOO_LARGE;
    occ = 1;
  } else if (use_l2) {
    // dynamic shared memory of the L2 variant = staging arrays of the cooperative tie replay (float + u16 per element);
    // what is left of the SM's 256 KB stays L1 cache for the workspace
    a.stage_cap = (ctx->stage_kb * 1024 / 6) & ~31;
    smem = (size_t)a.stage_cap * 6;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_l2, threads, smem));
  } else {
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_on, threads, smem));
  }
  if (occ < 1) return CCICP_E_TOO_LARGE;
  const int resident = use_cluster ? n_clusters * kCluster : occ * ctx->sm_count;

  const bool best_of = (prm->flags & CCICP_F_BEST_OF_SEEDS) && n_seeds > 1;
  int items_per_pair = 1, seeds_per_item = n_seeds;
  if (n_seeds > 1) {
    // >= CCICP_ITEMS_PER_CTA items per resident CTA: a CTA runs a whole number of items, so with 4 per CTA (65536 problems of config C5 =
    // 2560 items of 26 seeds on 592 CTAs) the launch lasted 5 item times where 4.3 were needed; the grid build an item
    // repeats is ~4 % of ONE seed's work
    long long want = ((long long)CCICP_ITEMS_PER_CTA * resident + n_pairs - 1) / n_pairs;
    if (want < 1) want = 1;
    if (want > n_seeds) want = n_seeds;
    seeds_per_item = (int)((n_seeds + want - 1) / want);
    items_per_pair = (n_seeds + seeds_per_item - 1) / seeds_per_item;
  }
  const long long n_items = (long long)n_pairs * items_per_pair;
  int grid = (int)(n_items < resident ? n_items : resident);
  if (use_cluster) grid = kCluster * (int)(n_items < n_clusters ? n_items : n_clusters);

  ccicp_result* kernel_out = d_out;
  if (best_of) {
    int rc = reserve(ctx, ctx->all_results, (size_t)n_pairs * n_seeds * sizeof(ccicp_result));
    if (rc) return rc;
    kernel_out = (ccicp_result*)ctx->all_results.p;
  }
  // `slot` separates launches that may be in flight together (chunked host path): own work counter, own scratch
  int rc = reserve(ctx, ctx->counter, 17 * 16);      // 16 launch slots + one for a deferred re-run
  if (rc) return rc;
  const size_t per_cta = replay_scratch_bytes(max_model);
  const size_t per_slot = per_cta * (size_t)resident;
  if (ctx->scratch.cap < per_slot * n_slots) {
    if (slot != 0) return CCICP_E_INVALID;      // sized by the first launch of a call
    rc = reserve(ctx, ctx->scratch, per_slot * n_slots);
    if (rc) return rc;
  }

  if (use_l2) {
    a.workspace_per_cta = workset_bytes(a.mcap, a.scap, a.ccap, true) + warm_bytes + (use_cluster ? 128 : 0);   // per cluster there
    const size_t ws_slot = a.workspace_per_cta * (size_t)(use_cluster ? n_clusters : resident);
    if (ctx->workspace.cap < ws_slot * n_slots) {
      if (slot != 0) return CCICP_E_INVALID;
      rc = reserve(ctx, ctx->workspace, ws_slot * n_slots);
      if (rc) return rc;
    }
    a.workspace = (char*)ctx->workspace.p + (size_t)(slot % n_slots) * ws_slot;
  }
  a.model_xy = d_model_xy; a.model_off = d_model_off; a.scene_xy = d_scene_xy; a.scene_off = d_scene_off;
  a.init_Rt = d_init_Rt; a.out = kernel_out; a.corr_idx = d_corr;
  a.model_id = d_model_id; a.scene_id = d_scene_id;
  int32_t* cnt = (int32_t*)ctx->counter.p + 4 * (slot & 15);      // {work, retry work, retry count, -} of this slot
  a.work_counter = cnt;
  a.scratch = (char*)ctx->scratch.p + (size_t)(slot % n_slots) * per_slot; a.scratch_per_cta = per_cta;
  a.pair_base = pair_base;
  a.n_pairs = n_pairs; a.n_seeds = n_seeds; a.seeds_per_item = seeds_per_item; a.items_per_pair = items_per_pair;
  a.max_iter = prm->max_iter; a.flags = prm->flags; a.min_delta = prm->min_delta; a.indist = prm->indist;
  a.trim_frac = prm->trim_frac;
  a.curv_tol = prm->curv_tol; a.feat_radius = prm->feat_radius; a.feat_k = prm->feat_k;
  a.total_scene = total_scene; a.max_model = max_model;
  if (diag) {
    const bool fresh = ctx->diag.p == nullptr;
    rc = reserve(ctx, ctx->diag, 32);
    if (rc) return rc;
    if (fresh) CK(cudaMemsetAsync(ctx->diag.p, 0, 32, st));
    a.diag = (unsigned long long*)ctx->diag.p;
  }

  CK(cudaMemsetAsync(cnt, 0, 16, st));
  if (slot == 0) CK(cudaEventRecord(ctx->ev[2], st));
  if (use_cluster) {
    ccfg.gridDim = dim3((unsigned)grid, 1, 1);
    CK(cudaLaunchKernelEx(&ccfg, icp_pair_cluster_kernel, a));
  } else if (use_l2) k_l2<<<grid, threads, smem, st>>>(a);
  else k_on<<<grid, threads, smem, st>>>(a);
  CK(cudaGetLastError());
  ctx->info.n_launches += 1;
  ctx->deferred.valid = false;
  if (retry && defer_retry) {        // the caller issues ONE re-run for the whole call (launch_retry) after its last chunk
    ctx->deferred.a = a; ctx->deferred.smem = smem; ctx->deferred.threads = threads; ctx->deferred.occ = occ;
    ctx->deferred.n_clusters = n_clusters; ctx->deferred.use_l2 = use_l2; ctx->deferred.use_cluster = use_cluster;
    ctx->deferred.out = kernel_out; ctx->deferred.valid = true;
  } else if (retry) {
    rc = launch_retry(ctx, a, kernel_out, (size_t)pair_base * n_seeds, (size_t)n_pairs * n_seeds, cnt, slot == 0, use_l2, use_cluster,
                      threads, smem, occ, n_clusters, st);
    if (rc) return rc;
  }
  if (best_of) {
    const int threads = 128, warps_per_block = threads / 32;
    select_best_seed_kernel<<<(n_pairs + warps_per_block - 1) / warps_per_block, threads, 0, st>>>(
        kernel_out, d_out, d_scene_off + (d_scene_id ? 0 : 2 * (size_t)pair_base), d_scene_id ? d_scene_id + pair_base : nullptr, n_pairs,
        n_seeds, prm->min_corr_frac,
        (prm->flags & CCICP_F_FIRST_BELOW) ? prm->score_threshold : -1.0);
    CK(cudaGetLastError());
    ctx->info.n_launches += 1;
  }
  CK(cudaEventRecord(ctx->ev[3], st));
  ctx->timed = true; ctx->timed_stream = (void*)st;
  cudaFuncAttributes fa = fa_on;
  if (use_l2) CK(cudaFuncGetAttributes(&fa, k_l2));
  ctx->info.grid = grid; ctx->info.block = threads; ctx->info.smem_bytes = (int)(smem + fa.sharedSizeBytes);
  if (use_cluster) CK(cudaFuncGetAttributes(&fa, icp_pair_cluster_kernel));
  ctx->info.regs_per_thread = fa.numRegs; ctx->info.blocks_per_sm = occ; ctx->info.kernel_variant = use_cluster ? 4 : use_l2 ? 1 : 0;
  ctx->info.sm_count = ctx->sm_count;
  return CCICP_OK;
}

bool offsets_ok(const int32_t* off, int32_t n_pairs) {
  if (off[0] < 0) return false;
  for (int64_t i = 0; i < 2 * (int64_t)n_pairs; i++)
    if (off[i + 1] < off[i]) return false;
  return true;
}

}  // namespace

extern "C" {

int ccicp_abi_version(void) { return CCICP_ABI_VERSION; }

const char* ccicp_strerror(int code) {
  switch (code) {
    case CCICP_OK: return "ok";
    case CCICP_E_INVALID: return "invalid argument";
    case CCICP_E_NO_DEVICE: return "no usable CUDA device (this library has no CPU path)";
    case CCICP_E_CUDA: return "CUDA runtime error";
    case CCICP_E_NOMEM: return "out of memory";
    case CCICP_E_TOO_LARGE: return "scan too large for the on-chip pair kernel";
    case CCICP_E_ARCH: return "device is not compute capability 10.x (library is built for sm_100a only)";
    default: return "unknown error";
  }
}

const char* ccicp_last_cuda_error(const ccicp_ctx* ctx) { return ctx ? ctx->last_err.c_str() : ""; }

void ccicp_default_params(ccicp_params* p) {
  if (!p) return;
  p->max_iter = 20;          // icp.cpp:27
  p->flags = 0;
  p->min_delta = 1e-6;       // icp.cpp:27
  p->indist = 5.0;           // icpTools.cpp:188
  p->min_corr_frac = 0.5;
  p->score_threshold = 0.002;  // MAX_SCORE, global_match.cpp:30
  p->reject_dist = 0.0;        // off; graph_slam uses 10 m / 0.2 rad (graphSlamTools.h:35-36)
  p->reject_rot = 0.0;
  p->trim_frac = 1.0;          // off (CCICP_F_TRIM is an extra the reference does not have)
  p->curv_tol = 0.05;          // CCICP_F_CURV_GATE (an extra, off by default)
  p->feat_radius = 0.5; p->feat_k = 5; p->reserved = 0;
}

int ccicp_create(ccicp_ctx** out, int device) {
  if (!out) return CCICP_E_INVALID;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) { cudaGetLastError(); return CCICP_E_NO_DEVICE; }
  if (device < 0 || device >= n) return CCICP_E_NO_DEVICE;
  ccicp_ctx* ctx = new (std::nothrow) ccicp_ctx();
  if (!ctx) return CCICP_E_NOMEM;
  ctx->device = device;
  { const char* e = getenv("CCICP_FORCE_L2"); ctx->force_l2 = e && e[0] == '1'; }
  { const char* e = getenv("CCICP_CLUSTER"); ctx->cluster = e && e[0] == '1'; }
  { const char* e = getenv("CCICP_CHUNK_PAIRS"); if (e && atoi(e) >= 16) ctx->chunk_pairs = atoi(e); }
  { const char* e = getenv("CCICP_CHUNK_RAMP"); if (e) ctx->chunk_ramp = atoi(e) != 0; }
  { const char* e = getenv("CCICP_STAGE_KB"); if (e && atoi(e) >= 1 && atoi(e) <= 216) ctx->stage_kb = atoi(e); }
  cudaDeviceProp prop;
  if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
    delete ctx; cudaGetLastError(); return CCICP_E_CUDA;
  }
  if (prop.major != 10) { delete ctx; return CCICP_E_ARCH; }
  ctx->sm_count = prop.multiProcessorCount;
  ctx->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; cudaGetLastError(); return CCICP_E_CUDA; }
  for (auto& ev : ctx->ev)
    if (cudaEventCreate(&ev) != cudaSuccess) { delete ctx; cudaGetLastError(); return CCICP_E_CUDA; }
  bool ok = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
  for (auto& ks : ctx->kstream) ok = ok && cudaStreamCreateWithFlags(&ks, cudaStreamNonBlocking) == cudaSuccess;
  for (auto& ev : ctx->chunk_ev) ok = ok && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess;
  const int mo = ctx->max_smem_optin;
  ok = ok && allow_max_smem(icp_pair_kernel, mo) && allow_max_smem(icp_pair_l2_kernel, mo) && allow_max_smem(icp_pair_exact_kernel, mo) &&
       allow_max_smem(icp_pair_l2_exact_kernel, mo) && allow_max_smem(icp_pair_trim_kernel, mo) && allow_max_smem(icp_pair_l2_trim_kernel, mo) &&
       allow_max_smem(icp_pair_diag_kernel, mo) && allow_max_smem(icp_pair_l2_diag_kernel, mo) && allow_max_smem(icp_pair_curv_kernel, mo) &&
       allow_max_smem(icp_pair_l2_curv_kernel, mo) && allow_max_smem(icp_pair_cluster_kernel, mo) && allow_max_smem(icp_pair_seeds_kernel, mo);
  if (!ok) { delete ctx; cudaGetLastError(); return CCICP_E_CUDA; }
  *out = ctx;
  return CCICP_OK;
}

int ccicp_destroy(ccicp_ctx* ctx) {
  if (!ctx) return CCICP_OK;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  DevBuf* bufs[] = {&ctx->f_xy, &ctx->f_off, &ctx->f_nrm, &ctx->f_curv, &ctx->t_xy, &ctx->t_off, &ctx->t_mid, &ctx->t_sid, &ctx->retry_list, &ctx->diag, &ctx->counter, &ctx->scratch, &ctx->all_results, &ctx->workspace, &ctx->d_model, &ctx->d_moff, &ctx->d_scene,
                    &ctx->d_soff, &ctx->d_init, &ctx->d_out, &ctx->d_corr, &ctx->c_keys, &ctx->c_vals, &ctx->c_keys_tmp,
                    &ctx->c_vals_tmp, &ctx->c_hist, &ctx->c_bitmap, &ctx->c_tile_base, &ctx->c_tile_seg, &ctx->c_group_base, &ctx->c_xyz, &ctx->c_off, &ctx->c_out_xyz,
                    &ctx->c_out_ga, &ctx->c_out_src, &ctx->c_out_cnt, &ctx->l_src, &ctx->l_soff, &ctx->l_tgt, &ctx->l_toff,
                    &ctx->l_dist, &ctx->l_idx, &ctx->l_info, &ctx->l_num, &ctx->l_status, &ctx->l_head, &ctx->l_next, &ctx->l_tabo};
  for (DevBuf* b : bufs) if (b->p) cudaFree(b->p);
  if (ctx->h_init.p) cudaFreeHost(ctx->h_init.p);
  if (ctx->h_out.p) cudaFreeHost(ctx->h_out.p);
  for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
  for (auto& ev : ctx->chunk_ev) if (ev) cudaEventDestroy(ev);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  for (auto& ks : ctx->kstream) if (ks) cudaStreamDestroy(ks);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return CCICP_OK;
}

int32_t ccicp_grid_cells(int32_t n_class_points) {
  const int g = grid_side(n_class_points);
  return g * g;
}

int ccicp_register_batch_device(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds, const double* d_model_xy,
                                const int32_t* d_model_off, const double* d_scene_xy, const int32_t* d_scene_off,
                                const double* d_init_Rt, const ccicp_params* params, ccicp_result* d_out,
                                int32_t* d_corr_idx, int32_t max_model, int32_t max_scene, void* stream) {
  return ccicp_register_batch_device_ex(ctx, n_pairs, n_seeds, d_model_xy, d_model_off, d_scene_xy, d_scene_off, d_init_Rt,
                                        params, d_out, d_corr_idx, max_model, max_scene, 0, stream);
}

int ccicp_register_batch_device_ex(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds, const double* d_model_xy,
                                   const int32_t* d_model_off, const double* d_scene_xy, const int32_t* d_scene_off,
                                   const double* d_init_Rt, const ccicp_params* params, ccicp_result* d_out,
                                   int32_t* d_corr_idx, int32_t max_model, int32_t max_scene, int32_t max_cells_hint,
                                   void* stream) {
  NvtxRange nvtx_("ccicp_register_batch_device_ex");
  if (!ctx || !params || n_pairs < 0 || n_seeds < 1 || max_model < 0 || max_scene < 0 || max_cells_hint < 0) return CCICP_E_INVALID;
  if (n_pairs > 0 && (!d_model_xy || !d_model_off || !d_scene_xy || !d_scene_off || !d_init_Rt || !d_out)) return CCICP_E_INVALID;
  if (d_corr_idx && n_seeds > 1 && (params->flags & CCICP_F_BEST_OF_SEEDS)) return CCICP_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  int64_t total_scene = 0;
  if (d_corr_idx && n_seeds > 1 && n_pairs > 0) {   // seed stride of corr_idx: needs the batch total
    int32_t last = 0;
    CK(cudaMemcpyAsync(&last, d_scene_off + 2 * (size_t)n_pairs, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    total_scene = last;
  }
  ctx->timed_copies = false;
  return launch_icp(ctx, 0, 0, n_pairs, n_seeds, d_model_xy, d_model_off, d_scene_xy, d_scene_off, d_init_Rt, params, d_out,
                    d_corr_idx, max_model, max_scene, max_cells_hint, total_scene, st);
}

int ccicp_stream_sync(ccicp_ctx* ctx, void* stream) {
  if (!ctx) return CCICP_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(stream ? (cudaStream_t)stream : ctx->stream));
  return CCICP_OK;
}

int ccicp_register_batch(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds, const double* model_xy,
                         const int32_t* model_off, const double* scene_xy, const int32_t* scene_off,
                         const double* init_Rt, const ccicp_params* params, ccicp_result* out, int32_t* corr_idx) {
  NvtxRange nvtx_("ccicp_register_batch");
  if (!ctx || !params || n_pairs < 0 || n_seeds < 1) return CCICP_E_INVALID;
  if (n_pairs == 0) return CCICP_OK;
  if (!model_off || !scene_off || !init_Rt || !out) return CCICP_E_INVALID;
  if (!offsets_ok(model_off, n_pairs) || !offsets_ok(scene_off, n_pairs)) return CCICP_E_INVALID;
  const bool best_of = (params->flags & CCICP_F_BEST_OF_SEEDS) && n_seeds > 1;
  if (corr_idx && best_of) return CCICP_E_INVALID;
  const int64_t m_base = model_off[0], s_base = scene_off[0];
  const int64_t n_model = model_off[2 * (size_t)n_pairs] - m_base, n_scene = scene_off[2 * (size_t)n_pairs] - s_base;
  if ((n_model > 0 && !model_xy) || (n_scene > 0 && !scene_xy)) return CCICP_E_INVALID;
  if (m_base != 0 || s_base != 0) return CCICP_E_INVALID;   // offsets are absolute into the arrays passed
  int max_model = 0, max_scene = 0, max_cells = 0;
  for (int32_t p = 0; p < n_pairs; p++) {
    const int a = model_off[2 * p + 1] - model_off[2 * p], b = model_off[2 * p + 2] - model_off[2 * p + 1];
    const int c = scene_off[2 * p + 2] - scene_off[2 * p];
    if (a + b > max_model) max_model = a + b;
    if (c > max_scene) max_scene = c;
    const int ga = grid_side(a), gb = grid_side(b);
    if (ga * ga + gb * gb > max_cells) max_cells = ga * ga + gb * gb;
  }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const size_t n_prob = (size_t)n_pairs * n_seeds, n_out = best_of ? (size_t)n_pairs : n_prob;
  const size_t off_bytes = (2 * (size_t)n_pairs + 1) * 4;
  int rc;
  if ((rc = reserve(ctx, ctx->d_model, (size_t)n_model * 16 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->d_scene, (size_t)n_scene * 16 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->d_moff, off_bytes))) return rc;
  if ((rc = reserve(ctx, ctx->d_soff, off_bytes))) return rc;
  if ((rc = reserve(ctx, ctx->d_init, n_prob * 48))) return rc;
  if ((rc = reserve(ctx, ctx->d_out, n_out * sizeof(ccicp_result)))) return rc;
  if (corr_idx && (rc = reserve(ctx, ctx->d_corr, (size_t)n_scene * n_seeds * 4 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->retry_list, n_prob * 4 + 16))) return rc;

  // Chunked pipeline: the H2D copy of chunk c+1 (copy stream) overlaps the kernel of chunk c; consecutive
  // chunks run on two compute streams so the tail of one kernel overlaps the head of the next.  The offsets
  // are absolute, so every chunk's kernel indexes the same device arrays with a different pair_base.
  int n_chunks = 1;
  int32_t bound[17] = {0};
  bound[1] = n_pairs;
  // chunk_pairs is meant for 1081-beam pairs (34.8 KB each): larger pairs get proportionally fewer per chunk, so that the
  // first kernel still starts after a short copy (C4's 1 MB pairs used to wait 10 ms for a first chunk of 512 pairs)
  int cp = ctx->chunk_pairs;
  if (n_pairs > 0) {
    const double bytes_per_pair = ((double)n_model + (double)n_scene) * 16.0 / (double)n_pairs;
    const double scaled = (double)ctx->chunk_pairs * 34816.0 / (bytes_per_pair > 1.0 ? bytes_per_pair : 1.0);
    if (scaled < (double)cp) cp = scaled < 16.0 ? 16 : (int)scaled;
  }
  if (n_seeds == 1 && !corr_idx && n_pairs >= 2 * cp) {
    // up to 4 chunk kernels are in flight (4 streams, 4 scratch slots).  The first chunks are short so that the first
    // kernel starts after a short copy; the copies outpace the kernels, so the GPU is full a few chunks later.
    n_chunks = 0;
    int32_t pos = 0;
    if (n_pairs >= 4 * cp && ctx->chunk_ramp)
      for (int sz : {cp / 6, cp / 3, 2 * cp / 3}) { if (sz >= 16) { pos += sz; bound[++n_chunks] = pos; } }
    int rest = (n_pairs - pos) / cp;
    if (rest > 16 - n_chunks) rest = 16 - n_chunks;
    if (rest < 1) rest = 1;
    for (int c = 1; c <= rest; c++) bound[n_chunks + c] = pos + (int32_t)((int64_t)(n_pairs - pos) * c / rest);
    n_chunks += rest;
  }
  // A pageable `out` would make every chunk's record copy wait on the host for that chunk's kernel (and a pageable
  // init array every chunk's upload), which serialises the pipeline: those two small arrays go through pinned staging.
  const double* init_src = init_Rt;
  ccicp_result* out_dst = out;
  if (n_chunks > 1) {
    if (!is_pinned(init_Rt)) {
      if ((rc = reserve_pinned(ctx, ctx->h_init, n_prob * 48))) return rc;
      memcpy(ctx->h_init.p, init_Rt, n_prob * 48);
      init_src = (const double*)ctx->h_init.p;
    }
    if (!is_pinned(out)) {
      if ((rc = reserve_pinned(ctx, ctx->h_out, n_out * sizeof(ccicp_result)))) return rc;
      out_dst = (ccicp_result*)ctx->h_out.p;
    }
  }
  cudaStream_t cs = ctx->copy_stream;
  CK(cudaEventRecord(ctx->ev[0], cs));
  CK(cudaMemcpyAsync(ctx->d_moff.p, model_off, off_bytes, cudaMemcpyHostToDevice, cs));
  CK(cudaMemcpyAsync(ctx->d_soff.p, scene_off, off_bytes, cudaMemcpyHostToDevice, cs));
  for (int c = 0; c < n_chunks; c++) {
    const int32_t p0 = bound[c], p1 = bound[c + 1];
    const int64_t m0 = model_off[2 * (size_t)p0], m1 = model_off[2 * (size_t)p1];
    const int64_t s0 = scene_off[2 * (size_t)p0], s1 = scene_off[2 * (size_t)p1];
    if (m1 > m0) CK(cudaMemcpyAsync((char*)ctx->d_model.p + m0 * 16, model_xy + 2 * m0, (size_t)(m1 - m0) * 16, cudaMemcpyHostToDevice, cs));
    if (s1 > s0) CK(cudaMemcpyAsync((char*)ctx->d_scene.p + s0 * 16, scene_xy + 2 * s0, (size_t)(s1 - s0) * 16, cudaMemcpyHostToDevice, cs));
    CK(cudaMemcpyAsync((char*)ctx->d_init.p + (size_t)p0 * n_seeds * 48, init_src + (size_t)p0 * n_seeds * 6,
                       (size_t)(p1 - p0) * n_seeds * 48, cudaMemcpyHostToDevice, cs));
    CK(cudaEventRecord(ctx->chunk_ev[c], cs));
    if (c == n_chunks - 1) CK(cudaEventRecord(ctx->ev[1], cs));
    cudaStream_t ks = n_chunks == 1 ? st : ctx->kstream[c & 3];
    CK(cudaStreamWaitEvent(ks, ctx->chunk_ev[c], 0));
    // several chunks: the exact-order re-run and the record copy are issued ONCE, after the last chunk kernel (two tiny
    // launches and a copy per chunk sat between the chunk kernels of a stream and cost 3 % of the end-to-end rate)
    const bool defer = n_chunks > 1;
    rc = launch_icp(ctx, p0, c, p1 - p0, n_seeds, (const double*)ctx->d_model.p, (const int32_t*)ctx->d_moff.p,
                    (const double*)ctx->d_scene.p, (const int32_t*)ctx->d_soff.p, (const double*)ctx->d_init.p, params,
                    (ccicp_result*)ctx->d_out.p, corr_idx ? (int32_t*)ctx->d_corr.p : nullptr, max_model, max_scene,
                    max_cells, n_scene, ks, n_chunks < 4 ? n_chunks : 4, nullptr, nullptr, defer);
    if (rc) { cudaDeviceSynchronize(); return rc; }
    if (defer) {
      CK(cudaEventRecord(ctx->chunk_ev[c], ks));          // (its H2D use is over: the kernel launch above waited on it)
      CK(cudaStreamWaitEvent(st, ctx->chunk_ev[c], 0));
      continue;
    }
    CK(cudaEventRecord(ctx->ev[4], ks));
    const size_t o0 = best_of ? (size_t)p0 : (size_t)p0 * n_seeds, o1 = best_of ? (size_t)p1 : (size_t)p1 * n_seeds;
    CK(cudaMemcpyAsync(out_dst + o0, (char*)ctx->d_out.p + o0 * sizeof(ccicp_result), (o1 - o0) * sizeof(ccicp_result),
                       cudaMemcpyDeviceToHost, ks));
    if (corr_idx && n_scene) CK(cudaMemcpyAsync(corr_idx, ctx->d_corr.p, (size_t)n_scene * n_seeds * 4, cudaMemcpyDeviceToHost, ks));
    CK(cudaEventRecord(ctx->ev[5], ks));
  }
  if (n_chunks > 1) {                 // st has waited for every chunk kernel
    const int launched = ctx->info.n_launches;
    if (ctx->deferred.valid) {
      int32_t* cnt = (int32_t*)ctx->counter.p + 4 * 16;
      CK(cudaMemsetAsync(cnt, 0, 16, st));
      rc = launch_retry(ctx, ctx->deferred.a, ctx->deferred.out, 0, n_prob, cnt, false, ctx->deferred.use_l2, ctx->deferred.use_cluster,
                        ctx->deferred.threads, ctx->deferred.smem, ctx->deferred.occ, ctx->deferred.n_clusters, st);
      if (rc) { cudaDeviceSynchronize(); return rc; }
    }
    (void)launched;
    CK(cudaEventRecord(ctx->ev[3], st));                // kernel time of the call: first chunk kernel .. end of the re-run
    CK(cudaEventRecord(ctx->ev[4], st));
    CK(cudaMemcpyAsync(out_dst, ctx->d_out.p, n_out * sizeof(ccicp_result), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(ctx->ev[5], st));
    ctx->timed_stream = (void*)st;
  }
  CK(cudaStreamSynchronize(cs));
  CK(cudaStreamSynchronize(st));
  if (out_dst != out) memcpy(out, out_dst, n_out * sizeof(ccicp_result));
  ctx->timed_copies = true;
  return CCICP_OK;
}

// ---- per-point features (north_star extra) -------------------------------------------------------------------------
int ccicp_scan_features(ccicp_ctx* ctx, int32_t n_clouds, const double* xy, const int32_t* off, int32_t feat_k,
                        double feat_radius, float* normal_xy, float* curvature) {
  NvtxRange nvtx_("ccicp_scan_features");
  if (!ctx || n_clouds < 0 || feat_k < 1 || !(feat_radius > 0.0)) return CCICP_E_INVALID;
  if (n_clouds == 0) return CCICP_OK;
  if (!off || off[0] != 0 || !offsets_ok(off, n_clouds) || !curvature) return CCICP_E_INVALID;
  const int64_t n = off[2 * (size_t)n_clouds];
  if (n == 0) return CCICP_OK;
  if (!xy) return CCICP_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int rc;
  const size_t off_bytes = (2 * (size_t)n_clouds + 1) * 4;
  if ((rc = reserve(ctx, ctx->f_xy, (size_t)n * 16))) return rc;
  if ((rc = reserve(ctx, ctx->f_off, off_bytes))) return rc;
  if ((rc = reserve(ctx, ctx->f_nrm, (size_t)n * 8))) return rc;
  if ((rc = reserve(ctx, ctx->f_curv, (size_t)n * 4))) return rc;
  CK(cudaMemcpyAsync(ctx->f_xy.p, xy, (size_t)n * 16, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->f_off.p, off, off_bytes, cudaMemcpyHostToDevice, st));
  CK(cudaEventRecord(ctx->ev[2], st));
  feat::scan_features_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const double2*)ctx->f_xy.p, (const int32_t*)ctx->f_off.p,
                                                                          2 * n_clouds, (int)n, feat_k, feat_radius * feat_radius,
                                                                          normal_xy ? (float2*)ctx->f_nrm.p : nullptr, (float*)ctx->f_curv.p);
  CK(cudaGetLastError());
  CK(cudaEventRecord(ctx->ev[3], st));
  ctx->timed = true; ctx->timed_copies = false; ctx->timed_stream = (void*)st;
  ctx->info.n_launches = 1; ctx->info.kernel_variant = 5; ctx->info.block = 256; ctx->info.grid = (int)((n + 255) / 256);
  if (normal_xy) CK(cudaMemcpyAsync(normal_xy, ctx->f_nrm.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(curvature, ctx->f_curv.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return CCICP_OK;
}

// ---- scan table: key-frame scans uploaded once, pairs as (model scan, scene scan) index lists ----------------------
int ccicp_scans_upload(ccicp_ctx* ctx, int32_t n_scans, const double* scan_xy, const int32_t* scan_off) {
  NvtxRange nvtx_("ccicp_scans_upload");
  if (!ctx || n_scans < 0) return CCICP_E_INVALID;
  ctx->t_scans = 0; ctx->t_max_pts = 0; ctx->t_max_cells = 0; ctx->t_size.clear();
  if (n_scans == 0) return CCICP_OK;
  if (!scan_off || scan_off[0] != 0 || !offsets_ok(scan_off, n_scans)) return CCICP_E_INVALID;
  const int64_t n_pts = scan_off[2 * (size_t)n_scans];
  if (n_pts > 0 && !scan_xy) return CCICP_E_INVALID;
  try { ctx->t_size.resize((size_t)n_scans); } catch (const std::bad_alloc&) { return CCICP_E_NOMEM; }
  int max_pts = 0, max_cells = 0;
  for (int32_t k = 0; k < n_scans; k++) {
    const int a = scan_off[2 * k + 1] - scan_off[2 * k], b = scan_off[2 * k + 2] - scan_off[2 * k + 1];
    ctx->t_size[k] = a + b;
    if (a + b > max_pts) max_pts = a + b;
    const int ga = grid_side(a), gb = grid_side(b);
    if (ga * ga + gb * gb > max_cells) max_cells = ga * ga + gb * gb;
  }
  if (max_pts > CCICP_MAX_POINTS - 1) return CCICP_E_TOO_LARGE;
  CK(cudaSetDevice(ctx->device));
  int rc;
  const size_t off_bytes = (2 * (size_t)n_scans + 1) * 4;
  if ((rc = reserve(ctx, ctx->t_xy, (size_t)n_pts * 16 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->t_off, off_bytes))) return rc;
  if (n_pts) CK(cudaMemcpyAsync(ctx->t_xy.p, scan_xy, (size_t)n_pts * 16, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->t_off.p, scan_off, off_bytes, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->t_scans = n_scans; ctx->t_max_pts = max_pts; ctx->t_max_cells = max_cells;
  return CCICP_OK;
}

int ccicp_register_scan_pairs(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds, const int32_t* model_scan,
                              const int32_t* scene_scan, const double* init_Rt, const ccicp_params* params, ccicp_result* out) {
  NvtxRange nvtx_("ccicp_register_scan_pairs");
  if (!ctx || !params || n_pairs < 0 || n_seeds < 1) return CCICP_E_INVALID;
  if (n_pairs == 0) return CCICP_OK;
  if (!model_scan || !scene_scan || !init_Rt || !out || ctx->t_scans <= 0) return CCICP_E_INVALID;
  for (int32_t p = 0; p < n_pairs; p++)
    if (model_scan[p] < 0 || model_scan[p] >= ctx->t_scans || scene_scan[p] < 0 || scene_scan[p] >= ctx->t_scans) return CCICP_E_INVALID;
  const bool best_of = (params->flags & CCICP_F_BEST_OF_SEEDS) && n_seeds > 1;
  const size_t n_prob = (size_t)n_pairs * n_seeds, n_out = best_of ? (size_t)n_pairs : n_prob;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int rc;
  if ((rc = reserve(ctx, ctx->t_mid, (size_t)n_pairs * 4))) return rc;
  if ((rc = reserve(ctx, ctx->t_sid, (size_t)n_pairs * 4))) return rc;
  if ((rc = reserve(ctx, ctx->d_init, n_prob * 48))) return rc;
  if ((rc = reserve(ctx, ctx->d_out, n_out * sizeof(ccicp_result)))) return rc;
  CK(cudaEventRecord(ctx->ev[0], st));
  CK(cudaMemcpyAsync(ctx->t_mid.p, model_scan, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->t_sid.p, scene_scan, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_init.p, init_Rt, n_prob * 48, cudaMemcpyHostToDevice, st));
  CK(cudaEventRecord(ctx->ev[1], st));
  rc = launch_icp(ctx, 0, 0, n_pairs, n_seeds, (const double*)ctx->t_xy.p, (const int32_t*)ctx->t_off.p, (const double*)ctx->t_xy.p,
                  (const int32_t*)ctx->t_off.p, (const double*)ctx->d_init.p, params, (ccicp_result*)ctx->d_out.p, nullptr,
                  ctx->t_max_pts, ctx->t_max_pts, ctx->t_max_cells, 0, st, 1, (const int32_t*)ctx->t_mid.p, (const int32_t*)ctx->t_sid.p);
  if (rc) { cudaDeviceSynchronize(); return rc; }
  CK(cudaEventRecord(ctx->ev[4], st));
  CK(cudaMemcpyAsync(out, ctx->d_out.p, n_out * sizeof(ccicp_result), cudaMemcpyDeviceToHost, st));
  CK(cudaEventRecord(ctx->ev[5], st));
  CK(cudaStreamSynchronize(st));
  ctx->timed_copies = true;
  return CCICP_OK;
}

void ccicp_pose_to_rt(const ccicp_pose* pose, double* Rt6) {
  if (pose && Rt6) ccicp_pose_math::pose_to_Rt(*pose, Rt6);
}

void ccicp_rt_to_pose(const double* Rt6, const ccicp_pose* init_pose, ccicp_pose* out) {
  if (Rt6 && init_pose && out) ccicp_pose_math::Rt_to_pose(Rt6, *init_pose, out);
}

// Batched CCICP::doICPMatch(vectors..., initPose), icpTools.cpp:123-218.
int ccicp_match_batch(ccicp_ctx* ctx, int32_t n_pairs, const double* model_xy, const int32_t* model_off,
                      const double* scene_xy, const int32_t* scene_off, const ccicp_pose* init_pose,
                      const ccicp_params* params, ccicp_pose* out_pose, ccicp_result* records) {
  NvtxRange nvtx_("ccicp_match_batch");
  if (!ctx || !params || n_pairs < 0) return CCICP_E_INVALID;
  if (n_pairs == 0) return CCICP_OK;
  if (!init_pose || !out_pose || !scene_off) return CCICP_E_INVALID;
  std::vector<double> init;
  std::vector<ccicp_result> tmp;
  try {
    init.resize((size_t)n_pairs * 6);
    if (!records) tmp.resize((size_t)n_pairs);
  } catch (const std::bad_alloc&) { return CCICP_E_NOMEM; }
  for (int32_t p = 0; p < n_pairs; p++) ccicp_pose_math::pose_to_Rt(init_pose[p], &init[(size_t)p * 6]);
  ccicp_result* rec = records ? records : tmp.data();
  const int rc = ccicp_register_batch(ctx, n_pairs, 1, model_xy, model_off, scene_xy, scene_off, init.data(), params, rec,
                                      nullptr);
  if (rc != CCICP_OK) return rc;
  for (int32_t p = 0; p < n_pairs; p++) {
    // the scene check comes first in the reference, before the Icp object exists (icpTools.cpp:179-187)
    if (scene_off[2 * (size_t)p + 2] - scene_off[2 * (size_t)p] < 5) { ccicp_pose_math::sentinel(&out_pose[p]); continue; }
    const double Rt[6] = {rec[p].R[0], rec[p].R[1], rec[p].R[2], rec[p].R[3], rec[p].t[0], rec[p].t[1]};
    ccicp_pose_math::Rt_to_pose(Rt, init_pose[p], &out_pose[p]);
    if (params->reject_dist > 0.0) {               // graphSlamTools.cpp:332-362
      const double x_diff = std::fabs(init_pose[p].px - out_pose[p].px), y_diff = std::fabs(init_pose[p].py - out_pose[p].py);
      double theta_diff = std::fabs(ccicp_pose_math::euler_ypr(init_pose[p].qx, init_pose[p].qy, init_pose[p].qz, init_pose[p].qw).yaw -
                                    ccicp_pose_math::euler_ypr(out_pose[p].qx, out_pose[p].qy, out_pose[p].qz, out_pose[p].qw).yaw);
      if (theta_diff > 2 * M_PI) theta_diff = theta_diff - 2 * M_PI;
      else if (theta_diff > M_PI) theta_diff = 2 * M_PI - theta_diff;
      if (x_diff > params->reject_dist || y_diff > params->reject_dist || theta_diff > params->reject_rot)
        rec[p].status |= CCICP_ST_EDGE_REJECTED;
    }
  }
  return CCICP_OK;
}

// graphSlamGetKNN (graphSlamTools.cpp:72-106)
int32_t ccicp_knn_candidates(const double* nodes_xy, int32_t n_nodes, const double* query_xy, int32_t query_idx, int32_t K,
                             int32_t* out) {
  if (!nodes_xy || !query_xy || !out || n_nodes <= 0 || K <= 0) return 0;
  K = std::min(n_nodes - 1, K);
  std::vector<std::pair<double, int32_t>> sc;
  for (int32_t i = 0; i < n_nodes - 2; i++) {
    const double dx = nodes_xy[2 * i] - query_xy[0], dy = nodes_xy[2 * i + 1] - query_xy[1];
    sc.emplace_back(std::sqrt(dx * dx + dy * dy), i);
  }
  std::stable_sort(sc.begin(), sc.end(), [](const std::pair<double, int32_t>& a, const std::pair<double, int32_t>& b) { return a.first < b.first; });
  int32_t n = 0;
  for (int32_t i = 0; i < K; i++)
    if (i < (int32_t)sc.size() && query_idx != sc[i].second) out[n++] = sc[i].second;
  return n;
}

// GA / NGA classification (CCICP::classifyPoints, icpTools.cpp:36-103), device buffers.
int ccicp_classify_points_device(ccicp_ctx* ctx, int32_t n_clouds, const float* d_xyz, const int32_t* d_cloud_off,
                                 int64_t n_points, float* d_out_xyz, uint8_t* d_out_ground_adj, int32_t* d_out_src_index,
                                 int32_t* d_out_count, void* stream) {
  NvtxRange nvtx_("ccicp_classify_points_device");
  if (!ctx || n_clouds < 0 || n_clouds > classify::kMaxClouds || n_points < 0 || n_points > 0xFFFFFFFFll) return CCICP_E_INVALID;
  if (n_clouds == 0) return CCICP_OK;
  if (!d_cloud_off || !d_out_count || (n_points > 0 && (!d_xyz || !d_out_xyz || !d_out_ground_adj))) return CCICP_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  int rc;
  const size_t nb = (size_t)(n_points > 0 ? n_points : 1) * 4;
  if ((rc = reserve(ctx, ctx->c_keys, nb))) return rc;
  if ((rc = reserve(ctx, ctx->c_vals, nb))) return rc;
  if ((rc = reserve(ctx, ctx->c_keys_tmp, nb))) return rc;
  if ((rc = reserve(ctx, ctx->c_vals_tmp, nb))) return rc;
  if ((rc = reserve(ctx, ctx->c_hist, classify::hist_words(n_points, n_clouds) * 4))) return rc;
  if ((rc = reserve(ctx, ctx->c_tile_base, ((size_t)n_clouds + 1) * 4))) return rc;
  if ((rc = reserve(ctx, ctx->c_group_base, ((size_t)n_clouds + 1) * 4))) return rc;
  if ((rc = reserve(ctx, ctx->c_tile_seg, ((size_t)ssort::max_tiles(n_points, n_clouds) + 1) * 4))) return rc;
  if ((rc = reserve(ctx, ctx->c_bitmap, (size_t)n_clouds * classify::kWordsPerCloud * 4))) return rc;
  classify::Scratch sc{(uint32_t*)ctx->c_keys.p, (uint32_t*)ctx->c_vals.p, (uint32_t*)ctx->c_keys_tmp.p,
                       (uint32_t*)ctx->c_vals_tmp.p, (uint32_t*)ctx->c_hist.p, (uint32_t*)ctx->c_bitmap.p,
                       (int32_t*)ctx->c_tile_base.p, (int32_t*)ctx->c_tile_seg.p, (int32_t*)ctx->c_group_base.p};
  CK(cudaEventRecord(ctx->ev[2], st));
  ctx->info.n_launches = classify::run(d_xyz, d_cloud_off, n_clouds, n_points, sc, d_out_xyz, d_out_ground_adj,
                                       d_out_src_index, d_out_count, st);
  CK(cudaGetLastError());
  CK(cudaEventRecord(ctx->ev[3], st));
  ctx->timed = true; ctx->timed_copies = false; ctx->timed_stream = (void*)st;
  ctx->info.kernel_variant = 2; ctx->info.block = 256; ctx->info.grid = ssort::max_tiles(n_points, n_clouds);
  return CCICP_OK;
}

int ccicp_classify_points(ccicp_ctx* ctx, int32_t n_clouds, const float* xyz, const int32_t* cloud_off, float* out_xyz,
                          uint8_t* out_ground_adj, int32_t* out_src_index, int32_t* out_count) {
  if (!ctx || n_clouds < 0 || n_clouds > classify::kMaxClouds) return CCICP_E_INVALID;
  if (n_clouds == 0) return CCICP_OK;
  if (!cloud_off || !out_count || cloud_off[0] != 0) return CCICP_E_INVALID;
  for (int32_t c = 0; c < n_clouds; c++) if (cloud_off[c + 1] < cloud_off[c]) return CCICP_E_INVALID;
  const int64_t n = cloud_off[n_clouds];
  if (n > 0 && (!xyz || !out_xyz || !out_ground_adj)) return CCICP_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int rc;
  const size_t np = (size_t)(n > 0 ? n : 1);
  if ((rc = reserve(ctx, ctx->c_xyz, np * 12))) return rc;
  if ((rc = reserve(ctx, ctx->c_off, ((size_t)n_clouds + 1) * 4))) return rc;
  if ((rc = reserve(ctx, ctx->c_out_xyz, np * 12))) return rc;
  if ((rc = reserve(ctx, ctx->c_out_ga, np))) return rc;
  if ((rc = reserve(ctx, ctx->c_out_src, np * 4))) return rc;
  if ((rc = reserve(ctx, ctx->c_out_cnt, (size_t)n_clouds * 4))) return rc;
  if (n > 0) CK(cudaMemcpyAsync(ctx->c_xyz.p, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->c_off.p, cloud_off, ((size_t)n_clouds + 1) * 4, cudaMemcpyHostToDevice, st));
  rc = ccicp_classify_points_device(ctx, n_clouds, (const float*)ctx->c_xyz.p, (const int32_t*)ctx->c_off.p, n,
                                    (float*)ctx->c_out_xyz.p, (uint8_t*)ctx->c_out_ga.p, (int32_t*)ctx->c_out_src.p,
                                    (int32_t*)ctx->c_out_cnt.p, (void*)st);
  if (rc) return rc;
  if (n > 0) {
    CK(cudaMemcpyAsync(out_xyz, ctx->c_out_xyz.p, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(out_ground_adj, ctx->c_out_ga.p, (size_t)n, cudaMemcpyDeviceToHost, st));
    if (out_src_index) CK(cudaMemcpyAsync(out_src_index, ctx->c_out_src.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  }
  CK(cudaMemcpyAsync(out_count, ctx->c_out_cnt.p, (size_t)n_clouds * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return CCICP_OK;
}

// 3-D Lu-Milios edge information (computeEdgeInformationLUM, graphSlamTools.cpp:108-214), one CTA per edge.
int ccicp_edge_information_lum(ccicp_ctx* ctx, int32_t n_edges, const float* source_xyz, const int32_t* source_off,
                               const float* target_xyz, const int32_t* target_off, double max_corr_distance, double* info,
                               int32_t* num_corr, uint32_t* status) {
  NvtxRange nvtx_("ccicp_edge_information_lum");
  if (!ctx || n_edges < 0) return CCICP_E_INVALID;
  if (n_edges == 0) return CCICP_OK;
  if (!source_off || !target_off || !info || !num_corr || !status || source_off[0] != 0 || target_off[0] != 0) return CCICP_E_INVALID;
  for (int32_t e = 0; e < n_edges; e++)
    if (source_off[e + 1] < source_off[e] || target_off[e + 1] < target_off[e]) return CCICP_E_INVALID;
  const size_t ns = (size_t)source_off[n_edges], nt = (size_t)target_off[n_edges];
  if ((ns && !source_xyz) || (nt && !target_xyz)) return CCICP_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int rc;
  const size_t offb = ((size_t)n_edges + 1) * 4;
  if ((rc = reserve(ctx, ctx->l_src, ns * 12 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->l_tgt, nt * 12 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->l_soff, offb))) return rc;
  if ((rc = reserve(ctx, ctx->l_toff, offb))) return rc;
  if ((rc = reserve(ctx, ctx->l_dist, ns * 4 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->l_idx, ns * 4 + 16))) return rc;
  if ((rc = reserve(ctx, ctx->l_info, (size_t)n_edges * 36 * 8))) return rc;
  if ((rc = reserve(ctx, ctx->l_num, (size_t)n_edges * 4))) return rc;
  if ((rc = reserve(ctx, ctx->l_status, (size_t)n_edges * 4))) return rc;
  if (ns) CK(cudaMemcpyAsync(ctx->l_src.p, source_xyz, ns * 12, cudaMemcpyHostToDevice, st));
  if (nt) CK(cudaMemcpyAsync(ctx->l_tgt.p, target_xyz, nt * 12, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->l_soff.p, source_off, offb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->l_toff.p, target_off, offb, cudaMemcpyHostToDevice, st));
  lum::Args a{};
  a.source_xyz = (const float*)ctx->l_src.p; a.source_off = (const int32_t*)ctx->l_soff.p;
  a.target_xyz = (const float*)ctx->l_tgt.p; a.target_off = (const int32_t*)ctx->l_toff.p;
  a.nn_dist = (float*)ctx->l_dist.p; a.nn_idx = (int32_t*)ctx->l_idx.p;
  a.info = (double*)ctx->l_info.p; a.num_corr = (int32_t*)ctx->l_num.p; a.status = (uint32_t*)ctx->l_status.p;
  a.max_corr_distance = max_corr_distance; a.n_edges = n_edges;
  const int grid = n_edges < ctx->sm_count * 4 ? n_edges : ctx->sm_count * 4;
  // clouds beyond one shared-memory tile: nearest neighbours through the 3-D hash grid (lum_grid.cuh) instead of the
  // O(n * m) sweep (CCICP_LUM_GRID=0 / 1 forces either)
  int max_t = 0;
  for (int32_t e = 0; e < n_edges; e++) max_t = std::max(max_t, target_off[e + 1] - target_off[e]);
  bool use_grid = max_t > lum::kTile;
  { const char* ev = getenv("CCICP_LUM_GRID"); if (ev) use_grid = ev[0] == '1'; }
  int n_launches = 1;
  CK(cudaEventRecord(ctx->ev[2], st));
  if (use_grid && ns > 0 && nt > 0) {
    std::vector<int64_t> tabo((size_t)n_edges + 1, 0);
    for (int32_t e = 0; e < n_edges; e++) {
      int64_t h = 64;
      while (h < 2 * (int64_t)(target_off[e + 1] - target_off[e])) h <<= 1;
      tabo[e + 1] = tabo[e] + h;
    }
    if ((rc = reserve(ctx, ctx->l_tabo, tabo.size() * 8))) return rc;
    if ((rc = reserve(ctx, ctx->l_head, (size_t)tabo[n_edges] * 4))) return rc;
    if ((rc = reserve(ctx, ctx->l_next, nt * 4 + 16))) return rc;
    CK(cudaMemcpyAsync(ctx->l_tabo.p, tabo.data(), tabo.size() * 8, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));                     // tabo is a local
    CK(cudaMemsetAsync(ctx->l_head.p, 0xFF, (size_t)tabo[n_edges] * 4, st));
    lum::GridArgs g{};
    g.source_xyz = a.source_xyz; g.source_off = a.source_off; g.target_xyz = a.target_xyz; g.target_off = a.target_off;
    g.table_off = (const int64_t*)ctx->l_tabo.p; g.head = (int32_t*)ctx->l_head.p; g.next = (int32_t*)ctx->l_next.p;
    g.nn_dist = a.nn_dist; g.nn_idx = a.nn_idx; g.n_edges = n_edges;
    g.inv_h = (float)(1.0 / (max_corr_distance * 1.001));
    g.gate = (float)(max_corr_distance * max_corr_distance);
    if ((double)g.gate < max_corr_distance * max_corr_distance) g.gate = nextafterf(g.gate, INFINITY);
    lum::lum_grid_build_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(g, (int)nt);
    lum::lum_grid_nn_kernel<<<(unsigned)((ns + 255) / 256), 256, 0, st>>>(g, (int)ns);
    lum::lum_edge_kernel<true><<<grid, lum::kThreads, 0, st>>>(a);
    n_launches = 3;
  } else {
    lum::lum_edge_kernel<false><<<grid, lum::kThreads, 0, st>>>(a);
  }
  CK(cudaGetLastError());
  CK(cudaEventRecord(ctx->ev[3], st));
  ctx->timed = true; ctx->timed_copies = false; ctx->timed_stream = (void*)st;
  ctx->info.n_launches = n_launches; ctx->info.grid = grid; ctx->info.block = lum::kThreads; ctx->info.kernel_variant = use_grid ? 6 : 3;
  CK(cudaMemcpyAsync(info, ctx->l_info.p, (size_t)n_edges * 36 * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(num_corr, ctx->l_num.p, (size_t)n_edges * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(status, ctx->l_status.p, (size_t)n_edges * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return CCICP_OK;
}

int ccicp_host_alloc(void** out, size_t bytes) {
  if (!out) return CCICP_E_INVALID;
  *out = nullptr;
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
  if (e != cudaSuccess) { cudaGetLastError(); *out = nullptr; return e == cudaErrorMemoryAllocation ? CCICP_E_NOMEM : CCICP_E_CUDA; }
  return CCICP_OK;
}

int ccicp_host_free(void* p) {
  if (!p) return CCICP_OK;
  cudaError_t e = cudaFreeHost(p);
  if (e != cudaSuccess) { cudaGetLastError(); return CCICP_E_CUDA; }
  return CCICP_OK;
}

int ccicp_get_diag_counts(ccicp_ctx* ctx, uint64_t out[3]) {
  if (!ctx || !out) return CCICP_E_INVALID;
  out[0] = out[1] = out[2] = 0;
  if (!ctx->diag.p) return CCICP_OK;
  CK(cudaSetDevice(ctx->device));
  CK(cudaDeviceSynchronize());
  unsigned long long h[4] = {0, 0, 0, 0};
  CK(cudaMemcpy(h, ctx->diag.p, 32, cudaMemcpyDeviceToHost));
  CK(cudaMemset(ctx->diag.p, 0, 32));
  out[0] = h[0]; out[1] = h[1]; out[2] = h[2];
  return CCICP_OK;
}

int ccicp_get_launch_info(ccicp_ctx* ctx, ccicp_launch_info* out) {
  if (!ctx || !out) return CCICP_E_INVALID;
  if (ctx->timed) {
    if (cudaEventQuery(ctx->ev[3]) == cudaSuccess) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]) == cudaSuccess) ctx->info.last_kernel_ms = ms;
      if (ctx->timed_copies && cudaEventQuery(ctx->ev[5]) == cudaSuccess) {
        if (cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]) == cudaSuccess) ctx->info.last_h2d_ms = ms;
        if (cudaEventElapsedTime(&ms, ctx->ev[4], ctx->ev[5]) == cudaSuccess) ctx->info.last_d2h_ms = ms;
      }
    }
    cudaGetLastError();
  }
  *out = ctx->info;
  return CCICP_OK;
}

}  // extern "C"

// ---- ccicp_register_batch_device_static: the allocation-free variant of ccicp_register_batch_device -------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch memory icp_pair_kernel
// needs -- its work counters, the list of problems for the exact-order re-run, and the lazily built kd-tree nodes + rank
// tables of the tie replay of every CTA of the launch -- lives in ONE file-scope device array of
// CCICP_STATIC_SCRATCH_BYTES, which every GPU that loads the library holds until the process exits.
//   [0, 256)                              counters {work, re-run work, re-run count} (cleared by every call)
//   [256, 256 + 4 * CCICP_STATIC_MAX_PAIRS)   problem list of the exact-order re-run (written before it is read)
//   [.., CCICP_STATIC_SCRATCH_BYTES)      replay_scratch_bytes(max_model) per CTA; the grid is capped to the CTAs that fit
// Every call shares the array: two calls that overlap on the GPU (different streams, no ordering) race on it.
static __device__ __align__(256) unsigned char g_static_scratch[CCICP_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
constexpr size_t kStaticListOff = 256;
constexpr size_t kStaticReplayOff = kStaticListOff + (size_t)CCICP_STATIC_MAX_PAIRS * 4;
static_assert(kStaticReplayOff % 256 == 0 && kStaticReplayOff < (size_t)CCICP_STATIC_SCRATCH_BYTES, "static scratch layout");
void* g_static_ptr[kStaticMaxDevices];      // device address of g_static_scratch on each device (the name is a host-side stand-in)
std::mutex g_static_mu;

int static_fail(ccicp_ctx* ctx, int code, const char* fmt, long long got, long long limit) {
  char buf[256];
  snprintf(buf, sizeof(buf), fmt, got, limit);
  ctx->last_err = std::string("ccicp_register_batch_device_static: ") + buf;
  return code;
}
}  // namespace

extern "C" int ccicp_register_batch_device_static(ccicp_ctx* ctx, int32_t n_pairs, const double* d_model_xy,
                                                  const int32_t* d_model_off, const double* d_scene_xy,
                                                  const int32_t* d_scene_off, const double* d_init_Rt,
                                                  const ccicp_params* params, ccicp_result* d_out, int32_t* d_corr_idx,
                                                  int32_t max_model, int32_t max_scene, void* stream) {
  NvtxRange nvtx_("ccicp_register_batch_device_static");
  if (!ctx || !params || n_pairs < 0 || max_model < 0 || max_scene < 0) return CCICP_E_INVALID;
  if (n_pairs > 0 && (!d_model_xy || !d_model_off || !d_scene_xy || !d_scene_off || !d_init_Rt || !d_out)) return CCICP_E_INVALID;
  const uint32_t unsupported = CCICP_F_EXACT_SUMS | CCICP_F_TRIM | CCICP_F_DIAG_COUNTS | CCICP_F_CURV_GATE | CCICP_F_BEST_OF_SEEDS | CCICP_F_FIRST_BELOW;
  if (params->flags & unsupported)
    return static_fail(ctx, CCICP_E_INVALID, "flags 0x%llx select a kernel variant this entry point does not offer (unsupported mask 0x%llx)",
                       (long long)params->flags, (long long)unsupported);
  if (n_pairs > CCICP_STATIC_MAX_PAIRS)
    return static_fail(ctx, CCICP_E_TOO_LARGE, "n_pairs %lld exceeds CCICP_STATIC_MAX_PAIRS %lld", n_pairs, CCICP_STATIC_MAX_PAIRS);
  if (max_model > CCICP_STATIC_MAX_POINTS)
    return static_fail(ctx, CCICP_E_TOO_LARGE, "max_model %lld exceeds CCICP_STATIC_MAX_POINTS %lld", max_model, CCICP_STATIC_MAX_POINTS);
  if (max_scene > CCICP_STATIC_MAX_POINTS)
    return static_fail(ctx, CCICP_E_TOO_LARGE, "max_scene %lld exceeds CCICP_STATIC_MAX_POINTS %lld", max_scene, CCICP_STATIC_MAX_POINTS);
  if (ctx->device < 0 || ctx->device >= kStaticMaxDevices) return CCICP_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  if (n_pairs == 0) return CCICP_OK;

  // set-up, once per device, with that device current: the DEVICE address of the array
  unsigned char* base = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[ctx->device]) {
      void* p = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
      if (e != cudaSuccess || !p) return fail_cuda(ctx, e != cudaSuccess ? e : cudaErrorInvalidSymbol, "cudaGetSymbolAddress(g_static_scratch)");
      g_static_ptr[ctx->device] = p;
    }
    base = (unsigned char*)g_static_ptr[ctx->device];
  }
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;

  IcpArgs a{};
  a.mcap = round_up(max_model < 32 ? 32 : max_model, 32);
  a.scap = round_up(max_scene < 32 ? 32 : max_scene, 32);
  a.ccap = max_cells_for(max_model);
  if (a.ccap < 64) a.ccap = 64;
  const bool retry = !(params->flags & CCICP_F_NO_EXACT_RETRY);
  cudaFuncAttributes fa_on, fa_ex;
  CK(cudaFuncGetAttributes(&fa_on, icp_pair_kernel));
  CK(cudaFuncGetAttributes(&fa_ex, icp_pair_exact_kernel));
  size_t static_smem = fa_on.sharedSizeBytes;
  if (retry && fa_ex.sharedSizeBytes > static_smem) static_smem = fa_ex.sharedSizeBytes;
  const size_t smem = workset_bytes(a.mcap, a.scap, a.ccap, false);
  if (smem + static_smem > (size_t)ctx->max_smem_optin)       // the L2 variant needs a per-CTA workspace of megabytes: not offered here
    return static_fail(ctx, CCICP_E_TOO_LARGE, "the pair's working set (%lld bytes) exceeds one SM's shared memory (%lld); use ccicp_register_batch_device",
                       (long long)(smem + static_smem), ctx->max_smem_optin);
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, icp_pair_kernel, kThreadsOnChip, smem));
  if (occ < 1) return CCICP_E_TOO_LARGE;
  const size_t per_cta = replay_scratch_bytes(max_model);
  const long long fit = (long long)(((size_t)CCICP_STATIC_SCRATCH_BYTES - kStaticReplayOff) / per_cta);
  long long grid = (long long)occ * ctx->sm_count;
  if (grid > fit) grid = fit;
  if (grid > n_pairs) grid = n_pairs;
  if (grid < 1) return CCICP_E_TOO_LARGE;

  int32_t* cnt = (int32_t*)base;
  a.model_xy = d_model_xy; a.model_off = d_model_off; a.scene_xy = d_scene_xy; a.scene_off = d_scene_off;
  a.init_Rt = d_init_Rt; a.out = d_out; a.corr_idx = d_corr_idx;
  a.work_counter = cnt;
  a.scratch = (char*)base + kStaticReplayOff; a.scratch_per_cta = per_cta;
  a.pair_base = 0; a.n_pairs = n_pairs; a.n_seeds = 1; a.seeds_per_item = 1; a.items_per_pair = 1;
  a.max_iter = params->max_iter; a.flags = params->flags; a.min_delta = params->min_delta; a.indist = params->indist;
  a.trim_frac = params->trim_frac;
  a.curv_tol = params->curv_tol; a.feat_radius = params->feat_radius; a.feat_k = params->feat_k;
  a.total_scene = 0; a.max_model = max_model;

  CK(cudaMemsetAsync(cnt, 0, 256, st));       // the array keeps the last call's contents: the counters must start at zero
  icp_pair_kernel<<<(unsigned)grid, kThreadsOnChip, smem, st>>>(a);
  CK(cudaGetLastError());
  if (retry) {
    int32_t* list = (int32_t*)(base + kStaticListOff);
    collect_retry_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, st>>>(d_out, 0, (int)n_pairs, list, cnt + 2);
    CK(cudaGetLastError());
    IcpArgs a2 = a;
    a2.item_list = list; a2.item_count = cnt + 2; a2.work_counter = cnt + 1;
    int occ2 = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, icp_pair_exact_kernel, kThreadsOnChip, smem));
    if (occ2 < 1) return CCICP_E_TOO_LARGE;
    long long grid2 = (long long)occ2 * ctx->sm_count;
    if (grid2 > grid) grid2 = grid;              // the replay scratch holds `grid` CTAs
    icp_pair_exact_kernel<<<(unsigned)grid2, kThreadsOnChip, smem, st>>>(a2);
    CK(cudaGetLastError());
  }
  return CCICP_OK;
}
