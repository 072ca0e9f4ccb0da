/* ccicp2d_b200.h -- C ABI of libccicp2d_b200.so: batched, B200-native (sm_100a)
 * two-class point-to-point 2-D ICP.
 *
 * This is the drop-in boundary for ONE hot path of the servos/SLAM reference:
 * the pair of calls at ccicp2d/src/icpTools.cpp:187-188
 *     IcpPointToPoint icp(refGA, refNGA, nGA, nNGA, 2);
 *     icp.fit(tarGA, tarNGA, nGA, nNGA, rot, trans, indist, 0);
 * i.e. Icp::Icp (ccicp2d/src/icp.cpp:26-70), Icp::fit / fitIterate
 * (icp.cpp:80-122), IcpPointToPoint::fitStep (icpPointToPoint.cpp:33-172),
 * kdtree::KDTree::n_nearest(q,1) (kdtree.cpp:378-391, 515-683),
 * IcpPointToPoint::getEdgeWeight (icpPointToPoint.cpp:233-316) and
 * getNumberCorrespondences (:317-320).  The reference has no FFI of its own;
 * its interface is the C++ class above.  the headers under include/ccicp2d/ re-declare that
 * class on top of this ABI so ccicp2d/src/icpTools.cpp compiles unchanged
 * (see INTEGRATION.md).
 *
 * Plain pointers and sizes only; no C++/torch types.  All functions return a
 * CCICP_E_* code (0 = OK) and never abort or exit.  There is no CPU fallback:
 * without a usable CUDA device every compute entry point fails with
 * CCICP_E_NO_DEVICE / CCICP_E_CUDA.
 */
#ifndef CCICP2D_B200_H
#define CCICP2D_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CCICP_ABI_VERSION 5

/* ---- call-level error codes ------------------------------------------- */
enum {
  CCICP_OK = 0,
  CCICP_E_INVALID = 1,     /* bad argument (NULL pointer, negative size, offsets not monotone) */
  CCICP_E_NO_DEVICE = 2,   /* no CUDA device / device index out of range                        */
  CCICP_E_CUDA = 3,        /* a CUDA runtime call failed; see ccicp_last_cuda_error              */
  CCICP_E_NOMEM = 4,       /* host or device allocation failed                                   */
  CCICP_E_TOO_LARGE = 5,   /* a scan exceeds the largest supported size (CCICP_MAX_POINTS)       */
  CCICP_E_ARCH = 6         /* device is not sm_100 (the library carries sm_100a code only)       */
};

/* Largest number of points per class-pair cloud (model or scene, GA+NGA) the
 * kernels accept.  The reference caps each class at 20000 points
 * (ccicp2d/include/ccicp2d/icpTools.h:21). */
#define CCICP_MAX_POINTS 65535

/* ---- per-problem status bits (ccicp_result.status) -------------------- */
enum {
  CCICP_ST_TOO_FEW_MODEL = 1u << 0, /* nGA+nNGA < 5 model points: reference ctor bails (icp.cpp:37-43); R,t returned unchanged */
  CCICP_ST_TOO_FEW_SCENE = 1u << 1, /* nGA+nNGA < 5 scene points: fit returns early (icp.cpp:100-103); R,t unchanged            */
  CCICP_ST_NO_CORR       = 1u << 2, /* no iteration produced a correspondence (fitStep -1, icpPointToPoint.cpp:128-131)        */
  CCICP_ST_REFLECTION    = 1u << 3, /* some iteration had det(H) < 0; like the reference (no det fix, icpPointToPoint.cpp:162)
                                       the improper V*U^T was applied                                                         */
  CCICP_ST_DEGENERATE    = 1u << 4, /* some iteration had a (nearly) rank-deficient H: |det H| <= 1e-9 ||H||_F^2 (one or two
                                       correspondences, collinear ones).  There the reference's SVD (matrix.cpp:582-810) and the
                                       default closed form pick their factor differently, so the library re-runs such a problem
                                       with the exact-order solve (CCICP_ST_EXACT_RETRY); the bit survives only where that
                                       re-run met det H == 0 exactly or the re-run is disabled (CCICP_F_NO_EXACT_RETRY)          */
  CCICP_ST_TIE_REPLAY    = 1u << 5, /* >= 1 query had several model points at the same float distance and was resolved by
                                       replaying the reference kd-tree traversal (kdtree.cpp:515-683)                          */
  CCICP_ST_CONVERGED     = 1u << 6, /* stopped because delta < min_delta (icp.cpp:119-120)                                     */
  CCICP_ST_TREE_OVERFLOW = 1u << 7, /* tie replay needed a deeper kd-tree than the replay stack supports; tie left to the
                                       grid's lowest-slot rule                                                                  */
  CCICP_ST_INFO_SINGULAR = 1u << 8, /* getEdgeWeight's 3x3 system was singular / ss == 0; info[] holds inf/nan like the ref    */
  CCICP_ST_NO_SEED_BELOW = 1u << 10, /* CCICP_F_FIRST_BELOW: no seed reached score / n_scene < score_threshold; the record is the
                                       arg-min seed instead (global_match.cpp:151-174 publishes nothing in that case)           */
  CCICP_ST_EDGE_REJECTED = 1u << 11, /* ccicp_match_batch with reject_dist > 0: the result moved too far from the initial pose     */
  CCICP_ST_EXACT_RETRY   = 1u << 12, /* this record comes from the exact-order re-run of a problem whose default-mode solve was
                                       flagged CCICP_ST_DEGENERATE: R,t bit-identical to the reference                          */
  CCICP_ST_HINT_TOO_SMALL = 1u << 9 /* max_cells_hint of ccicp_register_batch_device_ex was too small for this pair: not
                                       registered, R,t returned unchanged                                                      */
};

/* ---- flags (ccicp_params.flags) ---------------------------------------- */
enum {
  CCICP_F_INFO           = 1u << 0, /* also compute the Lu-Milios information matrix (getEdgeWeight) into result.info         */
  CCICP_F_INFO_FIXED_DY  = 1u << 1, /* use dy = ay - by instead of the reference's dy = ax - bx (icpPointToPoint.cpp:262)      */
  CCICP_F_BEST_OF_SEEDS  = 1u << 2, /* n_seeds > 1: write one record per pair, the arg-min score seed (ties: lowest seed)       */
  CCICP_F_FIRST_BELOW    = 1u << 5, /* with BEST_OF_SEEDS: global_matching's acceptance rule (global_match.cpp:131-151) -- the
                                       FIRST seed (lowest index) whose normalised score  score / n_scene  is below
                                       score_threshold wins, instead of the arg-min                                            */
  CCICP_F_TRIM           = 1u << 6, /* north_star extra, NO counterpart in the reference (its only outlier test is the fixed
                                       gate, icpPointToPoint.cpp:76,111): trimmed rejection, keep the params.trim_frac
                                       fraction of the gated correspondences per iteration.  Off in every parity run. */
  CCICP_F_CURV_GATE      = 1u << 9, /* north_star extra, NO counterpart on the reference's built path (its "cc" is the class
                                       constraint GA / NGA; its only normal code is the dead icpPointToPlane.cpp:276-348):
                                       curvature-gated matching.  Every model and scene point gets a curvature from its scan
                                       neighbours (ccicp_scan_features, params.feat_k / feat_radius); a model point is a
                                       candidate for a scene point only if the two curvatures differ by at most
                                       params.curv_tol, IN ADDITION to the class constraint.  Points without a feature are
                                       compatible with everything; float ties go to the lowest model index (no kd-tree replay:
                                       there is no reference traversal to follow).  Off in every parity run.              */
  CCICP_F_DIAG_COUNTS    = 1u << 7, /* diagnostics only: run the counting instantiation of the pair kernel (same algorithm and
                                       results, slower) and accumulate the work it did; read with ccicp_get_diag_counts.  Never
                                       set in a timed run.                                                                     */
  CCICP_F_NO_EXACT_RETRY = 1u << 8, /* diagnostics only: keep the default-mode record of problems flagged CCICP_ST_DEGENERATE    */
  CCICP_F_NO_TIE_REPLAY  = 1u << 3, /* diagnostics only: skip the kd-tree tie replay (ties -> lowest grid slot)                 */
  CCICP_F_EXACT_SUMS     = 1u << 4  /* evaluate centroids, H, the SVD (Matrix::svd, matrix.cpp:582-810) and the compose step in
                                       the reference's own sequential order: R,t of every iteration are then BIT-IDENTICAL to
                                       the reference (about 4x slower; default is a parallel sum + closed form, ~1e-13 m off)   */
};

typedef struct ccicp_params {
  int32_t  max_iter;   /* Icp::setMaxIterations (icp.h:48); reference default 20 (icp.cpp:27)      */
  uint32_t flags;      /* CCICP_F_*                                                                 */
  double   min_delta;  /* Icp::setMinDeltaParam (icp.h:51); default 1e-6; < -1 => fixed iterations */
  double   indist;     /* gate on the SQUARED float NN distance: keep iff (double)dis < indist
                          (icpPointToPoint.cpp:76,111); the facade passes 5 (icpTools.cpp:188)      */
  double   min_corr_frac; /* BEST_OF_SEEDS: a seed scores +inf unless num_corr >= frac * n_scene    */
  double   score_threshold; /* FIRST_BELOW: accept a seed iff score / n_scene < threshold
                               (MAX_SCORE 0.002, global_match.cpp:30)                               */
  double   reject_dist;   /* ccicp_match_batch: > 0 enables graph_slam's bad-match test: status gets
                             CCICP_ST_EDGE_REJECTED when |dx| or |dy| between initial and final pose exceeds this
                             (DIST_MOVE_THRESH 10 m) ...                                            */
  double   reject_rot;    /* ... or the wrapped |dyaw| exceeds this (ROT_MOVE_THRESH 0.2 rad)
                             (graph_slam/src/graphSlamTools.cpp:332-362, graphSlamTools.h:35-36)    */
  double   trim_frac;     /* CCICP_F_TRIM: fraction in (0, 1] of each iteration's gated correspondences that is kept
                             (those with the smallest squared distance); default 1 = the reference's behaviour     */
  double   curv_tol;      /* CCICP_F_CURV_GATE: largest |curvature(model) - curvature(scene)| of a candidate (default 0.05) */
  double   feat_radius;   /* features: neighbours farther than this from the point are ignored [m] (default 0.5)   */
  int32_t  feat_k;        /* features: index neighbours taken on each side (default 5; icpPointToPlane.cpp used 10 in all) */
  int32_t  reserved;
} ccicp_params;

/* One record per (pair, seed) problem -- 160 bytes. */
typedef struct ccicp_result {
  double   R[4];        /* row-major 2x2: model = R * scene + t                                      */
  double   t[2];
  double   info[9];     /* row-major 3x3 getEdgeWeight output (valid iff CCICP_F_INFO)               */
  double   last_delta;  /* return value of the last executed fitStep (-1 if it had no correspondences) */
  double   score;       /* mean squared float NN distance over the correspondences of the last
                           successful iteration (+inf if none)                                        */
  int32_t  num_corr;    /* getNumberCorrespondences(): count of the last successful fitStep          */
  int32_t  iters;       /* fitStep calls executed                                                    */
  uint32_t status;      /* CCICP_ST_*                                                                */
  int32_t  best_seed;   /* seed index this record belongs to                                         */
  int32_t  n_tie_replays; /* queries resolved by the kd-tree replay, summed over iterations         */
  int32_t  reserved;
} ccicp_result;

typedef struct ccicp_ctx ccicp_ctx;

/* ---- lifetime ----------------------------------------------------------- */
int  ccicp_abi_version(void);
int  ccicp_create(ccicp_ctx** out, int device);
int  ccicp_destroy(ccicp_ctx* ctx);
const char* ccicp_strerror(int code);
/* Text of the last CUDA error seen by this ctx (empty string if none). */
const char* ccicp_last_cuda_error(const ccicp_ctx* ctx);
void ccicp_default_params(ccicp_params* p);   /* max_iter 20, min_delta 1e-6, indist 5, flags 0 */

/* ---- the hot path, host buffers ----------------------------------------
 * Replaces, for every pair p (and every seed s of it), the two calls at
 * ccicp2d/src/icpTools.cpp:187-188.
 *
 * Layout (CSR over pairs, points as interleaved x,y doubles exactly as the
 * reference takes them, GA block first then NGA block):
 *   model_off[2*n_pairs+1] : pair p's GA model points are [off[2p], off[2p+1]),
 *                            its NGA model points [off[2p+1], off[2p+2])   (units: points)
 *   model_xy [2*model_off[2*n_pairs]]
 *   scene_off / scene_xy   : same for the scene ("template") clouds
 *   init_Rt  [n_pairs*n_seeds][6] : {r00,r01,r10,r11,t0,t1} initial R,t (fit's in/out R,t)
 *   out      [n_pairs*n_seeds]    (or [n_pairs] with CCICP_F_BEST_OF_SEEDS)
 *   corr_idx : NULL, or [n_seeds][scene_off[2*n_pairs]] int32 (seed-major; with n_seeds == 1
 *              simply one entry per scene point): per scene point the per-class model index
 *              matched in the LAST executed iteration, -1 if gated / class skipped
 *              (what KDTreeResult.idx holds, kdtree.h:31-35).
 * Ownership: every buffer is caller-owned and only read/written during the
 * call (the reference copies the model, icp.cpp:51-60, and borrows the scene).
 * Synchronous.  Host memory may be pageable or pinned (ccicp_host_alloc).     */
int ccicp_register_batch(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds,
                         const double* model_xy, const int32_t* model_off,
                         const double* scene_xy, const int32_t* scene_off,
                         const double* init_Rt, const ccicp_params* params,
                         ccicp_result* out, int32_t* corr_idx);

/* ---- the hot path, device-resident buffers -------------------------------
 * Same contract, but every pointer is a device pointer on ctx's device and the
 * work is enqueued on `stream` (a cudaStream_t cast to void*; NULL = the
 * ctx's own stream) without synchronising.  max_model / max_scene are the
 * largest per-pair point counts (GA+NGA) in the batch, max_cells_hint may be
 * 0.  Use ccicp_stream_sync (or your own event) before reading results.
 * A ctx serves ONE in-flight call at a time: its work counter, tie-replay scratch, L2 workspace and best-of-seeds
 * staging are per-ctx (and may be re-allocated by the next call), so a second call on the same ctx -- from
 * another thread or on another stream -- must wait until the first has completed.  Use one ctx per host thread /
 * stream (ctxs are cheap; several may share a device).  A pair whose real size exceeds max_model / max_scene (or
 * the cell hint) is not registered: status CCICP_ST_HINT_TOO_SMALL, R,t unchanged, corr_idx -1.               */
int ccicp_register_batch_device(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds,
                                const double* d_model_xy, const int32_t* d_model_off,
                                const double* d_scene_xy, const int32_t* d_scene_off,
                                const double* d_init_Rt, const ccicp_params* params,
                                ccicp_result* d_out, int32_t* d_corr_idx,
                                int32_t max_model, int32_t max_scene, void* stream);
/* Same with a hint for the shared-memory cell table: max over the pairs of ccicp_grid_cells(nGA) + ccicp_grid_cells(nNGA)
 * (model points per class).  0 = unknown: the worst case for max_model is assumed, which can cost one resident
 * CTA per SM.  The hint is checked on the device: a pair that needs more cells is not registered and is reported
 * with status CCICP_ST_HINT_TOO_SMALL (R,t unchanged).                                                          */
int ccicp_register_batch_device_ex(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds,
                                   const double* d_model_xy, const int32_t* d_model_off,
                                   const double* d_scene_xy, const int32_t* d_scene_off,
                                   const double* d_init_Rt, const ccicp_params* params,
                                   ccicp_result* d_out, int32_t* d_corr_idx,
                                   int32_t max_model, int32_t max_scene, int32_t max_cells_hint, void* stream);
/* ---- the allocation-free variant -----------------------------------------------------------------------------------
 * NO counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same contract and same records as
 * ccicp_register_batch_device with n_seeds = 1 (device pointers, enqueued on `stream`, 0 = the ctx's), but it calls no
 * cudaMalloc / cudaFree and no allocating library routine: the scratch memory of the pair kernel (work counters, the
 * problem list of the exact-order re-run, the kd-tree nodes and rank tables of every CTA's tie replay) is ONE file-scope
 * device array of CCICP_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until the process exits.
 * Limits, checked before anything is enqueued (CCICP_E_TOO_LARGE, the reason in ccicp_last_cuda_error):
 * n_pairs <= CCICP_STATIC_MAX_PAIRS, max_model and max_scene <= CCICP_STATIC_MAX_POINTS, and the pair's working set
 * must fit one SM's shared memory (the large-scan variant needs a per-CTA workspace of megabytes).  Flags that select
 * another kernel variant (EXACT_SUMS, TRIM, DIAG_COUNTS, CURV_GATE, BEST_OF_SEEDS, FIRST_BELOW) -> CCICP_E_INVALID.
 * The launch is capped to the CTAs whose replay scratch fits the array (about 220 for 1081-beam scans, 119 at 2048 points;
 * the records do not depend on the grid size).  EVERY call on a device shares the array: the counters are cleared on
 * the stream by each call, and two calls that overlap on the GPU -- different streams or ctxs without an ordering
 * between them -- race on it.                                                                                      */
#define CCICP_STATIC_SCRATCH_BYTES (32 << 20)
#define CCICP_STATIC_MAX_POINTS 2048
#define CCICP_STATIC_MAX_PAIRS 65536
int ccicp_register_batch_device_static(ccicp_ctx* ctx, int32_t n_pairs,
                                       const double* d_model_xy, const int32_t* d_model_off,
                                       const double* d_scene_xy, const int32_t* d_scene_off,
                                       const double* d_init_Rt, const ccicp_params* params,
                                       ccicp_result* d_out, int32_t* d_corr_idx,
                                       int32_t max_model, int32_t max_scene, void* stream);
/* ---- scan table: key-frame scans uploaded once, pairs as index lists ---------------------------------------------
 * Loop-closure batches reuse every key-frame scan many times: graph_slam registers each new key frame against its K = 3
 * nearest key frames plus the previous one (graph_slam/include/graph_slam/graphSlamTools.h:26-36,
 * graphSlamTools.cpp:72-106, 406-425), so a batch of P candidate pairs touches far fewer than 2 P distinct scans.
 * ccicp_scans_upload copies the scans to the device once (scan_off[2*n_scans+1]: GA block then NGA block per scan, as
 * model_off above; replaces the ctx's previous table); ccicp_register_scan_pairs then registers scan scene_scan[p]
 * (the "template") against scan model_scan[p] for every pair with the same contract as ccicp_register_batch, uploading
 * only the two index lists and the initial transforms.  No correspondence indices on this path.  Synchronous.      */
int ccicp_scans_upload(ccicp_ctx* ctx, int32_t n_scans, const double* scan_xy, const int32_t* scan_off);
int ccicp_register_scan_pairs(ccicp_ctx* ctx, int32_t n_pairs, int32_t n_seeds,
                              const int32_t* model_scan, const int32_t* scene_scan,
                              const double* init_Rt, const ccicp_params* params, ccicp_result* out);

/* ---- per-point features of a scan: normal and curvature (north_star extra, see CCICP_F_CURV_GATE) ----------------
 * Shape of the reference's dead computeNormal (ccicp2d/src/icpPointToPlane.cpp:276-338), in 2-D: PCA over the scan
 * neighbours.  Clouds in the CSR layout of the registration calls (xy FP64, off[2*n_clouds+1]: GA block then NGA block
 * per cloud).  Neighbours of point i: the points of its own class block within feat_k positions of i (a scan arrives in
 * beam order) and within feat_radius of it, i included; fewer than 3 -> curvature NaN, normal (0, 0).  With l1 >= l2
 * the eigenvalues of their scatter matrix: curvature = l2 / (l1 + l2) in [0, 0.5]; normal = unit eigenvector of l2
 * turned towards the origin of the scan frame.  normal_xy may be NULL.  Host buffers; synchronous.              */
int ccicp_scan_features(ccicp_ctx* ctx, int32_t n_clouds, const double* xy, const int32_t* off, int32_t feat_k,
                        double feat_radius, float* normal_xy, float* curvature);

int32_t ccicp_grid_cells(int32_t n_class_points);   /* cells of the table for a class of n model points */
int ccicp_stream_sync(ccicp_ctx* ctx, void* stream);

/* ---- the immediate caller, batched: CCICP::doICPMatch(model_GA, model_NGA, scene_GA, scene_NGA, initPose)
 * (ccicp2d/src/icpTools.cpp:123-218), SURVEY.md 8(f) rank 1.
 * A pose is geometry_msgs::Pose: position x,y,z + orientation quaternion x,y,z,w.  Per pair:
 *   R = rot(yaw(init.orientation)), t = (init.x, init.y)                        (icpTools.cpp:168-176)
 *   fewer than 5 scene points: out = {0,0,0, 0,0,0, 9999}, the reference's sentinel (icpTools.cpp:179-184),
 *                              status CCICP_ST_TOO_FEW_SCENE, nothing is registered
 *   otherwise the two calls at icpTools.cpp:187-188 with `params` (the reference passes indist 5 and keeps the
 *   Icp defaults max_iter 20 / min_delta 1e-6: ccicp_default_params), then
 *   out.x,y = t; out.z = init.z; out.orientation = quaternion(roll(init), pitch(init), atan2(R10, R00))
 *                                                                               (icpTools.cpp:195-212)
 * The yaw/roll/pitch <-> quaternion conversions restate ROS tf (tf::getYaw, tf::Matrix3x3::getEulerYPR,
 * tf::createQuaternionFromRPY; the reference depends on an un-vendored, un-pinned tf: ccicp2d/package.xml:10-28).
 * `records` may be NULL; otherwise it receives the per-pair ccicp_result as ccicp_register_batch would.      */
typedef struct ccicp_pose {
  double px, py, pz;
  double qx, qy, qz, qw;
} ccicp_pose;
int ccicp_match_batch(ccicp_ctx* ctx, int32_t n_pairs,
                      const double* model_xy, const int32_t* model_off,
                      const double* scene_xy, const int32_t* scene_off,
                      const ccicp_pose* init_pose, const ccicp_params* params,
                      ccicp_pose* out_pose, ccicp_result* records);
/* The pose conversions on their own (host arithmetic, exposed for tests and for callers that batch themselves). */
void ccicp_pose_to_rt(const ccicp_pose* pose, double* Rt6);
void ccicp_rt_to_pose(const double* Rt6, const ccicp_pose* init_pose, ccicp_pose* out);

/* ---- the step before the path: GA / NGA classification of segmented obstacle points --------------------
 * CCICP::classifyPoints (ccicp2d/src/icpTools.cpp:36-103; constants icpTools.h:24-26), SURVEY.md 8(f) rank 2,
 * batched over clouds.  Per cloud:
 *   - every point is binned on a 1200 x 1200 grid of 0.5 m cells centred on the origin:
 *       bind = floor(((double)coord + 300.0) / 0.5); points outside [0,1200) in x or y are dropped (:47-54)
 *   - the output lists the points bin by bin (bin index x*1200 + y ascending), points of one bin in input order;
 *     points of the outermost ring of bins are dropped (:72-77) but still count as "occupied" for their neighbours
 *   - ground_adj = 1 iff at least 2 (GRD_ADJ_THRESH) of the bin's 8 neighbour bins are empty (:79-98)
 * xyz: float [cloud_off[n_clouds]][3] (PointXYZGD x,y,z); cloud_off: int32 [n_clouds+1] (points).
 * out_xyz / out_ground_adj / out_src_index (may be NULL; index of the point inside its input cloud) are written
 * per cloud starting at cloud_off[c]; out_count[c] = points kept for cloud c.  NaN coordinates are dropped
 * (the reference's float -> int conversion of NaN is undefined).  Host buffers; synchronous.
 * At most 2048 clouds per call.                                                                              */
int ccicp_classify_points(ccicp_ctx* ctx, int32_t n_clouds, const float* xyz, const int32_t* cloud_off,
                          float* out_xyz, uint8_t* out_ground_adj, int32_t* out_src_index, int32_t* out_count);
/* Same, device-resident buffers, asynchronous on `stream` (NULL = ctx stream). */
int ccicp_classify_points_device(ccicp_ctx* ctx, int32_t n_clouds, const float* d_xyz, const int32_t* d_cloud_off,
                                 int64_t n_points, float* d_out_xyz, uint8_t* d_out_ground_adj,
                                 int32_t* d_out_src_index, int32_t* d_out_count, void* stream);

/* ---- the step after the path in graph_slam: 3-D Lu-Milios edge information ------------------------------
 * computeEdgeInformationLUM(source_trans, target, max_corr_distance) (graph_slam/src/graphSlamTools.cpp:108-214;
 * called with 0.75 at :302), SURVEY.md 8(f) rank 3, batched over edges.  source_xyz holds the source clouds ALREADY
 * transformed by the registration result (as the reference expects), target_xyz the target clouds, float x,y,z;
 * *_off are int32 [n_edges+1] in points.  info receives the 6x6 matrix row-major per edge (MM / s^2, or the identity
 * when s^2 < 1e-13 / not finite: status[e] = 1), num_corr the number of point pairs closer than max_corr_distance.
 * Nearest neighbours are exact in float squared distance (the reference uses PCL's KdTreeFLANN; ties -> lowest
 * target index here).  Host buffers; synchronous.                                                              */
int ccicp_edge_information_lum(ccicp_ctx* ctx, int32_t n_edges, const float* source_xyz, const int32_t* source_off,
                               const float* target_xyz, const int32_t* target_off, double max_corr_distance,
                               double* info, int32_t* num_corr, uint32_t* status);

/* Loop-closure candidate generator of graph_slam: graphSlamGetKNN (graph_slam/src/graphSlamTools.cpp:72-106), host
 * logic.  nodes_xy: [n_nodes][2] key-frame positions; the query node has index query_idx and position query_xy.
 * Like the reference: only nodes 0 .. n_nodes-3 are considered, K is clipped to n_nodes-1, candidates are ordered by
 * planar distance (ties: lower index first) and the query's own index is skipped.  Returns the number of candidate
 * indices written to out (<= K).                                                                                  */
int32_t ccicp_knn_candidates(const double* nodes_xy, int32_t n_nodes, const double* query_xy, int32_t query_idx,
                             int32_t K, int32_t* out);

/* ---- multi-device: the batch split over several GPUs of one process ---------------------------------------------
 * north_star: "the batch is sharded across the 8 GPUs of one box with only a final NCCL gather of transforms, since
 * pairs are independent".  The reference has no counterpart (single-threaded CPU); the consumer of such a call is
 * graph_slam's loop-closure candidate producer: graphSlamGetKNN (graph_slam/src/graphSlamTools.cpp:72-106) proposes the
 * candidate pairs, graphSlamAddEdgeToX (:406-425) registers each (the ICP call at :274).
 * Device g of G registers the contiguous pair range ccicp_shard_range(n_pairs, G, g) with its own ctx, driven by its
 * own host thread inside the calling process; there is no data-path collective.
 *   devices: CUDA device indices (a device may be listed twice: two ctxs share it); NULL, 0 = every visible device. */
typedef struct ccicp_multi ccicp_multi;
int  ccicp_multi_create(const int* devices, int n_devices, ccicp_multi** out);
int  ccicp_multi_destroy(ccicp_multi* m);
int  ccicp_multi_device_count(const ccicp_multi* m);
ccicp_ctx* ccicp_multi_ctx(ccicp_multi* m, int g);            /* the ctx of device slot g (launch info, errors)     */
const char* ccicp_multi_last_error(const ccicp_multi* m);
void ccicp_shard_range(int32_t n_pairs, int n_devices, int g, int32_t* begin, int32_t* end);  /* contiguous, balanced:
                                             sizes differ by at most one, larger shards first */
/* Host buffers: same arguments and result layout as ccicp_register_batch; every shard's records are copied straight
 * into the caller's out[] (that is the gather on this path).  corr_idx only with n_seeds == 1.  Synchronous.        */
int ccicp_register_batch_multi(ccicp_multi* m, int32_t n_pairs, int32_t n_seeds,
                               const double* model_xy, const int32_t* model_off,
                               const double* scene_xy, const int32_t* scene_off,
                               const double* init_Rt, const ccicp_params* params,
                               ccicp_result* out, int32_t* corr_idx);
/* Device-resident shards: every argument is an array with one entry per device slot.  Device g holds ITS shard as a CSR
 * of its own (offsets from 0; d_init_Rt[g] = the shard's transforms) and a record buffer d_out[g] sized for ALL n_pairs
 * (x n_seeds unless best-of) records.  Each device registers its shard into its slice of d_out[g]; the slices are then
 * exchanged so that EVERY d_out[g] holds every record: over NCCL (one grouped ncclBroadcast per shard over
 * NVLink/NVSwitch; libnccl.so.2 is bound at run time) or, when NCCL is unavailable or a device is listed twice, by peer
 * copies (ccicp_multi_gather_kind: 1 = NCCL, 2 = peer copies).  Asynchronous; ccicp_multi_sync waits for all devices. */
int ccicp_register_batch_multi_device(ccicp_multi* m, int32_t n_pairs, int32_t n_seeds,
                                      const double* const* d_model_xy, const int32_t* const* d_model_off,
                                      const double* const* d_scene_xy, const int32_t* const* d_scene_off,
                                      const double* const* d_init_Rt, const ccicp_params* params,
                                      ccicp_result* const* d_out, const int32_t* max_model, const int32_t* max_scene,
                                      const int32_t* max_cells_hint /* may be NULL */);
int ccicp_multi_sync(ccicp_multi* m);
int ccicp_multi_gather_kind(const ccicp_multi* m);

/* Pinned host memory for callers that want the H2D/D2H copies to overlap. */
int ccicp_host_alloc(void** out, size_t bytes);
int ccicp_host_free(void* p);

/* ---- introspection for benchmarks / tests -------------------------------- */
typedef struct ccicp_launch_info {
  int32_t grid, block, smem_bytes, regs_per_thread, blocks_per_sm, kernel_variant;
  int32_t n_launches;          /* kernels launched by the last register call */
  int32_t sm_count;
  float   last_kernel_ms;      /* device time of the last register call's ICP kernel(s), CUDA events on its stream;
                                  valid after a host-buffer call or after ccicp_stream_sync                      */
  float   last_h2d_ms, last_d2h_ms;
} ccicp_launch_info;
int ccicp_get_launch_info(ccicp_ctx* ctx, ccicp_launch_info* out);
/* Work counted by the CCICP_F_DIAG_COUNTS launches since the last call (then reset), synchronises the ctx streams:
 * out[0] = float distance evaluations of the NN search, out[1] = NN queries searched, out[2] = accepted (gated)
 * correspondences summed over the iterations.  The "useful work" of bench.py's roofline_useful comes from these. */
int ccicp_get_diag_counts(ccicp_ctx* ctx, uint64_t out[3]);

#ifdef __cplusplus
}
#endif
#endif /* CCICP2D_B200_H */
