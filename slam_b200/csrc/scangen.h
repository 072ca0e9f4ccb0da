/* scangen.h -- seeded synthetic 2-D lidar scan-pair generator (workload
 * generator for tests and bench.py).
 *
 * The reference has no scan generator or fixtures of its own (SURVEY.md
 * section 4): its live input is a ground-segmented Velodyne cloud projected to
 * xy (ccicp2d/src/icpTools.cpp:248-276).  The inner Icp API is a plain 2-D
 * two-class point-set API (ccicp2d/include/ccicp2d/icp.h:42,65), so Hokuyo
 * style 1081-beam scans drive it unchanged.
 *
 * Everything is counter-based (splitmix64 keyed by seed / scan / beam), so a
 * pair depends only on its own seed, never on batch order or thread count.
 */
#ifndef SLAM_B200_SCANGEN_H
#define SLAM_B200_SCANGEN_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct scangen_params {
  int32_t n_beams;        /* beams per scan (1081)                          */
  int32_t max_boxes;      /* 0..max_boxes box obstacles per world (4)       */
  int32_t class_rule;     /* 0: GA iff range jump to a neighbour beam > jump_thresh; 1: all NGA; 2: all GA */
  int32_t reserved;
  double fov_start_deg;   /* first beam angle (-135)                        */
  double fov_step_deg;    /* beam step (0.25)                               */
  double noise_sigma;     /* range noise sigma [m] (0.01)                   */
  double max_range;       /* beams beyond this are dropped (30)             */
  double room_rmin;       /* base room radius ~ U[room_rmin, room_rmax]     */
  double room_rmax;
  double off_xy;          /* truth offset tx,ty ~ U[-off_xy, off_xy]        */
  double off_th;          /* truth yaw ~ U[-off_th, off_th]                 */
  double init_sigma_xy;   /* init = truth + N(0, sigma) if > 0, else identity */
  double init_sigma_th;
  double jump_thresh;     /* class_rule 0 threshold [m] (0.04)              */
} scangen_params;

/* config: 1 = C1/C2/C5 odometry pairs, 3 = C3 loop-closure pairs, 4 = C4 dense 32768-beam */
void scangen_default(scangen_params* p, int32_t config);

/* Generates pairs with seeds seed0 .. seed0+n_pairs-1.
 *   model_xy, scene_xy : [n_pairs][n_beams][2] doubles; each pair's block holds its
 *                        GA points first, then its NGA points, compacted; tail unused
 *   counts             : [n_pairs][4] = {n_model_ga, n_model_nga, n_scene_ga, n_scene_nga}
 *   truth, init        : [n_pairs][3] = {tx, ty, yaw}:  model = R(yaw) * scene + t
 * Returns 0, or -1 on bad arguments. Thread-parallel over pairs (OpenMP). */
int32_t scangen_batch(const scangen_params* p, uint64_t seed0, int32_t n_pairs,
                      double* model_xy, double* scene_xy, int32_t* counts,
                      double* truth, double* init);

/* 256-seed global-matching hypotheses for config C5 (ranges mirror
 * global_matching/src/global_match.cpp:109-118): seed 0 = identity, seeds
 * 1.. = (U[-10,10], U[-10,10], U[0,2pi)) keyed by (pair_seed, k).
 * out: [n_seeds][3] = {x, y, yaw}. */
void scangen_hypotheses(uint64_t pair_seed, int32_t n_seeds, double* out);

/* Number of model points whose float-rounded coordinates coincide with an
 * earlier point of the same block (duplicates make kd-tree ties certain). */
int32_t scangen_count_float_duplicates(const double* xy, int32_t n);

#ifdef __cplusplus
}
#endif
#endif
