/* scangen.c -- see scangen.h.  Plain C11 + libm (+ OpenMP over pairs). */
#include "scangen.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define SG_PI 3.14159265358979323846
#define SG_MAX_SEG 64

typedef struct { double ax, ay, bx, by; } sg_seg;
typedef struct { sg_seg seg[SG_MAX_SEG]; int n; } sg_world;

/* ---- counter-based RNG ------------------------------------------------- */
static inline uint64_t sg_mix(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static inline uint64_t sg_key(uint64_t seed, uint64_t stream, uint64_t ctr) {
  return sg_mix(sg_mix(sg_mix(seed) ^ (stream * 0xD6E8FEB86659FD93ull)) + ctr);
}
/* uniform in (0,1): 53 random bits, never exactly 0 */
static inline double sg_u01(uint64_t seed, uint64_t stream, uint64_t ctr) {
  return ((double)(sg_key(seed, stream, ctr) >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}
static inline double sg_uni(uint64_t seed, uint64_t stream, uint64_t ctr, double lo, double hi) {
  return lo + (hi - lo) * sg_u01(seed, stream, ctr);
}
static inline double sg_gauss(uint64_t seed, uint64_t stream, uint64_t ctr) {
  double u1 = sg_u01(seed, stream, 2 * ctr);
  double u2 = sg_u01(seed, stream, 2 * ctr + 1);
  return sqrt(-2.0 * log(u1)) * cos(2.0 * SG_PI * u2);
}

enum { ST_WORLD = 1, ST_TRUTH = 2, ST_INIT = 3, ST_NOISE_M = 4, ST_NOISE_S = 5, ST_HYP = 6 };

/* ---- world ------------------------------------------------------------- */
static void sg_push(sg_world* w, double ax, double ay, double bx, double by) {
  if (w->n < SG_MAX_SEG) {
    w->seg[w->n].ax = ax; w->seg[w->n].ay = ay;
    w->seg[w->n].bx = bx; w->seg[w->n].by = by;
    w->n++;
  }
}

static void sg_build_world(const scangen_params* p, uint64_t seed, sg_world* w) {
  uint64_t c = 0;
  w->n = 0;
  int nv = 8 + (int)(sg_u01(seed, ST_WORLD, c++) * 7.0);  /* 8..14 */
  if (nv > 14) nv = 14;
  double rb = sg_uni(seed, ST_WORLD, c++, p->room_rmin, p->room_rmax);
  double vx[16], vy[16];
  for (int k = 0; k < nv; k++) {
    double a = 2.0 * SG_PI * ((double)k + sg_uni(seed, ST_WORLD, c++, -0.3, 0.3)) / (double)nv;
    double r = rb * sg_uni(seed, ST_WORLD, c++, 0.8, 1.0);
    vx[k] = r * cos(a); vy[k] = r * sin(a);
  }
  for (int k = 0; k < nv; k++) {
    int k2 = (k + 1) % nv;
    sg_push(w, vx[k], vy[k], vx[k2], vy[k2]);
  }
  int nb = (int)(sg_u01(seed, ST_WORLD, c++) * (double)(p->max_boxes + 1));
  if (nb > p->max_boxes) nb = p->max_boxes;
  /* keep obstacles clear of both sensor positions (origin and origin+offset) */
  double keepout = 1.0 + 1.5 * p->off_xy;
  for (int b = 0; b < 4; b++) {           /* always draw 4 boxes' numbers: stable counters */
    double d  = sg_uni(seed, ST_WORLD, c++, 0.35, 0.7) * rb;
    double a  = sg_uni(seed, ST_WORLD, c++, 0.0, 2.0 * SG_PI);
    double hw = sg_uni(seed, ST_WORLD, c++, 0.15, 0.75);
    double hh = sg_uni(seed, ST_WORLD, c++, 0.15, 0.75);
    double ro = sg_uni(seed, ST_WORLD, c++, 0.0, SG_PI);
    if (b >= nb) continue;
    if (d - 1.1 < keepout) continue;      /* too close to a sensor: skip this box */
    double cx = d * cos(a), cy = d * sin(a), cr = cos(ro), sr = sin(ro);
    double px[4] = { -hw, hw, hw, -hw }, py[4] = { -hh, -hh, hh, hh };
    double qx[4], qy[4];
    for (int k = 0; k < 4; k++) {
      qx[k] = cx + cr * px[k] - sr * py[k];
      qy[k] = cy + sr * px[k] + cr * py[k];
    }
    for (int k = 0; k < 4; k++) sg_push(w, qx[k], qy[k], qx[(k + 1) & 3], qy[(k + 1) & 3]);
  }
}

/* nearest hit range along a ray, or +inf */
static double sg_cast(const sg_world* w, double ox, double oy, double dx, double dy) {
  double best = INFINITY;
  for (int s = 0; s < w->n; s++) {
    double ex = w->seg[s].bx - w->seg[s].ax, ey = w->seg[s].by - w->seg[s].ay;
    double den = dx * ey - dy * ex;
    if (fabs(den) < 1e-12) continue;
    double fx = w->seg[s].ax - ox, fy = w->seg[s].ay - oy;
    double tr = (fx * ey - fy * ex) / den;   /* along the ray     */
    double ts = (fx * dy - fy * dx) / den;   /* along the segment */
    if (tr > 1e-6 && ts >= 0.0 && ts <= 1.0 && tr < best) best = tr;
  }
  return best;
}

/* One scan from sensor pose (sx,sy,sth): writes compacted GA-then-NGA points in
 * the sensor frame.  ranges/flags are caller scratch of n_beams entries. */
static void sg_scan(const scangen_params* p, const sg_world* w, uint64_t seed, uint64_t noise_stream,
                    double sx, double sy, double sth, double* ranges, unsigned char* ga,
                    double* out_xy, int32_t* n_ga, int32_t* n_nga) {
  const int nb = p->n_beams;
  for (int i = 0; i < nb; i++) {
    double a = (p->fov_start_deg + p->fov_step_deg * (double)i) * (SG_PI / 180.0);
    double r = sg_cast(w, sx, sy, cos(a + sth), sin(a + sth));
    if (r < INFINITY) r += p->noise_sigma * sg_gauss(seed, noise_stream, (uint64_t)i);
    if (!(r > 0.05) || r > p->max_range) r = INFINITY;   /* dropped beam */
    ranges[i] = r;
  }
  for (int i = 0; i < nb; i++) {
    unsigned char g = 0;
    if (p->class_rule == 0) {
      if (i > 0 && !(fabs(ranges[i] - ranges[i - 1]) <= p->jump_thresh)) g = 1;
      if (i + 1 < nb && !(fabs(ranges[i] - ranges[i + 1]) <= p->jump_thresh)) g = 1;
    } else if (p->class_rule == 2) {
      g = 1;
    }
    ga[i] = g;
  }
  int32_t k = 0;
  for (int pass = 1; pass >= 0; pass--) {     /* pass 1: GA block, pass 0: NGA block */
    for (int i = 0; i < nb; i++) {
      if (ranges[i] == INFINITY || ga[i] != (unsigned char)pass) continue;
      double a = (p->fov_start_deg + p->fov_step_deg * (double)i) * (SG_PI / 180.0);
      out_xy[2 * k] = ranges[i] * cos(a);
      out_xy[2 * k + 1] = ranges[i] * sin(a);
      k++;
    }
    if (pass == 1) *n_ga = k;
  }
  *n_nga = k - *n_ga;
}

void scangen_default(scangen_params* p, int32_t config) {
  memset(p, 0, sizeof(*p));
  p->n_beams = 1081; p->max_boxes = 4; p->class_rule = 0;
  p->fov_start_deg = -135.0; p->fov_step_deg = 0.25;
  p->noise_sigma = 0.01; p->max_range = 30.0;
  p->room_rmin = 4.0; p->room_rmax = 15.0;
  p->off_xy = 0.5; p->off_th = 0.1;
  p->init_sigma_xy = 0.0; p->init_sigma_th = 0.0;
  p->jump_thresh = 0.04;
  if (config == 3) {            /* loop-closure candidates: larger offsets, noisy init */
    p->room_rmin = 7.0;
    p->off_xy = 2.0; p->off_th = 0.3;
    p->init_sigma_xy = 0.3; p->init_sigma_th = 0.05;
  } else if (config == 4) {     /* dense 360 degree slices */
    p->n_beams = 32768; p->fov_start_deg = -180.0; p->fov_step_deg = 360.0 / 32768.0;
    p->noise_sigma = 0.005;
  }
}

int32_t scangen_batch(const scangen_params* p, uint64_t seed0, int32_t n_pairs,
                      double* model_xy, double* scene_xy, int32_t* counts,
                      double* truth, double* init) {
  if (!p || p->n_beams <= 0 || n_pairs < 0 || !model_xy || !scene_xy || !counts) return -1;
  const size_t stride = (size_t)p->n_beams * 2;
#pragma omp parallel
  {
    double* ranges = (double*)malloc(sizeof(double) * (size_t)p->n_beams);
    unsigned char* ga = (unsigned char*)malloc((size_t)p->n_beams);
    sg_world w;
#pragma omp for schedule(dynamic, 16)
    for (int32_t i = 0; i < n_pairs; i++) {
      uint64_t seed = seed0 + (uint64_t)i;
      sg_build_world(p, seed, &w);
      double tx = sg_uni(seed, ST_TRUTH, 0, -p->off_xy, p->off_xy);
      double ty = sg_uni(seed, ST_TRUTH, 1, -p->off_xy, p->off_xy);
      double th = sg_uni(seed, ST_TRUTH, 2, -p->off_th, p->off_th);
      sg_scan(p, &w, seed, ST_NOISE_M, 0.0, 0.0, 0.0, ranges, ga,
              model_xy + stride * i, &counts[4 * i + 0], &counts[4 * i + 1]);
      sg_scan(p, &w, seed, ST_NOISE_S, tx, ty, th, ranges, ga,
              scene_xy + stride * i, &counts[4 * i + 2], &counts[4 * i + 3]);
      if (truth) { truth[3 * i] = tx; truth[3 * i + 1] = ty; truth[3 * i + 2] = th; }
      if (init) {
        if (p->init_sigma_xy > 0.0 || p->init_sigma_th > 0.0) {
          init[3 * i]     = tx + p->init_sigma_xy * sg_gauss(seed, ST_INIT, 0);
          init[3 * i + 1] = ty + p->init_sigma_xy * sg_gauss(seed, ST_INIT, 1);
          init[3 * i + 2] = th + p->init_sigma_th * sg_gauss(seed, ST_INIT, 2);
        } else {
          init[3 * i] = init[3 * i + 1] = init[3 * i + 2] = 0.0;
        }
      }
    }
    free(ranges);
    free(ga);
  }
  return 0;
}

void scangen_hypotheses(uint64_t pair_seed, int32_t n_seeds, double* out) {
  for (int32_t k = 0; k < n_seeds; k++) {
    if (k == 0) { out[0] = out[1] = out[2] = 0.0; continue; }
    out[3 * k]     = sg_uni(pair_seed, ST_HYP, 3 * (uint64_t)k, -10.0, 10.0);
    out[3 * k + 1] = sg_uni(pair_seed, ST_HYP, 3 * (uint64_t)k + 1, -10.0, 10.0);
    out[3 * k + 2] = sg_uni(pair_seed, ST_HYP, 3 * (uint64_t)k + 2, 0.0, 2.0 * SG_PI);
  }
}

static int sg_cmp_u64(const void* a, const void* b) {
  uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
  return (x > y) - (x < y);
}

int32_t scangen_count_float_duplicates(const double* xy, int32_t n) {
  if (n <= 1) return 0;
  uint64_t* keys = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
  for (int32_t i = 0; i < n; i++) {
    float fx = (float)xy[2 * i] + 0.0f, fy = (float)xy[2 * i + 1] + 0.0f;  /* -0 -> +0 */
    uint32_t bx, by;
    memcpy(&bx, &fx, 4); memcpy(&by, &fy, 4);
    keys[i] = ((uint64_t)bx << 32) | by;
  }
  qsort(keys, (size_t)n, sizeof(uint64_t), sg_cmp_u64);
  int32_t dup = 0;
  for (int32_t i = 1; i < n; i++) dup += (keys[i] == keys[i - 1]);
  free(keys);
  return dup;
}
