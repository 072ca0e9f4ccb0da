/* oracle/ccicp_port.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's ccicp2d hot path, written from the
 * behaviour of the sources under /root/reference/ccicp2d/src (file:line cited
 * at each function).  It exists so that the parity tests still have a checker
 * where the compiled reference (oracle/_ref, built from the unmodified sources)
 * is absent, and as the readable statement of what "bit-exact" means here.
 * The product (slam_b200/) never links, imports or executes this file.
 *
 * Pinned: tests/test_oracle_cpu.py checks this port against oracle/_ref (same
 * seeds: kd-tree structure, nearest-neighbour answers incl. ties, per-iteration
 * R,t bit for bit) and against the committed fixtures in tests/golden/.
 *
 * Build: gcc -O3 -ffp-contract=off, x86-64 baseline (no FMA) -- the reference's
 * catkin Release flags (ccicp2d/CMakeLists.txt:18).  All float / double
 * expressions keep the reference's evaluation order.
 */
#include "ccicp_port.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ======================================================================== *
 * kd-tree  (kdtree2 as vendored: ccicp2d/src/kdtree.cpp, kdtree.h)
 * ======================================================================== */
#define KD_BUCKET 12 /* kdtree.h:132 */

typedef struct kd_node {
  int cut_dim;
  float cut_val, cut_val_left, cut_val_right;
  int l, u;
  float lo[2], hi[2]; /* box per coordinate */
  struct kd_node *left, *right;
} kd_node;

typedef struct kd_tree {
  int n;
  float* pts;  /* [n][2] original order ("the_data") */
  float* rear; /* [n][2] leaf order ("rearranged_data", kdtree.cpp:90-102) */
  int* ind;
  kd_node* root;
} kd_tree;

/* kdtree.cpp:235-269: min / max, walking two elements at a time */
static void kd_spread(const kd_tree* t, int c, int l, int u, float* lo, float* hi) {
  float smin = t->pts[2 * t->ind[l] + c], smax = smin;
  int i;
  for (i = l + 2; i <= u; i += 2) {
    float a = t->pts[2 * t->ind[i - 1] + c], b = t->pts[2 * t->ind[i] + c];
    if (a > b) { float s = a; a = b; b = s; }
    if (smin > a) smin = a;
    if (smax < b) smax = b;
  }
  if (i == u + 1) {
    float last = t->pts[2 * t->ind[u] + c];
    if (smin > last) smin = last;
    if (smax < last) smax = last;
  }
  *lo = smin; *hi = smax;
}

/* kdtree.cpp:294-317 */
static int kd_partition(kd_tree* t, int c, float alpha, int l, int u) {
  int lb = l, ub = u;
  while (lb < ub) {
    if (t->pts[2 * t->ind[lb] + c] <= alpha) lb++;
    else { int s = t->ind[lb]; t->ind[lb] = t->ind[ub]; t->ind[ub] = s; ub--; }
  }
  return (t->pts[2 * t->ind[lb] + c] <= alpha) ? lb : lb - 1;
}

/* kdtree.cpp:119-233 */
static kd_node* kd_build(kd_tree* t, int l, int u, const kd_node* parent) {
  if (u < l) return NULL;
  kd_node* nd = (kd_node*)calloc(1, sizeof(kd_node));
  if (u - l <= KD_BUCKET) {
    for (int i = 0; i < 2; i++) kd_spread(t, i, l, u, &nd->lo[i], &nd->hi[i]);
    nd->cut_dim = 0; nd->cut_val = 0.0f; nd->l = l; nd->u = u;
    return nd;
  }
  int c = -1;
  float maxspread = 0.0f;
  for (int i = 0; i < 2; i++) {
    if (parent == NULL || parent->cut_dim == i) kd_spread(t, i, l, u, &nd->lo[i], &nd->hi[i]);
    else { nd->lo[i] = parent->lo[i]; nd->hi[i] = parent->hi[i]; }
    float spread = nd->hi[i] - nd->lo[i];
    if (spread > maxspread) { maxspread = spread; c = i; }
  }
  if (c < 0) c = 0; /* > 13 coincident points: the reference reads out of bounds here */
  float sum = 0.0f;
  for (int k = l; k <= u; k++) sum += t->pts[2 * t->ind[k] + c];
  float average = sum / (float)(u - l + 1);
  int m = kd_partition(t, c, average, l, u);
  nd->cut_dim = c; nd->l = l; nd->u = u;
  nd->left = kd_build(t, l, m, nd);
  nd->right = kd_build(t, m + 1, u, nd);
  if (nd->right == NULL) {
    for (int i = 0; i < 2; i++) { nd->lo[i] = nd->left->lo[i]; nd->hi[i] = nd->left->hi[i]; }
    nd->cut_val = nd->left->hi[c];
    nd->cut_val_left = nd->cut_val_right = nd->cut_val;
  } else if (nd->left == NULL) {
    for (int i = 0; i < 2; i++) { nd->lo[i] = nd->right->lo[i]; nd->hi[i] = nd->right->hi[i]; }
    nd->cut_val = nd->right->hi[c];
    nd->cut_val_left = nd->cut_val_right = nd->cut_val;
  } else {
    nd->cut_val_right = nd->right->lo[c];
    nd->cut_val_left = nd->left->hi[c];
    nd->cut_val = (nd->cut_val_left + nd->cut_val_right) / 2.0;
    for (int i = 0; i < 2; i++) {
      nd->hi[i] = nd->left->hi[i] < nd->right->hi[i] ? nd->right->hi[i] : nd->left->hi[i]; /* std::max */
      nd->lo[i] = nd->right->lo[i] < nd->left->lo[i] ? nd->right->lo[i] : nd->left->lo[i]; /* std::min */
    }
  }
  return nd;
}

static void kd_free_nodes(kd_node* nd) {
  if (!nd) return;
  kd_free_nodes(nd->left); kd_free_nodes(nd->right); free(nd);
}

/* icp.cpp:51-60 (double -> float) + kdtree.cpp:72-106 */
static kd_tree* kd_create(const double* xy, int n) {
  kd_tree* t = (kd_tree*)calloc(1, sizeof(kd_tree));
  t->n = n;
  t->pts = (float*)malloc(sizeof(float) * 2 * (size_t)(n > 0 ? n : 1));
  t->rear = (float*)malloc(sizeof(float) * 2 * (size_t)(n > 0 ? n : 1));
  t->ind = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
  for (int i = 0; i < n; i++) { t->pts[2 * i] = (float)xy[2 * i]; t->pts[2 * i + 1] = (float)xy[2 * i + 1]; t->ind[i] = i; }
  t->root = n > 0 ? kd_build(t, 0, n - 1, NULL) : NULL;
  for (int i = 0; i < n; i++) { t->rear[2 * i] = t->pts[2 * t->ind[i]]; t->rear[2 * i + 1] = t->pts[2 * t->ind[i] + 1]; }
  return t;
}

static void kd_destroy(kd_tree* t) {
  if (!t) return;
  kd_free_nodes(t->root); free(t->pts); free(t->rear); free(t->ind); free(t);
}

typedef struct { float q[2]; float ballsize; int have; int idx; float dis; const kd_tree* t; } kd_search;

static float kd_dis_from_bnd(float x, float amin, float amax) {
  if (x > amax) return x - amax;
  else if (x < amin) return amin - x;
  return 0.0f;
}

/* kdtree.cpp:568-581 */
static int kd_box_in_range(const kd_node* nd, const kd_search* s) {
  float dis2 = 0.0f;
  for (int i = 0; i < 2; i++) {
    float d = kd_dis_from_bnd(s->q[i], nd->lo[i], nd->hi[i]);
    dis2 += d * d;
    if (dis2 > s->ballsize) return 0;
  }
  return 1;
}

/* kdtree.cpp:583-683 with nn == 1, rearranged data, no decorrelation */
static void kd_leaf(const kd_node* nd, kd_search* s) {
  float ballsize = s->ballsize;
  for (int i = nd->l; i <= nd->u; i++) {
    float dis = 0.0f;
    int early = 0;
    for (int k = 0; k < 2; k++) {
      float d = s->t->rear[2 * i + k] - s->q[k];
      dis += d * d;
      if (dis > ballsize) { early = 1; break; }
    }
    if (early) continue;
    /* first hit is pushed; every later surviving hit replaces it (dis <= ballsize) */
    s->have = 1; s->idx = s->t->ind[i]; s->dis = dis; ballsize = dis;
  }
  s->ballsize = ballsize;
}

/* kdtree.cpp:515-557 */
static void kd_search_node(const kd_node* nd, kd_search* s) {
  if (nd->left == NULL && nd->right == NULL) { kd_leaf(nd, s); return; }
  const kd_node *closer, *farther;
  float extra, qval = s->q[nd->cut_dim];
  if (qval < nd->cut_val) { closer = nd->left; farther = nd->right; extra = nd->cut_val_right - qval; }
  else { closer = nd->right; farther = nd->left; extra = qval - nd->cut_val_left; }
  if (closer) kd_search_node(closer, s);
  if (farther && extra * extra < s->ballsize)
    if (kd_box_in_range(farther, s)) kd_search_node(farther, s);
}

/* kdtree.cpp:378-391; returns 0 when nothing lies within 1e38 */
static int kd_nearest(const kd_tree* t, float qx, float qy, int* idx, float* dis) {
  kd_search s;
  s.q[0] = qx; s.q[1] = qy; s.ballsize = 1.0e38f; s.have = 0; s.idx = -1; s.dis = 0.0f; s.t = t;
  if (t->root) kd_search_node(t->root, &s);
  *idx = s.idx; *dis = s.dis;
  return s.have;
}

/* ======================================================================== *
 * 2x2 SVD  (Matrix::svd, ccicp2d/src/matrix.cpp:582-810; pythag :848-856)
 * Numerical-Recipes svdcmp specialised to m = n = 2; every expression keeps
 * the reference's order so U and V come out bit for bit.
 * ======================================================================== */
static double nr_sign(double a, double b) { return b >= 0.0 ? fabs(a) : -fabs(a); }
static double nr_sqr(double a) { return a == 0.0 ? 0.0 : a * a; }
static double nr_pythag(double a, double b) {
  double absa = fabs(a), absb = fabs(b);
  if (absa > absb) return absa * sqrt(1.0 + nr_sqr(absb / absa));
  return absb == 0.0 ? 0.0 : absb * sqrt(1.0 + nr_sqr(absa / absb));
}

void ccport_svd2(const double* H, double* Uo, double* Wo, double* Vo) {
  enum { N = 2 };
  double U[N][N] = {{H[0], H[1]}, {H[2], H[3]}}, V[N][N] = {{0, 0}, {0, 0}}, w[N], rv1[N];
  int flag, i, its, j, jj, k, l = 0, nm = 0;
  double anorm, c, f, g, h, s, scale, x, y, z;
  g = scale = anorm = 0.0;
  for (i = 0; i < N; i++) { /* Householder reduction to bidiagonal form (:597-646) */
    l = i + 1;
    rv1[i] = scale * g;
    g = s = scale = 0.0;
    for (k = i; k < N; k++) scale += fabs(U[k][i]);
    if (scale) {
      for (k = i; k < N; k++) { U[k][i] /= scale; s += U[k][i] * U[k][i]; }
      f = U[i][i];
      g = -nr_sign(sqrt(s), f);
      h = f * g - s;
      U[i][i] = f - g;
      for (j = l; j < N; j++) {
        for (s = 0.0, k = i; k < N; k++) s += U[k][i] * U[k][j];
        f = s / h;
        for (k = i; k < N; k++) U[k][j] += f * U[k][i];
      }
      for (k = i; k < N; k++) U[k][i] *= scale;
    }
    w[i] = scale * g;
    g = s = scale = 0.0;
    if (i != N - 1) {
      for (k = l; k < N; k++) scale += fabs(U[i][k]);
      if (scale) {
        for (k = l; k < N; k++) { U[i][k] /= scale; s += U[i][k] * U[i][k]; }
        f = U[i][l];
        g = -nr_sign(sqrt(s), f);
        h = f * g - s;
        U[i][l] = f - g;
        for (k = l; k < N; k++) rv1[k] = U[i][k] / h;
        for (j = l; j < N; j++) {
          for (s = 0.0, k = l; k < N; k++) s += U[j][k] * U[i][k];
          for (k = l; k < N; k++) U[j][k] += s * rv1[k];
        }
        for (k = l; k < N; k++) U[i][k] *= scale;
      }
    }
    { double a1 = anorm, a2 = fabs(w[i]) + fabs(rv1[i]); anorm = a1 > a2 ? a1 : a2; }
  }
  for (i = N - 1; i >= 0; i--) { /* right-hand transformations (:647-662) */
    if (i < N - 1) {
      if (g) {
        for (j = l; j < N; j++) V[j][i] = (U[i][j] / U[i][l]) / g;
        for (j = l; j < N; j++) {
          for (s = 0.0, k = l; k < N; k++) s += U[i][k] * V[k][j];
          for (k = l; k < N; k++) V[k][j] += s * V[k][i];
        }
      }
      for (j = l; j < N; j++) V[i][j] = V[j][i] = 0.0;
    }
    V[i][i] = 1.0;
    g = rv1[i];
    l = i;
  }
  for (i = N - 1; i >= 0; i--) { /* left-hand transformations (:663-678) */
    l = i + 1;
    g = w[i];
    for (j = l; j < N; j++) U[i][j] = 0.0;
    if (g) {
      g = 1.0 / g;
      for (j = l; j < N; j++) {
        for (s = 0.0, k = l; k < N; k++) s += U[k][i] * U[k][j];
        f = (s / U[i][i]) * g;
        for (k = i; k < N; k++) U[k][j] += f * U[k][i];
      }
      for (j = i; j < N; j++) U[j][i] *= g;
    } else for (j = i; j < N; j++) U[j][i] = 0.0;
    ++U[i][i];
  }
  for (k = N - 1; k >= 0; k--) { /* diagonalisation of the bidiagonal form (:679-765) */
    for (its = 0; its < 30; its++) {
      flag = 1;
      for (l = k; l >= 0; l--) {
        nm = l - 1;
        if ((double)(fabs(rv1[l]) + anorm) == anorm) { flag = 0; break; }
        if ((double)(fabs(w[nm]) + anorm) == anorm) break;   /* rv1[0] is always 0, so nm >= 0 here */
      }
      if (flag) {
        c = 0.0; s = 1.0;
        for (i = l; i <= k; i++) {
          f = s * rv1[i];
          rv1[i] = c * rv1[i];
          if ((double)(fabs(f) + anorm) == anorm) break;
          g = w[i];
          h = nr_pythag(f, g);
          w[i] = h;
          h = 1.0 / h;
          c = g * h;
          s = -f * h;
          for (j = 0; j < N; j++) {
            y = U[j][nm]; z = U[j][i];
            U[j][nm] = y * c + z * s;
            U[j][i] = z * c - y * s;
          }
        }
      }
      z = w[k];
      if (l == k) {
        if (z < 0.0) { w[k] = -z; for (j = 0; j < N; j++) V[j][k] = -V[j][k]; }
        break;
      }
      x = w[l];
      nm = k - 1;
      y = w[nm];
      g = rv1[nm];
      h = rv1[k];
      f = ((y - z) * (y + z) + (g - h) * (g + h)) / (2.0 * h * y);
      g = nr_pythag(f, 1.0);
      f = ((x - z) * (x + z) + h * ((y / (f + nr_sign(g, f))) - h)) / x;
      c = s = 1.0;
      for (j = l; j <= nm; j++) {
        i = j + 1;
        g = rv1[i];
        y = w[i];
        h = s * g;
        g = c * g;
        z = nr_pythag(f, h);
        rv1[j] = z;
        c = f / z;
        s = h / z;
        f = x * c + g * s;
        g = g * c - x * s;
        h = y * s;
        y *= c;
        for (jj = 0; jj < N; jj++) {
          x = V[jj][j]; z = V[jj][i];
          V[jj][j] = x * c + z * s;
          V[jj][i] = z * c - x * s;
        }
        z = nr_pythag(f, h);
        w[j] = z;
        if (z) { z = 1.0 / z; c = f * z; s = h * z; }
        f = c * g + s * y;
        x = c * y - s * g;
        for (jj = 0; jj < N; jj++) {
          y = U[jj][j]; z = U[jj][i];
          U[jj][j] = y * c + z * s;
          U[jj][i] = z * c - y * s;
        }
      }
      rv1[l] = 0.0;
      rv1[k] = f;
      w[k] = x;
    }
  }
  /* sort by decreasing singular value (insertion sort with inc = 1 for n = 2, :767-790) */
  if (w[0] < w[1]) {
    double sw = w[1]; w[1] = w[0]; w[0] = sw;
    for (k = 0; k < N; k++) { double a = U[k][1]; U[k][1] = U[k][0]; U[k][0] = a; a = V[k][1]; V[k][1] = V[k][0]; V[k][0] = a; }
  }
  for (k = 0; k < N; k++) { /* sign flips (:791-799) */
    int s2 = 0;
    for (i = 0; i < N; i++) if (U[i][k] < 0.0) s2++;
    for (j = 0; j < N; j++) if (V[j][k] < 0.0) s2++;
    if (s2 > (N + N) / 2) {
      for (i = 0; i < N; i++) U[i][k] = -U[i][k];
      for (j = 0; j < N; j++) V[j][k] = -V[j][k];
    }
  }
  for (i = 0; i < N; i++) for (j = 0; j < N; j++) { Uo[2 * i + j] = U[i][j]; Vo[2 * i + j] = V[i][j]; }
  Wo[0] = w[0]; Wo[1] = w[1];
}

/* ======================================================================== *
 * Icp / IcpPointToPoint
 * ======================================================================== */
typedef struct {
  kd_tree *tGA, *tNGA;
  int nmGA, nmNGA;
  int num_corr;          /* numCorr (icpPointToPoint.cpp:141) */
  int nc_saved;          /* rows of cModelPoints / cScenePoints */
  double *cm, *cs;       /* saved correspondences [n][2] (icpPointToPoint.cpp:145-146) */
} icp_state;

/* One class block of fitStep (icpPointToPoint.cpp:59-91 / 93-124). */
static void fit_class(const kd_tree* tree, int model_n, const double* T, int n, const double* Rt, double inDist,
                      double* pm, double* pt, double* mu_m, double* mu_t, int* inIDX,
                      int32_t* corr, float* dis_out, uint8_t* tie_out) {
  for (int i = 0; i < n; i++) {
    if (model_n <= 3) {
      if (corr) corr[i] = -1;
      if (dis_out) dis_out[i] = -1.0f;
      if (tie_out) tie_out[i] = 0;
      continue;
    }
    float qx = Rt[0] * T[i * 2 + 0] + Rt[1] * T[i * 2 + 1] + Rt[4];
    float qy = Rt[2] * T[i * 2 + 0] + Rt[3] * T[i * 2 + 1] + Rt[5];
    int idx; float dis;
    kd_nearest(tree, qx, qy, &idx, &dis);
    int keep = dis < inDist;
    if (corr) corr[i] = keep ? idx : -1;
    if (dis_out) dis_out[i] = dis;
    if (tie_out) {
      int shared = 0;
      for (int m = 0; m < model_n; m++) {
        float dx = tree->pts[2 * m] - qx, dy = tree->pts[2 * m + 1] - qy;
        float d = dx * dx;
        d += dy * dy;
        if (d == dis) shared++;
      }
      tie_out[i] = shared > 1;
    }
    if (keep) {
      int k = *inIDX;
      pm[2 * k] = tree->pts[2 * idx]; mu_m[0] += pm[2 * k];
      pm[2 * k + 1] = tree->pts[2 * idx + 1]; mu_m[1] += pm[2 * k + 1];
      pt[2 * k] = qx; mu_t[0] += pt[2 * k];
      pt[2 * k + 1] = qy; mu_t[1] += pt[2 * k + 1];
      *inIDX = k + 1;
    }
  }
}

/* IcpPointToPoint::fitStep (icpPointToPoint.cpp:33-172).  Rt = {r00,r01,r10,r11,t0,t1} in/out. */
static double fit_step(icp_state* st, const double* TGA, const double* TNGA, int nGA, int nNGA, double* Rt,
                       double inDist, int32_t* corr, float* dis, uint8_t* tie) {
  int total = nGA + nNGA, inIDX = 0;
  double* pm = (double*)calloc(2 * (size_t)(total > 0 ? total : 1), sizeof(double));
  double* pt = (double*)calloc(2 * (size_t)(total > 0 ? total : 1), sizeof(double));
  double mu_m[2] = {0, 0}, mu_t[2] = {0, 0};
  fit_class(st->tGA, st->nmGA, TGA, nGA, Rt, inDist, pm, pt, mu_m, mu_t, &inIDX, corr, dis, tie);
  fit_class(st->tNGA, st->nmNGA, TNGA, nNGA, Rt, inDist, pm, pt, mu_m, mu_t, &inIDX,
            corr ? corr + nGA : NULL, dis ? dis + nGA : NULL, tie ? tie + nGA : NULL);
  if (inIDX == 0) { free(pm); free(pt); return -1; }
  st->num_corr = inIDX;
  free(st->cm); free(st->cs);
  st->cm = pm; st->cs = pt; st->nc_saved = inIDX;
  mu_m[0] = mu_m[0] / (double)inIDX; mu_m[1] = mu_m[1] / (double)inIDX;
  mu_t[0] = mu_t[0] / (double)inIDX; mu_t[1] = mu_t[1] / (double)inIDX;
  /* H = ~q_t * q_m with q = p - ones*mu, k innermost and sequential (matrix.cpp:276-279) */
  double H[4] = {0, 0, 0, 0};
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 2; j++)
      for (int k = 0; k < inIDX; k++) H[2 * i + j] += (pt[2 * k + i] - mu_t[i]) * (pm[2 * k + j] - mu_m[j]);
  double U[4], W[2], V[4];
  ccport_svd2(H, U, W, V);
  double R_[4], t_[2];
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 2; j++) { double c = 0.0; for (int k = 0; k < 2; k++) c += V[2 * i + k] * U[2 * j + k]; R_[2 * i + j] = c; }
  for (int i = 0; i < 2; i++) { double c = 0.0; for (int k = 0; k < 2; k++) c += R_[2 * i + k] * mu_t[k]; t_[i] = mu_m[i] - c; }
  double Rn[4], tn[2];
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 2; j++) { double c = 0.0; for (int k = 0; k < 2; k++) c += R_[2 * i + k] * Rt[2 * k + j]; Rn[2 * i + j] = c; }
  for (int i = 0; i < 2; i++) { double c = 0.0; for (int k = 0; k < 2; k++) c += R_[2 * i + k] * Rt[4 + k]; tn[i] = c + t_[i]; }
  memcpy(Rt, Rn, sizeof(Rn)); Rt[4] = tn[0]; Rt[5] = tn[1];
  double nr = 0, nt = 0;
  for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) { double v = R_[2 * i + j] - (i == j ? 1.0 : 0.0); nr += v * v; }
  for (int i = 0; i < 2; i++) nt += t_[i] * t_[i];
  nr = sqrt(nr); nt = sqrt(nt);
  return nr < nt ? nt : nr;    /* std::max */
}

/* Matrix::solve with B = I (Gauss-Jordan, full pivoting; matrix.cpp:420-508) for a 3x3 */
static int gj_inverse3(double A[3][3], double B[3][3]) {
  int indxc[3], indxr[3], ipiv[3] = {0, 0, 0}, icol = 0, irow = 0;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) B[i][j] = i == j ? 1.0 : 0.0;
  for (int i = 0; i < 3; i++) {
    double big = 0.0;
    for (int j = 0; j < 3; j++)
      if (ipiv[j] != 1)
        for (int k = 0; k < 3; k++)
          if (ipiv[k] == 0 && fabs(A[j][k]) >= big) { big = fabs(A[j][k]); irow = j; icol = k; }
    ++ipiv[icol];
    if (irow != icol)
      for (int l = 0; l < 3; l++) {
        double tmp = A[irow][l]; A[irow][l] = A[icol][l]; A[icol][l] = tmp;
        tmp = B[irow][l]; B[irow][l] = B[icol][l]; B[icol][l] = tmp;
      }
    indxr[i] = irow; indxc[i] = icol;
    if (fabs(A[icol][icol]) < 1e-20) return 0;
    double pivinv = 1.0 / A[icol][icol];
    A[icol][icol] = 1.0;
    for (int l = 0; l < 3; l++) A[icol][l] *= pivinv;
    for (int l = 0; l < 3; l++) B[icol][l] *= pivinv;
    for (int ll = 0; ll < 3; ll++)
      if (ll != icol) {
        double dum = A[ll][icol];
        A[ll][icol] = 0.0;
        for (int l = 0; l < 3; l++) A[ll][l] -= A[icol][l] * dum;
        for (int l = 0; l < 3; l++) B[ll][l] -= B[icol][l] * dum;
      }
  }
  (void)indxr; (void)indxc;   /* the column unscrambling only touches A, which is discarded */
  return 1;
}

/* IcpPointToPoint::getEdgeWeight (icpPointToPoint.cpp:233-316), dy bug included (:262) */
static void edge_weight(const icp_state* st, double* eW) {
  int n = st->nc_saved;
  double sx = 0, sy = 0, xpy = 0, MZ[3] = {0, 0, 0};
  for (int i = 0; i < n; i++) {
    double ax = st->cm[2 * i], ay = st->cm[2 * i + 1], bx = st->cs[2 * i], by = st->cs[2 * i + 1];
    double x = (ax + bx) / 2.0, y = (ay + by) / 2.0;
    double dx = ax - bx, dy = ax - bx;
    sx += x; sy += y; xpy += x * x + y * y;
    MZ[0] += dx; MZ[1] += dy; MZ[2] += -y * dx + x * dy;
  }
  double MM[3][3] = {{(double)n, 0, -sy}, {0, (double)n, sx}, {-sy, sx, xpy}}, A[3][3], Mi[3][3];
  memcpy(A, MM, sizeof(A));
  if (!gj_inverse3(A, Mi)) { /* Matrix::inv leaves the matrix = eye when solve fails part-way; keep what we have */ }
  double D[3];
  for (int i = 0; i < 3; i++) { double c = 0.0; for (int k = 0; k < 3; k++) c += Mi[i][k] * MZ[k]; D[i] = c; }
  double ss = 0;
  for (int i = 0; i < n; i++) {
    double ax = st->cm[2 * i], ay = st->cm[2 * i + 1], bx = st->cs[2 * i], by = st->cs[2 * i + 1];
    double x = (ax + bx) / 2.0, y = (ay + by) / 2.0;
    double tx = (ax - bx - D[0] + y * D[2]), ty = (ay - by - D[1] - x * D[2]);
    ss += tx * tx + ty * ty;
  }
  ss = ss / (2 * n - 3);
  double sconst = (1.0 / ss);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) eW[3 * i + j] = MM[i][j] * sconst;
}

static const double dummy_pt[2] = {0, 0};
static const double* or_dummy(const double* p) { return p ? p : dummy_pt; }

static icp_state* icp_create(const double* mGA, int nmGA, const double* mNGA, int nmNGA) {
  icp_state* st = (icp_state*)calloc(1, sizeof(icp_state));
  st->nmGA = nmGA; st->nmNGA = nmNGA;
  st->tGA = kd_create(or_dummy(mGA), nmGA);     /* both trees are built unconditionally (icp.cpp:65,69) */
  st->tNGA = kd_create(or_dummy(mNGA), nmNGA);
  return st;
}
static void icp_destroy(icp_state* st) {
  kd_destroy(st->tGA); kd_destroy(st->tNGA); free(st->cm); free(st->cs); free(st);
}

int ccport_fit(const double* mGA, int nmGA, const double* mNGA, int nmNGA, const double* sGA, int nsGA,
               const double* sNGA, int nsNGA, int max_iter, double min_delta, double indist, double* Rt,
               int* num_corr, double* eW) {
  if (num_corr) *num_corr = 0;
  if (nmGA + nmNGA < 5) return CCPORT_TOO_FEW_MODEL;      /* icp.cpp:37-43 */
  if (nsGA + nsNGA < 5) return CCPORT_TOO_FEW_SCENE;      /* icp.cpp:100-103 */
  icp_state* st = icp_create(mGA, nmGA, mNGA, nmNGA);
  for (int it = 0; it < max_iter; it++)                   /* icp.cpp:116-122 */
    if (fit_step(st, or_dummy(sGA), or_dummy(sNGA), nsGA, nsNGA, Rt, indist, NULL, NULL, NULL) < min_delta) break;
  if (num_corr) *num_corr = st->num_corr;
  int status = CCPORT_OK;
  if (st->nc_saved == 0) status = CCPORT_NO_CORR;
  else if (eW) edge_weight(st, eW);
  icp_destroy(st);
  return status;
}

int ccport_trace(const double* mGA, int nmGA, const double* mNGA, int nmNGA, const double* sGA, int nsGA,
                 const double* sNGA, int nsNGA, int max_iter, double min_delta, double indist, const double* Rt0,
                 double* Rt_trace, int32_t* corr, float* dis, uint8_t* tie, int32_t* ncorr, double* delta,
                 int* iters_run, double* eW) {
  *iters_run = 0;
  memcpy(Rt_trace, Rt0, 6 * sizeof(double));
  if (nmGA + nmNGA < 5) return CCPORT_TOO_FEW_MODEL;
  if (nsGA + nsNGA < 5) return CCPORT_TOO_FEW_SCENE;
  icp_state* st = icp_create(mGA, nmGA, mNGA, nmNGA);
  const int ns = nsGA + nsNGA;
  double Rt[6];
  memcpy(Rt, Rt0, sizeof(Rt));
  int k = 0;
  for (; k < max_iter; k++) {
    memcpy(Rt_trace + 6 * k, Rt, sizeof(Rt));
    double d = fit_step(st, or_dummy(sGA), or_dummy(sNGA), nsGA, nsNGA, Rt, indist, corr + (size_t)k * ns,
                        dis ? dis + (size_t)k * ns : NULL, tie ? tie + (size_t)k * ns : NULL);
    delta[k] = d;
    int nc = 0;
    for (int i = 0; i < ns; i++) nc += corr[(size_t)k * ns + i] >= 0;
    ncorr[k] = nc;
    if (d < min_delta) { k++; break; }
  }
  *iters_run = k;
  memcpy(Rt_trace + 6 * k, Rt, sizeof(Rt));
  int status = CCPORT_OK;
  if (st->nc_saved == 0) status |= CCPORT_NO_CORR;
  else if (eW) edge_weight(st, eW);
  icp_destroy(st);
  return status;
}

static void dump_node(const kd_node* nd, float* rec, int* n, int max_nodes) {
  int leaf = nd->left == NULL && nd->right == NULL;
  if (*n < max_nodes) {
    float* r = rec + 10 * (size_t)*n;
    r[0] = leaf ? -1.0f : (float)nd->cut_dim; r[1] = nd->cut_val;
    r[2] = leaf ? 0.0f : nd->cut_val_left; r[3] = leaf ? 0.0f : nd->cut_val_right;
    r[4] = (float)nd->l; r[5] = (float)nd->u;
    r[6] = nd->lo[0]; r[7] = nd->hi[0]; r[8] = nd->lo[1]; r[9] = nd->hi[1];
  }
  (*n)++;
  if (nd->left) dump_node(nd->left, rec, n, max_nodes);
  if (nd->right) dump_node(nd->right, rec, n, max_nodes);
}

int ccport_tree(const double* pts, int n, int32_t* ind_out, float* nodes_out, int max_nodes) {
  if (n <= 0) return 0;
  kd_tree* t = kd_create(pts, n);
  for (int i = 0; i < n; i++) ind_out[i] = t->ind[i];
  int nn = 0;
  dump_node(t->root, nodes_out, &nn, max_nodes);
  kd_destroy(t);
  return nn > max_nodes ? -nn : nn;
}

void ccport_nn(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis) {
  kd_tree* t = kd_create(pts, n);
  for (int i = 0; i < nq; i++) { int id; float d; kd_nearest(t, q[2 * i], q[2 * i + 1], &id, &d); idx[i] = id; dis[i] = d; }
  kd_destroy(t);
}
