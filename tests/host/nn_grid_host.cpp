// Host build of the product's grid search (slam_b200/csrc/nn_grid.h): nn_search is run on the CPU over a host-built
// copy of the kernel's table layout (cells row-major, counting sort on cell id) and compared with a brute-force float
// arg-min.  Test scaffolding only.
// build: g++ -O2 -ffp-contract=off -fPIC -shared -o libnn_grid_host.so nn_grid_host.cpp
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>
#include "../../slam_b200/csrc/nn_grid.h"

using namespace ccicp;

extern "C" {

// Layout of one pair's model as build_grid leaves it.  model_xy: GA block then NGA block.  Outputs:
//   mpts[n][2] float sorted; morig[n] slot -> per-class original index; cstart[cells+2]; grid[2] as 12 floats/ints each
int nng_build(const double* model_xy, int nGA, int nNGA, float* mpts, uint16_t* morig, uint16_t* cstart,
              GridP* grid) {
  const int n = nGA + nNGA;
  std::vector<float2> pts(n > 0 ? n : 1);
  for (int i = 0; i < n; i++) { pts[i].x = (float)model_xy[2 * i]; pts[i].y = (float)model_xy[2 * i + 1]; }
  const int Gs[2] = {grid_side(nGA), grid_side(nNGA)};
  const int ncell = Gs[0] * Gs[0] + Gs[1] * Gs[1];
  for (int c = 0; c < 2; c++) {
    GridP g;
    memset(&g, 0, sizeof(g));
    const int first = c ? nGA : 0, cnt = c ? nNGA : nGA;
    g.G = Gs[c]; g.base = c ? Gs[0] * Gs[0] : 0; g.n = cnt; g.first = first;
    g.h = 1.0f; g.inv_h = 1.0f;
    if (g.G > 0) {
      float minx = pts[first].x, maxx = minx, miny = pts[first].y, maxy = miny;
      for (int i = first; i < first + cnt; i++) {
        minx = std::min(minx, pts[i].x); maxx = std::max(maxx, pts[i].x);
        miny = std::min(miny, pts[i].y); maxy = std::max(maxy, pts[i].y);
      }
      float ext = std::max(maxx - minx, maxy - miny);
      if (!(ext > 1.0e-6f)) ext = 1.0e-6f;
      g.minx = minx; g.miny = miny;
      g.cx = 0.5f * minx + 0.5f * maxx; g.cy = 0.5f * miny + 0.5f * maxy;
      g.h = ext / (float)g.G;
      g.inv_h = (float)g.G / ext;
    }
    grid[c] = g;
  }
  std::vector<int> cell(n > 0 ? n : 1, -1), cnt(ncell + 2, 0);
  for (int i = 0; i < n; i++) {
    const GridP& g = grid[i >= nGA];
    if (g.G == 0) continue;
    int ix, iy;
    cell_of(g, pts[i].x, pts[i].y, ix, iy);
    cell[i] = g.base + iy * g.G + ix;
    cnt[cell[i]]++;
  }
  std::vector<int> start(ncell + 2, 0);
  for (int c = 0; c <= ncell; c++) start[c + 1] = start[c] + cnt[c];
  std::vector<int> fill(start.begin(), start.end());
  std::vector<int> order(n > 0 ? n : 1, -1);        // slot -> global original index
  for (int i = 0; i < n; i++) if (cell[i] >= 0) order[fill[cell[i]]++] = i;
  for (int c = 0; c <= ncell; c++) cstart[c] = (uint16_t)start[c];
  for (int s = 0; s < n; s++) {
    const int i = order[s];
    if (i < 0) continue;
    mpts[2 * s] = pts[i].x; mpts[2 * s + 1] = pts[i].y;
    morig[s] = (uint16_t)(i - (i >= nGA ? nGA : 0));
  }
  return ncell;
}

// One query per entry: class cls[i], float query (qx, qy), warm slot (0xFFFF = none).
// dense != 0: the L2 variant's instantiation (doubling probe blocks + chord pruning).
// Outputs per query: slot (-1 none), dist, tie, evals (distance evaluations).
void nng_query(const float* mpts, const uint16_t* cstart, const GridP* grid, int dense, int nq, const uint8_t* cls,
               const float* qx, const float* qy, const uint16_t* warm, float gate_f, int32_t* slot, float* dist,
               uint8_t* tie, int32_t* evals) {
  const float2* mp = (const float2*)mpts;
  for (int i = 0; i < nq; i++) {
    const GridP& g = grid[cls[i]];
    float best = CUDART_INF_F; int bslot = -1, t = 0, ev = 0;
    if (g.G > 0) {
      if (dense) nn_search<true, true>(mp, cstart, g, qx[i], qy[i], gate_f, warm[i], best, bslot, t, ev);
      else nn_search<false, true>(mp, cstart, g, qx[i], qy[i], gate_f, warm[i], best, bslot, t, ev);
    }
    slot[i] = bslot; dist[i] = best; tie[i] = (uint8_t)t; evals[i] = ev;
  }
}

// Brute force over the class's slots (a skipped class occupies none, so the range comes from the table): float minimum, lowest slot reaching it, how many reach it.
void nng_brute(const float* mpts, const uint16_t* cstart, const GridP* grid, int nq, const uint8_t* cls, const float* qx, const float* qy,
               int32_t* slot, float* dist, int32_t* shared) {
  for (int i = 0; i < nq; i++) {
    const GridP& g = grid[cls[i]];
    float best = CUDART_INF_F; int bs = -1, sh = 0;
    if (g.G > 0) {
      for (int j = cstart[g.base]; j < cstart[g.base + g.G * g.G]; j++) {
        const float dx = mpts[2 * j] - qx[i], dy = mpts[2 * j + 1] - qy[i];
        const float d = dx * dx + dy * dy;
        if (d < best) { best = d; bs = j; sh = 1; }
        else if (d == best) sh++;
      }
    }
    slot[i] = bs; dist[i] = best; shared[i] = sh;
  }
}

int nng_gridp_size() { return (int)sizeof(GridP); }

}  // extern "C"
