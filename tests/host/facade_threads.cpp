// facade_threads.cpp -- SURVEY.md 8(b) "Threading": several host threads use the reference's class surface
// (IcpPointToPoint: construct, fit, getEdgeWeight) at the same time.  The reference itself cannot do this (matrix.cpp:26-31
// file-static scratch); the drop-in gives every host thread its own ccicp_ctx.  Each thread registers its own pairs
// `reps` times; every answer must equal the one a single thread computed first.  Test scaffolding.
// usage: facade_threads [threads] [pairs per thread] [reps]
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "ccicp2d/icpPointToPoint.h"
#include "../../slam_b200/csrc/scangen.h"

struct Answer { double t0, t1, yaw, eW[9]; int n; };

static Answer run_pair(const double* m, const int32_t* c, const double* s) {
  Matrix rot(2, 2), trans(2, 1);
  rot.val[0][0] = 1; rot.val[0][1] = 0; rot.val[1][0] = 0; rot.val[1][1] = 1;
  trans.val[0][0] = 0; trans.val[1][0] = 0;
  IcpPointToPoint icp(const_cast<double*>(m), const_cast<double*>(m) + 2 * c[0], c[0], c[1], 2);
  icp.setMaxIterations(30); icp.setMinDeltaParam(-2.0);
  icp.fit(const_cast<double*>(s), const_cast<double*>(s) + 2 * c[2], c[2], c[3], rot, trans, 5, 0);
  Answer a;
  std::memset(&a, 0, sizeof(a));           // (padding after `n`: the answers are compared with memcmp)
  a.t0 = trans.val[0][0]; a.t1 = trans.val[1][0]; a.yaw = std::atan2(rot.val[1][0], rot.val[0][0]);
  a.n = icp.getNumberCorrespondences();
  icp.getEdgeWeight(a.eW);
  return a;
}

int main(int argc, char** argv) {
  const int T = argc > 1 ? atoi(argv[1]) : 4, P = argc > 2 ? atoi(argv[2]) : 6, reps = argc > 3 ? atoi(argv[3]) : 5;
  const int N = T * P;
  scangen_params sp;
  scangen_default(&sp, 1);
  const int B = sp.n_beams;
  std::vector<double> m((size_t)N * B * 2), s((size_t)N * B * 2), truth(N * 3), init(N * 3);
  std::vector<int32_t> counts(N * 4);
  if (scangen_batch(&sp, 4242, N, m.data(), s.data(), counts.data(), truth.data(), init.data()) != 0) return 2;
  std::vector<Answer> want(N);
  for (int p = 0; p < N; p++) want[p] = run_pair(&m[(size_t)p * B * 2], &counts[4 * p], &s[(size_t)p * B * 2]);
  std::atomic<int> bad{0};
  std::vector<std::thread> th;
  for (int t = 0; t < T; t++)
    th.emplace_back([&, t] {
      for (int r = 0; r < reps; r++)
        for (int k = 0; k < P; k++) {
          const int p = t * P + k;
          const Answer a = run_pair(&m[(size_t)p * B * 2], &counts[4 * p], &s[(size_t)p * B * 2]);
          if (std::memcmp(&a, &want[p], sizeof(Answer)) != 0) bad++;
        }
    });
  for (auto& x : th) x.join();
  if (bad.load()) { std::fprintf(stderr, "FAIL: %d answers differ between threads\n", bad.load()); return 1; }
  std::printf("facade_threads OK: %d threads x %d pairs x %d reps, every answer identical to the single-thread run\n", T, P, reps);
  return 0;
}
