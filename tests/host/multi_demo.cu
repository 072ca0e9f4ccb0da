// multi_demo.cu -- a C++ host (no Python, no torch) driving several GPUs through the multi-device entry points of
// include/ccicp2d_b200.h: the batch is split over the device slots, the records are gathered, and both the host-buffer
// path and the device-resident path (NCCL or peer-copy gather) must reproduce the single-device call record for record.
// usage: multi_demo [devices, e.g. 0,1 or 0,0] [pairs]       (test scaffolding; built by slam_b200/build.py)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/ccicp2d_b200.h"
#include "../../slam_b200/csrc/scangen.h"

#define CHECK(x) do { int rc_ = (x); if (rc_ != 0) { fprintf(stderr, "FAIL %s -> %d (%s)\n", #x, rc_, ccicp_strerror(rc_)); return 1; } } while (0)
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "FAIL %s -> %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

static bool same_record(const ccicp_result& a, const ccicp_result& b) {
  return memcmp(a.R, b.R, sizeof(a.R)) == 0 && memcmp(a.t, b.t, sizeof(a.t)) == 0 && a.num_corr == b.num_corr &&
         a.iters == b.iters && a.status == b.status && a.score == b.score;
}

int main(int argc, char** argv) {
  std::vector<int> dev;
  if (argc > 1) { for (char* tok = strtok(argv[1], ","); tok; tok = strtok(nullptr, ",")) dev.push_back(atoi(tok)); }
  if (dev.empty()) { int n = 0; cudaGetDeviceCount(&n); if (n >= 2) dev = {0, 1}; else dev = {0, 0}; }
  const int P = argc > 2 ? atoi(argv[2]) : 97;            // odd on purpose: unequal shards
  const int G = (int)dev.size();

  scangen_params sp;
  scangen_default(&sp, 1);
  const int B = sp.n_beams;
  std::vector<double> m((size_t)P * B * 2), s((size_t)P * B * 2), truth(P * 3), init(P * 3);
  std::vector<int32_t> counts(P * 4);
  if (scangen_batch(&sp, 1000, P, m.data(), s.data(), counts.data(), truth.data(), init.data()) != 0) return 1;
  std::vector<double> mxy, sxy, Rt((size_t)P * 6);
  std::vector<int32_t> moff(2 * P + 1, 0), soff(2 * P + 1, 0);
  for (int p = 0; p < P; p++) {
    const int nm = counts[4 * p] + counts[4 * p + 1], ns = counts[4 * p + 2] + counts[4 * p + 3];
    mxy.insert(mxy.end(), m.begin() + (size_t)p * B * 2, m.begin() + (size_t)p * B * 2 + 2 * nm);
    sxy.insert(sxy.end(), s.begin() + (size_t)p * B * 2, s.begin() + (size_t)p * B * 2 + 2 * ns);
    moff[2 * p + 1] = moff[2 * p] + counts[4 * p]; moff[2 * p + 2] = moff[2 * p + 1] + counts[4 * p + 1];
    soff[2 * p + 1] = soff[2 * p] + counts[4 * p + 2]; soff[2 * p + 2] = soff[2 * p + 1] + counts[4 * p + 3];
    const double th = init[3 * p + 2];
    double* r = &Rt[(size_t)p * 6];
    r[0] = cos(th); r[1] = -sin(th); r[2] = sin(th); r[3] = cos(th); r[4] = init[3 * p]; r[5] = init[3 * p + 1];
  }
  ccicp_params prm;
  ccicp_default_params(&prm);
  prm.max_iter = 30; prm.min_delta = -2.0;

  // single device
  ccicp_ctx* ctx = nullptr;
  CHECK(ccicp_create(&ctx, dev[0]));
  std::vector<ccicp_result> ref(P), got(P);
  std::vector<int32_t> corr_ref(soff[2 * P]), corr_got(soff[2 * P], -9);
  CHECK(ccicp_register_batch(ctx, P, 1, mxy.data(), moff.data(), sxy.data(), soff.data(), Rt.data(), &prm, ref.data(), corr_ref.data()));
  CHECK(ccicp_destroy(ctx));

  // host path over G device slots
  ccicp_multi* mg = nullptr;
  CHECK(ccicp_multi_create(dev.data(), G, &mg));
  if (ccicp_multi_device_count(mg) != G) { fprintf(stderr, "FAIL device count\n"); return 1; }
  CHECK(ccicp_register_batch_multi(mg, P, 1, mxy.data(), moff.data(), sxy.data(), soff.data(), Rt.data(), &prm, got.data(), corr_got.data()));
  for (int p = 0; p < P; p++) if (!same_record(ref[p], got[p])) { fprintf(stderr, "FAIL host path: pair %d differs\n", p); return 1; }
  if (memcmp(corr_ref.data(), corr_got.data(), corr_ref.size() * 4) != 0) { fprintf(stderr, "FAIL host path: indices differ\n"); return 1; }

  // device-resident shards
  std::vector<double*> d_m(G), d_s(G), d_i(G);
  std::vector<int32_t*> d_mo(G), d_so(G);
  std::vector<ccicp_result*> d_out(G);
  std::vector<int32_t> mm(G, 0), ms(G, 0);
  for (int g = 0; g < G; g++) {
    int32_t b, e;
    ccicp_shard_range(P, G, g, &b, &e);
    const int np = e - b;
    std::vector<int32_t> lmo(2 * np + 1), lso(2 * np + 1);
    for (int k = 0; k <= 2 * np; k++) { lmo[k] = moff[2 * b + k] - moff[2 * b]; lso[k] = soff[2 * b + k] - soff[2 * b]; }
    for (int p = 0; p < np; p++) {
      mm[g] = mm[g] > lmo[2 * p + 2] - lmo[2 * p] ? mm[g] : lmo[2 * p + 2] - lmo[2 * p];
      ms[g] = ms[g] > lso[2 * p + 2] - lso[2 * p] ? ms[g] : lso[2 * p + 2] - lso[2 * p];
    }
    CU(cudaSetDevice(dev[g]));
    CU(cudaMalloc(&d_m[g], (size_t)lmo[2 * np] * 16 + 16)); CU(cudaMalloc(&d_s[g], (size_t)lso[2 * np] * 16 + 16));
    CU(cudaMalloc(&d_mo[g], lmo.size() * 4)); CU(cudaMalloc(&d_so[g], lso.size() * 4));
    CU(cudaMalloc(&d_i[g], (size_t)np * 48 + 48)); CU(cudaMalloc(&d_out[g], (size_t)P * sizeof(ccicp_result)));
    CU(cudaMemcpy(d_m[g], mxy.data() + 2 * (size_t)moff[2 * b], (size_t)lmo[2 * np] * 16, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_s[g], sxy.data() + 2 * (size_t)soff[2 * b], (size_t)lso[2 * np] * 16, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_mo[g], lmo.data(), lmo.size() * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_so[g], lso.data(), lso.size() * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_i[g], Rt.data() + (size_t)b * 6, (size_t)np * 48, cudaMemcpyHostToDevice));
    CU(cudaMemset(d_out[g], 0xEE, (size_t)P * sizeof(ccicp_result)));
  }
  for (int rep = 0; rep < 2; rep++) {
    CHECK(ccicp_register_batch_multi_device(mg, P, 1, d_m.data(), d_mo.data(), d_s.data(), d_so.data(), d_i.data(), &prm,
                                            d_out.data(), mm.data(), ms.data(), nullptr));
    CHECK(ccicp_multi_sync(mg));
  }
  for (int g = 0; g < G; g++) {
    CU(cudaSetDevice(dev[g]));
    CU(cudaMemcpy(got.data(), d_out[g], (size_t)P * sizeof(ccicp_result), cudaMemcpyDeviceToHost));
    for (int p = 0; p < P; p++) if (!same_record(ref[p], got[p])) { fprintf(stderr, "FAIL device path: slot %d, pair %d differs\n", g, p); return 1; }
  }
  const int kind = ccicp_multi_gather_kind(mg);
  printf("multi_demo OK: %d pairs over %d device slot(s) [", P, G);
  for (int g = 0; g < G; g++) printf("%s%d", g ? "," : "", dev[g]);
  printf("], host path and device path equal the single-device records; gather = %s%s%s\n",
         kind == 1 ? "nccl" : "peer copies", ccicp_multi_last_error(mg)[0] ? " ; note: " : "", ccicp_multi_last_error(mg));
  for (int g = 0; g < G; g++) { cudaSetDevice(dev[g]); cudaFree(d_m[g]); cudaFree(d_s[g]); cudaFree(d_mo[g]); cudaFree(d_so[g]); cudaFree(d_i[g]); cudaFree(d_out[g]); }
  CHECK(ccicp_multi_destroy(mg));
  return 0;
}
