// fp32_peak.cu -- measured issue peak of UN-FUSED FP32 (FMUL / FADD, no FMA) on every SM: the denominator of
// bench.py's roofline_useful ("148 SM x 128 lanes x 1.965 GHz = 37.2 T un-fused FP32 op/s; measure the
// real peak with a micro-benchmark").  The distance expression of the NN search is fl(fl(dx*dx) + fl(dy*dy)) with
// dx = fl(m - q): FADD/FMUL only, so FMA peak (2 flops per lane-cycle) is not reachable by construction.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o profiles/fp32_peak profiles/fp32_peak.cu
// run  : profiles/fp32_peak > profiles/fp32_unfused_peak.json
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int kAcc = 8;          // independent dependency chains per thread (FMUL/FADD latency 4 cycles, 4 warps/scheduler)
constexpr int kIters = 8192;

__global__ void __launch_bounds__(1024, 2) unfused_kernel(float* out, float a, float b) {
  float x[kAcc];
#pragma unroll
  for (int k = 0; k < kAcc; k++) x[k] = (float)(threadIdx.x + k) * 1e-3f;
#pragma unroll 1
  for (int i = 0; i < kIters; i++) {
#pragma unroll
    for (int k = 0; k < kAcc; k++) x[k] = __fmul_rn(x[k], a);
#pragma unroll
    for (int k = 0; k < kAcc; k++) x[k] = __fadd_rn(x[k], b);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < kAcc; k++) s += x[k];
  if (s == 123.456f) out[0] = s;     // never true: keeps the chain alive
}

__global__ void __launch_bounds__(1024, 2) fma_kernel(float* out, float a, float b) {
  float x[kAcc];
#pragma unroll
  for (int k = 0; k < kAcc; k++) x[k] = (float)(threadIdx.x + k) * 1e-3f;
#pragma unroll 1
  for (int i = 0; i < kIters; i++) {
#pragma unroll
    for (int k = 0; k < kAcc; k++) x[k] = __fmaf_rn(x[k], a, b);
#pragma unroll
    for (int k = 0; k < kAcc; k++) x[k] = __fmaf_rn(x[k], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < kAcc; k++) s += x[k];
  if (s == 123.456f) out[0] = s;
}

template <class K>
static double time_ms(K kern, int grid, float* d) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int w = 0; w < 3; w++) kern<<<grid, 1024>>>(d, 1.0000001f, 1e-9f);
  cudaDeviceSynchronize();
  double best = 1e30;
  for (int r = 0; r < 10; r++) {
    cudaEventRecord(e0);
    kern<<<grid, 1024>>>(d, 1.0000001f, 1e-9f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) { fprintf(stderr, "no device\n"); return 1; }
  float* d;
  cudaMalloc(&d, 4);
  const int grid = p.multiProcessorCount * 2;
  const double ops = (double)grid * 1024 * kIters * kAcc * 2;       // instructions (lane-ops) per launch
  const double ms_u = time_ms(unfused_kernel, grid, d), ms_f = time_ms(fma_kernel, grid, d);
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"device\": \"%s\", \"sms\": %d, \"sm_clock_khz_max\": %d, "
         "\"unfused_fp32_lane_ops_per_s\": %.6e, \"unfused_ms\": %.4f, "
         "\"fma_lane_instr_per_s\": %.6e, \"fma_flops_per_s\": %.6e, \"fma_ms\": %.4f, "
         "\"nominal_lane_ops_per_s\": %.6e, \"how\": \"%d CTAs x 1024 threads, %d independent chains per thread, %d x (8 FMUL + 8 FADD) "
         "per thread, -fmad=false, best of 10, CUDA events\"}\n",
         p.name, p.multiProcessorCount, clk, ops / (ms_u * 1e-3), ms_u, ops / (ms_f * 1e-3), 2 * ops / (ms_f * 1e-3), ms_f,
         (double)p.multiProcessorCount * 128 * clk * 1e3, grid, kAcc, kIters);
  cudaFree(d);
  return 0;
}
