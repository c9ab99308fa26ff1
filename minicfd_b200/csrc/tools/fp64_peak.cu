// tools/fp64_peak.cu -- FP64 pipe microbenchmarks on the B200 (measurement aid, not product code).
// Prints one JSON line: sustained DFMA TFLOP/s, and Gop/s of pow / exp / log / sqrt / div in FP64.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int kOp>
__global__ void bench(double *out, double seed, int iters) {
  double a0 = seed + threadIdx.x * 1e-9, a1 = a0 + 0.1, a2 = a0 + 0.2, a3 = a0 + 0.3;
  double a4 = a0 + 0.4, a5 = a0 + 0.5, a6 = a0 + 0.6, a7 = a0 + 0.7;
  const double m = 0.9999999, c = 1e-7;
  for (int i = 0; i < iters; ++i) {
    if (kOp == 0) {
      a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
      a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    } else if (kOp == 1) {
      a0 = pow(a0, 0.142857142857142857) + 0.5; a1 = pow(a1, 0.142857142857142857) + 0.5;
      a2 = pow(a2, 0.142857142857142857) + 0.5; a3 = pow(a3, 0.142857142857142857) + 0.5;
    } else if (kOp == 2) {
      a0 = exp(a0) * 0.3; a1 = exp(a1) * 0.3; a2 = exp(a2) * 0.3; a3 = exp(a3) * 0.3;
    } else if (kOp == 3) {
      a0 = log(a0) + 2.0; a1 = log(a1) + 2.0; a2 = log(a2) + 2.0; a3 = log(a3) + 2.0;
    } else if (kOp == 4) {
      a0 = sqrt(a0) + 0.5; a1 = sqrt(a1) + 0.5; a2 = sqrt(a2) + 0.5; a3 = sqrt(a3) + 0.5;
    } else {
      a0 = 1.0 / a0 + 0.5; a1 = 1.0 / a1 + 0.5; a2 = 1.0 / a2 + 0.5; a3 = 1.0 / a3 + 0.5;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// DMMA: mma.sync.aligned.m8n8k4 in FP64 (the only FP64 tensor-core shape): 8 x 8 x 4 = 256 multiply-adds per warp
// instruction.  Four independent accumulator tiles per warp, as the DFMA loop has eight independent chains.
__global__ void bench_dmma(double *out, double seed, int iters) {
  double a = seed + threadIdx.x * 1e-9, b = 1.0 - 1e-7 * threadIdx.x;
  double c0[2] = { 0, 0 }, c1[2] = { 0, 0 }, c2[2] = { 0, 0 }, c3[2] = { 0, 0 };
  for (int i = 0; i < iters; ++i) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c1[0]), "+d"(c1[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c2[0]), "+d"(c2[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c3[0]), "+d"(c3[1]) : "d"(a), "d"(b));
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = c0[0] + c0[1] + c1[0] + c1[1] + c2[0] + c2[1] + c3[0] + c3[1];
}

double run_dmma(double *d, int iters) {
  const int blocks = 148 * 8, threads = 256;
  bench_dmma<<<blocks, threads>>>(d, 1.0, iters);
  cudaDeviceSynchronize();
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  bench_dmma<<<blocks, threads>>>(d, 1.0, iters);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  // per warp and iteration: 4 instructions x 256 multiply-adds x 2 flop
  return double(blocks) * (threads / 32) * iters * 4.0 * 256.0 * 2.0 / (ms * 1e-3);
}

template <int kOp>
double run(double *d, int iters, int per_iter) {
  const int blocks = 148 * 8, threads = 256;
  bench<kOp><<<blocks, threads>>>(d, 1.0, iters);
  cudaDeviceSynchronize();
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  bench<kOp><<<blocks, threads>>>(d, 1.0, iters);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  return double(blocks) * threads * iters * per_iter / (ms * 1e-3);
}

int main() {
  double *d; cudaMalloc(&d, 148 * 8 * 256 * sizeof(double));
  double fma_s = run<0>(d, 20000, 8);
  double pow_s = run<1>(d, 500, 4), exp_s = run<2>(d, 2000, 4), log_s = run<3>(d, 2000, 4);
  double sqrt_s = run<4>(d, 2000, 4), div_s = run<5>(d, 2000, 4);
  double dmma = run_dmma(d, 20000);
  printf("{\"fp64_dmma_m8n8k4_tflops\": %.2f, ", dmma / 1e12);
  printf("\"fp64_fma_tflops\": %.2f, \"pow_gops\": %.1f, \"exp_gops\": %.1f, \"log_gops\": %.1f, \"sqrt_gops\": %.1f, \"div_gops\": %.1f}\n",
         2 * fma_s / 1e12, pow_s / 1e9, exp_s / 1e9, log_s / 1e9, sqrt_s / 1e9, div_s / 1e9);
  return 0;
}
