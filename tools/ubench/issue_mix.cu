// Micro-benchmark: does a non-FP64 instruction issued next to every DFMA cost extra time on sm_100a?
// Kernel<MIX>: 8 independent DFMA chains per thread; MIX extra independent ops per DFMA on other pipes
// (0: none, 1: one IMAD (fma-lite/int pipe), 2: one LDS broadcast load, 3: IMAD + LDS, 4: two IMADs + LDS).
#include <cstdio>
#include <cuda_runtime.h>
template <int MIX>
__global__ void __launch_bounds__(256) k(double* out, int iters, double a, double b, int s) {
  __shared__ double sh[256];
  sh[threadIdx.x] = threadIdx.x;
  __syncthreads();
  double x[8];
  for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i;
  int n[8];
  for (int i = 0; i < 8; ++i) n[i] = threadIdx.x * (i + 1);
  double acc = 0.0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      x[i] = fma(x[i], a, b);
      if (MIX == 1 || MIX >= 3) n[i] = n[i] * s + it;
      if (MIX >= 4) n[(i + 3) & 7] ^= n[i] >> 3;
      if (MIX == 2 || MIX >= 3) acc += 0 * 1, x[(i + 1) & 7] = x[(i + 1) & 7];
    }
    if (MIX == 2 || MIX >= 3) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { double v = sh[(it + i * 7 + n[0] * 0) & 255]; asm volatile("" :: "d"(v)); }
    }
  }
  double sum = acc;
  int ns = 0;
  for (int i = 0; i < 8; ++i) { sum += x[i]; ns += n[i]; }
  if (sum == 12345.678 || ns == 123456789) out[blockIdx.x * blockDim.x + threadIdx.x] = sum + ns;
}
template <int MIX>
float run(double* d, int iters) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MIX><<<148 * 8, 256>>>(d, iters, 1.0000001, 1e-9, 3);
  cudaEventRecord(e0);
  k<MIX><<<148 * 8, 256>>>(d, iters, 1.0000001, 1e-9, 3);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  double* d; cudaMalloc(&d, 148 * 8 * 256 * 8);
  const int iters = 20000;
  const double fl = 2.0 * 8 * iters * 148.0 * 8 * 256;
  float t0 = run<0>(d, iters), t1 = run<1>(d, iters), t2 = run<2>(d, iters), t3 = run<3>(d, iters), t4 = run<4>(d, iters);
  printf("DFMA only        %.3f ms  %.2f TFLOP/s\n", t0, fl / t0 * 1e-9);
  printf("+1 IMAD / DFMA   %.3f ms  x%.3f\n", t1, t1 / t0);
  printf("+1 LDS / DFMA    %.3f ms  x%.3f\n", t2, t2 / t0);
  printf("+IMAD+LDS / DFMA %.3f ms  x%.3f\n", t3, t3 / t0);
  printf("+2IMAD+LDS       %.3f ms  x%.3f\n", t4, t4 / t0);
  return 0;
}
