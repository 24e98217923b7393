// Micro-benchmark: FP64 pipe of one SM sub-partition on sm_100a -- dependent-issue latency of DFMA / DMUL / DADD and the
// number of independent chains per scheduler needed to saturate the pipe.
// Kernel<ILP>: ILP independent DFMA chains per thread; launched with W warps per SM (1 CTA per SM, W*32 threads).
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP, int OP>
__global__ void k(double* out, int iters, double a, double b, long long* cyc) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int i = 0; i < ILP; ++i) {
        if (OP == 0) x[i] = fma(x[i], a, b);
        if (OP == 1) x[i] = x[i] * a;
        if (OP == 2) x[i] = x[i] + b;
      }
  }
  long long t1 = clock64();
  double sum = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) sum += x[i];
  if (sum == 12345.678) out[threadIdx.x] = sum;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int ILP, int OP>
void run(double* d, long long* c, int warps) {
  const int iters = 2000;
  k<ILP, OP><<<148, warps * 32>>>(d, iters, 1.0000001, 1e-9, c);
  k<ILP, OP><<<148, warps * 32>>>(d, iters, 1.0000001, 1e-9, c);
  cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
  const double inst_per_warp = 8.0 * ILP * iters;
  // per scheduler: warps/4 warps each issuing inst_per_warp instructions in h cycles
  printf("op %d ILP %d warps/SM %2d: %.2f cycles per instruction per warp, %.3f FP64 warp-instr/cycle/scheduler\n", OP, ILP, warps,
         h / inst_per_warp, inst_per_warp * (warps / 4.0 > 1 ? warps / 4.0 : 1.0 * (warps >= 4 ? warps / 4.0 : warps / 4.0)) / h);
}
int main() {
  double* d; long long* c;
  cudaMalloc(&d, 1024 * 8); cudaMalloc(&c, 8);
  for (int op = 0; op < 3; ++op) { if (op == 0) run<1, 0>(d, c, 4); if (op == 1) run<1, 1>(d, c, 4); if (op == 2) run<1, 2>(d, c, 4); }
  int ws[] = {4, 8, 12, 16, 24, 32};
  for (int w : ws) run<1, 0>(d, c, w);
  for (int w : ws) run<2, 0>(d, c, w);
  for (int w : ws) run<4, 0>(d, c, w);
  run<8, 0>(d, c, 4); run<8, 0>(d, c, 8);
  return 0;
}
