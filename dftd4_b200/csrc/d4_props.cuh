// Property epilogues (reference: src/dftd4/disp.f90:138-188 get_properties, model/d4.f90:381-473 get_atomic_c6 energy
// branch, :478-549 get_polarizabilities) and the FP64 peak probes used as roofline denominators.
#pragma once
#include "d4_common.cuh"
#include "d4_kernels.cuh"

namespace d4k {

// alpha_i = sum_a gw_i[a] * aiw(1, a, z_i)          (d4.f90:538-546)
__global__ void alpha_kernel(int nat, const int* __restrict__ species, Tables tb, const double* __restrict__ gw,
                             double* __restrict__ alpha) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nat) return;
  const int s = species[i];
  double a = 0.0;
  const size_t o = gw_row(tb, i, s);  // D4S: gwvec(iref, iat, iat), the set against the atom's own species (d4s.f90:540-546)
  for (int r = 0; r < tb.nref[s]; ++r) a += gw[o + r] * tb.refalpha0[s * D4_MREF + r];
  alpha[i] = a;
}

// c6[i][j] = sum_ab gw_i[a] gw_j[b] R[z_i][z_j][a][b]   (d4.f90:455-470); inherently N x N output, properties only.
__global__ void c6_matrix_kernel(int nat, const int* __restrict__ species, Tables tb, const double* __restrict__ gw,
                                 double* __restrict__ c6) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= nat || j >= nat) return;
  const int zi = species[i], zj = species[j];
  const double* R = tb.refc6 + ((size_t)zi * tb.nid + zj) * D4_MREF * D4_MREF;
  double acc = 0.0;
  const size_t oi = gw_row(tb, i, zj), oj = gw_row(tb, j, zi);  // D4S: gwvec(:, iat, jat) and gwvec(:, jat, iat)
  for (int a = 0; a < tb.nref[zi]; ++a) {
    double u = 0.0;
    for (int b = 0; b < tb.nref[zj]; ++b) u += R[a * D4_MREF + b] * gw[oj + b];
    acc += gw[oi + a] * u;
  }
  c6[(size_t)i * nat + j] = acc;
}

// ---- roofline denominators ---------------------------------------------------------------------------------------
// Register-resident DFMA chains: 8 independent accumulators per thread, ITER x 8 FMAs.
__global__ void __launch_bounds__(256) dfma_peak_kernel(double* __restrict__ out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int it = 0; it < iters; ++it) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Register-resident DMMA (mma.sync.m8n8k4.f64) chains: 8 independent accumulator tiles per warp.
__global__ void __launch_bounds__(256) dmma_peak_kernel(double* __restrict__ out, int iters, double a, double b) {
  double c[8][2];
#pragma unroll
  for (int k = 0; k < 8; ++k) { c[k][0] = threadIdx.x + k; c[k][1] = k; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                   : "+d"(c[k][0]), "+d"(c[k][1])
                   : "d"(a), "d"(b));
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += c[k][0] + c[k][1];
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace d4k
