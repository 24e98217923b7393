// Blocked FP64 Cholesky + bordered (charge-constraint) solve for the EEQ matrix, sm_100a.
//
// Replaces the LAPACK dsytrf/dsytrs/dsytri + GEMM inside multicharge's `solve` (call site: reference
// src/dftd4/disp.f90:88).  The EEQ Coulomb matrix A (N x N) is symmetric positive definite; the constraint row/column
// of the reference's bordered (N+1) system is eliminated by a Schur complement on the host side of this file
// (two right-hand sides: X and 1).  Layout: row-major, leading dimension ld = N padded to a multiple of NB, only the
// lower triangle (j <= i) is referenced; the padding carries a unit diagonal.
//
// Right-looking, panel width NB = 128:
//   chol_diag_kernel : one CTA factors the 128x128 diagonal block in shared memory and inverts the factor in place
//   gemm_nt_kernel   : C = A*B^T (panel solve with the inverted factor) or C -= A*B^T (trailing update);
//                      128x128 output tiles, K = 128, FP64 tensor-core tiles (mma.sync.m8n8k4.f64 = DMMA),
//                      operands staged through shared memory with cp.async double buffering
//   solve_*_kernel   : block forward / backward substitution with the inverted diagonal blocks
#pragma once
#include "d4_common.cuh"

namespace d4k {

constexpr int NB = 128;          // panel width
constexpr int GK = 16;           // K-chunk of the GEMM pipeline
constexpr int GS = GK + 4;       // shared-memory row stride (doubles): conflict-free DMMA fragment loads
constexpr int DS = NB + 1;       // shared-memory stride of the diagonal-block kernel

__device__ __forceinline__ void dmma8x8x4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// ---------------------------------------------------------------------------------------------------------------
// Diagonal block: L = chol(A[kb,kb]) written back to A (lower), inv(L) written to invL (full 128x128, upper = 0).
// info[0] is set to (global row + 1) of the first non-positive pivot.
// Blocked inside the CTA with 32-wide sub-panels so that the 128 sequential pivots cost 4 x (one-warp 32x32
// factorisation + inversion) instead of 128 CTA-wide barriers:
//   for each sub-panel p: warp 0 factors D_p and inverts it (DI_p); all threads form X = A_panel DI_p^T and apply the
//   rank-32 update to the trailing lower triangle.
//   inv(L): M_pp = DI_p, M_pq = -M_pp sum_{t=q}^{p-1} L_pt M_tq for block distances 1, 2, 3 (kept transposed in the unused
//   upper blocks of the shared-memory tile).
constexpr int SB = 32;               // sub-panel width
constexpr int NSB = NB / SB;         // 4
constexpr int TS = SB + 1;           // stride of the 32-wide scratch tiles
__host__ __device__ inline size_t chol_diag_smem_bytes() {
  return sizeof(double) * ((size_t)NB * DS + (size_t)NSB * SB * TS + (size_t)(NB - SB) * TS) + 16;
}

__global__ void __launch_bounds__(512) chol_diag_kernel(double* __restrict__ A, int ld, int kb, double* __restrict__ invL,
                                                        int* __restrict__ info) {
  extern __shared__ double s[];          // [NB][DS] working tile
  double* DI = s + NB * DS;              // [NSB][SB][TS] inverses of the diagonal sub-blocks
  double* TMP = DI + NSB * SB * TS;      // [NB - SB][TS] scratch
  int* fail = reinterpret_cast<int*>(TMP + (NB - SB) * TS);
  __shared__ double sdinv[SB];           // reciprocal pivots of the current 32x32 sub-block
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* Ab = A + (size_t)kb * NB * ld + (size_t)kb * NB;
  for (int e = tid; e < NB * NB; e += blockDim.x) {
    const int i = e >> 7, j = e & (NB - 1);
    s[i * DS + j] = (j <= i) ? Ab[(size_t)i * ld + j] : 0.0;
  }
  if (tid == 0) *fail = 0;
  __syncthreads();

  for (int p = 0; p < NSB; ++p) {
    const int c0 = p * SB;
    double* Dp = s + c0 * DS + c0;       // D[r][c] = Dp[r * DS + c]
    double* Ip = DI + p * SB * TS;
    if (warp == 0) {
      // --- 32x32 Cholesky, lane = row, the row lives in registers; column k is broadcast with shuffles ---
      double row[SB];
#pragma unroll
      for (int j = 0; j < SB; ++j) row[j] = Dp[lane * DS + j];
      // branch-free: a single warp runs this dependent chain, so every divergent branch or long-latency operation is
      // on the critical path of the whole factorisation.  1/sqrt(piv) by hardware seed + two Newton steps (no DSQRT /
      // DDIV sequence); the update of columns j > k is predicated by a select instead of a branch.
      int bad_k = -1;
#pragma unroll
      for (int k = 0; k < SB; ++k) {
        const double piv = __shfl_sync(D4_FULL, row[k], k);
        if (bad_k < 0 && !(piv > 0.0)) bad_k = k;
        double dinv;
        asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(dinv) : "d"(piv));
        {
          double e = fma(-piv * dinv, dinv, 1.0);
          dinv = fma(dinv * e, fma(0.375, e, 0.5), dinv);
          e = fma(-piv * dinv, dinv, 1.0);
          dinv = fma(dinv * e, 0.5, dinv);
        }
        const double d = piv * dinv;
        if (lane == 0) sdinv[k] = dinv;                     // reused by the inversion below (no division there)
        const double lrk = lane == k ? d : row[k] * dinv;   // column k of L (rows >= k)
        row[k] = lrk;
#pragma unroll
        for (int j = k + 1; j < SB; ++j) {
          const double ljk = __shfl_sync(D4_FULL, lrk, j);
          row[j] = fma(-lrk, lane >= j ? ljk : 0.0, row[j]);
        }
      }
      if (bad_k >= 0) { if (lane == 0) { atomicCAS(info, 0, kb * NB + c0 + bad_k + 1); *fail = 1; } }
#pragma unroll
      for (int j = 0; j < SB; ++j) Dp[lane * DS + j] = (j <= lane) ? row[j] : 0.0;
      __syncwarp();
      // --- inverse of the factor, lane = column c: forward substitution D x = e_c, x in registers ---
      {
        double x[SB];
#pragma unroll
        for (int r = 0; r < SB; ++r) {
          double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;   // four partial sums: shorter dependent FMA chains
#pragma unroll
          for (int j = 0; j < r; ++j) {                     // x[j] = 0 for j < c
            const double dv = Dp[r * DS + j];
            if ((j & 3) == 0) a0 = fma(dv, x[j], a0);
            else if ((j & 3) == 1) a1 = fma(dv, x[j], a1);
            else if ((j & 3) == 2) a2 = fma(dv, x[j], a2);
            else a3 = fma(dv, x[j], a3);
          }
          const double acc = (a0 + a1) + (a2 + a3);
          const double dinv = sdinv[r];
          x[r] = (r == lane) ? dinv : (r > lane ? -acc * dinv : 0.0);
        }
#pragma unroll
        for (int r = 0; r < SB; ++r) Ip[r * TS + lane] = x[r];
      }
    }
    __syncthreads();
    if (*fail) return;
    const int r0 = c0 + SB, nr = NB - r0;   // rows below the sub-panel
    if (nr > 0) {
      // X = A_panel * DI^T : X[r][k] = sum_{j<=k} A[r][c0+j] * DI[k][j]
      for (int e = tid; e < nr * SB; e += blockDim.x) {
        const int r = e >> 5, k = e & 31;
        const double* ar = s + (r0 + r) * DS + c0;
        const double* ik = Ip + k * TS;
        // four partial sums: these dot products are dependent-FMA chains, the kernel is latency-bound
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int j = 0;
        for (; j + 3 <= k; j += 4) {
          a0 = fma(ar[j], ik[j], a0); a1 = fma(ar[j + 1], ik[j + 1], a1);
          a2 = fma(ar[j + 2], ik[j + 2], a2); a3 = fma(ar[j + 3], ik[j + 3], a3);
        }
        for (; j <= k; ++j) a0 = fma(ar[j], ik[j], a0);
        TMP[r * TS + k] = (a0 + a1) + (a2 + a3);
      }
      __syncthreads();
      for (int e = tid; e < nr * SB; e += blockDim.x) {
        const int r = e >> 5, k = e & 31;
        s[(r0 + r) * DS + c0 + k] = TMP[r * TS + k];
      }
      // trailing lower triangle: s[i][j] -= sum_k X[i][k] X[j][k]; 32 x 16 thread grid
      for (int i = (tid >> 4); i < nr; i += 32) {
        for (int j = (tid & 15); j <= i; j += 16) {
          const double* xi = TMP + i * TS;
          const double* xj = TMP + j * TS;
          double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
          for (int k = 0; k < SB; k += 4) {
            a0 = fma(xi[k], xj[k], a0); a1 = fma(xi[k + 1], xj[k + 1], a1);
            a2 = fma(xi[k + 2], xj[k + 2], a2); a3 = fma(xi[k + 3], xj[k + 3], a3);
          }
          s[(r0 + i) * DS + r0 + j] -= (a0 + a1) + (a2 + a3);
        }
      }
      __syncthreads();
    }
  }
  // ---- L back to the matrix ----
  for (int e = tid; e < NB * NB; e += blockDim.x) {
    const int i = e >> 7, j = e & (NB - 1);
    if (j <= i) Ab[(size_t)i * ld + j] = s[i * DS + j];
  }
  // ---- inv(L): off-diagonal blocks by block distance; M_pq is stored TRANSPOSED in the free upper block (q, p) ----
  for (int dist = 1; dist < NSB; ++dist) {
    const int npair = NSB - dist;            // pairs (p, q) = (q + dist, q)
    // S_pq = sum_{t=q}^{p-1} L_pt M_tq  -> TMP[pair]   (M_qq = DI_q; M_tq, t > q, transposed in block (q, t))
    for (int e = tid; e < npair * SB * SB; e += blockDim.x) {
      const int pr = e >> 10, i = (e >> 5) & 31, j = e & 31;
      const int q = pr, p = q + dist;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
      const double* Lrow = s + (p * SB + i) * DS;
      {  // t = q: M_qq lower triangular -> k >= j
        const double* Mq = DI + q * SB * TS + j;
        const double* Lq = Lrow + q * SB;
        int k = j;
        for (; k + 3 < SB; k += 4) {
          a0 = fma(Lq[k], Mq[k * TS], a0); a1 = fma(Lq[k + 1], Mq[(k + 1) * TS], a1);
          a2 = fma(Lq[k + 2], Mq[(k + 2) * TS], a2); a3 = fma(Lq[k + 3], Mq[(k + 3) * TS], a3);
        }
        for (; k < SB; ++k) a0 = fma(Lq[k], Mq[k * TS], a0);
      }
      for (int t = q + 1; t < p; ++t) {   // M_tq[k][j] = s[(q*SB + j)][t*SB + k]
        const double* Lt = Lrow + t * SB;
        const double* Mt = s + (q * SB + j) * DS + t * SB;
#pragma unroll
        for (int k = 0; k < SB; k += 4) {
          a0 = fma(Lt[k], Mt[k], a0); a1 = fma(Lt[k + 1], Mt[k + 1], a1);
          a2 = fma(Lt[k + 2], Mt[k + 2], a2); a3 = fma(Lt[k + 3], Mt[k + 3], a3);
        }
      }
      TMP[pr * SB * TS + i * TS + j] = (a0 + a1) + (a2 + a3);
    }
    __syncthreads();
    // M_pq = -M_pp S_pq, stored transposed: s[(q*SB + j)][p*SB + i] = M_pq[i][j]
    for (int e = tid; e < npair * SB * SB; e += blockDim.x) {
      const int pr = e >> 10, j = (e >> 5) & 31, i = e & 31;   // i fastest: consecutive threads write consecutive columns
      const int q = pr, p = q + dist;
      const double* Mp = DI + p * SB * TS;
      const double* S = TMP + pr * SB * TS;
      const double* mi = Mp + i * TS;
      const double* sj = S + j;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
      int k = 0;
      for (; k + 3 <= i; k += 4) {
        a0 = fma(mi[k], sj[k * TS], a0); a1 = fma(mi[k + 1], sj[(k + 1) * TS], a1);
        a2 = fma(mi[k + 2], sj[(k + 2) * TS], a2); a3 = fma(mi[k + 3], sj[(k + 3) * TS], a3);
      }
      for (; k <= i; ++k) a0 = fma(mi[k], sj[k * TS], a0);
      s[(q * SB + j) * DS + p * SB + i] = -((a0 + a1) + (a2 + a3));
    }
    __syncthreads();
  }
  double* Ib = invL + (size_t)kb * NB * NB;
  for (int e = tid; e < NB * NB; e += blockDim.x) {
    const int i = e >> 7, j = e & (NB - 1);
    const int p = i >> 5, q = j >> 5;
    double v = 0.0;
    if (p == q) v = (j <= i) ? DI[p * SB * TS + (i & 31) * TS + (j & 31)] : 0.0;
    else if (p > q) v = s[j * DS + i];
    Ib[e] = v;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// C(128x128 tile) = A*B^T (MODE 0) or C -= A*B^T (MODE 1); A, B, C row-major, K = 128.
//   MODE 0 (panel solve):      tile t: rows row0 + 128 t of the panel; A = C = panel tile, B = inverted diagonal factor
//   MODE 1 (trailing update):  blockIdx.x enumerates lower-triangle tile pairs (ti >= tj) of the trailing matrix
//   MODE 2 (rank-128 update):  C += P*P^T over all lower-triangle tile pairs of the whole matrix; P = invL argument,
//                              row-major [ld][128] (the Ewald phase matrix of the periodic EEQ build)
//   part (MODE 1 only): 0 = every tile pair, 1 = first tile column only (ti, 0), 2 = everything but the first column;
//                       the split lets the next diagonal block + panel start while the rest of the update runs;
//                       3 = the tile columns t0, t0 + W, ... only (multi-GPU factorisation: a rank updates the block
//                       columns it owns; grid = (ntile - t0, number of owned columns), CTAs past the diagonal exit)
template <int MODE>
__global__ void __launch_bounds__(256) gemm_nt_kernel(double* __restrict__ Amat, int ld, int kb, int ntile,
                                                      const double* __restrict__ invL, int part = 0, int t0 = 0, int W = 1) {
  extern __shared__ double sm[];  // 2 stages x (A tile [128][GS] + B tile [128][GS])
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int ti, tj;
  if (MODE == 0 || (MODE == 1 && part == 1)) {
    ti = blockIdx.x; tj = 0;
  } else if (MODE == 1 && part == 3) {
    tj = t0 + (int)blockIdx.y * W;
    ti = tj + (int)blockIdx.x;
    if (ti >= ntile) return;
  } else {
    // unrank the lower triangle: p = ti*(ti+1)/2 + tj
    const int p = blockIdx.x;
    ti = (int)floor((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
    while ((long long)(ti + 1) * (ti + 2) / 2 <= p) ++ti;
    while ((long long)ti * (ti + 1) / 2 > p) --ti;
    tj = p - ti * (ti + 1) / 2;
    if (MODE == 1 && part == 2) { ++ti; ++tj; }  // lower triangle of the tiles 1 .. rest-1
  }
  if (MODE == 2 && W > 1 && (tj % W) != t0) return;           // multi-GPU build: only the block columns of rank t0
  const size_t row0 = MODE == 2 ? 0 : (size_t)(kb + 1) * NB;  // first row below the diagonal block
  const double* Ag = Amat + (row0 + (size_t)ti * NB) * ld + (size_t)kb * NB;             // [128][K] panel rows of tile i
  const double* Bg;
  size_t ldb, lda = ld;
  double* Cg;
  if (MODE == 2) {
    Ag = invL + (size_t)ti * NB * NB; lda = NB;
    Bg = invL + (size_t)tj * NB * NB; ldb = NB;
    Cg = Amat + ((size_t)ti * NB) * ld + (size_t)tj * NB;
  } else if (MODE == 0) {
    Bg = invL + (size_t)kb * NB * NB; ldb = NB;
    Cg = Amat + (row0 + (size_t)ti * NB) * ld + (size_t)kb * NB;
  } else {
    Bg = Amat + (row0 + (size_t)tj * NB) * ld + (size_t)kb * NB; ldb = ld;
    Cg = Amat + (row0 + (size_t)ti * NB) * ld + row0 + (size_t)tj * NB;
  }
  (void)ntile;
  double* As[2] = {sm, sm + 2 * NB * GS};
  double* Bs[2] = {sm + NB * GS, sm + 3 * NB * GS};

  auto stage = [&](int st, int kc) {
    // 128 rows x 16 doubles = 128 x 8 chunks of 16 B per operand; 256 threads -> 4 chunks each per operand
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int c = tid + r * 256;
      const int row = c >> 3, part = c & 7;
      cp_async16(As[st] + row * GS + part * 2, Ag + (size_t)row * lda + kc * GK + part * 2);
      cp_async16(Bs[st] + row * GS + part * 2, Bg + (size_t)row * ldb + kc * GK + part * 2);
    }
    cp_async_commit();
  };

  // warp grid 2 (M) x 4 (N): warp tile 64 x 32 = 8 x 4 DMMA tiles
  const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;
  const int g = lane >> 2, q = lane & 3;
  double acc[8][4][2];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

  constexpr int NCH = NB / GK;
  stage(0, 0);
  for (int kc = 0; kc < NCH; ++kc) {
    if (kc + 1 < NCH) { stage((kc + 1) & 1, kc + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const double* a_s = As[kc & 1];
    const double* b_s = Bs[kc & 1];
#pragma unroll
    for (int k4 = 0; k4 < GK; k4 += 4) {
      double af[8], bf[4];
#pragma unroll
      for (int a = 0; a < 8; ++a) af[a] = a_s[(wm + a * 8 + g) * GS + k4 + q];
#pragma unroll
      for (int b = 0; b < 4; ++b) bf[b] = b_s[(wn + b * 8 + g) * GS + k4 + q];
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) dmma8x8x4(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
    __syncthreads();
  }
  // epilogue: lane holds C[g][2q], C[g][2q+1] of every 8x8 tile
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      double2* dst = reinterpret_cast<double2*>(Cg + (size_t)(wm + a * 8 + g) * ld + wn + b * 8 + 2 * q);
      if (MODE == 0) {
        *dst = make_double2(acc[a][b][0], acc[a][b][1]);
      } else if (MODE == 1) {
        double2 c = *dst;
        c.x -= acc[a][b][0];
        c.y -= acc[a][b][1];
        *dst = c;
      } else {
        double2 c = *dst;
        c.x += acc[a][b][0];
        c.y += acc[a][b][1];
        *dst = c;
      }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Trailing update C -= A B^T of the factorisation with 128 x 64 output tiles (the bulk of the N^3/3 flop).
// gemm_nt_kernel<1> holds a 128 x 128 tile per CTA: 64 accumulator doubles per thread -> 164 registers -> ONE CTA per
// SM, so the latency of its prologue (first operand chunk) and epilogue (C read-modify-write) is exposed.  Here a warp
// owns 32 x 32 (16 DMMA tiles, 32 accumulator doubles): <= 128 registers, two CTAs per SM that overlap each other's
// prologue / epilogue.  Tile (ti, tc): rows of block ti (128), columns of half-block tc (64) of the trailing matrix,
// lower triangle = tc <= 2 ti + 1.
//   part 0: every tile, blockIdx.x = ti (ti + 1) + tc          part 1: first block column only (2 x ntile CTAs)
//   part 2: everything but the first block column              part 3: block columns t0, t0 + W, ... (multi-GPU)
constexpr int UN = 64;
constexpr size_t UPD_SMEM = (size_t)2 * (NB + UN) * GS * sizeof(double);
__global__ void __launch_bounds__(256, 2) chol_update_kernel(double* __restrict__ Amat, int ld, int kb, int ntile, int part,
                                                             int t0, int W) {
  extern __shared__ double sm[];  // 2 stages x (A tile [128][GS] + B tile [64][GS])
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int ti, tc;
  if (part == 1) {
    ti = blockIdx.x >> 1; tc = blockIdx.x & 1;
  } else if (part == 3) {
    const int tj = t0 + (int)(blockIdx.y >> 1) * W;
    ti = tj + (int)blockIdx.x;
    if (ti >= ntile) return;
    tc = 2 * tj + (int)(blockIdx.y & 1);
  } else {
    const long long p = blockIdx.x;
    ti = (int)floor((sqrt(4.0 * (double)p + 1.0) - 1.0) * 0.5);
    while ((long long)(ti + 1) * (ti + 2) <= p) ++ti;
    while ((long long)ti * (ti + 1) > p) --ti;
    tc = (int)(p - (long long)ti * (ti + 1));
    if (part == 2) { ti += 1; tc += 2; }
  }
  const size_t row0 = (size_t)(kb + 1) * NB;
  const double* Ag = Amat + (row0 + (size_t)ti * NB) * ld + (size_t)kb * NB;   // [128][K]
  const double* Bg = Amat + (row0 + (size_t)tc * UN) * ld + (size_t)kb * NB;   // [64][K]
  double* Cg = Amat + (row0 + (size_t)ti * NB) * ld + row0 + (size_t)tc * UN;
  double* As[2] = {sm, sm + (NB + UN) * GS};
  double* Bs[2] = {sm + NB * GS, sm + (NB + UN) * GS + NB * GS};

  auto stage = [&](int st, int kc) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int c = tid + r * 256;
      const int row = c >> 3, part8 = c & 7;
      cp_async16(As[st] + row * GS + part8 * 2, Ag + (size_t)row * ld + kc * GK + part8 * 2);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int c = tid + r * 256;
      const int row = c >> 3, part8 = c & 7;
      cp_async16(Bs[st] + row * GS + part8 * 2, Bg + (size_t)row * ld + kc * GK + part8 * 2);
    }
    cp_async_commit();
  };

  // warp grid 4 (M) x 2 (N): warp tile 32 x 32 = 4 x 4 DMMA tiles
  const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;
  const int g = lane >> 2, q = lane & 3;
  double acc[4][4][2];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

  constexpr int NCH = NB / GK;
  stage(0, 0);
  for (int kc = 0; kc < NCH; ++kc) {
    if (kc + 1 < NCH) { stage((kc + 1) & 1, kc + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const double* a_s = As[kc & 1];
    const double* b_s = Bs[kc & 1];
#pragma unroll
    for (int k4 = 0; k4 < GK; k4 += 4) {
      double af[4], bf[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) af[a] = a_s[(wm + a * 8 + g) * GS + k4 + q];
#pragma unroll
      for (int b = 0; b < 4; ++b) bf[b] = b_s[(wn + b * 8 + g) * GS + k4 + q];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) dmma8x8x4(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
    __syncthreads();
  }
  // epilogue: all loads first (independent), then the stores
  double2 cv[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b)
      cv[a][b] = *reinterpret_cast<const double2*>(Cg + (size_t)(wm + a * 8 + g) * ld + wn + b * 8 + 2 * q);
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      cv[a][b].x -= acc[a][b][0];
      cv[a][b].y -= acc[a][b][1];
      *reinterpret_cast<double2*>(Cg + (size_t)(wm + a * 8 + g) * ld + wn + b * 8 + 2 * q) = cv[a][b];
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Substitution: forward AND backward sweep of all block rows in ONE launch (wavefront over the block rows; one launch
// per block row and sweep would be ~8 ms of launch latency at 10k atoms).  A CTA claims the next block row of the sweep
// from a ticket counter (ascending in the forward sweep, descending in the backward one), consumes y_k / x_k of the rows
// it depends on as soon as their owner publishes them (release / acquire flags in global memory) and publishes its own
// block.  Every row a CTA waits for was claimed before its own, i.e. by a CTA that is already running, so the waits
// cannot form a cycle whatever else shares the device.  A bounded spin turns a protocol error into an error code
// instead of a hang.
//   flags[0 .. nblk)        forward:  y_m published        flags[nblk .. 2 nblk)  backward: x_m published
//   b is read only; x receives y (forward) and is overwritten by the solution (backward).
__device__ __forceinline__ int ld_acquire_s32(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_s32(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool subst_wait(const int* flag, int* err) {
  __shared__ int s_bad;
  if (threadIdx.x == 0) {
    int bad = 0;
    long long spins = 0;
    while (ld_acquire_s32(flag) == 0) {
      if ((++spins & 0x3ff) == 0 && (spins > (1ll << 22) || ld_acquire_s32(err) != 0)) { atomicExch(err, 1); bad = 1; break; }
    }
    s_bad = bad;
  }
  __syncthreads();
  const bool ok = s_bad == 0;
  __syncthreads();  // s_bad is rewritten by the next wait
  return ok;
}

constexpr int SUBST_THREADS = 256;

// L2 prefetch of a 128 x 128 block (row stride ldm): issued before a wait so that the loads after it hit L2.
__device__ __forceinline__ void subst_prefetch_block(const double* __restrict__ M, size_t ldm) {
#pragma unroll
  for (int r = 0; r < (NB * NB * 8 / 128) / SUBST_THREADS; ++r) {
    const int line = threadIdx.x + r * SUBST_THREADS;          // 1024 lines of 128 B: 8 per row
    const double* p = M + (size_t)(line >> 3) * ldm + (line & 7) * 16;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
  }
}

// acc[r][row] -= sum_c M[row][c] * v[r][c] for the 128 rows of a row-major block M (row stride ldm); warp w owns rows
// 16 w .. 16 w + 15, lanes read 4 consecutive doubles of a row (coalesced), warp-shuffle reduction per row.
// STORE = true: out[r][row] = sum (used for y_m = invL_m * acc).  Eight rows (16 vector loads) are in flight per lane.
template <int NR, bool STORE>
__device__ __forceinline__ void subst_matvec_rows(const double* __restrict__ M, size_t ldm, const double (*v)[NB],
                                                  double (*acc)[NB], double (*out)[NB]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int RPW = NB / (SUBST_THREADS / 32);   // 16 rows per warp
  double vv[NR][4];
#pragma unroll
  for (int r = 0; r < NR; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q) vv[r][q] = v[r][lane * 4 + q];
#pragma unroll
  for (int r0 = 0; r0 < RPW; r0 += 8) {
    double2 a0[8], a1[8];
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) {
      const double2* src = reinterpret_cast<const double2*>(M + (size_t)(warp * RPW + r0 + rr) * ldm + lane * 4);
      a0[rr] = __ldcg(src); a1[rr] = __ldcg(src + 1);
    }
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) {
      const int row = warp * RPW + r0 + rr;
      double sum[NR];
#pragma unroll
      for (int r = 0; r < NR; ++r)
        sum[r] = warp_sum(a0[rr].x * vv[r][0] + a0[rr].y * vv[r][1] + a1[rr].x * vv[r][2] + a1[rr].y * vv[r][3]);
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < NR; ++r) {
          if (STORE) out[r][row] = sum[r]; else acc[r][row] -= sum[r];
        }
      }
    }
  }
}

// part[w][r][col] = sum over the 16 rows i of warp w of M[i][col] * v[r][i]  (transposed product).  Lanes own 4 columns,
// a warp reads whole rows (coalesced), 8 rows in flight; the caller adds the 8 partial sums.
template <int NR>
__device__ __forceinline__ void subst_matvec_cols(const double* __restrict__ M, size_t ldm, const double (*v)[NB],
                                                  double (*part)[NR][NB]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int RPW = NB / (SUBST_THREADS / 32);
  double sum[NR][4];
#pragma unroll
  for (int r = 0; r < NR; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q) sum[r][q] = 0.0;
#pragma unroll
  for (int r0 = 0; r0 < RPW; r0 += 8) {
    double2 a0[8], a1[8];
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) {
      const double2* src = reinterpret_cast<const double2*>(M + (size_t)(warp * RPW + r0 + rr) * ldm + lane * 4);
      a0[rr] = __ldcg(src); a1[rr] = __ldcg(src + 1);
    }
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) {
      const int i = warp * RPW + r0 + rr;
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        const double x = v[r][i];
        sum[r][0] = fma(a0[rr].x, x, sum[r][0]); sum[r][1] = fma(a0[rr].y, x, sum[r][1]);
        sum[r][2] = fma(a1[rr].x, x, sum[r][2]); sum[r][3] = fma(a1[rr].y, x, sum[r][3]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < NR; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q) part[warp][r][lane * 4 + q] = sum[r][q];
}

template <int NR>
__global__ void __launch_bounds__(SUBST_THREADS) chol_subst_kernel(const double* __restrict__ A, int ld, int nblk,
                                                                  const double* __restrict__ invL,
                                                                  const double* __restrict__ b, double* x, int ldb,
                                                                  int* flags, int* err) {
  __shared__ double acc[NR][NB];        // running right-hand side of the current block row
  __shared__ double vec[NR][NB];        // y_k / x_k being applied, then the row's result
  __shared__ double part[SUBST_THREADS / 32][NR][NB];    // per-warp partial sums of the transposed products
  __shared__ int s_row;
  const int tid = threadIdx.x;
  int* fwd = flags;
  int* bwd = flags + nblk;
  int* tickets = flags + 2 * nblk + 1;   // [0] forward, [1] backward sweep

  // Rows are claimed from a ticket counter, in sweep order: whoever waits for block row k waits for a CTA that claimed
  // its ticket earlier and is therefore already running -- no assumption about how many CTAs are co-resident.
  // ---------------- forward: y_m = invL_m (b_m - sum_{k<m} A[m,k] y_k) ----------------
  while (true) {
    if (tid == 0) s_row = atomicAdd(&tickets[0], 1);
    __syncthreads();
    const int m = s_row;
    __syncthreads();
    if (m >= nblk) break;
    if (tid < NB)
      for (int r = 0; r < NR; ++r) acc[r][tid] = b[(size_t)r * ldb + (size_t)m * NB + tid];
    __syncthreads();
    for (int k = 0; k < m; ++k) {
      subst_prefetch_block(A + (size_t)m * NB * ld + (size_t)k * NB, ld);
      if (k == m - 1) subst_prefetch_block(invL + (size_t)m * NB * NB, NB);
      if (!subst_wait(fwd + k, err)) return;
      if (tid < NB)
        for (int r = 0; r < NR; ++r) vec[r][tid] = __ldcg(x + (size_t)r * ldb + (size_t)k * NB + tid);
      __syncthreads();
      subst_matvec_rows<NR, false>(A + (size_t)m * NB * ld + (size_t)k * NB, ld, vec, acc, nullptr);
      __syncthreads();
    }
    subst_matvec_rows<NR, true>(invL + (size_t)m * NB * NB, NB, acc, nullptr, vec);
    __syncthreads();
    if (tid < NB)
      for (int r = 0; r < NR; ++r) x[(size_t)r * ldb + (size_t)m * NB + tid] = vec[r][tid];
    __threadfence();
    __syncthreads();
    if (tid == 0) st_release_s32(fwd + m, 1);
  }

  // ---------------- backward: x_m = invL_m^T (y_m - sum_{k>m} A[k,m]^T x_k) ----------------
  // (a CTA gets here only after every forward row has been claimed, so the owner of y_m is running)
  while (true) {
    if (tid == 0) s_row = atomicAdd(&tickets[1], 1);
    __syncthreads();
    const int m = nblk - 1 - s_row;
    __syncthreads();
    if (m < 0) break;
    if (!subst_wait(fwd + m, err)) return;   // y_m may come from another CTA
    if (tid < NB)
      for (int r = 0; r < NR; ++r) acc[r][tid] = __ldcg(x + (size_t)r * ldb + (size_t)m * NB + tid);  // y_m (own row)
    __syncthreads();
    for (int k = nblk - 1; k > m; --k) {
      subst_prefetch_block(A + (size_t)k * NB * ld + (size_t)m * NB, ld);
      if (k == m + 1) subst_prefetch_block(invL + (size_t)m * NB * NB, NB);
      if (!subst_wait(bwd + k, err)) return;
      if (tid < NB)
        for (int r = 0; r < NR; ++r) vec[r][tid] = __ldcg(x + (size_t)r * ldb + (size_t)k * NB + tid);
      __syncthreads();
      subst_matvec_cols<NR>(A + (size_t)k * NB * ld + (size_t)m * NB, ld, vec, part);
      __syncthreads();
      if (tid < NB)
        for (int r = 0; r < NR; ++r) {
          double t = 0.0;
#pragma unroll
          for (int w = 0; w < SUBST_THREADS / 32; ++w) t += part[w][r][tid];
          acc[r][tid] -= t;
        }
      __syncthreads();
    }
    subst_matvec_cols<NR>(invL + (size_t)m * NB * NB, NB, acc, part);   // (invL^T acc)[j] = sum_i invL[i][j] acc[i]
    __syncthreads();
    if (tid < NB)
      for (int r = 0; r < NR; ++r) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < SUBST_THREADS / 32; ++w) t += part[w][r][tid];
        x[(size_t)r * ldb + (size_t)m * NB + tid] = t;
      }
    __threadfence();
    __syncthreads();
    if (tid == 0) st_release_s32(bwd + m, 1);
  }
}

// Host side of the fused substitution: flags[2 nblk] + err[1] + tickets[2] live in `ctl` (2 nblk + 3 ints, zeroed here).
// Returns the number of launches, or -1 when the kernel cannot run on this device.
template <int NR>
inline int cholesky_solve_fused(const double* A, int ld, int nblk, const double* invL, const double* b, double* x, int ldb,
                                int* ctl, cudaStream_t st) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) return -1;
  const int grid = nblk < 2 * sms ? nblk : 2 * sms;   // more CTAs than block rows are never useful
  cudaMemsetAsync(ctl, 0, sizeof(int) * (2 * (size_t)nblk + 3), st);
  chol_subst_kernel<NR><<<grid, SUBST_THREADS, 0, st>>>(A, ld, nblk, invL, b, x, ldb, ctl, ctl + 2 * (size_t)nblk);
  return 1;
}

// Host driver: factor in place.  Returns the number of launches.
// With a side stream (look-ahead): the trailing update of step kb is split into its first tile column and the rest;
// diagonal block kb+1 and its panel solve run on the side stream as soon as that column is done, overlapping the bulk of
// the update, so the latency-bound single-CTA diagonal factorisation leaves the critical path of large matrices.
// ev[0]: column updated (main -> side), ev[1]: panel ready (side -> main).
inline int cholesky_factor(double* A, int ld, int nblk, double* invL, int* info, cudaStream_t st,
                           cudaStream_t side = nullptr, cudaEvent_t* ev = nullptr) {
  int launches = 0;
  const size_t diag_smem = chol_diag_smem_bytes();
  const size_t gemm_smem = (size_t)4 * NB * GS * sizeof(double);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)diag_smem);
    cudaFuncSetAttribute(gemm_nt_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem);
    cudaFuncSetAttribute(gemm_nt_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem);
    cudaFuncSetAttribute(chol_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UPD_SMEM);
    attr_done = true;
  }
  if (side && ev && nblk > 2) {
    // step 0 on the main stream; from then on diag + panel of step kb+1 live on the side stream
    chol_diag_kernel<<<1, 512, diag_smem, st>>>(A, ld, 0, invL, info);
    gemm_nt_kernel<0><<<nblk - 1, 256, gemm_smem, st>>>(A, ld, 0, nblk - 1, invL);
    launches += 2;
    for (int kb = 0; kb + 1 < nblk; ++kb) {
      const int rest = nblk - kb - 1;  // >= 1
      chol_update_kernel<<<2 * rest, 256, UPD_SMEM, st>>>(A, ld, kb, rest, 1, 0, 1);       // column kb+1
      cudaEventRecord(ev[0], st);
      cudaStreamWaitEvent(side, ev[0], 0);
      chol_diag_kernel<<<1, 512, diag_smem, side>>>(A, ld, kb + 1, invL, info);
      launches += 2;
      if (rest > 1) {
        gemm_nt_kernel<0><<<rest - 1, 256, gemm_smem, side>>>(A, ld, kb + 1, rest - 1, invL);  // panel kb+1
        chol_update_kernel<<<(rest - 1) * rest, 256, UPD_SMEM, st>>>(A, ld, kb, rest, 2, 0, 1);  // rest of update kb
        launches += 2;
      }
      cudaEventRecord(ev[1], side);
      cudaStreamWaitEvent(st, ev[1], 0);
    }
    return launches;
  }
  for (int kb = 0; kb < nblk; ++kb) {
    chol_diag_kernel<<<1, 512, diag_smem, st>>>(A, ld, kb, invL, info);
    ++launches;
    const int rest = nblk - kb - 1;
    if (rest > 0) {
      gemm_nt_kernel<0><<<rest, 256, gemm_smem, st>>>(A, ld, kb, rest, invL);
      chol_update_kernel<<<rest * (rest + 1), 256, UPD_SMEM, st>>>(A, ld, kb, rest, 0, 0, 1);
      launches += 2;
    }
  }
  return launches;
}

// A += P P^T on the lower-triangle tiles, P row-major [ld][128].  One launch.  world > 1: only the block columns
// rank, rank + world, ... (the ones this rank factors in the distributed Cholesky).
inline int rank128_update(double* A, int ld, int nblk, const double* P, cudaStream_t st, int rank = 0, int world = 1) {
  const size_t gemm_smem = (size_t)4 * NB * GS * sizeof(double);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(gemm_nt_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem);
    attr_done = true;
  }
  gemm_nt_kernel<2><<<nblk * (nblk + 1) / 2, 256, gemm_smem, st>>>(A, ld, 0, nblk, P, 0, rank, world);
  return 1;
}


}  // namespace d4k
