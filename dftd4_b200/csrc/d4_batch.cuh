// Batched small-molecule path (BASELINE config 3: tens of thousands of ~50-atom molecules, one launch per batch).
// One CTA evaluates one molecule end to end -- the whole of the reference driver src/dftd4/disp.f90:40-134 --
// with every intermediate in shared memory: D4 and EEQ coordination numbers, the EEQ Coulomb matrix with an
// in-shared-memory Cholesky factorisation (constraint by Schur complement, like the large path), reference weights,
// the rational-damping pair sum, the adjoint charge chain rule, the ATM sum (per-edge formulation of d4_atm.cuh over
// dense N x N C6(q=0) tables) and the CN chain rule.  Molecular only (no lattice), at most BATCH_MAXAT atoms per
// molecule, sharp cutoffs.  Tables are the per-species arrays of the union of all elements in the batch.
#pragma once
#include "d4_atm.cuh"
#include "d4_common.cuh"
#include "d4_eeq.cuh"
#include "d4_kernels.cuh"

namespace d4k {

constexpr int BATCH_MAXAT = 64;
constexpr int BATCH_THREADS = 256;
constexpr int BATCH_LDA = BATCH_MAXAT + 1;

struct BatchParams {
  double s6, s8, s9, a1, a2, alp;
  double cut2_2, cut3_2, cutcn_2;   // squared cutoffs
  int do2, do3, want_grad;
};

__host__ __device__ inline size_t batch_smem_bytes() {
  // xyz 3n | cn cne dxd q lam w2 dEdcn dEdq e z 10n | rhs 2(n+1) | A n*(n+1) | C0 n*n | DC0 n*n | 5 weight sets 35n | g 3n | sig 8 warps*6
  const size_t n = BATCH_MAXAT;
  return sizeof(double) * (3 * n + 10 * n + 2 * (n + 1) + n * BATCH_LDA + 2 * n * n + 35 * n + 3 * n + 64) + sizeof(int) * n;
}

__device__ __forceinline__ double block_sum(double v, double* scratch) {
  // scratch: >= 9 doubles of shared memory; all threads of the CTA must call
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < BATCH_THREADS / 32; ++w) t += scratch[w];
    scratch[8] = t;
  }
  __syncthreads();
  return scratch[8];
}

// Solve L L^T x = b in shared memory for two right-hand sides (b0, b1 overwritten by the solutions).
__device__ __forceinline__ void batch_chol_solve2(const double* A, int n, double* b0, double* b1) {
  const int tid = threadIdx.x;
  for (int k = 0; k < n; ++k) {  // forward
    __syncthreads();
    const double y0 = b0[k] / A[k * BATCH_LDA + k], y1 = b1[k] / A[k * BATCH_LDA + k];
    __syncthreads();
    if (tid == 0) { b0[k] = y0; b1[k] = y1; }
    for (int i = k + 1 + tid; i < n; i += BATCH_THREADS) {
      const double l = A[i * BATCH_LDA + k];
      b0[i] -= l * y0; b1[i] -= l * y1;
    }
  }
  for (int k = n - 1; k >= 0; --k) {  // backward: L^T x = y
    __syncthreads();
    const double x0 = b0[k] / A[k * BATCH_LDA + k], x1 = b1[k] / A[k * BATCH_LDA + k];
    __syncthreads();
    if (tid == 0) { b0[k] = x0; b1[k] = x1; }
    for (int i = tid; i < k; i += BATCH_THREADS) {
      const double l = A[k * BATCH_LDA + i];
      b0[i] -= l * x0; b1[i] -= l * x1;
    }
  }
  __syncthreads();
}

// status[m]: 0 ok, 1 EEQ matrix not positive definite.  energies[m]; gradient [total atoms][3] and sigma [nmol][9] may be null.
__global__ void __launch_bounds__(BATCH_THREADS) batch_kernel(int nmol, const long long* __restrict__ offsets,
                                                              const int* __restrict__ species_all, const double* __restrict__ xyz_all,
                                                              const double* __restrict__ charges, Tables tb, BatchParams bp,
                                                              double* __restrict__ energies, double* __restrict__ gradient,
                                                              double* __restrict__ sigma, int* __restrict__ status) {
  extern __shared__ double bsm[];
  const int m = blockIdx.x;
  if (m >= nmol) return;
  const long long at0 = offsets[m];
  const int n = (int)(offsets[m + 1] - at0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int N = BATCH_MAXAT, NW = BATCH_THREADS / 32;
  double* X = bsm;                 // [n][3]
  double* CN = X + 3 * N;          // D4 coordination number
  double* CNE = CN + N;            // raw EEQ coordination number
  double* DXD = CNE + N;           // dX/dCN' * dCN'/dCN_raw
  double* Q = DXD + N;
  double* LAM = Q + N;
  double* W2 = LAM + N;
  double* DCN = W2 + N;            // dE/dCN
  double* DQ = DCN + N;            // dE/dq
  double* EAT = DQ + N;
  double* ZV = EAT + N;            // A^-1 1
  double* B0 = ZV + N;             // [n+1] rhs / solution 0
  double* B1 = B0 + (N + 1);       // [n+1] rhs / solution 1
  double* A = B1 + (N + 1);        // [n][LDA] Coulomb matrix -> Cholesky factor
  double* C0 = A + N * BATCH_LDA;  // [n][n] C6(q=0)
  double* DC0 = C0 + N * N;        // [n][n] d C6(q=0)_ij / dCN_i
  double* GW = DC0 + N * N;        // 5 weight sets [n][7]
  double* GWDCN = GW + 7 * N;
  double* GWDQ = GWDCN + 7 * N;
  double* GW0 = GWDQ + 7 * N;
  double* GWDCN0 = GW0 + 7 * N;
  double* G = GWDCN0 + 7 * N;      // [n][3]
  double* SCR = G + 3 * N;         // [64] scratch: warp partials
  int* ZS = reinterpret_cast<int*>(SCR + 64);
  const int nid = tb.nid;
  const double kcn = 7.5, eeq_cut2 = 625.0, sqrtpi = EEQ_SQRTPI;

  for (int e = tid; e < 3 * n; e += BATCH_THREADS) { X[e] = xyz_all[3 * at0 + e]; G[e] = 0.0; }
  for (int e = tid; e < n; e += BATCH_THREADS) { ZS[e] = species_all[at0 + e]; DCN[e] = 0.0; DQ[e] = 0.0; EAT[e] = 0.0; }
  __syncthreads();

  // ---- 1. coordination numbers (mctc-lib erf count; ncoord.f90:36-75 and multicharge's own CN) ----
  for (int i = warp; i < n; i += NW) {
    const int zi = ZS[i];
    double a = 0.0, b = 0.0;
    for (int j = lane; j < n; j += 32) {
      if (j == i) continue;
      const int hi = max(i, j), lo = min(i, j);  // the reference forms xyz(iat) - xyz(jat) with jat <= iat
      const double r2 = exact_r2(__dsub_rn(X[3 * hi], X[3 * lo]), __dsub_rn(X[3 * hi + 1], X[3 * lo + 1]),
                                 __dsub_rn(X[3 * hi + 2], X[3 * lo + 2]));
      if (r2 < 1.0e-12) continue;
      const int zj = ZS[j];
      const double r = sqrt(r2), rc = tb.pair_rc[zi * nid + zj];
      const double cnt = 0.5 * (1.0 + erf(-kcn * (r - rc) / rc));
      if (!(r2 > bp.cutcn_2)) a += tb.pair_den[zi * nid + zj] * cnt;
      if (!(r2 > eeq_cut2)) b += cnt;
    }
    a = warp_sum(a); b = warp_sum(b);
    if (lane == 0) { CN[i] = a; CNE[i] = b; }
  }
  __syncthreads();

  // ---- 2. EEQ: right-hand sides, Coulomb matrix, Cholesky, charges (multicharge get_charges, disp.f90:88) ----
  for (int i = tid; i < n; i += BATCH_THREADS) {
    const double cut = 8.0, raw = CNE[i];
    const double dcut = exp(cut) / (exp(cut) + exp(raw));
    const double c = log(1.0 + exp(cut)) - log(1.0 + exp(cut - raw));
    const int s = ZS[i];
    const double tmp = tb.kcn[s] / sqrt(c + 1.0e-14);
    B0[i] = -tb.chi[s] + tmp * c;
    B1[i] = 1.0;
    DXD[i] = 0.5 * tmp * dcut;
  }
  for (int e = tid; e < n * n; e += BATCH_THREADS) {
    const int i = e / n, j = e - i * n;
    if (j > i) continue;
    double v;
    if (i == j) {
      v = tb.eeta[ZS[i]] + EEQ_SQRT2PI / tb.rad[ZS[i]];
    } else {
      const double dx = X[3 * j] - X[3 * i], dy = X[3 * j + 1] - X[3 * i + 1], dz = X[3 * j + 2] - X[3 * i + 2];
      const double r = sqrt(dx * dx + dy * dy + dz * dz);
      v = erf(tb.pair_gam[ZS[i] * nid + ZS[j]] * r) / r;
    }
    A[i * BATCH_LDA + j] = v;
  }
  __syncthreads();
  bool bad = false;
  for (int k = 0; k < n; ++k) {  // right-looking Cholesky, lower triangle in place
    const double piv = A[k * BATCH_LDA + k];
    if (!(piv > 0.0)) { bad = true; break; }  // uniform: every thread reads the same value
    const double dinv = 1.0 / sqrt(piv);
    __syncthreads();
    if (tid == 0) A[k * BATCH_LDA + k] = sqrt(piv);
    for (int i = k + 1 + tid; i < n; i += BATCH_THREADS) A[i * BATCH_LDA + k] *= dinv;
    __syncthreads();
    const int mrem = n - 1 - k;
    for (int e = tid; e < mrem * mrem; e += BATCH_THREADS) {
      const int i = k + 1 + e / mrem, j = k + 1 + e % mrem;
      if (j <= i) A[i * BATCH_LDA + j] -= A[i * BATCH_LDA + k] * A[j * BATCH_LDA + k];
    }
    __syncthreads();
  }
  if (bad) {
    if (tid == 0) { status[m] = 1; energies[m] = nan(""); }
    return;
  }
  batch_chol_solve2(A, n, B0, B1);
  const double qtot = charges ? charges[m] : 0.0;
  {
    double a = 0.0, b = 0.0;
    for (int i = tid; i < n; i += BATCH_THREADS) { a += B0[i]; b += B1[i]; }
    const double s0 = block_sum(a, SCR), s1 = block_sum(b, SCR);
    const double mu = (s0 - qtot) / s1;
    for (int i = tid; i < n; i += BATCH_THREADS) { Q[i] = B0[i] - mu * B1[i]; ZV[i] = B1[i]; }
    if (tid == 0) SCR[16] = s1;  // sum of A^-1 1, needed again by the adjoint solve
  }
  __syncthreads();

  // ---- 3. reference weights with q and with q = 0 (d4.f90:236-376, disp.f90:96,121) ----
  for (int i = tid; i < n; i += BATCH_THREADS)
    weights_atom(tb, ZS[i], tb.wf, CN[i], Q[i], GW + 7 * i, GWDCN + 7 * i, GWDQ + 7 * i, GW0 + 7 * i, GWDCN0 + 7 * i);
  __syncthreads();

  double sig[6] = {0, 0, 0, 0, 0, 0};  // per-thread virial partial (xx xy xz yy yz zz), reduced at the end

  // ---- 4. two-body rational damping with on-the-fly C6 (d4.f90:381-473, rational.f90:209-343); one warp per row ----
  if (bp.do2) {
    for (int i = warp; i < n; i += NW) {
      const int zi = ZS[i], nri = tb.nref[zi];
      double en = 0.0, dcn = 0.0, dq = 0.0, gx = 0.0, gy = 0.0, gz = 0.0;
      for (int j = lane; j < n; j += 32) {
        if (j == i) continue;
        const int hi = max(i, j), lo = min(i, j);
        const double sgn = i == hi ? 1.0 : -1.0;   // vec = x_hi - x_lo as the reference forms it; sign towards atom i
        const double ux = __dsub_rn(X[3 * hi], X[3 * lo]), uy = __dsub_rn(X[3 * hi + 1], X[3 * lo + 1]),
                     uz = __dsub_rn(X[3 * hi + 2], X[3 * lo + 2]);
        const double r2 = exact_r2(ux, uy, uz);
        if (r2 > bp.cut2_2 || r2 < 2.220446049250313e-16) continue;
        const int zj = ZS[j], nrj = tb.nref[zj];
        const double* R = tb.refc6 + ((size_t)zi * nid + zj) * D4_MREF * D4_MREF;
        double c6 = 0.0, c6cn = 0.0, c6q = 0.0;
        for (int a = 0; a < nri; ++a) {
          double t = 0.0;
          for (int b = 0; b < nrj; ++b) t += R[a * D4_MREF + b] * GW[7 * j + b];
          c6 += GW[7 * i + a] * t; c6cn += GWDCN[7 * i + a] * t; c6q += GWDQ[7 * i + a] * t;
        }
        const double rrij = tb.pair_rr[zi * nid + zj], r0 = tb.pair_r0[zi * nid + zj];
        const double r0_2 = r0 * r0, r0_6 = r0_2 * r0_2 * r0_2, r0_8 = r0_6 * r0_2;
        const double r4 = r2 * r2, r6 = r4 * r2;
        const double t6 = 1.0 / (r6 + r0_6), t8 = 1.0 / (r6 * r2 + r0_8);
        const double edisp = bp.s6 * t6 + bp.s8 * rrij * t8;
        en += -c6 * edisp * 0.5;
        if (bp.want_grad) {
          const double d6 = -6.0 * r4 * t6 * t6, d8 = -8.0 * r6 * t8 * t8;
          const double f = -c6 * (bp.s6 * d6 + bp.s8 * rrij * d8);
          dcn -= c6cn * edisp; dq -= c6q * edisp;
          const double vx = sgn * ux, vy = sgn * uy, vz = sgn * uz;
          gx += f * vx; gy += f * vy; gz += f * vz;
          const double h = 0.5 * f;
          sig[0] += h * vx * vx; sig[1] += h * vx * vy; sig[2] += h * vx * vz;
          sig[3] += h * vy * vy; sig[4] += h * vy * vz; sig[5] += h * vz * vz;
        }
      }
      en = warp_sum(en); dcn = warp_sum(dcn); dq = warp_sum(dq);
      gx = warp_sum(gx); gy = warp_sum(gy); gz = warp_sum(gz);
      if (lane == 0) { EAT[i] += en; DCN[i] += dcn; DQ[i] += dq; G[3 * i] += gx; G[3 * i + 1] += gy; G[3 * i + 2] += gz; }
    }
  }
  __syncthreads();

  // ---- 5. charge chain rule as an adjoint solve (disp.f90:115-118): A lambda = dE/dq with the constraint eliminated ----
  if (bp.want_grad) {
    double* T1 = C0;  // scratch right-hand side (C0 is filled later)
    for (int i = tid; i < n; i += BATCH_THREADS) { B0[i] = DQ[i]; T1[i] = 0.0; }
    __syncthreads();
    batch_chol_solve2(A, n, B0, T1);
    double a = 0.0;
    for (int i = tid; i < n; i += BATCH_THREADS) a += B0[i];
    const double sw = block_sum(a, SCR);
    const double mu = sw / SCR[16];
    for (int i = tid; i < n; i += BATCH_THREADS) { const double l = B0[i] - mu * ZV[i]; LAM[i] = l; W2[i] = l * DXD[i]; }
    __syncthreads();
    for (int i = warp; i < n; i += NW) {   // g_i -= sum_j (lambda_i q_j + lambda_j q_i) dA_ij/dr_i
      const int zi = ZS[i];
      double gx = 0.0, gy = 0.0, gz = 0.0;
      for (int j = lane; j < n; j += 32) {
        if (j == i) continue;
        const double rx = X[3 * i] - X[3 * j], ry = X[3 * i + 1] - X[3 * j + 1], rz = X[3 * i + 2] - X[3 * j + 2];
        const double gam = tb.pair_gam[zi * nid + ZS[j]];
        const double r2 = rx * rx + ry * ry + rz * rz, arg = gam * gam * r2, r1 = sqrt(r2);
        const double dtmp = 2.0 * gam * exp(-arg) / (sqrtpi * r2) - erf(sqrt(arg)) / (r2 * r1);
        const double f = -(LAM[i] * Q[j] + LAM[j] * Q[i]) * dtmp;
        gx += f * rx; gy += f * ry; gz += f * rz;
        const double h = 0.5 * f;
        sig[0] += h * rx * rx; sig[1] += h * rx * ry; sig[2] += h * rx * rz;
        sig[3] += h * ry * ry; sig[4] += h * ry * rz; sig[5] += h * rz * rz;
      }
      gx = warp_sum(gx); gy = warp_sum(gy); gz = warp_sum(gz);
      if (lane == 0) { G[3 * i] += gx; G[3 * i + 1] += gy; G[3 * i + 2] += gz; }
    }
    __syncthreads();
  }

  // ---- 6. ATM (atm.f90:235-453) with q = 0 weights: every triangle once per edge, one thread per undirected edge ----
  double eacc = 0.0;
  if (bp.do3) {
    // lower triangle (i > j): C6(q=0)_ij ; upper triangle (j > i): cbrt(R0_ij) ; DC0[i][j] = d C6_ij / d CN_i
    for (int e = tid; e < n * n; e += BATCH_THREADS) {
      const int i = e / n, j = e - i * n;
      if (i == j) continue;
      const int zi = ZS[i], zj = ZS[j], nri = tb.nref[zi], nrj = tb.nref[zj];
      const double* R = tb.refc6 + ((size_t)zi * nid + zj) * D4_MREF * D4_MREF;
      double c6 = 0.0, dc = 0.0;
      for (int a = 0; a < nri; ++a) {
        double t = 0.0;
        for (int b = 0; b < nrj; ++b) t += R[a * D4_MREF + b] * GW0[7 * j + b];
        c6 += GW0[7 * i + a] * t; dc += GWDCN0[7 * i + a] * t;
      }
      DC0[i * N + j] = dc;
      if (i > j) C0[i * N + j] = c6; else C0[i * N + j] = cbrt(tb.pair_r0[zi * nid + zj]);
    }
    __syncthreads();
    const int nedge = n * (n - 1) / 2;
    const bool alp16 = bp.alp == 16.0;
    for (int e = tid; e < nedge; e += BATCH_THREADS) {
      int i = (int)floor((sqrt(8.0 * e + 1.0) + 1.0) * 0.5);   // e = i (i - 1)/2 + j, j < i
      while (i * (i - 1) / 2 > e) --i;
      while ((i + 1) * i / 2 <= e) ++i;
      const int j = e - i * (i - 1) / 2;
      const double c6ij = C0[i * N + j], cr0ij = C0[j * N + i];
      double se = 0.0, sg = 0.0;
      for (int k = 0; k < n; ++k) {
        if (k == i || k == j) continue;
        // the three squared distances in the frame of the highest index (atm.f90:332-357)
        const int b = max(i, k), p1 = (k < i) ? j : i, p2 = (k < i) ? k : j;
        const double ax = __dsub_rn(X[3 * p1], X[3 * b]), ay = __dsub_rn(X[3 * p1 + 1], X[3 * b + 1]), az = __dsub_rn(X[3 * p1 + 2], X[3 * b + 2]);
        const double bx = __dsub_rn(X[3 * p2], X[3 * b]), by = __dsub_rn(X[3 * p2 + 1], X[3 * b + 1]), bz = __dsub_rn(X[3 * p2 + 2], X[3 * b + 2]);
        const double rA = exact_r2(ax, ay, az), rB = exact_r2(bx, by, bz);
        const double rC = exact_r2(__dsub_rn(bx, ax), __dsub_rn(by, ay), __dsub_rn(bz, az));
        const double eps = 2.220446049250313e-16;
        if (rA > bp.cut3_2 || rA < eps || rB > bp.cut3_2 || rB < eps || rC > bp.cut3_2 || rC < eps) continue;
        const double r2ij = (k < i) ? rA : rC, r2ik = (k < i) ? rB : rA, r2jk = (k < i) ? rC : rB;
        const double c6ik = C0[max(i, k) * N + min(i, k)], c6jk = C0[max(j, k) * N + min(j, k)];
        const double c0 = cr0ij * C0[min(i, k) * N + max(i, k)] * C0[min(j, k) * N + max(j, k)];
        const double C = fabs(c6ij * c6ik * c6jk);
        const double r2 = r2ij * r2ik * r2jk;
        const double p = atm_inv_sixth_root(r2);
        const double pp = p * p, u = pp * p, u2 = u * u, u3 = u2 * u, u5 = u3 * u2;
        const double v = c0 * p;
        double tmp;
        if (alp16) { const double v2 = v * v, v4 = v2 * v2, v8 = v4 * v4; tmp = v8 * v8; } else tmp = pow(v, bp.alp);
        const double d = fma(6.0, tmp, 1.0);
        const double q = atm_rsqrt(C * d * d);
        const double cf = C * q;
        const double bmc = r2jk - r2ik, bpc = r2jk + r2ik;
        const double ang = fma(0.375 * ((r2ij + bmc) * (r2ij - bmc) * (bpc - r2ij)), u5, u3);
        se += ang * cf;
        if (bp.want_grad) {
          const double tf = tmp * (cf * d * q), m2 = bmc * bmc;
          const double poly = r2ij * fma(r2ij, r2ij + bpc, fma(2.0, bpc * bpc, m2)) - 5.0 * (m2 * bpc);
          sg += cf * (0.375 * poly * u5 - 2.0 * bp.alp * ang * tf);
        }
      }
      se *= -bp.s9; sg *= -bp.s9;
      eacc -= se * (1.0 / 3.0);
      if (bp.want_grad) {
        const double vx = X[3 * j] - X[3 * i], vy = X[3 * j + 1] - X[3 * i + 1], vz = X[3 * j + 2] - X[3 * i + 2];
        const double r2ij = vx * vx + vy * vy + vz * vz;
        const double coef = sg / r2ij;
        const double gx = coef * vx, gy = coef * vy, gz = coef * vz;
        atomicAdd(&G[3 * i], -gx); atomicAdd(&G[3 * i + 1], -gy); atomicAdd(&G[3 * i + 2], -gz);
        atomicAdd(&G[3 * j], gx); atomicAdd(&G[3 * j + 1], gy); atomicAdd(&G[3 * j + 2], gz);
        sig[0] += gx * vx; sig[1] += gx * vy; sig[2] += gx * vz; sig[3] += gy * vy; sig[4] += gy * vz; sig[5] += gz * vz;
        atomicAdd(&DCN[i], -0.5 * se * DC0[i * N + j] / c6ij);
        atomicAdd(&DCN[j], -0.5 * se * DC0[j * N + i] / c6ij);
      }
    }
  }
  __syncthreads();

  // ---- 7. CN chain rule (ncoord.f90:78-117, disp.f90:127-130) fused with the EEQ CN' term of the charge chain rule ----
  if (bp.want_grad) {
    for (int i = warp; i < n; i += NW) {
      const int zi = ZS[i];
      double gx = 0.0, gy = 0.0, gz = 0.0;
      for (int j = lane; j < n; j += 32) {
        if (j == i) continue;
        const int hi = max(i, j), lo = min(i, j);
        const double sgn = i == hi ? 1.0 : -1.0;
        const double ux = __dsub_rn(X[3 * hi], X[3 * lo]), uy = __dsub_rn(X[3 * hi + 1], X[3 * lo + 1]),
                     uz = __dsub_rn(X[3 * hi + 2], X[3 * lo + 2]);
        const double r2 = exact_r2(ux, uy, uz);
        if (r2 < 1.0e-12) continue;
        const int zj = ZS[j];
        const double r1 = sqrt(r2), rc = tb.pair_rc[zi * nid + zj];
        const double ex = kcn * (r1 - rc) / rc;
        const double dc = -(kcn * exp(-ex * ex)) / (rc * sqrtpi);
        double f = 0.0;
        if (!(r2 > bp.cutcn_2)) f += (DCN[i] + DCN[j]) * tb.pair_den[zi * nid + zj] * dc;
        if (!(r2 > eeq_cut2)) f += (W2[i] + W2[j]) * dc;
        f /= r1;
        const double vx = sgn * ux, vy = sgn * uy, vz = sgn * uz;
        gx += f * vx; gy += f * vy; gz += f * vz;
        const double h = 0.5 * f;
        sig[0] += h * vx * vx; sig[1] += h * vx * vy; sig[2] += h * vx * vz;
        sig[3] += h * vy * vy; sig[4] += h * vy * vz; sig[5] += h * vz * vz;
      }
      gx = warp_sum(gx); gy = warp_sum(gy); gz = warp_sum(gz);
      if (lane == 0) { G[3 * i] += gx; G[3 * i + 1] += gy; G[3 * i + 2] += gz; }
    }
  }
  __syncthreads();

  // ---- 8. results ----
  for (int i = tid; i < n; i += BATCH_THREADS) eacc += EAT[i];
  const double etot = block_sum(eacc, SCR);
  if (tid == 0) { energies[m] = etot; status[m] = 0; }
  if (bp.want_grad) {
    if (gradient) for (int e = tid; e < 3 * n; e += BATCH_THREADS) gradient[3 * at0 + e] = G[e];
    if (sigma) {
      double s6[6];
      for (int c = 0; c < 6; ++c) s6[c] = block_sum(sig[c], SCR);
      if (tid == 0) {
        double* o = sigma + (size_t)m * 9;
        o[0] = s6[0]; o[1] = s6[1]; o[2] = s6[2]; o[3] = s6[1]; o[4] = s6[3]; o[5] = s6[4]; o[6] = s6[2]; o[7] = s6[4]; o[8] = s6[5];
      }
    }
  }
}

}  // namespace d4k
