// EEQ-2019 charge model on the device: Coulomb matrix build (molecular and Ewald/Wigner-Seitz periodic form),
// constraint elimination and the adjoint form of the charge chain rule.
//
// Reference call sites: src/dftd4/disp.f90:86-92 (get_charges) and :115-118 (d4_gemv(dqdr, dEdq, gradient) /
// d4_gemv(dqdL, dEdq, sigma)).  multicharge v0.5.0 itself is not part of the reference tree; formulas restated from
// the published model (DESIGN.md "EEQ").  Instead of dqdr(3,N,N) the chain rule is evaluated as
//     gradient += dqdr . dEdq  =  - sum_i lambda_i [ d(A q)_i/dr - dX_i/dr ],   [A 1; 1^T 0][lambda; mu] = [dEdq; 0]
// i.e. one extra solve with the same Cholesky factor and one O(N^2) pair pass; identical to round-off.
#pragma once
#include "d4_common.cuh"
#include "d4_kernels.cuh"
#include "d4_erfc_coef.inc"

namespace d4k {

#define EEQ_SQRT2PI 0.79788456080286535587989211986876373  // sqrt(2/pi)
#define EEQ_SQRTPI 1.77245385090551602729816748334114518

// Ewald / Wigner-Seitz data (device pointers)
struct EwaldDev {
  double alpha, volume;
  int ndir, nrec, nwsc;
  const double* dir;      // [ndir][3]
  const double* rec;      // [nrec][3]
  const double* rec_fac;  // [nrec]
  const double* wsc;      // [nwsc][3]
  const double* diag;     // [nid]   A_ii without geometry dependence on the atom (real-space self images + constants)
  const double* diag_ds;  // [nid][9] strain derivative of the self-image sum
  int nh;                 // reciprocal half set: one vector of every +G/-G pair (<= EEQ_NH)
  const double* gh;       // [nh][3]
  const double* fh;       // [nh] 4 pi/V exp(-G^2/4 alpha^2)/G^2
  // piecewise polynomials of the alpha part of the real-space terms in r^2 (eeq_table_kernel)
  const double* tabf;     // [tab_n][8]  erfc(alpha r)/r
  const double* tabg;     // [tab_n][8]  -(2 alpha/sqrt(pi) e^{-alpha^2 r^2} + erfc(alpha r)/r)/r^2
  int tab_n, tab_base;    // intervals; index of the first interval = (high word of r^2) >> 16
};
constexpr int EEQ_TAB_DEG = 7;        // degree of the pieces
constexpr int EEQ_TAB_EMIN = -2;      // first octave of r^2: 2^-2 (r = 0.5 bohr); closer pairs take the direct evaluation
constexpr int EEQ_TAB_MAXN = 16 * 24; // intervals the kernels reserve shared memory for (16 per octave of r^2)
constexpr int EEQ_NH = 64;   // phase matrix P[ld][128]: columns 0..63 sqrt(2 f_k) cos(G_k.r), 64..127 sqrt(2 f_k) sin(G_k.r)
constexpr int EEQ_TILE = 16; // pair tiles of the real-space kernels
#ifndef EEQ_DIR_XCUT
#define EEQ_DIR_XCUT 6.0     // real-space Ewald images with alpha * r beyond this are skipped (erfc(6) = 2.2e-17)
#endif

// ---- molecular -------------------------------------------------------------------------------------------------
// A[i][j] (j < i) = erf(gam r)/r, A[i][i] = eta_i + sqrt(2/pi)/rad_i, padding = identity.  One thread per element.
__global__ void eeq_build_0d_kernel(int nat, int ld, const double* __restrict__ xyz, const int* __restrict__ species,
                                    Tables tb, double* __restrict__ A) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= ld || j > i) return;
  double v;
  if (i >= nat) {
    v = (i == j) ? 1.0 : 0.0;
  } else if (i == j) {
    const int s = species[i];
    v = tb.eeta[s] + EEQ_SQRT2PI / tb.rad[s];
  } else {
    const double dx = xyz[3 * j] - xyz[3 * i], dy = xyz[3 * j + 1] - xyz[3 * i + 1], dz = xyz[3 * j + 2] - xyz[3 * i + 2];
    const double r2 = dx * dx + dy * dy + dz * dz;
    const double gam = tb.pair_gam[species[i] * tb.nid + species[j]];
    const double r = sqrt(r2);
    v = erf(gam * r) / r;
  }
  A[(size_t)i * ld + j] = v;
}

// ---- periodic ----------------------------------------------------------------------------------------------------
// Wigner-Seitz images of the pair vector rij (multicharge wignerseitz.f90 get_pairs): all of the 27 translations whose
// squared distance lies within tol = 0.01 of the minimum.  Returns a 27-bit mask.
__device__ __forceinline__ unsigned wsc_mask(double rx, double ry, double rz, const double* __restrict__ wsc,
                                             const float* __restrict__ wscf, int nwsc) {
  // FP32 screening (the FP32 error of a squared distance of order 1e4 is ~1e-3, the selection tolerance 1e-2): only the
  // images within 0.03 of the FP32 minimum can be selected; the selection itself is done on their FP64 distances.
  const float fx = (float)rx, fy = (float)ry, fz = (float)rz;
  float bestf = 3.0e38f;
  for (int t = 0; t < nwsc; ++t) {
    const float vx = fx - wscf[3 * t], vy = fy - wscf[3 * t + 1], vz = fz - wscf[3 * t + 2];
    bestf = fminf(bestf, fmaf(vz, vz, fmaf(vy, vy, vx * vx)));
  }
  const float lim = bestf + 0.03f + 1.0e-6f * bestf;
  unsigned cand = 0u;
  for (int t = 0; t < nwsc; ++t) {
    const float vx = fx - wscf[3 * t], vy = fy - wscf[3 * t + 1], vz = fz - wscf[3 * t + 2];
    cand |= (fmaf(vz, vz, fmaf(vy, vy, vx * vx)) <= lim ? 1u : 0u) << t;
  }
  double best = 1e300;
  for (unsigned mm = cand; mm; mm &= mm - 1) {
    const int t = __ffs(mm) - 1;
    const double vx = rx - wsc[3 * t], vy = ry - wsc[3 * t + 1], vz = rz - wsc[3 * t + 2];
    best = fmin(best, vx * vx + vy * vy + vz * vz);
  }
  unsigned m = 0;
  for (unsigned mm = cand; mm; mm &= mm - 1) {
    const int t = __ffs(mm) - 1;
    const double vx = rx - wsc[3 * t], vy = ry - wsc[3 * t + 1], vz = rz - wsc[3 * t + 2];
    const double r2 = vx * vx + vy * vy + vz * vz;
    if (fabs(r2 - best) <= 0.01) m |= 1u << t;
  }
  return m;
}

// Real-space Ewald sum over the 5x5x5 image block around the pair vector v (multicharge get_amat_dir_3d /
// get_damat_dir_3d): value a = sum_T (erf(gam r) - erf(alp r))/r, written as (erfc(alp r) - erfc(gam r))/r, and with
// GRAD the radial derivative sums dg = sum_T c (v+T), ds = sum_T c (v+T)(v+T)^T,
//   c = [2/sqrt(pi) (gam e^{-gam^2 r^2} - alp e^{-alp^2 r^2}) + (erfc(gam r) - erfc(alp r))/r] / r^2.
// The gam terms vanish to below 1e-17 once gam r > 6.2 (all but the nearest images).
// erfc(x) = ex * P(u) for 0 <= x <= 7 given ex = exp(-x^2): degree-18 polynomial in the mapped variable
// t = (x - 4)/(x + 4) (tools/gen_erfc.py, < 1e-15 relative); the reciprocal comes from the hardware seed + 2 Newton steps.
__device__ __forceinline__ double ewald_erfc_scaled(double x, double ex) {
  const double d = x + 4.0;
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  y = fma(y, fma(-d, y, 1.0), y);
  y = fma(y, fma(-d, y, 1.0), y);
  const double u = fma(D4_ERFC_A, (x - 4.0) * y, D4_ERFC_B);
  double p = d4_erfc_coef[D4_ERFC_DEG];
#pragma unroll
  for (int k = D4_ERFC_DEG - 1; k >= 0; --k) p = fma(p, u, d4_erfc_coef[k]);
  return p * ex;
}

// The image block is walked in two classes, decided per image from the FP32 distance (the class boundary is soft, both
// evaluations are accurate on either side of it):
//   NEAR  alp r <  4.5 : FP64, as described above (the gam terms only matter for the nearest images);
//   FAR   4.5 <= alp r <= 6 : erfc(alp r)/r <= 2e-10/r enters matrix elements of order 1e-2..1, i.e. it is needed to 7
//         digits -- evaluated in FP32 on the otherwise idle FP32 / SFU pipes (absolute error < 1e-18), summed in FP32
//         and added to the FP64 sum once.  For the 5x5x5 block of a cell whose alpha multicharge's search picks, more
//         than half of the contributing images are FAR.
// Every lane first collects its own NEAR / FAR image sets as bit masks and then walks them (the image index is data,
// not control flow), so lanes whose pair vectors differ do not serialise each other's images.
// ---- the alpha part of a NEAR image as a table lookup ------------------------------------------------------------------
// f(r^2) = erfc(alpha r)/r and g(r^2) = -(2 alpha/sqrt(pi) e^{-alpha^2 r^2} + erfc(alpha r)/r)/r^2 depend on the lattice only
// through alpha.  They are smooth in r^2 on the scale min(r, 1/(2 alpha^2 r)), so on intervals that cover 1/16 of an octave
// of r^2 (index = exponent and top four mantissa bits: no search) a degree-7 polynomial in the local variable
// s in [-1, 1] reproduces them to 3e-15 absolute where they are of order one and far below 1e-18 absolute where they are
// small (checked against scipy's erfc for alpha from 0.014 to 0.3): 8 fused multiply-adds and four 16-byte loads
// instead of exp + reciprocal + the degree-18 erfc polynomial (~150 instructions).  One thread per interval fits the
// piece through the eight Chebyshev nodes with the direct evaluation and converts it to monomials.
__global__ void eeq_table_kernel(double alpha, int nint, double* __restrict__ tabf, double* __restrict__ tabg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nint) return;
  constexpr int N = EEQ_TAB_DEG + 1;
  const int oct = i >> 4, m = i & 15;
  const double lo = ldexp(1.0 + m / 16.0, EEQ_TAB_EMIN + oct), hi = ldexp(1.0 + (m + 1) / 16.0, EEQ_TAB_EMIN + oct);
  const double mid = 0.5 * (lo + hi), half = 0.5 * (hi - lo), tsp = 2.0 / EEQ_SQRTPI;
  double xk[N], fk[N], gk[N];
  for (int k = 0; k < N; ++k) {
    xk[k] = cospi((k + 0.5) / N);
    const double r2 = fma(half, xk[k], mid), r1 = sqrt(r2), xa = alpha * r1;
    const double ex = exp(-xa * xa);
    const double ea = xa <= 7.0 ? ewald_erfc_scaled(xa, ex) : erfc(xa);
    fk[k] = ea / r1;
    gk[k] = -(tsp * alpha * ex + ea / r1) / r2;
  }
  for (int which = 0; which < 2; ++which) {
    const double* v = which ? gk : fk;
    double c[N];   // Chebyshev coefficients
    for (int j = 0; j < N; ++j) {
      double a = 0.0;
      for (int k = 0; k < N; ++k) a += v[k] * cospi(j * (k + 0.5) / N);
      c[j] = a * (j == 0 ? 1.0 : 2.0) / N;
    }
    // sum_j c_j T_j(s) -> monomials: T_0 = 1, T_1 = s, T_{j+1} = 2 s T_j - T_{j-1}
    double p[N] = {0}, t0[N] = {0}, t1[N] = {0}, t2[N];
    t0[0] = 1.0; t1[1] = 1.0;
    for (int d = 0; d < N; ++d) p[d] = c[0] * t0[d] + c[1] * t1[d];
    for (int j = 2; j < N; ++j) {
      for (int d = 0; d < N; ++d) t2[d] = (d > 0 ? 2.0 * t1[d - 1] : 0.0) - t0[d];
      for (int d = 0; d < N; ++d) { p[d] += c[j] * t2[d]; t0[d] = t1[d]; t1[d] = t2[d]; }
    }
    double* out = (which ? tabg : tabf) + (size_t)i * N;
    for (int d = 0; d < N; ++d) out[d] = p[d];
  }
}

// Piece index and local variable of r^2 (normal, positive): the interval is named by the exponent and the top four
// mantissa bits, s = 32 (mantissa in [1, 2)) - (33 + 2 m).
__device__ __forceinline__ int eeq_tab_locate(double r2, int base, double& sloc) {
  const int hi = __double2hiint(r2);
  const double t = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(r2));
  const int m = (hi >> 16) & 15;
  sloc = fma(t, 32.0, -(double)(33 + 2 * m));
  return (hi >> 16) - base;
}
__device__ __forceinline__ double eeq_tab_eval(const double* __restrict__ piece, double sloc) {
  const double2 c01 = reinterpret_cast<const double2*>(piece)[0], c23 = reinterpret_cast<const double2*>(piece)[1],
                c45 = reinterpret_cast<const double2*>(piece)[2], c67 = reinterpret_cast<const double2*>(piece)[3];
  double p = fma(c67.y, sloc, c67.x);
  p = fma(p, sloc, c45.y); p = fma(p, sloc, c45.x); p = fma(p, sloc, c23.y); p = fma(p, sloc, c23.x);
  p = fma(p, sloc, c01.y); p = fma(p, sloc, c01.x);
  return p;
}

constexpr int EEQ_MAXDIR = 128;   // images per block (multicharge: 125)
// Shared-memory tables of the real-space kernels: FP64 image vectors + Wigner-Seitz vectors, and FP32 copies -- the image
// block as float4 padded to EEQ_MAXDIR with far-away points (classified as neither NEAR nor FAR).
__host__ __device__ inline size_t eeq_dir_smem_bytes(int ndir, int nwsc, int tab_n) {
  return sizeof(double) * (size_t)(3 * ndir + 3 * nwsc + 2) + sizeof(float) * (size_t)(4 * EEQ_MAXDIR + 3 * nwsc + 3) +
         sizeof(double) * (size_t)tab_n * (EEQ_TAB_DEG + 1);
}
__device__ __forceinline__ void eeq_stage_tables(const EwaldDev& ew, double* sdir, double* swsc, float* sdirf, float* swscf,
                                                 double* stab, const double* __restrict__ tab, int tid, int nthreads) {
  for (int e = tid; e < ew.tab_n * (EEQ_TAB_DEG + 1); e += nthreads) stab[e] = tab[e];
  for (int e = tid; e < 3 * ew.ndir; e += nthreads) sdir[e] = ew.dir[e];
  for (int e = tid; e < 3 * ew.nwsc; e += nthreads) { swsc[e] = ew.wsc[e]; swscf[e] = (float)ew.wsc[e]; }
  for (int t = tid; t < EEQ_MAXDIR; t += nthreads) {
    float4 v = make_float4(1.0e18f, 1.0e18f, 1.0e18f, 0.0f);
    if (t < ew.ndir) v = make_float4((float)ew.dir[3 * t], (float)ew.dir[3 * t + 1], (float)ew.dir[3 * t + 2], 0.0f);
    reinterpret_cast<float4*>(sdirf)[t] = v;
  }
}
template <bool GRAD>
__device__ __forceinline__ void ewald_dir_sum(double vx, double vy, double vz, double gam, double alp,
                                              const double* __restrict__ sdir, const float* __restrict__ sdirf, int ndir,
                                              const double* __restrict__ stab, int tab_n, int tab_base, double wsw, double& a,
                                              double* dg, double* ds, unsigned& nimg) {
  const double gam2 = gam * gam, alp2 = alp * alp, tsp = 2.0 / EEQ_SQRTPI;
  const float vxf = (float)vx, vyf = (float)vy, vzf = (float)vz, alp2f = (float)alp2, alpf = (float)alp;
  unsigned nearm[EEQ_MAXDIR / 32], farm[EEQ_MAXDIR / 32];
  // (sdirf is padded to EEQ_MAXDIR images with far-away points: no bounds test, no branch in the classification)
#pragma unroll
  for (int w = 0; w < EEQ_MAXDIR / 32; ++w) {
    unsigned nm = 0u, fm = 0u;
#pragma unroll 4
    for (int b = 0; b < 32; ++b) {
      const float4 d = reinterpret_cast<const float4*>(sdirf)[32 * w + b];
      const float x = vxf + d.x, y = vyf + d.y, z = vzf + d.z;
      const float xa2 = alp2f * fmaf(z, z, fmaf(y, y, x * x));
      // (images beyond alp r = 6 add erfc(6)/r < 2e-17/r: dropped, like before)
      nm |= (xa2 < 20.25f ? 1u : 0u) << b;
      fm |= ((xa2 >= 20.25f && xa2 <= (float)(EEQ_DIR_XCUT * EEQ_DIR_XCUT)) ? 1u : 0u) << b;
    }
    nearm[w] = nm; farm[w] = fm;
    nimg += __popc(nm) + __popc(fm);   // contributing images (bench.py: algorithmic work of the periodic EEQ kernels)
  }
  // ---- NEAR images: FP64 ----
#pragma unroll
  for (int w = 0; w < EEQ_MAXDIR / 32; ++w) {
#pragma unroll 1
    for (unsigned m = nearm[w]; m; m &= m - 1) {
      const int t = 32 * w + __ffs(m) - 1;
      const double x = vx + sdir[3 * t], y = vy + sdir[3 * t + 1], z = vz + sdir[3 * t + 2];
      const double r2 = x * x + y * y + z * z;
      if (r2 < 2.220446049250313e-16) continue;  // r1 < sqrt(epsilon)
      const double xg2 = gam2 * r2;
      const bool nearg = xg2 < 38.5;                       // erfc(6.2) = 2e-18: only the nearest image of a close pair
      double sloc;
      const int piece = eeq_tab_locate(r2, tab_base, sloc);
      if (piece >= 0 && piece < tab_n && !nearg) {
        // the alpha part from its piecewise polynomial (eeq_table_kernel)
        const double v = eeq_tab_eval(stab + (size_t)piece * (EEQ_TAB_DEG + 1), sloc) * wsw;
        if (!GRAD) {
          a += v;
        } else {
          dg[0] += v * x; dg[1] += v * y; dg[2] += v * z;
          ds[0] += v * x * x; ds[1] += v * x * y; ds[2] += v * x * z; ds[3] += v * y * y; ds[4] += v * y * z; ds[5] += v * z * z;
        }
        continue;
      }
      // direct evaluation: close pairs (the gam terms matter) and distances outside the table
      const double xa2 = alp2 * r2;
      const double ri = rsqrt(r2), r1 = r2 * ri;
      double eg = 0.0, exg = 0.0;
      const double exa = exp(-xa2), ea = ewald_erfc_scaled(alp * r1, exa);
      if (nearg) { exg = exp(-xg2); eg = ewald_erfc_scaled(gam * r1, exg); }
      if (!GRAD) {
        a += (ea - eg) * ri * wsw;
      } else {
        const double c = (tsp * (gam * exg - alp * exa) + (eg - ea) * ri) * (ri * ri) * wsw;
        dg[0] += c * x; dg[1] += c * y; dg[2] += c * z;
        ds[0] += c * x * x; ds[1] += c * x * y; ds[2] += c * x * z; ds[3] += c * y * y; ds[4] += c * y * z; ds[5] += c * z * z;
      }
    }
  }
  // ---- FAR images: FP32 (the gam terms are zero to all digits out here) ----
  float fa = 0.0f, fg[3] = {0.0f, 0.0f, 0.0f}, fs[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
  for (int w = 0; w < EEQ_MAXDIR / 32; ++w) {
#pragma unroll 1
    for (unsigned m = farm[w]; m; m &= m - 1) {
      const int t = 32 * w + __ffs(m) - 1;
      const float4 d = reinterpret_cast<const float4*>(sdirf)[t];
      const float x = vxf + d.x, y = vyf + d.y, z = vzf + d.z;
      const float r2 = fmaf(z, z, fmaf(y, y, x * x));
      const float ri = rsqrtf(r2);
      // erfc(u) for u = alp r >= 4.5 from its asymptotic series, e^{-u^2}/(u sqrt(pi)) (1 - 1/(2u^2) + 3/(4u^4) - 15/(8u^6)
      // + 105/(16u^8)): the first dropped term is < 9e-6 relative, i.e. < 2e-17 absolute in the matrix element
      const float u2 = alp2f * r2, iu2 = __frcp_rn(u2);
      const float ex = __expf(-u2);
      const float ser = fmaf(fmaf(fmaf(fmaf(6.5625f, iu2, -1.875f), iu2, 0.75f), iu2, -0.5f), iu2, 1.0f);
      const float ea = ex * ser * (0.5641895835477563f / alpf) * ri;   // 1/(u sqrt(pi)) = ri / (alp sqrt(pi))
      if (!GRAD) {
        fa = fmaf(ea, ri, fa);
      } else {
        const float c = -(((float)tsp * alpf) * ex + ea * ri) * (ri * ri);
        fg[0] = fmaf(c, x, fg[0]); fg[1] = fmaf(c, y, fg[1]); fg[2] = fmaf(c, z, fg[2]);
        const float cx = c * x, cy = c * y;
        fs[0] = fmaf(cx, x, fs[0]); fs[1] = fmaf(cx, y, fs[1]); fs[2] = fmaf(cx, z, fs[2]);
        fs[3] = fmaf(cy, y, fs[3]); fs[4] = fmaf(cy, z, fs[4]); fs[5] = fmaf(c * z, z, fs[5]);
      }
    }
  }
  if (!GRAD) {
    a += (double)fa * wsw;
  } else {
#pragma unroll
    for (int c = 0; c < 3; ++c) dg[c] += (double)fg[c] * wsw;
#pragma unroll
    for (int c = 0; c < 6; ++c) ds[c] += (double)fs[c] * wsw;
  }
}

// unrank a lower-triangle tile index p = ti (ti + 1)/2 + tj
__device__ __forceinline__ void unrank_tile(long long p, int& ti, int& tj) {
  ti = (int)floor((sqrt(8.0 * (double)p + 1.0) - 1.0) * 0.5);
  while ((long long)(ti + 1) * (ti + 2) / 2 <= p) ++ti;
  while ((long long)ti * (ti + 1) / 2 > p) --ti;
  tj = (int)(p - (long long)ti * (ti + 1) / 2);
}

// Phase matrix of the reciprocal-space sum: sum_G f_G cos(G.(ri - rj)) = sum_{k in half set} 2 f_k (c_ik c_jk + s_ik s_jk)
// = (P P^T)_ij, so the reciprocal part of A is one rank-128 DMMA update (rank128_update) instead of 124 cosines per pair.
__global__ void eeq_phase_kernel(int nat, int ld, const double* __restrict__ xyz, EwaldDev ew, double* __restrict__ P) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (long long)ld * EEQ_NH) return;
  const int i = (int)(e / EEQ_NH), k = (int)(e % EEQ_NH);
  double c = 0.0, s = 0.0;
  if (i < nat && k < ew.nh) {
    const double gv = xyz[3 * i] * ew.gh[3 * k] + xyz[3 * i + 1] * ew.gh[3 * k + 1] + xyz[3 * i + 2] * ew.gh[3 * k + 2];
    sincos(gv, &s, &c);
    const double sf = sqrt(2.0 * ew.fh[k]);
    c *= sf; s *= sf;
  }
  P[(size_t)i * 2 * EEQ_NH + k] = c;
  P[(size_t)i * 2 * EEQ_NH + EEQ_NH + k] = s;
}

// Real-space part, diagonal and padding of the periodic Coulomb matrix; one thread per lower-triangle element, grid =
// lower-triangle 16x16 tiles only.
__global__ void __launch_bounds__(EEQ_TILE * EEQ_TILE, 3) eeq_build_3d_kernel(int nat, int ld, const double* __restrict__ xyz,
                                                                           const int* __restrict__ species, Tables tb,
                                                                           EwaldDev ew, double* __restrict__ A,
                                                                           long long tile_offset, int W, int rank,
                                                                           unsigned long long* __restrict__ counters) {
  extern __shared__ double sh[];
  double* sdir = sh;
  double* swsc = sdir + 3 * ew.ndir;
  double* stab = swsc + 3 * ew.nwsc + ((3 * ew.ndir + 3 * ew.nwsc) & 1);   // 16-byte aligned
  float* sdirf = reinterpret_cast<float*>(stab + (size_t)ew.tab_n * (EEQ_TAB_DEG + 1));
  float* swscf = sdirf + 4 * EEQ_MAXDIR;
  const int tid = threadIdx.x;
  eeq_stage_tables(ew, sdir, swsc, sdirf, swscf, stab, ew.tabf, tid, EEQ_TILE * EEQ_TILE);
  __syncthreads();
  int ti, tj;
  if (W > 0) {
    // distributed factorisation: this rank only needs the 128-wide block columns it owns (rank, rank + W, ...);
    // blockIdx.y = ordinal of the owned block column, blockIdx.x walks its 16x16 tiles from the diagonal block down
    const int bc = rank + (int)blockIdx.y * W;
    constexpr int TPB = 128 / EEQ_TILE;
    ti = bc * TPB + (int)(blockIdx.x / TPB);
    tj = bc * TPB + (int)(blockIdx.x % TPB);
    if (ti >= ld / EEQ_TILE) return;
  } else {
    unrank_tile(blockIdx.x + tile_offset, ti, tj);  // a rank builds a contiguous band of tile rows (tile_offset = first tile)
  }
  const int i = ti * EEQ_TILE + tid / EEQ_TILE, j = tj * EEQ_TILE + tid % EEQ_TILE;
  if (i >= ld || j > i) return;
  double v;
  if (i >= nat) {
    v = (i == j) ? 1.0 : 0.0;
  } else if (i == j) {
    v = ew.diag[species[i]];
  } else {
    const double rx = xyz[3 * i] - xyz[3 * j], ry = xyz[3 * i + 1] - xyz[3 * j + 1], rz = xyz[3 * i + 2] - xyz[3 * j + 2];
    const double gam = tb.pair_gam[species[i] * tb.nid + species[j]];
    const unsigned m = wsc_mask(rx, ry, rz, swsc, swscf, ew.nwsc);
    const double wsw = 1.0 / (double)__popc(m);
    v = 0.0;
    unsigned nimg = 0u;
    // every lane walks ITS OWN selected images (usually one): the image index is data, not control flow
    for (unsigned mm = m; mm; mm &= mm - 1) {
      const int t = __ffs(mm) - 1;
      ewald_dir_sum<false>(rx - swsc[3 * t], ry - swsc[3 * t + 1], rz - swsc[3 * t + 2], gam, ew.alpha, sdir, sdirf, ew.ndir, stab,
                           ew.tab_n, ew.tab_base, wsw, v, nullptr, nullptr, nimg);
    }
    if (counters) {
      const unsigned am = __activemask();   // taken once: the reduction and the leader election must see the same lanes
      const unsigned tot = __reduce_add_sync(am, nimg);
      if ((threadIdx.x & 31) == __ffs(am) - 1) atomicAdd(counters + 5, (unsigned long long)tot);
    }
  }
  A[(size_t)i * ld + j] = v;
}

// ---- constraint elimination --------------------------------------------------------------------------------------
// sol[0*ld + i] = (A^-1 X)_i, sol[1*ld + i] = (A^-1 1)_i.  q = x0 - mu x1 with mu = (sum x0 - Q)/sum x1.
// aux[0] = sum x1 (kept for the adjoint solve).  Single block.
__global__ void __launch_bounds__(1024) eeq_charges_kernel(int nat, int ld, const double* __restrict__ sol, double qtot,
                                                           double* __restrict__ q, double* __restrict__ aux) {
  __shared__ double s0[32], s1[32];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < nat; i += blockDim.x) { a += sol[i]; b += sol[ld + i]; }
  a = warp_sum(a); b = warp_sum(b);
  if ((threadIdx.x & 31) == 0) { s0[threadIdx.x >> 5] = a; s1[threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x < 32) {
    a = threadIdx.x < (blockDim.x >> 5) ? s0[threadIdx.x] : 0.0;
    b = threadIdx.x < (blockDim.x >> 5) ? s1[threadIdx.x] : 0.0;
    a = warp_sum(a); b = warp_sum(b);
    if (threadIdx.x == 0) { s0[0] = a; s1[0] = b; aux[0] = b; }
  }
  __syncthreads();
  const double mu = (s0[0] - qtot) / s1[0];
  for (int i = threadIdx.x; i < nat; i += blockDim.x) q[i] = sol[i] - mu * sol[ld + i];
}

// adjoint: w = A^-1 dEdq (in sol[0..]), z = A^-1 1 (zvec) -> lambda = w - (sum w / sum z) z ; w2 = lambda * dxdcn_cut
__global__ void __launch_bounds__(1024) eeq_lambda_kernel(int nat, const double* __restrict__ w, const double* __restrict__ zvec,
                                                          const double* __restrict__ aux, const double* __restrict__ dxdcn_cut,
                                                          double* __restrict__ lambda, double* __restrict__ w2) {
  __shared__ double s0[32];
  double a = 0.0;
  for (int i = threadIdx.x; i < nat; i += blockDim.x) a += w[i];
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) s0[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x < 32) {
    a = threadIdx.x < (blockDim.x >> 5) ? s0[threadIdx.x] : 0.0;
    a = warp_sum(a);
    if (threadIdx.x == 0) s0[0] = a;
  }
  __syncthreads();
  const double mu = s0[0] / aux[0];
  for (int i = threadIdx.x; i < nat; i += blockDim.x) {
    const double l = w[i] - mu * zvec[i];
    lambda[i] = l;
    w2[i] = l * dxdcn_cut[i];
  }
}

__global__ void eeq_pack_rhs_kernel(int nat, int ld, const double* __restrict__ src, double* __restrict__ rhs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ld) rhs[i] = i < nat ? src[i] : 0.0;
}

// ---- adjoint pair pass, molecular ---------------------------------------------------------------------------------
// g_i -= sum_j (lambda_i q_j + lambda_j q_i) dA_ij/dr_i ; sigma likewise.  One CTA per row, threads stride over j.
constexpr int EEQG_THREADS = 256;

__global__ void __launch_bounds__(EEQG_THREADS) eeq_grad_kernel(int nrows, const int* __restrict__ rows, int nat,
                                                                const double* __restrict__ xyz, const int* __restrict__ species,
                                                                Tables tb, const double* __restrict__ q,
                                                                const double* __restrict__ lambda, double* __restrict__ grad,
                                                                double* __restrict__ sig_part) {
  const int tid = threadIdx.x;
  const int i = rows[blockIdx.x];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  const int zs = species[i];
  const double li = lambda[i], qi = q[i];
  double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};  // gx gy gz sxx sxy sxz syy syz szz
  for (int j = tid; j < nat; j += EEQG_THREADS) {
    if (j == i) continue;
    const double rx = xi - xyz[3 * j], ry = yi - xyz[3 * j + 1], rz = zi - xyz[3 * j + 2];
    const double gam = tb.pair_gam[zs * tb.nid + species[j]];
    const double w = li * q[j] + lambda[j] * qi;
    const double r2 = rx * rx + ry * ry + rz * rz;
    const double arg = gam * gam * r2;
    const double r1 = sqrt(r2);
    const double dtmp = 2.0 * gam * exp(-arg) / (EEQ_SQRTPI * r2) - erf(sqrt(arg)) / (r2 * r1);
    const double f = -w * dtmp;
    acc[0] += f * rx; acc[1] += f * ry; acc[2] += f * rz;
    const double h = 0.5 * f;
    acc[3] += h * rx * rx; acc[4] += h * rx * ry; acc[5] += h * rx * rz;
    acc[6] += h * ry * ry; acc[7] += h * ry * rz; acc[8] += h * rz * rz;
  }
  __shared__ double red[EEQG_THREADS / 32][9];
  for (int c = 0; c < 9; ++c) acc[c] = warp_sum(acc[c]);
  if ((tid & 31) == 0) for (int c = 0; c < 9; ++c) red[tid >> 5][c] = acc[c];
  __syncthreads();
  if (tid == 0) {
    double t[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int w = 0; w < EEQG_THREADS / 32; ++w) for (int c = 0; c < 9; ++c) t[c] += red[w][c];
    grad[3 * i] += t[0]; grad[3 * i + 1] += t[1]; grad[3 * i + 2] += t[2];
    double* sp = sig_part + (size_t)i * 6;
    for (int c = 0; c < 6; ++c) sp[c] += t[3 + c];
  }
}

// ---- periodic adjoint pass --------------------------------------------------------------------------------------
// Reciprocal space is separable: with the structure factors over the half set
//   Qc_k = sum_j q_j cos(G_k.r_j), Qs_k, Lc_k = sum_j lambda_j cos(G_k.r_j), Ls_k,  S0 = sum_j lambda_j q_j
// the pair sum  g_i -= sum_j (lambda_i q_j + lambda_j q_i) dA^rec_ij/dr_i  collapses to O(N nh):
//   g_i += 2 sum_k f_k G_k [ s_ik (lambda_i Qc_k + q_i Lc_k) - c_ik (lambda_i Qs_k + q_i Ls_k) ]
//   sigma -= 2 sum_k f_k M_k (Lc_k Qc_k + Ls_k Qs_k - S0),   M_k = (2/G^2 + 1/(2 alpha^2)) G G^T - 1
// sf[4*EEQ_NH + 1]: Qc | Qs | Lc | Ls | S0.  One block per k (block nh computes S0); cos/sin come from the phase matrix.
__global__ void __launch_bounds__(256) eeq_sfac_kernel(int nat, EwaldDev ew, const double* __restrict__ P,
                                                       const double* __restrict__ q, const double* __restrict__ lambda,
                                                       double* __restrict__ sf) {
  __shared__ double red[8][4];
  const int k = blockIdx.x, tid = threadIdx.x;
  double a[4] = {0, 0, 0, 0};
  if (k < ew.nh) {
    const double inv = 1.0 / sqrt(2.0 * ew.fh[k]);
    for (int j = tid; j < nat; j += 256) {
      const double c = P[(size_t)j * 2 * EEQ_NH + k] * inv, s = P[(size_t)j * 2 * EEQ_NH + EEQ_NH + k] * inv;
      a[0] += q[j] * c; a[1] += q[j] * s; a[2] += lambda[j] * c; a[3] += lambda[j] * s;
    }
  } else {
    for (int j = tid; j < nat; j += 256) a[0] += lambda[j] * q[j];
  }
  for (int c = 0; c < 4; ++c) a[c] = warp_sum(a[c]);
  if ((tid & 31) == 0) for (int c = 0; c < 4; ++c) red[tid >> 5][c] = a[c];
  __syncthreads();
  if (tid == 0) {
    double t[4] = {0, 0, 0, 0};
    for (int w = 0; w < 8; ++w) for (int c = 0; c < 4; ++c) t[c] += red[w][c];
    if (k < ew.nh) { for (int c = 0; c < 4; ++c) sf[c * EEQ_NH + k] = t[c]; }
    else sf[4 * EEQ_NH] = t[0];
  }
}

// One warp per own row: reciprocal gradient, self-image strain term; row 0 of rank 0 also adds the reciprocal virial.
__global__ void __launch_bounds__(256) eeq_grad_rec_kernel(int nrows, const int* __restrict__ rows, const int* __restrict__ species,
                                                           EwaldDev ew, const double* __restrict__ P, const double* __restrict__ q,
                                                           const double* __restrict__ lambda, const double* __restrict__ sf,
                                                           int add_sigma, double* __restrict__ grad, double* __restrict__ sig_part) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= nrows) return;
  const int i = rows[r];
  const double li = lambda[i], qi = q[i];
  double g[3] = {0, 0, 0}, sg[6] = {0, 0, 0, 0, 0, 0};
  const bool do_sigma = add_sigma && r == 0;
  for (int k = lane; k < ew.nh; k += 32) {
    const double f = ew.fh[k], inv = 1.0 / sqrt(2.0 * f);
    const double c = P[(size_t)i * 2 * EEQ_NH + k] * inv, s = P[(size_t)i * 2 * EEQ_NH + EEQ_NH + k] * inv;
    const double Qc = sf[k], Qs = sf[EEQ_NH + k], Lc = sf[2 * EEQ_NH + k], Ls = sf[3 * EEQ_NH + k];
    const double gx = ew.gh[3 * k], gy = ew.gh[3 * k + 1], gz = ew.gh[3 * k + 2];
    const double t = 2.0 * f * (s * (li * Qc + qi * Lc) - c * (li * Qs + qi * Ls));
    g[0] += t * gx; g[1] += t * gy; g[2] += t * gz;
    if (do_sigma) {
      const double g2 = gx * gx + gy * gy + gz * gz, f2 = 2.0 / g2 + 0.5 / (ew.alpha * ew.alpha);
      const double w = -2.0 * f * (Lc * Qc + Ls * Qs - sf[4 * EEQ_NH]);
      sg[0] += w * (f2 * gx * gx - 1.0); sg[1] += w * f2 * gx * gy; sg[2] += w * f2 * gx * gz;
      sg[3] += w * (f2 * gy * gy - 1.0); sg[4] += w * f2 * gy * gz; sg[5] += w * (f2 * gz * gz - 1.0);
    }
  }
  for (int c = 0; c < 3; ++c) g[c] = warp_sum(g[c]);
  for (int c = 0; c < 6; ++c) sg[c] = warp_sum(sg[c]);
  if (lane == 0) {
    atomicAdd(&grad[3 * i], g[0]); atomicAdd(&grad[3 * i + 1], g[1]); atomicAdd(&grad[3 * i + 2], g[2]);
    // self-image strain term: sigma -= lambda_i q_i dS_ii
    const double* d = ew.diag_ds + 9 * species[i];
    const double f = li * qi;
    double* sp = sig_part + (size_t)i * 6;
    atomicAdd(&sp[0], sg[0] - f * d[0]); atomicAdd(&sp[1], sg[1] - f * d[1]); atomicAdd(&sp[2], sg[2] - f * d[2]);
    atomicAdd(&sp[3], sg[3] - f * d[4]); atomicAdd(&sp[4], sg[4] - f * d[5]); atomicAdd(&sp[5], sg[5] - f * d[8]);
  }
}

// Real-space part: every unordered pair once (lower-triangle 16x16 tiles, tile p handled by rank p % world).
//   g_i -= w dg, g_j += w dg, sigma -= w ds,  w = lambda_i q_j + lambda_j q_i
__global__ void __launch_bounds__(EEQ_TILE * EEQ_TILE, 3) eeq_grad_dir_kernel(int nat, long long tile0, int tile_stride,
                                                                           const double* __restrict__ xyz, const int* __restrict__ species,
                                                                           Tables tb, EwaldDev ew, const double* __restrict__ q,
                                                                           const double* __restrict__ lambda, double* __restrict__ grad,
                                                                           double* __restrict__ sig_part) {
  extern __shared__ double sh[];
  double* sdir = sh;
  double* swsc = sdir + 3 * ew.ndir;
  double* stab = swsc + 3 * ew.nwsc + ((3 * ew.ndir + 3 * ew.nwsc) & 1);   // 16-byte aligned
  float* sdirf = reinterpret_cast<float*>(stab + (size_t)ew.tab_n * (EEQ_TAB_DEG + 1));
  float* swscf = sdirf + 4 * EEQ_MAXDIR;
  __shared__ double colg[EEQ_TILE][EEQ_TILE][3];
  __shared__ double sred[EEQ_TILE * EEQ_TILE / 32][6];
  const int tid = threadIdx.x, ty = tid / EEQ_TILE, tx = tid % EEQ_TILE;
  eeq_stage_tables(ew, sdir, swsc, sdirf, swscf, stab, ew.tabg, tid, EEQ_TILE * EEQ_TILE);
  __syncthreads();
  int ti, tj;
  unrank_tile(tile0 + (long long)blockIdx.x * tile_stride, ti, tj);
  const int i = ti * EEQ_TILE + ty, j = tj * EEQ_TILE + tx;
  double dg[3] = {0, 0, 0}, ds[6] = {0, 0, 0, 0, 0, 0};
  if (i < nat && j < i) {
    const double rx = xyz[3 * i] - xyz[3 * j], ry = xyz[3 * i + 1] - xyz[3 * j + 1], rz = xyz[3 * i + 2] - xyz[3 * j + 2];
    const double gam = tb.pair_gam[species[i] * tb.nid + species[j]];
    const double w = lambda[i] * q[j] + lambda[j] * q[i];
    const unsigned m = wsc_mask(rx, ry, rz, swsc, swscf, ew.nwsc);
    const double wsw = w / (double)__popc(m);
    double dummy = 0.0;
    unsigned nimg = 0u;
    for (unsigned mm = m; mm; mm &= mm - 1) {
      const int t = __ffs(mm) - 1;
      ewald_dir_sum<true>(rx - swsc[3 * t], ry - swsc[3 * t + 1], rz - swsc[3 * t + 2], gam, ew.alpha, sdir, sdirf, ew.ndir, stab,
                          ew.tab_n, ew.tab_base, wsw, dummy, dg, ds, nimg);
    }
  }
  // column sums (atom j): over ty through shared memory; row sums (atom i): over tx inside the 16-lane half warp
  for (int c = 0; c < 3; ++c) colg[ty][tx][c] = dg[c];
  for (int c = 0; c < 3; ++c) {
    double v = dg[c];
    for (int o = EEQ_TILE / 2; o > 0; o >>= 1) v += __shfl_xor_sync(D4_FULL, v, o);
    dg[c] = v;
  }
  for (int c = 0; c < 6; ++c) ds[c] = warp_sum(ds[c]);
  if ((tid & 31) == 0) for (int c = 0; c < 6; ++c) sred[tid >> 5][c] = ds[c];
  __syncthreads();
  if (tx == 0 && i < nat) { atomicAdd(&grad[3 * i], -dg[0]); atomicAdd(&grad[3 * i + 1], -dg[1]); atomicAdd(&grad[3 * i + 2], -dg[2]); }
  if (ty == 0 && j < nat) {
    double v[3] = {0, 0, 0};
    for (int y = 0; y < EEQ_TILE; ++y) for (int c = 0; c < 3; ++c) v[c] += colg[y][tx][c];
    atomicAdd(&grad[3 * j], v[0]); atomicAdd(&grad[3 * j + 1], v[1]); atomicAdd(&grad[3 * j + 2], v[2]);
  }
  if (tid < 6) {
    double v = 0.0;
    for (int w = 0; w < EEQ_TILE * EEQ_TILE / 32; ++w) v += sred[w][tid];
    const int i0 = min(ti * EEQ_TILE, nat - 1);
    atomicAdd(&sig_part[(size_t)i0 * 6 + tid], -v);
  }
}

}  // namespace d4k
