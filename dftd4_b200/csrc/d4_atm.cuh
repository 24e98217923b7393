// Axilrod-Teller-Muto three-body stage (reference: src/dftd4/damping/atm.f90:235-453, C6 from model/d4.f90:381-473
// with q = 0 weights, disp.f90:120-126).
//
// Formulation (DESIGN.md "ATM").  The reference loops i >= j >= k over atoms and all image pairs with triple_scale.
// Here an edge (i, j') of the disp3 list is OWNED by its higher-index atom i (the atom whose frame the reference
// evaluates the triangle in, atm.f90:332-357); self-image edges (j == i) are owned in both directions with weight 1/2,
// which reproduces triple_scale for two and three equal indices entry by entry.  Energy, gradient, virial and dE/dCN
// of a triangle decompose into per-edge terms, and per edge only two scalars have to be summed over the third atoms:
//     SE = sum_k' dE0                 (energy)          SG = sum_k' c9 (-dang_ij fdmp + ang dfdmp)   (radial force)
// A triangle a < b < c is visited twice: once at its highest atom c, where both centre edges (c,a), (c,b) are owned
// and ONE evaluation yields the sums of both ("two-edge visit"), and once at the middle atom b for edge (b,a) with the
// third atom above the centre ("one-edge visit").
//
// Kernels (one evaluation = four launches):
//   row_own_partition_kernel  rows of the disp3 list reordered [owned partners | partners above the centre]
//   atm_prep_kernel           one 64-byte record per list entry: vector from the centre, r^2, C6(centre, entry),
//                             cbrt(R0), atom / translation / flags -- everything a visit needs about the third atom
//                             that does not depend on the edge
//   atm_stream_kernel         the hot kernel.  A WARP is the unit of work: 32 consecutive owned edges of one centre
//                             (one per lane, in registers) against the stream of all later entries of the row.  Warps
//                             are autonomous -- no CTA barrier anywhere: every warp owns a private double-buffered
//                             shared-memory ring that it fills asynchronously, completion on an mbarrier: one
//                             cp.async.bulk (TMA bulk copy) for the chunk's contiguous records, 16-byte cp.async
//                             copies (two lanes per row) for the gathered precontracted C6 rows (W0).  Work items are claimed from a global counter, heaviest rows first.
//                             Per-edge sums go to a per-entry accumulator (eacc) with fire-and-forget reductions.
//   atm_finalize_kernel       per owned entry: forces on both edge atoms, virial, dE/dCN, energy from (SE, SG).
//
// C6_jk is contracted on the fly: c6_jk = gw0_j . W0_k[species(j)] with W0_k = R gw0_k precontracted per atom
// (c6_rows_kernel); the lane keeps c6_ij * gw0_j in registers.
//
// Bit-exact acceptance: the cutoff tests use the reference's un-fused association whenever a squared distance lies
// within 1e-6 relative of the cutoff (integer compare on the high word; everything else cannot be changed by the last
// ulp); when additionally the third atom has a higher index than the centre (the reference then works in the third
// atom's frame) the three tests are redone in that frame (atm_frame_check), so counts and energies agree with the
// reference even for cutoff-coincident lattice distances.
#pragma once
#include "d4_common.cuh"
#include "d4_kernels.cuh"

namespace d4k {

struct AtmParams {
  double s9, a1, a2, alp, cutoff, cutoff2, width, eps2;
  double lat[9];   // lattice[3*k+c] (zero when molecular)
  int rep[3];      // repetitions of the disp3 translation list
};

// Everything a visit needs about list entry k' of centre i (written by atm_prep_kernel, streamed by atm_stream_kernel).
struct __align__(16) AtmRec {
  double vx, vy, vz;  // fl(x_k' - x_i), formed like the reference (atm.f90:332)
  double r2;          // un-fused |v|^2
  double c6;          // C6(i, k') with q = 0 weights
  double cr0;         // cbrt(R0(z_i, z_k)), or R0^(16/3) when alp = 16 (the R0 slot of the centre's C6 row)
  int atom;           // home atom of the entry
  unsigned meta;      // translation index (20 bit) | species (8 bit) << 20 | flags << 28
  double sw;          // smooth-cutoff factor of |v| (1 when the cutoff is sharp)
};
static_assert(sizeof(AtmRec) == 64, "AtmRec is streamed in 64-byte units");
constexpr unsigned ATM_META_NEAR = 1u << 28;  // r2 within 1e-10 relative of cutoff^2
constexpr unsigned ATM_META_SELF = 1u << 29;  // image of the centre itself
__device__ __forceinline__ int atm_meta_trans(unsigned m) { return (int)(m & 0xfffffu); }
__device__ __forceinline__ int atm_meta_species(unsigned m) { return (int)((m >> 20) & 0xffu); }

constexpr int ATM_WARPS = 8;        // warps per CTA of the streaming kernel's default shape (each one autonomous)
constexpr int ATM_MINB = 2;         // ... and the resident CTAs per SM its register budget is sized for
constexpr int ATM_STAGES = 2;       // ring depth of a warp's shared-memory stream
#ifndef ATM_NJ
#define ATM_NJ 2                    // owned edges per lane of the streaming kernel
#endif
constexpr int ATM_KC_MAX = 16;      // third atoms per chunk (the launch picks kc <= ATM_KC_MAX from the smem budget)

__host__ __device__ inline size_t atm_warp_smem_bytes(int tsp, int kc) {
  return (size_t)ATM_STAGES * kc * (sizeof(AtmRec) + sizeof(double) * tsp);
}

// W0[a][soff[s] + b] = sum_a' gw0_a[a'] R[z_a][s][a'][b]  (b < nref_s),  W0[a][soff[s] + nref_s] = cbrt(R0(z_a, s))
// (r0_pow16: R0(z_a, s)^(16/3));
// W0cn the same with d gw0 / dCN (R0 slot = 0).  Row stride tsp >= ts + 7 (the 7-term contraction may read past a
// species block; the extra entries are multiplied by zero weights and must be finite).
__global__ void c6_rows_kernel(int nat, int ts, int tsp, const int* __restrict__ species, Tables tb,
                               const double* __restrict__ gw0, const double* __restrict__ gwdcn0,
                               double* __restrict__ W0, double* __restrict__ W0cn, int r0_pow16) {
  const int a = blockIdx.x;
  if (a >= nat) return;
  const int za = species[a], nid = tb.nid, nra = tb.nref[za];
  for (int e = threadIdx.x; e < tsp; e += blockDim.x) {
    double u = 0.0, ucn = 0.0;
    if (e < ts) {
      int s = 0;
      while (tb.soff[s + 1] <= e) ++s;
      const int b = e - tb.soff[s];
      if (b < tb.nref[s]) {
        const size_t ga = gw_row(tb, a, s);  // D4S: the atom's weights against partner species s
        for (int r = 0; r < nra; ++r) {
          const double rc6 = tb.refc6[(((size_t)za * nid + s) * D4_MREF + r) * D4_MREF + b];
          u += gw0[ga + r] * rc6;
          ucn += gwdcn0[ga + r] * rc6;
        }
      } else {
        u = cbrt(tb.pair_r0[za * nid + s]);  // the kernel needs (R0ij R0ik R0jk)^(1/3) ...
        if (r0_pow16) { u *= u; u *= u; u *= u; u *= u; }   // ... or, for alp = 16, its 16th power (atm_lockstep)
      }
    }
    W0[(size_t)a * tsp + e] = u;
    W0cn[(size_t)a * tsp + e] = ucn;
  }
}

// Translation vector of an integer triple, formed like cutoff.f90:144-145: (l1*n1 + l2*n2) + l3*n3 per component.
__device__ __forceinline__ double atm_trans(const double* lat, int c, int n1, int n2, int n3) {
  return __dadd_rn(__dadd_rn(__dmul_rn(lat[c], (double)n1), __dmul_rn(lat[3 + c], (double)n2)),
                   __dmul_rn(lat[6 + c], (double)n3));
}

// The three cutoff tests of triangle (centre i, j + T(tj), k + T(tk)) as the reference performs them when k is the
// highest index: iat = k at home, the other two atoms at translations -tk and tj - tk (atm.f90:332-357).
__device__ __noinline__ bool atm_frame_check(int i, int j, int tj, int k, int tk, const double* __restrict__ xyz,
                                             const int* __restrict__ triples, const AtmParams& ap) {
  const int a1 = -triples[3 * tk], a2 = -triples[3 * tk + 1], a3 = -triples[3 * tk + 2];
  const int b1 = triples[3 * tj] + a1, b2 = triples[3 * tj + 1] + a2, b3 = triples[3 * tj + 2] + a3;
  if (abs(b1) > ap.rep[0] || abs(b2) > ap.rep[1] || abs(b3) > ap.rep[2]) return false;  // not in the reference's list
  double v1[3], v2[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const double xk = xyz[3 * k + c];
    v1[c] = __dsub_rn(__dadd_rn(xyz[3 * i + c], atm_trans(ap.lat, c, a1, a2, a3)), xk);
    v2[c] = __dsub_rn(__dadd_rn(xyz[3 * j + c], atm_trans(ap.lat, c, b1, b2, b3)), xk);
  }
  const double ra = exact_r2(v1[0], v1[1], v1[2]), rb = exact_r2(v2[0], v2[1], v2[2]);
  const double rc = exact_r2(__dsub_rn(v2[0], v1[0]), __dsub_rn(v2[1], v1[1]), __dsub_rn(v2[2], v1[2]));
  return !(ra > ap.cutoff2 || ra < ap.eps2) && !(rb > ap.cutoff2 || rb < ap.eps2) && !(rc > ap.cutoff2 || rc < ap.eps2);
}

// x^(-1/2) and x^(-1/6) for positive normal x: hardware seed (2^-22) + one third-order correction -> ~1e-16 relative.
// No special-case handling (the ATM arguments are products of squared distances / C6 coefficients).
__device__ __forceinline__ double atm_rsqrt(double x) {
  double q;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(q) : "d"(x));
  const double e = fma(-x, q * q, 1.0);                 // (1 - e)^(-1/2) = 1 + e/2 + 3 e^2/8 + O(e^3), e ~ 5e-7
  return fma(q, e * fma(0.375, e, 0.5), q);
}
__device__ __forceinline__ double atm_inv_sixth_root(double x) {
  float lg, sd;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"((float)x));   // x is a product of squared distances: fp32 range
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(sd) : "f"(lg * (-1.0f / 6.0f)));
  double p = (double)sd;
  const double p2 = p * p, p4 = p2 * p2;
  const double e = fma(-x, p4 * p2, 1.0);               // (1 - e)^(-1/6) = 1 + e/6 + 7 e^2/72 + O(e^3), e ~ 3e-6
  return fma(p, e * fma(7.0 / 72.0, e, 1.0 / 6.0), p);
}

// Damped Axilrod-Teller-Muto term of one (centre; j', k') evaluation, division-free (atm.f90:360-399 restated).
//   a = r2ij, b = r2ik, c = r2jk;  C = c6ij c6ik c6jk;  c0 = (R0ij R0ik R0jk)^(1/3);  rbar^2 = a b c
//   X = a - b + c, Y = a + b - c, Z = -a + b + c,  P = X Y Z
//   ang  = (3/8) P / rbar^5 + 1 / rbar^3                                                            (atm.f90:366-368)
//   e0   = ang fdmp sqrt(C),   fdmp = 1 / (1 + 6 (R0 / rbar)^(alp/3))                               (atm.f90:362-364)
// Radial-force factor of edge ij (the reference's dGij = gg_ij vec_ij / r2ij summed with the c9 sign, atm.f90:376-383):
//   r d(ang)/dr_ij = -(3/4) u^5 t_a - ... collected:  gg_ij = sqrt(C) fdmp [ 3 u^3 - (3/4) u^5 t_a - 2 alp ang tmp fdmp ]
//   with t_a = a (2 a Z - X Y) - (5/2) P  (from dP/da = 2 a Z - X Y) and likewise t_b = b (2 b X - Y Z) - (5/2) P for
//   edge ik: both edges share every expensive intermediate.
struct AtmShared {
  double cf, ang, base, s2, m25P, X, Y, Z, XY, YZ;
};
template <bool ALP16>
__device__ __forceinline__ void atm_common(double a, double b, double c, double C, double c0, double alp2, double alp,
                                           int want_grad, AtmShared& S) {
  const double r2 = a * b * c;
  const double p = atm_inv_sixth_root(r2);    // r2^(-1/6)
  const double p2 = p * p, u = p2 * p, u2 = u * u, u3 = u2 * u, u5 = u3 * u2;   // u = 1/rbar
  const double v = c0 * p;                    // (R0 / rbar)^(1/3)
  double tmp;
  if (ALP16) {
    const double v2 = v * v, v4 = v2 * v2, v8 = v4 * v4;
    tmp = v8 * v8;
  } else {
    tmp = pow(v, alp);
  }
  const double d = fma(6.0, tmp, 1.0);        // 1 / fdmp
  const double q = atm_rsqrt(C * d * d);      // sqrt(C) fdmp = C q ;  fdmp = C d q^2
  S.cf = C * q;
  const double m = c - b, s = b + c;
  S.X = a + m; S.Y = a - m; S.Z = s - a;
  S.YZ = S.Y * S.Z; S.XY = S.X * S.Y;
  const double P = S.X * S.YZ;
  S.ang = fma(0.375 * u5, P, u3);
  if (want_grad) {
    const double tf = tmp * (S.cf * d * q);   // tmp * fdmp
    S.base = fma(3.0, u3, -(alp2 * (S.ang * tf)));
    S.s2 = -0.75 * u5;
    S.m25P = -2.5 * P;
  }
}

// The same algebra for NJ independent visits of one lane in LOCK STEP: every source line advances all NJ dependency
// chains by one operation, so that consecutive FP64 instructions of the warp are independent.  (Measured on B200,
// tools/ubench/fp64_latency.cu: a dependent DFMA issues every 8 cycles; the FP64 pipe of a scheduler takes the next
// instruction after 2.2 cycles when it comes from the SAME warp and after 3 cycles when it comes from another one --
// single-chain warps saturate at 72 % of the pipe no matter how many are resident, two chains reach 86 %, four 100 %.)
//   a[m] = r2ij, b = r2ik (shared), c[m] = r2jk, C[m] = c6ij c6ik c6jk, R[m]: ALP16 ? (R0ij R0ik R0jk)^(16/3)
//   : (R0ij R0ik R0jk)^(1/3).  TWO: also the radial factor of the streamed edge ik.  ok[m] = false yields exact zeros.
template <bool ALP16, bool GRAD, bool TWO, int NJ>
__device__ __forceinline__ void atm_lockstep(const double (&a)[NJ], double b, const double (&c)[NJ], const double (&C)[NJ],
                                             const double (&R)[NJ], double alp2, double alp, const bool (&ok)[NJ],
                                             double (&e0)[NJ], double (&ggj)[NJ], double (&ggk)[NJ]) {
  double r2[NJ], p[NJ], x2[NJ], x4[NJ], e[NJ], u3[NJ], u5[NJ], tmp[NJ], d[NJ], q[NJ], cf[NJ];
  double X[NJ], Y[NJ], Z[NJ], XY[NJ], YZ[NJ], P[NJ], ang[NJ];
  float lg[NJ];
#pragma unroll
  for (int m = 0; m < NJ; ++m) r2[m] = a[m] * b;
#pragma unroll
  for (int m = 0; m < NJ; ++m) r2[m] *= c[m];
#pragma unroll
  for (int m = 0; m < NJ; ++m) asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg[m]) : "f"((float)r2[m]));
#pragma unroll
  for (int m = 0; m < NJ; ++m) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(lg[m]) : "f"(lg[m] * (-1.0f / 6.0f)));
#pragma unroll
  for (int m = 0; m < NJ; ++m) { p[m] = (double)lg[m]; const double t = c[m] - b; X[m] = a[m] + t; Y[m] = a[m] - t; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { x2[m] = p[m] * p[m]; Z[m] = (b + c[m]) - a[m]; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { x4[m] = x2[m] * x2[m]; XY[m] = X[m] * Y[m]; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { x2[m] *= x4[m]; YZ[m] = Y[m] * Z[m]; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { e[m] = fma(-r2[m], x2[m], 1.0); P[m] = X[m] * YZ[m]; }   // (1 - e)^(-1/6) = 1 + e/6 + 7 e^2/72
#pragma unroll
  for (int m = 0; m < NJ; ++m) e[m] *= fma(7.0 / 72.0, e[m], 1.0 / 6.0);
#pragma unroll
  for (int m = 0; m < NJ; ++m) p[m] = fma(p[m], e[m], p[m]);                              // r2^(-1/6), ~1e-16
  if (ALP16) {
    // tmp = (R0 / rbar)^(16/3) = R p^16;  u3 = p^9, u5 = p^15 from the same squarings
#pragma unroll
    for (int m = 0; m < NJ; ++m) x2[m] = p[m] * p[m];
#pragma unroll
    for (int m = 0; m < NJ; ++m) x4[m] = x2[m] * x2[m];
#pragma unroll
    for (int m = 0; m < NJ; ++m) { e[m] = x4[m] * x4[m]; x2[m] *= x4[m]; }   // p^8, p^6
#pragma unroll
    for (int m = 0; m < NJ; ++m) { u3[m] = e[m] * p[m]; tmp[m] = e[m] * e[m]; }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { u5[m] = u3[m] * x2[m]; tmp[m] *= R[m]; }
  } else {
#pragma unroll
    for (int m = 0; m < NJ; ++m) {
      const double pp = p[m] * p[m], u = pp * p[m], uu = u * u;
      u3[m] = uu * u; u5[m] = u3[m] * uu;
      tmp[m] = pow(R[m] * p[m], alp);
    }
  }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { d[m] = fma(6.0, tmp[m], 1.0); ang[m] = fma(0.375 * u5[m], P[m], u3[m]); }
#pragma unroll
  for (int m = 0; m < NJ; ++m) r2[m] = C[m] * d[m];
#pragma unroll
  for (int m = 0; m < NJ; ++m) r2[m] *= d[m];
#pragma unroll
  for (int m = 0; m < NJ; ++m) asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(q[m]) : "d"(r2[m]));
#pragma unroll
  for (int m = 0; m < NJ; ++m) e[m] = q[m] * q[m];
#pragma unroll
  for (int m = 0; m < NJ; ++m) e[m] = fma(-r2[m], e[m], 1.0);                             // (1 - e)^(-1/2) = 1 + e/2 + 3 e^2/8
#pragma unroll
  for (int m = 0; m < NJ; ++m) e[m] *= fma(0.375, e[m], 0.5);
#pragma unroll
  for (int m = 0; m < NJ; ++m) q[m] = fma(q[m], e[m], q[m]);
#pragma unroll
  for (int m = 0; m < NJ; ++m) cf[m] = ok[m] ? C[m] * q[m] : 0.0;                         // sqrt(C) fdmp (0 for a rejected slot)
#pragma unroll
  for (int m = 0; m < NJ; ++m) e0[m] = ang[m] * cf[m];
  if (GRAD) {
    double base[NJ], ta[NJ], tb[NJ];
#pragma unroll
    for (int m = 0; m < NJ; ++m) { base[m] = (d[m] * q[m]) * (cf[m] * tmp[m]); ta[m] = fma(a[m] + a[m], Z[m], -XY[m]); }   // tmp fdmp
#pragma unroll
    for (int m = 0; m < NJ; ++m) { base[m] *= ang[m]; P[m] *= -2.5; if (TWO) tb[m] = fma(b + b, X[m], -YZ[m]); }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { base[m] = fma(3.0, u3[m], -(alp2 * base[m])); ta[m] = fma(a[m], ta[m], P[m]); u5[m] *= -0.75; }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { if (TWO) tb[m] = fma(b, tb[m], P[m]); ta[m] = fma(u5[m], ta[m], base[m]); }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { if (TWO) tb[m] = fma(u5[m], tb[m], base[m]); ggj[m] = cf[m] * ta[m]; }
#pragma unroll
    for (int m = 0; m < NJ; ++m) ggk[m] = TWO ? cf[m] * tb[m] : 0.0;
  } else {
#pragma unroll
    for (int m = 0; m < NJ; ++m) { ggj[m] = 0.0; ggk[m] = 0.0; }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Rows of the disp3 list reordered [owned partners | partners above the centre], keeping the list order inside both
// parts.  A pair (j', k') of two owned partners is then visited ONCE -- by the lane of the entry that comes first in
// the row.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) row_own_partition_kernel(int nrows, const int* __restrict__ rows,
                                                                const Point* __restrict__ pts,
                                                                const long long* __restrict__ offsets,
                                                                const unsigned* __restrict__ idx_in,
                                                                unsigned* __restrict__ idx_out, int* __restrict__ nown,
                                                                int* __restrict__ nblk, int edges_per_item) {
  __shared__ int s_o[8], s_u[8];
  const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (r >= nrows) return;
  const int i = rows[r];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  int no = 0;
  for (int c0 = 0; c0 < n; c0 += 256) {
    const int c = c0 + tid;
    const bool own = c < n && point_atom(pts[idx_in[beg + c]].w) <= i;
    const unsigned m = __ballot_sync(D4_FULL, own);
    if (lane == 0) s_o[warp] = __popc(m);
    __syncthreads();
    for (int w = 0; w < 8; ++w) no += s_o[w];
    __syncthreads();
  }
  int bo = 0, bu = no;
  for (int c0 = 0; c0 < n; c0 += 256) {
    const int c = c0 + tid;
    unsigned v = 0;
    bool own = false, oth = false;
    if (c < n) { v = idx_in[beg + c]; own = point_atom(pts[v].w) <= i; oth = !own; }
    const unsigned mo = __ballot_sync(D4_FULL, own), mu = __ballot_sync(D4_FULL, oth);
    if (lane == 0) { s_o[warp] = __popc(mo); s_u[warp] = __popc(mu); }
    __syncthreads();
    int po = bo, pu = bu, to = 0, tu = 0;
    for (int w = 0; w < 8; ++w) {
      if (w < warp) { po += s_o[w]; pu += s_u[w]; }
      to += s_o[w]; tu += s_u[w];
    }
    const unsigned below = (1u << lane) - 1u;
    if (own) idx_out[beg + po + __popc(mo & below)] = v;
    if (oth) idx_out[beg + pu + __popc(mu & below)] = v;
    bo += to; bu += tu;
    __syncthreads();
  }
  if (tid == 0) {
    nown[r] = no;
    if (nblk) nblk[r] = (no + edges_per_item - 1) / edges_per_item;   // work items (blocks of owned edges) of the row
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Per-entry records + the compact atom-index array the stream's copy engine walks.  One CTA per row.
// ---------------------------------------------------------------------------------------------------------------
template <bool SMOOTH>
__global__ void __launch_bounds__(256) atm_prep_kernel(int nrows, const int* __restrict__ rows, const double* __restrict__ xyz,
                                                       const int* __restrict__ species, const Point* __restrict__ pts,
                                                       const long long* __restrict__ offsets, const unsigned* __restrict__ idx,
                                                       Tables tb, const double* __restrict__ gw0, const double* __restrict__ W0,
                                                       int tsp, double cutoff, double cutoff2, double width,
                                                       AtmRec* __restrict__ rec, int* __restrict__ atom3) {
  const int r = blockIdx.x;
  if (r >= nrows) return;
  const int i = rows[r];
  const int spi = species[i];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  const double near_tol = cutoff2 * 1.0e-10;
  const double* UI = W0 + (size_t)i * tsp;
  for (int e = threadIdx.x; e < n; e += blockDim.x) {
    const Point p = load_point(pts + idx[beg + e]);
    AtmRec q;
    q.vx = __dsub_rn(p.x, xi); q.vy = __dsub_rn(p.y, yi); q.vz = __dsub_rn(p.z, zi);
    q.r2 = exact_r2(q.vx, q.vy, q.vz);
    const int k = point_atom(p.w), zk = point_species(p.w);
    const int so = tb.soff[zk], nrk = tb.nref[zk];
    double c6 = 0.0;
    const size_t gk = gw_row(tb, k, spi);
    for (int b = 0; b < nrk; ++b) c6 += UI[so + b] * gw0[gk + b];
    q.c6 = c6;
    q.cr0 = UI[so + nrk];
    q.atom = k;
    q.meta = (unsigned)point_trans(p.w) | ((unsigned)zk << 20) | (fabs(q.r2 - cutoff2) <= near_tol ? ATM_META_NEAR : 0u) |
             (k == i ? ATM_META_SELF : 0u);
    q.sw = 1.0;
    if (SMOOTH) { double dsw; smooth_cutoff(sqrt(q.r2), cutoff, width, q.sw, dsw); }
    rec[beg + e] = q;
    atom3[beg + e] = k;
  }
}

// ---- mbarrier / bulk-copy primitives (PTX; no tensor map needed for 1-D bulk copies) ----------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
// 16-byte asynchronous copy (LDGSTS) and the arrive that makes an mbarrier wait for the thread's copies issued so far;
// .noinc: the arrival is part of the barrier's initial count
__device__ __forceinline__ void cp_async_16(unsigned dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc(unsigned bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// The streaming kernel.  A lane carries NJ consecutive owned edges of the centre (a warp: 32 NJ edges) against the stream
// of later entries: the third atom's record and C6 row are read from shared memory ONCE per NJ visits, and so are the
// loop control, the warp reduction of the streamed edge's sums and its RED -- per visit roughly half of the
// instructions of the NJ = 1 form are not FP64, and with NJ = 2 the two visits of a lane are independent dependency
// chains.  D4S = true (src/dftd4/model/d4s.f90): the weights of j in c6_jk depend on the species of k, so they are
// re-read per visit (L1-resident) instead of living in registers; everything else is shared with D4.
// ---------------------------------------------------------------------------------------------------------------
template <bool SMOOTH, bool ALP16, bool D4S, bool GRAD, int NJ, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) atm_stream_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const long long* __restrict__ offsets, const int* __restrict__ nown, const int* __restrict__ blk_off,
    const AtmRec* __restrict__ rec, const int* __restrict__ atom3, const int* __restrict__ triples, Tables tb,
    const double* __restrict__ gw0, const double* __restrict__ W0, const __grid_constant__ AtmParams ap, int tsp, int kc,
    double2* __restrict__ eacc, unsigned long long* __restrict__ counters, int* __restrict__ work_counter) {
  extern __shared__ __align__(128) unsigned char atm_smem[];
  __shared__ __align__(8) unsigned long long s_bar[WARPS][ATM_STAGES];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t stage_bytes = (size_t)kc * (sizeof(AtmRec) + sizeof(double) * tsp);
  unsigned char* const wbase = atm_smem + (size_t)warp * ATM_STAGES * stage_bytes;
  const unsigned rec_bytes = (unsigned)(kc * sizeof(AtmRec));
  const unsigned row_bytes = (unsigned)(tsp * sizeof(double));
  const unsigned bar0 = smem_u32(&s_bar[warp][0]);
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < ATM_STAGES; ++s) mbar_init(bar0 + 8u * s, 33u);   // lane 0's expect_tx + one copy-arrive per lane
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  unsigned phase_bits = 0;  // bit s = parity the next wait on stage s expects

  const int total_items = blk_off[nrows];
  const double near_tol = ap.cutoff2 * 1.0e-10;
  // integer window on the high word of r2jk inside which the cutoff decision is redone with the reference's rounding
  const int hi_lo = __double2hiint(ap.cutoff2 * (1.0 - 1.0e-6)), hi_hi = __double2hiint(ap.cutoff2 * (1.0 + 1.0e-6));
  const int hi_eps = __double2hiint(ap.eps2) + 1;
  const unsigned fast_span = (unsigned)(hi_lo - hi_eps);   // hi_eps <= hw < hi_lo: accepted without a second look
  const double alp2 = 2.0 * ap.alp;
  unsigned long long visits = 0;
  constexpr int EPB = 32 * NJ;   // edges per work item

  while (true) {
    int item = 0;
    if (lane == 0) item = atomicAdd(work_counter, 1);
    item = __shfl_sync(D4_FULL, item, 0);
    if (item >= total_items) break;
    // heaviest first: rows are ascending in atom index (= number of owned edges); walk the items from the back, the
    // blocks of a row from its first (which streams the whole row) to its last
    const int t = total_items - 1 - item;
    int lo = 0, hi = nrows - 1;   // largest r with blk_off[r] <= t
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (blk_off[mid] <= t) lo = mid; else hi = mid - 1;
    }
    const int r = lo;
    const int blk = blk_off[r + 1] - 1 - t;
    const int i = rows[r];
    const int spi = species[i];
    const long long beg = offsets[r];
    const int n = (int)(offsets[r + 1] - beg);
    const int no = nown[r];
    const AtmRec* const rrow = rec + beg;
    const int* const arow = atom3 + beg;

    // ---- this lane's edges: row positions pj0 + m (neighbours in the row = neighbours in space) ----
    const int pj0 = blk * EPB + lane * NJ;
    double vjx[NJ], vjy[NJ], vjz[NJ], r2ij[NJ], rij0[NJ], swij[NJ], se[NJ], sg[NJ];
    double gjs[NJ][D4_MREF];
    double c6ij_e[NJ];   // D4S only: c6ij kept apart from the weights
    int jat[NJ];         // D4S only: the partner atom (its weights are re-read per visit)
    int so_j[NJ], so_r[NJ];   // offsets of the partner's species block / of its R0 slot in a C6 row
    int nrj[NJ];         // D4S only
    unsigned fspan[NJ];  // fast-accept span of the edge in one-edge visits (0: every visit takes the careful path)
    unsigned cnt[NJ];    // accepted visits of the current phase (two-edge, then one-edge)
    unsigned selfmask = 0u;   // bit m: half-weight self-image edge
    unsigned nks = 0u;        // accepted two-edge visits with a self-image third atom
#pragma unroll
    for (int m = 0; m < NJ; ++m) {
      const int pj = pj0 + m;
      // an empty slot sits far away (its distance to every third atom fails the fast test) with a positive C6 product
      vjx[m] = 1.0e30; vjy[m] = 0.0; vjz[m] = 0.0; r2ij[m] = 1.0; rij0[m] = 1.0; swij[m] = 1.0; se[m] = 0.0; sg[m] = 0.0;
      c6ij_e[m] = 0.0; jat[m] = 0; so_j[m] = 0; so_r[m] = 0; nrj[m] = 0; fspan[m] = fast_span; cnt[m] = 0u;
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b) gjs[m][b] = b == 0 ? 1.0 : 0.0;
      if (pj < no) {
        const AtmRec q = rrow[pj];
        vjx[m] = q.vx; vjy[m] = q.vy; vjz[m] = q.vz; r2ij[m] = q.r2;
        const int j = q.atom;
        const int zj = atm_meta_species(q.meta);
        const int nr = tb.nref[zj];
        so_j[m] = tb.soff[zj]; so_r[m] = so_j[m] + nr; nrj[m] = nr;
        const size_t gj = gw_row(tb, j, spi);
#pragma unroll
        for (int b = 0; b < D4_MREF; ++b) gjs[m][b] = b < nr ? gw0[gj + b] * q.c6 : 0.0;
        c6ij_e[m] = q.c6; jat[m] = j;
        rij0[m] = q.cr0;
        if (q.meta & ATM_META_SELF) selfmask |= 1u << m;
        if (q.meta & ATM_META_NEAR) fspan[m] = 0u;   // edge at the cutoff: the frame decides (one-edge visits)
        swij[m] = q.sw;
      }
    }
    bool two_phase = true;   // counters hold two-edge visits until the stream leaves the owned part of the row
    auto fold_two_edge_counts = [&]() {
#pragma unroll
      for (int m = 0; m < NJ; ++m) {
        visits += (unsigned long long)cnt[m] * (4u - ((selfmask >> m) & 1u));   // edge (2 or 1) + streamed edge (2)
        cnt[m] = 0u;
      }
      visits -= nks;   // a self-image streamed edge counts 1
      two_phase = false;
    };
    // warp sums of a visit's second-edge pair (e0, ggk): lanes 0-15 fold the first value, lanes 16-31 the second
    // (`mine` = the value this half keeps, `other` = the one it hands to the other half), then one RED per value
    const bool hi16 = (lane & 16) != 0;
    double* const eacc_row = reinterpret_cast<double*>(eacc + beg) + (hi16 ? 1 : 0);

    // ---- stream the row from the entry behind the block's first edge (earlier owned partners took this block's
    //      edges as their second edge already) ----
    const int s0 = blk * EPB + 1;
    const int nchunk = (n - s0 + kc - 1) / kc;
    // A stage = the chunk's records (contiguous in the row: ONE bulk copy, counted in bytes by the stage's mbarrier) + the
    // precontracted C6 row of every third atom (gathered by atom index: two lanes per row, 16-byte asynchronous copies,
    // each lane's copies counted by one arrive).
    const int ppr = (int)(row_bytes >> 4), pph = (ppr + 1) >> 1;   // 16-byte pieces per C6 row / per lane
    auto issue = [&](int c, int atom_l) {
      const int st = c % ATM_STAGES;
      const int pos = s0 + c * kc, cnt_ = min(kc, n - pos);
      unsigned char* const sb = wbase + (size_t)st * stage_bytes;
      const unsigned bar = bar0 + 8u * st;
      if (lane == 0) {
        mbar_expect_tx(bar, (unsigned)cnt_ * (unsigned)sizeof(AtmRec));
        bulk_g2s(smem_u32(sb), rrow + pos, (unsigned)cnt_ * (unsigned)sizeof(AtmRec), bar);
      }
      const int row = lane >> 1;
      if (row < cnt_) {
        const int p0 = (lane & 1) * pph, p1 = min(ppr, p0 + pph);
        const unsigned dst = smem_u32(sb + rec_bytes + (size_t)row * row_bytes) + 16u * p0;
        const char* src = reinterpret_cast<const char*>(W0 + (size_t)atom_l * tsp) + 16 * p0;
        for (int p = 0; p < p1 - p0; ++p) cp_async_16(dst + 16u * p, src + 16 * p);
      }
      cp_async_arrive_noinc(bar);
    };
    auto chunk_atom = [&](int c) {  // atom index of the entry whose C6 row this lane helps to copy in chunk c
      const int pos = s0 + c * kc + (lane >> 1);
      return (c < nchunk && (lane >> 1) < kc && pos < n) ? arow[pos] : 0;
    };
    int atom_next = chunk_atom(0);
    if (nchunk > 0) issue(0, atom_next);
    atom_next = chunk_atom(1);

    for (int c = 0; c < nchunk; ++c) {
      if (c + 1 < nchunk) issue(c + 1, atom_next);   // its stage was drained by chunk c - 1 (syncwarp below)
      atom_next = chunk_atom(c + 2);
      const int st = c % ATM_STAGES;
      mbar_wait(bar0 + 8u * st, (phase_bits >> st) & 1u);
      phase_bits ^= 1u << st;
      const unsigned char* const sb = wbase + (size_t)st * stage_bytes;
      const double2* const R2 = reinterpret_cast<const double2*>(sb);        // records as 4 x double2
      const double* const KW = reinterpret_cast<const double*>(sb + rec_bytes);
      const int pos0 = s0 + c * kc, cnt_c = min(kc, n - pos0);
      const int k2end = max(0, min(cnt_c, no - pos0));

      // ---- two-edge visits: third atom owned by the centre as well (the centre is the highest index: its frame) ----
      for (int kk = 0; kk < k2end; ++kk) {
        const double2 ga = R2[4 * kk], gb = R2[4 * kk + 1];
        const int kpos = pos0 + kk;
        double cs[NJ], swjk[NJ];
        bool ok[NJ], any = false;
#pragma unroll
        for (int m = 0; m < NJ; ++m) {
          const double dx = ga.x - vjx[m], dy = ga.y - vjy[m], dz = gb.x - vjz[m];
          cs[m] = fma(dz, dz, fma(dy, dy, dx * dx));
          const int hw = __double2hiint(cs[m]);
          ok[m] = (unsigned)(hw - hi_eps) < fast_span;
          if (!ok[m] && hw <= hi_hi) {   // (almost never) decide with the reference's rounding
            const double ex = exact_r2(dx, dy, dz);
            ok[m] = !(ex > ap.cutoff2 || ex < ap.eps2);
            cs[m] = ex;
          }
          ok[m] = ok[m] && kpos > pj0 + m;
          swjk[m] = 1.0;
          any = any || ok[m];
        }
        const double2 gd = R2[4 * kk + 3];   // {atom | meta, sw}
        const unsigned kmeta = (unsigned)__double2hiint(gd.x);
        if (SMOOTH) {
          any = false;
#pragma unroll
          for (int m = 0; m < NJ; ++m) {
            double dsw;
            smooth_cutoff(sqrt(cs[m]), ap.cutoff, ap.width, swjk[m], dsw);
            ok[m] = ok[m] && (swij[m] * gd.y * swjk[m] > 0.0);
            any = any || ok[m];
          }
        }
        if (!__any_sync(D4_FULL, any)) continue;
        const double2 gc = R2[4 * kk + 2];   // {c6ik, R0ik term}
        const double* const Wrow = KW + (size_t)kk * tsp;
        double C[NJ], R[NJ], e0[NJ], ggj[NJ], ggk[NJ];
#pragma unroll
        for (int m = 0; m < NJ; ++m) {
          const double* Wk = Wrow + so_j[m];
          double c6p;
          if (D4S) {
            const double* gjk = gw0 + ((size_t)jat[m] * tb.nsp + atm_meta_species(kmeta)) * D4_MREF;  // weights of j against species(k)
            c6p = 0.0;
            for (int b = 0; b < nrj[m]; ++b) c6p = fma(gjk[b], Wk[b], c6p);
            c6p *= c6ij_e[m];
          } else {
            c6p = gjs[m][0] * Wk[0];
#pragma unroll
            for (int b = 1; b < D4_MREF; ++b) c6p = fma(gjs[m][b], Wk[b], c6p);
          }
          // a rejected slot evaluates a harmless triangle (third side = the cutoff, positive C6 product: finite
          // everywhere) and contributes exact zeros
          C[m] = fabs(c6p * gc.x);
          if (D4S) C[m] = ok[m] ? C[m] : 1.0;   // (D4: an empty slot carries the weight vector (1, 0, ...))
          cs[m] = ok[m] ? cs[m] : ap.cutoff2;
          R[m] = rij0[m] * gc.y * Wrow[so_r[m]];
        }
        atm_lockstep<ALP16, GRAD, true, NJ>(r2ij, gb.y, cs, C, R, alp2, ap.alp, ok, e0, ggj, ggk);
        double ka = 0.0, kb = 0.0;
#pragma unroll
        for (int m = 0; m < NJ; ++m) {
          if (SMOOTH) {  // each edge's sums carry the cutoff factors of the OTHER two sides
            const double fj = gd.y * swjk[m], fk = swij[m] * swjk[m];
            se[m] = fma(e0[m], fj, se[m]); sg[m] = fma(ggj[m], fj, sg[m]);
            ka = fma(e0[m], fk, ka); kb = fma(ggk[m], fk, kb);
          } else {
            se[m] += e0[m]; sg[m] += ggj[m];
            ka += e0[m]; kb += ggk[m];
          }
          cnt[m] += ok[m] ? 1u : 0u;
        }
        if (kmeta & ATM_META_SELF) {   // (warp-uniform, rare)
#pragma unroll
          for (int m = 0; m < NJ; ++m) nks += ok[m] ? 1u : 0u;
        }
        double keep = (hi16 ? kb : ka) + __shfl_xor_sync(D4_FULL, hi16 ? ka : kb, 16);
        keep += __shfl_xor_sync(D4_FULL, keep, 8);
        keep += __shfl_xor_sync(D4_FULL, keep, 4);
        keep += __shfl_xor_sync(D4_FULL, keep, 2);
        keep += __shfl_xor_sync(D4_FULL, keep, 1);
        if ((lane & 15) == 0) atomicAdd(eacc_row + 2 * (size_t)kpos, keep);
      }
      if (two_phase && k2end < cnt_c) fold_two_edge_counts();

      // ---- one-edge visits: third atom above the centre (the reference evaluates the triangle in ITS frame) ----
      for (int kk = k2end; kk < cnt_c; ++kk) {
        const double2 ga = R2[4 * kk], gb = R2[4 * kk + 1];
        const double2 gd = R2[4 * kk + 3];
        const unsigned kmeta = (unsigned)__double2hiint(gd.x);
        const bool knear = (kmeta & ATM_META_NEAR) != 0;   // warp-uniform
        double cs[NJ], swjk[NJ];
        bool ok[NJ], any = false;
#pragma unroll
        for (int m = 0; m < NJ; ++m) {
          const double dx = ga.x - vjx[m], dy = ga.y - vjy[m], dz = gb.x - vjz[m];
          cs[m] = fma(dz, dz, fma(dy, dy, dx * dx));
          const int hw = __double2hiint(cs[m]);
          ok[m] = (unsigned)(hw - hi_eps) < fspan[m] && !knear;
          if (!ok[m] && hw <= hi_hi) {
            const double ex = exact_r2(dx, dy, dz);
            if (knear || fspan[m] == 0u || fabs(ex - ap.cutoff2) <= near_tol) {
              const AtmRec q = rrow[pj0 + m];
              ok[m] = atm_frame_check(i, q.atom, atm_meta_trans(q.meta), __double2loint(gd.x), atm_meta_trans(kmeta), xyz, triples, ap);
            } else {
              ok[m] = !(ex > ap.cutoff2 || ex < ap.eps2);
            }
            cs[m] = ex;
          }
          swjk[m] = 1.0;
          if (SMOOTH) {
            double dsw;
            smooth_cutoff(sqrt(cs[m]), ap.cutoff, ap.width, swjk[m], dsw);
            ok[m] = ok[m] && (swij[m] * gd.y * swjk[m] > 0.0);
          }
          any = any || ok[m];
        }
        if (!__any_sync(D4_FULL, any)) continue;
        const double2 gc = R2[4 * kk + 2];
        const double* const Wrow = KW + (size_t)kk * tsp;
        double C[NJ], R[NJ], e0[NJ], gg[NJ], unused[NJ];
#pragma unroll
        for (int m = 0; m < NJ; ++m) {
          const double* Wk = Wrow + so_j[m];
          double c6p;
          if (D4S) {
            const double* gjk = gw0 + ((size_t)jat[m] * tb.nsp + atm_meta_species(kmeta)) * D4_MREF;
            c6p = 0.0;
            for (int b = 0; b < nrj[m]; ++b) c6p = fma(gjk[b], Wk[b], c6p);
            c6p *= c6ij_e[m];
          } else {
            c6p = gjs[m][0] * Wk[0];
#pragma unroll
            for (int b = 1; b < D4_MREF; ++b) c6p = fma(gjs[m][b], Wk[b], c6p);
          }
          C[m] = fabs(c6p * gc.x);
          if (D4S) C[m] = ok[m] ? C[m] : 1.0;
          cs[m] = ok[m] ? cs[m] : ap.cutoff2;
          R[m] = rij0[m] * gc.y * Wrow[so_r[m]];
        }
        atm_lockstep<ALP16, GRAD, false, NJ>(r2ij, gb.y, cs, C, R, alp2, ap.alp, ok, e0, gg, unused);
#pragma unroll
        for (int m = 0; m < NJ; ++m) {
          if (SMOOTH) { const double f = gd.y * swjk[m]; se[m] = fma(e0[m], f, se[m]); sg[m] = fma(gg[m], f, sg[m]); }
          else { se[m] += e0[m]; sg[m] += gg[m]; }
          cnt[m] += ok[m] ? 1u : 0u;
        }
      }
      __syncwarp();   // the stage may be refilled from here on
    }
    if (two_phase) fold_two_edge_counts();   // the stream never left the owned part

    // ---- this block's own sums join the per-entry accumulators (finalised by atm_finalize_kernel) ----
#pragma unroll
    for (int m = 0; m < NJ; ++m) {
      if (pj0 + m < no) {
        double* acc = reinterpret_cast<double*>(eacc + beg + pj0 + m);
        atomicAdd(acc, se[m]);
        atomicAdd(acc + 1, sg[m]);
        visits += (unsigned long long)cnt[m] * (2u - ((selfmask >> m) & 1u));   // one-edge visits: the lane's edge only
      }
    }
  }

  const unsigned long long v = warp_sum_u64(visits);
  if (lane == 0 && v) atomicAdd(&counters[4], v);
}

// ---------------------------------------------------------------------------------------------------------------
// Per owned entry: forces on both edge atoms, virial, dE/dCN and energy from the two sums.  One CTA per row.
// ---------------------------------------------------------------------------------------------------------------
template <bool SMOOTH>
__global__ void __launch_bounds__(256) atm_finalize_kernel(
    int nrows, const int* __restrict__ rows, const int* __restrict__ species, const long long* __restrict__ offsets,
    const int* __restrict__ nown, const AtmRec* __restrict__ rec, Tables tb, const double* __restrict__ gw0,
    const double* __restrict__ gwdcn0, const double* __restrict__ W0, const double* __restrict__ W0cn, int tsp, double s9,
    double cutoff, double width, const double2* __restrict__ eacc, double* __restrict__ e_atom, double* __restrict__ dEdcn,
    double* __restrict__ grad, double* __restrict__ sig_part, int want_grad, double* __restrict__ e3, int e3_ld) {
  __shared__ double s_acc[8][11];
  const int r = blockIdx.x;
  if (r >= nrows) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int i = rows[r];
  const int spi = species[i];
  const long long beg = offsets[r];
  const int no = nown[r];
  const double* UI = W0 + (size_t)i * tsp;
  const double* UIcn = W0cn + (size_t)i * tsp;
  double out[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // e, dcn_i, g[3], s[6] of the centre
  for (int pj = tid; pj < no; pj += 256) {
    const AtmRec q = rec[beg + pj];
    const double2 acc = eacc[beg + pj];
    const int j = q.atom;
    const int zj = atm_meta_species(q.meta);
    const int so_j = tb.soff[zj], nrj = tb.nref[zj];
    double swij = 1.0, dswij = 0.0;
    const double rij = sqrt(q.r2);
    if (SMOOTH) smooth_cutoff(rij, cutoff, width, swij, dswij);
    const double wedge = ((q.meta & ATM_META_SELF) ? 0.5 : 1.0) * -s9;
    const double se = acc.x * wedge, sg = acc.y * wedge;
    const double eedge = swij * se;  // sum over k of dE0 * sw, weighted
    out[0] += -eedge * (1.0 / 3.0);
    if (e3) {  // pair-resolved energies (rational.f90:487-606): a sixth of the triangle energy per directed pair
      atomicAdd(&e3[(size_t)i * e3_ld + j], -eedge * (1.0 / 6.0));
      atomicAdd(&e3[(size_t)j * e3_ld + i], -eedge * (1.0 / 6.0));
    }
    if (want_grad) {
      double c6ij = 0.0, dci = 0.0, dcj = 0.0;
      const size_t gj = gw_row(tb, j, spi);
      for (int b = 0; b < nrj; ++b) {
        const double g = gw0[gj + b];
        c6ij += UI[so_j + b] * g;
        dci += UIcn[so_j + b] * g;
        dcj += UI[so_j + b] * gwdcn0[gj + b];
      }
      double coef = swij * sg / q.r2;
      if (SMOOTH) coef -= se * dswij / rij;
      const double gx = coef * q.vx, gy = coef * q.vy, gz = coef * q.vz;  // dGij summed over k
      out[2] -= gx; out[3] -= gy; out[4] -= gz;
      atomicAdd(&grad[3 * j], gx); atomicAdd(&grad[3 * j + 1], gy); atomicAdd(&grad[3 * j + 2], gz);
      out[5] += gx * q.vx; out[6] += gx * q.vy; out[7] += gx * q.vz;
      out[8] += gy * q.vy; out[9] += gy * q.vz; out[10] += gz * q.vz;
      out[1] += -0.5 * eedge * dci / c6ij;
      atomicAdd(&dEdcn[j], -0.5 * eedge * dcj / c6ij);
    }
  }
#pragma unroll
  for (int c = 0; c < 11; ++c) {
    const double t = warp_sum(out[c]);
    if (lane == 0) s_acc[warp][c] = t;
  }
  __syncthreads();
  if (tid == 0) {
    double t[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int w = 0; w < 8; ++w)
      for (int c = 0; c < 11; ++c) t[c] += s_acc[w][c];
    atomicAdd(&e_atom[i], t[0]);
    if (want_grad) {
      atomicAdd(&dEdcn[i], t[1]);
      atomicAdd(&grad[3 * i], t[2]); atomicAdd(&grad[3 * i + 1], t[3]); atomicAdd(&grad[3 * i + 2], t[4]);
      double* sp = sig_part + (size_t)i * 6;
      for (int c = 0; c < 6; ++c) sp[c] += t[5 + c];
    }
  }
}

}  // namespace d4k
