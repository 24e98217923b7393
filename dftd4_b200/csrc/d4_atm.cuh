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
//                             shared-memory ring that it fills with cp.async.bulk (TMA bulk copies completing on an
//                             mbarrier): one copy for the chunk's records, one per third atom for its precontracted
//                             C6 row (W0).  Work items are claimed from a global counter, heaviest rows first.
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
  double cr0;         // cbrt(R0(z_i, z_k))
  int atom;           // home atom of the entry
  unsigned meta;      // translation index (20 bit) | species (8 bit) << 20 | flags << 28
  double sw;          // smooth-cutoff factor of |v| (1 when the cutoff is sharp)
};
static_assert(sizeof(AtmRec) == 64, "AtmRec is streamed in 64-byte units");
constexpr unsigned ATM_META_NEAR = 1u << 28;  // r2 within 1e-10 relative of cutoff^2
constexpr unsigned ATM_META_SELF = 1u << 29;  // image of the centre itself
__device__ __forceinline__ int atm_meta_trans(unsigned m) { return (int)(m & 0xfffffu); }
__device__ __forceinline__ int atm_meta_species(unsigned m) { return (int)((m >> 20) & 0xffu); }

constexpr int ATM_WARPS = 8;        // warps per CTA of the streaming kernel (each one autonomous)
constexpr int ATM_STAGES = 2;       // ring depth of a warp's shared-memory stream
constexpr int ATM_KC_MAX = 16;      // third atoms per chunk (the launch picks kc <= ATM_KC_MAX from the smem budget)
#ifndef ATM_MINB
#define ATM_MINB 3                  // resident CTAs per SM the register budget is sized for
#endif

__host__ __device__ inline size_t atm_warp_smem_bytes(int tsp, int kc) {
  return (size_t)ATM_STAGES * kc * (sizeof(AtmRec) + sizeof(double) * tsp);
}

// W0[a][soff[s] + b] = sum_a' gw0_a[a'] R[z_a][s][a'][b]  (b < nref_s),  W0[a][soff[s] + nref_s] = cbrt(R0(z_a, s));
// W0cn the same with d gw0 / dCN (R0 slot = 0).  Row stride tsp >= ts + 7 (the 7-term contraction may read past a
// species block; the extra entries are multiplied by zero weights and must be finite).
__global__ void c6_rows_kernel(int nat, int ts, int tsp, const int* __restrict__ species, Tables tb,
                               const double* __restrict__ gw0, const double* __restrict__ gwdcn0,
                               double* __restrict__ W0, double* __restrict__ W0cn) {
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
        u = cbrt(tb.pair_r0[za * nid + s]);  // the kernel needs (R0ij R0ik R0jk)^(1/3)
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
                                                                int* __restrict__ nblk) {
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
    if (nblk) nblk[r] = (no + 31) >> 5;   // work items (32-edge blocks) of the row
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
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// The streaming kernel.  D4S = true (src/dftd4/model/d4s.f90): the weights of j in c6_jk depend on the species of k,
// so they are re-read per visit (L1-resident) instead of living in registers; everything else is shared with D4.
// ---------------------------------------------------------------------------------------------------------------
template <bool SMOOTH, bool ALP16, bool D4S>
__global__ void __launch_bounds__(ATM_WARPS * 32, ATM_MINB) atm_stream_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const long long* __restrict__ offsets, const int* __restrict__ nown, const int* __restrict__ blk_off,
    const AtmRec* __restrict__ rec, const int* __restrict__ atom3, const int* __restrict__ triples, Tables tb,
    const double* __restrict__ gw0, const double* __restrict__ W0, const __grid_constant__ AtmParams ap, int tsp, int kc,
    double2* __restrict__ eacc, unsigned long long* __restrict__ counters, int* __restrict__ work_counter, int want_grad) {
  extern __shared__ __align__(128) unsigned char atm_smem[];
  __shared__ __align__(8) unsigned long long s_bar[ATM_WARPS][ATM_STAGES];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t stage_bytes = (size_t)kc * (sizeof(AtmRec) + sizeof(double) * tsp);
  unsigned char* const wbase = atm_smem + (size_t)warp * ATM_STAGES * stage_bytes;
  const unsigned rec_bytes = (unsigned)(kc * sizeof(AtmRec));
  const unsigned row_bytes = (unsigned)(tsp * sizeof(double));
  const unsigned bar0 = smem_u32(&s_bar[warp][0]);
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < ATM_STAGES; ++s) mbar_init(bar0 + 8u * s, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  unsigned phase_bits = 0;  // bit s = parity the next wait on stage s expects

  const int total_items = blk_off[nrows];
  const double near_tol = ap.cutoff2 * 1.0e-10;
  // integer window on the high word of r2jk inside which the cutoff decision is redone with the reference's rounding
  const int hi_lo = __double2hiint(ap.cutoff2 * (1.0 - 1.0e-6)), hi_hi = __double2hiint(ap.cutoff2 * (1.0 + 1.0e-6));
  const int hi_eps = __double2hiint(ap.eps2) + 1;
  const double alp2 = 2.0 * ap.alp;
  unsigned long long visits = 0;

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

    // ---- this lane's edge ----
    const int pj = blk * 32 + lane;
    const bool valid = pj < no;
    double vjx = 0, vjy = 0, vjz = 0, r2ij = 1.0, a2x = 2.0, cr0ij = 1.0, swij = 1.0;
    double gjs[D4_MREF];
    double c6ij_e = 0.0;  // D4S only: c6ij kept apart from the weights
    int j = 0, tj = 0, so_j = 0, nrj = 0;
    unsigned vinc = 2u;   // ordered-triple count per visit: 2 for a regular edge, 1 for a half-weight self-image edge
    bool near_ij = false;
#pragma unroll
    for (int b = 0; b < D4_MREF; ++b) gjs[b] = 0.0;
    if (valid) {
      const AtmRec q = rrow[pj];
      vjx = q.vx; vjy = q.vy; vjz = q.vz; r2ij = q.r2; a2x = q.r2 + q.r2;
      j = q.atom; tj = atm_meta_trans(q.meta);
      const int zj = atm_meta_species(q.meta);
      so_j = tb.soff[zj]; nrj = tb.nref[zj];
      const size_t gj = gw_row(tb, j, spi);
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b)
        if (b < nrj) gjs[b] = gw0[gj + b] * q.c6;
      c6ij_e = q.c6;
      cr0ij = q.cr0;
      vinc = (q.meta & ATM_META_SELF) ? 1u : 2u;
      near_ij = (q.meta & ATM_META_NEAR) != 0;
      swij = q.sw;
    }
    double se = 0.0, sg = 0.0;
    unsigned nvis = 0;
    // warp sums of a visit's second-edge pair (e0, ggk): lanes 0-15 fold the first value, lanes 16-31 the second
    // (`mine` = the value this half keeps, `other` = the one it hands to the other half), then one RED per value
    const bool hi16 = (lane & 16) != 0;
    double* const eacc_row = reinterpret_cast<double*>(eacc + beg) + (hi16 ? 1 : 0);

    // ---- stream the row from the entry behind the block's first edge (earlier owned partners took this block's
    //      edges as their second edge already) ----
    const int s0 = blk * 32 + 1;
    const int nchunk = (n - s0 + kc - 1) / kc;
    auto issue = [&](int c, int atom_l) {
      const int st = c % ATM_STAGES;
      const int pos = s0 + c * kc, cnt = min(kc, n - pos);
      unsigned char* const sb = wbase + (size_t)st * stage_bytes;
      const unsigned bar = bar0 + 8u * st;
      if (lane == 0) mbar_expect_tx(bar, (unsigned)cnt * (unsigned)(sizeof(AtmRec)) + (unsigned)cnt * row_bytes);
      __syncwarp();
      if (lane == 0) bulk_g2s(smem_u32(sb), rrow + pos, (unsigned)cnt * (unsigned)sizeof(AtmRec), bar);
      if (lane < cnt) bulk_g2s(smem_u32(sb + rec_bytes + (size_t)lane * row_bytes), W0 + (size_t)atom_l * tsp, row_bytes, bar);
    };
    auto chunk_atom = [&](int c) {  // atom index of the entry this lane copies the C6 row of in chunk c
      const int pos = s0 + c * kc + lane;
      return (c < nchunk && lane < kc && pos < n) ? arow[pos] : 0;
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
      const int pos0 = s0 + c * kc, cnt = min(kc, n - pos0);
      const int k2end = max(0, min(cnt, no - pos0));

      // ---- two-edge visits: third atom owned by the centre as well ----
      for (int kk = 0; kk < k2end; ++kk) {
        const double2 ga = R2[4 * kk], gb = R2[4 * kk + 1];
        const double dx = ga.x - vjx, dy = ga.y - vjy, dz = gb.x - vjz;
        double r2jk = fma(dz, dz, fma(dy, dy, dx * dx));
        const int hw = __double2hiint(r2jk);
        bool ok = valid && (pos0 + kk > pj) && hw < hi_lo;
        if (hw >= hi_lo || hw < hi_eps) {   // (almost never) decide with the reference's rounding
          if (hw <= hi_hi) {
            const double ex = exact_r2(dx, dy, dz);
            ok = valid && (pos0 + kk > pj) && !(ex > ap.cutoff2 || ex < ap.eps2);
            r2jk = ex;
          }
        }
        const double2 gd = R2[4 * kk + 3];   // {atom | meta, sw}
        const unsigned kmeta = (unsigned)__double2hiint(gd.x);
        double swjk = 1.0;
        if (SMOOTH) {
          double dsw;
          smooth_cutoff(sqrt(r2jk), ap.cutoff, ap.width, swjk, dsw);
          ok = ok && (swij * gd.y * swjk > 0.0);
        }
        if (!__any_sync(D4_FULL, ok)) continue;
        const double2 gc = R2[4 * kk + 2];   // {c6ik, cr0ik}
        const double* Wk = KW + (size_t)kk * tsp + so_j;
        double c6p;
        if (D4S) {
          const double* gjk = gw0 + ((size_t)j * tb.nsp + atm_meta_species(kmeta)) * D4_MREF;  // weights of j against species(k)
          c6p = 0.0;
          for (int b = 0; b < nrj; ++b) c6p = fma(gjk[b], Wk[b], c6p);
          c6p *= c6ij_e;
        } else {
          c6p = gjs[0] * Wk[0];
#pragma unroll
          for (int b = 1; b < D4_MREF; ++b) c6p = fma(gjs[b], Wk[b], c6p);
        }
        AtmShared S;
        atm_common<ALP16>(r2ij, gb.y, r2jk, fabs(c6p * gc.x), cr0ij * gc.y * Wk[nrj], alp2, ap.alp, want_grad, S);
        double e0 = S.ang * S.cf, ggj = 0.0, ggk = 0.0;
        if (want_grad) {
          const double ta = fma(r2ij, fma(a2x, S.Z, -S.XY), S.m25P);
          const double b2x = gb.y + gb.y;
          const double tb_ = fma(gb.y, fma(b2x, S.X, -S.YZ), S.m25P);
          ggj = S.cf * fma(S.s2, ta, S.base);
          ggk = S.cf * fma(S.s2, tb_, S.base);
        }
        // rejected lanes contribute zeros (selects, no divergent region)
        double ka = ok ? e0 : 0.0, kb = ok ? ggk : 0.0, ja = ka, jb = ok ? ggj : 0.0;
        if (SMOOTH) {  // each edge's sums carry the cutoff factors of the OTHER two sides
          const double fj = gd.y * swjk, fk = swij * swjk;
          ja *= fj; jb *= fj; ka *= fk; kb *= fk;
        }
        se += ja; sg += jb;
        nvis += ok ? vinc + ((kmeta & ATM_META_SELF) ? 1u : 2u) : 0u;
        double keep = (hi16 ? kb : ka) + __shfl_xor_sync(D4_FULL, hi16 ? ka : kb, 16);
        keep += __shfl_xor_sync(D4_FULL, keep, 8);
        keep += __shfl_xor_sync(D4_FULL, keep, 4);
        keep += __shfl_xor_sync(D4_FULL, keep, 2);
        keep += __shfl_xor_sync(D4_FULL, keep, 1);
        if ((lane & 15) == 0) atomicAdd(eacc_row + 2 * (size_t)(pos0 + kk), keep);
      }

      // ---- one-edge visits: third atom above the centre (the reference evaluates the triangle in ITS frame) ----
      for (int kk = k2end; kk < cnt; ++kk) {
        const double2 ga = R2[4 * kk], gb = R2[4 * kk + 1];
        const double dx = ga.x - vjx, dy = ga.y - vjy, dz = gb.x - vjz;
        double r2jk = fma(dz, dz, fma(dy, dy, dx * dx));
        const int hw = __double2hiint(r2jk);
        const double2 gd = R2[4 * kk + 3];
        const unsigned kmeta = (unsigned)__double2hiint(gd.x);
        bool ok = valid && hw < hi_lo;
        if ((hw >= hi_lo && hw <= hi_hi) || hw < hi_eps || ((kmeta & ATM_META_NEAR) && hw <= hi_hi) || (near_ij && hw <= hi_hi)) {
          if (valid) {
            const double ex = exact_r2(dx, dy, dz);
            if ((kmeta & ATM_META_NEAR) || near_ij || fabs(ex - ap.cutoff2) <= near_tol)
              ok = atm_frame_check(i, j, tj, __double2loint(gd.x), atm_meta_trans(kmeta), xyz, triples, ap);
            else
              ok = !(ex > ap.cutoff2 || ex < ap.eps2);
            r2jk = ex;
          }
        }
        double swjk = 1.0;
        if (SMOOTH) {
          double dsw;
          smooth_cutoff(sqrt(r2jk), ap.cutoff, ap.width, swjk, dsw);
          ok = ok && (swij * gd.y * swjk > 0.0);
        }
        if (!__any_sync(D4_FULL, ok)) continue;
        const double2 gc = R2[4 * kk + 2];
        const double* Wk = KW + (size_t)kk * tsp + so_j;
        double c6p;
        if (D4S) {
          const double* gjk = gw0 + ((size_t)j * tb.nsp + atm_meta_species(kmeta)) * D4_MREF;
          c6p = 0.0;
          for (int b = 0; b < nrj; ++b) c6p = fma(gjk[b], Wk[b], c6p);
          c6p *= c6ij_e;
        } else {
          c6p = gjs[0] * Wk[0];
#pragma unroll
          for (int b = 1; b < D4_MREF; ++b) c6p = fma(gjs[b], Wk[b], c6p);
        }
        AtmShared S;
        atm_common<ALP16>(r2ij, gb.y, r2jk, fabs(c6p * gc.x), cr0ij * gc.y * Wk[nrj], alp2, ap.alp, want_grad, S);
        double e0 = S.ang * S.cf, gg = 0.0;
        if (want_grad) gg = S.cf * fma(S.s2, fma(r2ij, fma(a2x, S.Z, -S.XY), S.m25P), S.base);
        if (SMOOTH) { const double f = gd.y * swjk; e0 *= f; gg *= f; }
        if (ok) { se += e0; sg += gg; nvis += vinc; }
      }
      __syncwarp();   // the stage may be refilled from here on
    }

    // ---- this block's own sums join the per-entry accumulators (finalised by atm_finalize_kernel) ----
    if (valid) {
      double* acc = reinterpret_cast<double*>(eacc + beg + pj);
      atomicAdd(acc, se);
      atomicAdd(acc + 1, sg);
    }
    visits += nvis;
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
