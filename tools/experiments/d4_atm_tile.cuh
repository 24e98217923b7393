// Axilrod-Teller-Muto three-body stage as an N-body TILE kernel: one evaluation per triangle.
// (reference: src/dftd4/damping/atm.f90:235-453 with the q = 0 C6 of model/d4.f90:381-473, disp.f90:120-126)
//
// The streaming kernel of d4_atm.cuh walks the neighbour row of one centre atom and therefore sees a triangle twice
// (the sums of its third edge belong to another row).  Here the home atoms are sorted along a Z-curve and cut into
// TILES of 32; lattice images of a tile are tiles too.  A CTA takes a pair (I, J) -- I a home tile, J an image tile
// not above it in tile order -- and streams every tile K not above J that is within reach of both, in sub-tiles of 8
// third atoms:
//     lane  = atom j of J            warp = NI centres i of I          loop = third atoms k of K
// With that mapping all three edges of a triangle have a local accumulator:
//     edge (i, j)   registers of the lane (one pair of sums per centre of the warp), live for the whole work item
//     edge (j, k)   lane-private slot of the warp in shared memory; the warps' slots are folded per sub-tile
//     edge (i, k)   transposed butterfly over the lanes, then a warp-private slot in shared memory
// and everything pairwise a visit needs is a TABLE entry built once per sub-tile and shared by the 32 centres:
//     (j, k): r^2, C6, R0 term, flags      (i, k): the same     (i, j): registers
// so a visit is the damped ATM algebra of three table entries: no distance test, no C6 contraction, no second visit.
// The warp runs its NI centres in lock step (independent dependency chains, see atm_lockstep in d4_atm.cuh).
// Per-edge sums land in dense 32 x 32 blocks (one plane per sum), one per (home tile, image tile) pair, with fire-and-
// forget reductions in runs of consecutive doubles;
// atm_tile_finalize_kernel turns them into forces, virial, dE/dCN and energy exactly like atm_finalize_kernel.
//
// Which triangles: the reference sums over atom indices i >= j >= k and all translation pairs with triple_scale
// (atm.f90:324-357, 457-483).  The sum is over translation classes of triangles and triple_scale only depends on
// which atoms coincide, so any total order of the atoms gives the same sum; here the order is the position on the
// Z-curve ("rank").  Bit-exact acceptance: a pair within 1e-10 (relative) of the cutoff is flagged NEAR and every
// triangle that contains one is decided by atm_frame_check3, which redoes the three tests in the frame the reference
// uses (highest ATOM index at home, translations from its list, un-fused arithmetic).
#pragma once
#include "d4_atm.cuh"

namespace d4k {

constexpr int ATM_TK = 8;                 // third atoms per staged sub-tile
constexpr unsigned PF_OK = 1u;            // passes the cutoff test, or is NEAR (then the frame decides)
constexpr unsigned PF_NEAR = 2u;          // |r^2 - cutoff^2| <= 1e-10 cutoff^2
constexpr unsigned PF_SAME = 4u;          // both points are images of the same atom

struct TileSet {
  int nat, ntile, ntr;
  const double *tx, *ty, *tz;   // [ntile * 32] home positions in rank order (padding slots: 0)
  const int *tat, *tsp;         // atom index (-1: padding slot), species
  const double *bmin, *bmax;    // [ntile][3] bounding boxes of the tiles
  const double* trans;          // [ntr][3] disp3 translations, reference order (origin first)
  const int* triples;           // [ntr][3] their integer multiples
  const int* tmap;              // dense map (n1, n2, n3) -> translation index, |n_c| <= rep_c
  int rep[3];
  const int* poff;              // [ntile + 1] first pair of every home tile
  const int *pairB, *pairT;     // [npairs] partner tile and translation index, ascending in (B, T) per home tile
};

// ---- Z-curve keys of the home atoms ----------------------------------------------------------------------------
__device__ __forceinline__ unsigned spread10(unsigned v) {
  v &= 0x3ffu;
  v = (v | (v << 16)) & 0x030000ffu;
  v = (v | (v << 8)) & 0x0300f00fu;
  v = (v | (v << 4)) & 0x030c30c3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
__global__ void tile_keys_kernel(int nat, const double* __restrict__ xyz, double lox, double loy, double loz, double sx,
                                 double sy, double sz, unsigned* __restrict__ keys, unsigned* __restrict__ ids) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= nat) return;
  const unsigned cx = (unsigned)min(1023.0, fmax(0.0, (xyz[3 * a] - lox) * sx));
  const unsigned cy = (unsigned)min(1023.0, fmax(0.0, (xyz[3 * a + 1] - loy) * sy));
  const unsigned cz = (unsigned)min(1023.0, fmax(0.0, (xyz[3 * a + 2] - loz) * sz));
  keys[a] = spread10(cx) | (spread10(cy) << 1) | (spread10(cz) << 2);
  ids[a] = (unsigned)a;
}

// One warp per tile: positions / species / atom index in rank order and the tile's bounding box.
__global__ void tile_gather_kernel(int nat, int ntile, const unsigned* __restrict__ ids, const double* __restrict__ xyz,
                                   const int* __restrict__ species, double* __restrict__ tx, double* __restrict__ ty,
                                   double* __restrict__ tz, int* __restrict__ tat, int* __restrict__ tsp,
                                   double* __restrict__ bmin, double* __restrict__ bmax, const double* __restrict__ gw0,
                                   const double* __restrict__ W0, int wstride, double* __restrict__ tgw,
                                   double* __restrict__ tW0) {
  const int tile = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (tile >= ntile) return;
  const int r = tile * 32 + lane;
  const bool valid = r < nat;
  const int a = valid ? (int)ids[r] : -1;
  double x = 0.0, y = 0.0, z = 0.0;
  if (valid) { x = xyz[3 * a]; y = xyz[3 * a + 1]; z = xyz[3 * a + 2]; }
  tx[r] = x; ty[r] = y; tz[r] = z; tat[r] = a; tsp[r] = valid ? species[a] : 0;
  // weights and C6 rows in tile order (a tile's rows are then one contiguous block for the bulk copies)
  for (int l = 0; l < 32; ++l) {
    const int al = __shfl_sync(D4_FULL, a, l);
    const size_t rl = (size_t)tile * 32 + l;
    if (lane < D4_MREF) tgw[rl * D4_MREF + lane] = al >= 0 ? gw0[(size_t)al * D4_MREF + lane] : 0.0;
    for (int e = lane; e < wstride; e += 32) tW0[rl * wstride + e] = al >= 0 ? W0[(size_t)al * wstride + e] : 0.0;
  }
  double lo[3] = {valid ? x : 1e300, valid ? y : 1e300, valid ? z : 1e300};
  double hi[3] = {valid ? x : -1e300, valid ? y : -1e300, valid ? z : -1e300};
#pragma unroll
  for (int c = 0; c < 3; ++c)
    for (int o = 16; o > 0; o >>= 1) {
      lo[c] = fmin(lo[c], __shfl_xor_sync(D4_FULL, lo[c], o));
      hi[c] = fmax(hi[c], __shfl_xor_sync(D4_FULL, hi[c], o));
    }
  if (lane < 3) { bmin[3 * tile + lane] = lo[lane]; bmax[3 * tile + lane] = hi[lane]; }
}

// Squared distance between box A and box B shifted by t.
__device__ __forceinline__ double box_dist2(const double* __restrict__ bmin, const double* __restrict__ bmax, int A, int B,
                                            double t0, double t1, double t2) {
  const double t[3] = {t0, t1, t2};
  double d2 = 0.0;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const double d = fmax(0.0, fmax(bmin[3 * B + c] + t[c] - bmax[3 * A + c], bmin[3 * A + c] - (bmax[3 * B + c] + t[c])));
    d2 = fma(d, d, d2);
  }
  return d2;
}

// Pairs (A, B <= A, translation) whose boxes come within the cutoff.  One CTA per home tile; FILL = false counts,
// FILL = true writes them in ascending (B, t) order behind poff[A].
template <bool FILL>
__global__ void __launch_bounds__(256) tile_pair_kernel(int ntile, int ntr, const double* __restrict__ bmin,
                                                        const double* __restrict__ bmax, const double* __restrict__ trans,
                                                        double reach2, int* __restrict__ counts, const int* __restrict__ poff,
                                                        int* __restrict__ pairB, int* __restrict__ pairT) {
  __shared__ int s_w[8];
  __shared__ int s_base;
  const int A = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (A >= ntile) return;
  const long long ncand = (long long)(A + 1) * ntr;
  if (tid == 0) s_base = FILL ? poff[A] : 0;
  __syncthreads();
  for (long long c0 = 0; c0 < ncand; c0 += 256) {
    const long long c = c0 + tid;
    bool hit = false;
    int B = 0, t = 0;
    if (c < ncand) {
      B = (int)(c / ntr); t = (int)(c - (long long)B * ntr);
      hit = box_dist2(bmin, bmax, A, B, trans[3 * t], trans[3 * t + 1], trans[3 * t + 2]) <= reach2;
    }
    const unsigned m = __ballot_sync(D4_FULL, hit);
    if (lane == 0) s_w[warp] = __popc(m);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < 8; ++w) { if (w < warp) before += s_w[w]; total += s_w[w]; }
    if (FILL && hit) {
      const int pos = s_base + before + __popc(m & ((1u << lane) - 1u));
      pairB[pos] = B; pairT[pos] = t;
    }
    __syncthreads();
    if (tid == 0) s_base += total;
    __syncthreads();
  }
  if (!FILL && tid == 0) counts[A] = s_base;
}

// ---- exact acceptance of a triangle with a NEAR side -----------------------------------------------------------------
// Points: (a0 at home), (a1 + n1), (a2 + n2); n = integer multiples of the lattice vectors.  The reference visits the
// triangle with its highest atom index at home (ties: the first such point -- that pairs the two representatives of a
// class with a repeated atom one to one with the reference's), the other two at translations of ITS list, and tests
// vij, vik and vjk = vik - vij with un-fused arithmetic (atm.f90:332-357).
__device__ __noinline__ bool atm_frame_check3(int a0, int a1, int a2, int n10, int n11, int n12, int n20, int n21, int n22,
                                              const double* __restrict__ xyz, const AtmParams& ap) {
  const int at[3] = {a0, a1, a2};
  const int nn[3][3] = {{0, 0, 0}, {n10, n11, n12}, {n20, n21, n22}};
  int h = 0;
  if (at[1] > at[h]) h = 1;
  if (at[2] > at[h]) h = 2;
  const int o1 = h == 0 ? 1 : 0, o2 = h == 2 ? 1 : 2;
  int m1[3], m2[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    m1[c] = nn[o1][c] - nn[h][c];
    m2[c] = nn[o2][c] - nn[h][c];
    if (abs(m1[c]) > ap.rep[c] || abs(m2[c]) > ap.rep[c]) return false;   // not in the reference's translation list
  }
  double v1[3], v2[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const double xh = xyz[3 * at[h] + c];
    v1[c] = __dsub_rn(__dadd_rn(xyz[3 * at[o1] + c], atm_trans(ap.lat, c, m1[0], m1[1], m1[2])), xh);
    v2[c] = __dsub_rn(__dadd_rn(xyz[3 * at[o2] + c], atm_trans(ap.lat, c, m2[0], m2[1], m2[2])), xh);
  }
  const double ra = exact_r2(v1[0], v1[1], v1[2]), rb = exact_r2(v2[0], v2[1], v2[2]);
  const double rc = exact_r2(__dsub_rn(v2[0], v1[0]), __dsub_rn(v2[1], v1[1]), __dsub_rn(v2[2], v1[2]));
  return !(ra > ap.cutoff2 || ra < ap.eps2) && !(rb > ap.cutoff2 || rb < ap.eps2) && !(rc > ap.cutoff2 || rc < ap.eps2);
}

__device__ __forceinline__ unsigned pair_flags(double r2, double cut2, double eps2, double tol) {
  const bool near = fabs(r2 - cut2) <= tol;
  const bool ok = !(r2 > cut2 || r2 < eps2);
  return ((ok || near) ? PF_OK : 0u) | (near ? PF_NEAR : 0u);
}

// ---- the damped ATM term of NJ independent triangles in lock step, all three radial factors --------------------------
//   a = r2ij, b = r2ik, c = r2jk;  C = c6ij c6ik c6jk (times triple_scale^2, see the caller);  R: ALP16 ?
//   (R0ij R0ik R0jk)^(16/3) : (R0ij R0ik R0jk)^(1/3).  ok[m] = false yields exact zeros.
//   e0 = ang fdmp sqrt(C);  gg_e = sqrt(C) fdmp [ 3 u^3 - (3/4) u^5 t_e - 2 alp ang tmp fdmp ],
//   t_a = a (2 a Z - X Y) - (5/2) P,  t_b = b (2 b X - Y Z) - (5/2) P,  t_c = c (2 c Y - X Z) - (5/2) P
//   with X = a - b + c, Y = a + b - c, Z = -a + b + c, P = X Y Z (atm.f90:366-399, division-free restatement).
template <bool ALP16, bool GRAD, int NJ>
__device__ __forceinline__ void atm_lockstep3(const double (&a)[NJ], const double (&b)[NJ], const double (&c)[NJ],
                                              const double (&C)[NJ], const double (&R)[NJ], double alp2, double alp,
                                              const bool (&ok)[NJ], double (&e0)[NJ], double (&ga)[NJ], double (&gb)[NJ],
                                              double (&gc)[NJ]) {
  double r2[NJ], p[NJ], x2[NJ], x4[NJ], e[NJ], u3[NJ], u5[NJ], tmp[NJ], d[NJ], q[NJ], cf[NJ];
  double X[NJ], Y[NJ], Z[NJ], XY[NJ], YZ[NJ], XZ[NJ], P[NJ], ang[NJ];
  float lg[NJ];
#pragma unroll
  for (int m = 0; m < NJ; ++m) r2[m] = a[m] * b[m];
#pragma unroll
  for (int m = 0; m < NJ; ++m) r2[m] *= c[m];
#pragma unroll
  for (int m = 0; m < NJ; ++m) asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg[m]) : "f"((float)r2[m]));
#pragma unroll
  for (int m = 0; m < NJ; ++m) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(lg[m]) : "f"(lg[m] * (-1.0f / 6.0f)));
#pragma unroll
  for (int m = 0; m < NJ; ++m) { p[m] = (double)lg[m]; const double t = c[m] - b[m]; X[m] = a[m] + t; Y[m] = a[m] - t; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { x2[m] = p[m] * p[m]; Z[m] = (b[m] + c[m]) - a[m]; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { x4[m] = x2[m] * x2[m]; XY[m] = X[m] * Y[m]; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { x2[m] *= x4[m]; YZ[m] = Y[m] * Z[m]; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) { e[m] = fma(-r2[m], x2[m], 1.0); P[m] = X[m] * YZ[m]; }   // (1 - e)^(-1/6) = 1 + e/6 + 7 e^2/72
#pragma unroll
  for (int m = 0; m < NJ; ++m) { e[m] *= fma(7.0 / 72.0, e[m], 1.0 / 6.0); XZ[m] = X[m] * Z[m]; }
#pragma unroll
  for (int m = 0; m < NJ; ++m) p[m] = fma(p[m], e[m], p[m]);                              // r2^(-1/6), ~1e-16
  if (ALP16) {
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
    double base[NJ], ta[NJ], tb[NJ], tc[NJ];
#pragma unroll
    for (int m = 0; m < NJ; ++m) { base[m] = (d[m] * q[m]) * (cf[m] * tmp[m]); ta[m] = fma(a[m] + a[m], Z[m], -XY[m]); }   // tmp fdmp
#pragma unroll
    for (int m = 0; m < NJ; ++m) { base[m] *= ang[m]; P[m] *= -2.5; tb[m] = fma(b[m] + b[m], X[m], -YZ[m]); }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { base[m] = fma(3.0, u3[m], -(alp2 * base[m])); tc[m] = fma(c[m] + c[m], Y[m], -XZ[m]); u5[m] *= -0.75; }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { ta[m] = fma(a[m], ta[m], P[m]); tb[m] = fma(b[m], tb[m], P[m]); tc[m] = fma(c[m], tc[m], P[m]); }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { ta[m] = fma(u5[m], ta[m], base[m]); tb[m] = fma(u5[m], tb[m], base[m]); tc[m] = fma(u5[m], tc[m], base[m]); }
#pragma unroll
    for (int m = 0; m < NJ; ++m) { ga[m] = cf[m] * ta[m]; gb[m] = cf[m] * tb[m]; gc[m] = cf[m] * tc[m]; }
  } else {
#pragma unroll
    for (int m = 0; m < NJ; ++m) { ga[m] = 0.0; gb[m] = 0.0; gc[m] = 0.0; }
  }
}

// Shared-memory layout of the tile kernel (one CTA): fixed part, then the centre tile's C6 rows, then two stages of raw
// third-atom tiles (positions, atom / species ids, weights, C6 rows -- contiguous in tile order, so a stage is filled
// by five bulk copies completing on the stage's mbarrier while the previous tile is being worked on).
template <int WARPS, int TK>
struct AtmTileSmem {
  unsigned long long bar[2];
  double r2jk[TK][32], c6jk[TK][32], rjk[TK][32];   // (j, k) table: lane-indexed rows
  double r2ik[32][TK], c6ik[32][TK], rik[32][TK];   // (i, k) table
  double2 ikacc[32][TK];                            // sums of edge (i, k): rows owned by the centre's warp
  double2 jkacc[WARPS][TK][32];                     // sums of edge (j, k): one slab per warp, folded per sub-tile
  unsigned char fjk[TK][32], fik[32][TK];
  int item;
};
struct AtmTileStage {   // offsets (bytes) inside a stage
  static constexpr int X = 0, Y = 256, Z = 512, AT = 768, SP = 896, GW = 1024, W = 1024 + 32 * D4_MREF * 8;
  __host__ __device__ static size_t bytes(int tsp) { return (size_t)W + 32 * (size_t)tsp * 8; }
};
template <int WARPS, int TK>
__host__ __device__ inline size_t atm_tile_smem_bytes(int tsp) {
  return ((sizeof(AtmTileSmem<WARPS, TK>) + 127) & ~(size_t)127) + 32 * (size_t)tsp * 8 + 2 * AtmTileStage::bytes(tsp);
}

// ---------------------------------------------------------------------------------------------------------------
// The tile kernel.  NI centres per warp (lock step), 32 / NI warps per CTA, one CTA per SM (persistent, work items
// claimed from a counter, heaviest first).  tgw / tW0: weights and C6 rows permuted into tile order.
// ---------------------------------------------------------------------------------------------------------------
template <bool ALP16, bool GRAD, int NI, int TK>
__global__ void __launch_bounds__(1024 / NI, 1) atm_tile_kernel(
    const TileSet ts, const Tables tb, const double* __restrict__ tgw, const double* __restrict__ tW0, int tsp,
    const __grid_constant__ AtmParams ap, const double* __restrict__ xyz, double* __restrict__ accE, double* __restrict__ accG,
    int npairs, int rank, int world, int* __restrict__ work_counter, unsigned long long* __restrict__ counters,
    int* __restrict__ err, int dbg) {
  constexpr int WARPS = 32 / NI, NT = WARPS * 32;
  extern __shared__ __align__(128) unsigned char tile_smem_raw[];
  AtmTileSmem<WARPS, TK>& S = *reinterpret_cast<AtmTileSmem<WARPS, TK>*>(tile_smem_raw);
  double* const WA = reinterpret_cast<double*>(tile_smem_raw + ((sizeof(AtmTileSmem<WARPS, TK>) + 127) & ~(size_t)127));
  unsigned char* const stage0 = reinterpret_cast<unsigned char*>(WA + 32 * (size_t)tsp);
  const size_t stage_bytes = AtmTileStage::bytes(tsp);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double near_tol = ap.cutoff2 * 1.0e-10;
  const double reach2 = ap.cutoff2 * (1.0 + 1.0e-9) + 1.0e-9;
  const double alp2 = 2.0 * ap.alp;
  const int nitems = (npairs - rank + world - 1) / world;   // items rank, rank + world, ... of the reversed pair order
  unsigned long long visits = 0;
  const unsigned bar0 = smem_u32(&S.bar[0]);
  unsigned phase_bits = 0;   // bit s: parity the next wait on stage s expects (kept identical in every thread)

  // accumulators start at zero; the flush steps zero them again
  for (int e = tid; e < 32 * TK; e += NT) (&S.ikacc[0][0])[e] = make_double2(0.0, 0.0);
  for (int e = tid; e < WARPS * TK * 32; e += NT) (&S.jkacc[0][0][0])[e] = make_double2(0.0, 0.0);
  if (tid == 0) {
    mbar_init(bar0, 1u); mbar_init(bar0 + 8u, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // raw data of tile Ct into stage st (one elected thread; the tile's rows are contiguous in tile order)
  auto issue_stage = [&](int st, int Ct) {
    unsigned char* const sb = stage0 + (size_t)st * stage_bytes;
    const unsigned bar = bar0 + 8u * st;
    const size_t r0 = (size_t)Ct * 32;
    mbar_expect_tx(bar, (unsigned)stage_bytes);
    bulk_g2s(smem_u32(sb + AtmTileStage::X), ts.tx + r0, 256u, bar);
    bulk_g2s(smem_u32(sb + AtmTileStage::Y), ts.ty + r0, 256u, bar);
    bulk_g2s(smem_u32(sb + AtmTileStage::Z), ts.tz + r0, 256u, bar);
    bulk_g2s(smem_u32(sb + AtmTileStage::AT), ts.tat + r0, 128u, bar);
    bulk_g2s(smem_u32(sb + AtmTileStage::SP), ts.tsp + r0, 128u, bar);
    bulk_g2s(smem_u32(sb + AtmTileStage::GW), tgw + r0 * D4_MREF, 32u * D4_MREF * 8u, bar);
    bulk_g2s(smem_u32(sb + AtmTileStage::W), tW0 + r0 * tsp, 32u * (unsigned)tsp * 8u, bar);
  };

  while (true) {
    if (tid == 0) S.item = atomicAdd(work_counter, 1);
    __syncthreads();
    const int item = S.item;
    if (item >= nitems) break;
    const int p = npairs - 1 - (item * world + rank);   // heaviest first: high tiles own the longest streams
    int A;
    {
      int lo = 0, hi = ts.ntile - 1;   // largest A with poff[A] <= p
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (ts.poff[mid] <= p) lo = mid; else hi = mid - 1;
      }
      A = lo;
    }
    const int B = ts.pairB[p], T = ts.pairT[p];
    const double tJx = ts.trans[3 * T], tJy = ts.trans[3 * T + 1], tJz = ts.trans[3 * T + 2];
    const int nJ0 = ts.triples[3 * T], nJ1 = ts.triples[3 * T + 1], nJ2 = ts.triples[3 * T + 2];
    const int q0 = ts.poff[A], q1 = ts.poff[A + 1];

    // next tile K (not above J) within reach of J, searching from q
    auto next_k = [&](int q) {
      for (; q < q1; ++q) {
        const int Ct = ts.pairB[q];
        if (Ct > B) return q1;
        const int Tk = ts.pairT[q];
        if (box_dist2(ts.bmin, ts.bmax, B, Ct, ts.trans[3 * Tk] - tJx, ts.trans[3 * Tk + 1] - tJy, ts.trans[3 * Tk + 2] - tJz) <= reach2)
          return q;
      }
      return q1;
    };
    int q = next_k(q0);
    int stg = 0;
    if (q < q1 && tid == 0) issue_stage(0, ts.pairB[q]);

    // the centre tile's C6 rows (all 32: the (i, k) table is built by every thread)
    for (int e = tid; e < 32 * tsp; e += NT) WA[e] = tW0[(size_t)A * 32 * tsp + e];

    // ---- this lane's atom j of J (image position formed like the reference: fl(x_j + T)) ----
    const int rj = B * 32 + lane;
    const int aj = ts.tat[rj];
    const bool vj = aj >= 0;
    const double xj = __dadd_rn(ts.tx[rj], tJx), yj = __dadd_rn(ts.ty[rj], tJy), zj = __dadd_rn(ts.tz[rj], tJz);
    const int zsj = ts.tsp[rj];
    const int so_j = tb.soff[zsj], nrj = tb.nref[zsj];
    double gj[D4_MREF];
#pragma unroll
    for (int b = 0; b < D4_MREF; ++b) gj[b] = (vj && b < nrj) ? tgw[(size_t)rj * D4_MREF + b] : 0.0;

    // ---- the warp's centres i of I against j: registers for the whole work item ----
    double r2ij[NI], c6ij[NI], rij[NI], se_ij[NI], sg_ij[NI];
    unsigned fij[NI];
    int ai[NI];
#pragma unroll
    for (int m = 0; m < NI; ++m) {
      const int li = warp * NI + m, ri = A * 32 + li;
      ai[m] = ts.tat[ri];
      const double vx = __dsub_rn(xj, ts.tx[ri]), vy = __dsub_rn(yj, ts.ty[ri]), vz = __dsub_rn(zj, ts.tz[ri]);
      r2ij[m] = exact_r2(vx, vy, vz);
      unsigned f = pair_flags(r2ij[m], ap.cutoff2, ap.eps2, near_tol);
      // order on the Z-curve: j not above i (ties = images of the centre itself, the origin image excluded by eps)
      const bool order = B < A || lane <= li;
      if (!(vj && ai[m] >= 0 && order)) f = 0u;
      if (B == A && lane == li) f |= PF_SAME;
      fij[m] = f;
      c6ij[m] = 1.0; rij[m] = 1.0; se_ij[m] = 0.0; sg_ij[m] = 0.0;
      if (f & PF_OK) {
        const double* Wi = tW0 + (size_t)ri * tsp + so_j;
        double c6 = gj[0] * Wi[0];
#pragma unroll
        for (int b = 1; b < D4_MREF; ++b) c6 = fma(gj[b], Wi[b], c6);
        c6ij[m] = c6; rij[m] = Wi[nrj];
      } else {
        r2ij[m] = ap.cutoff2;   // harmless value for the lock-step algebra
      }
    }
    unsigned nvis = 0u;
    bool any_ij = false;
#pragma unroll
    for (int m = 0; m < NI; ++m) any_ij = any_ij || (fij[m] & PF_OK);
    const bool warp_has_ij = __any_sync(D4_FULL, any_ij);
    __syncthreads();   // WA complete

    // ---- stream the tiles K not above J ----
    while (q < q1) {
      const int Ct = ts.pairB[q], Tk = ts.pairT[q];
      const int qn = next_k(q + 1);
      if (qn < q1 && tid == 0) issue_stage(stg ^ 1, ts.pairB[qn]);   // that stage was drained before the last barrier
      const double tKx = ts.trans[3 * Tk], tKy = ts.trans[3 * Tk + 1], tKz = ts.trans[3 * Tk + 2];
      const int nK0 = ts.triples[3 * Tk], nK1 = ts.triples[3 * Tk + 1], nK2 = ts.triples[3 * Tk + 2];
      // block of edge (j, k): the higher tile (B) at home, the other one at the relative translation
      int pb = -1;
      {
        const int d0 = nK0 - nJ0, d1 = nK1 - nJ1, d2 = nK2 - nJ2;
        if (abs(d0) <= ts.rep[0] && abs(d1) <= ts.rep[1] && abs(d2) <= ts.rep[2]) {
          const int tr = ts.tmap[((d0 + ts.rep[0]) * (2 * ts.rep[1] + 1) + (d1 + ts.rep[1])) * (2 * ts.rep[2] + 1) + (d2 + ts.rep[2])];
          if (tr >= 0) {
            int lo = ts.poff[B], hi = ts.poff[B + 1] - 1;
            const long long key = (long long)Ct * ts.ntr + tr;
            while (lo <= hi) {
              const int mid = (lo + hi) >> 1;
              const long long km = (long long)ts.pairB[mid] * ts.ntr + ts.pairT[mid];
              if (km == key) { pb = mid; break; }
              if (km < key) lo = mid + 1; else hi = mid - 1;
            }
          }
        }
      }
      const bool same_jk_tile = Ct == B;
      mbar_wait(bar0 + 8u * stg, (phase_bits >> stg) & 1u);
      phase_bits ^= 1u << stg;
      const unsigned char* const sb = stage0 + (size_t)stg * stage_bytes;
      const double* const kx = reinterpret_cast<const double*>(sb + AtmTileStage::X);
      const double* const ky = reinterpret_cast<const double*>(sb + AtmTileStage::Y);
      const double* const kz = reinterpret_cast<const double*>(sb + AtmTileStage::Z);
      const int* const kat = reinterpret_cast<const int*>(sb + AtmTileStage::AT);
      const int* const ksp = reinterpret_cast<const int*>(sb + AtmTileStage::SP);
      const double* const kgw = reinterpret_cast<const double*>(sb + AtmTileStage::GW);
      const double* const kW = reinterpret_cast<const double*>(sb + AtmTileStage::W);

      for (int sub = 0; sub < 32 / TK; ++sub) {
        // ---- build the (j, k) and (i, k) tables of this sub-tile from the staged tile ----
        bool built_ok = false;
        for (int kk = warp; kk < TK; kk += WARPS) {   // third atom kk against every j
          const int lk = sub * TK + kk;
          const int ak = kat[lk];
          const double xk = __dadd_rn(kx[lk], tKx), yk = __dadd_rn(ky[lk], tKy), zk = __dadd_rn(kz[lk], tKz);
          const double r2 = exact_r2(__dsub_rn(xk, xj), __dsub_rn(yk, yj), __dsub_rn(zk, zj));
          unsigned f = pair_flags(r2, ap.cutoff2, ap.eps2, near_tol);
          const bool order = !same_jk_tile || lk <= lane;   // k not above j; the same point (same atom, same image) fails eps
          if (!(vj && ak >= 0 && order)) f = 0u;
          if (same_jk_tile && lk == lane) f |= PF_SAME;
          double c6 = 1.0, rr = 1.0;
          if (f & PF_OK) {
            const double* Wk = kW + (size_t)lk * tsp + so_j;
            c6 = gj[0] * Wk[0];
#pragma unroll
            for (int b = 1; b < D4_MREF; ++b) c6 = fma(gj[b], Wk[b], c6);
            rr = Wk[nrj];
            built_ok = true;
          }
          S.r2jk[kk][lane] = (f & PF_OK) ? r2 : ap.cutoff2;
          S.c6jk[kk][lane] = c6; S.rjk[kk][lane] = rr; S.fjk[kk][lane] = (unsigned char)f;
        }
        for (int e = tid; e < 32 * TK; e += NT) {   // entry (i, kk)
          const int li = e / TK, kk = e - li * TK, lk = sub * TK + kk;
          const int ri = A * 32 + li;
          const int ac = ts.tat[ri], ak = kat[lk];
          const double xk = __dadd_rn(kx[lk], tKx), yk = __dadd_rn(ky[lk], tKy), zk = __dadd_rn(kz[lk], tKz);
          const double r2 = exact_r2(__dsub_rn(xk, ts.tx[ri]), __dsub_rn(yk, ts.ty[ri]), __dsub_rn(zk, ts.tz[ri]));
          unsigned f = pair_flags(r2, ap.cutoff2, ap.eps2, near_tol);
          if (!(ac >= 0 && ak >= 0)) f = 0u;
          if (Ct == A && lk == li) f |= PF_SAME;
          double c6 = 1.0, rr = 1.0;
          if (f & PF_OK) {
            const int zk_ = ksp[lk];
            const int so_k = tb.soff[zk_], nrk = tb.nref[zk_];
            const double* Wi = WA + (size_t)li * tsp + so_k;
            const double* gk = kgw + (size_t)lk * D4_MREF;
            c6 = 0.0;
            for (int b = 0; b < nrk; ++b) c6 = fma(gk[b], Wi[b], c6);
            rr = Wi[nrk];
          }
          S.r2ik[li][kk] = (f & PF_OK) ? r2 : ap.cutoff2;
          S.c6ik[li][kk] = c6; S.rik[li][kk] = rr; S.fik[li][kk] = (unsigned char)f;
        }
        // tables ready; nothing to do when no (j, k) pair of the sub-tile is inside the cutoff
        if (!__syncthreads_or(built_ok ? 1 : 0)) continue;
        if (pb < 0 && tid == 0) atomicExch(err, 1);   // an accepted pair outside the translation list: caller falls back

        // ---- visits: NI centres in lock step per third atom ----
        if (warp_has_ij && !(dbg & 2)) {
          for (int kk = 0; kk < TK; ++kk) {
            const unsigned fk = S.fjk[kk][lane];
            unsigned fi[NI];
            bool ok[NI], any = false, near = false;
#pragma unroll
            for (int m = 0; m < NI; ++m) {
              fi[m] = S.fik[warp * NI + m][kk];
              ok[m] = (fij[m] & fk & fi[m] & PF_OK) != 0u;
              any = any || ok[m];
              near = near || (ok[m] && ((fij[m] | fk | fi[m]) & PF_NEAR));
            }
            if (!__any_sync(D4_FULL, any)) continue;
            if (__any_sync(D4_FULL, near)) {   // (rare) a side at the cutoff: the reference's frame decides
#pragma unroll
              for (int m = 0; m < NI; ++m)
                if (ok[m] && ((fij[m] | fk | fi[m]) & PF_NEAR))
                  ok[m] = atm_frame_check3(ai[m], aj, kat[sub * TK + kk], nJ0, nJ1, nJ2, nK0, nK1, nK2, xyz, ap);
            }
            const double c = S.r2jk[kk][lane], c6c = S.c6jk[kk][lane], rc = S.rjk[kk][lane];
            double b[NI], cc[NI], C[NI], R[NI], e0[NI], ga[NI], gb[NI], gc[NI];
#pragma unroll
            for (int m = 0; m < NI; ++m) {
              const int li = warp * NI + m;
              b[m] = S.r2ik[li][kk]; cc[m] = c;
              C[m] = fabs(c6ij[m] * S.c6ik[li][kk] * c6c);
              R[m] = rij[m] * S.rik[li][kk] * rc;
            }
            // triple_scale (atm.f90:457-483): 1/2 when two of the atoms coincide, 1/6 when all three do; every output is
            // linear in sqrt(C)
            unsigned w6[NI];   // 6 * triple_scale: the visit's share of the ordered-triple count
#pragma unroll
            for (int m = 0; m < NI; ++m) {
              const unsigned s = (fij[m] & PF_SAME) + (fk & PF_SAME) + (fi[m] & PF_SAME);
              w6[m] = 6u;
              if (s) { const bool all3 = s >= 2u * PF_SAME; C[m] *= all3 ? (1.0 / 36.0) : 0.25; w6[m] = all3 ? 1u : 3u; }
            }
            atm_lockstep3<ALP16, GRAD, NI>(r2ij, b, cc, C, R, alp2, ap.alp, ok, e0, ga, gb, gc);
            double ske = 0.0, skg = 0.0;
#pragma unroll
            for (int m = 0; m < NI; ++m) {
              se_ij[m] += e0[m]; sg_ij[m] += ga[m];
              ske += e0[m]; skg += gc[m];
              nvis += ok[m] ? w6[m] : 0u;
            }
            {   // edge (j, k): lane-private slot of this warp
              double2 t = S.jkacc[warp][kk][lane];
              t.x += ske; t.y += skg;
              S.jkacc[warp][kk][lane] = t;
            }
            // edge (i, k): 2 NI values summed over the lanes by a transposed butterfly (every round halves the values a
            // lane carries), then added to the warp-private slot
            if (NI == 4) {
              const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
              double v0 = h16 ? e0[2] : e0[0], v1 = h16 ? gb[2] : gb[0], v2 = h16 ? e0[3] : e0[1], v3 = h16 ? gb[3] : gb[1];
              v0 += __shfl_xor_sync(D4_FULL, h16 ? e0[0] : e0[2], 16);
              v1 += __shfl_xor_sync(D4_FULL, h16 ? gb[0] : gb[2], 16);
              v2 += __shfl_xor_sync(D4_FULL, h16 ? e0[1] : e0[3], 16);
              v3 += __shfl_xor_sync(D4_FULL, h16 ? gb[1] : gb[3], 16);
              double w0 = h8 ? v2 : v0, w1 = h8 ? v3 : v1;
              w0 += __shfl_xor_sync(D4_FULL, h8 ? v0 : v2, 8);
              w1 += __shfl_xor_sync(D4_FULL, h8 ? v1 : v3, 8);
              double z = h4 ? w1 : w0;
              z += __shfl_xor_sync(D4_FULL, h4 ? w0 : w1, 4);
              z += __shfl_xor_sync(D4_FULL, z, 2);
              z += __shfl_xor_sync(D4_FULL, z, 1);
              if ((lane & 3) == 0) {   // lane bits (16, 8, 4) = (centre pair, centre of the pair, which sum)
                const int m = ((lane >> 4) & 1) * 2 + ((lane >> 3) & 1);
                double* slot = reinterpret_cast<double*>(&S.ikacc[warp * NI + m][kk]) + ((lane >> 2) & 1);
                *slot += z;
              }
            } else {
              const bool h16 = lane & 16, h8 = lane & 8;
              double v0 = h16 ? e0[1] : e0[0], v1 = h16 ? gb[1] : gb[0];
              v0 += __shfl_xor_sync(D4_FULL, h16 ? e0[0] : e0[1], 16);
              v1 += __shfl_xor_sync(D4_FULL, h16 ? gb[0] : gb[1], 16);
              double z = h8 ? v1 : v0;
              z += __shfl_xor_sync(D4_FULL, h8 ? v0 : v1, 8);
              z += __shfl_xor_sync(D4_FULL, z, 4);
              z += __shfl_xor_sync(D4_FULL, z, 2);
              z += __shfl_xor_sync(D4_FULL, z, 1);
              if ((lane & 7) == 0) {
                const int m = (lane >> 4) & 1;
                double* slot = reinterpret_cast<double*>(&S.ikacc[warp * NI + m][kk]) + ((lane >> 3) & 1);
                *slot += z;
              }
            }
          }
        }
        __syncthreads();
        // ---- fold the sub-tile's (j, k) slabs and flush both shared accumulators (zeroed for the next sub-tile) ----
        // (thread <-> (row, kk) with kk fastest: a warp's reductions fall into runs of TK consecutive doubles)
        for (int e = tid; e < TK * 32; e += NT) {
          const int j = e / TK, kk = e - j * TK;
          double sx = 0.0, sy = 0.0;
#pragma unroll
          for (int w = 0; w < WARPS; ++w) {
            const double2 t = S.jkacc[w][kk][j];
            sx += t.x; sy += t.y;
            S.jkacc[w][kk][j] = make_double2(0.0, 0.0);
          }
          if ((sx != 0.0 || sy != 0.0) && pb >= 0 && !(dbg & 1)) {
            const size_t o = (size_t)pb * 1024 + (size_t)j * 32 + (sub * TK + kk);
            atomicAdd(accE + o, sx);
            if (GRAD) atomicAdd(accG + o, sy);
          }
        }
        for (int e = tid; e < 32 * TK; e += NT) {
          const int li = e / TK, kk = e - li * TK;
          const double2 t = S.ikacc[li][kk];
          if ((t.x != 0.0 || t.y != 0.0) && !(dbg & 1)) {
            S.ikacc[li][kk] = make_double2(0.0, 0.0);
            const size_t o = (size_t)q * 1024 + (size_t)li * 32 + (sub * TK + kk);
            atomicAdd(accE + o, t.x);
            if (GRAD) atomicAdd(accG + o, t.y);
          }
        }
        // (the next sub-tile's table build touches neither accumulator; its barrier orders this flush before the
        //  next visits)
      }
      __syncthreads();   // every thread is done with the stage: it may be refilled by the next iteration's copy
      q = qn;
      stg ^= 1;
    }

    visits += nvis;
    // ---- sums of the edges (i, j) ----
#pragma unroll
    for (int m = 0; m < NI; ++m) {
      if (se_ij[m] != 0.0 || sg_ij[m] != 0.0) {
        const size_t o = (size_t)p * 1024 + (size_t)(warp * NI + m) * 32 + lane;
        atomicAdd(accE + o, se_ij[m]);
        if (GRAD) atomicAdd(accG + o, sg_ij[m]);
      }
    }
    __syncthreads();
  }

  const unsigned long long v = warp_sum_u64(visits);
  if (lane == 0 && v) atomicAdd(&counters[4], v);
}

// ---------------------------------------------------------------------------------------------------------------
// Per (home tile, image tile) block: forces on both atoms of every edge, virial, dE/dCN and energy from the two sums
// (same algebra as atm_finalize_kernel; triple_scale is already inside the sums).  One CTA per block.
// ---------------------------------------------------------------------------------------------------------------
template <bool GRAD>
__global__ void __launch_bounds__(256) atm_tile_finalize_kernel(const TileSet ts, int npairs, const Tables tb,
                                                                const double* __restrict__ gw0, const double* __restrict__ gwdcn0,
                                                                const double* __restrict__ W0, const double* __restrict__ W0cn,
                                                                int tsp, double s9, const double* __restrict__ accE,
                                                                const double* __restrict__ accG,
                                                                double* __restrict__ e_atom, double* __restrict__ dEdcn,
                                                                double* __restrict__ grad, double* __restrict__ sig_part) {
  const int p = blockIdx.x;
  if (p >= npairs) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int A;
  {
    int lo = 0, hi = ts.ntile - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (ts.poff[mid] <= p) lo = mid; else hi = mid - 1;
    }
    A = lo;
  }
  const int B = ts.pairB[p], T = ts.pairT[p];
  const int rj = B * 32 + lane, j = ts.tat[rj];
  const double xj = __dadd_rn(ts.tx[rj], ts.trans[3 * T]), yj = __dadd_rn(ts.ty[rj], ts.trans[3 * T + 1]),
               zj = __dadd_rn(ts.tz[rj], ts.trans[3 * T + 2]);
  const int zsj = ts.tsp[rj];
  const int so_j = tb.soff[zsj], nrj = tb.nref[zsj];
  for (int li = warp; li < 32; li += 8) {
    const int ri = A * 32 + li, i = ts.tat[ri];
    if (i < 0) continue;   // (warp-uniform)
    const size_t o = (size_t)p * 1024 + (size_t)li * 32 + lane;
    const double2 a = make_double2(accE[o], GRAD ? accG[o] : 0.0);
    double out[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // e, dcn_i, g[3], s[6] of the centre
    if (j >= 0 && (a.x != 0.0 || a.y != 0.0)) {
      const double vx = __dsub_rn(xj, ts.tx[ri]), vy = __dsub_rn(yj, ts.ty[ri]), vz = __dsub_rn(zj, ts.tz[ri]);
      const double r2 = exact_r2(vx, vy, vz);
      const double se = a.x * -s9, sg = a.y * -s9;
      out[0] = -se * (1.0 / 3.0);
      if (GRAD) {
        const double* UI = W0 + (size_t)i * tsp;
        const double* UIcn = W0cn + (size_t)i * tsp;
        double c6ij = 0.0, dci = 0.0, dcj = 0.0;
        const size_t gj = (size_t)j * D4_MREF;
        for (int b = 0; b < nrj; ++b) {
          const double g = gw0[gj + b];
          c6ij += UI[so_j + b] * g;
          dci += UIcn[so_j + b] * g;
          dcj += UI[so_j + b] * gwdcn0[gj + b];
        }
        const double coef = sg / r2;
        const double gx = coef * vx, gy = coef * vy, gz = coef * vz;  // dGij summed over the third atoms
        out[2] = -gx; out[3] = -gy; out[4] = -gz;
        atomicAdd(&grad[3 * j], gx); atomicAdd(&grad[3 * j + 1], gy); atomicAdd(&grad[3 * j + 2], gz);
        out[5] = gx * vx; out[6] = gx * vy; out[7] = gx * vz;
        out[8] = gy * vy; out[9] = gy * vz; out[10] = gz * vz;
        out[1] = -0.5 * se * dci / c6ij;
        atomicAdd(&dEdcn[j], -0.5 * se * dcj / c6ij);
      }
    }
#pragma unroll
    for (int c = 0; c < 11; ++c) out[c] = warp_sum(out[c]);
    if (lane == 0) {
      if (out[0] != 0.0) atomicAdd(&e_atom[i], out[0]);
      if (GRAD) {
        atomicAdd(&dEdcn[i], out[1]);
        atomicAdd(&grad[3 * i], out[2]); atomicAdd(&grad[3 * i + 1], out[3]); atomicAdd(&grad[3 * i + 2], out[4]);
        double* sp = sig_part + (size_t)i * 6;
        for (int c = 0; c < 6; ++c) atomicAdd(&sp[c], out[5 + c]);
      }
    }
  }
}

}  // namespace d4k
