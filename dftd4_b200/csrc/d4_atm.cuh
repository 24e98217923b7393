// Axilrod-Teller-Muto three-body kernel (reference: src/dftd4/damping/atm.f90:235-453, C6 from model/d4.f90:381-473
// with q = 0 weights, disp.f90:120-126).
//
// Formulation (DESIGN.md "ATM").  The reference loops i >= j >= k over atoms and all image pairs with triple_scale.
// Here every accepted triangle is visited once per EDGE (atm_kernel, below) or -- the default, atm_two_edge_kernel at
// the end of this file -- once for the two edges of its highest-index atom plus once for the third edge.  An edge (i, j') is owned by its higher-index atom i (the
// atom whose frame the reference evaluates the triangle in, atm.f90:332-357); self-image edges (j == i) are owned in
// both directions with weight 1/2, which reproduces triple_scale for two and three equal indices entry by entry.
// One CTA per centre atom i; a pass holds one owned edge per thread in registers and streams ALL neighbours k' of i
// through shared memory.  A visit only needs two scalar accumulators per edge,
//     SE = sum_k' dE0                 (energy)          SG = sum_k' c9 (-dang_ij fdmp + ang dfdmp)   (radial force)
// because energy, gradient, virial and dE/dCN of a triangle decompose into per-edge terms; forces on the two edge
// atoms, the virial and dE/dCN are applied once per edge after the stream, so the inner loop carries no atomics.
// C6_jk is contracted on the fly: c6_jk = gw0_j . W0_k[species(j)] with W0_k = R gw0_k precontracted per atom
// (c6_rows_kernel) and staged per chunk in shared memory; the lane keeps gw0_j in registers.
//
// Bit-exact acceptance: the three cutoff tests are evaluated in the centre's frame with the reference's association;
// when the third atom has a higher index than the centre (the reference then works in the third atom's frame) and one
// of the three squared distances lies within 1e-10 relative of the cutoff, the tests are redone in that frame
// (atm_frame_check), so counts and energies agree with the reference even for cutoff-coincident lattice distances.
#pragma once
#include "d4_common.cuh"
#include "d4_kernels.cuh"

namespace d4k {

#ifndef ATM_THREADS_N
#define ATM_THREADS_N 256
#endif
#ifndef ATM_KC_N
#define ATM_KC_N 128
#endif
constexpr int ATM_THREADS = ATM_THREADS_N;  // = edges per pass
constexpr int ATM_WARPS = ATM_THREADS / 32;
constexpr int ATM_KC = ATM_KC_N;            // neighbours per staged chunk
constexpr int ATM_GEO = 7;        // {kx,ky} {kz,r2ik} {c6ik,cbrt R0ik} ksw
#ifndef ATM_UNROLL2
#define ATM_UNROLL2 0             // 1: two third atoms per inner iteration in the hot configuration
#endif
#ifndef ATM_PENDING
#define ATM_PENDING 0             // 1: fold a visit's second-edge sums at the start of the next executed visit
#endif
#ifndef ATM_CYCLIC
#define ATM_CYCLIC 0              // 0: triangular pairing inside a pass's own block of owned entries (two-edge kernel)
#endif
#ifndef ATM_MINB
#define ATM_MINB 3                // resident CTAs per SM the register budget is sized for
#endif

struct AtmParams {
  double s9, a1, a2, alp, cutoff, cutoff2, width, eps2;
  double lat[9];   // lattice[3*k+c] (zero when molecular)
  int rep[3];      // repetitions of the disp3 translation list
};

__host__ __device__ inline size_t atm_smem_bytes(int tsp) {
  // UI, UIcn [tsp] | geometry [ATM_GEO][KC] | W rows [KC][tsp] | ebuf [2*T] u32 | katom, ktr, kflag [KC] int
  return sizeof(double) * ((size_t)2 * tsp + (size_t)ATM_GEO * ATM_KC + (size_t)ATM_KC * tsp) +  // tsp even: 16-B aligned
         sizeof(unsigned) * 2 * ATM_THREADS + sizeof(int) * 3 * ATM_KC;
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

// Damped Axilrod-Teller-Muto term of one (edge, third atom) visit: e0 = ang * fdmp * sqrt(C) and, with want_grad, the
// radial-force factor of edge ij (atm.f90:360-399 restated division-free; C = c6ij c6ik c6jk, c0 = (R0ij R0ik R0jk)^(1/3)).
template <bool ALP16>
__device__ __forceinline__ void atm_core(double r2ij, double r2ik, double r2jk, double C, double c0, double alp,
                                         int want_grad, double& e0, double& gg) {
  const double r2 = r2ij * r2ik * r2jk;
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
  const double cf = C * q;
  const double bmc = r2jk - r2ik, bpc = r2jk + r2ik;
  const double ang = fma(0.375 * ((r2ij + bmc) * (r2ij - bmc) * (bpc - r2ij)), u5, u3);
  e0 = ang * cf;
  gg = 0.0;
  if (want_grad) {
    const double tf = tmp * (cf * d * q);     // tmp * fdmp
    const double m2 = bmc * bmc;
    const double poly = r2ij * fma(r2ij, r2ij + bpc, fma(2.0, bpc * bpc, m2)) - 5.0 * (m2 * bpc);
    const double dang = -0.375 * poly * u5;
    gg = cf * (-dang - 2.0 * alp * ang * tf);
  }
}

// Two-edge variant of atm_core for a centre with BOTH partners owned: one evaluation yields the radial-force factors
// of edge ij (ggj) and of edge ik (ggk); energy and damping terms are shared (the triangle term is symmetric in j, k).
template <bool ALP16>
__device__ __forceinline__ void atm_core2(double r2ij, double r2ik, double r2jk, double C, double c0, double alp,
                                          int want_grad, double& e0, double& ggj, double& ggk) {
  const double r2 = r2ij * r2ik * r2jk;
  const double p = atm_inv_sixth_root(r2);
  const double p2 = p * p, u = p2 * p, u2 = u * u, u3 = u2 * u, u5 = u3 * u2;
  const double v = c0 * p;
  double tmp;
  if (ALP16) {
    const double v2 = v * v, v4 = v2 * v2, v8 = v4 * v4;
    tmp = v8 * v8;
  } else {
    tmp = pow(v, alp);
  }
  const double d = fma(6.0, tmp, 1.0);
  const double q = atm_rsqrt(C * d * d);
  const double cf = C * q;
  const double bmc = r2jk - r2ik, bpc = r2jk + r2ik;
  const double ang = fma(0.375 * ((r2ij + bmc) * (r2ij - bmc) * (bpc - r2ij)), u5, u3);
  e0 = ang * cf;
  ggj = 0.0; ggk = 0.0;
  if (want_grad) {
    const double tf = tmp * (cf * d * q);
    const double damp = 2.0 * alp * ang * tf;
    const double m2 = bmc * bmc;
    const double polyj = r2ij * fma(r2ij, r2ij + bpc, fma(2.0, bpc * bpc, m2)) - 5.0 * (m2 * bpc);
    const double bmj = r2jk - r2ij, bpj = r2jk + r2ij, mj2 = bmj * bmj;
    const double polyk = r2ik * fma(r2ik, r2ik + bpj, fma(2.0, bpj * bpj, mj2)) - 5.0 * (mj2 * bpj);
    const double s = 0.375 * u5;
    ggj = cf * fma(s, polyj, -damp);
    ggk = cf * fma(s, polyk, -damp);
  }
}

// D4S = true (src/dftd4/model/d4s.f90): the weights of j in c6_jk depend on the species of k, so they are re-read per
// visit (L1-resident) instead of living in registers; everything else is shared with the D4 path.
template <bool SMOOTH, bool ALP16, bool D4S>
__global__ void __launch_bounds__(ATM_THREADS, ATM_MINB) atm_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const Point* __restrict__ pts, const long long* __restrict__ offsets, const unsigned* __restrict__ idx,
    const int* __restrict__ triples, Tables tb, const double* __restrict__ gw0, const double* __restrict__ gwdcn0,
    const double* __restrict__ W0, const double* __restrict__ W0cn, const __grid_constant__ AtmParams ap, int tsp,
    double* __restrict__ e_atom, double* __restrict__ dEdcn, double* __restrict__ grad, double* __restrict__ sig_part,
    unsigned long long* __restrict__ counters, int want_grad, double* __restrict__ e3, int e3_ld) {
  extern __shared__ double sm[];
  double* UI = sm;                        // [tsp]  W0 row of the centre
  double* UIcn = UI + tsp;                // [tsp]
  double2* GA = reinterpret_cast<double2*>(UIcn + tsp);  // [KC] {kx, ky}
  double2* GB = GA + ATM_KC;              // [KC] {kz, r2ik}
  double2* GC = GB + ATM_KC;              // [KC] {c6ik, cbrt(R0ik)}
  double* ksw = reinterpret_cast<double*>(GC + ATM_KC);  // [KC]
  double* KW = ksw + ATM_KC;              // [KC][tsp]
  unsigned* ebuf = reinterpret_cast<unsigned*>(KW + (size_t)ATM_KC * tsp);  // [2*ATM_THREADS]
  int* katom = reinterpret_cast<int*>(ebuf + 2 * ATM_THREADS);
  int* ktr = katom + ATM_KC;
  int* kflag = ktr + ATM_KC;
  __shared__ int s_wcnt[ATM_WARPS];
  __shared__ double s_acc[ATM_WARPS][12];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = nrows - 1 - blockIdx.x;  // rows are ascending in atom index: heaviest centres (most owned edges) first
  const int i = rows[r];
  const int spi = species[i];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi_ = xyz[3 * i + 2];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  const double near_tol = ap.cutoff2 * 1.0e-10;

  for (int e = tid; e < tsp; e += ATM_THREADS) { UI[e] = W0[(size_t)i * tsp + e]; UIcn[e] = W0cn[(size_t)i * tsp + e]; }
  if (lane < 12) s_acc[warp][lane] = 0.0;
  __syncthreads();

  unsigned long long visits = 0;
  int pos = 0, cnt = 0;  // next list position to scan, owned edges waiting in ebuf (both uniform across the CTA)
  while (true) {
    // ---- refill the edge buffer: owned = partner index below the centre's, or a self image ----
    while (cnt < ATM_THREADS && pos < n) {
      const int c = pos + tid;
      bool own = false;
      if (c < n) own = point_atom(pts[idx[beg + c]].w) <= i;
      const unsigned m = __ballot_sync(D4_FULL, own);
      if (lane == 0) s_wcnt[warp] = __popc(m);
      __syncthreads();
      int base = cnt, tot = 0;
      for (int w = 0; w < ATM_WARPS; ++w) { if (w < warp) base += s_wcnt[w]; tot += s_wcnt[w]; }
      if (own) ebuf[base + __popc(m & ((1u << lane) - 1u))] = (unsigned)c;
      cnt += tot;
      pos += ATM_THREADS;
      __syncthreads();
    }
    if (cnt == 0) break;
    const int ne = min(cnt, ATM_THREADS);

    // ---- this thread's edge: geometry, C6-scaled weights gjs = c6ij * gw0_j, cbrt(R0ij) ----
    const bool valid = tid < ne;
    double vjx = 0, vjy = 0, vjz = 0, r2ij = 1.0, cr0ij = 1.0, swij = 1.0;
    double gjs[D4_MREF];
    double c6ij_e = 0.0;  // D4S only: c6ij kept apart from the weights
    int j = 0, tj = 0, so_j = 0, nrj = 0;
    unsigned vinc = 2u;  // ordered-triple count per visit: 2 for a regular edge, 1 for a half-weight self-image edge
    bool near_ij = false;
#pragma unroll
    for (int b = 0; b < D4_MREF; ++b) gjs[b] = 0.0;
    if (valid) {
      const Point p = load_point(pts + idx[beg + ebuf[tid]]);
      vjx = __dsub_rn(p.x, xi); vjy = __dsub_rn(p.y, yi); vjz = __dsub_rn(p.z, zi_);
      r2ij = exact_r2(vjx, vjy, vjz);
      j = point_atom(p.w); tj = point_trans(p.w);
      const int zj = point_species(p.w);
      so_j = tb.soff[zj]; nrj = tb.nref[zj];
      double c6ij = 0.0;
      const size_t gj = gw_row(tb, j, spi);
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b)
        if (b < nrj) { gjs[b] = gw0[gj + b]; c6ij += UI[so_j + b] * gjs[b]; }
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b) gjs[b] *= c6ij;
      c6ij_e = c6ij;
      cr0ij = UI[so_j + nrj];
      vinc = (j == i) ? 1u : 2u;
      near_ij = fabs(r2ij - ap.cutoff2) <= near_tol;
      if (SMOOTH) { double dsw; smooth_cutoff(sqrt(r2ij), ap.cutoff, ap.width, swij, dsw); }
    }
    const double cut2v = valid ? ap.cutoff2 : -1.0;
    double se = 0.0, sg = 0.0;  // sums of  c6-product-scaled  ang*fdmp  and radial-force terms (without -s9)

    // ---- stream all neighbours k' of the centre in chunks ----
    for (int c0 = 0; c0 < n; c0 += ATM_KC) {
      const int kc = min(ATM_KC, n - c0);
      __syncthreads();
      if (tid < kc) {
        const Point p = load_point(pts + idx[beg + c0 + tid]);
        const double vx = __dsub_rn(p.x, xi), vy = __dsub_rn(p.y, yi), vz = __dsub_rn(p.z, zi_);
        const double r2 = exact_r2(vx, vy, vz);
        const int k = point_atom(p.w), zk = point_species(p.w);
        const int so = tb.soff[zk], nrk = tb.nref[zk];
        double c6 = 0.0;
        const size_t gk = gw_row(tb, k, spi);
        for (int b = 0; b < nrk; ++b) c6 += UI[so + b] * gw0[gk + b];
        GA[tid] = make_double2(vx, vy); GB[tid] = make_double2(vz, r2); GC[tid] = make_double2(c6, UI[so + nrk]);
        katom[tid] = k; ktr[tid] = point_trans(p.w);
        kflag[tid] = (k > i ? 1 : 0) | (fabs(r2 - ap.cutoff2) <= near_tol ? 2 : 0) | (zk << 8);
        if (SMOOTH) { double sw, dsw; smooth_cutoff(sqrt(r2), ap.cutoff, ap.width, sw, dsw); ksw[tid] = sw; }
      }
      for (int e = tid; e < kc * tsp; e += ATM_THREADS) {
        const int kk = e / tsp, o = e - kk * tsp;
        KW[e] = W0[(size_t)point_atom(pts[idx[beg + c0 + kk]].w) * tsp + o];
      }
      __syncthreads();

      int kk = 0;
#if ATM_UNROLL2
      // hot configuration: two third atoms per iteration -- two independent dependency chains per lane hide the FP64
      // latency with fewer resident warps, and the loop / address overhead is paid once per pair
      if (!SMOOTH && !D4S) {
        for (; kk + 1 < kc; kk += 2) {
          const double2 ga0 = GA[kk], gb0 = GB[kk], ga1 = GA[kk + 1], gb1 = GB[kk + 1];
          const double r2jk0 = exact_r2(__dsub_rn(ga0.x, vjx), __dsub_rn(ga0.y, vjy), __dsub_rn(gb0.x, vjz));
          const double r2jk1 = exact_r2(__dsub_rn(ga1.x, vjx), __dsub_rn(ga1.y, vjy), __dsub_rn(gb1.x, vjz));
          bool ok0 = !(r2jk0 > cut2v || r2jk0 < ap.eps2), ok1 = !(r2jk1 > cut2v || r2jk1 < ap.eps2);
          const int kf0 = kflag[kk], kf1 = kflag[kk + 1];
          if ((kf0 | kf1) & 1) {
            if ((kf0 & 1) && valid && ((kf0 & 2) || near_ij || fabs(r2jk0 - ap.cutoff2) <= near_tol))
              ok0 = atm_frame_check(i, j, tj, katom[kk], ktr[kk], xyz, triples, ap);
            if ((kf1 & 1) && valid && ((kf1 & 2) || near_ij || fabs(r2jk1 - ap.cutoff2) <= near_tol))
              ok1 = atm_frame_check(i, j, tj, katom[kk + 1], ktr[kk + 1], xyz, triples, ap);
          }
          if (!__any_sync(D4_FULL, ok0 || ok1)) continue;
          const double2 gc0 = GC[kk], gc1 = GC[kk + 1];
          const double* Wk0 = KW + (size_t)kk * tsp + so_j;
          const double* Wk1 = Wk0 + tsp;
          double c6p0 = 0.0, c6p1 = 0.0;
#pragma unroll
          for (int b = 0; b < D4_MREF; ++b) { c6p0 = fma(gjs[b], Wk0[b], c6p0); c6p1 = fma(gjs[b], Wk1[b], c6p1); }
          double e00, gg0, e01, gg1;
          atm_core<ALP16>(r2ij, gb0.y, r2jk0, fabs(c6p0 * gc0.x), cr0ij * gc0.y * Wk0[nrj], ap.alp, want_grad, e00, gg0);
          atm_core<ALP16>(r2ij, gb1.y, r2jk1, fabs(c6p1 * gc1.x), cr0ij * gc1.y * Wk1[nrj], ap.alp, want_grad, e01, gg1);
          if (ok0) { se += e00; sg += gg0; visits += vinc; }
          if (ok1) { se += e01; sg += gg1; visits += vinc; }
        }
      }
#endif
      for (; kk < kc; ++kk) {
        const double2 ga = GA[kk], gb = GB[kk];
        const double dx = __dsub_rn(ga.x, vjx), dy = __dsub_rn(ga.y, vjy), dz = __dsub_rn(gb.x, vjz);
        const double r2jk = exact_r2(dx, dy, dz);
        bool ok = !(r2jk > cut2v || r2jk < ap.eps2);   // cut2v < 0 for idle lanes
        const int kf = kflag[kk];
        if (kf & 1) {  // third atom above the centre: the reference evaluates this triangle in the third atom's frame
          if (valid && ((kf & 2) || near_ij || fabs(r2jk - ap.cutoff2) <= near_tol))
            ok = atm_frame_check(i, j, tj, katom[kk], ktr[kk], xyz, triples, ap);
        }
        double swjk = 1.0;
        if (SMOOTH) {
          double dsw;
          smooth_cutoff(sqrt(r2jk), ap.cutoff, ap.width, swjk, dsw);
          ok = ok && (swij * ksw[kk] * swjk > 0.0);
        }
        if (!__any_sync(D4_FULL, ok)) continue;
        const double2 gc = GC[kk];
        const double* Wk = KW + (size_t)kk * tsp + so_j;
        double c6p = 0.0;  // c6ij * c6jk
        if (D4S) {
          const double* gjk = gw0 + ((size_t)j * tb.nsp + (kf >> 8)) * D4_MREF;  // weights of j against species(k)
          for (int b = 0; b < nrj; ++b) c6p = fma(gjk[b], Wk[b], c6p);
          c6p *= c6ij_e;
        } else {
#pragma unroll
          for (int b = 0; b < D4_MREF; ++b) c6p = fma(gjs[b], Wk[b], c6p);
        }
        double e0, gg;
        atm_core<ALP16>(r2ij, gb.y, r2jk, fabs(c6p * gc.x), cr0ij * gc.y * Wk[nrj], ap.alp, want_grad, e0, gg);
        if (SMOOTH) { const double f = ksw[kk] * swjk; e0 *= f; gg *= f; }
        if (ok) { se += e0; sg += gg; visits += vinc; }
      }
    }

    // ---- apply the per-edge sums once ----
    double out[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // e, dcn_i, g[3], s[6] of the centre
    if (valid) {
      double dswij = 0.0;
      const double rij = sqrt(r2ij);
      if (SMOOTH) smooth_cutoff(rij, ap.cutoff, ap.width, swij, dswij);
      const double wedge = (j == i ? 0.5 : 1.0) * -ap.s9;
      se *= wedge; sg *= wedge;
      const double eedge = swij * se;  // sum over k of dE0 * sw, weighted
      out[0] = -eedge * (1.0 / 3.0);
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
        double coef = swij * sg / r2ij;
        if (SMOOTH) coef -= se * dswij / rij;
        const double gx = coef * vjx, gy = coef * vjy, gz = coef * vjz;  // dGij summed over k
        out[2] = -gx; out[3] = -gy; out[4] = -gz;
        atomicAdd(&grad[3 * j], gx); atomicAdd(&grad[3 * j + 1], gy); atomicAdd(&grad[3 * j + 2], gz);
        out[5] = gx * vjx; out[6] = gx * vjy; out[7] = gx * vjz;
        out[8] = gy * vjy; out[9] = gy * vjz; out[10] = gz * vjz;
        out[1] = -0.5 * eedge * dci / c6ij;
        atomicAdd(&dEdcn[j], -0.5 * eedge * dcj / c6ij);
      }
    }
#pragma unroll
    for (int c = 0; c < 11; ++c) {
      const double t = warp_sum(out[c]);
      if (lane == 0) s_acc[warp][c] += t;
    }
    // ---- drop the processed edges ----
    __syncthreads();
    const int rem = cnt - ne;
    unsigned keep = 0;
    if (tid < rem) keep = ebuf[ne + tid];
    __syncthreads();
    if (tid < rem) ebuf[tid] = keep;
    cnt = rem;
    __syncthreads();
  }

  // ---- CTA reduction of the centre's own quantities ----
  const unsigned long long v = warp_sum_u64(visits);
  if (lane == 0) s_acc[warp][11] = __longlong_as_double((long long)v);
  __syncthreads();
  if (tid == 0) {
    double t[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    unsigned long long vt = 0;
    for (int w = 0; w < ATM_WARPS; ++w) {
      for (int c = 0; c < 11; ++c) t[c] += s_acc[w][c];
      vt += (unsigned long long)__double_as_longlong(s_acc[w][11]);
    }
    atomicAdd(&e_atom[i], t[0]);
    atomicAdd(&counters[4], vt);
    if (want_grad) {
      atomicAdd(&dEdcn[i], t[1]);
      atomicAdd(&grad[3 * i], t[2]); atomicAdd(&grad[3 * i + 1], t[3]); atomicAdd(&grad[3 * i + 2], t[4]);
      double* sp = sig_part + (size_t)i * 6;
      for (int c = 0; c < 6; ++c) sp[c] += t[5 + c];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Two-edge scheme (the default for every configuration).  row_own_partition_kernel reorders every row of the
// disp3 list into [owned partners | partners above the centre], keeping the cell order inside both parts.  The edges of
// a pass are then 256 consecutive owned entries, and a pair (j', k') of two owned partners is visited ONCE -- by the
// thread of the entry that comes first in the row -- producing the sums of both centre edges: the second edge's two
// scalars are reduced over the warp (k' is warp-uniform) and added to a per-entry accumulator in global memory that the
// owner of that entry reads when it finalises its edge.  Partners above the centre are streamed as in atm_kernel.
// A triangle a < b < c costs one two-edge visit at c and one single-edge visit at b instead of three visits.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) row_own_partition_kernel(int nrows, const int* __restrict__ rows,
                                                                const Point* __restrict__ pts,
                                                                const long long* __restrict__ offsets,
                                                                const unsigned* __restrict__ idx_in,
                                                                unsigned* __restrict__ idx_out, int* __restrict__ nown) {
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
  if (tid == 0) nown[r] = no;
}

static_assert(ATM_THREADS % ATM_KC == 0 && (!ATM_CYCLIC || (ATM_THREADS & (ATM_THREADS - 1)) == 0), "a pass must start on a chunk boundary");

template <bool SMOOTH, bool ALP16, bool D4S>
__global__ void __launch_bounds__(ATM_THREADS, ATM_MINB) atm_two_edge_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const Point* __restrict__ pts, const long long* __restrict__ offsets, const unsigned* __restrict__ idx,
    const int* __restrict__ nown, const int* __restrict__ triples, Tables tb, const double* __restrict__ gw0,
    const double* __restrict__ gwdcn0, const double* __restrict__ W0, const double* __restrict__ W0cn,
    const __grid_constant__ AtmParams ap,
    int tsp, double2* __restrict__ eacc, double* __restrict__ e_atom, double* __restrict__ dEdcn,
    double* __restrict__ grad, double* __restrict__ sig_part, unsigned long long* __restrict__ counters, int want_grad,
    double* __restrict__ e3, int e3_ld) {
  extern __shared__ double sm[];
  double* UI = sm;
  double* UIcn = UI + tsp;
  double2* GA = reinterpret_cast<double2*>(UIcn + tsp);
  double2* GB = GA + ATM_KC;
  double2* GC = GB + ATM_KC;
  double* ksw = reinterpret_cast<double*>(GC + ATM_KC);           // [KC] (same layout as atm_kernel)
  double* KW = ksw + ATM_KC;
  int* katom = reinterpret_cast<int*>(reinterpret_cast<unsigned*>(KW + (size_t)ATM_KC * tsp) + 2 * ATM_THREADS);
  int* ktr = katom + ATM_KC;
  int* kflag = ktr + ATM_KC;
  __shared__ double s_acc[ATM_WARPS][12];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = nrows - 1 - blockIdx.x;
  const int i = rows[r];
  const int spi = species[i];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi_ = xyz[3 * i + 2];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  const int no = nown[r];
  const double near_tol = ap.cutoff2 * 1.0e-10;

  for (int e = tid; e < tsp; e += ATM_THREADS) { UI[e] = W0[(size_t)i * tsp + e]; UIcn[e] = W0cn[(size_t)i * tsp + e]; }
  for (int e = tid; e < no; e += ATM_THREADS) eacc[beg + e] = make_double2(0.0, 0.0);
  if (lane < 12) s_acc[warp][lane] = 0.0;
  __syncthreads();

  unsigned long long visits = 0;
  for (int p0 = 0; p0 < no; p0 += ATM_THREADS) {
    // ---- this thread's edge ----
    const int pj = p0 + tid;
    const bool valid = pj < no;
    const int w0 = p0 + warp * 32;  // row position of the warp's first edge
    double vjx = 0, vjy = 0, vjz = 0, r2ij = 1.0, cr0ij = 1.0, swij = 1.0;
    double gjs[D4_MREF];
    double c6ij_e = 0.0;  // D4S only: c6ij kept apart from the weights
    int j = 0, tj = 0, so_j = 0, nrj = 0;
    unsigned vinc = 2u;
    bool near_ij = false;
#pragma unroll
    for (int b = 0; b < D4_MREF; ++b) gjs[b] = 0.0;
    if (valid) {
      const Point p = load_point(pts + idx[beg + pj]);
      vjx = __dsub_rn(p.x, xi); vjy = __dsub_rn(p.y, yi); vjz = __dsub_rn(p.z, zi_);
      r2ij = exact_r2(vjx, vjy, vjz);
      j = point_atom(p.w); tj = point_trans(p.w);
      const int zj = point_species(p.w);
      so_j = tb.soff[zj]; nrj = tb.nref[zj];
      double c6ij = 0.0;
      const size_t gj = gw_row(tb, j, spi);
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b)
        if (b < nrj) { gjs[b] = gw0[gj + b]; c6ij += UI[so_j + b] * gjs[b]; }
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b) gjs[b] *= c6ij;
      c6ij_e = c6ij;
      cr0ij = UI[so_j + nrj];
      vinc = (j == i) ? 1u : 2u;
      near_ij = fabs(r2ij - ap.cutoff2) <= near_tol;
      if (SMOOTH) { double dsw; smooth_cutoff(sqrt(r2ij), ap.cutoff, ap.width, swij, dsw); }
    }
    double se = 0.0, sg = 0.0;
    unsigned nvis = 0;  // ordered triples of this pass (a row holds far fewer than 2^30 entries)
    // warp sums of a visit's second-edge pair (e0, ggk): lanes 0-15 fold the first value, lanes 16-31 the second
    // (`mine` = the value this half keeps, `other` = the one it hands to the other half), then one RED per value
    const bool hi = (lane & 16) != 0;
    double* const eacc_row = reinterpret_cast<double*>(eacc + beg) + (hi ? 1 : 0);  // this half-warp's component
    auto fold2 = [&](double mine, double other, int q) {
      double keep = mine + __shfl_xor_sync(D4_FULL, other, 16);
      keep += __shfl_xor_sync(D4_FULL, keep, 8);
      keep += __shfl_xor_sync(D4_FULL, keep, 4);
      keep += __shfl_xor_sync(D4_FULL, keep, 2);
      keep += __shfl_xor_sync(D4_FULL, keep, 1);
      if ((lane & 15) == 0) atomicAdd(eacc_row + 2 * (size_t)q, keep);
    };
#if ATM_PENDING
    double pend = 0.0, pend_x = 0.0;  // sums of the last executed visit, folded at the start of the next one
    int pend_q = p0;                  // (adding the initial zeros to the pass's first entry is a no-op)
#endif

    // ---- stream the partners from the pass's own position on (earlier owned partners took this pass's edges as their
    //      second edge already) ----
    for (int c0 = p0; c0 < n; c0 += ATM_KC) {
      const int kc = min(ATM_KC, n - c0);
      __syncthreads();
      if (tid < kc) {
        const Point p = load_point(pts + idx[beg + c0 + tid]);
        const double vx = __dsub_rn(p.x, xi), vy = __dsub_rn(p.y, yi), vz = __dsub_rn(p.z, zi_);
        const double r2 = exact_r2(vx, vy, vz);
        const int k = point_atom(p.w), zk = point_species(p.w);
        const int so = tb.soff[zk], nrk = tb.nref[zk];
        double c6 = 0.0;
        const size_t gk = gw_row(tb, k, spi);
        for (int b = 0; b < nrk; ++b) c6 += UI[so + b] * gw0[gk + b];
        GA[tid] = make_double2(vx, vy); GB[tid] = make_double2(vz, r2); GC[tid] = make_double2(c6, UI[so + nrk]);
        katom[tid] = k; ktr[tid] = point_trans(p.w);
        kflag[tid] = (k > i ? 1 : 0) | (fabs(r2 - ap.cutoff2) <= near_tol ? 2 : 0) | (k == i ? 4 : 0) | (zk << 8);
        if (SMOOTH) { double sw, dsw; smooth_cutoff(sqrt(r2), ap.cutoff, ap.width, sw, dsw); ksw[tid] = sw; }
      }
      for (int e = tid; e < kc * tsp; e += ATM_THREADS) {
        const int kk = e / tsp, o = e - kk * tsp;
        KW[e] = W0[(size_t)point_atom(pts[idx[beg + c0 + kk]].w) * tsp + o];
      }
      __syncthreads();

      // two-edge visits.  Entries behind the pass's 256-entry block pair with every edge of the pass; inside the block a
      // pair is taken by the entry from which the other lies 1..128 positions ahead cyclically (128: the lower entry),
      // so every warp of the pass gets the same share of its own block (a triangular rule leaves the upper warps idle
      // at the barriers of the block's chunks).
      const int k2end = min(kc, no - c0);
      auto visit2 = [&](int kk, bool okpos) {
        const double2 ga = GA[kk], gb = GB[kk];
        const double dx = __dsub_rn(ga.x, vjx), dy = __dsub_rn(ga.y, vjy), dz = __dsub_rn(gb.x, vjz);
        const double r2jk = exact_r2(dx, dy, dz);
        bool ok = valid && !(r2jk > ap.cutoff2 || r2jk < ap.eps2) && okpos;
        double swjk = 1.0;
        if (SMOOTH) {
          double dsw;
          smooth_cutoff(sqrt(r2jk), ap.cutoff, ap.width, swjk, dsw);
          ok = ok && (swij * ksw[kk] * swjk > 0.0);
        }
        if (!__any_sync(D4_FULL, ok)) return;
#if ATM_PENDING
        fold2(pend, pend_x, pend_q);  // the previous visit's second-edge sums: their shuffle chain overlaps this visit
#endif
        const double2 gc = GC[kk];
        const double* Wk = KW + (size_t)kk * tsp + so_j;
        const int kf = kflag[kk];
        double c6p = 0.0;
        if (D4S) {
          const double* gjk = gw0 + ((size_t)j * tb.nsp + (kf >> 8)) * D4_MREF;  // weights of j against species(k)
          for (int b = 0; b < nrj; ++b) c6p = fma(gjk[b], Wk[b], c6p);
          c6p *= c6ij_e;
        } else {
#pragma unroll
          for (int b = 0; b < D4_MREF; ++b) c6p = fma(gjs[b], Wk[b], c6p);
        }
        double e0, ggj, ggk;
        atm_core2<ALP16>(r2ij, gb.y, r2jk, fabs(c6p * gc.x), cr0ij * gc.y * Wk[nrj], ap.alp, want_grad, e0, ggj, ggk);
        // rejected lanes contribute zeros (selects, no divergent region)
        double ka = ok ? e0 : 0.0, kb = ok ? ggk : 0.0, ja = ka, jb = ok ? ggj : 0.0;
        if (SMOOTH) {  // each edge's sums carry the cutoff factors of the OTHER two sides
          const double fj = ksw[kk] * swjk, fk = swij * swjk;
          ja *= fj; jb *= fj; ka *= fk; kb *= fk;
        }
        se += ja; sg += jb;
        nvis += ok ? vinc + ((kf & 4) ? 1u : 2u) : 0u;
#if ATM_PENDING
        pend = hi ? kb : ka; pend_x = hi ? ka : kb; pend_q = c0 + kk;
#else
        fold2(hi ? kb : ka, hi ? ka : kb, c0 + kk);
#endif
      };
      if (c0 - p0 >= ATM_THREADS) {
        for (int kk = 0; kk < k2end; ++kk) visit2(kk, true);
      } else {
#if ATM_CYCLIC
        constexpr int HALF = ATM_THREADS / 2;
        for (int kk = 0; kk < k2end; ++kk) {
          const int t = (c0 + kk - w0) & (ATM_THREADS - 1);   // distance ahead of the warp's first edge
          if (t < 1 || t > HALF + 31) continue;
          const int d = (t - lane) & (ATM_THREADS - 1);       // distance ahead of this lane's edge
          visit2(kk, d >= 1 && (d < HALF || (d == HALF && pj < c0 + kk)));
        }
#else
        for (int kk = max(0, w0 + 1 - c0); kk < k2end; ++kk) visit2(kk, c0 + kk > pj);
#endif
      }
      // single-edge visits: partners above the centre
      for (int kk = max(0, no - c0); kk < kc; ++kk) {
        const double2 ga = GA[kk], gb = GB[kk];
        const double dx = __dsub_rn(ga.x, vjx), dy = __dsub_rn(ga.y, vjy), dz = __dsub_rn(gb.x, vjz);
        const double r2jk = exact_r2(dx, dy, dz);
        bool ok = valid && !(r2jk > ap.cutoff2 || r2jk < ap.eps2);
        const int kf = kflag[kk];
        if (kf & 1) {
          if (valid && ((kf & 2) || near_ij || fabs(r2jk - ap.cutoff2) <= near_tol))
            ok = atm_frame_check(i, j, tj, katom[kk], ktr[kk], xyz, triples, ap);
        }
        double swjk = 1.0;
        if (SMOOTH) {
          double dsw;
          smooth_cutoff(sqrt(r2jk), ap.cutoff, ap.width, swjk, dsw);
          ok = ok && (swij * ksw[kk] * swjk > 0.0);
        }
        if (!__any_sync(D4_FULL, ok)) continue;
        const double2 gc = GC[kk];
        const double* Wk = KW + (size_t)kk * tsp + so_j;
        double c6p = 0.0;
        if (D4S) {
          const double* gjk = gw0 + ((size_t)j * tb.nsp + (kf >> 8)) * D4_MREF;
          for (int b = 0; b < nrj; ++b) c6p = fma(gjk[b], Wk[b], c6p);
          c6p *= c6ij_e;
        } else {
#pragma unroll
          for (int b = 0; b < D4_MREF; ++b) c6p = fma(gjs[b], Wk[b], c6p);
        }
        double e0, gg;
        atm_core<ALP16>(r2ij, gb.y, r2jk, fabs(c6p * gc.x), cr0ij * gc.y * Wk[nrj], ap.alp, want_grad, e0, gg);
        if (SMOOTH) { const double f = ksw[kk] * swjk; e0 *= f; gg *= f; }
        if (ok) { se += e0; sg += gg; nvis += vinc; }
      }
    }
#if ATM_PENDING
    fold2(pend, pend_x, pend_q);
#endif
    __syncthreads();  // every second-edge sum aimed at this pass's entries has been issued

    visits += nvis;
    // ---- apply the per-edge sums once ----
    double out[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (valid) {
      double* acc = reinterpret_cast<double*>(eacc + beg + pj);
      se += atomicAdd(acc, 0.0);      // read at L2, where the reductions landed
      sg += atomicAdd(acc + 1, 0.0);
      double dswij = 0.0;
      const double rij = sqrt(r2ij);
      if (SMOOTH) smooth_cutoff(rij, ap.cutoff, ap.width, swij, dswij);
      const double wedge = (j == i ? 0.5 : 1.0) * -ap.s9;
      se *= wedge; sg *= wedge;
      const double eedge = swij * se;
      out[0] = -eedge * (1.0 / 3.0);
      if (e3) {
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
        double coef = swij * sg / r2ij;
        if (SMOOTH) coef -= se * dswij / rij;
        const double gx = coef * vjx, gy = coef * vjy, gz = coef * vjz;
        out[2] = -gx; out[3] = -gy; out[4] = -gz;
        atomicAdd(&grad[3 * j], gx); atomicAdd(&grad[3 * j + 1], gy); atomicAdd(&grad[3 * j + 2], gz);
        out[5] = gx * vjx; out[6] = gx * vjy; out[7] = gx * vjz;
        out[8] = gy * vjy; out[9] = gy * vjz; out[10] = gz * vjz;
        out[1] = -0.5 * eedge * dci / c6ij;
        atomicAdd(&dEdcn[j], -0.5 * eedge * dcj / c6ij);
      }
    }
#pragma unroll
    for (int c = 0; c < 11; ++c) {
      const double t = warp_sum(out[c]);
      if (lane == 0) s_acc[warp][c] += t;
    }
  }

  const unsigned long long v = warp_sum_u64(visits);
  if (lane == 0) s_acc[warp][11] = __longlong_as_double((long long)v);
  __syncthreads();
  if (tid == 0) {
    double t[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    unsigned long long vt = 0;
    for (int w = 0; w < ATM_WARPS; ++w) {
      for (int c = 0; c < 11; ++c) t[c] += s_acc[w][c];
      vt += (unsigned long long)__double_as_longlong(s_acc[w][11]);
    }
    atomicAdd(&e_atom[i], t[0]);
    atomicAdd(&counters[4], vt);
    if (want_grad) {
      atomicAdd(&dEdcn[i], t[1]);
      atomicAdd(&grad[3 * i], t[2]); atomicAdd(&grad[3 * i + 1], t[3]); atomicAdd(&grad[3 * i + 2], t[4]);
      double* sp = sig_part + (size_t)i * 6;
      for (int c = 0; c < 6; ++c) sp[c] += t[5 + c];
    }
  }
}

}  // namespace d4k
