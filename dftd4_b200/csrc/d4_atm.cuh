// Axilrod-Teller-Muto three-body kernel (reference: src/dftd4/damping/atm.f90:235-453, C6 from model/d4.f90:381-473
// with q = 0 weights, disp.f90:120-126).
//
// Formulation (DESIGN.md "ATM").  The reference loops i >= j >= k over atoms and all image pairs with triple_scale.
// Here every accepted triangle is visited once per EDGE.  An edge (i, j') is owned by its higher-index atom i (the
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

constexpr int ATM_THREADS = 256;  // = edges per pass
constexpr int ATM_WARPS = ATM_THREADS / 32;
constexpr int ATM_KC = 64;        // neighbours per staged chunk
constexpr int ATM_GEO = 7;        // kx ky kz kr2 kc6 kr0 ksw

struct AtmParams {
  double s9, a1, a2, alp, cutoff, cutoff2, width, eps2;
  double lat[9];   // lattice[3*k+c] (zero when molecular)
  int rep[3];      // repetitions of the disp3 translation list
};

__host__ __device__ inline size_t atm_smem_bytes(int tsp) {
  // UI, UIcn [tsp] | geometry [ATM_GEO][KC] | W rows [KC][tsp] | ebuf [2*T] u32 | katom, ktr, kflag [KC] int
  return sizeof(double) * ((size_t)2 * tsp + (size_t)ATM_GEO * ATM_KC + (size_t)ATM_KC * tsp) +
         sizeof(unsigned) * 2 * ATM_THREADS + sizeof(int) * 3 * ATM_KC;
}

// W0[a][soff[s] + b] = sum_a' gw0_a[a'] R[z_a][s][a'][b]  (b < nref_s),  W0[a][soff[s] + nref_s] = R0(z_a, s);
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
        for (int r = 0; r < nra; ++r) {
          const double rc6 = tb.refc6[(((size_t)za * nid + s) * D4_MREF + r) * D4_MREF + b];
          u += gw0[(size_t)a * D4_MREF + r] * rc6;
          ucn += gwdcn0[(size_t)a * D4_MREF + r] * rc6;
        }
      } else {
        u = tb.pair_r0[za * nid + s];
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

template <bool SMOOTH, bool ALP16>
__global__ void __launch_bounds__(ATM_THREADS, 2) atm_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const Point* __restrict__ pts, const long long* __restrict__ offsets, const unsigned* __restrict__ idx,
    const int* __restrict__ triples, Tables tb, const double* __restrict__ gw0, const double* __restrict__ gwdcn0,
    const double* __restrict__ W0, const double* __restrict__ W0cn, AtmParams ap, int tsp,
    double* __restrict__ e_atom, double* __restrict__ dEdcn, double* __restrict__ grad, double* __restrict__ sig_part,
    unsigned long long* __restrict__ counters, int want_grad, double* __restrict__ e3, int e3_ld) {
  extern __shared__ double sm[];
  double* UI = sm;                        // [tsp]  W0 row of the centre
  double* UIcn = UI + tsp;                // [tsp]
  double* GEO = UIcn + tsp;               // [ATM_GEO][ATM_KC]
  double* KW = GEO + ATM_GEO * ATM_KC;    // [ATM_KC][tsp]
  unsigned* ebuf = reinterpret_cast<unsigned*>(KW + (size_t)ATM_KC * tsp);  // [2*ATM_THREADS]
  int* katom = reinterpret_cast<int*>(ebuf + 2 * ATM_THREADS);
  int* ktr = katom + ATM_KC;
  int* kflag = ktr + ATM_KC;
  __shared__ int s_wcnt[ATM_WARPS];
  __shared__ double s_red[ATM_WARPS][12];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = nrows - 1 - blockIdx.x;  // rows are ascending in atom index: heaviest centres (most owned edges) first
  const int i = rows[r];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi_ = xyz[3 * i + 2];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  const double near_tol = ap.cutoff2 * 1.0e-10;

  for (int e = tid; e < tsp; e += ATM_THREADS) { UI[e] = W0[(size_t)i * tsp + e]; UIcn[e] = W0cn[(size_t)i * tsp + e]; }
  __syncthreads();

  double* kx = GEO, *ky = GEO + ATM_KC, *kz = GEO + 2 * ATM_KC, *kr2 = GEO + 3 * ATM_KC, *kc6 = GEO + 4 * ATM_KC,
        *kr0 = GEO + 5 * ATM_KC, *ksw = GEO + 6 * ATM_KC;
  const double alp3 = ap.alp / 3.0;

  double acc_e = 0.0, acc_dcn = 0.0, acc_g[3] = {0, 0, 0}, acc_s[6] = {0, 0, 0, 0, 0, 0};
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

    // ---- this thread's edge ----
    const bool valid = tid < ne;
    double vjx = 0, vjy = 0, vjz = 0, r2ij = 1.0, c6ij = 0.0, r0ij = 1.0, swij = 1.0, wedge = 0.0;
    double gj[D4_MREF];
    int j = 0, tj = 0, so_j = 0, nrj = 0;
    unsigned vinc = 2u;  // ordered-triple count per visit: 2 for a regular edge, 1 for a half-weight self-image edge
    bool near_ij = false;
#pragma unroll
    for (int b = 0; b < D4_MREF; ++b) gj[b] = 0.0;
    if (valid) {
      const Point p = load_point(pts + idx[beg + ebuf[tid]]);
      vjx = __dsub_rn(p.x, xi); vjy = __dsub_rn(p.y, yi); vjz = __dsub_rn(p.z, zi_);
      r2ij = exact_r2(vjx, vjy, vjz);
      j = point_atom(p.w); tj = point_trans(p.w);
      const int zj = point_species(p.w);
      so_j = tb.soff[zj]; nrj = tb.nref[zj];
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b)
        if (b < nrj) { gj[b] = gw0[(size_t)j * D4_MREF + b]; c6ij += UI[so_j + b] * gj[b]; }
      r0ij = UI[so_j + nrj];
      wedge = (j == i) ? 0.5 : 1.0;
      vinc = (j == i) ? 1u : 2u;
      near_ij = fabs(r2ij - ap.cutoff2) <= near_tol;
      if (SMOOTH) { double dsw; smooth_cutoff(sqrt(r2ij), ap.cutoff, ap.width, swij, dsw); }
    }
    double se = 0.0, sg = 0.0;

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
        for (int b = 0; b < nrk; ++b) c6 += UI[so + b] * gw0[(size_t)k * D4_MREF + b];
        kx[tid] = vx; ky[tid] = vy; kz[tid] = vz; kr2[tid] = r2; kc6[tid] = c6; kr0[tid] = UI[so + nrk];
        katom[tid] = k; ktr[tid] = point_trans(p.w);
        kflag[tid] = (k > i ? 1 : 0) | (fabs(r2 - ap.cutoff2) <= near_tol ? 2 : 0);
        if (SMOOTH) { double sw, dsw; smooth_cutoff(sqrt(r2), ap.cutoff, ap.width, sw, dsw); ksw[tid] = sw; }
      }
      for (int e = tid; e < kc * tsp; e += ATM_THREADS) {
        const int kk = e / tsp, o = e - kk * tsp;
        KW[e] = W0[(size_t)point_atom(pts[idx[beg + c0 + kk]].w) * tsp + o];
      }
      __syncthreads();

      for (int kk = 0; kk < kc; ++kk) {
        const double dx = __dsub_rn(kx[kk], vjx), dy = __dsub_rn(ky[kk], vjy), dz = __dsub_rn(kz[kk], vjz);
        const double r2jk = exact_r2(dx, dy, dz);
        bool ok = valid && !(r2jk > ap.cutoff2 || r2jk < ap.eps2);
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
        const double* Wk = KW + (size_t)kk * tsp + so_j;
        double c6jk = 0.0;
#pragma unroll
        for (int b = 0; b < D4_MREF; ++b) c6jk = fma(gj[b], Wk[b], c6jk);
        const double r0 = r0ij * kr0[kk] * Wk[nrj];
        const double r2ik = kr2[kk];
        const double c9 = -ap.s9 * sqrt(fabs(c6ij * kc6[kk] * c6jk));
        const double r2 = r2ij * r2ik * r2jk;
        const double u = rsqrt(r2), u2 = u * u, u3 = u2 * u, u5 = u3 * u2;
        const double y = r0 * u;
        double tmp;
        if (ALP16) {
          const double y2 = y * y;
          tmp = y2 * y2 * y * cbrt(y);
        } else {
          tmp = pow(y, alp3);
        }
        const double fdmp = 1.0 / (1.0 + 6.0 * tmp);
        const double ang = fma(0.375 * (r2ij + r2jk - r2ik) * (r2ij - r2jk + r2ik) * (-r2ij + r2jk + r2ik), u5, u3);
        double e0 = ang * fdmp * c9;
        double gg = 0.0;
        if (want_grad) {
          const double dfdmp = -2.0 * ap.alp * tmp * fdmp * fdmp;
          const double dmk = r2jk - r2ik, spk = r2jk + r2ik;
          const double dang = -0.375 * (r2ij * r2ij * r2ij + r2ij * r2ij * spk +
                                        r2ij * (3.0 * r2jk * r2jk + 2.0 * r2jk * r2ik + 3.0 * r2ik * r2ik) -
                                        5.0 * dmk * dmk * spk) * u5;
          gg = c9 * (-dang * fdmp + ang * dfdmp);
        }
        if (SMOOTH) { const double f = ksw[kk] * swjk; e0 *= f; gg *= f; }
        if (ok) { se += e0; sg += gg; visits += vinc; }
      }
    }

    // ---- apply the per-edge sums once ----
    if (valid) {
      double dswij = 0.0;
      const double rij = sqrt(r2ij);
      if (SMOOTH) smooth_cutoff(rij, ap.cutoff, ap.width, swij, dswij);
      const double eedge = wedge * swij * se;  // sum over k of dE0 * sw, weighted
      acc_e -= eedge * (1.0 / 3.0);
      if (e3) {  // pair-resolved energies (rational.f90:487-606): a sixth of the triangle energy per directed pair
        atomicAdd(&e3[(size_t)i * e3_ld + j], -eedge * (1.0 / 6.0));
        atomicAdd(&e3[(size_t)j * e3_ld + i], -eedge * (1.0 / 6.0));
      }
      if (want_grad) {
        double dci = 0.0, dcj = 0.0;
        for (int b = 0; b < nrj; ++b) {
          dci += UIcn[so_j + b] * gw0[(size_t)j * D4_MREF + b];
          dcj += UI[so_j + b] * gwdcn0[(size_t)j * D4_MREF + b];
        }
        double coef = swij * sg / r2ij;
        if (SMOOTH) coef -= se * dswij / rij;
        coef *= wedge;
        const double gx = coef * vjx, gy = coef * vjy, gz = coef * vjz;  // dGij summed over k
        acc_g[0] -= gx; acc_g[1] -= gy; acc_g[2] -= gz;
        atomicAdd(&grad[3 * j], gx); atomicAdd(&grad[3 * j + 1], gy); atomicAdd(&grad[3 * j + 2], gz);
        acc_s[0] += gx * vjx; acc_s[1] += gx * vjy; acc_s[2] += gx * vjz;
        acc_s[3] += gy * vjy; acc_s[4] += gy * vjz; acc_s[5] += gz * vjz;
        acc_dcn -= 0.5 * eedge * dci / c6ij;
        atomicAdd(&dEdcn[j], -0.5 * eedge * dcj / c6ij);
      }
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
  double vals[11] = {acc_e, acc_dcn, acc_g[0], acc_g[1], acc_g[2], acc_s[0], acc_s[1], acc_s[2], acc_s[3], acc_s[4], acc_s[5]};
  for (int c = 0; c < 11; ++c) vals[c] = warp_sum(vals[c]);
  const unsigned long long v = warp_sum_u64(visits);
  if (lane == 0) {
    for (int c = 0; c < 11; ++c) s_red[warp][c] = vals[c];
    s_red[warp][11] = __longlong_as_double((long long)v);
  }
  __syncthreads();
  if (tid == 0) {
    double t[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    unsigned long long vt = 0;
    for (int w = 0; w < ATM_WARPS; ++w) {
      for (int c = 0; c < 11; ++c) t[c] += s_red[w][c];
      vt += (unsigned long long)__double_as_longlong(s_red[w][11]);
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
