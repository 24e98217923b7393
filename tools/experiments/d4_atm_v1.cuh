// Round-1 two-edge ATM kernel (CTA per centre, lockstep chunk staging), kept for A/B runs against the streaming
// kernels of d4_atm.cuh; selected with D4CU_ATM_SCHEME=1.  Not the default path.
#pragma once
#include "d4_atm.cuh"

namespace d4k {

constexpr int ATM1_THREADS = 256;  // = edges per pass
constexpr int ATM1_WARPS = ATM1_THREADS / 32;
constexpr int ATM1_KC = 128;       // neighbours per staged chunk
constexpr int ATM1_GEO = 7;

__host__ __device__ inline size_t atm1_smem_bytes(int tsp) {
  return sizeof(double) * ((size_t)2 * tsp + (size_t)ATM1_GEO * ATM1_KC + (size_t)ATM1_KC * tsp) +
         sizeof(unsigned) * 2 * ATM1_THREADS + sizeof(int) * 3 * ATM1_KC;
}

// Damped Axilrod-Teller-Muto term of one (edge, third atom) visit: e0 = ang * fdmp * sqrt(C) and, with want_grad, the
// radial-force factor of edge ij (atm.f90:360-399 restated division-free; C = c6ij c6ik c6jk, c0 = (R0ij R0ik R0jk)^(1/3)).
template <bool ALP16>
__device__ __forceinline__ void atm_core_v1(double r2ij, double r2ik, double r2jk, double C, double c0, double alp,
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
__device__ __forceinline__ void atm_core2_v1(double r2ij, double r2ik, double r2jk, double C, double c0, double alp,
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

static_assert(ATM1_THREADS % ATM1_KC == 0, "a pass must start on a chunk boundary");

template <bool SMOOTH, bool ALP16, bool D4S>
__global__ void __launch_bounds__(ATM1_THREADS, 3) atm_two_edge_kernel(
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
  double2* GB = GA + ATM1_KC;
  double2* GC = GB + ATM1_KC;
  double* ksw = reinterpret_cast<double*>(GC + ATM1_KC);           // [KC] (same layout as atm_kernel)
  double* KW = ksw + ATM1_KC;
  int* katom = reinterpret_cast<int*>(reinterpret_cast<unsigned*>(KW + (size_t)ATM1_KC * tsp) + 2 * ATM1_THREADS);
  int* ktr = katom + ATM1_KC;
  int* kflag = ktr + ATM1_KC;
  __shared__ double s_acc[ATM1_WARPS][12];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = nrows - 1 - blockIdx.x;
  const int i = rows[r];
  const int spi = species[i];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi_ = xyz[3 * i + 2];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  const int no = nown[r];
  const double near_tol = ap.cutoff2 * 1.0e-10;

  for (int e = tid; e < tsp; e += ATM1_THREADS) { UI[e] = W0[(size_t)i * tsp + e]; UIcn[e] = W0cn[(size_t)i * tsp + e]; }
  for (int e = tid; e < no; e += ATM1_THREADS) eacc[beg + e] = make_double2(0.0, 0.0);
  if (lane < 12) s_acc[warp][lane] = 0.0;
  __syncthreads();

  unsigned long long visits = 0;
  for (int p0 = 0; p0 < no; p0 += ATM1_THREADS) {
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

    // ---- stream the partners from the pass's own position on (earlier owned partners took this pass's edges as their
    //      second edge already) ----
    for (int c0 = p0; c0 < n; c0 += ATM1_KC) {
      const int kc = min(ATM1_KC, n - c0);
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
      for (int e = tid; e < kc * tsp; e += ATM1_THREADS) {
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
        atm_core2_v1<ALP16>(r2ij, gb.y, r2jk, fabs(c6p * gc.x), cr0ij * gc.y * Wk[nrj], ap.alp, want_grad, e0, ggj, ggk);
        // rejected lanes contribute zeros (selects, no divergent region)
        double ka = ok ? e0 : 0.0, kb = ok ? ggk : 0.0, ja = ka, jb = ok ? ggj : 0.0;
        if (SMOOTH) {  // each edge's sums carry the cutoff factors of the OTHER two sides
          const double fj = ksw[kk] * swjk, fk = swij * swjk;
          ja *= fj; jb *= fj; ka *= fk; kb *= fk;
        }
        se += ja; sg += jb;
        nvis += ok ? vinc + ((kf & 4) ? 1u : 2u) : 0u;
        fold2(hi ? kb : ka, hi ? ka : kb, c0 + kk);
      };
      if (c0 - p0 >= ATM1_THREADS) {
        for (int kk = 0; kk < k2end; ++kk) visit2(kk, true);
      } else {
        for (int kk = max(0, w0 + 1 - c0); kk < k2end; ++kk) visit2(kk, c0 + kk > pj);
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
        atm_core_v1<ALP16>(r2ij, gb.y, r2jk, fabs(c6p * gc.x), cr0ij * gc.y * Wk[nrj], ap.alp, want_grad, e0, gg);
        if (SMOOTH) { const double f = ksw[kk] * swjk; e0 *= f; gg *= f; }
        if (ok) { se += e0; sg += gg; nvis += vinc; }
      }
    }
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
    for (int w = 0; w < ATM1_WARPS; ++w) {
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
