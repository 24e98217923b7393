// Axilrod-Teller-Muto three-body kernel, warp-per-edge formulation with lane compaction (hot configuration: D4 model,
// sharp cutoff, alp = 16).  Reference: src/dftd4/damping/atm.f90:235-453, C6 from model/d4.f90:381-473 with q = 0
// weights (disp.f90:120-126).  Same edge decomposition, edge ownership, per-edge application and frame-exact
// acceptance as atm_kernel (d4_atm.cuh) -- what changes is the mapping of the (edge, third atom) visits onto lanes.
//
// atm_kernel keeps one edge per lane and streams the third atoms k' uniformly: only ~47 % of the (j', k') pairs inside
// the centre's cutoff sphere are within the cutoff of each other, so half of the lanes of every visit idle.  Here a
// WARP owns one edge (i, j') at a time and its LANES run over the third atoms of the staged chunk:
//   test phase   32 k' per round: r2(j'k') with the reference's rounding, cutoff test, ballot -> the accepted k' are
//                compacted into a per-warp list in shared memory (index + r2jk)
//   heavy phase  full warps walk the list: C6 contraction, damping, angular term, radial force -- every lane does
//                accepted work (the last batch of a list is the only partial one)
// Edge data (geometry, c6ij-scaled weights of j') are warp-uniform and come from a per-pass edge table in shared memory;
// the per-edge sums SE, SG are reduced across the warp once per (edge, chunk) and accumulated in that table.
#pragma once
#include "d4_atm.cuh"

namespace d4k {

constexpr int A2_THREADS = 256;
constexpr int A2_WARPS = A2_THREADS / 32;
#ifndef A2_EPP
#define A2_EPP 96   // edges per pass (edge-table rows)
#endif
#ifndef A2_KC
#define A2_KC 128   // third atoms per staged chunk
#endif
#ifndef A2_MINB
#define A2_MINB 3   // resident CTAs per SM the register budget is sized for
#endif

// row stride of the precontracted C6 rows for this kernel: odd (in doubles), so that lanes gathering the same column
// of 16 consecutive rows hit 16 distinct bank pairs (a 64-bit gather then needs its minimum of 2 wavefronts)
__host__ __device__ inline int atm2_row_stride(int ts) { return (ts + D4_MREF + 1) | 1; }

__host__ __device__ inline size_t atm2_smem_bytes(int tsp) {
  // double2: GA, GB, GC [KC] | double4 EV [EPP] -> 16-B aligned block first
  // double : ECR, ESE, ESG [EPP] | EG [EPP][7] | UI, UIcn [tsp] | KW [KC][tsp] | LR [WARPS][KC]
  // int    : katom, ktr, kflag [KC] | EJ, ETJ, ESO, ENF [EPP] | LK [WARPS][KC] | ebuf [2*THREADS]
  return sizeof(double) * ((size_t)6 * A2_KC + (size_t)4 * A2_EPP + (size_t)(3 + D4_MREF) * A2_EPP + (size_t)2 * tsp +
                           (size_t)A2_KC * tsp + (size_t)A2_WARPS * A2_KC) +
         sizeof(int) * ((size_t)3 * A2_KC + (size_t)4 * A2_EPP + (size_t)A2_WARPS * A2_KC + (size_t)2 * A2_THREADS);
}

template <bool ALP16>
__global__ void __launch_bounds__(A2_THREADS, A2_MINB) atm2_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const Point* __restrict__ pts, const long long* __restrict__ offsets, const unsigned* __restrict__ idx,
    const int* __restrict__ triples, Tables tb, const double* __restrict__ gw0, const double* __restrict__ gwdcn0,
    const double* __restrict__ W0, const double* __restrict__ W0cn, AtmParams ap, int tsp,
    double* __restrict__ e_atom, double* __restrict__ dEdcn, double* __restrict__ grad, double* __restrict__ sig_part,
    unsigned long long* __restrict__ counters, int want_grad, double* __restrict__ e3, int e3_ld) {
  extern __shared__ __align__(16) double sm2[];
  double* sm = sm2;
  double2* GA = reinterpret_cast<double2*>(sm);   // [KC] {kx, ky}
  double2* GB = GA + A2_KC;                       // [KC] {kz, r2ik}
  double2* GC = GB + A2_KC;                       // [KC] {c6ik, cbrt(R0ik)}
  double4* EV = reinterpret_cast<double4*>(GC + A2_KC);  // [EPP] {vjx, vjy, vjz, r2ij}
  double* ECR = reinterpret_cast<double*>(EV + A2_EPP);  // [EPP] cbrt(R0ij)
  double* ESE = ECR + A2_EPP;                     // [EPP] sum_k ang * fdmp * sqrt(C)   (energy, without -s9)
  double* ESG = ESE + A2_EPP;                     // [EPP] sum_k radial-force term
  double* EG = ESG + A2_EPP;                      // [EPP][7] c6ij * gw0_j
  double* UI = EG + (size_t)A2_EPP * D4_MREF;     // [tsp] W0 row of the centre
  double* UIcn = UI + tsp;                        // [tsp]
  double* KW = UIcn + tsp;                        // [KC][tsp]
  double* LR = KW + (size_t)A2_KC * tsp;          // [WARPS][KC] r2jk of the listed third atoms
  int* katom = reinterpret_cast<int*>(LR + (size_t)A2_WARPS * A2_KC);
  int* ktr = katom + A2_KC;
  int* kflag = ktr + A2_KC;
  int* EJ = kflag + A2_KC;                        // [EPP] atom index of j'
  int* ETJ = EJ + A2_EPP;                         // [EPP] translation index of j'
  int* ESO = ETJ + A2_EPP;                        // [EPP] soff[species(j')]
  int* ENF = ESO + A2_EPP;                        // [EPP] nref_j | near_ij << 8 | vinc << 16
  int* LK = ENF + A2_EPP;                         // [WARPS][KC] chunk positions of the listed third atoms
  unsigned* ebuf = reinterpret_cast<unsigned*>(LK + (size_t)A2_WARPS * A2_KC);  // [2*THREADS]
  __shared__ int s_wcnt[A2_WARPS];
  __shared__ double s_acc[A2_WARPS][12];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  const int r = nrows - 1 - blockIdx.x;  // rows ascend in atom index: heaviest centres (most owned edges) first
  const int i = rows[r];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi_ = xyz[3 * i + 2];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  const double near_tol = ap.cutoff2 * 1.0e-10;
  const double cut2 = ap.cutoff2, eps2 = ap.eps2;
  double* lr = LR + (size_t)warp * A2_KC;
  int* lk = LK + (size_t)warp * A2_KC;

  for (int e = tid; e < tsp; e += A2_THREADS) { UI[e] = W0[(size_t)i * tsp + e]; UIcn[e] = W0cn[(size_t)i * tsp + e]; }
  if (lane < 12) s_acc[warp][lane] = 0.0;
  __syncthreads();

  unsigned long long visits = 0;
  int pos = 0, cnt = 0;  // next list position to scan, owned edges waiting in ebuf (both uniform across the CTA)
  while (true) {
    // ---- refill the edge buffer: owned = partner index below the centre's, or a self image ----
    while (cnt < A2_EPP && pos < n) {
      const int c = pos + tid;
      bool own = false;
      if (c < n) own = point_atom(pts[idx[beg + c]].w) <= i;
      const unsigned m = __ballot_sync(D4_FULL, own);
      if (lane == 0) s_wcnt[warp] = __popc(m);
      __syncthreads();
      int base = cnt, tot = 0;
      for (int w = 0; w < A2_WARPS; ++w) { if (w < warp) base += s_wcnt[w]; tot += s_wcnt[w]; }
      if (own) ebuf[base + __popc(m & lt_mask)] = (unsigned)c;
      cnt += tot;
      pos += A2_THREADS;
      __syncthreads();
    }
    if (cnt == 0) break;
    const int ne = min(cnt, A2_EPP);

    // ---- edge table of this pass: one thread per edge ----
    if (tid < ne) {
      const Point p = load_point(pts + idx[beg + ebuf[tid]]);
      const double vjx = __dsub_rn(p.x, xi), vjy = __dsub_rn(p.y, yi), vjz = __dsub_rn(p.z, zi_);
      const double r2ij = exact_r2(vjx, vjy, vjz);
      const int j = point_atom(p.w), zj = point_species(p.w);
      const int so_j = tb.soff[zj], nrj = tb.nref[zj];
      double g[D4_MREF], c6ij = 0.0;
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b) {
        g[b] = b < nrj ? gw0[(size_t)j * D4_MREF + b] : 0.0;
        if (b < nrj) c6ij += UI[so_j + b] * g[b];
      }
#pragma unroll
      for (int b = 0; b < D4_MREF; ++b) EG[tid * D4_MREF + b] = g[b] * c6ij;
      EV[tid] = make_double4(vjx, vjy, vjz, r2ij);
      ECR[tid] = UI[so_j + nrj];
      ESE[tid] = 0.0; ESG[tid] = 0.0;
      EJ[tid] = j; ETJ[tid] = point_trans(p.w); ESO[tid] = so_j;
      ENF[tid] = nrj | (fabs(r2ij - cut2) <= near_tol ? 0x100 : 0) | ((j == i ? 1 : 2) << 16);
    }

    // ---- stream all neighbours k' of the centre in chunks ----
    for (int c0 = 0; c0 < n; c0 += A2_KC) {
      const int kc = min(A2_KC, n - c0);
      __syncthreads();  // previous chunk fully consumed (and the edge table written, first chunk)
      if (tid < kc) {
        const Point p = load_point(pts + idx[beg + c0 + tid]);
        const double vx = __dsub_rn(p.x, xi), vy = __dsub_rn(p.y, yi), vz = __dsub_rn(p.z, zi_);
        const double r2 = exact_r2(vx, vy, vz);
        const int k = point_atom(p.w), zk = point_species(p.w);
        const int so = tb.soff[zk], nrk = tb.nref[zk];
        double c6 = 0.0;
        for (int b = 0; b < nrk; ++b) c6 += UI[so + b] * gw0[(size_t)k * D4_MREF + b];
        GA[tid] = make_double2(vx, vy); GB[tid] = make_double2(vz, r2); GC[tid] = make_double2(c6, UI[so + nrk]);
        katom[tid] = k; ktr[tid] = point_trans(p.w);
        kflag[tid] = (k > i ? 1 : 0) | (fabs(r2 - cut2) <= near_tol ? 2 : 0);
      }
      for (int e = tid; e < kc * tsp; e += A2_THREADS) {
        const int kk = e / tsp, o = e - kk * tsp;
        KW[e] = W0[(size_t)point_atom(pts[idx[beg + c0 + kk]].w) * tsp + o];
      }
      __syncthreads();

      for (int e = warp; e < ne; e += A2_WARPS) {
        // warp-uniform edge data (broadcast loads)
        const double4 ev = EV[e];
        const int enf = ENF[e];
        const int nrj = enf & 0xff, so_j = ESO[e];
        const bool near_ij = (enf & 0x100) != 0;
        const unsigned vinc = (unsigned)(enf >> 16);

        // ---- test phase: lanes over the chunk's third atoms, accepted ones compacted into the warp's list ----
        int nl = 0;
        for (int k0 = 0; k0 < kc; k0 += 32) {
          const int kk = k0 + lane;
          bool cand = false;
          double r2jk = 0.0;
          if (kk < kc) {
            const double2 ga = GA[kk], gb = GB[kk];
            r2jk = exact_r2(__dsub_rn(ga.x, ev.x), __dsub_rn(ga.y, ev.y), __dsub_rn(gb.x, ev.z));
            cand = !(r2jk > cut2 || r2jk < eps2);
            const int kf = kflag[kk];
            // third atom above the centre and a distance on the cutoff: the reference's frame decides (heavy phase)
            if ((kf & 1) && ((kf & 2) || near_ij || fabs(r2jk - cut2) <= near_tol)) cand = true;
          }
          const unsigned m = __ballot_sync(D4_FULL, cand);
          if (cand) {
            const int p = nl + __popc(m & lt_mask);
            lk[p] = kk; lr[p] = r2jk;
          }
          nl += __popc(m);
        }
        __syncwarp();

        // ---- heavy phase: full warps over the list ----
        if (nl > 0) {
          const double r2ij = ev.w, cr0ij = ECR[e];
          double gjs[D4_MREF];
#pragma unroll
          for (int b = 0; b < D4_MREF; ++b) gjs[b] = b < nrj ? EG[e * D4_MREF + b] : 0.0;
          double se = 0.0, sg = 0.0;
          unsigned vis = 0;
          for (int b0 = 0; b0 < nl; b0 += 32) {
            const int p = min(b0 + lane, nl - 1);
            const bool act = b0 + lane < nl;
            const int kk = lk[p];
            const double r2jk = lr[p];
            bool ok = !(r2jk > cut2 || r2jk < eps2);
            const int kf = kflag[kk];
            if ((kf & 1) && ((kf & 2) || near_ij || fabs(r2jk - cut2) <= near_tol))
              ok = atm_frame_check(i, EJ[e], ETJ[e], katom[kk], ktr[kk], xyz, triples, ap);
            ok = ok && act;
            const double2 gc = GC[kk];
            const double r2ik = GB[kk].y;
            const double* Wk = KW + (size_t)kk * tsp + so_j;
            double c6p = 0.0;  // c6ij * c6jk
#pragma unroll
            for (int b = 0; b < D4_MREF; ++b)
              if (b < nrj) c6p = fma(gjs[b], Wk[b], c6p);
            const double C = fabs(c6p * gc.x);          // c6ij c6ik c6jk
            const double c0 = cr0ij * gc.y * Wk[nrj];   // (R0ij R0ik R0jk)^(1/3)
            const double r2 = r2ij * r2ik * r2jk;
            const double pq = atm_inv_sixth_root(r2);   // r2^(-1/6)
            const double p2 = pq * pq, u = p2 * pq, u2 = u * u, u3 = u2 * u, u5 = u3 * u2;   // u = 1/rbar
            const double v = c0 * pq;                   // (R0 / rbar)^(1/3)
            double tmp;
            if (ALP16) {
              const double v2 = v * v, v4 = v2 * v2, v8 = v4 * v4;
              tmp = v8 * v8;
            } else {
              tmp = pow(v, ap.alp);
            }
            const double d = fma(6.0, tmp, 1.0);        // 1 / fdmp
            const double q = atm_rsqrt(C * d * d);      // sqrt(C) fdmp = C q ;  fdmp = C d q^2
            const double cf = C * q;
            const double bmc = r2jk - r2ik, bpc = r2jk + r2ik;
            const double ang = fma(0.375 * ((r2ij + bmc) * (r2ij - bmc) * (bpc - r2ij)), u5, u3);
            const double e0 = ang * cf;
            double gg = 0.0;
            if (want_grad) {
              const double tf = tmp * (cf * d * q);     // tmp * fdmp
              const double m2 = bmc * bmc;
              const double poly = r2ij * fma(r2ij, r2ij + bpc, fma(2.0, bpc * bpc, m2)) - 5.0 * (m2 * bpc);
              const double dang = -0.375 * poly * u5;
              gg = cf * (-dang - 2.0 * ap.alp * ang * tf);
            }
            if (ok) { se += e0; sg += gg; vis += vinc; }
          }
          se = warp_sum(se); sg = warp_sum(sg);
          if (lane == 0) { ESE[e] += se; ESG[e] += sg; }  // edge e belongs to this warp for the whole pass
          visits += vis;
        }
        __syncwarp();  // the list is rewritten by the next edge
      }
    }
    __syncthreads();

    // ---- apply the per-edge sums once ----
    double out[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // e, dcn_i, g[3], s[6] of the centre
    if (tid < ne) {
      const double4 ev = EV[tid];
      const int j = EJ[tid], so_j = ESO[tid], nrj = ENF[tid] & 0xff;
      const double wedge = (j == i ? 0.5 : 1.0) * -ap.s9;
      const double se = ESE[tid] * wedge, sg = ESG[tid] * wedge;
      out[0] = -se * (1.0 / 3.0);
      if (e3) {  // pair-resolved energies (rational.f90:487-606): a sixth of the triangle energy per directed pair
        atomicAdd(&e3[(size_t)i * e3_ld + j], -se * (1.0 / 6.0));
        atomicAdd(&e3[(size_t)j * e3_ld + i], -se * (1.0 / 6.0));
      }
      if (want_grad) {
        double c6ij = 0.0, dci = 0.0, dcj = 0.0;
        for (int b = 0; b < nrj; ++b) {
          const double g = gw0[(size_t)j * D4_MREF + b];
          c6ij += UI[so_j + b] * g;
          dci += UIcn[so_j + b] * g;
          dcj += UI[so_j + b] * gwdcn0[(size_t)j * D4_MREF + b];
        }
        const double coef = sg / ev.w;
        const double gx = coef * ev.x, gy = coef * ev.y, gz = coef * ev.z;  // dGij summed over k
        out[2] = -gx; out[3] = -gy; out[4] = -gz;
        atomicAdd(&grad[3 * j], gx); atomicAdd(&grad[3 * j + 1], gy); atomicAdd(&grad[3 * j + 2], gz);
        out[5] = gx * ev.x; out[6] = gx * ev.y; out[7] = gx * ev.z;
        out[8] = gy * ev.y; out[9] = gy * ev.z; out[10] = gz * ev.z;
        out[1] = -0.5 * se * dci / c6ij;
        atomicAdd(&dEdcn[j], -0.5 * se * dcj / c6ij);
      }
    }
    if (warp * 32 < ne) {  // warps without edges in this pass contribute zeros
#pragma unroll
      for (int c = 0; c < 11; ++c) {
        const double t = warp_sum(out[c]);
        if (lane == 0) s_acc[warp][c] += t;
      }
    }
    // ---- drop the processed edges ----
    __syncthreads();
    const int rem = cnt - ne;
    unsigned keep0 = 0, keep1 = 0;
    if (tid < rem) keep0 = ebuf[ne + tid];
    if (tid + A2_THREADS < rem) keep1 = ebuf[ne + tid + A2_THREADS];
    __syncthreads();
    if (tid < rem) ebuf[tid] = keep0;
    if (tid + A2_THREADS < rem) ebuf[tid + A2_THREADS] = keep1;
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
    for (int w = 0; w < A2_WARPS; ++w) {
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
