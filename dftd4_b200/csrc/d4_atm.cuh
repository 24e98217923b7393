// Axilrod-Teller-Muto three-body kernel (reference: src/dftd4/damping/atm.f90:235-453, C6 from model/d4.f90:381-473
// with q = 0 weights, disp.f90:120-126).
//
// Formulation (see DESIGN.md "ATM"): the reference's i>=j>=k loop with triple_scale equals one sixth of the sum over
// ordered, distinct (home atom, image, image) triples.  Every triangle is visited here once per EDGE: the CTA of
// centre atom i owns the undirected edges (i, j') selected by a parity rule, each lane of a warp holds one edge, and
// the warp streams all neighbours k' of i (shared-memory chunks).  A visit only needs two scalar accumulators per
// edge,
//     SE = sum_k' dE0                 (energy)          SG = sum_k' c9 (-dang_ij fdmp + ang dfdmp)   (radial force)
// because energy, gradient, virial and dE/dCN of a triangle decompose into per-edge terms.  Forces on the two edge
// atoms, the virial and dE/dCN are applied once per edge after the stream, so the inner loop carries no atomics.
// C6_jk is contracted on the fly from a per-lane table U_j[species][ref] (shared memory, lane-interleaved) and the
// broadcast q=0 weights of k.
#pragma once
#include "d4_common.cuh"
#include "d4_kernels.cuh"

namespace d4k {

constexpr int ATM_WARPS = 8;
constexpr int ATM_THREADS = ATM_WARPS * 32;
constexpr int ATM_KC = 256;    // neighbours per staged chunk
constexpr int ATM_EC = 2048;   // edges per pass
constexpr int ATM_KROWS = 7 + D4_MREF;  // kx ky kz kr2 kc6 kr0 ksw + 7 weights

struct AtmParams {
  double s9, a1, a2, alp, cutoff, cutoff2, width, eps2;
  int alp16;
};

__host__ __device__ inline size_t atm_smem_bytes(int ts) {
  // UI, UIcn [ts] | chunk [ATM_KROWS][KC] | UJ [warps][ts][32] | SE, SG [EC] | eidx [EC] u32 | ksoff, knref [KC] int | misc
  return sizeof(double) * ((size_t)2 * ts + (size_t)ATM_KROWS * ATM_KC + (size_t)ATM_WARPS * ts * 32 + 2 * (size_t)ATM_EC + 64) +
         sizeof(unsigned) * ATM_EC + sizeof(int) * 2 * ATM_KC;
}

__device__ __forceinline__ bool atm_owns_edge(int i, int j, const int* __restrict__ triples, int t) {
  if (i != j) return (i > j) != (((i + j) & 1) != 0);
  const int a = triples[3 * t], b = triples[3 * t + 1], c = triples[3 * t + 2];
  return a > 0 || (a == 0 && (b > 0 || (b == 0 && c > 0)));
}

template <bool SMOOTH>
__global__ void __launch_bounds__(ATM_THREADS) atm_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const Point* __restrict__ pts, const long long* __restrict__ offsets, const unsigned* __restrict__ idx,
    const int* __restrict__ triples, Tables tb, const double* __restrict__ gw0, const double* __restrict__ gwdcn0,
    AtmParams ap, int ts, double* __restrict__ e_atom, double* __restrict__ dEdcn, double* __restrict__ grad,
    double* __restrict__ sig_part, unsigned long long* __restrict__ counters, int want_grad, double* __restrict__ e3,
    int e3_ld) {
  extern __shared__ double sm[];
  double* UI = sm;                       // [ts]
  double* UIcn = UI + ts;                // [ts]
  double* KC = UIcn + ts;                // [ATM_KROWS][ATM_KC]
  double* UJ = KC + ATM_KROWS * ATM_KC;  // [warps][ts][32]
  double* SE = UJ + (size_t)ATM_WARPS * ts * 32;
  double* SG = SE + ATM_EC;
  double* red = SG + ATM_EC;             // [64] scratch
  unsigned* eidx = reinterpret_cast<unsigned*>(red + 64);
  int* ksoff = reinterpret_cast<int*>(eidx + ATM_EC);
  int* knref = ksoff + ATM_KC;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = nrows - 1 - blockIdx.x;  // heaviest (longest lists tend to sit anywhere; reverse order is harmless)
  const int i = rows[r];
  const int nid = tb.nid;
  const int zi = species[i];
  const int nri = tb.nref[zi];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi_ = xyz[3 * i + 2];
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);

  // centre tables: UI[soff[s]+b] = sum_a gw0_i[a] R[zi][s][a][b]; UI[soff[s]+nref_s] = R0(zi,s)
  for (int e = tid; e < ts; e += ATM_THREADS) {
    int s = 0;
    while (tb.soff[s + 1] <= e) ++s;
    const int b = e - tb.soff[s];
    double u = 0.0, ucn = 0.0;
    if (b < tb.nref[s]) {
      for (int a = 0; a < nri; ++a) {
        const double rc6 = tb.refc6[(((size_t)zi * nid + s) * D4_MREF + a) * D4_MREF + b];
        u += gw0[(size_t)i * D4_MREF + a] * rc6;
        ucn += gwdcn0[(size_t)i * D4_MREF + a] * rc6;
      }
    } else {
      u = tb.pair_r0[zi * nid + s];
    }
    UI[e] = u; UIcn[e] = ucn;
  }
  __syncthreads();

  double* kx = KC, *ky = KC + ATM_KC, *kz = KC + 2 * ATM_KC, *kr2 = KC + 3 * ATM_KC, *kc6 = KC + 4 * ATM_KC,
        *kr0 = KC + 5 * ATM_KC, *ksw = KC + 6 * ATM_KC, *kgw = KC + 7 * ATM_KC;
  double* UJw = UJ + (size_t)warp * ts * 32;
  const double alp3 = ap.alp / 3.0;

  double acc_e = 0.0, acc_dcn = 0.0, acc_g[3] = {0, 0, 0}, acc_s[6] = {0, 0, 0, 0, 0, 0};
  unsigned long long visits = 0;

  int pos = 0;  // next list position to scan for owned edges
  while (pos < n) {
    // ---- collect up to ATM_EC owned edges (list order preserved) ----
    __shared__ int s_ne, s_pos, s_wcnt[ATM_WARPS];
    if (tid == 0) { s_ne = 0; }
    __syncthreads();
    int ne = 0;
    while (pos < n && ne + ATM_THREADS <= ATM_EC) {
      const int c = pos + tid;
      bool own = false;
      if (c < n) {
        const Point p = load_point(pts + idx[beg + c]);
        own = atm_owns_edge(i, point_atom(p.w), triples, point_trans(p.w));
      }
      const unsigned m = __ballot_sync(D4_FULL, own);
      if (lane == 0) s_wcnt[warp] = __popc(m);
      __syncthreads();
      int base = ne;
      for (int w = 0; w < warp; ++w) base += s_wcnt[w];
      if (own) eidx[base + __popc(m & ((1u << lane) - 1u))] = (unsigned)c;
      int tot = 0;
      for (int w = 0; w < ATM_WARPS; ++w) tot += s_wcnt[w];
      ne += tot;
      pos += ATM_THREADS;
      __syncthreads();
    }
    if (pos > n) pos = n;
    (void)s_ne; (void)s_pos;
    for (int e = tid; e < ne; e += ATM_THREADS) { SE[e] = 0.0; SG[e] = 0.0; }
    __syncthreads();
    if (ne == 0) continue;

    // ---- stream all neighbours k' of the centre in chunks ----
    for (int c0 = 0; c0 < n; c0 += ATM_KC) {
      const int kc = min(ATM_KC, n - c0);
      for (int kk = tid; kk < kc; kk += ATM_THREADS) {
        const Point p = load_point(pts + idx[beg + c0 + kk]);
        const double vx = __dsub_rn(p.x, xi), vy = __dsub_rn(p.y, yi), vz = __dsub_rn(p.z, zi_);
        const double r2 = exact_r2(vx, vy, vz);
        const int k = point_atom(p.w), zk = point_species(p.w);
        const int so = tb.soff[zk], nrk = tb.nref[zk];
        double c6 = 0.0;
        for (int b = 0; b < D4_MREF; ++b) {
          const double g = b < nrk ? gw0[(size_t)k * D4_MREF + b] : 0.0;
          kgw[b * ATM_KC + kk] = g;
          if (b < nrk) c6 += UI[so + b] * g;
        }
        kx[kk] = vx; ky[kk] = vy; kz[kk] = vz; kr2[kk] = r2; kc6[kk] = c6; kr0[kk] = UI[so + nrk];
        ksoff[kk] = so; knref[kk] = nrk;
        if (SMOOTH) {
          double sw, dsw;
          smooth_cutoff(sqrt(r2), ap.cutoff, ap.width, sw, dsw);
          ksw[kk] = sw;
        }
      }
      __syncthreads();

      for (int g0 = warp * 32; g0 < ne; g0 += ATM_THREADS) {
        const int e = g0 + lane;
        const bool valid = e < ne;
        // edge constants (recomputed per chunk: ~2% of the visit work, saves 100 KB of shared memory)
        double vjx = 0, vjy = 0, vjz = 0, r2ij = 1.0, c6ij = 0.0, r0ij = 1.0, swij = 1.0;
        {
          const Point p = load_point(pts + idx[beg + (valid ? eidx[e] : 0u)]);
          vjx = __dsub_rn(p.x, xi); vjy = __dsub_rn(p.y, yi); vjz = __dsub_rn(p.z, zi_);
          r2ij = exact_r2(vjx, vjy, vjz);
          const int j = point_atom(p.w), zj = point_species(p.w);
          const int so = tb.soff[zj], nrj = tb.nref[zj];
          double gj[D4_MREF];
#pragma unroll
          for (int b = 0; b < D4_MREF; ++b) gj[b] = b < nrj ? gw0[(size_t)j * D4_MREF + b] : 0.0;
          for (int b = 0; b < nrj; ++b) c6ij += UI[so + b] * gj[b];
          r0ij = UI[so + nrj];
          // U_j table, lane-interleaved: UJw[o*32 + lane]
          for (int s = 0; s < nid; ++s) {
            const int sb = tb.soff[s], nrs = tb.nref[s];
            const double* R = tb.refc6 + ((size_t)zj * nid + s) * D4_MREF * D4_MREF;
            for (int b = 0; b < nrs; ++b) {
              double u = 0.0;
#pragma unroll
              for (int a = 0; a < D4_MREF; ++a) u += gj[a] * R[a * D4_MREF + b];
              UJw[(sb + b) * 32 + lane] = u;
            }
            UJw[(sb + nrs) * 32 + lane] = tb.pair_r0[zj * nid + s];
          }
          if (SMOOTH) {
            double dsw;
            smooth_cutoff(sqrt(r2ij), ap.cutoff, ap.width, swij, dsw);
          }
        }
        __syncwarp();
        double se = 0.0, sg = 0.0;
        for (int kk = 0; kk < kc; ++kk) {
          const double dx = __dsub_rn(kx[kk], vjx), dy = __dsub_rn(ky[kk], vjy), dz = __dsub_rn(kz[kk], vjz);
          const double r2jk = exact_r2(dx, dy, dz);
          bool ok = valid && !(r2jk > ap.cutoff2 || r2jk < ap.eps2);
          double swjk = 1.0;
          if (SMOOTH) {
            double dsw;
            smooth_cutoff(sqrt(r2jk), ap.cutoff, ap.width, swjk, dsw);
            ok = ok && (swij * ksw[kk] * swjk > 0.0);
          }
          if (!__any_sync(D4_FULL, ok)) continue;
          const int so = ksoff[kk], nrk = knref[kk];
          double c6jk = 0.0;
          for (int b = 0; b < nrk; ++b) c6jk += UJw[(so + b) * 32 + lane] * kgw[b * ATM_KC + kk];
          const double r0jk = UJw[(so + nrk) * 32 + lane];
          const double r2ik = kr2[kk];
          const double c9 = -ap.s9 * sqrt(fabs(c6ij * kc6[kk] * c6jk));
          const double r0 = r0ij * kr0[kk] * r0jk;
          const double r2 = r2ij * r2ik * r2jk;
          const double r1 = sqrt(r2), r3 = r2 * r1, r5 = r3 * r2;
          double tmp;
          if (ap.alp16) {
            const double y = r0 / r1, y2 = y * y;
            tmp = y2 * y2 * y * cbrt(y);
          } else {
            tmp = pow(r0 / r1, alp3);
          }
          const double fdmp = 1.0 / (1.0 + 6.0 * tmp);
          const double ang = 0.375 * (r2ij + r2jk - r2ik) * (r2ij - r2jk + r2ik) * (-r2ij + r2jk + r2ik) / r5 + 1.0 / r3;
          const double dfdmp = -2.0 * ap.alp * tmp * fdmp * fdmp;
          const double dang = -0.375 * (r2ij * r2ij * r2ij + r2ij * r2ij * (r2jk + r2ik) +
                                        r2ij * (3.0 * r2jk * r2jk + 2.0 * r2jk * r2ik + 3.0 * r2ik * r2ik) -
                                        5.0 * (r2jk - r2ik) * (r2jk - r2ik) * (r2jk + r2ik)) / r5;
          double e0 = ang * fdmp * c9;
          double gg = c9 * (-dang * fdmp + ang * dfdmp);
          if (SMOOTH) { const double f = ksw[kk] * swjk; e0 *= f; gg *= f; }
          if (ok) { se += e0; sg += gg; ++visits; }
        }
        if (valid) { SE[e] += se; SG[e] += sg; }
        __syncwarp();
      }
      __syncthreads();
    }

    // ---- apply the per-edge sums once ----
    for (int e = tid; e < ne; e += ATM_THREADS) {
      const Point p = load_point(pts + idx[beg + eidx[e]]);
      const double vx = __dsub_rn(p.x, xi), vy = __dsub_rn(p.y, yi), vz = __dsub_rn(p.z, zi_);
      const double r2ij = exact_r2(vx, vy, vz);
      const int j = point_atom(p.w), zj = point_species(p.w);
      const int so = tb.soff[zj], nrj = tb.nref[zj];
      double swij = 1.0, dswij = 0.0;
      const double rij = sqrt(r2ij);
      if (SMOOTH) smooth_cutoff(rij, ap.cutoff, ap.width, swij, dswij);
      const double eedge = swij * SE[e];       // sum over k of dE0*sw
      acc_e -= eedge * (1.0 / 3.0);
      if (e3) {  // pair-resolved energies (rational.f90:487-606): a sixth of the triangle energy per directed pair
        atomicAdd(&e3[(size_t)i * e3_ld + j], -eedge * (1.0 / 6.0));
        atomicAdd(&e3[(size_t)j * e3_ld + i], -eedge * (1.0 / 6.0));
      }
      if (want_grad) {
        double c6ij = 0.0, dci = 0.0, dcj = 0.0;
        for (int b = 0; b < nrj; ++b) {
          const double g = gw0[(size_t)j * D4_MREF + b];
          c6ij += UI[so + b] * g;
          dci += UIcn[so + b] * g;
          dcj += UI[so + b] * gwdcn0[(size_t)j * D4_MREF + b];
        }
        double coef = swij * SG[e] / r2ij;
        if (SMOOTH) coef -= SE[e] * dswij / rij;
        const double gx = coef * vx, gy = coef * vy, gz = coef * vz;  // dGij summed over k
        acc_g[0] -= gx; acc_g[1] -= gy; acc_g[2] -= gz;
        atomicAdd(&grad[3 * j], gx); atomicAdd(&grad[3 * j + 1], gy); atomicAdd(&grad[3 * j + 2], gz);
        acc_s[0] += gx * vx; acc_s[1] += gx * vy; acc_s[2] += gx * vz;
        acc_s[3] += gy * vy; acc_s[4] += gy * vz; acc_s[5] += gz * vz;
        acc_dcn -= 0.5 * eedge * dci / c6ij;
        atomicAdd(&dEdcn[j], -0.5 * eedge * dcj / c6ij);
      }
    }
    __syncthreads();
  }

  // ---- CTA reduction of the centre's own quantities ----
  double vals[11] = {acc_e, acc_dcn, acc_g[0], acc_g[1], acc_g[2], acc_s[0], acc_s[1], acc_s[2], acc_s[3], acc_s[4], acc_s[5]};
  __shared__ double s_red[ATM_WARPS][12];
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
    e_atom[i] += t[0];
    atomicAdd(&counters[4], 2ull * vt);
    if (want_grad) {
      atomicAdd(&dEdcn[i], t[1]);
      atomicAdd(&grad[3 * i], t[2]); atomicAdd(&grad[3 * i + 1], t[3]); atomicAdd(&grad[3 * i + 2], t[4]);
      double* sp = sig_part + (size_t)i * 6;
      for (int c = 0; c < 6; ++c) sp[c] += t[5 + c];
    }
  }
}

}  // namespace d4k
