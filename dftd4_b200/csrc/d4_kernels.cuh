// Neighbour lists over lattice images, coordination numbers, reference weights, the rational-damping pair kernel and
// the CN chain-rule kernel.  All "directed": every atom walks its own full list and accumulates only its own
// quantities, so the kernels are atomics-free and bitwise reproducible.
#pragma once
#include "d4_common.cuh"

namespace d4k {

// Species-level tables living in global memory (tiny, L1/L2 resident)
struct Tables {
  int nid;
  double ga, gc, wf;
  const double *zeff, *eta, *refcn, *refq, *refc6, *refalpha0;  // [nid], [nid], [nid][7], [nid][7], [nid][nid][7][7], [nid][7]
  const int *nref, *ngw;                                        // [nid], [nid][7]
  const double *chi, *eeta, *kcn, *rad;                         // EEQ-2019 [nid]
  const double *pair_rc;    // [nid][nid] rcov_i + rcov_j
  const double *pair_den;   // [nid][nid] EN factor of the D4 coordination number
  const double *pair_rr;    // [nid][nid] 3 r4r2_i r4r2_j                         (rational.f90:283)
  const double *pair_r0;    // [nid][nid] a1 sqrt(rr) + a2                        (rational.f90:284)
  const double *pair_gam;   // [nid][nid] 1/sqrt(rad_i^2 + rad_j^2)               (EEQ)
  const int *soff;          // [nid+1] offsets of the packed per-species rows used by the ATM kernel
  // D4S model (src/dftd4/model/d4s.f90): the reference's gwvec(iref, iat, jat) depends on jat only through its
  // species, so every atom carries nsp = nid weight sets (one per partner species); D4: nsp = 1, wfpair = nullptr.
  int nsp;
  const double *wfpair;     // [nid][nid] pair-specific Gaussian weighting factor (D4S), else nullptr
};
// Offset of the 7 weights atom a uses against a partner of species sp.
__device__ __forceinline__ size_t gw_row(const Tables& tb, int a, int sp) {
  return ((size_t)a * tb.nsp + (tb.nsp > 1 ? sp : 0)) * D4_MREF;
}

// ---------------------------------------------------------------------------------------------------------------
// neighbour lists
// ---------------------------------------------------------------------------------------------------------------
// Sort key = cell index, refined by `fine` Morton bits per dimension inside the cell (fine = 0: plain cell order).  The
// refinement only orders the points of a cell along a Z-curve, so that consecutive entries of a neighbour row form
// compact patches (the ATM stream takes 32 consecutive entries as the edges of one warp).
__global__ void point_keys_kernel(int nat, long long npts, const double* __restrict__ xyz, const double* __restrict__ trans,
                                  Grid g, int fine, unsigned* __restrict__ keys, unsigned* __restrict__ pids) {
  const long long pid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pid >= npts) return;
  const int t = (int)(pid / nat), j = (int)(pid - (long long)t * nat);
  const double x = __dadd_rn(xyz[3 * j], trans[3 * t]);
  const double y = __dadd_rn(xyz[3 * j + 1], trans[3 * t + 1]);
  const double z = __dadd_rn(xyz[3 * j + 2], trans[3 * t + 2]);
  const bool in = x >= g.lox && x <= g.hix && y >= g.loy && y <= g.hiy && z >= g.loz && z <= g.hiz;
  unsigned key = (unsigned)(g.nx * g.ny * g.nz) << (3 * fine);
  if (in) {
    const int cx = grid_coord(x, g.ox, g.inv_h, g.nx), cy = grid_coord(y, g.oy, g.inv_h, g.ny),
              cz = grid_coord(z, g.oz, g.inv_h, g.nz);
    key = (unsigned)((cz * g.ny + cy) * g.nx + cx) << (3 * fine);
    if (fine > 0) {
      const int nsub = 1 << fine;
      const int sx = min(nsub - 1, max(0, (int)(((x - g.ox) * g.inv_h - cx) * nsub)));
      const int sy = min(nsub - 1, max(0, (int)(((y - g.oy) * g.inv_h - cy) * nsub)));
      const int sz = min(nsub - 1, max(0, (int)(((z - g.oz) * g.inv_h - cz) * nsub)));
      unsigned m = 0;
      for (int b = 0; b < fine; ++b)
        m |= (unsigned)((sx >> b) & 1) << (3 * b) | (unsigned)((sy >> b) & 1) << (3 * b + 1) | (unsigned)((sz >> b) & 1) << (3 * b + 2);
      key |= m;
    }
  }
  keys[pid] = key;
  pids[pid] = (unsigned)pid;
}

// cell_start[c] = first sorted position with key >= c, c = 0..ncell (cell_start[ncell] = number of valid points)
__global__ void cell_bounds_kernel(const unsigned* __restrict__ keys, long long npts, int ncell, int fine,
                                   unsigned* __restrict__ cell_start) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > ncell) return;
  const unsigned target = (unsigned)c << (3 * fine);
  long long lo = 0, hi = npts;
  while (lo < hi) {
    const long long mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  cell_start[c] = (unsigned)lo;
}

__global__ void make_points_kernel(int nat, long long npts, const double* __restrict__ xyz, const double* __restrict__ trans,
                                   const int* __restrict__ species, const unsigned* __restrict__ keys,
                                   const unsigned* __restrict__ pids, unsigned ncell, Point* __restrict__ pts) {
  const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npts || keys[k] >= ncell) return;
  const unsigned pid = pids[k];
  const int t = (int)(pid / (unsigned)nat), j = (int)(pid - (unsigned)t * (unsigned)nat);
  double4 v;
  v.x = __dadd_rn(xyz[3 * j], trans[3 * t]);
  v.y = __dadd_rn(xyz[3 * j + 1], trans[3 * t + 1]);
  v.z = __dadd_rn(xyz[3 * j + 2], trans[3 * t + 2]);
  v.w = __longlong_as_double((long long)pack_point(j, species[j], t));
  *reinterpret_cast<double4*>(pts + k) = v;
}

// One warp per row (= home atom rows[r]).  FILL=false: counts[r]; FILL=true: idx[offsets[r] ...] = sorted point index.
// Accept iff !(r2 > rc2 || r2 < eps2) with the reference's rounding (rational.f90:290-292, atm.f90:332-334).
template <bool FILL>
__global__ void __launch_bounds__(256) nlist_kernel(int nrows, const int* __restrict__ rows, const double* __restrict__ xyz,
                                                    const Point* __restrict__ pts, const unsigned* __restrict__ cell_start,
                                                    Grid g, double rc2, double eps2, const double* __restrict__ trans,
                                                    int* __restrict__ counts,
                                                    const long long* __restrict__ offsets, unsigned* __restrict__ idx) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= nrows) return;
  const int i = rows[r];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  const int cx = grid_coord(xi, g.ox, g.inv_h, g.nx), cy = grid_coord(yi, g.oy, g.inv_h, g.ny),
            cz = grid_coord(zi, g.oz, g.inv_h, g.nz);
  const int x0 = max(cx - g.reach, 0), x1 = min(cx + g.reach, g.nx - 1);
  const double rc = sqrt(rc2) * (1.0 + 1e-12) + 1e-9;
  int n = 0;
  const long long base = FILL ? offsets[r] : 0;
  for (int z = max(cz - g.reach, 0); z <= min(cz + g.reach, g.nz - 1); ++z) {
    // distance from the atom to the z-slab / y-row of cells: prune rows of cells that cannot reach
    const double zl = g.oz + z * g.h, zh = zl + g.h;
    const double dz = zi < zl ? zl - zi : (zi > zh ? zi - zh : 0.0);
    if (dz > rc) continue;
    for (int y = max(cy - g.reach, 0); y <= min(cy + g.reach, g.ny - 1); ++y) {
      const double yl = g.oy + y * g.h, yh = yl + g.h;
      const double dy = yi < yl ? yl - yi : (yi > yh ? yi - yh : 0.0);
      if (dy * dy + dz * dz > rc * rc) continue;
      const size_t row = ((size_t)z * g.ny + y) * g.nx;
      const unsigned beg = cell_start[row + x0], end = cell_start[row + x1 + 1];
      for (unsigned k0 = beg; k0 < end; k0 += 32) {
        const unsigned k = k0 + lane;
        bool ok = false;
        if (k < end) {
          const Point p = load_point(pts + k);
          double r2 = exact_r2(__dsub_rn(xi, p.x), __dsub_rn(yi, p.y), __dsub_rn(zi, p.z));
          r2 = frame_r2(r2, rc2, i, xi, yi, zi, p, xyz, trans);
          ok = !(r2 > rc2 || r2 < eps2);
        }
        const unsigned m = __ballot_sync(D4_FULL, ok);
        if (FILL && ok) idx[base + n + __popc(m & ((1u << lane) - 1u))] = k;
        n += __popc(m);
      }
    }
  }
  if (!FILL && lane == 0) counts[r] = n;
}

// Stable reorder of every row of a neighbour list into distance shells around the row atom (16 shells of cutoff/16),
// keeping the cell order inside a shell.  The ATM kernel takes 256 consecutive entries as the edges of one pass: edges
// of one shell accept the same fraction of third atoms, so the warps of a pass finish together, while 32 consecutive
// entries of a shell still form a compact patch (whole-row skips stay effective).  Rows longer than NSORT_CAP are left
// in cell order (only costs balance).  One CTA per row; warp w compacts shells w and w + 8 in list order.
constexpr int NSORT_CAP = 8192;
constexpr int NSORT_BINS = 16;  // maximum; the launch picks nbins <= NSORT_BINS
__global__ void __launch_bounds__(256) row_shell_sort_kernel(int nrows, const int* __restrict__ rows,
                                                             const double* __restrict__ xyz, const Point* __restrict__ pts,
                                                             const long long* __restrict__ offsets, unsigned* __restrict__ idx,
                                                             double cutoff, int nbins) {
  __shared__ unsigned vals[NSORT_CAP];
  __shared__ unsigned char bins[NSORT_CAP];
  __shared__ int hist[NSORT_BINS + 1];
  const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long beg = offsets[r];
  const int n = (int)(offsets[r + 1] - beg);
  if (n > NSORT_CAP || n < 2) return;
  const int i = rows[r];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  if (tid <= NSORT_BINS) hist[tid] = 0;
  __syncthreads();
  const double scale = nbins / cutoff;
  for (int e = tid; e < n; e += 256) {
    const unsigned v = idx[beg + e];
    const Point p = load_point(pts + v);
    const double dx = p.x - xi, dy = p.y - yi, dz = p.z - zi;
    const int b = min(nbins - 1, (int)(sqrt(dx * dx + dy * dy + dz * dz) * scale));
    vals[e] = v;
    bins[e] = (unsigned char)b;
    atomicAdd(&hist[b + 1], 1);
  }
  __syncthreads();
  if (tid == 0) for (int b = 0; b < nbins; ++b) hist[b + 1] += hist[b];
  __syncthreads();
  for (int b = warp; b < nbins; b += 8) {
    int out = hist[b];
    if (hist[b + 1] == out) continue;
    for (int e0 = 0; e0 < n; e0 += 32) {
      const int e = e0 + lane;
      const bool hit = e < n && bins[e] == b;
      const unsigned m = __ballot_sync(D4_FULL, hit);
      if (hit) idx[beg + out + __popc(m & ((1u << lane) - 1u))] = vals[e];
      out += __popc(m);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// coordination numbers: D4 flavour (erf count x EN factor, cutoff cn) and EEQ flavour (erf count, 25 bohr) in one pass
// over the CN list.  mctc-lib mctc_ncoord (absent from the reference tree); call sites src/dftd4/ncoord.f90:66-73 and
// multicharge get_charges.  One warp per atom.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cn_kernel(int nat, const double* __restrict__ xyz, const int* __restrict__ species,
                                                 const Point* __restrict__ pts, const long long* __restrict__ offsets,
                                                 const unsigned* __restrict__ idx, const double* __restrict__ trans, Tables tb,
                                                 double cn_cut2, double eeq_cut2,
                                                 double kcn, double* __restrict__ cn, double* __restrict__ cn_eeq_raw,
                                                 unsigned long long* __restrict__ counters) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= nat) return;
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  const int zs = species[i];
  double a = 0.0, b = 0.0;
  unsigned long long na = 0, nb = 0;
  for (long long e = offsets[i] + lane; e < offsets[i + 1]; e += 32) {
    const Point p = load_point(pts + idx[e]);
    const double vx = __dsub_rn(xi, p.x), vy = __dsub_rn(yi, p.y), vz = __dsub_rn(zi, p.z);
    const double r2 = exact_r2(vx, vy, vz);
    const int js = point_species(p.w);
    const double r = sqrt(r2);
    const double rc = tb.pair_rc[zs * tb.nid + js];
    const double cnt = 0.5 * (1.0 + erf(-kcn * (r - rc) / rc));
    if (!(frame_r2(r2, cn_cut2, i, xi, yi, zi, p, xyz, trans) > cn_cut2)) { a += tb.pair_den[zs * tb.nid + js] * cnt; ++na; }
    if (!(frame_r2(r2, eeq_cut2, i, xi, yi, zi, p, xyz, trans) > eeq_cut2)) { b += cnt; ++nb; }
  }
  a = warp_sum(a); b = warp_sum(b);
  na = warp_sum_u64(na); nb = warp_sum_u64(nb);
  if (lane == 0) {
    cn[i] = a;
    cn_eeq_raw[i] = b;
    atomicAdd(&counters[0], na);
    atomicAdd(&counters[1], nb);
  }
}

// EEQ right-hand side (multicharge get_vrhs) with the logarithmic CN cut (mctc-lib cut_coordination_number, cn_max = 8)
// rhs layout: rhs[0*ld + i] = X_i, rhs[1*ld + i] = 1 (constraint column), padding rows = 0
__global__ void eeq_rhs_kernel(int nat, int ld, const int* __restrict__ species, Tables tb, const double* __restrict__ cn_raw,
                               double* __restrict__ rhs, double* __restrict__ dxdcn_cut) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ld) return;
  if (i >= nat) { rhs[i] = 0.0; rhs[ld + i] = 0.0; return; }
  const double cut = 8.0;
  const double raw = cn_raw[i];
  const double dcut = exp(cut) / (exp(cut) + exp(raw));
  const double c = log(1.0 + exp(cut)) - log(1.0 + exp(cut - raw));
  const int s = species[i];
  const double tmp = tb.kcn[s] / sqrt(c + 1.0e-14);
  rhs[i] = -tb.chi[s] + tmp * c;
  rhs[ld + i] = 1.0;
  dxdcn_cut[i] = 0.5 * tmp * dcut;  // dX_i/dCN'_i * dCN'_i/dCN_raw_i
}

// ---------------------------------------------------------------------------------------------------------------
// reference weights, src/dftd4/model/d4.f90:236-376.  One thread per atom.
// Outputs (row stride 7): gw/gwdcn/gwdq with the EEQ charge, gw0/gwdcn0 with q = 0 (disp.f90:120-122).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double zeta_d(double a, double c, double qref, double qmod) {
  return qmod < 0.0 ? exp(a) : exp(a * (1.0 - exp(c * (1.0 - qref / qmod))));
}
__device__ __forceinline__ double dzeta_d(double a, double c, double qref, double qmod) {
  return qmod < 0.0 ? 0.0 : -a * c * exp(c * (1.0 - qref / qmod)) * zeta_d(a, c, qref, qmod) * qref / (qmod * qmod);
}
__device__ __forceinline__ bool exceptional(double v) { return isnan(v) || isinf(v); }

// Weights of one atom (species s, coordination number cni, charge qi) with Gaussian weighting factor wf0; outputs are
// 7-vectors (zero beyond nref).
__device__ __forceinline__ void weights_atom(const Tables& tb, int s, double wf0, double cni, double qi, double* __restrict__ gw,
                                             double* __restrict__ gwdcn, double* __restrict__ gwdq, double* __restrict__ gw0,
                                             double* __restrict__ gwdcn0) {
  const int nr = tb.nref[s];
  const double zi = tb.zeff[s], gi = tb.eta[s] * tb.gc;
  const double eps_norm = 1.4916681462400413e-154;  // sqrt(tiny(1.0_wp))
  double norm = 0.0, dnorm = 0.0, maxcn = -1e300;
  for (int r = 0; r < nr; ++r) {
    const double cr = tb.refcn[s * D4_MREF + r];
    maxcn = fmax(maxcn, cr);
    for (int ig = 1; ig <= tb.ngw[s * D4_MREF + r]; ++ig) {
      const double wf = ig * wf0;
      const double gwk = exp(-wf * ((cni - cr) * (cni - cr)));
      norm += gwk;
      dnorm += 2 * wf * (cr - cni) * gwk;
    }
  }
  norm = fabs(norm) > eps_norm ? 1.0 / norm : 0.0;
  for (int r = 0; r < D4_MREF; ++r) {
    double o_gw = 0.0, o_dcn = 0.0, o_dq = 0.0, o_gw0 = 0.0, o_dcn0 = 0.0;
    if (r < nr) {
      const double cr = tb.refcn[s * D4_MREF + r];
      double expw = 0.0, expd = 0.0;
      for (int ig = 1; ig <= tb.ngw[s * D4_MREF + r]; ++ig) {
        const double wf = ig * wf0;
        const double gwk = exp(-wf * ((cni - cr) * (cni - cr)));
        expw += gwk;
        expd += 2 * wf * (cr - cni) * gwk;
      }
      double gwk = expw * norm;
      if (exceptional(gwk) || norm == 0.0) gwk = fabs(maxcn - cr) < 1e-12 ? 1.0 : 0.0;
      double dgwk = norm * (expd - expw * dnorm * norm);
      if (exceptional(dgwk) || norm == 0.0) dgwk = 0.0;
      const double qr = tb.refq[s * D4_MREF + r] + zi;
      const double zt = zeta_d(tb.ga, gi, qr, qi + zi), zt0 = zeta_d(tb.ga, gi, qr, zi);
      o_gw = gwk * zt;
      o_dq = gwk * dzeta_d(tb.ga, gi, qr, qi + zi);
      o_dcn = dgwk * zt;
      o_gw0 = gwk * zt0;
      o_dcn0 = dgwk * zt0;
    }
    gw[r] = o_gw; gwdcn[r] = o_dcn; gwdq[r] = o_dq; gw0[r] = o_gw0; gwdcn0[r] = o_dcn0;
  }
}

__global__ void weights_kernel(int nat, const int* __restrict__ species, Tables tb, const double* __restrict__ cn,
                               const double* __restrict__ q, double* __restrict__ gw, double* __restrict__ gwdcn,
                               double* __restrict__ gwdq, double* __restrict__ gw0, double* __restrict__ gwdcn0) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nat) return;
  const int s = species[i];
  for (int sp = 0; sp < tb.nsp; ++sp) {  // d4s.f90:275-340: one set per partner species
    const size_t o = ((size_t)i * tb.nsp + sp) * D4_MREF;
    const double wf0 = tb.wfpair ? tb.wfpair[s * tb.nid + sp] : tb.wf;
    weights_atom(tb, s, wf0, cn[i], q[i], gw + o, gwdcn + o, gwdq + o, gw0 + o, gwdcn0 + o);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// two-body rational (Becke-Johnson) damping: src/dftd4/damping/rational.f90:209-343 with the C6 interpolation of
// src/dftd4/model/d4.f90:381-473 evaluated on the fly (no N x N arrays).
// One warp per row i; the warp first contracts gw_i / gwdcn_i / gwdq_i with the reference C6 block of every partner
// species (U tables in shared memory), then each lane handles list entries: c6_ij = U[z_j] . gw_j  (<= 7 FMA).
// Per-atom outputs (own quantities only): energy, dE/dCN, dE/dq, gradient, and a 6-component virial partial.
// ---------------------------------------------------------------------------------------------------------------
constexpr int PAIR_WARPS = 8;
template <bool GRAD, bool SMOOTH>
__global__ void __launch_bounds__(PAIR_WARPS * 32) pair_kernel(
    int nrows, const int* __restrict__ rows, const double* __restrict__ xyz, const int* __restrict__ species,
    const Point* __restrict__ pts, const long long* __restrict__ offsets, const unsigned* __restrict__ idx, Tables tb,
    const double* __restrict__ gw, const double* __restrict__ gwdcn, const double* __restrict__ gwdq, double s6, double s8,
    double cutoff, double width, double* __restrict__ e_atom, double* __restrict__ dEdcn, double* __restrict__ dEdq,
    double* __restrict__ grad, double* __restrict__ sig_part, unsigned long long* __restrict__ counters,
    double* __restrict__ e2, int e2_ld) {
  extern __shared__ double smem[];  // per warp: U, Ucn, Uq  [nid][7]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * PAIR_WARPS + warp;
  if (r >= nrows) return;
  const int i = rows[r];
  const int nid = tb.nid;
  double* U = smem + (size_t)warp * 3 * nid * D4_MREF;
  double* Ucn = U + nid * D4_MREF;
  double* Uq = Ucn + nid * D4_MREF;
  const int zs = species[i];
  const int nri = tb.nref[zs];
  for (int e = lane; e < nid * D4_MREF; e += 32) {
    const int s = e / D4_MREF, b = e % D4_MREF;
    double u = 0.0, ucn = 0.0, uq = 0.0;
    const size_t gi = gw_row(tb, i, s);  // D4S: the centre's weights against partner species s
    for (int a = 0; a < nri; ++a) {
      const double rc6 = tb.refc6[(((size_t)zs * nid + s) * D4_MREF + a) * D4_MREF + b];
      u += gw[gi + a] * rc6;
      if (GRAD) {
        ucn += gwdcn[gi + a] * rc6;
        uq += gwdq[gi + a] * rc6;
      }
    }
    U[e] = u; Ucn[e] = ucn; Uq[e] = uq;
  }
  __syncwarp();
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  double en = 0.0, dcn = 0.0, dq = 0.0, gx = 0.0, gy = 0.0, gz = 0.0;
  double sxx = 0.0, sxy = 0.0, sxz = 0.0, syy = 0.0, syz = 0.0, szz = 0.0;
  unsigned long long cnt = 0;
  for (long long e = offsets[r] + lane; e < offsets[r + 1]; e += 32) {
    const Point p = load_point(pts + idx[e]);
    const double vx = __dsub_rn(xi, p.x), vy = __dsub_rn(yi, p.y), vz = __dsub_rn(zi, p.z);
    const double r2 = exact_r2(vx, vy, vz);
    double sw = 1.0, dsw = 0.0, rr1 = 0.0;
    if (SMOOTH) {
      rr1 = sqrt(r2);
      smooth_cutoff(rr1, cutoff, width, sw, dsw);
      if (sw <= 0.0) continue;
    }
    ++cnt;
    const int j = point_atom(p.w), js = point_species(p.w);
    const double* gj = gw + gw_row(tb, j, zs);  // D4S: the partner's weights against the centre's species
    const int nrj = tb.nref[js];
    double c6 = 0.0, c6cn = 0.0, c6q = 0.0;
    for (int b = 0; b < nrj; ++b) {
      const double g = gj[b];
      c6 += U[js * D4_MREF + b] * g;
      if (GRAD) {
        c6cn += Ucn[js * D4_MREF + b] * g;
        c6q += Uq[js * D4_MREF + b] * g;
      }
    }
    const double rrij = tb.pair_rr[zs * nid + js], r0 = tb.pair_r0[zs * nid + js];
    const double r0_2 = r0 * r0, r0_6 = r0_2 * r0_2 * r0_2, r0_8 = r0_6 * r0_2;
    const double r4 = r2 * r2, r6 = r4 * r2;
    const double t6 = 1.0 / (r6 + r0_6), t8 = 1.0 / (r6 * r2 + r0_8);
    const double edisp0 = s6 * t6 + s8 * rrij * t8;
    const double edisp = sw * edisp0;
    en += -c6 * edisp * 0.5;
    // pair-resolved energies (rational.f90:401-483): e2[i][j] receives this directed half, e2[j][i] its mirror
    if (!GRAD && e2) atomicAdd(&e2[(size_t)i * e2_ld + j], -c6 * edisp * 0.5);
    if (GRAD) {
      const double d6 = -6.0 * r4 * t6 * t6, d8 = -8.0 * r6 * t8 * t8;
      double gdisp = sw * (s6 * d6 + s8 * rrij * d8);
      if (SMOOTH) gdisp += dsw * edisp0 / rr1;
      const double f = -c6 * gdisp;
      dcn -= c6cn * edisp;
      dq -= c6q * edisp;
      // self images (j == i) carry no force in the reference (rational.f90:317-323); here +t and -t cancel exactly
      gx += f * vx; gy += f * vy; gz += f * vz;
      const double h = 0.5 * f;
      sxx += h * vx * vx; sxy += h * vx * vy; sxz += h * vx * vz;
      syy += h * vy * vy; syz += h * vy * vz; szz += h * vz * vz;
    }
  }
  en = warp_sum(en);
  cnt = warp_sum_u64(cnt);
  if (GRAD) {
    dcn = warp_sum(dcn); dq = warp_sum(dq);
    gx = warp_sum(gx); gy = warp_sum(gy); gz = warp_sum(gz);
    sxx = warp_sum(sxx); sxy = warp_sum(sxy); sxz = warp_sum(sxz);
    syy = warp_sum(syy); syz = warp_sum(syz); szz = warp_sum(szz);
  }
  if (lane == 0) {
    e_atom[i] += en;
    atomicAdd(&counters[2], cnt);
    if (GRAD) {
      dEdcn[i] += dcn;
      dEdq[i] += dq;
      grad[3 * i] += gx; grad[3 * i + 1] += gy; grad[3 * i + 2] += gz;
      double* sp = sig_part + (size_t)i * 6;
      sp[0] += sxx; sp[1] += sxy; sp[2] += sxz; sp[3] += syy; sp[4] += syz; sp[5] += szz;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// CN chain rule: mctc-lib add_coordination_number_derivs (call site src/dftd4/disp.f90:127-130) for dE/dCN of the D4
// coordination number, fused with the EEQ coordination-number term of the charge chain rule
// (weights w2_i = lambda_i dX_i/dCN'_i dCN'_i/dCN_i).  One warp per row.
//   g_i += sum_j' [ (w1_i + w1_j) den dcount_D4 + (w2_i + w2_j) dcount_EEQ ] * rij / r
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cn_deriv_kernel(int nrows, const int* __restrict__ rows, const double* __restrict__ xyz,
                                                       const int* __restrict__ species, const Point* __restrict__ pts,
                                                       const long long* __restrict__ offsets, const unsigned* __restrict__ idx,
                                                       Tables tb, double cn_cut2, double eeq_cut2, double kcn,
                                                       const double* __restrict__ w1, const double* __restrict__ w2,
                                                       const double* __restrict__ trans,
                                                       const int* __restrict__ triples, int rep3x, int rep3y, int rep3z,
                                                       double* __restrict__ grad, double* __restrict__ sig_part) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= nrows) return;
  const int i = rows[r];
  const double xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  const int zs = species[i];
  const double w1i = w1[i], w2i = w2[i];
  const double sqrtpi = 1.77245385090551602729816748334114518;
  double gx = 0.0, gy = 0.0, gz = 0.0, sxx = 0.0, sxy = 0.0, sxz = 0.0, syy = 0.0, syz = 0.0, szz = 0.0;
  for (long long e = offsets[i] + lane; e < offsets[i + 1]; e += 32) {  // the CN list is indexed by atom
    const Point p = load_point(pts + idx[e]);
    const double vx = __dsub_rn(xi, p.x), vy = __dsub_rn(yi, p.y), vz = __dsub_rn(zi, p.z);
    const double r2 = exact_r2(vx, vy, vz);
    const int j = point_atom(p.w), js = point_species(p.w);
    const double r1 = sqrt(r2);
    const double rc = tb.pair_rc[zs * tb.nid + js];
    const double ex = kcn * (r1 - rc) / rc;
    const double dc = -(kcn * exp(-ex * ex)) / (rc * sqrtpi);
    double f = 0.0;
    // disp.f90:124-128 walks the disp3 translation list with the cn cutoff: images outside that list are dropped
    bool in3 = true;
    if (triples) {
      const int t = point_trans(p.w);
      in3 = abs(triples[3 * t]) <= rep3x && abs(triples[3 * t + 1]) <= rep3y && abs(triples[3 * t + 2]) <= rep3z;
    }
    if (!(frame_r2(r2, cn_cut2, i, xi, yi, zi, p, xyz, trans) > cn_cut2) && in3) f += (w1i + w1[j]) * tb.pair_den[zs * tb.nid + js] * dc;
    if (!(frame_r2(r2, eeq_cut2, i, xi, yi, zi, p, xyz, trans) > eeq_cut2)) f += (w2i + w2[j]) * dc;
    f /= r1;
    gx += f * vx; gy += f * vy; gz += f * vz;
    const double h = 0.5 * f;
    sxx += h * vx * vx; sxy += h * vx * vy; sxz += h * vx * vz;
    syy += h * vy * vy; syz += h * vy * vz; szz += h * vz * vz;
  }
  gx = warp_sum(gx); gy = warp_sum(gy); gz = warp_sum(gz);
  sxx = warp_sum(sxx); sxy = warp_sum(sxy); sxz = warp_sum(sxz);
  syy = warp_sum(syy); syz = warp_sum(syz); szz = warp_sum(szz);
  if (lane == 0) {
    grad[3 * i] += gx; grad[3 * i + 1] += gy; grad[3 * i + 2] += gz;
    double* sp = sig_part + (size_t)i * 6;
    sp[0] += sxx; sp[1] += sxy; sp[2] += sxz; sp[3] += syy; sp[4] += syz; sp[5] += szz;
  }
}

// deterministic final reduction: out[0] = sum e_atom, out[1..9] = sigma (full 3x3 from the 6 partial components)
__global__ void __launch_bounds__(1024) final_reduce_kernel(int nat, const double* __restrict__ e_atom,
                                                            const double* __restrict__ sig_part, const double* __restrict__ sig_extra,
                                                            double* __restrict__ out) {
  __shared__ double sh[32][7];
  double acc[7] = {0, 0, 0, 0, 0, 0, 0};
  for (int i = threadIdx.x; i < nat; i += blockDim.x) {
    acc[0] += e_atom[i];
    for (int c = 0; c < 6; ++c) acc[1 + c] += sig_part[(size_t)i * 6 + c];
  }
  for (int c = 0; c < 7; ++c) acc[c] = warp_sum(acc[c]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) for (int c = 0; c < 7; ++c) sh[warp][c] = acc[c];
  __syncthreads();
  if (warp == 0) {
    for (int c = 0; c < 7; ++c) {
      double v = lane < (blockDim.x >> 5) ? sh[lane][c] : 0.0;
      v = warp_sum(v);
      acc[c] = v;
    }
    if (lane == 0) {
      out[0] = acc[0];
      const double s[9] = {acc[1], acc[2], acc[3], acc[2], acc[4], acc[5], acc[3], acc[5], acc[6]};
      for (int c = 0; c < 9; ++c) out[1 + c] = s[c] + (sig_extra ? sig_extra[c] : 0.0);
    }
  }
}

}  // namespace d4k
