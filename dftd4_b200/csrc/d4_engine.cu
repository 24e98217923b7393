// d4cu engine: host orchestration of the sm_100a DFT-D4 pipeline (include/d4cu.h).
// Stage order follows the reference driver src/dftd4/disp.f90:40-134.
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/d4cu.h"
#include "d4_atm.cuh"
#include "d4_batch.cuh"
#include "d4_chol.cuh"
#include "d4_common.cuh"
#include "d4_eeq.cuh"
#include "d4_host.hpp"
#include "d4_kernels.cuh"
#include "d4_props.cuh"

using namespace d4k;

namespace {

thread_local std::string g_last_error;
int fail(const std::string& msg) {
  g_last_error = msg;
  return 1;
}
#define CUDA_OK(call)                                                                             \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call);           \
  } while (0)

template <class T>
struct DBuf {
  T* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = n + n / 8 + 64;
    cudaError_t e = cudaMalloc(&p, cap * sizeof(T));
    if (e != cudaSuccess) cap = 0;
    return e;
  }
  ~DBuf() { if (p) cudaFree(p); }
};

template <class T>
cudaError_t upload_vec(DBuf<T>& d, const std::vector<T>& h, cudaStream_t st) {
  cudaError_t e = d.ensure(std::max<size_t>(h.size(), 1));
  if (e != cudaSuccess) return e;
  if (h.empty()) return cudaSuccess;
  return cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, st);
}

// ---- NCCL through dlopen: no link-time dependency, single-GPU use never touches it --------------------------
struct Nccl {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, const void*, int) = nullptr;  // ncclUniqueId passed by value: 128-byte struct
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  bool load(std::string& err) {
    if (lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) { err = "NCCL library not found (dlopen libnccl.so.2)"; return false; }
    GetUniqueId = reinterpret_cast<int (*)(void*)>(dlsym(lib, "ncclGetUniqueId"));
    AllReduce = reinterpret_cast<int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t)>(dlsym(lib, "ncclAllReduce"));
    CommDestroy = reinterpret_cast<int (*)(void*)>(dlsym(lib, "ncclCommDestroy"));
    Broadcast = reinterpret_cast<int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t)>(dlsym(lib, "ncclBroadcast"));
    GroupStart = reinterpret_cast<int (*)()>(dlsym(lib, "ncclGroupStart"));
    GroupEnd = reinterpret_cast<int (*)()>(dlsym(lib, "ncclGroupEnd"));
    if (!GetUniqueId || !AllReduce || !CommDestroy || !Broadcast || !GroupStart || !GroupEnd || !dlsym(lib, "ncclCommInitRank")) {
      err = "NCCL symbols missing";
      return false;
    }
    return true;
  }
};
Nccl g_nccl;
struct NcclUid { char b[128]; };
using CommInitRankFn = int (*)(void**, int, NcclUid, int);

// ---- one neighbour list (cutoff class) ------------------------------------------------------------------------
struct NList {
  d4::Translations tr;       // host, reference order
  DBuf<double> d_trans;
  DBuf<int> d_triples;
  DBuf<unsigned> keys, keys2, pids, pids2, cell_start;
  DBuf<Point> pts;
  DBuf<int> counts;
  DBuf<long long> offsets;
  DBuf<unsigned> idx;
  DBuf<unsigned char> cub_tmp;
  long long total = 0, npts = 0;
  int nrows = 0;
  double cutoff = 0.0;
};

}  // namespace

struct d4cu_engine {
  int device = 0;
  cudaStream_t st = nullptr;
  cudaStream_t st_side = nullptr;        // look-ahead stream of the Cholesky factorisation (d4_chol.cuh)
  cudaEvent_t ev_chol[2] = {nullptr, nullptr};
  std::mutex mtx;
  // problem
  int nat = 0, nid = 0, ld = 0, nblk = 0, ts = 0;
  bool any_periodic = false, uploaded = false, computed = false;
  double charge = 0.0, lattice[9] = {0}, param[6] = {0}, cutoff[5] = {0};
  double ga = 3, gc = 2, wf = 6;
  std::vector<double> h_xyz;
  std::vector<int> h_id;
  // sharding
  int rank = 0, world = 1;
  void* comm = nullptr;
  std::vector<int> h_rows;
  // device state
  DBuf<double> d_xyz, d_tab_d;   // all double tables packed
  DBuf<int> d_id, d_tab_i, d_rows, d_ident;
  Tables tb{};
  EwaldDev ew{};
  d4::EwaldSetup ewh;
  DBuf<double> d_ew, d_ewtab;
  NList lcn, l2, l3;
  DBuf<double> cn, cn_eeq, dxdcn_cut, q, gw, gwdcn, gwdq, gw0, gwdcn0, e_atom, dEdcn, dEdq, grad, sig_part, lambda, w2, W0, W0cn, P, sfac, eacc;
  DBuf<unsigned> idx3own;   // disp3 rows reordered [owned partners | others] for the two-edge ATM kernel
  DBuf<int> nown3, nblk3, blkoff3, atom3, work3;   // owned entries / 32-edge blocks per row, their scan, atom of every entry
  DBuf<double> rec3;                               // AtmRec per disp3 list entry (64 B)
  DBuf<double> A, invL, rhs, sol, zvec, aux, out, redbuf;
  DBuf<int> info;
  DBuf<double> panel;     // packed block column of L travelling between ranks (cholesky_factor_dist)
  DBuf<int> ctl2, ctl1;   // flags + error word of the fused substitution kernels (2-RHS charge solve, 1-RHS adjoint solve)
  DBuf<unsigned long long> counters;
  DBuf<long long> scan_tmp_ll;
  cudaEvent_t ev[12] = {nullptr};
  cudaEvent_t evA[10] = {nullptr}, evB[10] = {nullptr};   // begin / end of every stage (stages run in a rank-dependent order)
  bool st_used[10] = {false};
  float ms[10] = {0};
  // Multi-GPU "EEQ owner" mode: rank 0 alone builds, factors and solves the EEQ system while the other ranks work on
  // the three-body rows; rank 0 takes a smaller share of those rows (atm_v0 of every atm_v0 + 16 (world - 1) row blocks),
  // adapted from the measured stage times of the previous evaluation.
  std::vector<int> h_rows3;
  DBuf<int> d_rows3;
  int atm_v0 = 16, atm_v0_rows = -1, atm_rows_nat = -1;   // h_rows3 was dealt for (atm_v0_rows, atm_rows_nat, rank, world)
  DBuf<double> bal;
  int launches = 0;
  bool want_grad = true;
};

namespace {

int upload_tables(d4cu_engine* E, const d4cu_model* m) {
  const int nid = m->nid;
  E->nid = nid; E->ga = m->ga; E->gc = m->gc; E->wf = m->wf;
  std::vector<double> d;
  std::vector<int> iv;
  auto push_d = [&](const double* p, size_t n) { size_t o = d.size(); d.insert(d.end(), p, p + n); return o; };
  auto push_i = [&](const int* p, size_t n) { size_t o = iv.size(); iv.insert(iv.end(), p, p + n); return o; };
  const size_t o_zeff = push_d(m->zeff, nid), o_eta = push_d(m->eta, nid), o_refcn = push_d(m->refcn, nid * 7),
               o_refq = push_d(m->refq, nid * 7), o_refc6 = push_d(m->refc6, (size_t)nid * nid * 49),
               o_alpha0 = push_d(m->refalpha0, nid * 7), o_chi = push_d(m->eeq_chi, nid), o_eeta = push_d(m->eeq_eta, nid),
               o_kcn = push_d(m->eeq_kcn, nid), o_rad = push_d(m->eeq_rad, nid);
  const size_t o_wfp = m->wfpair ? push_d(m->wfpair, (size_t)nid * nid) : 0;
  std::vector<double> rc(nid * nid), den(nid * nid), rr(nid * nid), r0(nid * nid), gam(nid * nid);
  const double k4 = 4.10451, k5 = 19.08857, k6 = 2.0 * 11.28174 * 11.28174;  // mctc-lib ncoord dftd4 flavour
  for (int a = 0; a < nid; ++a)
    for (int b = 0; b < nid; ++b) {
      rc[a * nid + b] = m->rcov[a] + m->rcov[b];
      const double dEN = std::fabs(m->en[a] - m->en[b]) + k5;
      den[a * nid + b] = k4 * std::exp(-(dEN * dEN) / k6);
      rr[a * nid + b] = 3 * m->r4r2[a] * m->r4r2[b];                               // rational.f90:283
      r0[a * nid + b] = E->param[3] * std::sqrt(rr[a * nid + b]) + E->param[4];    // rational.f90:284
      gam[a * nid + b] = 1.0 / std::sqrt(m->eeq_rad[a] * m->eeq_rad[a] + m->eeq_rad[b] * m->eeq_rad[b]);
    }
  const size_t o_rc = push_d(rc.data(), rc.size()), o_den = push_d(den.data(), den.size()), o_rr = push_d(rr.data(), rr.size()),
               o_r0 = push_d(r0.data(), r0.size()), o_gam = push_d(gam.data(), gam.size());
  std::vector<int> soff(nid + 1, 0);
  for (int s = 0; s < nid; ++s) soff[s + 1] = soff[s] + m->nref[s] + 1;
  E->ts = soff[nid];
  const size_t o_nref = push_i(m->nref, nid), o_ngw = push_i(m->ngw, nid * 7), o_soff = push_i(soff.data(), nid + 1);
  CUDA_OK(upload_vec(E->d_tab_d, d, E->st));
  CUDA_OK(upload_vec(E->d_tab_i, iv, E->st));
  Tables& t = E->tb;
  const double* D = E->d_tab_d.p;
  const int* I = E->d_tab_i.p;
  t.nid = nid; t.ga = m->ga; t.gc = m->gc; t.wf = m->wf;
  t.zeff = D + o_zeff; t.eta = D + o_eta; t.refcn = D + o_refcn; t.refq = D + o_refq; t.refc6 = D + o_refc6;
  t.refalpha0 = D + o_alpha0; t.chi = D + o_chi; t.eeta = D + o_eeta; t.kcn = D + o_kcn; t.rad = D + o_rad;
  t.pair_rc = D + o_rc; t.pair_den = D + o_den; t.pair_rr = D + o_rr; t.pair_r0 = D + o_r0; t.pair_gam = D + o_gam;
  t.nref = I + o_nref; t.ngw = I + o_ngw; t.soff = I + o_soff;
  t.nsp = m->wfpair ? nid : 1;
  t.wfpair = m->wfpair ? D + o_wfp : nullptr;
  return 0;
}

// Host part of the periodic EEQ: per-species diagonal element and its strain derivative.  The self-image list follows
// multicharge's get_pairs, which stores the position inside the compacted distance array (the origin is skipped), so
// the image actually used is the translation one slot before the nearest one; for lattice vectors the Ewald sum is
// invariant under that substitution up to its truncation.
int upload_ewald(d4cu_engine* E, const d4cu_model* m) {
  d4::make_ewald_setup(E->lattice, E->ewh);
  const d4::EwaldSetup& S = E->ewh;
  const int nid = m->nid, nw = S.wsc.count(), nd = S.dir.count(), nr = (int)S.rec_fac.size();
  const double thr = 1.4901161193847656e-08, tol = 0.01, sqrtpi = 1.77245385090551602729816748334114518;
  std::vector<int> sel;
  {
    std::vector<double> dist;
    for (int t = 0; t < nw; ++t) {
      const double* v = &S.wsc.trans[3 * (size_t)t];
      const double r2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
      if (r2 < thr) continue;
      dist.push_back(r2);
    }
    double best = 1e300;
    for (double x : dist) best = std::min(best, x);
    for (size_t c = 0; c < dist.size(); ++c)
      if (std::fabs(dist[c] - best) <= tol) sel.push_back((int)c);
  }
  std::vector<double> diag(nid, 0.0), diag_ds((size_t)nid * 9, 0.0);
  const double alp = S.alpha, alp2 = alp * alp;
  for (int s = 0; s < nid; ++s) {
    const double rad = m->eeq_rad[s];
    const double gam = 1.0 / std::sqrt(2.0 * rad * rad), gam2 = gam * gam;
    const double wsw = 1.0 / (double)sel.size();
    double acc = 0.0, ds[9] = {0};
    for (int c : sel) {
      const double* t = &S.wsc.trans[3 * (size_t)c];
      for (int d = 0; d < nd; ++d) {
        const double v[3] = {t[0] + S.dir.trans[3 * (size_t)d], t[1] + S.dir.trans[3 * (size_t)d + 1], t[2] + S.dir.trans[3 * (size_t)d + 2]};
        const double r1 = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        if (r1 < thr) continue;
        acc += (std::erf(gam * r1) / r1 - std::erf(alp * r1) / r1) * wsw;
        const double r2 = r1 * r1;
        const double gtmp = +2 * gam * std::exp(-r2 * gam2) / (sqrtpi * r2) - std::erf(r1 * gam) / (r2 * r1);
        const double atmp = -2 * alp * std::exp(-r2 * alp2) / (sqrtpi * r2) + std::erf(r1 * alp) / (r2 * r1);
        for (int a = 0; a < 3; ++a)
          for (int b = 0; b < 3; ++b) ds[3 * a + b] += (gtmp + atmp) * v[a] * v[b] * wsw;
      }
      for (int k = 0; k < nr; ++k) {
        const double* g = &S.rec[3 * (size_t)k];
        const double g2 = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
        if (g2 < thr) continue;
        const double gv = t[0] * g[0] + t[1] * g[1] + t[2] * g[2];
        // (the value term sum_k f_k cos(G.T) = sum_k f_k is added on the device by the phase-matrix product B B^T)
        const double c2 = S.rec_fac[k] * std::cos(gv) * wsw;
        for (int a = 0; a < 3; ++a)
          for (int b = 0; b < 3; ++b) ds[3 * a + b] += c2 * ((2.0 / g2 + 0.5 / alp2) * g[a] * g[b] - (a == b ? 1.0 : 0.0));
      }
    }
    diag[s] = acc + (m->eeq_eta[s] + EEQ_SQRT2PI / rad - 2 * alp / sqrtpi);
    for (int a = 0; a < 9; ++a) diag_ds[(size_t)s * 9 + a] = ds[a];
  }
  // half set of the reciprocal vectors (one of every +G/-G pair: first non-zero multiple positive)
  std::vector<double> gh, fh;
  for (int k = 0; k < nr; ++k) {
    const int* t = &S.rec_triples[3 * (size_t)k];
    const bool pos = t[0] > 0 || (t[0] == 0 && (t[1] > 0 || (t[1] == 0 && t[2] > 0)));
    if (!pos || S.rec_fac[k] == 0.0) continue;
    gh.insert(gh.end(), &S.rec[3 * (size_t)k], &S.rec[3 * (size_t)k] + 3);
    fh.push_back(S.rec_fac[k]);
  }
  if (fh.size() > EEQ_NH) return fail("internal error: more than 64 reciprocal half-vectors");
  const int nh = (int)fh.size();
  std::vector<double> pack;
  auto push = [&](const std::vector<double>& v) { size_t o = pack.size(); pack.insert(pack.end(), v.begin(), v.end()); return o; };
  const size_t o_dir = push(S.dir.trans), o_rec = push(S.rec), o_fac = push(S.rec_fac), o_wsc = push(S.wsc.trans),
               o_diag = push(diag), o_ds = push(diag_ds), o_gh = push(gh), o_fh = push(fh);
  CUDA_OK(upload_vec(E->d_ew, pack, E->st));
  EwaldDev& w = E->ew;
  w.alpha = S.alpha; w.volume = S.volume; w.ndir = nd; w.nrec = nr; w.nwsc = nw;
  w.dir = E->d_ew.p + o_dir; w.rec = E->d_ew.p + o_rec; w.rec_fac = E->d_ew.p + o_fac; w.wsc = E->d_ew.p + o_wsc;
  w.diag = E->d_ew.p + o_diag; w.diag_ds = E->d_ew.p + o_ds;
  w.nh = nh; w.gh = E->d_ew.p + o_gh; w.fh = E->d_ew.p + o_fh;
  // piecewise polynomials of the alpha part of the real-space terms: r^2 from 2^EEQ_TAB_EMIN to beyond (4.5 / alpha)^2, the
  // NEAR class of ewald_dir_sum (16 intervals per octave)
  {
    const double r2max = (4.6 / S.alpha) * (4.6 / S.alpha);
    int noct = 1;
    while (std::ldexp(1.0, EEQ_TAB_EMIN + noct) <= r2max) ++noct;
    int nint = 16 * noct;
    if (nint > EEQ_TAB_MAXN) nint = EEQ_TAB_MAXN;   // (larger distances then take the direct evaluation)
    CUDA_OK(E->d_ewtab.ensure(2 * (size_t)nint * (EEQ_TAB_DEG + 1)));
    w.tabf = E->d_ewtab.p; w.tabg = E->d_ewtab.p + (size_t)nint * (EEQ_TAB_DEG + 1);
    w.tab_n = nint; w.tab_base = (1023 + EEQ_TAB_EMIN) << 4;
    eeq_table_kernel<<<(nint + 63) / 64, 64, 0, E->st>>>(S.alpha, nint, E->d_ewtab.p, E->d_ewtab.p + (size_t)nint * (EEQ_TAB_DEG + 1));
    CUDA_OK(cudaGetLastError());
  }
  return 0;
}

// Build one neighbour list for the given rows.
int build_list(d4cu_engine* E, NList& L, double cutoff, double eps2, const int* d_rows, int nrows, int fine = 0) {
  cudaStream_t st = E->st;
  const int nat = E->nat;
  L.cutoff = cutoff;
  L.nrows = nrows;
  d4::lattice_points_cutoff(E->lattice, E->any_periodic, cutoff, L.tr);
  const int T = L.tr.count();
  const long long npts = (long long)nat * T;
  if (npts >= (1ll << 31)) return fail("image-expanded point set exceeds 2^31 entries");
  if (T >= (1 << 24)) return fail("more than 2^24 lattice translations");
  L.npts = npts;
  CUDA_OK(upload_vec(L.d_trans, L.tr.trans, st));
  CUDA_OK(upload_vec(L.d_triples, L.tr.triples, st));
  // bounding box of the home atoms
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int i = 0; i < nat; ++i)
    for (int c = 0; c < 3; ++c) {
      lo[c] = std::min(lo[c], E->h_xyz[3 * (size_t)i + c]);
      hi[c] = std::max(hi[c], E->h_xyz[3 * (size_t)i + c]);
    }
  Grid g;
  const double margin = cutoff * (1.0 + 1e-9) + 1e-6;
  g.lox = lo[0] - margin; g.loy = lo[1] - margin; g.loz = lo[2] - margin;
  g.hix = hi[0] + margin; g.hiy = hi[1] + margin; g.hiz = hi[2] + margin;
  g.ox = g.lox; g.oy = g.loy; g.oz = g.loz;
  // cell edge: cutoff/2 (reach 2) but never more than ~2M cells
  double h = cutoff * 0.5 * (1.0 + 1e-9);
  int reach = 2;
  auto ncells = [&](double hh) {
    return (long long)(std::floor((g.hix - g.ox) / hh) + 1) * (long long)(std::floor((g.hiy - g.oy) / hh) + 1) *
           (long long)(std::floor((g.hiz - g.oz) / hh) + 1);
  };
  while (ncells(h) > (1ll << 21)) { h *= 2.0; reach = std::max(1, reach / 2); }
  if (reach < 1) reach = 1;
  if (h * reach < cutoff) reach = (int)std::ceil(cutoff / h);
  g.h = h; g.inv_h = 1.0 / h; g.reach = reach;
  g.nx = (int)std::floor((g.hix - g.ox) / h) + 1;
  g.ny = (int)std::floor((g.hiy - g.oy) / h) + 1;
  g.nz = (int)std::floor((g.hiz - g.oz) / h) + 1;
  const int ncell = g.nx * g.ny * g.nz;

  CUDA_OK(L.keys.ensure(npts)); CUDA_OK(L.keys2.ensure(npts)); CUDA_OK(L.pids.ensure(npts)); CUDA_OK(L.pids2.ensure(npts));
  CUDA_OK(L.cell_start.ensure(ncell + 2)); CUDA_OK(L.pts.ensure(npts));
  int bits = 1;
  while ((1ll << bits) <= ncell) ++bits;
  while (fine > 0 && bits + 3 * fine > 31) --fine;
  point_keys_kernel<<<(unsigned)((npts + 255) / 256), 256, 0, st>>>(nat, npts, E->d_xyz.p, L.d_trans.p, g, fine, L.keys.p, L.pids.p);
  bits += 3 * fine;
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, L.keys.p, L.keys2.p, L.pids.p, L.pids2.p, (int)npts, 0, bits, st);
  CUDA_OK(L.cub_tmp.ensure(tmp_bytes));
  CUDA_OK(cub::DeviceRadixSort::SortPairs(L.cub_tmp.p, tmp_bytes, L.keys.p, L.keys2.p, L.pids.p, L.pids2.p, (int)npts, 0, bits, st));
  cell_bounds_kernel<<<(ncell + 1 + 255) / 256, 256, 0, st>>>(L.keys2.p, npts, ncell, fine, L.cell_start.p);
  make_points_kernel<<<(unsigned)((npts + 255) / 256), 256, 0, st>>>(nat, npts, E->d_xyz.p, L.d_trans.p, E->d_id.p, L.keys2.p,
                                                                    L.pids2.p, (unsigned)ncell << (3 * fine), L.pts.p);
  CUDA_OK(L.counts.ensure(nrows + 1)); CUDA_OK(L.offsets.ensure(nrows + 2));
  if (nrows == 0) {   // a rank without rows (fewer row blocks than ranks): empty list, but every later collective is reached
    CUDA_OK(cudaMemsetAsync(L.offsets.p, 0, 2 * sizeof(long long), st));
    L.total = 0;
    CUDA_OK(L.idx.ensure(1));
    E->launches += 4;
    return 0;
  }
  const double rc2 = cutoff * cutoff;
  const int nb = (nrows * 32 + 255) / 256;
  nlist_kernel<false><<<nb, 256, 0, st>>>(nrows, d_rows, E->d_xyz.p, L.pts.p, L.cell_start.p, g, rc2, eps2, L.d_trans.p, L.counts.p, nullptr, nullptr);
  CUDA_OK(cudaMemsetAsync(L.counts.p + nrows, 0, sizeof(int), st));
  tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, L.counts.p, L.offsets.p, nrows + 1, st);
  CUDA_OK(L.cub_tmp.ensure(tmp_bytes));
  CUDA_OK(cub::DeviceScan::ExclusiveSum(L.cub_tmp.p, tmp_bytes, L.counts.p, L.offsets.p, nrows + 1, st));
  long long total = 0;
  CUDA_OK(cudaMemcpyAsync(&total, L.offsets.p + nrows, sizeof(long long), cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  L.total = total;
  CUDA_OK(L.idx.ensure((size_t)std::max<long long>(total, 1)));
  nlist_kernel<true><<<nb, 256, 0, st>>>(nrows, d_rows, E->d_xyz.p, L.pts.p, L.cell_start.p, g, rc2, eps2, L.d_trans.p, nullptr, L.offsets.p, L.idx.p);
  E->launches += 7;
  CUDA_OK(cudaGetLastError());
  return 0;
}

int allreduce(d4cu_engine* E, double* buf, size_t n) {
  if (E->world <= 1) return 0;
  const int rc = g_nccl.AllReduce(buf, buf, n, /*ncclFloat64*/ 8, /*ncclSum*/ 0, E->comm, E->st);
  if (rc != 0) return fail("ncclAllReduce failed with code " + std::to_string(rc));
  return 0;
}


// Multi-GPU Cholesky of the (replicated) EEQ matrix: block column k belongs to rank k % W.  Per step the owner factors
// the diagonal block and solves the panel, the panel (= block column k of L) and inv(L_kk) are broadcast over NVLink,
// every rank stores them into its own copy of A / invL and applies the rank-128 update to the block columns IT owns.
// All ranks end with the complete off-diagonal factor and all inverted diagonal blocks, which is what the (replicated)
// substitution kernels read; trailing-update flops and C-tile traffic are divided by W.  Used for large matrices only
// (the per-step broadcast costs more than it saves when a trailing update takes < ~100 us per rank).
int cholesky_factor_dist(d4cu_engine* E, int* launches) {
  cudaStream_t st = E->st;
  const int ld = E->ld, nblk = E->nblk, W = E->world, rank = E->rank;
  const size_t diag_smem = chol_diag_smem_bytes();
  const size_t gemm_smem = (size_t)4 * NB * GS * sizeof(double);
  cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)diag_smem);
  cudaFuncSetAttribute(gemm_nt_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem);
  cudaFuncSetAttribute(chol_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UPD_SMEM);
  CUDA_OK(E->panel.ensure((size_t)ld * NB));
  double* A = E->A.p;
  // Look-ahead: diagonal block, panel solve and the broadcast of step kb live on the high-priority side stream; the rank
  // that owns block column kb + 1 updates that column first and signals the side stream, so factorisation + broadcast of
  // the next panel overlap the bulk of the trailing update on the main stream.  All NCCL calls of the factorisation are
  // issued on the side stream in the same order on every rank; the main stream holds no collective in this phase.
  cudaStream_t sd = E->st_side;
  cudaEvent_t ev_col = E->ev_chol[0], ev_pan = E->ev_chol[1];
  CUDA_OK(cudaEventRecord(ev_col, st));   // the matrix build (and its band exchange) precedes the first panel
  for (int kb = 0; kb < nblk; ++kb) {
    const int owner = d4cu_chol_column_owner(kb, W), rest = nblk - kb - 1;
    double* col = A + (size_t)(kb + 1) * NB * ld + (size_t)kb * NB;   // first panel tile of block column kb
    // ---- side stream: factor + solve + exchange block column kb ----
    CUDA_OK(cudaStreamWaitEvent(sd, ev_col, 0));
    if (rank == owner) {
      chol_diag_kernel<<<1, 512, diag_smem, sd>>>(A, ld, kb, E->invL.p, E->info.p);
      ++*launches;
      if (rest > 0) {
        gemm_nt_kernel<0><<<rest, 256, gemm_smem, sd>>>(A, ld, kb, rest, E->invL.p);
        ++*launches;
        CUDA_OK(cudaMemcpy2DAsync(E->panel.p, NB * sizeof(double), col, (size_t)ld * sizeof(double), NB * sizeof(double),
                                  (size_t)rest * NB, cudaMemcpyDeviceToDevice, sd));
      }
    }
    if (g_nccl.GroupStart() != 0) return fail("ncclGroupStart failed");
    double* ik = E->invL.p + (size_t)kb * NB * NB;
    if (g_nccl.Broadcast(ik, ik, (size_t)NB * NB, /*ncclFloat64*/ 8, owner, E->comm, sd) != 0) return fail("ncclBroadcast(invL) failed");
    if (rest > 0 && g_nccl.Broadcast(E->panel.p, E->panel.p, (size_t)rest * NB * NB, 8, owner, E->comm, sd) != 0)
      return fail("ncclBroadcast(panel) failed");
    if (g_nccl.GroupEnd() != 0) return fail("ncclGroupEnd failed");
    if (rest > 0 && rank != owner)
      CUDA_OK(cudaMemcpy2DAsync(col, (size_t)ld * sizeof(double), E->panel.p, NB * sizeof(double), NB * sizeof(double),
                                (size_t)rest * NB, cudaMemcpyDeviceToDevice, sd));
    CUDA_OK(cudaEventRecord(ev_pan, sd));
    // ---- main stream: trailing update of the owned block columns, the next panel's column first ----
    CUDA_OK(cudaStreamWaitEvent(st, ev_pan, 0));
    if (rest > 0) {
      // owned tile columns of the trailing matrix: absolute block column kb + 1 + tj with (kb + 1 + tj) % W == rank
      int t0 = d4cu_chol_first_owned(kb, rank, W);
      if (t0 == 0) {   // this rank factors block column kb + 1 next
        chol_update_kernel<<<2 * rest, 256, UPD_SMEM, st>>>(A, ld, kb, rest, 1, 0, 1);
        ++*launches;
        t0 = W;
      }
      CUDA_OK(cudaEventRecord(ev_col, st));
      if (t0 < rest) {
        const int ncols = (rest - t0 + W - 1) / W;
        for (int c0 = 0; c0 < ncols; c0 += 16384) {   // grid.y limit (two half-block tiles per column)
          const int nc = std::min(16384, ncols - c0);
          dim3 grid((unsigned)(rest - (t0 + c0 * W)), (unsigned)(2 * nc));
          chol_update_kernel<<<grid, 256, UPD_SMEM, st>>>(A, ld, kb, rest, 3, t0 + c0 * W, W);
          ++*launches;
        }
      }
    }
  }
  // a failed pivot is seen by its owner only: make the verdict collective (max over ranks)
  if (g_nccl.AllReduce(E->info.p, E->info.p, 1, /*ncclInt32*/ 2, /*ncclMax*/ 2, E->comm, st) != 0) return fail("ncclAllReduce(info) failed");
  return 0;
}

enum { ST_LISTS = 0, ST_CN, ST_BUILD, ST_SOLVE, ST_WEIGHTS, ST_PAIR, ST_ADJ, ST_ATM, ST_CHAIN, ST_TOTAL };
inline void stage_begin(d4cu_engine* E, int k) { cudaEventRecord(E->evA[k], E->st); E->st_used[k] = true; }
inline void stage_end(d4cu_engine* E, int k) { cudaEventRecord(E->evB[k], E->st); }

// Whether the EEQ system of a multi-GPU run is handled by rank 0 alone (the others then start with their three-body rows)
// or factored column-cyclically by all ranks (cholesky_factor_dist): D4CU_CHOL_DIST = 1 forces the distributed
// factorisation, 0 the owner; default: owner below 16384 rows (the factorisation fits beside the other ranks' three-body
// work), distributed above.
inline bool eeq_distributed(const d4cu_engine* E) {
  static const int dist_mode = [] { const char* v = std::getenv("D4CU_CHOL_DIST"); return v ? std::atoi(v) : -1; }();
  return E->world > 1 && (dist_mode == 1 || (dist_mode < 0 && E->ld >= 16384));
}
inline bool eeq_owner_mode(const d4cu_engine* E) { return E->world > 1 && !eeq_distributed(E); }

// Coordination numbers (D4 + EEQ flavours), replicated on every rank.
int stage_cn(d4cu_engine* E) {
  cudaStream_t st = E->st;
  const int nat = E->nat;
  const double cutcn = E->cutoff[2];
  const double eeq_cut = 25.0, kcn = 7.5;
  CUDA_OK(E->cn.ensure(nat)); CUDA_OK(E->cn_eeq.ensure(nat)); CUDA_OK(E->dxdcn_cut.ensure(nat)); CUDA_OK(E->q.ensure(nat));
  CUDA_OK(E->counters.ensure(8));
  CUDA_OK(cudaMemsetAsync(E->counters.p, 0, 8 * sizeof(unsigned long long), st));
  stage_begin(E, ST_CN);
  cn_kernel<<<(nat * 32 + 255) / 256, 256, 0, st>>>(nat, E->d_xyz.p, E->d_id.p, E->lcn.pts.p, E->lcn.offsets.p, E->lcn.idx.p,
                                                    E->lcn.d_trans.p, E->tb, cutcn * cutcn, eeq_cut * eeq_cut, kcn, E->cn.p, E->cn_eeq.p, E->counters.p);
  ++E->launches;
  stage_end(E, ST_CN);
  return 0;
}

// Reference weights: the q = 0 set only needs the coordination numbers (with_q = false: q is read from a zeroed buffer
// and the q-dependent outputs are rewritten later), the full call follows the charges.
int stage_weights(d4cu_engine* E, bool with_q) {
  cudaStream_t st = E->st;
  const int nat = E->nat;
  const size_t n7 = (size_t)nat * E->tb.nsp * D4_MREF;  // D4S: one weight set per partner species
  CUDA_OK(E->gw.ensure(n7)); CUDA_OK(E->gwdcn.ensure(n7)); CUDA_OK(E->gwdq.ensure(n7)); CUDA_OK(E->gw0.ensure(n7)); CUDA_OK(E->gwdcn0.ensure(n7));
  if (!with_q) CUDA_OK(cudaMemsetAsync(E->q.p, 0, sizeof(double) * nat, st));
  if (with_q) stage_begin(E, ST_WEIGHTS);
  weights_kernel<<<(nat + 127) / 128, 128, 0, st>>>(nat, E->d_id.p, E->tb, E->cn.p, E->q.p, E->gw.p, E->gwdcn.p, E->gwdq.p, E->gw0.p, E->gwdcn0.p);
  ++E->launches;
  if (with_q) stage_end(E, ST_WEIGHTS);
  return 0;
}

// EEQ right-hand side, matrix build, factorisation, charges.  Owner mode: only rank 0 builds / factors / solves, q is
// broadcast; otherwise replicated (one GPU) or distributed over the ranks (cholesky_factor_dist).
int stage_eeq(d4cu_engine* E) {
  cudaStream_t st = E->st;
  const int nat = E->nat, ld = E->ld;
  const bool owner = eeq_owner_mode(E), mine = !owner || E->rank == 0;
  // ---- EEQ: build, factor, solve ----
  CUDA_OK(E->A.ensure((size_t)ld * ld)); CUDA_OK(E->invL.ensure((size_t)E->nblk * NB * NB));
  CUDA_OK(E->rhs.ensure(2 * (size_t)ld)); CUDA_OK(E->sol.ensure(2 * (size_t)ld)); CUDA_OK(E->zvec.ensure(ld));
  CUDA_OK(E->aux.ensure(8)); CUDA_OK(E->info.ensure(1));
  CUDA_OK(cudaMemsetAsync(E->info.p, 0, sizeof(int), st));
  eeq_rhs_kernel<<<(ld + 255) / 256, 256, 0, st>>>(nat, ld, E->d_id.p, E->tb, E->cn_eeq.p, E->rhs.p, E->dxdcn_cut.p);
  const bool dist = eeq_distributed(E);
  stage_begin(E, ST_BUILD);
  if (E->any_periodic) {   // phase matrix of the reciprocal-space sums: every rank needs it (adjoint pass)
    CUDA_OK(E->P.ensure((size_t)ld * 2 * EEQ_NH));
    eeq_phase_kernel<<<(unsigned)(((long long)ld * EEQ_NH + 255) / 256), 256, 0, st>>>(nat, ld, E->d_xyz.p, E->ew, E->P.p);
  }
  if (mine) {
    if (E->any_periodic) {
      // real-space Ewald sum per lower-triangle element, then the reciprocal-space sum as one rank-128 DMMA update
      const long long nt = ld / EEQ_TILE;
      const size_t sh = eeq_dir_smem_bytes(E->ew.ndir, E->ew.nwsc, E->ew.tab_n);
      // One GPU and owner mode (rank 0): the whole lower triangle.  Distributed factorisation: see below.
      const long long count = nt * (nt + 1) / 2;
      if (dist) {
        // distributed factorisation: a rank only ever touches the block columns it owns -> build exactly those, no exchange
        const int ncol = (E->nblk - E->rank + E->world - 1) / E->world;
        if (ncol > 0) {
          dim3 grid((unsigned)((nt - (long long)E->rank * (NB / EEQ_TILE)) * (NB / EEQ_TILE)), (unsigned)ncol);
          eeq_build_3d_kernel<<<grid, EEQ_TILE * EEQ_TILE, sh, st>>>(nat, ld, E->d_xyz.p, E->d_id.p, E->tb, E->ew, E->A.p, 0, E->world, E->rank, E->counters.p);
        }
      } else if (count > 0) {
        eeq_build_3d_kernel<<<(unsigned)count, EEQ_TILE * EEQ_TILE, sh, st>>>(nat, ld, E->d_xyz.p, E->d_id.p, E->tb, E->ew, E->A.p, 0, 0, 0, E->counters.p);
      }
      E->launches += 1 + rank128_update(E->A.p, ld, E->nblk, E->P.p, st, dist ? E->rank : 0, dist ? E->world : 1);
    } else {
      dim3 blk(16, 16), grd((ld + 15) / 16, (ld + 15) / 16);
      eeq_build_0d_kernel<<<grd, blk, 0, st>>>(nat, ld, E->d_xyz.p, E->d_id.p, E->tb, E->A.p);
    }
  }
  E->launches += 2;
  stage_end(E, ST_BUILD);
  stage_begin(E, ST_SOLVE);
  if (mine) {
    static const bool no_lookahead = [] { const char* v = std::getenv("D4CU_CHOL_LOOKAHEAD"); return v && std::atoi(v) == 0; }();
    if (dist) {
      if (cholesky_factor_dist(E, &E->launches)) return 1;
    } else {
      E->launches += cholesky_factor(E->A.p, ld, E->nblk, E->invL.p, E->info.p, st, no_lookahead ? nullptr : E->st_side, E->ev_chol);
    }
  }
  // substitution: forward and backward sweep of all block rows in one wavefront launch
  CUDA_OK(E->ctl2.ensure(2 * (size_t)E->nblk + 3)); CUDA_OK(E->ctl1.ensure(2 * (size_t)E->nblk + 3));
  CUDA_OK(cudaMemsetAsync(E->ctl1.p + 2 * (size_t)E->nblk, 0, sizeof(int), st));
  CUDA_OK(cudaMemsetAsync(E->ctl2.p + 2 * (size_t)E->nblk, 0, sizeof(int), st));
  if (mine) {
    if (cholesky_solve_fused<2>(E->A.p, ld, E->nblk, E->invL.p, E->rhs.p, E->sol.p, ld, E->ctl2.p, st) < 0)
      return fail("fused substitution kernel cannot run on this device");
    ++E->launches;
    eeq_charges_kernel<<<1, 1024, 0, st>>>(nat, ld, E->sol.p, E->charge, E->q.p, E->aux.p);
    CUDA_OK(cudaMemcpyAsync(E->zvec.p, E->sol.p + ld, sizeof(double) * ld, cudaMemcpyDeviceToDevice, st));
    ++E->launches;
  }
  stage_end(E, ST_SOLVE);   // (owner mode: the caller broadcasts q once rank 0 has also done its three-body share)
  // (the pivot / substitution verdicts are read by eeq_check_status after the stream has drained)
  return 0;
}

// Verdict of the factorisation and the charge substitution (device flags), read once the stream is idle.
int eeq_check_status(d4cu_engine* E) {
  int info = 0, serr = 0;
  CUDA_OK(cudaMemcpy(&info, E->info.p, sizeof(int), cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(&serr, E->ctl2.p + 2 * (size_t)E->nblk, sizeof(int), cudaMemcpyDeviceToHost));
  if (info != 0)
    return fail("EEQ Coulomb matrix is not positive definite (pivot " + std::to_string(info) + "): Cholesky factorization failed");
  if (serr != 0) return fail("EEQ substitution kernel timed out waiting for a block row (internal error)");
  return 0;
}

// Owner mode: the charges and the verdicts of the factorisation (info) and of the charge substitution (error word of
// ctl2) exist on rank 0 only.  Hand all three to every rank in one group, so that eeq_check_status reaches the same
// verdict everywhere (the distributed factorisation gets there with an all-reduce, cholesky_factor_dist).
int eeq_owner_publish(d4cu_engine* E) {
  if (!eeq_owner_mode(E)) return 0;
  int* serr = E->ctl2.p + 2 * (size_t)E->nblk;
  if (g_nccl.GroupStart() != 0) return fail("ncclGroupStart failed");
  if (g_nccl.Broadcast(E->q.p, E->q.p, (size_t)E->nat, /*ncclFloat64*/ 8, 0, E->comm, E->st) != 0 ||
      g_nccl.Broadcast(E->info.p, E->info.p, 1, /*ncclInt32*/ 2, 0, E->comm, E->st) != 0 ||
      g_nccl.Broadcast(serr, serr, 1, /*ncclInt32*/ 2, 0, E->comm, E->st) != 0)
    return fail("ncclBroadcast(charges) failed");
  if (g_nccl.GroupEnd() != 0) return fail("ncclGroupEnd failed");
  return 0;
}

// CN -> EEQ -> weights in one go (properties; single-GPU order of the evaluation).
int stage_cn_eeq_weights(d4cu_engine* E) {
  if (stage_cn(E)) return 1;
  if (stage_eeq(E)) return 1;
  if (eeq_owner_publish(E)) return 1;
  CUDA_OK(cudaStreamSynchronize(E->st));
  if (eeq_check_status(E)) return 1;
  return stage_weights(E, true);
}

}  // namespace

// =================================================================================================================
extern "C" {

const char* d4cu_last_error(void) { return g_last_error.c_str(); }
void d4cu_set_last_error(const char* msg) { g_last_error = msg ? msg : ""; }

int d4cu_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int d4cu_engine_create(d4cu_engine** out, int device) {
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0)
    return fail("no CUDA device available: the D4 engine has no CPU fallback");
  if (device < 0 || device >= n) return fail("invalid CUDA device index");
  CUDA_OK(cudaSetDevice(device));
  d4cu_engine* E = new d4cu_engine();
  E->device = device;
  CUDA_OK(cudaStreamCreateWithFlags(&E->st, cudaStreamNonBlocking));
  {
    // highest priority: the single-CTA diagonal factorisation must be dispatched ahead of the queued CTAs of the
    // trailing update it overlaps with
    int prio_lo = 0, prio_hi = 0;
    CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CUDA_OK(cudaStreamCreateWithPriority(&E->st_side, cudaStreamNonBlocking, prio_hi));
  }
  for (auto& e : E->ev) CUDA_OK(cudaEventCreate(&e));
  for (auto& e : E->evA) CUDA_OK(cudaEventCreate(&e));
  for (auto& e : E->evB) CUDA_OK(cudaEventCreate(&e));
  for (auto& e : E->ev_chol) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  *out = E;
  return 0;
}

void d4cu_engine_destroy(d4cu_engine* E) {
  if (!E) return;
  cudaSetDevice(E->device);
  cudaStreamSynchronize(E->st);
  if (E->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(E->comm);
  for (auto& e : E->ev) if (e) cudaEventDestroy(e);
  for (auto& e : E->evA) if (e) cudaEventDestroy(e);
  for (auto& e : E->evB) if (e) cudaEventDestroy(e);
  for (auto& e : E->ev_chol) if (e) cudaEventDestroy(e);
  if (E->st_side) cudaStreamDestroy(E->st_side);
  if (E->st) cudaStreamDestroy(E->st);
  delete E;
}

int d4cu_partition(int nat, int rank, int world, int block, int* n_blocks, int* first_block, int* stride) {
  if (world < 1 || rank < 0 || rank >= world || block < 1) return 1;
  const int nb = (nat + block - 1) / block;
  if (n_blocks) *n_blocks = nb;
  if (first_block) *first_block = rank;
  if (stride) *stride = world;
  return 0;
}

int d4cu_chol_column_owner(int block_column, int world) { return world > 0 ? block_column % world : 0; }
int d4cu_chol_first_owned(int kb, int rank, int world) {
  if (world < 1) return 0;
  return ((rank - (kb + 1)) % world + world) % world;
}

int d4cu_nccl_unique_id(void* uid128) {
  std::string err;
  if (!g_nccl.load(err)) return fail(err);
  if (g_nccl.GetUniqueId(uid128) != 0) return fail("ncclGetUniqueId failed");
  return 0;
}

int d4cu_engine_set_comm(d4cu_engine* E, int rank, int world, const void* uid128) {
  if (!E) return fail("engine is NULL");
  std::lock_guard<std::mutex> lk(E->mtx);
  if (world <= 1) { E->rank = 0; E->world = 1; return 0; }
  std::string err;
  if (!g_nccl.load(err)) return fail(err);
  CUDA_OK(cudaSetDevice(E->device));
  NcclUid uid;
  std::memcpy(uid.b, uid128, 128);
  auto init = reinterpret_cast<CommInitRankFn>(dlsym(g_nccl.lib, "ncclCommInitRank"));
  void* comm = nullptr;
  const int rc = init(&comm, world, uid, rank);
  if (rc != 0) return fail("ncclCommInitRank failed with code " + std::to_string(rc));
  E->comm = comm; E->rank = rank; E->world = world;
  E->uploaded = false;
  E->atm_v0 = 16; E->atm_v0_rows = -1;   // the owner-mode three-body rows were dealt for the old rank / world
  return 0;
}

int d4cu_upload(d4cu_engine* E, int nat, const int* id, const double* xyz, double charge, const double* lattice,
                const int* periodic, const d4cu_model* model, const double* param, const double* cutoff) {
  if (!E) return fail("engine is NULL");
  if (nat <= 0) return fail("structure has no atoms");
  std::lock_guard<std::mutex> lk(E->mtx);
  CUDA_OK(cudaSetDevice(E->device));
  E->nat = nat;
  E->charge = charge;
  E->any_periodic = periodic && (periodic[0] || periodic[1] || periodic[2]);
  // 1-d / 2-d periodic input: the reference generates translations along all three lattice vectors as soon as any
  // direction is periodic (cutoff.f90:192-199), wraps all three fractional coordinates (utils.f90:36-44) and multicharge
  // switches to its Ewald matrices on any(periodic) as well -- so such a structure is evaluated as the 3-d periodic one.
  if (E->any_periodic && !lattice) return fail("periodic structure without lattice");
  std::memset(E->lattice, 0, sizeof(E->lattice));
  if (lattice) std::memcpy(E->lattice, lattice, sizeof(E->lattice));
  std::memcpy(E->param, param, sizeof(E->param));
  std::memcpy(E->cutoff, cutoff, sizeof(E->cutoff));
  for (int i = 0; i < nat; ++i)
    if (id[i] < 0 || id[i] >= model->nid) return fail("species index out of range");
  if (model->nid > 255) return fail("more than 255 species");
  E->h_xyz.assign(xyz, xyz + 3 * (size_t)nat);
  E->h_id.assign(id, id + nat);
  CUDA_OK(upload_vec(E->d_xyz, E->h_xyz, E->st));
  CUDA_OK(upload_vec(E->d_id, E->h_id, E->st));
  if (upload_tables(E, model)) return 1;
  if (E->any_periodic && upload_ewald(E, model)) return 1;
  // rows of this rank: blocks of 64 atoms dealt round-robin
  E->h_rows.clear();
  const int block = D4CU_ROW_BLOCK;
  for (int b = E->rank; b * block < nat; b += E->world)
    for (int i = b * block; i < std::min(nat, (b + 1) * block); ++i) E->h_rows.push_back(i);
  CUDA_OK(upload_vec(E->d_rows, E->h_rows, E->st));
  {
    std::vector<int> ident(nat);
    for (int i = 0; i < nat; ++i) ident[i] = i;
    CUDA_OK(upload_vec(E->d_ident, ident, E->st));
    CUDA_OK(cudaStreamSynchronize(E->st));
  }
  E->ld = ((nat + NB - 1) / NB) * NB;
  E->nblk = E->ld / NB;
  E->uploaded = true;
  E->computed = false;
  return 0;
}

static int compute_impl(d4cu_engine* E, int want_gradient, double* d_e2, double* d_e3) {
  if (!E->uploaded) return fail("d4cu_compute called before d4cu_upload");
  CUDA_OK(cudaSetDevice(E->device));
  cudaStream_t st = E->st;
  const int nat = E->nat, nid = E->nid, ld = E->ld;
  const int nrows = (int)E->h_rows.size();
  const double s6 = E->param[0], s8 = E->param[1], s9 = E->param[2], alp = E->param[5];
  const double cut2 = E->cutoff[0], cut3 = E->cutoff[1], cutcn = E->cutoff[2], w2 = E->cutoff[3], w3 = E->cutoff[4];
  const double eeq_cut = 25.0, kcn = 7.5;
  const bool grad = want_gradient != 0;
  E->want_grad = grad;
  E->launches = 0;
  const double dbl_eps = 2.220446049250313e-16;
  const bool do2 = !(std::fabs(s6) < dbl_eps && std::fabs(s8) < dbl_eps);  // rational.f90:110
  const bool do3 = !(std::fabs(s9) < dbl_eps);                             // atm.f90:91

  for (bool& u : E->st_used) u = false;
  const bool owner = eeq_owner_mode(E);
  // rows of the three-body stage: the pair / chain-rule rows, except in owner mode where rank 0 takes a smaller share
  // (it also carries the EEQ system): atm_v0 of every atm_v0 + 16 (world - 1) row blocks, the rest dealt round-robin
  const int* d_rows3 = E->d_rows.p;
  int nrows3 = nrows;
  if (owner) {
    if (E->atm_v0_rows != E->atm_v0 || E->atm_rows_nat != nat) {
      E->h_rows3.clear();
      const int block = D4CU_ROW_BLOCK, v0 = E->atm_v0, V = v0 + 16 * (E->world - 1);
      for (int b = 0; b * block < nat; ++b) {
        const int slot = b % V;
        const int own = slot < v0 ? 0 : 1 + (slot - v0) % (E->world - 1);
        if (own == E->rank)
          for (int i = b * block; i < std::min(nat, (b + 1) * block); ++i) E->h_rows3.push_back(i);
      }
      CUDA_OK(upload_vec(E->d_rows3, E->h_rows3, st));
      E->atm_v0_rows = E->atm_v0; E->atm_rows_nat = nat;
    }
    d_rows3 = E->d_rows3.p;
    nrows3 = (int)E->h_rows3.size();
  }

  CUDA_OK(cudaEventRecord(E->ev[0], st));
  stage_begin(E, ST_LISTS);
  // neighbour lists: CN list for every atom (CN and weights are replicated on every rank), disp2/disp3 for own rows
  const int* d_ident = E->d_ident.p;
  // disp.f90:128 walks the disp3 translation list with the cn cutoff; when that list is a superset of the cn list the
  // accepted set equals the plain cn list, otherwise (disp3 < cn under PBC) images are dropped exactly like the reference.
  const double cn_list_cut = std::max(cutcn, eeq_cut);
  if (build_list(E, E->lcn, cn_list_cut, 1.0e-12, d_ident, nat)) return 1;
  if (do2 && build_list(E, E->l2, cut2, dbl_eps, E->d_rows.p, nrows)) return 1;
  // disp3 rows are refined along a Z-curve inside the cells: consecutive entries (the edges of one warp of the ATM stream)
  // then form a compact patch, which is what the warp-level acceptance of a third atom depends on
  static const int fine3 = [] { const char* v = std::getenv("D4CU_SORT_FINE"); return v ? std::max(0, std::min(3, std::atoi(v))) : 2; }();
  if (do3 && build_list(E, E->l3, cut3, dbl_eps, d_rows3, nrows3, fine3)) return 1;
  stage_end(E, ST_LISTS);

  if (stage_cn(E)) return 1;

  // ---- accumulators: [e_atom | dEdcn | dEdq | grad | sig_part] ----
  CUDA_OK(E->e_atom.ensure(nat)); CUDA_OK(E->dEdcn.ensure(nat)); CUDA_OK(E->dEdq.ensure(nat)); CUDA_OK(E->grad.ensure(3 * (size_t)nat));
  CUDA_OK(E->sig_part.ensure(6 * (size_t)nat)); CUDA_OK(E->lambda.ensure(nat)); CUDA_OK(E->w2.ensure(nat)); CUDA_OK(E->out.ensure(16));
  CUDA_OK(cudaMemsetAsync(E->e_atom.p, 0, sizeof(double) * nat, st));
  CUDA_OK(cudaMemsetAsync(E->dEdcn.p, 0, sizeof(double) * nat, st));
  CUDA_OK(cudaMemsetAsync(E->dEdq.p, 0, sizeof(double) * nat, st));
  CUDA_OK(cudaMemsetAsync(E->grad.p, 0, sizeof(double) * 3 * nat, st));
  CUDA_OK(cudaMemsetAsync(E->sig_part.p, 0, sizeof(double) * 6 * nat, st));
  CUDA_OK(cudaMemsetAsync(E->w2.p, 0, sizeof(double) * nat, st));

  // ---- two-body ----
  auto run_pair = [&]() -> int {
    stage_begin(E, ST_PAIR);
  if (do2 && nrows > 0) {
    const size_t sh = sizeof(double) * PAIR_WARPS * 3 * nid * D4_MREF;
    const int nb = (nrows + PAIR_WARPS - 1) / PAIR_WARPS;
    const bool smooth = w2 > 0.0;
#define LAUNCH_PAIR(G, S)                                                                                              \
  pair_kernel<G, S><<<nb, PAIR_WARPS * 32, sh, st>>>(nrows, E->d_rows.p, E->d_xyz.p, E->d_id.p, E->l2.pts.p, E->l2.offsets.p, \
                                                     E->l2.idx.p, E->tb, E->gw.p, E->gwdcn.p, E->gwdq.p, s6, s8, cut2, w2,     \
                                                     E->e_atom.p, E->dEdcn.p, E->dEdq.p, E->grad.p, E->sig_part.p, E->counters.p, \
                                                     d_e2, nat)
    if (sh > 48 * 1024) {
      cudaFuncSetAttribute(pair_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
      cudaFuncSetAttribute(pair_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
      cudaFuncSetAttribute(pair_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
      cudaFuncSetAttribute(pair_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
    }
    if (grad) { if (smooth) LAUNCH_PAIR(true, true); else LAUNCH_PAIR(true, false); }
    else { if (smooth) LAUNCH_PAIR(false, true); else LAUNCH_PAIR(false, false); }
#undef LAUNCH_PAIR
    ++E->launches;
  }
    stage_end(E, ST_PAIR);
    return 0;
  };

  // ---- charge chain rule (adjoint EEQ) ----
  auto run_adjoint = [&]() -> int {
    stage_begin(E, ST_ADJ);
  if (grad) {
    if (allreduce(E, E->dEdq.p, nat)) return 1;
    if (!owner || E->rank == 0) {   // (owner mode: the factor lives on rank 0)
      eeq_pack_rhs_kernel<<<(ld + 255) / 256, 256, 0, st>>>(nat, ld, E->dEdq.p, E->rhs.p);
      if (cholesky_solve_fused<1>(E->A.p, ld, E->nblk, E->invL.p, E->rhs.p, E->sol.p, ld, E->ctl1.p, st) < 0)
        return fail("fused substitution kernel cannot run on this device");
      E->launches += 2;
      eeq_lambda_kernel<<<1, 1024, 0, st>>>(nat, E->sol.p, E->zvec.p, E->aux.p, E->dxdcn_cut.p, E->lambda.p, E->w2.p);
      ++E->launches;
    }
    if (owner) {
      if (g_nccl.GroupStart() != 0) return fail("ncclGroupStart failed");
      if (g_nccl.Broadcast(E->lambda.p, E->lambda.p, (size_t)nat, /*ncclFloat64*/ 8, 0, E->comm, st) != 0 ||
          g_nccl.Broadcast(E->w2.p, E->w2.p, (size_t)nat, /*ncclFloat64*/ 8, 0, E->comm, st) != 0 ||
          g_nccl.Broadcast(E->ctl1.p + 2 * (size_t)E->nblk, E->ctl1.p + 2 * (size_t)E->nblk, 1, /*ncclInt32*/ 2, 0, E->comm, st) != 0)
        return fail("ncclBroadcast(adjoint charges) failed");
      if (g_nccl.GroupEnd() != 0) return fail("ncclGroupEnd failed");
    }
    if (E->any_periodic) {
      CUDA_OK(E->sfac.ensure(4 * EEQ_NH + 8));
      eeq_sfac_kernel<<<E->ew.nh + 1, 256, 0, st>>>(nat, E->ew, E->P.p, E->q.p, E->lambda.p, E->sfac.p);
      if (nrows > 0)
        eeq_grad_rec_kernel<<<(nrows * 32 + 255) / 256, 256, 0, st>>>(nrows, E->d_rows.p, E->d_id.p, E->ew, E->P.p, E->q.p, E->lambda.p,
                                                                     E->sfac.p, E->rank == 0 ? 1 : 0, E->grad.p, E->sig_part.p);
      const long long nt = (nat + EEQ_TILE - 1) / EEQ_TILE, ntl = nt * (nt + 1) / 2;
      const long long mine = (ntl - E->rank + E->world - 1) / E->world;  // tiles rank, rank + world, ...
      const size_t sh = eeq_dir_smem_bytes(E->ew.ndir, E->ew.nwsc, E->ew.tab_n);
      if (mine > 0)
        eeq_grad_dir_kernel<<<(unsigned)mine, EEQ_TILE * EEQ_TILE, sh, st>>>(nat, E->rank, E->world, E->d_xyz.p, E->d_id.p, E->tb, E->ew,
                                                                            E->q.p, E->lambda.p, E->grad.p, E->sig_part.p);
      E->launches += 3;
    } else if (nrows > 0) {
      eeq_grad_kernel<<<nrows, EEQG_THREADS, 0, st>>>(nrows, E->d_rows.p, nat, E->d_xyz.p, E->d_id.p, E->tb, E->q.p,
                                                           E->lambda.p, E->grad.p, E->sig_part.p);
      ++E->launches;
    }
  }
    stage_end(E, ST_ADJ);
    return 0;
  };

  // ---- three-body ----
  auto run_atm = [&]() -> int {
    stage_begin(E, ST_ATM);
  if (do3 && nrows3 > 0) {
    AtmParams ap;
    ap.s9 = s9; ap.a1 = E->param[3]; ap.a2 = E->param[4]; ap.alp = alp; ap.cutoff = cut3; ap.cutoff2 = cut3 * cut3;
    ap.width = w3; ap.eps2 = dbl_eps;
    for (int c = 0; c < 9; ++c) ap.lat[c] = E->any_periodic ? E->lattice[c] : 0.0;
    for (int c = 0; c < 3; ++c) ap.rep[c] = E->l3.tr.rep[c];
    const int tsp = (E->ts + D4_MREF + 1) & ~1;  // row stride of the precontracted C6 rows (see c6_rows_kernel)
    CUDA_OK(E->W0.ensure((size_t)nat * tsp)); CUDA_OK(E->W0cn.ensure((size_t)nat * tsp));
    const bool a16 = alp == 16.0, smooth3 = w3 > 0.0, d4s = E->tb.nsp > 1;
    c6_rows_kernel<<<nat, 64, 0, st>>>(nat, E->ts, tsp, E->d_id.p, E->tb, E->gw0.p, E->gwdcn0.p, E->W0.p, E->W0cn.p,
                                       (a16 && !d4s) ? 1 : 0);
    const size_t tot = (size_t)std::max<long long>(E->l3.total, 1);
    CUDA_OK(E->idx3own.ensure(tot)); CUDA_OK(E->nown3.ensure(nrows3)); CUDA_OK(E->eacc.ensure(2 * tot));
    CUDA_OK(E->nblk3.ensure(nrows3 + 1)); CUDA_OK(E->blkoff3.ensure(nrows3 + 2));
    row_own_partition_kernel<<<nrows3, 256, 0, st>>>(nrows3, d_rows3, E->l3.pts.p, E->l3.offsets.p, E->l3.idx.p,
                                                    E->idx3own.p, E->nown3.p, E->nblk3.p, 32 * ATM_NJ);
    E->launches += 2;
    {
      // ---- records -> autonomous warps over (row, block of owned edges) items -> per-edge finalisation ----
      if (E->l3.tr.count() >= (1 << 20)) return fail("more than 2^20 lattice translations inside the three-body cutoff");
      CUDA_OK(E->rec3.ensure(tot * (sizeof(AtmRec) / sizeof(double)))); CUDA_OK(E->atom3.ensure(tot)); CUDA_OK(E->work3.ensure(1));
      CUDA_OK(cudaMemsetAsync(E->nblk3.p + nrows3, 0, sizeof(int), st));
      size_t tmp_bytes = 0;
      cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, E->nblk3.p, E->blkoff3.p, nrows3 + 1, st);
      CUDA_OK(E->l3.cub_tmp.ensure(tmp_bytes));
      CUDA_OK(cub::DeviceScan::ExclusiveSum(E->l3.cub_tmp.p, tmp_bytes, E->nblk3.p, E->blkoff3.p, nrows3 + 1, st));
      CUDA_OK(cudaMemsetAsync(E->eacc.p, 0, sizeof(double) * 2 * tot, st));
      CUDA_OK(cudaMemsetAsync(E->work3.p, 0, sizeof(int), st));
      AtmRec* rec = reinterpret_cast<AtmRec*>(E->rec3.p);
      if (smooth3)
        atm_prep_kernel<true><<<nrows3, 256, 0, st>>>(nrows3, d_rows3, E->d_xyz.p, E->d_id.p, E->l3.pts.p, E->l3.offsets.p, E->idx3own.p,
                                                     E->tb, E->gw0.p, E->W0.p, tsp, cut3, cut3 * cut3, w3, rec, E->atom3.p);
      else
        atm_prep_kernel<false><<<nrows3, 256, 0, st>>>(nrows3, d_rows3, E->d_xyz.p, E->d_id.p, E->l3.pts.p, E->l3.offsets.p, E->idx3own.p,
                                                      E->tb, E->gw0.p, E->W0.p, tsp, cut3, cut3 * cut3, w3, rec, E->atom3.p);
      const int warps = ATM_WARPS, minb = ATM_MINB;
      // shared memory: every warp owns ATM_STAGES chunks of kc (record + C6 row); kc from the budget of the CTAs per SM
      static const int kc_env = [] { const char* v = std::getenv("D4CU_ATM_KC"); return v ? std::atoi(v) : 0; }();
      int kc = 0;
      for (int mb = minb; mb >= 1 && kc < 4; --mb) {
        const size_t budget = ((size_t)227 * 1024 / mb - 2048) / warps;
        kc = (int)std::min<size_t>(ATM_KC_MAX, budget / atm_warp_smem_bytes(tsp, 1));
      }
      if (kc < 1) return fail("too many species/reference systems for the ATM kernel's shared-memory stream");
      if (kc_env > 0) kc = std::min(kc, kc_env);
      const size_t sh = warps * atm_warp_smem_bytes(tsp, kc);
      int sms = 0;
      CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, E->device));
#define LAUNCH_ATMS_NJ(S, A, D, G, J, W, B)                                                                           \
  do {                                                                                                                \
    cudaFuncSetAttribute(atm_stream_kernel<S, A, D, G, J, W, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh); \
    int per_sm = 1;                                                                                                   \
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, atm_stream_kernel<S, A, D, G, J, W, B>, W * 32, sh);        \
    atm_stream_kernel<S, A, D, G, J, W, B><<<sms * std::max(per_sm, 1), W * 32, sh, st>>>(nrows3, d_rows3,          \
        E->d_xyz.p, E->d_id.p, E->l3.offsets.p, E->nown3.p, E->blkoff3.p, rec, E->atom3.p, E->l3.d_triples.p, E->tb,  \
        E->gw0.p, E->W0.p, ap, tsp, kc, reinterpret_cast<double2*>(E->eacc.p), E->counters.p, E->work3.p);            \
  } while (0)
#define LAUNCH_ATMS(S, A, D)                                                                                          \
  do {                                                                                                                \
    if (grad) LAUNCH_ATMS_NJ(S, A, D, true, ATM_NJ, ATM_WARPS, ATM_MINB);                                             \
    else LAUNCH_ATMS_NJ(S, A, D, false, ATM_NJ, ATM_WARPS, ATM_MINB);                                                 \
  } while (0)
      if (d4s) { if (smooth3) LAUNCH_ATMS(true, false, true); else LAUNCH_ATMS(false, false, true); }
      else if (smooth3) { if (a16) LAUNCH_ATMS(true, true, false); else LAUNCH_ATMS(true, false, false); }
      else { if (a16) LAUNCH_ATMS(false, true, false); else LAUNCH_ATMS(false, false, false); }
#undef LAUNCH_ATMS_NJ
#undef LAUNCH_ATMS
      if (smooth3)
        atm_finalize_kernel<true><<<nrows3, 256, 0, st>>>(nrows3, d_rows3, E->d_id.p, E->l3.offsets.p, E->nown3.p, rec, E->tb, E->gw0.p,
            E->gwdcn0.p, E->W0.p, E->W0cn.p, tsp, s9, cut3, w3, reinterpret_cast<const double2*>(E->eacc.p), E->e_atom.p,
            E->dEdcn.p, E->grad.p, E->sig_part.p, grad ? 1 : 0, d_e3, nat);
      else
        atm_finalize_kernel<false><<<nrows3, 256, 0, st>>>(nrows3, d_rows3, E->d_id.p, E->l3.offsets.p, E->nown3.p, rec, E->tb, E->gw0.p,
            E->gwdcn0.p, E->W0.p, E->W0cn.p, tsp, s9, cut3, w3, reinterpret_cast<const double2*>(E->eacc.p), E->e_atom.p,
            E->dEdcn.p, E->grad.p, E->sig_part.p, grad ? 1 : 0, d_e3, nat);
      E->launches += 4;
    }
  }
    stage_end(E, ST_ATM);
    return 0;
  };

  // ---- order of the stages ----
  // one GPU, or the EEQ system on all ranks: CN -> EEQ -> weights -> pair -> adjoint -> three-body.
  // owner mode: the other ranks run their three-body rows (which only need the coordination numbers) while rank 0
  // builds, factors and solves and then runs its smaller three-body share; q, and later the adjoint charges, reach
  // the other ranks by broadcast.
  if (owner) {
    if (stage_weights(E, false)) return 1;            // q = 0 weights: all the three-body stage needs
    if (E->rank != 0 && run_atm()) return 1;
    if (stage_eeq(E)) return 1;                       // (work on rank 0 only)
    if (E->rank == 0 && run_atm()) return 1;          // rank 0's smaller share, sized to end with the others
    if (eeq_owner_publish(E)) return 1;               // q + the factorisation / substitution verdicts
    if (stage_weights(E, true)) return 1;
    if (run_pair()) return 1;
    if (grad && run_adjoint()) return 1;
  } else {
    if (stage_eeq(E)) return 1;
    if (stage_weights(E, true)) return 1;
    if (run_pair()) return 1;
    if (grad && run_adjoint()) return 1;
    if (run_atm()) return 1;
  }

  // ---- CN chain rule ----
  stage_begin(E, ST_CHAIN);
  if (grad) {
    if (allreduce(E, E->dEdcn.p, nat)) return 1;
    // reference quirk (disp.f90:124-128): translations of the disp3 list, cutoff cn.  With disp3's repetitions covering
    // cn's this is the plain cn list; otherwise restrict to the images the disp3 list contains.
    double cn_eff2 = cutcn * cutcn;
    const int* d_tri = nullptr;
    int rep3[3] = {0, 0, 0};
    if (E->any_periodic) {
      d4::Translations t3;
      d4::lattice_points_cutoff(E->lattice, true, cut3, t3);
      const auto& rc = E->lcn.tr.rep;
      if (t3.rep[0] < rc[0] || t3.rep[1] < rc[1] || t3.rep[2] < rc[2]) {
        d_tri = E->lcn.d_triples.p;
        rep3[0] = t3.rep[0]; rep3[1] = t3.rep[1]; rep3[2] = t3.rep[2];
      }
    }
    // own rows only; the CN list is indexed by atom id
    if (nrows > 0) {
      cn_deriv_kernel<<<(nrows * 32 + 255) / 256, 256, 0, st>>>(nrows, E->d_rows.p, E->d_xyz.p, E->d_id.p, E->lcn.pts.p, E->lcn.offsets.p,
                                                               E->lcn.idx.p, E->tb, cn_eff2, eeq_cut * eeq_cut, kcn, E->dEdcn.p, E->w2.p,
                                                               E->lcn.d_trans.p, d_tri, rep3[0], rep3[1], rep3[2], E->grad.p, E->sig_part.p);
      ++E->launches;
    }
  }
  final_reduce_kernel<<<1, 1024, 0, st>>>(nat, E->e_atom.p, E->sig_part.p, nullptr, E->out.p);
  ++E->launches;
  {
    const unsigned long long p3 = (unsigned long long)(do3 ? E->l3.total : 0);
    CUDA_OK(cudaMemcpyAsync(E->counters.p + 3, &p3, sizeof(p3), cudaMemcpyHostToDevice, st));
  }
  if (E->world > 1) {
    if (g_nccl.AllReduce(E->counters.p + 2, E->counters.p + 2, 4, /*ncclUint64*/ 5, 0, E->comm, st) != 0)
      return fail("ncclAllReduce(counters) failed");
    if (allreduce(E, E->out.p, 10)) return 1;
    if (grad && allreduce(E, E->grad.p, 3 * (size_t)nat)) return 1;
  }
  stage_end(E, ST_CHAIN);
  CUDA_OK(cudaEventRecord(E->ev[10], st));
  int subst_err = 0;
  if (grad) CUDA_OK(cudaMemcpyAsync(&subst_err, E->ctl1.p + 2 * (size_t)E->nblk, sizeof(int), cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  CUDA_OK(cudaGetLastError());
  if (eeq_check_status(E)) return 1;
  if (subst_err != 0) return fail("EEQ adjoint substitution kernel timed out waiting for a block row (internal error)");
  for (int k = 0; k < 9; ++k) {
    E->ms[k] = 0.0f;
    if (E->st_used[k]) cudaEventElapsedTime(&E->ms[k], E->evA[k], E->evB[k]);
  }
  cudaEventElapsedTime(&E->ms[9], E->ev[0], E->ev[10]);
  // owner mode: rebalance the three-body rows for the next evaluation from this one's stage times -- rank 0 should
  // finish its EEQ work + three-body share when the others finish theirs
  if (owner && do3) {
    CUDA_OK(E->bal.ensure(E->world + 1));
    std::vector<double> h(E->world + 1, 0.0);
    h[E->rank] = E->ms[ST_ATM];
    if (E->rank == 0) h[E->world] = E->ms[ST_BUILD] + E->ms[ST_SOLVE];
    CUDA_OK(cudaMemcpyAsync(E->bal.p, h.data(), sizeof(double) * h.size(), cudaMemcpyHostToDevice, st));
    if (allreduce(E, E->bal.p, h.size())) return 1;
    CUDA_OK(cudaMemcpyAsync(h.data(), E->bal.p, sizeof(double) * h.size(), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    double atm_total = 0.0;
    for (int r = 0; r < E->world; ++r) atm_total += h[r];
    const double t_eeq = h[E->world];
    if (atm_total > 0.0) {
      const double a0 = std::max(0.0, (atm_total - (E->world - 1) * t_eeq) / E->world);   // rank 0's share of the ATM time
      const double f0 = std::min(a0 / atm_total, 1.0 / E->world);
      int v0 = (int)std::lround(f0 * 16.0 * (E->world - 1) / std::max(1e-12, 1.0 - f0));
      v0 = std::max(0, std::min(16, v0));
      E->atm_v0 = v0;   // (takes effect at the next compute: every rank computed the same value)
    }
  }
  E->computed = true;
  return 0;
}

int d4cu_compute(d4cu_engine* E, int want_gradient) {
  if (!E) return fail("engine is NULL");
  std::lock_guard<std::mutex> lk(E->mtx);
  return compute_impl(E, want_gradient, nullptr, nullptr);
}

int d4cu_pairwise(d4cu_engine* E, double* e2, double* e3) {
  if (!E) return fail("engine is NULL");
  std::lock_guard<std::mutex> lk(E->mtx);
  if (!E->uploaded) return fail("d4cu_pairwise called before d4cu_upload");
  if (E->world > 1) return fail("pair-resolved energies are a single-GPU property evaluation");
  CUDA_OK(cudaSetDevice(E->device));
  const size_t nn = (size_t)E->nat * E->nat;
  DBuf<double> d2, d3;
  CUDA_OK(d2.ensure(nn));
  CUDA_OK(d3.ensure(nn));
  CUDA_OK(cudaMemsetAsync(d2.p, 0, sizeof(double) * nn, E->st));
  CUDA_OK(cudaMemsetAsync(d3.p, 0, sizeof(double) * nn, E->st));
  if (compute_impl(E, 0, d2.p, d3.p)) return 1;
  if (e2) CUDA_OK(cudaMemcpyAsync(e2, d2.p, sizeof(double) * nn, cudaMemcpyDeviceToHost, E->st));
  if (e3) CUDA_OK(cudaMemcpyAsync(e3, d3.p, sizeof(double) * nn, cudaMemcpyDeviceToHost, E->st));
  CUDA_OK(cudaStreamSynchronize(E->st));
  return 0;
}

// Real-space Ewald images (alpha r <= 6) the periodic EEQ matrix build of the last compute summed, over all pairs i > j.
int d4cu_last_ewald_images(d4cu_engine* E, long long* images) {
  if (!E || !images) return fail("engine or output is NULL");
  std::lock_guard<std::mutex> lk(E->mtx);
  if (!E->computed) return fail("d4cu_last_ewald_images called before d4cu_compute");
  CUDA_OK(cudaSetDevice(E->device));
  unsigned long long v = 0;
  CUDA_OK(cudaMemcpy(&v, E->counters.p + 5, sizeof(v), cudaMemcpyDeviceToHost));
  *images = (long long)v;
  return 0;
}

int d4cu_download(d4cu_engine* E, double* energy, double* gradient, double* sigma, d4cu_counts* counts) {
  if (!E) return fail("engine is NULL");
  std::lock_guard<std::mutex> lk(E->mtx);
  if (!E->computed) return fail("d4cu_download called before d4cu_compute");
  CUDA_OK(cudaSetDevice(E->device));
  double out[10];
  CUDA_OK(cudaMemcpyAsync(out, E->out.p, sizeof(out), cudaMemcpyDeviceToHost, E->st));
  if (gradient) {
    if (!E->want_grad) return fail("gradient requested but the last compute was energy-only");
    CUDA_OK(cudaMemcpyAsync(gradient, E->grad.p, sizeof(double) * 3 * (size_t)E->nat, cudaMemcpyDeviceToHost, E->st));
  }
  unsigned long long c[8];
  CUDA_OK(cudaMemcpyAsync(c, E->counters.p, sizeof(c), cudaMemcpyDeviceToHost, E->st));
  CUDA_OK(cudaStreamSynchronize(E->st));
  if (energy) *energy = out[0];
  if (sigma) {
    if (!E->want_grad) return fail("sigma requested but the last compute was energy-only");
    std::memcpy(sigma, out + 1, 9 * sizeof(double));
  }
  if (counts) {
    counts->pcn = (long long)c[0]; counts->peeq = (long long)c[1]; counts->p2 = (long long)c[2];
    counts->p3 = (long long)c[3]; counts->t3 = (long long)c[4];
  }
  return 0;
}

int d4cu_evaluate(d4cu_engine* E, int nat, const int* id, const double* xyz, double charge, const double* lattice,
                  const int* periodic, const d4cu_model* model, const double* param, const double* cutoff, double* energy,
                  double* gradient, double* sigma, d4cu_counts* counts) {
  if (d4cu_upload(E, nat, id, xyz, charge, lattice, periodic, model, param, cutoff)) return 1;
  if (d4cu_compute(E, (gradient || sigma) ? 1 : 0)) return 1;
  return d4cu_download(E, energy, gradient, sigma, counts);
}

int d4cu_batch_evaluate(d4cu_engine* E, int nmol, const long long* offsets, const int* id, const double* xyz,
                        const double* charges, const d4cu_model* model, const double* param, const double* cutoff,
                        double* energies, double* gradient, double* sigma, int* status) {
  if (!E) return fail("engine is NULL");
  if (nmol <= 0) return fail("batch has no molecules");
  std::lock_guard<std::mutex> lk(E->mtx);
  CUDA_OK(cudaSetDevice(E->device));
  cudaStream_t st = E->st;
  const long long total = offsets[nmol];
  for (int m = 0; m < nmol; ++m) {
    const long long n = offsets[m + 1] - offsets[m];
    if (n < 1 || n > BATCH_MAXAT)
      return fail("batched evaluation supports 1.." + std::to_string(BATCH_MAXAT) + " atoms per molecule (molecule " +
                  std::to_string(m) + " has " + std::to_string(n) + ")");
  }
  if (cutoff[3] > 0.0 || cutoff[4] > 0.0) return fail("batched evaluation supports sharp cutoffs only");
  if (model->wfpair) return fail("batched evaluation implements the D4 model only (D4S: use the per-structure API)");
  if (model->nid > 255) return fail("more than 255 species");
  for (long long a = 0; a < total; ++a)
    if (id[a] < 0 || id[a] >= model->nid) return fail("species index out of range");
  std::memcpy(E->param, param, sizeof(E->param));
  std::memcpy(E->cutoff, cutoff, sizeof(E->cutoff));
  if (upload_tables(E, model)) return 1;
  E->uploaded = false; E->computed = false;  // the table upload invalidates a resident single-structure state
  DBuf<long long> d_off;
  DBuf<int> d_idb, d_status;
  DBuf<double> d_x, d_q, d_e, d_g, d_s;
  CUDA_OK(d_off.ensure(nmol + 1)); CUDA_OK(d_idb.ensure(total)); CUDA_OK(d_status.ensure(nmol)); CUDA_OK(d_x.ensure(3 * total));
  CUDA_OK(d_e.ensure(nmol));
  const bool grad = gradient || sigma;
  if (grad) { CUDA_OK(d_g.ensure(3 * total)); CUDA_OK(d_s.ensure(9 * (size_t)nmol)); }
  if (charges) { CUDA_OK(d_q.ensure(nmol)); CUDA_OK(cudaMemcpyAsync(d_q.p, charges, sizeof(double) * nmol, cudaMemcpyHostToDevice, st)); }
  CUDA_OK(cudaMemcpyAsync(d_off.p, offsets, sizeof(long long) * (nmol + 1), cudaMemcpyHostToDevice, st));
  CUDA_OK(cudaMemcpyAsync(d_idb.p, id, sizeof(int) * total, cudaMemcpyHostToDevice, st));
  CUDA_OK(cudaMemcpyAsync(d_x.p, xyz, sizeof(double) * 3 * total, cudaMemcpyHostToDevice, st));
  BatchParams bp;
  bp.s6 = param[0]; bp.s8 = param[1]; bp.s9 = param[2]; bp.a1 = param[3]; bp.a2 = param[4]; bp.alp = param[5];
  bp.cut2_2 = cutoff[0] * cutoff[0]; bp.cut3_2 = cutoff[1] * cutoff[1]; bp.cutcn_2 = cutoff[2] * cutoff[2];
  const double dbl_eps = 2.220446049250313e-16;
  bp.do2 = !(std::fabs(param[0]) < dbl_eps && std::fabs(param[1]) < dbl_eps);
  bp.do3 = !(std::fabs(param[2]) < dbl_eps);
  bp.want_grad = grad ? 1 : 0;
  const size_t sh = batch_smem_bytes();
  cudaFuncSetAttribute(batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
  CUDA_OK(cudaEventRecord(E->ev[0], st));
  batch_kernel<<<nmol, BATCH_THREADS, sh, st>>>(nmol, d_off.p, d_idb.p, d_x.p, charges ? d_q.p : nullptr, E->tb, bp, d_e.p,
                                                grad ? d_g.p : nullptr, grad ? d_s.p : nullptr, d_status.p);
  CUDA_OK(cudaEventRecord(E->ev[10], st));
  E->launches = 1;
  CUDA_OK(cudaMemcpyAsync(energies, d_e.p, sizeof(double) * nmol, cudaMemcpyDeviceToHost, st));
  if (gradient) CUDA_OK(cudaMemcpyAsync(gradient, d_g.p, sizeof(double) * 3 * total, cudaMemcpyDeviceToHost, st));
  if (sigma) CUDA_OK(cudaMemcpyAsync(sigma, d_s.p, sizeof(double) * 9 * (size_t)nmol, cudaMemcpyDeviceToHost, st));
  std::vector<int> h_status(nmol, 0);
  CUDA_OK(cudaMemcpyAsync(h_status.data(), d_status.p, sizeof(int) * nmol, cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  CUDA_OK(cudaGetLastError());
  for (auto& v : E->ms) v = 0.0f;
  cudaEventElapsedTime(&E->ms[9], E->ev[0], E->ev[10]);
  int nbad = 0;
  for (int m = 0; m < nmol; ++m) { if (status) status[m] = h_status[m]; nbad += h_status[m] != 0; }
  if (nbad && !status) return fail("EEQ Coulomb matrix is not positive definite for " + std::to_string(nbad) + " molecule(s) of the batch");
  return 0;
}

int d4cu_last_timing(d4cu_engine* E, float* ms10, int* launches) {
  if (!E) return fail("engine is NULL");
  if (ms10) std::memcpy(ms10, E->ms, sizeof(E->ms));
  if (launches) *launches = E->launches;
  return 0;
}

long long d4cu_pair_list(d4cu_engine* E, int which, int* i_idx, int* j_idx, int* t_idx) {
  if (!E || !E->computed) { fail("no computed state"); return -1; }
  std::lock_guard<std::mutex> lk(E->mtx);
  cudaSetDevice(E->device);
  NList& L = which == 0 ? E->lcn : (which == 1 ? E->l2 : E->l3);
  const long long total = L.total;
  if (!i_idx) return total;
  std::vector<long long> off(L.nrows + 1);
  std::vector<unsigned> idx((size_t)std::max<long long>(total, 1));
  cudaMemcpy(off.data(), L.offsets.p, sizeof(long long) * (L.nrows + 1), cudaMemcpyDeviceToHost);
  if (total > 0) cudaMemcpy(idx.data(), L.idx.p, sizeof(unsigned) * total, cudaMemcpyDeviceToHost);
  std::vector<Point> pts((size_t)std::max<long long>(L.npts, 1));
  cudaMemcpy(pts.data(), L.pts.p, sizeof(Point) * L.npts, cudaMemcpyDeviceToHost);
  // rows the list was built for: every atom (cn), this rank's rows (disp2), and for disp3 the owner-mode share when that
  // mode dealt the three-body rows separately
  const std::vector<int>* rows = which == 0 ? nullptr : (which == 2 && eeq_owner_mode(E) ? &E->h_rows3 : &E->h_rows);
  if (rows && (size_t)L.nrows > rows->size()) { fail("pair list and row set disagree"); return -1; }
  for (int r = 0; r < L.nrows; ++r)
    for (long long e = off[r]; e < off[r + 1]; ++e) {
      const unsigned long long w = pts[idx[e]].w;
      i_idx[e] = rows ? (*rows)[r] : r;
      j_idx[e] = (int)(w & 0xffffffffull);
      t_idx[e] = (int)((w >> 40) & 0xffffff);
    }
  return total;
}

int d4cu_properties(d4cu_engine* E, double* cn, double* q, double* c6, double* alpha) {
  if (!E) return fail("engine is NULL");
  std::lock_guard<std::mutex> lk(E->mtx);
  if (!E->uploaded) return fail("d4cu_properties called before d4cu_upload");
  CUDA_OK(cudaSetDevice(E->device));
  cudaStream_t st = E->st;
  const int nat = E->nat;
  E->launches = 0;
  CUDA_OK(cudaEventRecord(E->ev[0], st));
  const double cn_list_cut = std::max(E->cutoff[2], 25.0);
  if (build_list(E, E->lcn, cn_list_cut, 1.0e-12, E->d_ident.p, nat)) return 1;
  CUDA_OK(cudaEventRecord(E->ev[1], st));
  if (stage_cn_eeq_weights(E)) return 1;
  DBuf<double> d_alpha, d_c6;
  CUDA_OK(d_alpha.ensure(nat));
  alpha_kernel<<<(nat + 127) / 128, 128, 0, st>>>(nat, E->d_id.p, E->tb, E->gw.p, d_alpha.p);
  ++E->launches;
  if (c6) {
    CUDA_OK(d_c6.ensure((size_t)nat * nat));
    dim3 blk(16, 16), grd((nat + 15) / 16, (nat + 15) / 16);
    c6_matrix_kernel<<<grd, blk, 0, st>>>(nat, E->d_id.p, E->tb, E->gw.p, d_c6.p);
    ++E->launches;
    CUDA_OK(cudaMemcpyAsync(c6, d_c6.p, sizeof(double) * (size_t)nat * nat, cudaMemcpyDeviceToHost, st));
  }
  if (cn) CUDA_OK(cudaMemcpyAsync(cn, E->cn.p, sizeof(double) * nat, cudaMemcpyDeviceToHost, st));
  if (q) CUDA_OK(cudaMemcpyAsync(q, E->q.p, sizeof(double) * nat, cudaMemcpyDeviceToHost, st));
  if (alpha) CUDA_OK(cudaMemcpyAsync(alpha, d_alpha.p, sizeof(double) * nat, cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  CUDA_OK(cudaGetLastError());
  return 0;
}

int d4cu_measure_fp64_peak(int device, double* dfma_tflops, double* dmma_tflops) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return fail("no CUDA device available");
  CUDA_OK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_OK(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
  double* d_out = nullptr;
  CUDA_OK(cudaMalloc(&d_out, sizeof(double) * blocks * threads));
  cudaEvent_t a, b;
  CUDA_OK(cudaEventCreate(&a));
  CUDA_OK(cudaEventCreate(&b));
  double best[2] = {0, 0};
  for (int which = 0; which < 2; ++which)
    for (int rep = 0; rep < 4; ++rep) {
      CUDA_OK(cudaEventRecord(a));
      if (which == 0) dfma_peak_kernel<<<blocks, threads>>>(d_out, iters, 0.999999, 1e-9);
      else dmma_peak_kernel<<<blocks, threads>>>(d_out, iters, 0.999999, 1e-9);
      CUDA_OK(cudaEventRecord(b));
      CUDA_OK(cudaEventSynchronize(b));
      float ms = 0;
      CUDA_OK(cudaEventElapsedTime(&ms, a, b));
      // DFMA: 8 FMA (2 flop) per thread per iteration; DMMA: 8 tiles x (8*8*4*2 flop) per warp per iteration
      const double flop = which == 0 ? 2.0 * 8 * iters * (double)blocks * threads
                                     : 8.0 * 512 * iters * (double)blocks * (threads / 32);
      if (rep > 0) best[which] = std::max(best[which], flop / (ms * 1e-3) / 1e12);
    }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(d_out);
  CUDA_OK(cudaGetLastError());
  if (dfma_tflops) *dfma_tflops = best[0];
  if (dmma_tflops) *dmma_tflops = best[1];
  return 0;
}

// ---- d4cu_batch_noalloc: the batched kernel for callers that cannot allocate device memory (include/d4cu.h) --------
// Everything d4cu_batch_evaluate keeps in DBufs it allocates (device copy of the offsets, model tables, a status array
// when the caller has none) lives in one fixed `static __device__` array instead.  On the host the array's name is only
// a handle: the kernels get the device address cudaGetSymbolAddress returns, looked up once per device below.
}  // extern "C"

namespace {

constexpr size_t na_round(size_t b) { return (b + 255) / 256 * 256; }
constexpr int NA_MAX_MOL = D4CU_NOALLOC_MAX_MOLECULES, NA_MAX_SP = D4CU_NOALLOC_MAX_SPECIES;
// doubles: zeff eta chi eeta kcn rad [S] | refcn refq refalpha0 [7S] | refc6 [49 S^2] | pair_rc den rr r0 gam [S^2]
constexpr size_t NA_TAB_D = 6 * (size_t)NA_MAX_SP + 21 * (size_t)NA_MAX_SP + 54 * (size_t)NA_MAX_SP * NA_MAX_SP;
constexpr size_t NA_TAB_I = 9 * (size_t)NA_MAX_SP + 1;   // nref [S] | ngw [7S] | soff [S+1]
constexpr size_t NA_OFF_FLAGS = 0;                       // unsigned[2]: failed molecules, species indices out of range
constexpr size_t NA_OFF_OFFSETS = 256;
constexpr size_t NA_OFF_STATUS = NA_OFF_OFFSETS + na_round(sizeof(long long) * (NA_MAX_MOL + 1));
constexpr size_t NA_OFF_TAB_I = NA_OFF_STATUS + na_round(sizeof(int) * NA_MAX_MOL);
constexpr size_t NA_OFF_TAB_D = NA_OFF_TAB_I + na_round(sizeof(int) * NA_TAB_I);
constexpr size_t NA_BYTES = NA_OFF_TAB_D + na_round(sizeof(double) * NA_TAB_D);
static_assert(NA_BYTES <= D4CU_NOALLOC_SCRATCH_BYTES, "d4cu.h states a smaller scratch array than the layout needs");

static __device__ __align__(256) unsigned char g_noalloc_scratch[D4CU_NOALLOC_SCRATCH_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_noalloc_mtx;                                   // one call at a time: all calls share the array
unsigned char* g_noalloc_addr[NA_MAX_DEVICES] = {nullptr};  // device address of g_noalloc_scratch, per device

__global__ void noalloc_check_species_kernel(long long total, const int* __restrict__ species, int nid, unsigned* __restrict__ flags) {
  const long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (a < total && (species[a] < 0 || species[a] >= nid)) atomicAdd(&flags[1], 1u);
}
__global__ void noalloc_count_failed_kernel(int nmol, const int* __restrict__ status, unsigned* __restrict__ flags) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m < nmol && status[m] != 0) atomicAdd(&flags[0], 1u);
}

struct DeviceRestore {   // put the caller's current device back on every return path
  int prev = -1;
  DeviceRestore() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
  ~DeviceRestore() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

extern "C" {

int d4cu_batch_noalloc(int device, void* stream, int nmol, const long long* offsets, int nspecies, const int* species_numbers,
                       const int* d_species, const double* d_xyz, const double* d_charges, double ga, double gc, double wf,
                       const double* param, const double* cutoff, double* d_energies, double* d_gradient, double* d_sigma,
                       int* d_status) {
  if (!offsets || !species_numbers || !d_species || !d_xyz || !param || !cutoff || !d_energies)
    return fail("d4cu_batch_noalloc: a required pointer is NULL");
  if ((d_gradient == nullptr) != (d_sigma == nullptr))
    return fail("d4cu_batch_noalloc: pass d_gradient and d_sigma together, or both NULL for energies only");
  if (nmol <= 0) return fail("batch has no molecules");
  if (nmol > NA_MAX_MOL)
    return fail("d4cu_batch_noalloc: " + std::to_string(nmol) + " molecules exceed the fixed scratch array (at most " +
                std::to_string(NA_MAX_MOL) + " per call; split the batch)");
  if (nspecies <= 0 || nspecies > NA_MAX_SP)
    return fail("d4cu_batch_noalloc: " + std::to_string(nspecies) + " species, the fixed scratch array holds tables for 1.." +
                std::to_string(NA_MAX_SP));
  if (offsets[0] != 0) return fail("d4cu_batch_noalloc: offsets[0] must be 0");
  for (int m = 0; m < nmol; ++m) {
    const long long n = offsets[m + 1] - offsets[m];
    if (n < 1 || n > BATCH_MAXAT)
      return fail("batched evaluation supports 1.." + std::to_string(BATCH_MAXAT) + " atoms per molecule (molecule " +
                  std::to_string(m) + " has " + std::to_string(n) + ")");
  }
  if (cutoff[3] > 0.0 || cutoff[4] > 0.0) return fail("batched evaluation supports sharp cutoffs only");
  if (!(wf > 0.0)) return fail("batched evaluation implements the D4 model only (D4S: use the per-structure API)");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail("no CUDA device available");
  if (device < 0 || device >= ndev || device >= NA_MAX_DEVICES) return fail("d4cu_batch_noalloc: no such device");
  const long long total = offsets[nmol];

  // per-species model tables on the host (host memory only), d4.f90:71-231
  d4::Structure uni;
  uni.species_z.assign(species_numbers, species_numbers + nspecies);
  d4::HostModel hm;
  std::string err;
  if (!hm.build(uni, ga, gc, wf, err)) return fail(err);
  const d4cu_model mv = hm.view();
  const int S = nspecies;
  std::vector<double> td;
  std::vector<int> ti;
  td.reserve(27 * (size_t)S + 54 * (size_t)S * S);
  auto push_d = [&](const double* p, size_t n) { const size_t o = td.size(); td.insert(td.end(), p, p + n); return o; };
  auto push_i = [&](const int* p, size_t n) { const size_t o = ti.size(); ti.insert(ti.end(), p, p + n); return o; };
  const size_t o_zeff = push_d(mv.zeff, S), o_eta = push_d(mv.eta, S), o_chi = push_d(mv.eeq_chi, S), o_eeta = push_d(mv.eeq_eta, S),
               o_kcn = push_d(mv.eeq_kcn, S), o_rad = push_d(mv.eeq_rad, S), o_refcn = push_d(mv.refcn, 7 * (size_t)S),
               o_refq = push_d(mv.refq, 7 * (size_t)S), o_alpha0 = push_d(mv.refalpha0, 7 * (size_t)S),
               o_refc6 = push_d(mv.refc6, 49 * (size_t)S * S);
  std::vector<double> pr(5 * (size_t)S * S);
  double *rc = pr.data(), *den = rc + S * S, *rr = den + S * S, *r0 = rr + S * S, *gam = r0 + S * S;
  const double k4 = 4.10451, k5 = 19.08857, k6 = 2.0 * 11.28174 * 11.28174;  // mctc-lib ncoord, dftd4 flavour
  for (int a = 0; a < S; ++a)
    for (int b = 0; b < S; ++b) {
      rc[a * S + b] = mv.rcov[a] + mv.rcov[b];
      const double dEN = std::fabs(mv.en[a] - mv.en[b]) + k5;
      den[a * S + b] = k4 * std::exp(-(dEN * dEN) / k6);
      rr[a * S + b] = 3 * mv.r4r2[a] * mv.r4r2[b];                      // rational.f90:283
      r0[a * S + b] = param[3] * std::sqrt(rr[a * S + b]) + param[4];   // rational.f90:284
      gam[a * S + b] = 1.0 / std::sqrt(mv.eeq_rad[a] * mv.eeq_rad[a] + mv.eeq_rad[b] * mv.eeq_rad[b]);
    }
  const size_t o_pair = push_d(pr.data(), pr.size());
  std::vector<int> soff(S + 1, 0);
  for (int s = 0; s < S; ++s) soff[s + 1] = soff[s] + mv.nref[s] + 1;
  const size_t o_nref = push_i(mv.nref, S), o_ngw = push_i(mv.ngw, 7 * (size_t)S), o_soff = push_i(soff.data(), S + 1);
  if (td.size() > NA_TAB_D || ti.size() > NA_TAB_I) return fail("d4cu_batch_noalloc: model tables exceed the scratch array");

  std::lock_guard<std::mutex> lk(g_noalloc_mtx);
  DeviceRestore restore;
  CUDA_OK(cudaSetDevice(device));
  if (!g_noalloc_addr[device]) {   // first call on this device
    void* p = nullptr;
    CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) return fail("d4cu_batch_noalloc: cudaGetSymbolAddress returned no address");
    CUDA_OK(cudaFuncSetAttribute(batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)batch_smem_bytes()));
    g_noalloc_addr[device] = static_cast<unsigned char*>(p);
  }
  unsigned char* base = g_noalloc_addr[device];
  unsigned* d_flags = reinterpret_cast<unsigned*>(base + NA_OFF_FLAGS);
  long long* d_off = reinterpret_cast<long long*>(base + NA_OFF_OFFSETS);
  int* d_stat = d_status ? d_status : reinterpret_cast<int*>(base + NA_OFF_STATUS);
  int* d_ti = reinterpret_cast<int*>(base + NA_OFF_TAB_I);
  double* D = reinterpret_cast<double*>(base + NA_OFF_TAB_D);
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  // the array still holds the previous call's data: counters start from zero, offsets and tables are overwritten
  CUDA_OK(cudaMemsetAsync(d_flags, 0, 2 * sizeof(unsigned), st));
  CUDA_OK(cudaMemcpyAsync(d_off, offsets, sizeof(long long) * (nmol + 1), cudaMemcpyHostToDevice, st));
  CUDA_OK(cudaMemcpyAsync(d_ti, ti.data(), sizeof(int) * ti.size(), cudaMemcpyHostToDevice, st));
  CUDA_OK(cudaMemcpyAsync(D, td.data(), sizeof(double) * td.size(), cudaMemcpyHostToDevice, st));
  noalloc_check_species_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(total, d_species, S, d_flags);
  unsigned h_flags[2] = {0, 0};
  CUDA_OK(cudaMemcpyAsync(h_flags, d_flags, sizeof(h_flags), cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  CUDA_OK(cudaGetLastError());
  if (h_flags[1]) return fail("species index out of range");

  Tables t{};
  t.nid = S; t.ga = ga; t.gc = gc; t.wf = wf;
  t.zeff = D + o_zeff; t.eta = D + o_eta; t.refcn = D + o_refcn; t.refq = D + o_refq; t.refc6 = D + o_refc6;
  t.refalpha0 = D + o_alpha0; t.chi = D + o_chi; t.eeta = D + o_eeta; t.kcn = D + o_kcn; t.rad = D + o_rad;
  t.pair_rc = D + o_pair; t.pair_den = t.pair_rc + S * S; t.pair_rr = t.pair_den + S * S; t.pair_r0 = t.pair_rr + S * S;
  t.pair_gam = t.pair_r0 + S * S;
  t.nref = d_ti + o_nref; t.ngw = d_ti + o_ngw; t.soff = d_ti + o_soff;
  t.nsp = 1; t.wfpair = nullptr;
  BatchParams bp;
  bp.s6 = param[0]; bp.s8 = param[1]; bp.s9 = param[2]; bp.a1 = param[3]; bp.a2 = param[4]; bp.alp = param[5];
  bp.cut2_2 = cutoff[0] * cutoff[0]; bp.cut3_2 = cutoff[1] * cutoff[1]; bp.cutcn_2 = cutoff[2] * cutoff[2];
  const double dbl_eps = 2.220446049250313e-16;
  bp.do2 = !(std::fabs(param[0]) < dbl_eps && std::fabs(param[1]) < dbl_eps);
  bp.do3 = !(std::fabs(param[2]) < dbl_eps);
  bp.want_grad = d_gradient ? 1 : 0;
  batch_kernel<<<nmol, BATCH_THREADS, batch_smem_bytes(), st>>>(nmol, d_off, d_species, d_xyz, d_charges, t, bp, d_energies,
                                                                d_gradient, d_sigma, d_stat);
  noalloc_count_failed_kernel<<<(nmol + 255) / 256, 256, 0, st>>>(nmol, d_stat, d_flags);
  CUDA_OK(cudaMemcpyAsync(h_flags, d_flags, sizeof(h_flags), cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  CUDA_OK(cudaGetLastError());
  if (h_flags[0] && !d_status)
    return fail("EEQ Coulomb matrix is not positive definite for " + std::to_string(h_flags[0]) + " molecule(s) of the batch");
  return 0;
}

}  // extern "C"
