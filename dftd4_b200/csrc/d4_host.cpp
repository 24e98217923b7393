// Host-side model construction, damping-parameter lookup and lattice translation lists.
// Compiled without FMA contraction (-ffp-contract=off) so that translation vectors and wrapped coordinates carry the
// same bits a stock gfortran build of the reference produces.
#include "d4_host.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace {
#include "d4_tables_data.inc"
#include "d4_wfpair_data.inc"
#include "eeq2019_params.inc"

constexpr double kPi = 3.14159265358979323846264338327950288;

// mctc-lib mctc_io_convert: CODATA-2018 Bohr radius in Angstrom
inline double aatoau() { return 1.0 / (5.29177210903e-11 * 1.0e10); }

// src/dftd4/model/utils.f90:45-59
inline double zeta_fn(double a, double c, double qref, double qmod) {
  return qmod < 0.0 ? std::exp(a) : std::exp(a * (1.0 - std::exp(c * (1.0 - qref / qmod))));
}

// Casimir-Polder trapezoid weights, src/dftd4/model/utils.f90:82-122
struct TrapzWeights {
  double w[d4::kNfreq];
  TrapzWeights() {
    const double f[d4::kNfreq] = {0.000001, 0.050000, 0.100000, 0.200000, 0.300000, 0.400000, 0.500000, 0.600000,
                                  0.700000, 0.800000, 0.900000, 1.000000, 1.200000, 1.400000, 1.600000, 1.800000,
                                  2.000000, 2.500000, 3.000000, 4.000000, 5.000000, 7.500000, 10.00000};
    for (int i = 0; i < d4::kNfreq; ++i) {
      const double lo = i > 0 ? f[i] - f[i - 1] : 0.0;
      const double hi = i + 1 < d4::kNfreq ? f[i + 1] - f[i] : 0.0;
      w[i] = 0.5 * (i == 0 ? hi : (i + 1 == d4::kNfreq ? lo : lo + hi));
    }
  }
};

inline void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - b[1] * a[2];
  c[1] = a[2] * b[0] - b[2] * a[0];
  c[2] = a[0] * b[1] - b[0] * a[1];
}
inline double len3(const double* a) { return std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }
inline double det3(const double* a) {
  return a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
}
}  // namespace

// ---- EEQ-2019 parameter rows as a run-time input --------------------------------------------------------------------
// multicharge's param/eeq2019.f90 is not part of the reference tree; the compiled table (eeq2019_params.inc) restates
// Z = 1..86.  Rows given at run time override the compiled ones and add the elements the table lacks (Z = 87..103):
// d4cu_set_eeq_params / d4cu_load_eeq_params, or a file named by the environment variable D4CU_EEQ_PARAMS (read on the
// first model build).  File format: one row "Z chi eta kcn rad" per line, '#' starts a comment.
namespace {
struct EeqOverride {
  std::mutex mtx;
  bool env_done = false;
  bool have[104] = {false};
  double row[104][4] = {{0}};
};
EeqOverride& eeq_override() {
  static EeqOverride o;
  return o;
}
int eeq_load_file_locked(EeqOverride& o, const char* path, std::string& err) {
  FILE* f = std::fopen(path, "r");
  if (!f) { err = std::string("cannot open EEQ parameter file '") + path + "'"; return 1; }
  char line[512];
  int lineno = 0, rows = 0;
  while (std::fgets(line, sizeof(line), f)) {
    ++lineno;
    if (char* h = std::strchr(line, '#')) *h = 0;
    int z = 0;
    double v[4];
    const int n = std::sscanf(line, "%d %lf %lf %lf %lf", &z, &v[0], &v[1], &v[2], &v[3]);
    if (n <= 0) continue;   // blank / comment
    if (n != 5 || z < 1 || z > 103 || !(v[3] > 0.0)) {
      std::fclose(f);
      err = std::string("EEQ parameter file '") + path + "': malformed row at line " + std::to_string(lineno) +
            " (expected: Z chi eta kcn rad, 1 <= Z <= 103, rad > 0)";
      return 1;
    }
    o.have[z] = true;
    for (int c = 0; c < 4; ++c) o.row[z][c] = v[c];
    ++rows;
  }
  std::fclose(f);
  if (rows == 0) { err = std::string("EEQ parameter file '") + path + "' holds no rows"; return 1; }
  return 0;
}
}  // namespace

namespace d4 {
int eeq_params_set(int z, const double* row4) {
  if (z < 1 || z > 103 || !row4 || !(row4[3] > 0.0)) return 1;
  EeqOverride& o = eeq_override();
  std::lock_guard<std::mutex> lk(o.mtx);
  o.have[z] = true;
  for (int c = 0; c < 4; ++c) o.row[z][c] = row4[c];
  return 0;
}
int eeq_params_load(const char* path, std::string& err) {
  EeqOverride& o = eeq_override();
  std::lock_guard<std::mutex> lk(o.mtx);
  return eeq_load_file_locked(o, path, err);
}
void eeq_params_clear() {
  EeqOverride& o = eeq_override();
  std::lock_guard<std::mutex> lk(o.mtx);
  for (int z = 0; z < 104; ++z) o.have[z] = false;
}
// Row in effect for element z: 0 ok, 1 no row (compiled table ends at EEQ2019_MAXZ and nothing was supplied), 2 the
// file named by D4CU_EEQ_PARAMS could not be read (err says why).
int eeq_params_get(int z, double* row4, std::string& err) {
  if (z < 1 || z > 103) return 1;
  EeqOverride& o = eeq_override();
  std::lock_guard<std::mutex> lk(o.mtx);
  if (!o.env_done) {
    o.env_done = true;
    if (const char* p = std::getenv("D4CU_EEQ_PARAMS"))
      if (*p && eeq_load_file_locked(o, p, err)) { o.env_done = false; return 2; }
  }
  if (o.have[z]) { for (int c = 0; c < 4; ++c) row4[c] = o.row[z][c]; return 0; }
  if (z > EEQ2019_MAXZ) return 1;
  row4[0] = eeq2019_chi[z]; row4[1] = eeq2019_eta[z]; row4[2] = eeq2019_kcn[z]; row4[3] = eeq2019_rad[z];
  return 0;
}
}  // namespace d4

namespace d4 {

// ---------------------------------------------------------------------------------------------------------------
void Structure::assign(int natoms, const int* nums, const double* pos, const double* chrg, const double* lat,
                       const bool* per) {
  nat = natoms;
  numbers.assign(nums, nums + natoms);
  xyz.assign(pos, pos + 3 * static_cast<size_t>(natoms));
  id.resize(natoms);
  species_z.clear();
  for (int i = 0; i < natoms; ++i) {
    auto it = std::find(species_z.begin(), species_z.end(), nums[i]);
    if (it == species_z.end()) {
      species_z.push_back(nums[i]);
      id[i] = static_cast<int>(species_z.size()) - 1;
    } else {
      id[i] = static_cast<int>(it - species_z.begin());
    }
  }
  charge = chrg ? *chrg : 0.0;  // api.f90:233-239
  lattice.fill(0.0);
  periodic = {{0, 0, 0}};
  if (lat) {
    std::copy(lat, lat + 9, lattice.begin());
    // a lattice without an explicit periodic flag array is fully periodic (mctc-lib `new`; pinned by the NH3
    // crystal golden of python/dftd4/test_interface.py:555-654)
    for (int k = 0; k < 3; ++k) periodic[k] = per ? (per[k] ? 1 : 0) : 1;
  }
  any_periodic = periodic[0] || periodic[1] || periodic[2];
  ++revision;
  wrap_to_central_cell();
}

void Structure::update(const double* pos, const double* lat) {
  std::copy(pos, pos + 3 * static_cast<size_t>(nat), xyz.begin());
  if (lat) std::copy(lat, lat + 9, lattice.begin());
  ++revision;
  wrap_to_central_cell();
}

void Structure::wrap_to_central_cell() {
  if (!any_periodic) return;
  // Fortran lattice(c,k) == lattice[3*k+c]; fractional = inverse * xyz; adjugate/determinant inverse (mctc matinv_3x3)
  double a[3][3], inv[3][3];
  for (int k = 0; k < 3; ++k)
    for (int c = 0; c < 3; ++c) a[c][k] = lattice[3 * k + c];
  const double det = a[0][0] * a[1][1] * a[2][2] - a[0][0] * a[1][2] * a[2][1] - a[0][1] * a[1][0] * a[2][2] +
                     a[0][1] * a[1][2] * a[2][0] + a[0][2] * a[1][0] * a[2][1] - a[0][2] * a[1][1] * a[2][0];
  const double di = 1.0 / det;
  inv[0][0] = +di * (a[1][1] * a[2][2] - a[1][2] * a[2][1]);
  inv[1][0] = -di * (a[1][0] * a[2][2] - a[1][2] * a[2][0]);
  inv[2][0] = +di * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
  inv[0][1] = -di * (a[0][1] * a[2][2] - a[0][2] * a[2][1]);
  inv[1][1] = +di * (a[0][0] * a[2][2] - a[0][2] * a[2][0]);
  inv[2][1] = -di * (a[0][0] * a[2][1] - a[0][1] * a[2][0]);
  inv[0][2] = +di * (a[0][1] * a[1][2] - a[0][2] * a[1][1]);
  inv[1][2] = -di * (a[0][0] * a[1][2] - a[0][2] * a[1][0]);
  inv[2][2] = +di * (a[0][0] * a[1][1] - a[0][1] * a[1][0]);
  constexpr double eps = 1.0e-14;
  for (int i = 0; i < nat; ++i) {
    double* x = &xyz[3 * static_cast<size_t>(i)];
    double f[3];
    for (int r = 0; r < 3; ++r) {
      const double v = inv[r][0] * x[0] + inv[r][1] * x[1] + inv[r][2] * x[2];
      double o = v;
      if (v < (0.0 - eps)) o = v + std::ceil(-v);
      if (v > (1.0 + eps)) o = v - std::floor(v);
      if (std::fabs(v - 1.0) < eps) o = v - 1.0;
      f[r] = o;
    }
    const double nx = a[0][0] * f[0] + a[0][1] * f[1] + a[0][2] * f[2];
    const double ny = a[1][0] * f[0] + a[1][1] * f[1] + a[1][2] * f[2];
    const double nz = a[2][0] * f[0] + a[2][1] * f[1] + a[2][2] * f[2];
    x[0] = nx; x[1] = ny; x[2] = nz;
  }
}

bool Structure::has_cold_fusion() const {
  // O(N log N) replacement of the reference's O(N^2) loop: sort along x, then only compare atoms whose x differs
  // by less than the threshold.  Same predicate: norm2(xj - xi) < 1e-9, no periodic images.
  constexpr double thr = 1.0e-9;
  std::vector<int> order(nat);
  for (int i = 0; i < nat; ++i) order[i] = i;
  std::sort(order.begin(), order.end(), [&](int p, int q) { return xyz[3 * size_t(p)] < xyz[3 * size_t(q)]; });
  for (int p = 0; p < nat; ++p) {
    const double* xi = &xyz[3 * size_t(order[p])];
    for (int q = p + 1; q < nat; ++q) {
      const double* xj = &xyz[3 * size_t(order[q])];
      if (xj[0] - xi[0] >= thr) break;
      const double d[3] = {xj[0] - xi[0], xj[1] - xi[1], xj[2] - xi[2]};
      if (len3(d) < thr) return true;
    }
  }
  return false;
}

// ---------------------------------------------------------------------------------------------------------------
bool HostModel::build(const Structure& mol, double ga_, double gc_, double wf_, std::string& err) {
  nid = static_cast<int>(mol.species_z.size());
  species_z = mol.species_z;
  ga = ga_; gc = gc_; wf = wf_;
  d4s = wf_ < 0.0;
  for (int z : species_z) {
    if (z > 103 && z < 112) {  // d4.f90:100-106
      static const char* sym[] = {"Rf", "Db", "Sg", "Bh", "Hs", "Mt", "Ds", "Rg"};
      err = std::string("Structure contains unsupported element '") + sym[z - 104] + "'";
      return false;
    }
    if (z < 1 || z > D4T_MAXZ) {
      err = "Structure contains invalid atomic number " + std::to_string(z);
      return false;
    }
    double row[4];
    std::string perr;
    const int prc = eeq_params_get(z, row, perr);
    if (prc == 2) { err = perr; return false; }
    if (prc != 0) {
      err = "EEQ-2019 parameters for element Z=" + std::to_string(z) +
            " are not available (multicharge's parameter file is not part of the reference tree): supply the row at run "
            "time with D4CU_EEQ_PARAMS=<file of 'Z chi eta kcn rad' rows> or d4cu_set_eeq_params";
      return false;
    }
  }
  const size_t n = nid;
  rcov.assign(n, 0); en.assign(n, 0); zeff.assign(n, 0); eta.assign(n, 0); r4r2.assign(n, 0);
  nref.assign(n, 0); refcn.assign(n * kMref, 0); refq.assign(n * kMref, 0); ngw.assign(n * kMref, 1);
  refalpha0.assign(n * kMref, 0); aiw.assign(n * kMref * kNfreq, 0.0);
  eeq_chi.assign(n, 0); eeq_eta.assign(n, 0); eeq_kcn.assign(n, 0); eeq_rad.assign(n, 0);
  for (size_t s = 0; s < n; ++s) {
    const int z = species_z[s];
    rcov[s] = (4.0 / 3.0) * (aatoau() * d4t_covrad2009_angstrom[z]);          // covrad.f90:38,64-65
    en[s] = d4t_pauling_en[z];
    zeff[s] = d4t_zeff[z];
    eta[s] = d4t_hardness[z];
    r4r2[s] = std::sqrt(0.5 * (d4t_r4_over_r2[z] * std::sqrt(static_cast<double>(z))));  // r4r2.f90:69-71
    const int nr = d4t_refn[z];
    nref[s] = nr;
    // reference.f90:145-161 (Gaussian centres = refcovcn) and :223-239 (EEQ reference charges = clsq)
    for (int r = 0; r < nr; ++r) {
      refcn[s * kMref + r] = d4t_refcovcn[z * kMref + r];
      refq[s * kMref + r] = d4t_clsq[z * kMref + r];
    }
    // reference.f90:179-205: multiplicity from how many references share the same rounded refcn
    int bucket[20] = {1};
    auto slot = [&](int r) { return std::min(static_cast<int>(std::lround(d4t_refcn[z * kMref + r])), 19); };
    for (int r = 0; r < nr; ++r) ++bucket[slot(r)];
    for (int r = 0; r < nr; ++r) {
      const int c = bucket[slot(r)];
      ngw[s * kMref + r] = c * (c + 1) / 2;
    }
    // reference.f90:331-365: charge-scaled reference polarisabilities
    for (int r = 0; r < nr; ++r) {
      const int sys = d4t_refsys[z * kMref + r];
      if (sys == 0) continue;
      const double zsec = d4t_zeff[sys];
      const double zt = zeta_fn(ga, d4t_hardness[sys] * gc, zsec, d4t_clsh[z * kMref + r] + zsec);
      for (int w = 0; w < kNfreq; ++w) {
        const double sec = (d4t_sscale[sys] * d4t_secaiw[sys * kNfreq + w]) * zt;
        const double val = d4t_ascale[z * kMref + r] *
                           (d4t_alphaiw[(z * kMref + r) * kNfreq + w] - d4t_hcount[z * kMref + r] * sec);
        aiw[(s * kMref + r) * kNfreq + w] = std::max(val, 0.0);
      }
      refalpha0[s * kMref + r] = aiw[(s * kMref + r) * kNfreq];
    }
    {
      double row[4];
      std::string perr;
      eeq_params_get(z, row, perr);   // (checked above)
      eeq_chi[s] = row[0]; eeq_eta[s] = row[1]; eeq_kcn[s] = row[2]; eeq_rad[s] = row[3];
    }
  }
  // d4s.f90:114-120: pair-specific weighting factors by atomic number
  wfpair.clear();
  if (d4s) {
    wfpair.assign(n * n, 0.0);
    for (size_t a = 0; a < n; ++a)
      for (size_t b = 0; b < n; ++b) wfpair[a * n + b] = d4t_wfpair[species_z[a] * (D4T_MAXZ + 1) + species_z[b]];
  }
  // d4.f90:215-229: reference C6 by Casimir-Polder integration
  static const TrapzWeights tw;
  const double thopi = 3.0 / kPi;
  refc6.assign(n * n * kMref * kMref, 0.0);
  for (size_t a = 0; a < n; ++a)
    for (size_t b = 0; b <= a; ++b)
      for (int ra = 0; ra < nref[a]; ++ra)
        for (int rb = 0; rb < nref[b]; ++rb) {
          double acc = 0.0;
          for (int w = 0; w < kNfreq; ++w)
            acc += (aiw[(a * kMref + ra) * kNfreq + w] * aiw[(b * kMref + rb) * kNfreq + w]) * tw.w[w];
          const double c6 = thopi * acc;
          refc6[((a * n + b) * kMref + ra) * kMref + rb] = c6;
          refc6[((b * n + a) * kMref + rb) * kMref + ra] = c6;
        }
  return true;
}

d4cu_model HostModel::view() const {
  d4cu_model m;
  m.nid = nid; m.ga = ga; m.gc = gc; m.wf = wf;
  m.rcov = rcov.data(); m.en = en.data(); m.zeff = zeff.data(); m.eta = eta.data(); m.r4r2 = r4r2.data();
  m.nref = nref.data(); m.refcn = refcn.data(); m.refq = refq.data(); m.ngw = ngw.data(); m.refc6 = refc6.data();
  m.refalpha0 = refalpha0.data();
  m.eeq_chi = eeq_chi.data(); m.eeq_eta = eeq_eta.data(); m.eeq_kcn = eeq_kcn.data(); m.eeq_rad = eeq_rad.data();
  m.wfpair = d4s ? wfpair.data() : nullptr;
  return m;
}

// ---------------------------------------------------------------------------------------------------------------
bool load_rational_damping(const std::string& method, bool atm, DampingParam& out) {
  // param.f90:263-265: drop everything after '/', lower-case
  std::string name = method.substr(0, method.find('/'));
  while (!name.empty() && name.back() == ' ') name.pop_back();
  for (char& ch : name)
    if (ch >= 'A' && ch <= 'Z') ch = static_cast<char>(ch - 'A' + 'a');
  int fid = 0;
  for (int i = 0; i < D4T_NALIAS; ++i)
    if (name == d4t_alias[i].name) { fid = d4t_alias[i].id; break; }
  if (fid == 0) return false;
  // param.f90:288-301: with ATM prefer the BJ-ATM fit, without prefer the BJ fit, fall back to the other table
  const double* prefer = atm ? d4t_param_bjatm[fid] : d4t_param_bj[fid];
  const double* other = atm ? d4t_param_bj[fid] : d4t_param_bjatm[fid];
  const double* row = prefer[0] != 0.0 ? prefer : (other[0] != 0.0 ? other : nullptr);
  if (!row) return false;
  out.s6 = row[1]; out.s8 = row[2]; out.a1 = row[3]; out.a2 = row[4]; out.alp = row[5];
  out.s9 = atm ? 1.0 : 0.0;  // api.f90:605-611
  return true;
}

// ---------------------------------------------------------------------------------------------------------------
void lattice_points_rep(const double* lat, const int rep[3], bool origin, Translations& out) {
  out.trans.clear();
  out.triples.clear();
  out.rep = {{rep[0], rep[1], rep[2]}};
  // cutoff.f90:137-151: non-negative multiples outermost, sign loops (+ then -) innermost
  for (int ix = 0; ix <= rep[0]; ++ix)
    for (int iy = 0; iy <= rep[1]; ++iy)
      for (int iz = 0; iz <= rep[2]; ++iz) {
        if (!origin && ix == 0 && iy == 0 && iz == 0) continue;
        for (int jx = 1; jx >= (ix > 0 ? -1 : 1); jx -= 2)
          for (int jy = 1; jy >= (iy > 0 ? -1 : 1); jy -= 2)
            for (int jz = 1; jz >= (iz > 0 ? -1 : 1); jz -= 2) {
              for (int c = 0; c < 3; ++c) {
                const double t1 = (lat[0 + c] * ix) * jx;
                const double t2 = (lat[3 + c] * iy) * jy;
                const double t3 = (lat[6 + c] * iz) * jz;
                out.trans.push_back((t1 + t2) + t3);
              }
              out.triples.push_back(ix * jx);
              out.triples.push_back(iy * jy);
              out.triples.push_back(iz * jz);
            }
      }
}

void lattice_points_cutoff(const double* lat, bool any_periodic, double rthr, Translations& out) {
  if (!any_periodic) {  // cutoff.f90:193-195
    out.trans.assign(3, 0.0);
    out.triples.assign(3, 0);
    out.rep = {{0, 0, 0}};
    return;
  }
  // cutoff.f90:207-241
  double nx[3], ny[3], nz[3];
  cross3(lat + 3, lat + 6, nx);
  cross3(lat + 6, lat + 0, ny);
  cross3(lat + 0, lat + 3, nz);
  const double lx = len3(nx), ly = len3(ny), lz = len3(nz);
  for (int k = 0; k < 3; ++k) { nx[k] /= lx; ny[k] /= ly; nz[k] /= lz; }
  const double cos10 = nx[0] * lat[0] + nx[1] * lat[1] + nx[2] * lat[2];
  const double cos21 = ny[0] * lat[3] + ny[1] * lat[4] + ny[2] * lat[5];
  const double cos32 = nz[0] * lat[6] + nz[1] * lat[7] + nz[2] * lat[8];
  const int rep[3] = {static_cast<int>(std::ceil(std::fabs(rthr / cos10))),
                      static_cast<int>(std::ceil(std::fabs(rthr / cos21))),
                      static_cast<int>(std::ceil(std::fabs(rthr / cos32)))};
  lattice_points_rep(lat, rep, true, out);
}

// ---------------------------------------------------------------------------------------------------------------
namespace {
double ewald_balance(double alpha, double rlen, double dlen, double vol) {
  auto rec = [&](double g) { return 4.0 * kPi * (std::exp(-0.25 * g * g / (alpha * alpha)) / (vol * g * g)); };
  auto dir = [&](double r) { return std::erfc(alpha * r) / r; };
  return (rec(4 * rlen) - rec(5 * rlen)) - (dir(2 * dlen) - dir(3 * dlen));
}
}  // namespace

void make_ewald_setup(const double* lat, EwaldSetup& out) {
  out.volume = std::fabs(det3(lat));
  double rec[9];
  {
    const double v = det3(lat);
    cross3(lat + 3, lat + 6, rec + 0);
    cross3(lat + 6, lat + 0, rec + 3);
    cross3(lat + 0, lat + 3, rec + 6);
    for (double& x : rec) x = 2.0 * kPi * x / v;
  }
  double rlen = DBL_MAX, dlen = DBL_MAX;
  for (int k = 0; k < 3; ++k) {
    rlen = std::min(rlen, rec[3 * k] * rec[3 * k] + rec[3 * k + 1] * rec[3 * k + 1] + rec[3 * k + 2] * rec[3 * k + 2]);
    dlen = std::min(dlen, lat[3 * k] * lat[3 * k] + lat[3 * k + 1] * lat[3 * k + 1] + lat[3 * k + 2] * lat[3 * k + 2]);
  }
  rlen = std::sqrt(rlen);
  dlen = std::sqrt(dlen);
  // multicharge ewald.f90 search_alpha: doubling bracket from 1e-8, then bisection to |diff| <= 1e-8 (<= 30 steps);
  // start value and tolerance are pinned by the NH3 crystal goldens (tests/golden), see DESIGN.md.
  constexpr double tol = 1.0e-8, alpha0 = 1.0e-8;
  double alpha = alpha0, lo = 0.0, hi = 0.0;
  int stat = 0;
  double diff = ewald_balance(alpha, rlen, dlen, out.volume);
  while (diff < -tol && alpha <= DBL_MAX) { alpha *= 2.0; diff = ewald_balance(alpha, rlen, dlen, out.volume); }
  if (alpha > DBL_MAX) stat = 1; else if (alpha == alpha0) stat = 2;
  if (stat == 0) {
    lo = 0.5 * alpha;
    while (diff < tol && alpha <= DBL_MAX) { alpha *= 2.0; diff = ewald_balance(alpha, rlen, dlen, out.volume); }
    if (alpha > DBL_MAX) stat = 3;
  }
  if (stat == 0) {
    hi = alpha;
    alpha = (lo + hi) * 0.5;
    int it = 0;
    diff = ewald_balance(alpha, rlen, dlen, out.volume);
    while (std::fabs(diff) > tol && it <= 30) {
      if (diff < 0) lo = alpha; else hi = alpha;
      alpha = (lo + hi) * 0.5;
      diff = ewald_balance(alpha, rlen, dlen, out.volume);
      ++it;
    }
    if (it > 30) stat = 4;
  }
  if (stat != 0) alpha = 0.25;
  out.alpha = alpha;
  const int rep2[3] = {2, 2, 2};
  lattice_points_rep(lat, rep2, true, out.dir);
  Translations g;
  lattice_points_rep(rec, rep2, false, g);
  out.rec = g.trans;
  out.rec_triples = g.triples;
  out.rec_fac.resize(g.count());
  const double fac = 4.0 * kPi / out.volume;
  for (int i = 0; i < g.count(); ++i) {
    const double* v = &g.trans[3 * size_t(i)];
    const double g2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    out.rec_fac[i] = g2 < 1.4901161193847656e-08 ? 0.0 : fac * (std::exp(-0.25 * g2 / (alpha * alpha)) / g2);
  }
  const int rep1[3] = {1, 1, 1};
  // wignerseitz.f90: get_lattice_points(periodic, lattice, thr = sqrt(epsilon), trans) -> rep = 1 -> 27 images
  lattice_points_rep(lat, rep1, true, out.wsc);
}

}  // namespace d4
