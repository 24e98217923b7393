// libdftd4 C-API (include/dftd4.h) served by the sm_100a engine.
//
// Host-side replacement of the reference's bind(C) layer src/dftd4/api.f90:98-941: same 22 symbols, handle ownership,
// NULL-handle behaviour and error strings.  Every compute entry point forwards to the d4cu_* C ABI (include/d4cu.h);
// there is no CPU implementation of the hot path behind it -- without a CUDA device the calls fail through the error
// handle.  D4 and D4S models (api.f90:351-431, 433-472) are both served by the engine.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/d4cu.h"
#include "../../include/dftd4.h"
#include "d4_host.hpp"

struct _dftd4_error {
  bool set = false;
  std::string message;
};
struct _dftd4_structure {
  d4::Structure mol;
};
struct _dftd4_model {
  d4::HostModel model;
  d4cu_engine* engine = nullptr;
  ~_dftd4_model() {
    if (engine) d4cu_engine_destroy(engine);
  }
};
struct _dftd4_param {
  d4::DampingParam p;
};

namespace {

// mctc_env fatal_error: the first error stays (api.f90 never clears a handle between calls either)
void fatal_error(dftd4_error err, const std::string& msg) {
  err->set = true;
  err->message = msg;
}

int pick_device() {
  // one process per GPU: D4CU_DEVICE, else torchrun's LOCAL_RANK, else device 0
  for (const char* name : {"D4CU_DEVICE", "LOCAL_RANK"}) {
    const char* v = std::getenv(name);
    if (v && *v) {
      const int d = std::atoi(v);
      const int n = d4cu_device_count();
      if (n > 0) return d % n;
    }
  }
  return 0;
}

bool ensure_engine(dftd4_error err, dftd4_model disp) {
  if (disp->engine) return true;
  if (d4cu_engine_create(&disp->engine, pick_device()) != 0) {
    fatal_error(err, std::string("CUDA engine unavailable: ") + d4cu_last_error());
    disp->engine = nullptr;
    return false;
  }
  return true;
}

// api.f90:669-690 style argument checks shared by the evaluation entry points
bool check_handles(dftd4_error err, dftd4_structure mol, dftd4_model disp, bool need_param, dftd4_param param) {
  if (!mol) { fatal_error(err, "Molecular structure data is missing"); return false; }
  if (!disp) { fatal_error(err, "Dispersion model is missing"); return false; }
  if (need_param && !param) { fatal_error(err, "Damping parameters are missing"); return false; }
  if (mol->mol.nat <= 0) { fatal_error(err, "Invalid molecular structure data provided"); return false; }
  // the model tables are per species of the structure it was built from (d4.f90:128-162)
  if (disp->model.species_z != mol->mol.species_z) {
    fatal_error(err, "Dispersion model does not match the molecular structure (different species)");
    return false;
  }
  return true;
}

int evaluate(dftd4_model disp, const d4::Structure& mol, const d4::DampingParam& p, double* energy, double* gradient,
             double* sigma) {
  const d4::HostModel& m = disp->model;
  const d4cu_model view = m.view();
  const double param[6] = {p.s6, p.s8, p.s9, p.a1, p.a2, p.alp};
  const double cutoff[5] = {m.disp2, m.disp3, m.cn, m.width2, m.width3};
  return d4cu_evaluate(disp->engine, mol.nat, mol.id.data(), mol.xyz.data(), mol.charge,
                       mol.any_periodic ? mol.lattice.data() : nullptr, mol.periodic.data(), &view, param, cutoff, energy,
                       gradient, sigma, nullptr);
}

}  // namespace

// ---- library ------------------------------------------------------------------------------------------------------
int dftd4_get_version(void) { return 40200; }  // version.f90:26-29 -> 4.2.0

// ---- error handle -------------------------------------------------------------------------------------------------
dftd4_error dftd4_new_error(void) { return new _dftd4_error(); }

int dftd4_check_error(dftd4_error error) {
  if (!error) return 2;
  return error->set ? 1 : 0;
}

void dftd4_get_error(dftd4_error error, char* buffer, const int* buffersize) {
  if (!error || !buffer || !error->set) return;
  // api.f90:894-904: copy min(len-1, len_trim(message)) characters and terminate
  const long long max_length = buffersize ? *buffersize : 2147483645LL;
  long long n = static_cast<long long>(error->message.size());
  if (n > max_length - 1) n = max_length - 1;
  if (n < 0) n = 0;
  std::memcpy(buffer, error->message.data(), static_cast<size_t>(n));
  buffer[n] = '\0';
}

void dftd4_delete_error(dftd4_error* error) {
  if (error && *error) {
    delete *error;
    *error = nullptr;
  }
}

// ---- structure ----------------------------------------------------------------------------------------------------
dftd4_structure dftd4_new_structure(dftd4_error error, const int natoms, const int* numbers, const double* positions,
                                    const double* charge, const double* lattice, const bool* periodic) {
  if (!error) return nullptr;
  if (natoms <= 0 || !numbers || !positions) {
    fatal_error(error, "Invalid molecular structure data provided");
    return nullptr;
  }
  dftd4_structure mol = new _dftd4_structure();
  mol->mol.assign(natoms, numbers, positions, charge, lattice, periodic);  // wraps into the central cell
  if (mol->mol.has_cold_fusion()) fatal_error(error, "Too close interatomic distances found");
  return mol;  // like api.f90:240-244 the handle is returned even when verification failed
}

void dftd4_delete_structure(dftd4_structure* mol) {
  if (mol && *mol) {
    delete *mol;
    *mol = nullptr;
  }
}

void dftd4_update_structure(dftd4_error error, dftd4_structure mol, const double* positions, const double* lattice) {
  if (!error) return;
  if (!mol) { fatal_error(error, "Molecular structure data is missing"); return; }
  if (mol->mol.nat <= 0 || mol->mol.species_z.empty() || !positions) {
    fatal_error(error, "Invalid molecular structure data provided");
    return;
  }
  mol->mol.update(positions, lattice);
  if (mol->mol.has_cold_fusion()) fatal_error(error, "Too close interatomic distances found");
}

// ---- dispersion model ---------------------------------------------------------------------------------------------
dftd4_model dftd4_custom_d4_model(dftd4_error error, dftd4_structure mol, double ga, double gc, double wf) {
  if (!error) return nullptr;
  if (!mol) { fatal_error(error, "Molecular structure data is missing"); return nullptr; }
  std::unique_ptr<_dftd4_model> disp(new _dftd4_model());
  std::string msg;
  if (!disp->model.build(mol->mol, ga, gc, wf, msg)) {
    fatal_error(error, msg);
    return nullptr;
  }
  return disp.release();
}

dftd4_model dftd4_new_d4_model(dftd4_error error, dftd4_structure mol) {
  return dftd4_custom_d4_model(error, mol, 3.0, 2.0, 6.0);  // d4.f90:59-65
}

// D4S (api.f90:351-428 -> model/d4s.f90:68-228): pair-specific Gaussian weighting factors instead of wf.  The host
// model selects it with wf < 0 (new_d4s_model takes no wf).
dftd4_model dftd4_custom_d4s_model(dftd4_error error, dftd4_structure mol, double ga, double gc) {
  return dftd4_custom_d4_model(error, mol, ga, gc, -1.0);
}

dftd4_model dftd4_new_d4s_model(dftd4_error error, dftd4_structure mol) {
  return dftd4_custom_d4s_model(error, mol, 3.0, 2.0);  // d4s.f90:59-63
}

void dftd4_delete_model(dftd4_model* disp) {
  if (disp && *disp) {
    delete *disp;
    *disp = nullptr;
  }
}

void dftd4_set_model_realspace_cutoff_smooth(dftd4_error error, dftd4_model disp, double disp2, double disp3, double cn,
                                             double width2, double width3) {
  if (!error) return;
  if (!disp) { fatal_error(error, "Dispersion model is missing"); return; }
  disp->model.disp2 = disp2; disp->model.disp3 = disp3; disp->model.cn = cn;
  disp->model.width2 = width2; disp->model.width3 = width3;
}

void dftd4_set_model_realspace_cutoff(dftd4_error error, dftd4_model disp, double disp2, double disp3, double cn) {
  // api.f90:495-519 builds realspace_cutoff(disp2, disp3, cn): widths fall back to their default 0
  dftd4_set_model_realspace_cutoff_smooth(error, disp, disp2, disp3, cn, 0.0, 0.0);
}

// ---- damping parameters -------------------------------------------------------------------------------------------
dftd4_param dftd4_new_rational_damping(dftd4_error error, double s6, double s8, double s9, double a1, double a2,
                                       double alp) {
  if (!error) return nullptr;
  dftd4_param p = new _dftd4_param();
  p->p.s6 = s6; p->p.s8 = s8; p->p.s9 = s9; p->p.a1 = a1; p->p.a2 = a2; p->p.alp = alp;
  return p;
}

dftd4_param dftd4_load_rational_damping(dftd4_error error, char* method, bool mbd) {
  if (!error) return nullptr;
  const std::string name = method ? method : "";
  d4::DampingParam p;
  if (!d4::load_rational_damping(name, mbd, p)) {
    fatal_error(error, "Functional '" + name + "' not known");
    return nullptr;
  }
  dftd4_param out = new _dftd4_param();
  out->p = p;
  return out;
}

void dftd4_delete_param(dftd4_param* param) {
  if (param && *param) {
    delete *param;
    *param = nullptr;
  }
}

// ---- evaluations --------------------------------------------------------------------------------------------------
void dftd4_get_dispersion(dftd4_error error, dftd4_structure mol, dftd4_model disp, dftd4_param param, double* energy,
                          double* gradient, double* sigma) {
  if (!error) return;
  if (!check_handles(error, mol, disp, true, param)) return;
  if (!ensure_engine(error, disp)) return;
  double e = 0.0;
  if (evaluate(disp, mol->mol, param->p, &e, gradient, sigma) != 0) {
    fatal_error(error, d4cu_last_error());
    return;
  }
  if (energy) *energy = e;
}

void dftd4_get_properties(dftd4_error error, dftd4_structure mol, dftd4_model disp, double* cn, double* charges,
                          double* c6, double* alpha) {
  if (!error) return;
  if (!check_handles(error, mol, disp, false, nullptr)) return;
  if (!ensure_engine(error, disp)) return;
  const d4::HostModel& m = disp->model;
  const d4cu_model view = m.view();
  const double param[6] = {1.0, 0.0, 0.0, 0.0, 0.0, 16.0};
  const double cutoff[5] = {m.disp2, m.disp3, m.cn, m.width2, m.width3};
  const d4::Structure& s = mol->mol;
  if (d4cu_upload(disp->engine, s.nat, s.id.data(), s.xyz.data(), s.charge, s.any_periodic ? s.lattice.data() : nullptr,
                  s.periodic.data(), &view, param, cutoff) != 0 ||
      d4cu_properties(disp->engine, cn, charges, c6, alpha) != 0)
    fatal_error(error, d4cu_last_error());
}

void dftd4_get_numerical_hessian(dftd4_error error, dftd4_structure mol, dftd4_model disp, dftd4_param param,
                                 double* hess) {
  if (!error) return;
  if (!check_handles(error, mol, disp, true, param)) return;
  if (!ensure_engine(error, disp)) return;
  // numdiff.f90:35-79: central differences of the analytic gradient, step 1e-4, displaced geometries are not re-wrapped
  const int nat = mol->mol.nat;
  const double step = 1.0e-4;
  // Device loop for small molecules: the 6 N displaced geometries are one batch, evaluated by ONE launch of the batched
  // kernel (one CTA per displaced geometry) instead of 6 N engine evaluations (the reference runs them in an OpenMP loop).
  if (!mol->mol.any_periodic && nat <= D4CU_BATCH_MAXAT && !disp->model.d4s && disp->model.width2 <= 0.0 &&
      disp->model.width3 <= 0.0) {
    const d4::HostModel& m = disp->model;
    const d4cu_model view = m.view();
    const d4::DampingParam& p = param->p;
    const double prm[6] = {p.s6, p.s8, p.s9, p.a1, p.a2, p.alp};
    const double cutoff[5] = {m.disp2, m.disp3, m.cn, m.width2, m.width3};
    const int nmol = 6 * nat;
    const size_t n3 = 3 * static_cast<size_t>(nat);
    std::vector<long long> offsets(nmol + 1);
    std::vector<int> ids(static_cast<size_t>(nmol) * nat), status(nmol, 0);
    std::vector<double> xyz(static_cast<size_t>(nmol) * n3), charges(nmol, mol->mol.charge), en(nmol), g(xyz.size());
    for (int k = 0; k <= nmol; ++k) offsets[k] = static_cast<long long>(k) * nat;
    for (int k = 0; k < nmol; ++k) {
      std::copy(mol->mol.id.begin(), mol->mol.id.end(), ids.begin() + static_cast<size_t>(k) * nat);
      std::copy(mol->mol.xyz.begin(), mol->mol.xyz.end(), xyz.begin() + k * n3);
      xyz[k * n3 + k / 2] += (k % 2 == 0 ? step : -step);   // geometries 2c / 2c+1: coordinate c displaced by +/- step
    }
    if (d4cu_batch_evaluate(disp->engine, nmol, offsets.data(), ids.data(), xyz.data(), charges.data(), &view, prm, cutoff,
                            en.data(), g.data(), nullptr, status.data()) != 0) {
      fatal_error(error, d4cu_last_error());
      return;
    }
    for (int k = 0; k < nmol; ++k)
      if (status[k] != 0) { fatal_error(error, "EEQ charge solve failed for a displaced geometry"); return; }
    for (size_t c = 0; c < n3; ++c) {
      double* row = hess + c * n3;   // Fortran hessian(:, :, ix, iat) == C hess[iat][ix][jat][jx]
      const double* gl = g.data() + (2 * c) * n3;
      const double* gr = g.data() + (2 * c + 1) * n3;
      for (size_t q = 0; q < n3; ++q) row[q] = (gl[q] - gr[q]) / (2 * step);
    }
    return;
  }
  d4::Structure displ = mol->mol;
  std::vector<double> gl(3 * static_cast<size_t>(nat)), gr(gl.size());
  double el, er;
  for (int iat = 0; iat < nat; ++iat)
    for (int ix = 0; ix < 3; ++ix) {
      const size_t c = 3 * static_cast<size_t>(iat) + ix;
      displ.xyz[c] = mol->mol.xyz[c] + step;
      if (evaluate(disp, displ, param->p, &el, gl.data(), nullptr) != 0) { fatal_error(error, d4cu_last_error()); return; }
      displ.xyz[c] = mol->mol.xyz[c] - step;
      if (evaluate(disp, displ, param->p, &er, gr.data(), nullptr) != 0) { fatal_error(error, d4cu_last_error()); return; }
      displ.xyz[c] = mol->mol.xyz[c];
      // Fortran hessian(:, :, ix, iat) == C hess[iat][ix][jat][jx]
      double* row = hess + c * 3 * static_cast<size_t>(nat);
      for (size_t k = 0; k < gl.size(); ++k) row[k] = (gl[k] - gr[k]) / (2 * step);
    }
}

void dftd4_get_pairwise_dispersion(dftd4_error error, dftd4_structure mol, dftd4_model disp, dftd4_param param,
                                   double* pair_energy2, double* pair_energy3) {
  if (!error) return;
  if (!check_handles(error, mol, disp, true, param)) return;
  if (!ensure_engine(error, disp)) return;
  const d4::HostModel& m = disp->model;
  const d4cu_model view = m.view();
  const d4::DampingParam& p = param->p;
  const double prm[6] = {p.s6, p.s8, p.s9, p.a1, p.a2, p.alp};
  const double cutoff[5] = {m.disp2, m.disp3, m.cn, m.width2, m.width3};
  const d4::Structure& s = mol->mol;
  if (d4cu_upload(disp->engine, s.nat, s.id.data(), s.xyz.data(), s.charge, s.any_periodic ? s.lattice.data() : nullptr,
                  s.periodic.data(), &view, prm, cutoff) != 0 ||
      d4cu_pairwise(disp->engine, pair_energy2, pair_energy3) != 0)
    fatal_error(error, d4cu_last_error());
}

// ---- bridge to the d4cu ABI (include/d4cu.h, "bridge" section) -----------------------------------------------------
extern "C" d4cu_engine* d4cu_engine_of_model(void* handle) {
  dftd4_model disp = static_cast<dftd4_model>(handle);
  if (!disp) return nullptr;
  if (!disp->engine && d4cu_engine_create(&disp->engine, pick_device()) != 0) disp->engine = nullptr;
  return disp->engine;
}

extern "C" int d4cu_upload_handles(void* hmol, void* hdisp, void* hparam) {
  dftd4_structure mol = static_cast<dftd4_structure>(hmol);
  dftd4_model disp = static_cast<dftd4_model>(hdisp);
  dftd4_param param = static_cast<dftd4_param>(hparam);
  if (!mol || !disp || !param) return 1;
  d4cu_engine* E = d4cu_engine_of_model(disp);
  if (!E) return 1;
  const d4::HostModel& m = disp->model;
  const d4cu_model view = m.view();
  const d4::DampingParam& p = param->p;
  const double prm[6] = {p.s6, p.s8, p.s9, p.a1, p.a2, p.alp};
  const double cutoff[5] = {m.disp2, m.disp3, m.cn, m.width2, m.width3};
  const d4::Structure& s = mol->mol;
  return d4cu_upload(E, s.nat, s.id.data(), s.xyz.data(), s.charge, s.any_periodic ? s.lattice.data() : nullptr,
                     s.periodic.data(), &view, prm, cutoff);
}

// Coordinates held by a structure handle, i.e. after wrap_to_central_cell (api.f90:242,303): what the engine sees.
extern "C" int d4cu_structure_positions(void* hmol, double* xyz_out) {
  dftd4_structure mol = static_cast<dftd4_structure>(hmol);
  if (!mol || !xyz_out) return 1;
  std::memcpy(xyz_out, mol->mol.xyz.data(), sizeof(double) * 3 * static_cast<size_t>(mol->mol.nat));
  return 0;
}

extern "C" int d4cu_batch_numbers(int nmol, const long long* offsets, const int* numbers, const double* xyz, const double* charges,
                                  double ga, double gc, double wf, const double* param, const double* cutoff, double* energies,
                                  double* gradient, double* sigma, int* status, float* kernel_ms) {
  static std::mutex mtx;
  static d4cu_engine* engine = nullptr;
  std::lock_guard<std::mutex> lk(mtx);
  if (nmol <= 0 || !offsets || !numbers || !xyz || !energies) return 1;
  if (!engine && d4cu_engine_create(&engine, pick_device()) != 0) { engine = nullptr; return 1; }
  const long long total = offsets[nmol];
  // union of the elements, species ids in order of first appearance (like mctc-lib's structure_type)
  d4::Structure uni;
  std::vector<int> id(total);
  for (long long a = 0; a < total; ++a) {
    auto it = std::find(uni.species_z.begin(), uni.species_z.end(), numbers[a]);
    if (it == uni.species_z.end()) { uni.species_z.push_back(numbers[a]); id[a] = (int)uni.species_z.size() - 1; }
    else id[a] = (int)(it - uni.species_z.begin());
  }
  d4::HostModel m;
  std::string err;
  if (!m.build(uni, ga, gc, wf, err)) { d4cu_set_last_error(err.c_str()); return 1; }
  const d4cu_model view = m.view();
  const int rc = d4cu_batch_evaluate(engine, nmol, offsets, id.data(), xyz, charges, &view, param, cutoff, energies, gradient, sigma, status);
  if (rc == 0 && kernel_ms) { float ms[10]; int nl; d4cu_last_timing(engine, ms, &nl); *kernel_ms = ms[9]; }
  return rc;
}

extern "C" int d4cu_param_values(void* hparam, double* out6) {
  dftd4_param param = static_cast<dftd4_param>(hparam);
  if (!param || !out6) return 1;
  const d4::DampingParam& p = param->p;
  out6[0] = p.s6; out6[1] = p.s8; out6[2] = p.s9; out6[3] = p.a1; out6[4] = p.a2; out6[5] = p.alp;
  return 0;
}

// ---- EEQ-2019 parameter rows as a run-time input (include/d4cu.h) ---------------------------------------------------
extern "C" int d4cu_set_eeq_params(int z, double chi, double eta, double kcn, double rad) {
  const double row[4] = {chi, eta, kcn, rad};
  if (d4::eeq_params_set(z, row)) {
    d4cu_set_last_error("d4cu_set_eeq_params: need 1 <= Z <= 103 and rad > 0");
    return 1;
  }
  return 0;
}
extern "C" int d4cu_load_eeq_params(const char* path) {
  std::string err;
  if (!path || d4::eeq_params_load(path, err)) {
    d4cu_set_last_error(path ? err.c_str() : "d4cu_load_eeq_params: path is NULL");
    return 1;
  }
  return 0;
}
extern "C" void d4cu_clear_eeq_params(void) { d4::eeq_params_clear(); }
extern "C" int d4cu_get_eeq_params(int z, double* row4) {
  std::string err;
  const int rc = row4 ? d4::eeq_params_get(z, row4, err) : 1;
  if (rc == 2) d4cu_set_last_error(err.c_str());
  return rc;
}
// The row a model handle carries for its species of atomic number z (what the engine uploads): 0 ok, 1 not in the model.
extern "C" int d4cu_model_eeq_params(void* hmodel, int z, double* row4) {
  dftd4_model disp = static_cast<dftd4_model>(hmodel);
  if (!disp || !row4) return 1;
  const d4::HostModel& m = disp->model;
  for (int s = 0; s < m.nid; ++s)
    if (m.species_z[s] == z) {
      row4[0] = m.eeq_chi[s]; row4[1] = m.eeq_eta[s]; row4[2] = m.eeq_kcn[s]; row4[3] = m.eeq_rad[s];
      return 0;
    }
  return 1;
}
