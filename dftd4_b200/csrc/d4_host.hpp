// Host-side objects behind the dftd4 C-API handles and the helpers the CUDA engine needs on the host
// (model tables, damping-parameter database, lattice translations).  C++17, no CUDA in this header.
#pragma once
#include <array>
#include <string>
#include <vector>

#include "../../include/d4cu.h"

namespace d4 {

constexpr int kMref = D4CU_MREF;
constexpr int kNfreq = 23;

// mol: what the reference keeps in mctc-lib's structure_type (numbers -> species ids in order of first appearance).
struct Structure {
  int nat = 0;
  std::vector<int> numbers;    // atomic number per atom
  std::vector<int> id;         // species index per atom
  std::vector<int> species_z;  // atomic number per species
  std::vector<double> xyz;     // [nat][3], wrapped into the central cell when periodic
  std::array<double, 9> lattice{};  // lattice[3*k+c]: component c of vector k
  std::array<int, 3> periodic{{0, 0, 0}};
  bool any_periodic = false;
  double charge = 0.0;
  unsigned long long revision = 0;  // bumped by every update, lets the engine skip redundant uploads

  void assign(int natoms, const int* nums, const double* pos, const double* chrg, const double* lat, const bool* per);
  void update(const double* pos, const double* lat);
  // src/dftd4/utils.f90:29-65
  void wrap_to_central_cell();
  // src/dftd4/api.f90:925-938 : true when two atoms are closer than 1e-9 bohr
  bool has_cold_fusion() const;
};

// The reference's d4_model (src/dftd4/model/d4.f90:39-55) flattened per species, plus its EEQ-2019 parameters.
struct HostModel {
  int nid = 0;
  double ga = 3.0, gc = 2.0, wf = 6.0;
  bool d4s = false;             // D4S model (src/dftd4/model/d4s.f90): pair-specific weighting factors replace wf
  std::vector<double> wfpair;   // [nid][nid], D4S only
  std::vector<double> rcov, en, zeff, eta, r4r2, refcn, refq, refc6, refalpha0, aiw;
  std::vector<int> nref, ngw, species_z;
  std::vector<double> eeq_chi, eeq_eta, eeq_kcn, eeq_rad;
  // realspace cutoffs travel with the model handle in the C-API (api.f90:76-82)
  double disp2 = 60.0, disp3 = 40.0, cn = 30.0, width2 = 0.0, width3 = 0.0;

  // src/dftd4/model/d4.f90:71-231 (D4) and model/d4s.f90:68-228 (D4S, requested with wf_ < 0: new_d4s_model has
  // no wf argument); returns false and fills err on unsupported input
  bool build(const Structure& mol, double ga_, double gc_, double wf_, std::string& err);
  d4cu_model view() const;
};

// EEQ-2019 parameter rows (chi, eta, kcn, rad) as a run-time input: see d4_host.cpp / include/d4cu.h.
int eeq_params_set(int z, const double* row4);
int eeq_params_load(const char* path, std::string& err);
void eeq_params_clear();
int eeq_params_get(int z, double* row4, std::string& err);   // 0 ok, 1 no row for z, 2 D4CU_EEQ_PARAMS unreadable

struct DampingParam {
  double s6 = 1.0, s8 = 0.0, s9 = 1.0, a1 = 0.0, a2 = 0.0, alp = 16.0;
};

// src/dftd4/param.f90:248-303 + api.f90:582-621.  false when the functional is unknown.
bool load_rational_damping(const std::string& method, bool atm, DampingParam& out);

// Reference-ordered lattice translations (src/dftd4/cutoff.f90:118-241), bit-identical arithmetic.
// trans[3*t+c]; triples[3*t+c] = signed integer multiples of the lattice vectors.
struct Translations {
  std::vector<double> trans;
  std::vector<int> triples;
  std::array<int, 3> rep{{0, 0, 0}};
  int count() const { return static_cast<int>(trans.size() / 3); }
};
void lattice_points_cutoff(const double* lattice, bool any_periodic, double rthr, Translations& out);
void lattice_points_rep(const double* lattice, const int rep[3], bool origin, Translations& out);

// EEQ under periodic boundary conditions (multicharge v0.5.0, not in the reference tree):
// Ewald splitting parameter, real/reciprocal translation sets and the 27-image Wigner-Seitz search set.
struct EwaldSetup {
  double alpha = 0.0, volume = 0.0;
  Translations dir;            // 5x5x5 real-space images incl. origin
  std::vector<double> rec;     // 124 reciprocal vectors [n][3]
  std::vector<int> rec_triples; // their integer multiples of the reciprocal lattice vectors [n][3]
  std::vector<double> rec_fac; // 4 pi / V * exp(-G^2/(4 alpha^2)) / G^2 per reciprocal vector
  Translations wsc;            // 27 images
};
void make_ewald_setup(const double* lattice, EwaldSetup& out);

}  // namespace d4
