/*
 * d4cu.h -- thin C ABI of the B200 (sm_100a) DFT-D4 engine.
 *
 * This is the drop-in boundary for the reference's hot path `get_dispersion`
 * (reference: src/dftd4/disp.f90:40-134): a Fortran host keeps its own structure_type /
 * d4_model / rational_damping_param objects and hands their plain arrays to d4cu_evaluate()
 * through iso_c_binding (INTEGRATION.md shows the ~60-line shim).  No torch types, no C++ types.
 *
 * All lengths in Bohr, energies in Hartree.  Arrays are C-contiguous:
 *   xyz[nat][3], lattice[3][3] with lattice[k][:] = k-th lattice vector (= Fortran lattice(:,k)),
 *   gradient[nat][3], sigma[3][3].
 * Every function returns 0 on success; on failure a message is available from d4cu_last_error().
 * There is no CPU fallback: without a CUDA device every compute entry point fails loudly.
 */
#ifndef D4CU_H
#define D4CU_H

#ifdef __cplusplus
extern "C" {
#endif

#define D4CU_MREF 7 /* max reference systems per element (reference.f90:74-93) */

/* Per-species tables of a D4 model: the members of the reference's d4_model
 * (src/dftd4/model/d4.f90:39-55, model/type.f90:29-84) plus the EEQ-2019 parameters
 * held by its mchrg member (multicharge new_eeq2019_model, call site d4.f90:190). */
typedef struct d4cu_model {
  int nid;              /* number of species */
  double ga, gc, wf;    /* charge scaling height / steepness, CN weighting factor */
  const double *rcov;   /* [nid] covalent radii (covrad.f90:64-65) */
  const double *en;     /* [nid] Pauling EN (en.f90) */
  const double *zeff;   /* [nid] effective nuclear charge (zeff.f90) */
  const double *eta;    /* [nid] chemical hardness (hardness.f90) */
  const double *r4r2;   /* [nid] sqrt(0.5*r4/r2*sqrt(Z)) (r4r2.f90:69-71) */
  const int *nref;      /* [nid] number of reference systems */
  const double *refcn;  /* [nid][7] reference CN (d4%cn) */
  const double *refq;   /* [nid][7] reference charge (d4%q) */
  const int *ngw;       /* [nid][7] Gaussian multiplicity (d4%ngw) */
  const double *refc6;  /* [nid][nid][7][7] reference C6 (d4%c6(iref,jref,isp,jsp) at [isp][jsp][iref][jref]) */
  const double *refalpha0; /* [nid][7] static reference polarisability aiw(1,iref,isp); only for get_properties */
  const double *eeq_chi, *eeq_eta, *eeq_kcn, *eeq_rad; /* [nid] EEQ-2019 parameters */
  const double *wfpair; /* D4S model (src/dftd4/model/d4s.f90:39-55): [nid][nid] pair-specific Gaussian weighting
                         * factors d4%wf(isp, jsp) = get_wfpair_val (d4s.f90:114-120); NULL selects the D4 model (wf) */
} d4cu_model;

/* Integer interaction counts: the "neighbour counts / pair lists" of the hot path. */
typedef struct d4cu_counts {
  long long pcn;   /* directed (i, j-image) entries with 1e-12 <= r^2 <= cn^2            */
  long long peeq;  /* same with the EEQ coordination-number cutoff (25 bohr)              */
  long long p2;    /* directed entries with eps <= r^2 <= disp2^2 (and sw > 0)            */
  long long p3;    /* directed entries with eps <= r^2 <= disp3^2                         */
  long long t3;    /* ordered distinct triples (i, j-image, k-image), all three in range */
} d4cu_counts;

typedef struct d4cu_engine d4cu_engine;

/* Create / destroy an engine bound to one CUDA device (one per model handle; owns a stream and workspaces). */
int d4cu_engine_create(d4cu_engine **engine, int device);
void d4cu_engine_destroy(d4cu_engine *engine);
const char *d4cu_last_error(void);
void d4cu_set_last_error(const char *msg); /* used by the host layers above the engine */
int d4cu_device_count(void);

/*
 * One evaluation of the hot path (disp.f90:40-134) with HOST buffers.
 *   id[nat]       0-based species index of every atom
 *   periodic[3]   0/1; lattice may be NULL when nothing is periodic
 *   param[6]      s6, s8, s9, a1, a2, alp           (rational.f90:32-53)
 *   cutoff[5]     disp2, disp3, cn, width2, width3  (cutoff.f90:43-61)
 *   gradient, sigma, counts may be NULL (energy only when both gradient and sigma are NULL).
 */
int d4cu_evaluate(d4cu_engine *engine, int nat, const int *id, const double *xyz, double charge, const double *lattice,
                  const int *periodic, const d4cu_model *model, const double *param, const double *cutoff,
                  double *energy, double *gradient, double *sigma, d4cu_counts *counts);

/* The same evaluation split so that a caller (bench.py `value`) can time the device work with inputs resident in
 * HBM: upload = H2D of the structure + host-side translation lists; compute = all kernels, results stay on the
 * device; download = D2H of energy / gradient / sigma / counts. */
int d4cu_upload(d4cu_engine *engine, int nat, const int *id, const double *xyz, double charge, const double *lattice,
                const int *periodic, const d4cu_model *model, const double *param, const double *cutoff);
int d4cu_compute(d4cu_engine *engine, int want_gradient);
int d4cu_download(d4cu_engine *engine, double *energy, double *gradient, double *sigma, d4cu_counts *counts);

/* Properties (disp.f90:138-188): cn[nat], q[nat], c6[nat*nat] (nullable), alpha[nat] after an upload. */
int d4cu_properties(d4cu_engine *engine, double *cn, double *q, double *c6, double *alpha);
/* Pair-resolved energies (disp.f90:192-255): e2[nat*nat], e3[nat*nat] after an upload. */
int d4cu_pairwise(d4cu_engine *engine, double *e2, double *e3);

/* Pair list of the last compute for the cutoff class `which` (0 cn, 1 disp2, 2 disp3): directed entries
 * (i, j, translation index into the reference-ordered translation list of that cutoff).  Call with out==NULL to
 * get the entry count. */
long long d4cu_pair_list(d4cu_engine *engine, int which, int *i_idx, int *j_idx, int *t_idx);

/* Timing of the last compute in milliseconds (CUDA events on the engine stream):
 * [0] lists [1] cn [2] eeq build [3] eeq factor+solve [4] weights [5] pair [6] eeq adjoint [7] atm [8] cn chain
 * [9] total.  `launches` receives the number of kernels launched by the last compute. */
int d4cu_last_timing(d4cu_engine *engine, float *ms10, int *launches);

/* Real-space Ewald images (alpha r <= 6) summed by the periodic EEQ matrix build of the last compute, over all pairs
 * i > j and ranks: the unit of work of the eeq_build / eeq_adjoint roofline lines of bench.py (0 for molecules). */
int d4cu_last_ewald_images(d4cu_engine *engine, long long *images);

/* Batched small molecules, one launch per batch (extension; the reference API has no batched entry -- BASELINE config 3
 * would be nmol calls of dftd4_get_dispersion).  Molecule m owns atoms offsets[m] .. offsets[m+1]-1 of id / xyz /
 * gradient; at most 64 atoms per molecule, molecular (no lattice), sharp cutoffs.  `model` holds the per-species tables
 * of the union of all elements in the batch, id[] indexes into it.  charges (total charge per molecule), gradient
 * [total][3], sigma [nmol][9] and status [nmol] (0 ok, 1 EEQ failure) may be NULL.  Each molecule is evaluated by one
 * CTA entirely in shared memory (d4_batch.cuh). */
#define D4CU_BATCH_MAXAT 64
int d4cu_batch_evaluate(d4cu_engine *engine, int nmol, const long long *offsets, const int *id, const double *xyz,
                        const double *charges, const d4cu_model *model, const double *param, const double *cutoff,
                        double *energies, double *gradient, double *sigma, int *status);
/* The same from atomic numbers: builds the union-species D4 model (ga, gc, wf) and a process-wide engine on the device
 * picked like the dftd4 C-API does (D4CU_DEVICE / LOCAL_RANK / 0).  kernel_ms (nullable) receives the kernel time. */
int d4cu_batch_numbers(int nmol, const long long *offsets, const int *numbers, const double *xyz, const double *charges,
                       double ga, double gc, double wf, const double *param, const double *cutoff, double *energies,
                       double *gradient, double *sigma, int *status, float *kernel_ms);

/*
 * Batched small molecules for callers that must not allocate device memory.  (No counterpart in the reference: its API
 * has neither a batched entry nor a device.)
 *
 * Who needs this: a host program that handed all of the GPU's memory to its own pool when it started (an MD engine, a
 * quantum-chemistry package or an ML framework with an arena allocator), where a cudaMalloc made behind its back fails or
 * fragments the pool; and a plug-in running under a rule that forbids cudaMalloc / cudaFree in the step loop because
 * both stall every stream of the device while they run.  d4cu_batch_evaluate / d4cu_batch_numbers allocate staging
 * buffers and device copies of the model tables on every call; this entry point allocates nothing, on the device or
 * through any library: no cudaMalloc, no cudaFree, no engine object, no stream or event creation.
 *
 * The caller owns the bulk data and passes DEVICE pointers (prefix d_), valid on `device`:
 *   d_species[total]   index of every atom into species_numbers (0 .. nspecies-1)
 *   d_xyz[total][3]    Bohr;   d_charges[nmol] total charge per molecule, or NULL for neutral molecules
 *   d_energies[nmol];  d_gradient[total][3] and d_sigma[nmol][9] both NULL for energies only;  d_status[nmol] or NULL
 * and HOST pointers for the small descriptors: offsets[nmol+1] (molecule m owns atoms offsets[m] .. offsets[m+1]-1),
 * species_numbers[nspecies] (atomic number of every species), param[6], cutoff[5] as for d4cu_batch_evaluate.
 * stream is a cudaStream_t passed as void* (NULL: the default stream).
 *
 * What the other entry points keep in freshly allocated memory lives here in ONE fixed array compiled into the library
 * (D4CU_NOALLOC_SCRATCH_BYTES of `static __device__` storage; every GPU on which the library is loaded holds it until
 * the process exits): the device copy of offsets, the per-species model tables the kernel reads, a status array for
 * when d_status is NULL, and two counters (species indices out of range, molecules whose EEQ matrix was not positive
 * definite).  Hence the limits, rejected with a message in d4cu_last_error() before anything is launched:
 *   nmol <= D4CU_NOALLOC_MAX_MOLECULES,  nspecies <= D4CU_NOALLOC_MAX_SPECIES,  1 .. D4CU_BATCH_MAXAT atoms per molecule,
 * molecular, sharp cutoffs, D4 model (wf > 0), like d4cu_batch_evaluate.
 *
 * The array is shared by every call on a device and keeps what the previous call left in it.  Each call overwrites the
 * tables and offsets it uses and zeroes the two counters before use.  Two calls whose GPU work overlapped would race
 * on the array; the entry point therefore takes a process-wide lock and returns only after `stream` has finished its
 * work, so calls made through it never overlap.  Do not give the array's contents any meaning between calls.
 *
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device current)
 * and raises the kernel's shared-memory limit; make that first call before you start any stream capture on the device.
 * The calling thread's current device is restored on return.  Returns 0, or 1 with d4cu_last_error(): invalid
 * arguments, a species index out of range (nothing evaluated), or -- only when d_status is NULL -- a failed molecule.
 */
#define D4CU_NOALLOC_MAX_MOLECULES 65536
#define D4CU_NOALLOC_MAX_SPECIES 32
#define D4CU_NOALLOC_SCRATCH_BYTES (1280 * 1024)
int d4cu_batch_noalloc(int device, void *stream, int nmol, const long long *offsets, int nspecies,
                       const int *species_numbers, const int *d_species, const double *d_xyz, const double *d_charges,
                       double ga, double gc, double wf, const double *param, const double *cutoff, double *d_energies,
                       double *d_gradient, double *d_sigma, int *d_status);

/* Multi-GPU (one process per GPU): shard the i-rows of the pair / ATM / CN stages over `world` ranks and all-reduce
 * the accumulators with NCCL over NVLink.  uid = 128-byte ncclUniqueId obtained on rank 0 by d4cu_nccl_unique_id and
 * broadcast by the host program. */
int d4cu_nccl_unique_id(void *uid128);
int d4cu_engine_set_comm(d4cu_engine *engine, int rank, int world, const void *uid128);
/* Row partition used by rank `rank` of `world` for a structure of nat atoms: half-open [begin,end) blocks dealt
 * round-robin (host-only helper, also exercised by the CPU tests).  The engine deals blocks of D4CU_ROW_BLOCK atoms:
 * the ATM cost of a row grows with the atom index (edge ownership), small blocks keep the ranks balanced. */
#define D4CU_ROW_BLOCK 8
int d4cu_partition(int nat, int rank, int world, int block, int *n_blocks, int *first_block, int *stride);
/* Multi-GPU EEQ factorisation (large matrices): block column k of the Cholesky factor belongs to rank k % world.
 * d4cu_chol_first_owned returns the first tile column tj >= 0 of the trailing matrix of step kb (absolute block column
 * kb + 1 + tj) owned by `rank`; its further columns follow with stride `world`.  Host-only helpers (CPU tests). */
int d4cu_chol_column_owner(int block_column, int world);
int d4cu_chol_first_owned(int kb, int rank, int world);

/* Measured FP64 peaks for the roofline denominators: a register-resident DFMA chain on all SMs (vector pipe) and a
 * DMMA (mma.sync f64) chain; returns TFLOP/s. */
int d4cu_measure_fp64_peak(int device, double *dfma_tflops, double *dmma_tflops);

/* ---- bridge to the dftd4.h handles (extension; the 22 reference symbols are untouched) ----------------------------
 * A host program that holds dftd4_structure / dftd4_model / dftd4_param handles can reach the engine behind a model
 * handle to (a) keep the structure resident and time the device work alone (upload / compute / download),
 * (b) attach an NCCL communicator for multi-GPU sharding, (c) read counts, pair lists and stage timings.
 * Handles are passed as void* so that this header does not depend on dftd4.h. */
d4cu_engine *d4cu_engine_of_model(void *dftd4_model_handle); /* created on first use; NULL + d4cu_last_error on failure */
int d4cu_upload_handles(void *dftd4_structure_handle, void *dftd4_model_handle, void *dftd4_param_handle);
/* s6, s8, s9, a1, a2, alp held by a dftd4_param handle (needed by the batched entry, which takes plain values). */
int d4cu_param_values(void *dftd4_param_handle, double *out6);
/* Coordinates [nat][3] held by a structure handle (after the API's wrap into the central cell, api.f90:242,303). */
int d4cu_structure_positions(void *dftd4_structure_handle, double *xyz_out);

/* ---- EEQ-2019 parameter rows as a run-time input -------------------------------------------------------------------
 * The charge model's per-element table (chi, eta, kcn, rad) lives in multicharge (param/eeq2019.f90), which is not part
 * of the reference tree; the compiled table restates Z = 1..86.  Rows supplied at run time override compiled rows and add
 * the elements the table lacks (Z = 87..103; the reference evaluates them, test/unit/test_dftd4.f90:901-1044):
 *   d4cu_set_eeq_params   one row;   d4cu_load_eeq_params   a text file, one "Z chi eta kcn rad" row per line ('#' starts a
 *   comment);   the environment variable D4CU_EEQ_PARAMS names such a file, read when the first model is built.
 * Rows take effect for models created afterwards.  d4cu_get_eeq_params: row in effect (returns 1 when there is none,
 * 2 when the D4CU_EEQ_PARAMS file is unreadable); d4cu_model_eeq_params: row a dftd4_model handle carries for element z. */
int d4cu_set_eeq_params(int z, double chi, double eta, double kcn, double rad);
int d4cu_load_eeq_params(const char *path);
void d4cu_clear_eeq_params(void);
int d4cu_get_eeq_params(int z, double *row4);
int d4cu_model_eeq_params(void *dftd4_model_handle, int z, double *row4);

#ifdef __cplusplus
}
#endif
#endif
