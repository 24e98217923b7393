"""d4cu_batch_noalloc: the batched small-molecule kernel for callers that may not allocate device memory.

Some host programs cannot let this library call cudaMalloc: an MD engine or ML framework that gave the whole GPU to its
own memory pool at start-up (a second allocator then fails or fragments the pool), or a plug-in whose step loop must
not contain cudaMalloc / cudaFree because both stall every stream on the device.  For them include/d4cu.h offers one
entry point that takes the caller's device buffers and keeps all of its own working data (offsets, model tables, a
status array, two counters) in a fixed array compiled into the library.

The test plays such a caller: torch owns every device buffer, the library is reached through the C ABI with raw
pointers, and the results must equal those of the allocating entry point (d4cu_batch_numbers, host buffers) within
the tolerances of the other parity tests -- energy 1e-10 relative, gradient / virial 1e-9 absolute.  Two calls run back
to back on different batches, so the second one finds the first one's tables, offsets and counters still in the shared
array and has to cope with them.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the parity
tests.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

E_RTOL = 1e-10   # relative, energy
G_ATOL = 1e-9    # absolute, gradient and virial (Hartree/Bohr, Hartree)

MAX_MOLECULES = 65536   # D4CU_NOALLOC_MAX_MOLECULES
MAX_SPECIES = 32        # D4CU_NOALLOC_MAX_SPECIES
CUTOFF = (60.0, 40.0, 30.0, 0.0, 0.0)
MODEL = (3.0, 2.0, 6.0)

_vp = C.c_void_p
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_lp = C.POINTER(C.c_longlong)


@pytest.fixture(scope="module")
def entry(d4lib):
    fn = d4lib.lib.d4cu_batch_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_int, _vp, C.c_int, _lp, C.c_int, _ip, _vp, _vp, _vp, C.c_double, C.c_double, C.c_double, _dp, _dp,
                   _vp, _vp, _vp, _vp]
    d4lib.lib.d4cu_last_error.restype = C.c_char_p
    return fn


def _species(numbers):
    """Species list in order of first appearance and the per-atom index into it."""
    uniq, first = np.unique(numbers, return_index=True)
    order = uniq[np.argsort(first)].astype(np.int32)
    lookup = {int(z): s for s, z in enumerate(order)}
    return order, np.array([lookup[int(z)] for z in numbers], dtype=np.int32)


def _call_noalloc(entry, torch, w, param, charges, grad=True, with_status=True):
    """Play the caller that owns all device memory: torch tensors in, torch tensors out, raw pointers in between."""
    dev = torch.device("cuda:0")
    offsets = np.ascontiguousarray(w["offsets"], dtype=np.int64)
    nmol, total = len(offsets) - 1, int(offsets[-1])
    species_numbers, species = _species(w["numbers"])
    d_species = torch.from_numpy(species).to(dev)
    d_xyz = torch.from_numpy(np.ascontiguousarray(w["positions"], dtype=np.float64)).to(dev)
    d_q = None if charges is None else torch.from_numpy(np.ascontiguousarray(charges, dtype=np.float64)).to(dev)
    # outputs start out poisoned so that anything the kernel leaves unwritten shows up
    d_e = torch.full((nmol,), float("nan"), dtype=torch.float64, device=dev)
    d_g = torch.full((total, 3), float("nan"), dtype=torch.float64, device=dev) if grad else None
    d_s = torch.full((nmol, 9), float("nan"), dtype=torch.float64, device=dev) if grad else None
    d_st = torch.full((nmol,), -7, dtype=torch.int32, device=dev) if with_status else None
    p = np.array(param, dtype=np.float64)
    cut = np.array(CUTOFF, dtype=np.float64)
    torch.cuda.synchronize()
    ptr = lambda t: None if t is None else t.data_ptr()
    rc = entry(0, torch.cuda.current_stream().cuda_stream or None, nmol, offsets.ctypes.data_as(_lp),
               len(species_numbers), species_numbers.ctypes.data_as(_ip), ptr(d_species), ptr(d_xyz), ptr(d_q),
               MODEL[0], MODEL[1], MODEL[2], p.ctypes.data_as(_dp), cut.ctypes.data_as(_dp), ptr(d_e), ptr(d_g), ptr(d_s),
               ptr(d_st))
    host = lambda t: None if t is None else t.cpu().numpy()
    return rc, dict(energy=host(d_e), gradient=host(d_g), virial=None if d_s is None else host(d_s).reshape(nmol, 3, 3),
                    status=host(d_st))


def _assert_same(out, ref):
    assert np.all(np.isfinite(out["energy"]))
    assert np.max(np.abs(out["energy"] - ref["energy"]) / np.abs(ref["energy"])) <= E_RTOL
    if ref["gradient"] is not None and out["gradient"] is not None:
        assert np.max(np.abs(out["gradient"] - ref["gradient"])) <= G_ATOL
        assert np.max(np.abs(out["virial"] - ref["virial"])) <= G_ATOL
    if out["status"] is not None:
        assert np.array_equal(out["status"], ref["status"])


def test_noalloc_entry_matches_allocating_entry_on_two_calls_in_a_row(d4lib, entry):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.fail("this test needs a CUDA device; there is no CPU path to fall back to")
    from dftd4_b200 import workloads
    from dftd4_b200.batch import get_dispersion_batch

    param = (1.0, 2.02929367, 1.0, 0.40868035, 4.53807137, 16.0)   # B3LYP-D4 (param.f90:384-386), s9 = 1, alp = 16
    # first batch: 24 drug-like conformers of 7 elements, two of them charged
    a = workloads.conformers(24, 7)
    qa = np.zeros(24)
    qa[3], qa[5] = 1.0, -1.0
    ref_a = get_dispersion_batch(a["offsets"], a["numbers"], a["positions"], param, charges=qa, grad=True)
    # second batch: fewer, different molecules from another seed, restricted to H/C/N/O so that the species tables
    # change size and order; neutral, and without a caller-side status array (status then lives in the shared array)
    b = workloads.conformers(9, 11)
    keep = np.isin(b["numbers"], [1, 6, 7, 8])
    b["numbers"] = np.where(keep, b["numbers"], 6).astype(np.int32)
    ref_b = get_dispersion_batch(b["offsets"], b["numbers"], b["positions"], param, charges=None, grad=True)

    rc, out_a = _call_noalloc(entry, torch, a, param, qa)
    assert rc == 0, d4lib.lib.d4cu_last_error()
    rc, out_b = _call_noalloc(entry, torch, b, param, None, with_status=False)
    assert rc == 0, d4lib.lib.d4cu_last_error()
    _assert_same(out_a, ref_a)
    _assert_same(out_b, ref_b)
    # and once more the first batch, energies only: nothing of batch b may leak into it
    rc, out_a2 = _call_noalloc(entry, torch, a, param, qa, grad=False)
    assert rc == 0, d4lib.lib.d4cu_last_error()
    _assert_same(out_a2, ref_a)
    assert np.array_equal(out_a2["energy"], out_a["energy"])   # same kernel, same inputs: bitwise


def test_noalloc_entry_refuses_what_the_fixed_array_cannot_hold(d4lib, entry):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.fail("this test needs a CUDA device; there is no CPU path to fall back to")
    dev = torch.device("cuda:0")
    param = np.array((1.0, 2.02929367, 1.0, 0.40868035, 4.53807137, 16.0))
    cut = np.array(CUTOFF)
    buf = torch.zeros(64, dtype=torch.float64, device=dev)
    e_out = torch.zeros(8, dtype=torch.float64, device=dev)
    ids = torch.zeros(64, dtype=torch.int32, device=dev)

    def call(nmol, offsets, numbers):
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        numbers = np.ascontiguousarray(numbers, dtype=np.int32)
        return entry(0, None, nmol, offsets.ctypes.data_as(_lp), len(numbers), numbers.ctypes.data_as(_ip), ids.data_ptr(),
                     buf.data_ptr(), None, 3.0, 2.0, 6.0, param.ctypes.data_as(_dp), cut.ctypes.data_as(_dp), e_out.data_ptr(),
                     None, None, None)

    # more molecules than the array has offsets for: refused before anything is read or launched
    nmol = MAX_MOLECULES + 1
    assert call(nmol, np.arange(nmol + 1), [1]) == 1
    assert b"at most 65536" in d4lib.lib.d4cu_last_error()
    # more species than the array has table rows for
    assert call(1, [0, 2], np.arange(1, MAX_SPECIES + 2)) == 1
    assert b"species" in d4lib.lib.d4cu_last_error()
    # a molecule beyond the kernel's 64 atoms
    assert call(1, [0, 65], [1]) == 1
    assert b"atoms per molecule" in d4lib.lib.d4cu_last_error()
    # a species index outside the species list is found on the device and nothing is evaluated
    ids[1] = 5
    buf[:6] = torch.tensor([0.0, 0.0, 0.0, 0.0, 0.0, 1.4], dtype=torch.float64)
    assert call(1, [0, 2], [1]) == 1
    assert b"species index out of range" in d4lib.lib.d4cu_last_error()
    # and the entry point still works afterwards (the counter it tripped is cleared by the next call)
    ids[1] = 0
    e = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
    offsets = np.array([0, 2], dtype=np.int64)
    numbers = np.array([1], dtype=np.int32)
    rc = entry(0, None, 1, offsets.ctypes.data_as(_lp), 1, numbers.ctypes.data_as(_ip), ids.data_ptr(), buf.data_ptr(), None,
               3.0, 2.0, 6.0, param.ctypes.data_as(_dp), cut.ctypes.data_as(_dp), e.data_ptr(), None, None, None)
    assert rc == 0, d4lib.lib.d4cu_last_error()
    from dftd4_b200.batch import get_dispersion_batch
    ref = get_dispersion_batch([0, 2], [1, 1], [[0.0, 0.0, 0.0], [0.0, 0.0, 1.4]], tuple(param), grad=False)
    assert abs(e.item() - ref["energy"][0]) <= E_RTOL * abs(ref["energy"][0])
