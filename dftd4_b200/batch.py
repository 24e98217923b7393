"""Batched small-molecule evaluation (extension of the reference interface; BASELINE config 3).

The reference API has no batched entry: evaluating 50 000 conformers means 50 000 x (dftd4_new_structure,
dftd4_new_d4_model, dftd4_get_dispersion) -- python/dftd4/interface.py:237-380 per molecule.  `get_dispersion_batch`
takes the molecules concatenated (offsets like a CSR row pointer) and evaluates the whole batch in ONE kernel launch
(one CTA per molecule, include/d4cu.h d4cu_batch_numbers).  Results match per-molecule calls of
DispersionModel.get_dispersion to round-off; there is no CPU fallback.
"""
import ctypes as C

import numpy as np

from .library import lib

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_lp = C.POINTER(C.c_longlong)

lib.d4cu_batch_numbers.restype = C.c_int
lib.d4cu_batch_numbers.argtypes = [C.c_int, _lp, _ip, _dp, _dp, C.c_double, C.c_double, C.c_double, _dp, _dp, _dp, _dp, _dp,
                                   _ip, C.POINTER(C.c_float)]
lib.d4cu_last_error.restype = C.c_char_p

DEFAULT_CUTOFF = (60.0, 40.0, 30.0, 0.0, 0.0)


def get_dispersion_batch(offsets, numbers, positions, param, charges=None, grad=True, model=(3.0, 2.0, 6.0),
                         cutoff=DEFAULT_CUTOFF):
    """offsets[nmol+1], numbers[total], positions[total,3] (Bohr); param = DampingParam or (s6, s8, s9, a1, a2, alp).

    Returns dict(energy[nmol], gradient[total,3], virial[nmol,3,3], status[nmol], kernel_ms)."""
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    numbers = np.ascontiguousarray(numbers, dtype=np.int32)
    positions = np.ascontiguousarray(positions, dtype=np.float64)
    nmol = len(offsets) - 1
    total = int(offsets[-1])
    if numbers.shape != (total,) or positions.shape != (total, 3):
        raise ValueError("Dimension missmatch between offsets, numbers and positions")
    p = np.array(param.as_tuple() if hasattr(param, "as_tuple") else param, dtype=np.float64)
    if p.shape != (6,):
        raise ValueError("damping parameters must be (s6, s8, s9, a1, a2, alp)")
    cut = np.array(cutoff, dtype=np.float64)
    q = None if charges is None else np.ascontiguousarray(charges, dtype=np.float64)
    e = np.zeros(nmol)
    g = np.zeros((total, 3)) if grad else None
    s = np.zeros((nmol, 3, 3)) if grad else None
    st = np.zeros(nmol, dtype=np.int32)
    ms = C.c_float(0.0)
    rc = lib.d4cu_batch_numbers(nmol, offsets.ctypes.data_as(_lp), numbers.ctypes.data_as(_ip), positions.ctypes.data_as(_dp),
                                None if q is None else q.ctypes.data_as(_dp), model[0], model[1], model[2],
                                p.ctypes.data_as(_dp), cut.ctypes.data_as(_dp), e.ctypes.data_as(_dp),
                                None if g is None else g.ctypes.data_as(_dp), None if s is None else s.ctypes.data_as(_dp),
                                st.ctypes.data_as(_ip), C.byref(ms))
    if rc != 0:
        raise RuntimeError((lib.d4cu_last_error() or b"batched evaluation failed").decode())
    return dict(energy=e, gradient=g, virial=s, status=st, kernel_ms=float(ms.value))
