"""Host-side mirror of the reference's Python API python/dftd4/interface.py:32-474.

Same classes (Structure, DampingParam, DispersionModel), argument meaning, result dictionaries and error behaviour
(ValueError for malformed input, RuntimeError(message) for library errors), bound to the CUDA library through
dftd4_b200.library instead of the reference's CFFI module.  All quantities in atomic units (Bohr, Hartree).
"""
import ctypes as C
from typing import Optional

import numpy as np

from . import library

_dp = C.POINTER(C.c_double)


class Structure:
    """Molecular structure data (interface.py:32-146): fixed number of atoms and immutable atomic numbers."""

    _mol = None

    def __init__(self, numbers: np.ndarray, positions: np.ndarray, charge: Optional[float] = None,
                 lattice: Optional[np.ndarray] = None, periodic: Optional[np.ndarray] = None):
        numbers = np.asarray(numbers)
        positions = np.asarray(positions)
        if positions.size % 3 != 0:
            raise ValueError("Expected triples of cartesian coordinates")
        if 3 * numbers.size != positions.size:
            raise ValueError("Dimension mismatch between numbers and positions")
        self._natoms = len(numbers)
        _numbers = np.ascontiguousarray(numbers, dtype="i4")
        _positions = np.ascontiguousarray(positions, dtype=float)
        _charge = _ref(charge)
        if lattice is not None:
            lattice = np.asarray(lattice)
            if lattice.size != 9:
                raise ValueError("Invalid lattice provided")
            _lattice = np.ascontiguousarray(lattice, dtype=float)
        else:
            _lattice = None
        if periodic is not None:
            periodic = np.asarray(periodic)
            if periodic.size != 3:
                raise ValueError("Invalid periodicity provided")
            _periodic = np.ascontiguousarray(periodic, dtype=bool)
        else:
            _periodic = None
        self._mol = library.new_structure(
            self._natoms,
            _numbers.ctypes.data_as(C.POINTER(C.c_int)),
            _cast(_positions),
            _charge,
            _cast(_lattice),
            None if _periodic is None else _periodic.ctypes.data_as(C.POINTER(C.c_bool)),
        )

    def __len__(self):
        return self._natoms

    def update(self, positions: np.ndarray, lattice: Optional[np.ndarray] = None) -> None:
        """Update coordinates and (optionally) lattice parameters, both in Bohr (interface.py:110-145)."""
        positions = np.asarray(positions)
        if 3 * len(self) != positions.size:
            raise ValueError("Dimension mismatch for positions")
        _positions = np.ascontiguousarray(positions, dtype=float)
        if lattice is not None:
            lattice = np.asarray(lattice)
            if lattice.size != 9:
                raise ValueError("Invalid lattice provided")
            _lattice = np.ascontiguousarray(lattice, dtype=float)
        else:
            _lattice = None
        library.update_structure(self._mol, _cast(_positions), _cast(_lattice))


class DampingParam:
    """Rational (Becke-Johnson) damping parameters (interface.py:148-235): by method name or explicit values."""

    _param = None

    def __init__(self, **kwargs):
        if "method" in kwargs and kwargs["method"] is None:
            del kwargs["method"]
        if not kwargs:
            raise TypeError("Method name or complete damping parameter set required")
        if "method" in kwargs:
            self._param = self.load_param(**kwargs)
        else:
            self._param = self.new_param(**kwargs)

    @staticmethod
    def load_param(method, atm=True):
        return library.load_rational_damping(method.encode(), bool(atm))

    @staticmethod
    def new_param(*, s6=1.0, s8, s9=1.0, a1, a2, alp=16.0):
        return library.new_rational_damping(s6, s8, s9, a1, a2, alp)

    def as_tuple(self):
        """(s6, s8, s9, a1, a2, alp) of this parameter set (extension: include/d4cu.h d4cu_param_values)."""
        import ctypes as C
        out = (C.c_double * 6)()
        library.lib.d4cu_param_values.restype = C.c_int
        library.lib.d4cu_param_values.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        if library.lib.d4cu_param_values(self._param.ptr, out) != 0:
            raise RuntimeError("Damping parameters are missing")
        return tuple(float(x) for x in out)


class DispersionModel(Structure):
    """D4 dispersion model bound to the structure it was created from (interface.py:237-474)."""

    _disp = None

    def __init__(self, numbers: np.ndarray, positions: np.ndarray, charge: Optional[float] = None,
                 lattice: Optional[np.ndarray] = None, periodic: Optional[np.ndarray] = None, model: str = "d4", **kwargs):
        Structure.__init__(self, numbers, positions, charge, lattice, periodic)
        if model.lower().replace(" ", "") == "d4":
            if "ga" in kwargs or "gc" in kwargs or "wf" in kwargs:
                self._disp = library.custom_d4_model(self._mol, kwargs.get("ga", 3.0), kwargs.get("gc", 2.0),
                                                     kwargs.get("wf", 6.0))
            else:
                self._disp = library.new_d4_model(self._mol)
        elif model.lower().replace(" ", "") == "d4s":
            if "ga" in kwargs or "gc" in kwargs:
                self._disp = library.custom_d4s_model(self._mol, kwargs.get("ga", 3.0), kwargs.get("gc", 2.0))
            else:
                self._disp = library.new_d4s_model(self._mol)
        else:
            raise ValueError(f"Unknown dispersion model '{model}'.")

    def set_realspace_cutoff(self, disp2: float, disp3: float, cn: float, width2: float = 0.0,
                             width3: float = 0.0) -> None:
        library.set_model_realspace_cutoff(self._disp, disp2, disp3, cn, width2, width3)

    def get_dispersion(self, param: DampingParam, grad: bool) -> dict:
        """Energy (and gradient + virial when grad) -- the hot path, interface.py:320-380."""
        _energy = np.array(0.0)
        if grad:
            _gradient = np.zeros((len(self), 3))
            _sigma = np.zeros((3, 3))
        else:
            _gradient = None
            _sigma = None
        library.get_dispersion(self._mol, self._disp, param._param, _cast(_energy), _cast(_gradient), _cast(_sigma))
        results = dict(energy=_energy)
        if _gradient is not None:
            results.update(gradient=_gradient)
        if _sigma is not None:
            results.update(virial=_sigma)
        return results

    def get_properties(self) -> dict:
        """CN, EEQ charges, C6 matrix and static polarisabilities (interface.py:382-432)."""
        n = len(self)
        _c6, _cn, _charges, _alpha = np.zeros((n, n)), np.zeros(n), np.zeros(n), np.zeros(n)
        library.get_properties(self._mol, self._disp, _cast(_cn), _cast(_charges), _cast(_c6), _cast(_alpha))
        return {
            "coordination numbers": _cn,
            "partial charges": _charges,
            "c6 coefficients": _c6,
            "polarizabilities": _alpha,
        }

    def get_pairwise_dispersion(self, param: DampingParam) -> dict:
        """Pair-resolved two- and three-body energies (interface.py:434-474)."""
        n = len(self)
        _pair_disp2, _pair_disp3 = np.zeros((n, n)), np.zeros((n, n))
        library.get_pairwise_dispersion(self._mol, self._disp, param._param, _cast(_pair_disp2), _cast(_pair_disp3))
        return {
            "additive pairwise energy": _pair_disp2,
            "non-additive pairwise energy": _pair_disp3,
        }

    def get_numerical_hessian(self, param: DampingParam) -> np.ndarray:
        """hess[iat][ix][jat][jx] by central differences of the analytic gradient (dftd4.h:218-221)."""
        n = len(self)
        _hess = np.zeros((n, 3, n, 3))
        library.get_numerical_hessian(self._mol, self._disp, param._param, _cast(_hess))
        return _hess


def _cast(array):
    """Pointer to a contiguous float64 array, NULL for None (interface.py:477-483)."""
    return None if array is None else array.ctypes.data_as(_dp)


def _ref(value):
    """Pointer to a double, NULL for None (interface.py:486-492)."""
    return None if value is None else C.byref(C.c_double(value))
