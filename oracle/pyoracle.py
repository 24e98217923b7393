"""ctypes front-end of the CPU oracle (oracle/d4ref.c).  TEST INFRASTRUCTURE ONLY -- never imported by dftd4_b200."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

DEFAULT_MODEL = (3.0, 2.0, 6.0)  # ga, gc, wf  (src/dftd4/model/d4.f90:56-63)
DEFAULT_CUTOFF = (60.0, 40.0, 30.0, 0.0, 0.0)  # disp2, disp3, cn, width2, width3 (src/dftd4/cutoff.f90:27-59)


def build(force=False):
    so = os.path.join(_HERE, "libd4ref.so")
    src = os.path.join(_HERE, "d4ref.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libd4ref.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.d4ref_num_threads.restype = C.c_int
    return _LIB


def use_lapack(enable=True):
    """Hand scipy's LAPACK / BLAS (OpenBLAS) entry points to the oracle: the EEQ linear algebra then runs through
    dsytrf / dsytri / dsytrs / dgemm like multicharge's `solve` (d4ref_set_lapack).  Returns True when active."""
    L = lib()
    if not enable:
        L.d4ref_set_lapack(None, None, None, None)
        return False
    try:
        from scipy.linalg import cython_blas, cython_lapack
    except Exception:  # noqa: BLE001
        return False
    get = C.pythonapi.PyCapsule_GetPointer
    get.restype = C.c_void_p
    get.argtypes = [C.py_object, C.c_char_p]
    name = C.pythonapi.PyCapsule_GetName
    name.restype = C.c_char_p
    name.argtypes = [C.py_object]
    ptrs = []
    for mod, fn in ((cython_lapack, "dsytrf"), (cython_lapack, "dsytrs"), (cython_lapack, "dsytri"), (cython_blas, "dgemm")):
        cap = mod.__pyx_capi__[fn]
        ptrs.append(C.c_void_p(get(cap, name(cap))))
    L.d4ref_set_lapack(*ptrs)
    return bool(L.d4ref_has_lapack())


def set_num_threads(n):
    """OpenMP threads of the oracle and threads of scipy's OpenBLAS (torchrun exports OMP_NUM_THREADS=1)."""
    lib().d4ref_set_num_threads(C.c_int(int(n)))
    try:
        import threadpoolctl
        threadpoolctl.threadpool_limits(limits=int(n))
    except Exception:  # noqa: BLE001
        pass
    return num_threads()


def eeq_times():
    """Seconds of the last EEQ stage: matrix build (N^2), factorisation + charges (N^3), derivative matrices (N^2),
    inverse + dq/dr products (N^3)."""
    out = np.zeros(4)
    lib().d4ref_get_eeq_times(_dp(out))
    return dict(zip(("build", "factor", "dbuild", "inverse_gemm"), out.tolist()))


def set_eeq_param(z, chi, eta, kcn, rad):
    """EEQ-2019 row of element z supplied at run time (rad <= 0 clears it)."""
    lib().d4ref_set_eeq_param(C.c_int(z), C.c_double(chi), C.c_double(eta), C.c_double(kcn), C.c_double(rad))


def row_busy():
    """Core-seconds inside the (sampled) rows of the last pair / ATM loops (d4ref_get_row_busy)."""
    out = np.zeros(2)
    lib().d4ref_get_row_busy(_dp(out))
    return dict(disp2=float(out[0]), disp3=float(out[1]))


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_int))


def _prep(numbers, positions, lattice, periodic):
    numbers = np.ascontiguousarray(numbers, dtype=np.int32)
    positions = np.ascontiguousarray(positions, dtype=np.float64).reshape(-1, 3)
    lat = None if lattice is None else np.ascontiguousarray(lattice, dtype=np.float64).reshape(3, 3)
    per = None if periodic is None else np.ascontiguousarray(periodic, dtype=np.int32)
    return numbers, positions, lat, per


def load_rational_damping(method, atm=True):
    """(s6, s8, s9, a1, a2, alp) or None -- src/dftd4/param.f90:248-303."""
    out = np.zeros(6)
    ok = lib().d4ref_load_rational_damping(method.encode(), C.c_int(1 if atm else 0), _dp(out))
    return tuple(out) if ok else None


def wrap_to_central_cell(positions, lattice, periodic=(1, 1, 1)):
    x = np.array(positions, dtype=np.float64).reshape(-1, 3).copy()
    lat = np.ascontiguousarray(lattice, dtype=np.float64)
    per = np.ascontiguousarray(periodic, dtype=np.int32)
    lib().d4ref_wrap_to_central_cell(C.c_int(len(x)), _dp(x), _dp(lat), _ip(per))
    return x


def get_dispersion(numbers, positions, param, charge=0.0, lattice=None, periodic=None, model=DEFAULT_MODEL,
                   cutoff=DEFAULT_CUTOFF, grad=True, explicit_dqdr=True, sample_stride=1):
    """param = (s6, s8, s9, a1, a2, alp).  Returns dict(energy, gradient, sigma, counts, times)."""
    numbers, positions, lat, per = _prep(numbers, positions, lattice, periodic)
    n = len(numbers)
    e = C.c_double(0.0)
    g = np.zeros((n, 3)) if grad else None
    s = np.zeros((3, 3)) if grad else None
    counts = np.zeros(8, dtype=np.int64)
    times = np.zeros(8)
    err = C.create_string_buffer(512)
    m = np.array(model, dtype=np.float64)
    p = np.array(param, dtype=np.float64)
    c = np.array(cutoff, dtype=np.float64)
    L = lib()
    L.d4ref_set_sample_stride(C.c_int(sample_stride))
    rc = L.d4ref_get_dispersion(C.c_int(n), _ip(numbers), _dp(positions), C.c_double(charge or 0.0), _dp(lat), _ip(per),
                                _dp(m), _dp(p), _dp(c), C.c_int(1 if explicit_dqdr else 0), C.byref(e), _dp(g), _dp(s),
                                counts.ctypes.data_as(C.POINTER(C.c_longlong)), _dp(times), err)
    L.d4ref_set_sample_stride(C.c_int(1))
    if rc:
        raise RuntimeError(err.value.decode() or "d4ref error %d" % rc)
    return dict(energy=e.value, gradient=g, sigma=s,
                counts=dict(pcn=int(counts[0]), p2=int(counts[1]), n3=int(counts[2]), t3=int(counts[3]), peeq=int(counts[4])),
                times=dict(zip(("cn", "eeq", "c6", "disp2", "dq", "c6_q0", "disp3", "dcn"), times.tolist())))


def get_properties(numbers, positions, charge=0.0, lattice=None, periodic=None, model=DEFAULT_MODEL, cutoff_cn=30.0):
    numbers, positions, lat, per = _prep(numbers, positions, lattice, periodic)
    n = len(numbers)
    cn, q, c6, alpha = np.zeros(n), np.zeros(n), np.zeros((n, n)), np.zeros(n)
    err = C.create_string_buffer(512)
    m = np.array(model, dtype=np.float64)
    rc = lib().d4ref_get_properties(C.c_int(n), _ip(numbers), _dp(positions), C.c_double(charge or 0.0), _dp(lat), _ip(per),
                                    _dp(m), C.c_double(cutoff_cn), _dp(cn), _dp(q), _dp(c6), _dp(alpha), err)
    if rc:
        raise RuntimeError(err.value.decode() or "d4ref error %d" % rc)
    return dict(cn=cn, charges=q, c6=c6, alpha=alpha)


def get_pairwise(numbers, positions, param, charge=0.0, lattice=None, periodic=None, model=DEFAULT_MODEL,
                 cutoff=DEFAULT_CUTOFF):
    numbers, positions, lat, per = _prep(numbers, positions, lattice, periodic)
    n = len(numbers)
    e2, e3 = np.zeros((n, n)), np.zeros((n, n))
    err = C.create_string_buffer(512)
    m = np.array(model, dtype=np.float64)
    p = np.array(param, dtype=np.float64)
    c = np.array(cutoff, dtype=np.float64)
    rc = lib().d4ref_get_pairwise(C.c_int(n), _ip(numbers), _dp(positions), C.c_double(charge or 0.0), _dp(lat), _ip(per),
                                  _dp(m), _dp(p), _dp(c), _dp(e2), _dp(e3), err)
    if rc:
        raise RuntimeError(err.value.decode() or "d4ref error %d" % rc)
    return dict(e2=e2, e3=e3)


def set_eeq_opts(alpha0=1.0e-8, tol=1.0e-8, quirk=1):
    lib().d4ref_set_eeq_opts(C.c_double(alpha0), C.c_double(tol), C.c_int(quirk))


def num_threads():
    return lib().d4ref_num_threads()


def pair_list(positions, lattice, periodic, cutoff, eps2):
    """Directed (i, j, translation index) entries passing the reference cutoff test, brute force."""
    _, positions, lat, per = _prep(np.ones(len(positions), dtype=np.int32), positions, lattice, periodic)
    n = len(positions)
    L = lib()
    L.d4ref_pair_list.restype = C.c_longlong
    ntr = C.c_int(0)
    args = (C.c_int(n), _dp(positions), _dp(lat), _ip(per), C.c_double(cutoff), C.c_double(eps2))
    total = L.d4ref_pair_list(*args, C.c_longlong(0), None, None, None, C.byref(ntr))
    i, j, t = (np.zeros(max(total, 1), dtype=np.int32) for _ in range(3))
    L.d4ref_pair_list(*args, C.c_longlong(total), _ip(i), _ip(j), _ip(t), C.byref(ntr))
    return i[:total], j[:total], t[:total], ntr.value


def set_skip_eeq(flag):
    """bench.py timing aid: evaluate with q = 0 and no dq chain rule (see d4ref_set_skip_eeq)."""
    lib().d4ref_set_skip_eeq(C.c_int(1 if flag else 0))
