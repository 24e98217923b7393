"""Thin ctypes binding of libdftd4.so -- mirrors the reference's CFFI layer python/dftd4/library.py:34-168.

Same function names, same handle life-cycle (a fresh error handle per call, RuntimeError(message) on failure,
deleters attached to the returned handles).  The reference builds an out-of-line CFFI extension against
include/dftd4.h; here the identical 22 C symbols are bound with ctypes so the package needs no compile step in Python.
The shared library holds the CUDA engine: importing this module fails loudly when it is missing.
"""
import ctypes as C
import functools
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBPATH = os.path.join(_HERE, "lib", "libdftd4.so")


def _load():
    path = os.environ.get("DFTD4_B200_LIB", _LIBPATH)
    if not os.path.exists(path):
        raise ImportError(
            "dftd4_b200: %s not found -- run `python -m dftd4_b200.build` (nvcc, sm_100a); there is no CPU fallback" % path)
    return C.CDLL(path)


lib = _load()

_vp = C.c_void_p
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_bp = C.POINTER(C.c_bool)


def _sig(name, restype, *argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


# the 22 symbols of include/dftd4.h (reference include/dftd4.h:65-230)
_sig("dftd4_get_version", C.c_int)
_sig("dftd4_new_error", _vp)
_sig("dftd4_check_error", C.c_int, _vp)
_sig("dftd4_get_error", None, _vp, C.c_char_p, _ip)
_sig("dftd4_delete_error", None, C.POINTER(_vp))
_sig("dftd4_new_structure", _vp, _vp, C.c_int, _ip, _dp, _dp, _dp, _bp)
_sig("dftd4_delete_structure", None, C.POINTER(_vp))
_sig("dftd4_update_structure", None, _vp, _vp, _dp, _dp)
_sig("dftd4_new_d4_model", _vp, _vp, _vp)
_sig("dftd4_new_d4s_model", _vp, _vp, _vp)
_sig("dftd4_custom_d4_model", _vp, _vp, _vp, C.c_double, C.c_double, C.c_double)
_sig("dftd4_custom_d4s_model", _vp, _vp, _vp, C.c_double, C.c_double)
_sig("dftd4_delete_model", None, C.POINTER(_vp))
_sig("dftd4_set_model_realspace_cutoff", None, _vp, _vp, C.c_double, C.c_double, C.c_double)
_sig("dftd4_set_model_realspace_cutoff_smooth", None, _vp, _vp, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double)
_sig("dftd4_new_rational_damping", _vp, _vp, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double)
_sig("dftd4_load_rational_damping", _vp, _vp, C.c_char_p, C.c_bool)
_sig("dftd4_delete_param", None, C.POINTER(_vp))
_sig("dftd4_get_properties", None, _vp, _vp, _vp, _dp, _dp, _dp, _dp)
_sig("dftd4_get_dispersion", None, _vp, _vp, _vp, _vp, _dp, _dp, _dp)
_sig("dftd4_get_numerical_hessian", None, _vp, _vp, _vp, _vp, _dp)
_sig("dftd4_get_pairwise_dispersion", None, _vp, _vp, _vp, _vp, _dp, _dp)


def get_api_version() -> str:
    """API version of the loaded library, e.g. '4.2.0' (library.py:39-52)."""
    v = lib.dftd4_get_version()
    return "{}.{}.{}".format(v // 10000, v % 10000 // 100, v % 100)


class _Handle:
    """Owning wrapper of an opaque library handle (the reference uses ffi.gc for the same purpose)."""

    __slots__ = ("ptr", "_deleter")

    def __init__(self, ptr, deleter):
        self.ptr = ptr
        self._deleter = deleter

    def __bool__(self):
        return bool(self.ptr)

    def __del__(self):
        if self.ptr and self._deleter is not None:
            p = _vp(self.ptr)
            try:
                self._deleter(C.byref(p))
            except Exception:
                pass
            self.ptr = None


def _raw(h):
    return h.ptr if isinstance(h, _Handle) else h


def new_error():
    return _Handle(lib.dftd4_new_error(), lib.dftd4_delete_error)


def error_check(func):
    """Run a library function with a fresh error handle; raise RuntimeError(message) on failure (library.py:67-81)."""

    @functools.wraps(func)
    def handle_error(*args, **kwargs):
        err = new_error()
        value = func(err.ptr, *[_raw(a) for a in args], **kwargs)
        if lib.dftd4_check_error(err.ptr):
            buf = C.create_string_buffer(512)
            lib.dftd4_get_error(err.ptr, buf, None)
            raise RuntimeError(buf.value.decode())
        return value

    return handle_error


def new_structure(natoms, numbers, positions, charge, lattice, periodic):
    return _Handle(error_check(lib.dftd4_new_structure)(natoms, numbers, positions, charge, lattice, periodic),
                   lib.dftd4_delete_structure)


def new_d4_model(mol):
    return _Handle(error_check(lib.dftd4_new_d4_model)(mol), lib.dftd4_delete_model)


def new_d4s_model(mol):
    return _Handle(error_check(lib.dftd4_new_d4s_model)(mol), lib.dftd4_delete_model)


def custom_d4_model(mol, ga, gc, wf):
    return _Handle(error_check(lib.dftd4_custom_d4_model)(mol, ga, gc, wf), lib.dftd4_delete_model)


def custom_d4s_model(mol, ga, gc):
    return _Handle(error_check(lib.dftd4_custom_d4s_model)(mol, ga, gc), lib.dftd4_delete_model)


def new_rational_damping(s6, s8, s9, a1, a2, alp):
    return _Handle(error_check(lib.dftd4_new_rational_damping)(s6, s8, s9, a1, a2, alp), lib.dftd4_delete_param)


def load_rational_damping(method, mbd):
    return _Handle(error_check(lib.dftd4_load_rational_damping)(method, mbd), lib.dftd4_delete_param)


def set_model_realspace_cutoff(disp, disp2, disp3, cn, width2=0.0, width3=0.0) -> None:
    """library.py:158-163: always the smooth entry point, widths default to 0."""
    return error_check(lib.dftd4_set_model_realspace_cutoff_smooth)(disp, disp2, disp3, cn, width2, width3)


update_structure = error_check(lib.dftd4_update_structure)
get_dispersion = error_check(lib.dftd4_get_dispersion)
get_numerical_hessian = error_check(lib.dftd4_get_numerical_hessian)
get_pairwise_dispersion = error_check(lib.dftd4_get_pairwise_dispersion)
get_properties = error_check(lib.dftd4_get_properties)
