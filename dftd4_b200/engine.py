"""ctypes binding of the thin d4cu_* C ABI (include/d4cu.h): device-resident evaluation, stage timings, interaction
counts, pair lists, NCCL sharding and the FP64 peak probes.  Used by bench.py and the parity tests; the product entry
point for users stays dftd4_b200.interface (the dftd4 API)."""
import ctypes as C

import numpy as np

from .library import lib

_vp = C.c_void_p
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)

STAGES = ("lists", "cn", "eeq_build", "eeq_solve", "weights", "pair", "eeq_adjoint", "atm", "cn_chain", "total")


class Counts(C.Structure):
    _fields_ = [("pcn", C.c_longlong), ("peeq", C.c_longlong), ("p2", C.c_longlong), ("p3", C.c_longlong),
                ("t3", C.c_longlong)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


lib.d4cu_last_error.restype = C.c_char_p
lib.d4cu_device_count.restype = C.c_int
lib.d4cu_engine_of_model.restype = _vp
lib.d4cu_engine_of_model.argtypes = [_vp]
lib.d4cu_upload_handles.restype = C.c_int
lib.d4cu_upload_handles.argtypes = [_vp, _vp, _vp]
lib.d4cu_structure_positions.restype = C.c_int
lib.d4cu_structure_positions.argtypes = [_vp, _dp]
lib.d4cu_compute.restype = C.c_int
lib.d4cu_compute.argtypes = [_vp, C.c_int]
lib.d4cu_download.restype = C.c_int
lib.d4cu_download.argtypes = [_vp, _dp, _dp, _dp, C.POINTER(Counts)]
lib.d4cu_last_timing.restype = C.c_int
lib.d4cu_last_timing.argtypes = [_vp, C.POINTER(C.c_float), _ip]
lib.d4cu_last_ewald_images.restype = C.c_int
lib.d4cu_last_ewald_images.argtypes = [_vp, C.POINTER(C.c_longlong)]
lib.d4cu_pair_list.restype = C.c_longlong
lib.d4cu_pair_list.argtypes = [_vp, C.c_int, _ip, _ip, _ip]
lib.d4cu_nccl_unique_id.restype = C.c_int
lib.d4cu_nccl_unique_id.argtypes = [C.c_char_p]
lib.d4cu_engine_set_comm.restype = C.c_int
lib.d4cu_engine_set_comm.argtypes = [_vp, C.c_int, C.c_int, C.c_char_p]
lib.d4cu_partition.restype = C.c_int
lib.d4cu_partition.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _ip, _ip, _ip]
lib.d4cu_measure_fp64_peak.restype = C.c_int
lib.d4cu_measure_fp64_peak.argtypes = [C.c_int, _dp, _dp]


def _check(rc):
    if rc != 0:
        raise RuntimeError((lib.d4cu_last_error() or b"d4cu error").decode())


def device_count():
    return lib.d4cu_device_count()


class ResidentEvaluation:
    """Engine behind a DispersionModel with the structure resident in HBM: upload() once, compute() many times."""

    def __init__(self, model, param):
        self.model, self.param = model, param
        self.engine = lib.d4cu_engine_of_model(model._disp.ptr)
        if not self.engine:
            _check(1)

    def set_comm(self, rank, world, uid):
        _check(lib.d4cu_engine_set_comm(self.engine, rank, world, uid))

    def upload(self):
        _check(lib.d4cu_upload_handles(self.model._mol.ptr, self.model._disp.ptr, self.param._param.ptr))

    def positions(self):
        """Coordinates the structure handle holds (wrapped into the central cell under PBC)."""
        xyz = np.zeros((len(self.model), 3))
        _check(lib.d4cu_structure_positions(self.model._mol.ptr, xyz.ctypes.data_as(_dp)))
        return xyz

    def compute(self, grad=True):
        _check(lib.d4cu_compute(self.engine, 1 if grad else 0))

    def download(self, grad=True):
        n = len(self.model)
        e = C.c_double(0.0)
        g = np.zeros((n, 3)) if grad else None
        s = np.zeros((3, 3)) if grad else None
        c = Counts()
        _check(lib.d4cu_download(self.engine, C.byref(e), None if g is None else g.ctypes.data_as(_dp),
                                 None if s is None else s.ctypes.data_as(_dp), C.byref(c)))
        return dict(energy=e.value, gradient=g, virial=s, counts=c.as_dict())

    def timing(self):
        ms = (C.c_float * 10)()
        n = C.c_int(0)
        _check(lib.d4cu_last_timing(self.engine, ms, C.byref(n)))
        return dict(zip(STAGES, [float(x) for x in ms])), n.value

    def ewald_images(self):
        """Real-space Ewald images the periodic EEQ build of the last compute summed (all pairs i > j)."""
        n = C.c_longlong(0)
        _check(lib.d4cu_last_ewald_images(self.engine, C.byref(n)))
        return n.value

    def pair_list(self, which):
        """Directed (i, j, translation index) entries of cutoff class which = 0 cn / 1 disp2 / 2 disp3."""
        total = lib.d4cu_pair_list(self.engine, which, None, None, None)
        if total < 0:
            _check(1)
        i, j, t = (np.zeros(max(total, 1), dtype=np.int32) for _ in range(3))
        lib.d4cu_pair_list(self.engine, which, i.ctypes.data_as(_ip), j.ctypes.data_as(_ip), t.ctypes.data_as(_ip))
        return i[:total], j[:total], t[:total]


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    _check(lib.d4cu_nccl_unique_id(buf))
    return buf.raw


ROW_BLOCK = 8  # include/d4cu.h D4CU_ROW_BLOCK


def partition(nat, rank, world, block=ROW_BLOCK):
    nb, first, stride = C.c_int(), C.c_int(), C.c_int()
    if lib.d4cu_partition(nat, rank, world, block, C.byref(nb), C.byref(first), C.byref(stride)) != 0:
        raise ValueError("invalid partition request")
    rows = []
    for b in range(first.value, nb.value, stride.value):
        rows.extend(range(b * block, min(nat, (b + 1) * block)))
    return np.array(rows, dtype=np.int64)


def measure_fp64_peak(device=0):
    a, b = C.c_double(0.0), C.c_double(0.0)
    _check(lib.d4cu_measure_fp64_peak(device, C.byref(a), C.byref(b)))
    return dict(dfma_tflops=a.value, dmma_tflops=b.value)
