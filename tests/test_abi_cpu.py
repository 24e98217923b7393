"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol the headers declare, and
the handle / error conventions of the reference C-API (src/dftd4/api.f90) hold.  No compute calls (no GPU here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header, prefix):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(%s[a-z0-9_]+)\s*\(" % prefix, text)) - {"dftd4_delete"})


def test_exports_all_declared_symbols(d4lib):
    names = _declared("dftd4.h", "dftd4_") + _declared("d4cu.h", "d4cu_")
    assert len(_declared("dftd4.h", "dftd4_")) == 22  # the reference surface, include/dftd4.h:65-230
    missing = [n for n in names if not hasattr(d4lib.lib, n)]
    assert not missing, missing


def test_version(d4lib):
    assert d4lib.lib.dftd4_get_version() == 40200
    assert d4lib.get_api_version() == "4.2.0"


def test_error_handle_convention(d4lib):
    lib = d4lib.lib
    assert lib.dftd4_check_error(None) == 2  # api.f90:146-167
    err = C.c_void_p(lib.dftd4_new_error())
    assert lib.dftd4_check_error(err) == 0
    # NULL structure -> message through the handle (api.f90:327-333)
    assert lib.dftd4_new_d4_model(err, None) is None
    assert lib.dftd4_check_error(err) == 1
    buf = C.create_string_buffer(64)
    lib.dftd4_get_error(err, buf, None)
    assert buf.value == b"Molecular structure data is missing"
    small = C.create_string_buffer(10)
    lib.dftd4_get_error(err, small, C.byref(C.c_int(10)))
    assert small.value == b"Molecular"  # min(len-1, len_trim) characters, NUL-terminated (api.f90:894-904)
    lib.dftd4_delete_error(C.byref(err))
    assert err.value is None  # deleters reset the caller's handle (api.f90:127-142)
    # NULL error handle: silent no-op, NULL result (api.f90:222)
    assert lib.dftd4_new_structure(None, 1, None, None, None, None, None) is None


def test_interface_input_validation(d4lib):
    """python/dftd4/test_interface.py:42-102 of the reference."""
    from dftd4_b200.interface import DampingParam, DispersionModel, Structure
    rng = np.random.default_rng(0)
    numbers = np.array([6, 7, 6, 7, 6, 6, 6, 8, 7, 6, 8, 7, 6, 6])
    positions = rng.uniform(-5, 5, size=(14, 3))
    with pytest.raises(ValueError, match="Dimension mismatch"):
        Structure(numbers, positions[:7])
    with pytest.raises(ValueError, match="Expected triples"):
        Structure(numbers, rng.uniform(size=7))
    with pytest.raises(ValueError, match="Invalid lattice"):
        Structure(numbers, positions, lattice=rng.uniform(size=(2, 3)))
    with pytest.raises(ValueError, match="Invalid periodicity"):
        Structure(numbers, positions, periodic=np.array([True, False]))
    with pytest.raises(RuntimeError, match="Too close interatomic distances found"):
        Structure(numbers, np.zeros((14, 3)))
    mol = Structure(numbers, positions)
    assert len(mol) == 14
    with pytest.raises(ValueError, match="Dimension mismatch for positions"):
        mol.update(positions[:7])
    with pytest.raises(RuntimeError, match="Too close interatomic distances found"):
        mol.update(np.zeros((14, 3)))
    with pytest.raises(TypeError):
        DampingParam()
    with pytest.raises(TypeError):
        DampingParam(s8=1.0, a1=0.4)
    with pytest.raises(RuntimeError, match="Functional 'abc' not known"):
        DampingParam(method="abc")
    with pytest.raises(ValueError, match="Unknown dispersion model"):
        DispersionModel(numbers, positions, model="d5")
    with pytest.raises(RuntimeError, match="unsupported element"):
        DispersionModel(np.array([104, 1]), positions[:2])  # Z = 104..111 (d4.f90:100-106)
    DampingParam(method="PBE0/def2-QZVP")  # lower-cased, basis suffix dropped (param.f90:263-265)


def test_damping_database_matches_oracle(d4lib, oracle):
    """The library's functional database (param.f90 tables) against the oracle's, via the energy-free route: both are
    generated from the reference source by tools/gen_tables.py, so they must agree for every known name."""
    from dftd4_b200.interface import DampingParam
    for name in ("pbe", "pbe0", "b3lyp", "r2scan", "tpss", "scan", "b97m", "blyp", "m06l", "tpssh", "lh20t"):
        for atm in (True, False):
            assert oracle.load_rational_damping(name, atm) is not None
            DampingParam(method=name, atm=atm)


def test_no_cpu_fallback_without_device(d4lib):
    """Without a CUDA device every compute entry point must fail loudly through the error handle."""
    from dftd4_b200 import engine
    from dftd4_b200.interface import DampingParam, DispersionModel
    if engine.device_count() > 0:
        pytest.skip("a CUDA device is present")
    disp = DispersionModel(np.array([8, 1, 1]), np.array([[0, 0, -0.7], [1.4, 0, 0.4], [-1.4, 0, 0.4]]))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        disp.get_dispersion(DampingParam(method="pbe"), grad=True)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        disp.get_properties()


def test_wrap_matches_oracle(d4lib, oracle):
    """Structure wrapping (utils.f90:29-65) in the library's host code is bit-identical to the oracle's restatement:
    checked through the cold-fusion predicate on a structure that only collides after wrapping."""
    from dftd4_b200.interface import Structure
    lat = np.array([[7.0, 0.0, 0.0], [1.0, 8.0, 0.0], [0.5, 0.5, 9.0]])
    pos = np.array([[0.3, 0.4, 0.5], [0.3 + 7.0 + 1.0, 0.4 + 8.0, 0.5]])  # atom 2 = atom 1 + a + b
    with pytest.raises(RuntimeError, match="Too close"):
        Structure(np.array([1, 1]), pos, lattice=lat)
    w = oracle.wrap_to_central_cell(pos, lat)
    assert np.allclose(w[0], w[1], atol=1e-12)


def _amf3():
    """AmF3 of the reference's actinide golden (test/unit/test_dftd4.f90:901-919): geometry only."""
    numbers = np.array([95, 9, 9, 9])
    positions = np.array([[-1.13163973200000, -2.17446990100000, +1.10012477100000],
                          [-4.66377948900000, -3.12947883400000, -0.36987606800000],
                          [-0.19032564300000, +1.36339950600000, -0.36521789300000],
                          [+1.46283310800000, -4.75734549200000, -0.36503081000000]])
    return numbers, positions


def test_eeq_parameter_rows_as_runtime_input(d4lib, tmp_path):
    """EEQ-2019 rows for the elements the compiled table lacks (Z = 87..103, reference test_dftd4.f90:901-1044) and
    overrides of compiled rows arrive at run time: d4cu_load_eeq_params / d4cu_set_eeq_params / D4CU_EEQ_PARAMS.
    Without a row the model build fails loudly; with one the row reaches the tables the engine uploads."""
    from dftd4_b200.interface import DispersionModel
    lib = d4lib.lib
    lib.d4cu_get_eeq_params.argtypes = [C.c_int, C.POINTER(C.c_double)]
    lib.d4cu_model_eeq_params.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
    lib.d4cu_set_eeq_params.argtypes = [C.c_int] + [C.c_double] * 4
    lib.d4cu_load_eeq_params.argtypes = [C.c_char_p]
    lib.d4cu_last_error.restype = C.c_char_p
    numbers, positions = _amf3()
    row = (C.c_double * 4)()
    lib.d4cu_clear_eeq_params()
    assert lib.d4cu_get_eeq_params(9, row) == 0 and abs(row[0] - 1.57540015) < 1e-12   # compiled row of F
    assert lib.d4cu_get_eeq_params(95, row) == 1                                       # no row for Am
    with pytest.raises(RuntimeError, match="EEQ-2019 parameters for element Z=95 are not available"):
        DispersionModel(numbers, positions)
    # a synthetic file: Am added, F overridden (values are placeholders, not multicharge's)
    f = tmp_path / "eeq.txt"
    f.write_text("# Z chi eta kcn rad\n95 0.61 0.05 0.021 1.9   # Am\n 9 1.50 0.30 0.100 0.7\n")
    assert lib.d4cu_load_eeq_params(str(f).encode()) == 0
    try:
        disp = DispersionModel(numbers, positions)
        assert lib.d4cu_model_eeq_params(disp._disp.ptr, 95, row) == 0
        assert list(row) == [0.61, 0.05, 0.021, 1.9]
        assert lib.d4cu_model_eeq_params(disp._disp.ptr, 9, row) == 0
        assert list(row) == [1.50, 0.30, 0.100, 0.7]
        assert lib.d4cu_set_eeq_params(9, 1.6, 0.31, 0.11, 0.71) == 0
        disp2 = DispersionModel(numbers, positions)
        assert lib.d4cu_model_eeq_params(disp2._disp.ptr, 9, row) == 0 and row[0] == 1.6
    finally:
        lib.d4cu_clear_eeq_params()
    # malformed input is an error, not a guess
    bad = tmp_path / "bad.txt"
    bad.write_text("95 0.61 0.05\n")
    assert lib.d4cu_load_eeq_params(str(bad).encode()) == 1
    assert b"malformed row at line 1" in lib.d4cu_last_error()
    assert lib.d4cu_load_eeq_params(str(tmp_path / "missing.txt").encode()) == 1


def test_compiled_eeq_table_is_frozen_and_sane(d4lib, oracle):
    """The compiled EEQ-2019 table is not under /root/reference (multicharge is un-vendored): freeze it by checksum so
    that an accidental edit of one digit cannot pass silently, check every row for plausibility (finite, radius > 0),
    and -- oracle-independent physics -- that every element Z = 1..86 gives a neutral, finite diatomic with hydrogen
    (X-H at 3 bohr: charges sum to zero, |q| < 1.5) through the oracle's EEQ solve."""
    import hashlib
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dftd4_b200", "csrc", "eeq2019_params.inc")
    text = open(path).read()
    arrs = []
    for name in ("chi", "eta", "kcn", "rad"):
        body = re.search(r"eeq2019_%s\[EEQ2019_MAXZ \+ 1\] = \{(.*?)\};" % name, text, re.S).group(1)
        vals = [float(x) for x in re.findall(r"-?\d+\.\d+(?:[eE][-+]?\d+)?", body)]
        assert len(vals) == 87 and vals[0] == 0.0
        arrs.append(vals)
    assert hashlib.sha256(repr(arrs).encode()).hexdigest() == "3bc88057ace93d05238ee2c6126f5524d55e16af48a286c73efedf4ed9c3d866"
    lib = d4lib.lib
    lib.d4cu_get_eeq_params.argtypes = [C.c_int, C.POINTER(C.c_double)]
    lib.d4cu_clear_eeq_params()
    row = (C.c_double * 4)()
    for z in range(1, 87):
        assert lib.d4cu_get_eeq_params(z, row) == 0
        assert list(row) == [arrs[0][z], arrs[1][z], arrs[2][z], arrs[3][z]]
        assert all(np.isfinite(list(row))) and row[3] > 0.0 and 0.0 < row[0] < 3.0
        out = oracle.get_properties(np.array([z, 1], dtype=np.int32), np.array([[0.0, 0.0, 0.0], [0.0, 0.0, 3.0]]))
        q = out["charges"]
        assert abs(q.sum()) < 1e-12 and np.all(np.isfinite(q)) and np.max(np.abs(q)) < 1.5, (z, q)
