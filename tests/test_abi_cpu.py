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
