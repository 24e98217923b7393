#!/usr/bin/env python3
"""Harvest the reference's OWN golden vectors for the D4 hot path into tests/golden/*.json.

The reference cannot be built or imported here (Fortran + un-vendored dependencies), so its test-suite is the only
source of pinned numbers.  This script does not compute anything: it

  * executes /root/reference/python/dftd4/test_interface.py and test_qcschema.py against RECORDING stubs of
    `dftd4.interface` / `dftd4.qcschema` / `pytest.approx`, which capture (inputs, call, expected value, tolerance)
    for every `approx(result) == expected` assertion the reference makes;
  * reads the CLI goldens app/*.json together with their input files and the flags from app/meson.build;
  * copies the inline Fortran goldens of test/unit/test_dftd4.f90 (AmF3, actinides).

Run once in the build container:  python tools/make_golden.py   (needs /root/reference; the GPU box does not have it).
"""
import json
import os
import re
import sys
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

SYMBOLS = ("h he li be b c n o f ne na mg al si p s cl ar k ca sc ti v cr mn fe co ni cu zn ga ge as se br kr rb sr y zr "
           "nb mo tc ru rh pd ag cd in sn sb te i xe cs ba la ce pr nd pm sm eu gd tb dy ho er tm yb lu hf ta w re os ir "
           "pt au hg tl pb bi po at rn fr ra ac th pa u np pu am cm bk cf es fm md no lr rf db sg bh hs mt ds rg cn nh fl "
           "mc lv ts og").split()
Z = {s: i + 1 for i, s in enumerate(SYMBOLS)}
# mctc-lib unit conversion (CODATA 2018), used by the reference for xyz/mol/gen inputs
AATOAU = 1.0 / (5.29177210903e-11 * 1.0e10)

cases = []


def tolist(x):
    return np.asarray(x, dtype=float).tolist()


# ------------------------------------------------------------------------------------------------------------------
# recording stubs
# ------------------------------------------------------------------------------------------------------------------
class Token:
    def __init__(self, case, key):
        self.case, self.key = case, key


class Approx:
    def __init__(self, val, abs=None, rel=None):
        self.val, self.abs = val, abs

    def _rec(self, other):
        if isinstance(self.val, Token):
            c = dict(self.val.case)
            c["quantity"] = self.val.key
            c["expected"] = tolist(other) if np.ndim(other) else float(other)
            c["tol"] = self.abs
            cases.append(c)
        return True

    def __eq__(self, other):
        return self._rec(other)

    def __ne__(self, other):
        return True


class Result(dict):
    def __init__(self, case):
        super().__init__()
        self.case = case

    def get(self, key):
        return Token(self.case, key)

    def __getitem__(self, key):
        return Token(self.case, key)


class DampingParam:
    def __init__(self, **kw):
        if not kw:
            raise TypeError("no args")
        if "method" in kw and len(kw) > 1 and any(k in kw for k in ("s8", "a1", "a2")):
            raise TypeError("method and params")
        if "method" not in kw:
            for k in ("s8", "a1", "a2"):
                if k not in kw:
                    raise TypeError(k)
        self.kw = kw


class Structure:
    def __init__(self, numbers, positions, charge=None, lattice=None, periodic=None):
        if len(numbers) == 0 or np.asarray(positions).size != 3 * len(numbers):
            raise ValueError("bad input")
        self.base = dict(numbers=[int(z) for z in numbers], positions=tolist(positions),
                         charge=None if charge is None else float(charge),
                         lattice=None if lattice is None else tolist(lattice),
                         periodic=None if periodic is None else [bool(p) for p in periodic])

    def update(self, positions, lattice=None):
        self.base["positions"] = tolist(positions)


class DispersionModel(Structure):
    def __init__(self, numbers, positions, charge=None, lattice=None, periodic=None, model="d4", **kw):
        super().__init__(numbers, positions, charge, lattice, periodic)
        if model not in ("d4", "d4s"):
            raise ValueError("model")
        self.base["model"] = model
        self.base["model_args"] = kw
        self.base["cutoff"] = None

    def set_realspace_cutoff(self, **kw):
        self.base["cutoff"] = kw

    def _case(self, call, param=None, **kw):
        c = json.loads(json.dumps(self.base))
        c["call"] = call
        if param is not None:
            c["param"] = param.kw
        c.update(kw)
        return c

    def get_dispersion(self, param, grad):
        return Result(self._case("get_dispersion", param, grad=bool(grad)))

    def get_properties(self):
        return Result(self._case("get_properties"))

    def get_pairwise_dispersion(self, param):
        return Result(self._case("get_pairwise_dispersion", param))

    def get_numerical_hessian(self, param):
        return Result(self._case("get_numerical_hessian", param))


class Raises:
    def __init__(self, *a, **k):
        pass

    def __enter__(self):
        return self

    def __exit__(self, et, ev, tb):
        return True


def run_interface_tests():
    src = open(os.path.join(REF, "python", "dftd4", "test_interface.py")).read()
    iface = types.ModuleType("dftd4.interface")
    iface.DampingParam, iface.DispersionModel, iface.Structure = DampingParam, DispersionModel, Structure
    pkg = types.ModuleType("dftd4")
    pkg.interface = iface
    fake_pytest = types.ModuleType("pytest")
    fake_pytest.approx, fake_pytest.raises = Approx, Raises
    saved = {k: sys.modules.get(k) for k in ("dftd4", "dftd4.interface", "pytest")}
    sys.modules.update({"dftd4": pkg, "dftd4.interface": iface, "pytest": fake_pytest})
    try:
        ns = {}
        exec(compile(src, "test_interface.py", "exec"), ns)
        for name, fn in list(ns.items()):
            if name.startswith("test_") and callable(fn):
                n0 = len(cases)
                try:
                    fn()
                except Exception as e:  # error-path tests do not produce goldens
                    print("  (skipped %s: %s)" % (name, type(e).__name__))
                for c in cases[n0:]:
                    c["source"] = "python/dftd4/test_interface.py::" + name
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def run_qcschema_tests():
    src = open(os.path.join(REF, "python", "dftd4", "test_qcschema.py")).read()

    class AtomicInput(dict):
        def __init__(self, **kw):
            super().__init__(**kw)

    class QResult:
        def __init__(self, case):
            self.success = True
            self.return_result = Token(case, "return_result")
            self.extras = {"dftd4": Result(case)}

    def run_qcschema(inp):
        mol = inp["molecule"]
        spec = inp.get("specification", inp)
        kw = spec["keywords"]
        case = dict(numbers=[Z[s.lower()] for s in mol["symbols"]], positions=tolist(mol["geometry"]),
                    charge=float(mol.get("molecular_charge", 0.0)), lattice=None, periodic=None,
                    real=mol.get("real"), driver=spec["driver"], method=spec["model"]["method"], keywords=kw,
                    call="qcschema")
        return QResult(case)

    qcs = types.ModuleType("dftd4.qcschema")
    qcs.run_qcschema = run_qcschema
    qcs.qcel_v1 = types.SimpleNamespace(AtomicInput=AtomicInput)
    qcs.qcel_v2 = None
    pkg = types.ModuleType("dftd4")
    pkg.qcschema = qcs

    class Mark:
        @staticmethod
        def skipif(*a, **k):
            return None

    fake_pytest = types.ModuleType("pytest")
    fake_pytest.approx = Approx
    fake_pytest.mark = Mark
    fake_pytest.param = lambda *a, **k: a[0]
    fake_pytest.fixture = lambda *a, **k: (lambda f: f)
    saved = {k: sys.modules.get(k) for k in ("dftd4", "dftd4.qcschema", "pytest")}
    sys.modules.update({"dftd4": pkg, "dftd4.qcschema": qcs, "pytest": fake_pytest})
    try:
        ns = {}
        exec(compile(src, "test_qcschema.py", "exec"), ns)
        for name, fn in list(ns.items()):
            if name.startswith("test_") and callable(fn) and "error" not in name:
                n0 = len(cases)
                try:
                    fn(1)
                except Exception as e:
                    print("  (partial %s: %s)" % (name, type(e).__name__))
                for c in cases[n0:]:
                    c["source"] = "python/dftd4/test_qcschema.py::" + name
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


# ------------------------------------------------------------------------------------------------------------------
# CLI goldens: parse the four input formats by hand (mctc-lib readers are not in the tree)
# ------------------------------------------------------------------------------------------------------------------
def read_tmol(path):
    numbers, pos, lat, per = [], [], None, None
    sec = None
    for line in open(path):
        t = line.split()
        if not t:
            continue
        if t[0].startswith("$"):
            sec = t[0]
            if sec == "$periodic":
                per = [True] * int(t[1])
            continue
        if sec == "$coord":
            pos.append([float(x) for x in t[:3]])
            numbers.append(Z[t[3].lower()])
        elif sec == "$lattice":
            lat = (lat or []) + [[float(x) for x in t[:3]]]
    return numbers, pos, lat, per


def read_xyz(path):
    lines = open(path).read().split("\n")
    n = int(lines[0])
    numbers, pos = [], []
    for l in lines[2:2 + n]:
        t = l.split()
        numbers.append(Z[t[0].lower()])
        pos.append([float(x) * AATOAU for x in t[1:4]])
    return numbers, pos, None, None


def read_mol(path):
    lines = open(path).read().split("\n")
    n = int(lines[3][:3])
    numbers, pos = [], []
    for l in lines[4:4 + n]:
        numbers.append(Z[l[31:34].strip().lower()])
        pos.append([float(l[0:10]) * AATOAU, float(l[10:20]) * AATOAU, float(l[20:30]) * AATOAU])
    return numbers, pos, None, None


def read_gen(path):
    lines = [l for l in open(path).read().split("\n") if l.strip()]
    n, kind = lines[0].split()[:2]
    n = int(n)
    assert kind.upper() == "C"
    syms = lines[1].split()
    numbers, pos = [], []
    for l in lines[2:2 + n]:
        t = l.split()
        numbers.append(Z[syms[int(t[1]) - 1].lower()])
        pos.append([float(x) * AATOAU for x in t[2:5]])
    return numbers, pos, None, None


def cli_cases():
    app = os.path.join(REF, "app")
    spec = [
        ("01-energy.json", "01-ammonia.tmol", read_tmol, dict(method="tpssh")),
        ("02-gradient.json", "02-nitralin.mol", read_mol, dict(method="scan", grad=True)),
        ("03-properties.json", "03-lenalidomid.gen", read_gen, dict(properties=True)),
        ("04-pair-analysis.json", "04-caffeine.xyz", read_xyz, dict(method="b3lyp", pair=True)),
        ("05-r2scan3c.json", "04-caffeine.xyz", read_xyz, dict(method="r2scan-3c")),
        ("06-r2scan3c-mod.json", "04-caffeine.xyz", read_xyz, dict(method="r2scan-3c", zeta=[4.0, 5.0], mbdscale=1.0)),
    ]
    for js, inp, reader, opt in spec:
        numbers, pos, lat, per = reader(os.path.join(app, inp))
        gold = json.load(open(os.path.join(app, js)))
        gold.pop("version", None)
        cases.append(dict(source="app/%s (input app/%s, flags from app/meson.build)" % (js, inp), call="cli",
                          numbers=numbers, positions=pos, lattice=lat, periodic=per, charge=None, options=opt,
                          expected=gold, tol=1e-9))


def fortran_inline_cases():
    """test/unit/test_dftd4.f90: AmF3 (:901-919) and actinides (:971-1002), D4 model only."""
    src = open(os.path.join(REF, "test", "unit", "test_dftd4.f90")).read()

    def grab(sub):
        m = re.search(r"subroutine %s\(error\)(.*?)end subroutine %s" % (sub, sub), src, re.S)
        body = m.group(1)
        nat = int(re.search(r"nat = (\d+)", body).group(1))
        num = [int(x) for x in re.search(r"num\(nat\) = \[(.*?)\]", body, re.S).group(1).replace("&", " ").split(",")]
        xyz = [float(x.replace("_wp", "")) for x in
               re.findall(r"[-+]?\d+\.\d+(?:[eE][-+]?\d+)?_wp", re.search(r"xyz\(3, nat\) = reshape\(\[(.*?)\]", body, re.S).group(1))]
        ref = float(re.search(r"ref = ([-+0-9.eE]+)_wp", body).group(1))
        par = {k: float(re.search(r"\b%s = ([-+0-9.eE]+)_wp" % k, body).group(1)) for k in ("s6", "s8", "s9", "a1", "a2", "alp")}
        return nat, num, xyz, ref, par

    for sub in ("test_tpsshd4atm_amf3", "test_actinides_d4"):
        try:
            nat, num, xyz, ref, par = grab(sub)
        except Exception as e:
            print("  (fortran inline %s not parsed: %s)" % (sub, e))
            continue
        cases.append(dict(source="test/unit/test_dftd4.f90::" + sub, call="get_dispersion", grad=False,
                          numbers=num, positions=np.reshape(xyz, (nat, 3)).tolist(),
                          lattice=None, periodic=None, charge=None, model="d4", model_args={}, cutoff=None,
                          param=par, quantity="energy", expected=ref, tol=2.2e-14,
                          note="needs EEQ parameters for Z>86 (multicharge, not in the reference tree)"))


def main():
    os.makedirs(OUT, exist_ok=True)
    run_interface_tests()
    run_qcschema_tests()
    cli_cases()
    fortran_inline_cases()
    with open(os.path.join(OUT, "reference_goldens.json"), "w") as f:
        json.dump(dict(generator="tools/make_golden.py", reference="dftd4 v4.2.0 (/root/reference)", cases=cases), f, indent=1)
    print("wrote %d golden cases" % len(cases))
    for c in cases:
        print("  %-60s %-24s %s" % (c["source"][-60:], c.get("quantity", c["call"]), c.get("model", "")))


if __name__ == "__main__":
    main()
