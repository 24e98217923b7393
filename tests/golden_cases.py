"""Turn the harvested reference goldens (tests/golden/reference_goldens.json) into concrete evaluation requests.

Each request is a dict: kind ('dispersion' | 'properties' | 'pairwise'), numbers, positions, charge, lattice, periodic,
model (ga, gc, wf), param spec, cutoff (disp2, disp3, cn, width2, width3), and a list of checks
(quantity, expected ndarray/float, tol).  `resolve_param(spec, loader)` maps a param spec onto (s6,s8,s9,a1,a2,alp)
using a loader with the semantics of dftd4_load_rational_damping (reference src/dftd4/api.f90:582-621).
The same requests drive the oracle tests (CPU) and the CUDA parity tests (GPU).
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_CUTOFF = (60.0, 40.0, 30.0, 0.0, 0.0)
EEQ_MAXZ = 86


def _load():
    with open(os.path.join(HERE, "golden", "reference_goldens.json")) as f:
        return json.load(f)["cases"]


def _cutoff(c):
    if not c:
        return DEFAULT_CUTOFF
    return (c.get("disp2", 60.0), c.get("disp3", 40.0), c.get("cn", 30.0), c.get("width2", 0.0), c.get("width3", 0.0))


def _model(args):
    args = args or {}
    return (float(args.get("ga", 3.0)), float(args.get("gc", 2.0)), float(args.get("wf", 6.0)))


def resolve_param(spec, loader):
    """spec: {'method': name, 'atm': bool} or explicit {s6,s8,s9,a1,a2,alp}; python/dftd4/interface.py:179-233."""
    if "method" in spec and spec["method"] is not None:
        p = loader(spec["method"], bool(spec.get("atm", True)))
        if p is None:
            raise KeyError(spec["method"])
        return tuple(p)
    return (float(spec.get("s6", 1.0)), float(spec["s8"]), float(spec.get("s9", 1.0)), float(spec["a1"]),
            float(spec["a2"]), float(spec.get("alp", 16.0)))


def requests(include_unsupported=False):
    out = []
    for i, c in enumerate(_load()):
        base = dict(id="%02d-%s" % (i, c["source"].split("::")[-1].split(" ")[0].replace("/", "_")),
                    source=c["source"], numbers=np.array(c["numbers"], dtype=np.int32),
                    positions=np.array(c["positions"], dtype=np.float64), charge=c.get("charge") or 0.0,
                    lattice=None if c.get("lattice") is None else np.array(c["lattice"], dtype=np.float64),
                    periodic=None if c.get("periodic") is None else np.array(c["periodic"], dtype=np.int32),
                    cutoff=DEFAULT_CUTOFF, model=_model(None), mask=None, supported=True, why="")
        if max(c["numbers"]) > EEQ_MAXZ:
            base["supported"], base["why"] = False, "EEQ parameters for Z>86 are not in the reference tree"
        call = c["call"]
        if call in ("get_dispersion", "get_properties", "get_pairwise_dispersion"):
            base["model"] = _model(c.get("model_args"))
            if c.get("model", "d4") == "d4s":
                # D4S (src/dftd4/model/d4s.f90) has no wf argument; this repo's engines select it with wf < 0
                base["model"] = (base["model"][0], base["model"][1], -1.0)
            base["cutoff"] = _cutoff(c.get("cutoff"))
            q = c["quantity"]
            if call == "get_dispersion":
                base.update(kind="dispersion", param=c["param"], grad=bool(c.get("grad", False)),
                            checks=[(q, c["expected"], c["tol"])])
            elif call == "get_properties":
                base.update(kind="properties", checks=[(q, np.array(c["expected"]), c["tol"])])
            else:
                base.update(kind="pairwise", param=c["param"], checks=[(q, np.array(c["expected"]), c["tol"])])
        elif call == "qcschema":
            kw = c.get("keywords") or {}
            level = kw.get("level_hint", "d4").lower()
            method = c["method"].split("-")
            if method[-1].lower() == level:
                method.pop()
            method = "-".join(method) or None
            tweaks = dict(kw.get("params_tweaks", {"method": method}))
            base["model"] = (float(tweaks.pop("ga", 3.0)), float(tweaks.pop("gc", 2.0)), float(tweaks.pop("wf", 6.0)))
            if level == "d4s":
                base["model"] = (base["model"][0], base["model"][1], -1.0)
            if c.get("real") is not None:
                base["mask"] = np.array(c["real"], dtype=bool)
            grad = c["driver"] == "gradient"
            q = c["quantity"]
            if q == "return_result":
                q = "gradient" if grad else "energy"
            kind = "properties" if q == "partial charges" else "dispersion"
            base.update(kind=kind, param=tweaks, grad=grad, checks=[(q, np.array(c["expected"]), c["tol"])])
        elif call == "cli":
            opt = c["options"]
            exp = c["expected"]
            n = len(c["numbers"])
            if opt.get("properties"):
                base.update(kind="properties", checks=[
                    ("coordination numbers", np.array(exp["coordination numbers"]), c["tol"]),
                    ("partial charges", np.array(exp["partial charges"]), c["tol"]),
                    ("c6 coefficients", np.array(exp["c6 coefficients"]).reshape(n, n), c["tol"]),
                    ("polarizabilities", np.array(exp["polarizabilities"]), c["tol"])])
            else:
                spec = dict(method=opt["method"], atm=True)
                if opt["method"] == "r2scan-3c":
                    # app/driver.f90:135-144: r2SCAN-3c sets ga=2, gc=1 and s9=2 unless overridden on the command line
                    ga, gc = opt.get("zeta", [2.0, 1.0])
                    base["model"] = (float(ga), float(gc), 6.0)
                    spec["s9_override"] = float(opt.get("mbdscale", 2.0))
                if opt.get("pair"):
                    base.update(kind="pairwise", param=spec, checks=[
                        ("additive pairwise energy", np.array(exp["additive pairwise energy"]).reshape(n, n), c["tol"]),
                        ("non-additive pairwise energy", np.array(exp["non-additive pairwise energy"]).reshape(n, n), c["tol"]),
                        ("energy_sum", exp["energy"], c["tol"])])
                else:
                    checks = [("energy", exp["energy"], c["tol"])]
                    if "gradient" in exp:
                        checks.append(("gradient", np.array(exp["gradient"]).reshape(n, 3), c["tol"]))
                        checks.append(("virial", np.array(exp["virial"]).reshape(3, 3), c["tol"]))
                    base.update(kind="dispersion", param=spec, grad="gradient" in exp, checks=checks)
        else:
            continue
        if base["supported"] or include_unsupported:
            out.append(base)
    return out


def resolve(req, loader):
    """Concrete (numbers, positions, param6) honouring ghost-atom masks and the CLI's s9 override."""
    spec = dict(req.get("param") or {})
    s9o = spec.pop("s9_override", None)
    p = None
    if req["kind"] != "properties":
        p = list(resolve_param(spec, loader))
        if s9o is not None:
            p[2] = s9o
        p = tuple(p)
    numbers, positions = req["numbers"], req["positions"]
    if req["mask"] is not None:
        numbers, positions = numbers[req["mask"]], positions[req["mask"]]
    return numbers, positions, p
