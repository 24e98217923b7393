"""Stage-by-stage comparison CUDA vs oracle for one synthetic test case (debug aid, run on the GPU box)."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_parity_gpu as T
from helpers import make_model
from oracle import pyoracle as O
O.build()
kind, seed = sys.argv[1], int(sys.argv[2])
req = T._full_request(kind, seed)
model, param = make_model(req)
from dftd4_b200.engine import ResidentEvaluation
ev = ResidentEvaluation(model, param); ev.upload(); xyz = ev.positions()
p = O.load_rational_damping(req["param"]["method"], True)
pr = model.get_properties()
po = O.get_properties(req["numbers"], xyz, req["charge"], req["lattice"], None, req["model"], req["cutoff"][2])
print("cn   max diff", np.max(np.abs(pr["coordination numbers"] - po["cn"])))
print("q    max diff", np.max(np.abs(pr["partial charges"] - po["charges"])), pr["partial charges"][:3], po["charges"][:3])
print("c6   max rel ", np.max(np.abs(pr["c6 coefficients"] - po["c6"]) / np.abs(po["c6"])))
pw = model.get_pairwise_dispersion(param)
ow = O.get_pairwise(req["numbers"], xyz, p, req["charge"], req["lattice"], None, req["model"], req["cutoff"])
e2, e3 = pw["additive pairwise energy"], pw["non-additive pairwise energy"]
print("e2 sum", e2.sum(), ow["e2"].sum(), "maxdiff", np.max(np.abs(e2 - ow["e2"])))
print("e3 sum", e3.sum(), ow["e3"].sum(), "maxdiff", np.max(np.abs(e3 - ow["e3"])))
r = model.get_dispersion(param, grad=True)
o = O.get_dispersion(req["numbers"], xyz, p, req["charge"], req["lattice"], None, req["model"], req["cutoff"], grad=True)
print("E", r["energy"], o["energy"], (r["energy"] - o["energy"]) / o["energy"])
print("counts", ev.compute(True) or ev.download(True)["counts"], o["counts"])
