"""Dump energy / gradient / virial of a workload to an .npz (used to compare library variants)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dftd4_b200 import workloads
from dftd4_b200.interface import DampingParam, DispersionModel
w = workloads.get(sys.argv[1])
m = DispersionModel(w["numbers"], w["positions"], lattice=w["lattice"])
r = m.get_dispersion(DampingParam(method=w["method"]), grad=True)
np.savez(sys.argv[2], energy=float(r["energy"]), gradient=r["gradient"], virial=r["virial"])
print(sys.argv[2], "%.16e" % float(r["energy"]))
