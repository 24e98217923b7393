"""dftd4_b200 -- the DFT-D4 energy / gradient / virial hot path as hand-written sm_100a CUDA behind the dftd4 API.

`dftd4_b200.interface` mirrors python/dftd4/interface.py of the reference (Structure, DampingParam, DispersionModel);
`dftd4_b200.library` binds the 22 C symbols of include/dftd4.h; `dftd4_b200.engine` binds the thin d4cu_* C ABI
(include/d4cu.h) used for device-resident timing and multi-GPU sharding.  The shared library is built in-tree by
`python -m dftd4_b200.build`; nothing here falls back to a CPU implementation.
"""
__version__ = "4.2.0"  # API-compatible with the reference release (src/dftd4/version.f90:26-29)
