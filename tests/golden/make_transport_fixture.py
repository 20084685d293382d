"""Generates tests/golden/reference_thermo_es80.npz in the build container from the one Cantera-made
coefficient file the reference ships:
  examples/dfLowMachFoam/notorch/threeD_reactingTGV/H2/cvodeIntegrator/thermo_ES80_H2-7-16.txt
(binary layout read by src_gpu/dfThermo.cu:361-435: int N; W[N]; N x 15 NASA [Tmid, high7, low7];
N x 5 viscosity, N x 5 conductivity, N x N x 5 binary-diffusion fits in ln T).  It is the only
Cantera OUTPUT in the tree at this boundary, so it pins the mechanism compiler's molecular weights
and NASA-7 tables and supplies real transport fits.  usage: python tests/golden/make_transport_fixture.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))
from dfchem_host import read_thermo_file  # noqa: E402  (pure NumPy)

SRC = "/root/reference/examples/dfLowMachFoam/notorch/threeD_reactingTGV/H2/cvodeIntegrator/thermo_ES80_H2-7-16.txt"

if __name__ == "__main__":
    d = read_thermo_file(SRC)
    out = os.path.join(ROOT, "tests", "golden", "reference_thermo_es80.npz")
    np.savez_compressed(out, **d)
    print(out, {k: v.shape for k, v in d.items()})
