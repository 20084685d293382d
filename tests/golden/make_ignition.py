"""Generates tests/golden/ignition_es80.npz in the build container: the temperature history of the
reference's C1 case (df0DFoam 0-D H2/air constant-pressure ignition, ES80 mechanism, T0 = 1000 K,
1 atm, Y from examples/df0DFoam/zeroD_cubicReactor/H2/cvodeIntegrator/0, deltaT = 1e-6 s) computed by
the converged oracle's operator-split loop (oracle/integrate.py::zero_d_const_pressure, which follows
df0DFoam.C:106-111, YEqn.H:15, EEqn.H:5).  usage: python tests/golden/make_ignition.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))
from dfchem_host import synth  # noqa: E402
from oracle import integrate as OI  # noqa: E402
from oracle import mech as OM  # noqa: E402

NSTEPS = 400


def main():
    m = OM.deserialize(synth.load_blob("es80"))
    Y0 = np.zeros(m.N)
    Y0[m.species.index("H2")] = 0.02555366261652
    Y0[m.species.index("O2")] = 0.227143668370244
    Y0[m.species.index("N2")] = 0.7473026689382
    Ts = OI.zero_d_const_pressure(m, 1000.0, 101325.0, Y0, 1e-6, NSTEPS, rtol=1e-10, atol=1e-24)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ignition_es80.npz"), T=Ts, dt=1e-6, Y0=Y0,
                        species=np.array(m.species))
    print("T(end)", Ts[-1], "argmax dT", int(np.argmax(np.diff(Ts))))


if __name__ == "__main__":
    main()
