"""pytest configuration: `gpu` marker, import paths, shared helpers."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "deepflame-dev_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

REFERENCE_MECH = "/root/reference/mechanisms"
HAVE_REFERENCE = os.path.isdir(REFERENCE_MECH)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Native artefacts (library, mechanism compiler, emulator); built once per session."""
    import __graft_entry__ as ge

    ge.build()
    return ge


def golden(name):
    import numpy as np

    return np.load(os.path.join(ROOT, "tests", "golden", f"solve_{name}.npz"))


def rr_parity_error(RR, RR_ref, Y, rho, dt, rtol_ref=1e-9, atol_ref=1e-15):
    """Parity metric (DESIGN.md 'Parity'): |RR - RR_ref| / (1e-6*|RR_ref| + floor), where the
    absolute floor is the RR error the reference's own default CVODE tolerances admit,
    floor = rho/dt * (rtol_ref*|Y| + atol_ref)  (dfChemistryModel.C:57-58).  <= 1 passes."""
    import numpy as np

    floor = (rho / dt)[None, :] * (rtol_ref * np.abs(Y) + atol_ref)
    return np.abs(RR - RR_ref) / (1e-6 * np.abs(RR_ref) + floor)
