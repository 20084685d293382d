"""Parity with the REFERENCE's own device kernels for the correctThermo path (a12 / a18 of SURVEY section 8).

oracle/_ref/libdfthermo_ref.so is src_gpu/dfThermo.cu:8-357 (kernels + constant tables) compiled for sm_100a by
oracle/ref_thermo/build_ref.sh and driven in the launch order of dfThermo::correctThermo (:588-601).  The reference caps
these kernels at 7 species (NUM_SPECIES, :11), i.e. at the ES80 mechanism, whose Cantera-made coefficient file is the
fixture tests/golden/reference_thermo_es80.npz.  Tolerances: T 1e-6 K (the reference stops Newton at |dT/T| < 1e-7),
psi, rho 1e-9, mu, alpha, rhoD 1e-9 relative."""
import os
import sys

import pytest
from conftest import ROOT

REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libdfthermo_ref.so")


def test_reference_recipe_is_committed_and_output_ignored():
    assert os.path.exists(os.path.join(ROOT, "oracle", "ref_thermo", "build_ref.sh"))
    assert os.path.exists(os.path.join(ROOT, "oracle", "ref_thermo", "ref_thermo_driver.cu"))
    assert "oracle/_ref/" in open(os.path.join(ROOT, ".gitignore")).read()


@pytest.mark.gpu
def test_thermo_and_transport_match_reference_device_kernels(built):
    if not os.path.exists(REF_LIB):
        pytest.skip("oracle/_ref/libdfthermo_ref.so not built (needs /root/reference at build time)")
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import ref_thermo_compare

    r = ref_thermo_compare.compare(n=200_003, reps=1)
    assert r["max_abs_dT_K"] < 1e-6, r
    assert r["max_rel_psi"] < 1e-9 and r["max_rel_rho"] < 1e-9, r
    assert r["max_rel_mu"] < 1e-9 and r["max_rel_alpha"] < 1e-9 and r["max_rel_rhoD"] < 1e-9, r
