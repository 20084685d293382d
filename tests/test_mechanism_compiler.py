"""Host mechanism compiler: C++ (dfchem_mechc) vs the Python twin, blob round trip, dialect errors."""
import os
import subprocess

import numpy as np
import pytest
from conftest import HAVE_REFERENCE, REFERENCE_MECH

from oracle import mech as OM

TINY = """
units: {length: cm, quantity: mol, activation-energy: cal/mol}
phases:
- name: gas
  thermo: ideal-gas
  elements: [H, O, N]
  species: [H2, O2, H, OH, H2O, N2]
species:
- name: H2
  composition: {H: 2}
  thermo: {model: NASA7, temperature-ranges: [200.0, 1000.0, 3500.0], data: [[2.34, 7.98e-03, -1.9e-05, 2.0e-08, -7.3e-12, -917.9, 0.68], [3.33, -4.9e-05, 4.9e-07, -1.7e-10, 2.0e-14, -950.1, -3.2]]}
- name: O2
  composition: {O: 2}
  thermo: {model: NASA7, temperature-ranges: [200.0, 1000.0, 3500.0], data: [[3.78, -2.99e-03, 9.8e-06, -9.6e-09, 3.2e-12, -1063.9, 3.65], [3.28, 1.48e-03, -7.5e-07, 2.0e-10, -2.1e-14, -1088.4, 5.45]]}
- name: H
  composition: {H: 1}
  thermo: {model: NASA7, temperature-ranges: [200.0, 1000.0, 3500.0], data: [[2.5, 0, 0, 0, 0, 25473.6, -0.446], [2.5, 0, 0, 0, 0, 25473.6, -0.446]]}
- name: OH
  composition: {H: 1, O: 1}
  thermo: {model: NASA7, temperature-ranges: [200.0, 1000.0, 3500.0], data: [[3.99, -2.4e-03, 4.6e-06, -3.8e-09, 1.3e-12, 3615.0, -0.10], [3.09, 5.4e-04, 1.2e-07, -8.8e-11, 1.1e-14, 3858.6, 4.47]]}
- name: H2O
  composition: {H: 2, O: 1}
  thermo: {model: NASA7, temperature-ranges: [200.0, 1000.0, 3500.0], data: [[4.19, -2.03e-03, 6.5e-06, -5.4e-09, 1.7e-12, -30293.7, -0.849], [3.03, 2.17e-03, -1.6e-07, -9.7e-11, 1.6e-14, -30004.2, 4.96]]}
- name: N2
  composition: {N: 2}
  thermo: {model: NASA7, temperature-ranges: [300.0, 5000.0], data: [[3.29, 1.4e-03, -3.9e-06, 5.6e-09, -2.4e-12, -1020.8, 3.95]]}
reactions:
- equation: H + O2 <=> OH + OH  # made-up test reaction
  rate-constant: {A: 2.0e+14, b: 0.0, Ea: 16000.0}
- equation: 2 H + M <=> H2 + M
  type: three-body
  rate-constant: {A: 1.0e+18, b: -1.0, Ea: 0.0}
  efficiencies: {H2: 2.5, H2O: 12.0}
- equation: H + OH (+M) <=> H2O (+M)
  type: falloff
  low-P-rate-constant: {A: 2.2e+22, b: -2.0, Ea: 0.0}
  high-P-rate-constant: {A: 1.0e+14, b: 0.0, Ea: "1.0 kcal/mol"}
  Troe: {A: 0.7, T3: 100.0, T1: 2000.0}
- equation: H2 + O2 => 2 OH
  rate-constant: {A: 1.7e+13, b: 0.0, Ea: 47780.0}
  orders: {H2: 0.5, O2: 1.25}
"""


def test_tiny_mechanism_cpp_matches_python(built, tmp_path):
    y = tmp_path / "tiny.yaml"
    y.write_text(TINY)
    out = tmp_path / "tiny.blob"
    subprocess.run([built.MECHC, str(y), str(out)], check=True)
    m = OM.load_mechanism(str(y))
    assert out.read_bytes() == OM.serialize(m)
    assert m.N == 6 and m.R == 4
    assert list(m.kind) == [OM.ELEMENTARY, OM.THREE_BODY, OM.FALLOFF_TROE, OM.ELEMENTARY]
    assert list(m.reversible) == [1, 1, 1, 0]
    # unit rules: bimolecular A in cm3/mol/s -> m3/kmol/s is 1e-3 per order above one
    assert np.isclose(m.A[0], 2.0e14 * 1e-3)
    assert np.isclose(m.A[1], 1.0e18 * 1e-6)            # 2 H + M: order 3
    assert np.isclose(m.A0[2], 2.2e22 * 1e-6) and np.isclose(m.A[2], 1.0e14 * 1e-3)
    assert np.isclose(m.Ta[2], 1000.0 * 4.184 * 1000 / OM.GAS_CONSTANT)  # explicit unit overrides block
    assert np.isclose(m.A[3], 1.7e13 * (1e-3) ** 0.75)  # orders 0.5 + 1.25
    assert np.isinf(m.troe[2, 3])                        # 3-parameter Troe
    assert m.eff[1, m.species.index("H2O")] == 12.0 and m.eff[1, m.species.index("N2")] == 1.0
    m2 = OM.deserialize(out.read_bytes())
    assert np.array_equal(m2.r_ord, m.r_ord) and m2.species == m.species


@pytest.mark.skipif(not HAVE_REFERENCE, reason="reference mechanisms only exist in the build container")
@pytest.mark.parametrize("rel", ["H2/ES80_H2-7-16.yaml", "H2/H2_Li.yaml", "H2/H2_AR.yaml", "H2/H2_Ja.yaml",
                                 "CH4/drm19.yaml", "CH4/gri30.yaml", "C7H16/C7_oneStep.yaml",
                                 "C2H5OH/ethanol66.yaml"])
def test_reference_mechanisms_cpp_matches_python(built, tmp_path, rel):
    out = tmp_path / "m.blob"
    src = os.path.join(REFERENCE_MECH, rel)
    subprocess.run([built.MECHC, src, str(out)], check=True)
    assert out.read_bytes() == OM.serialize(OM.load_mechanism(src))


@pytest.mark.skipif(not HAVE_REFERENCE, reason="reference mechanisms only exist in the build container")
def test_committed_blobs_are_current(built):
    from dfchem_host import synth

    for name, rel in built.MECH_FILES.items():
        m = OM.load_mechanism(os.path.join(REFERENCE_MECH, rel))
        assert synth.load_blob(name) == OM.serialize(m), name


def test_sizes_of_committed_blobs():
    from dfchem_host import synth

    expect = {"es80": (7, 16), "h2_li": (9, 21), "drm19": (21, 84), "gri30": (36, 219), "c7_onestep": (5, 1),
              "ethanol66": (34, 69)}
    for name, (N, R) in expect.items():
        m = OM.deserialize(synth.load_blob(name))
        assert (m.N, m.R) == (N, R)


def test_unsupported_dialect_is_an_error(built, tmp_path):
    bad = TINY.replace("type: three-body", "type: pressure-dependent-Arrhenius")
    y = tmp_path / "bad.yaml"
    y.write_text(bad)
    r = subprocess.run([built.MECHC, str(y), str(tmp_path / "o.blob")], capture_output=True)
    assert r.returncode != 0 and b"error" in r.stderr
    r = subprocess.run([built.MECHC, str(tmp_path / "missing.yaml")], capture_output=True)
    assert r.returncode != 0
