"""Mechanism classes the six shipped blobs do not reach: (a) the case-local mechanisms of the reference's examples/
(H2_burke 13 species, Burke2012_s9r23, C2H4_r10 with non-elementary orders) through both front ends, the table builder
and the emulated kernel; (b) synthetic mechanisms that land in the NP = 40 / 48 / 64 kernel instantiations
(csrc/dispatch.hpp), which no reference mechanism needs; (c) the real-fluid case is rejected, not mis-read."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
from conftest import HAVE_REFERENCE

from oracle import kinetics as K
from oracle import mech as OM

P = ctypes.POINTER(ctypes.c_double)
CASE_LOCAL = {
    "H2_burke": "examples/dfHighSpeedFoam/oneD_detH2WaterMist/H2_burke.yaml",
    "Burke2012_s9r23": "examples/dfLowMachFoam/paddle/twoD_HIT_flame/H2/Burke2012_s9r23.yaml",
    "C2H4_r10": "examples/dfHighSpeedFoam/twoD_detC2H4WaterMist/C2H4_r10.yaml",
}
N2_LIKE = "[[3.298677, 1.4082404e-03, -3.963222e-06, 5.641515e-09, -2.444854e-12, {h}, 3.950372], [2.92664, 1.4879768e-03, -5.68476e-07, 1.0097038e-10, -6.753351e-15, {h2}, 5.980528]]"


def synthetic_mechanism(n_active):
    """A made-up mechanism with n_active reacting species A0..A{n-1} plus inert N2: isomerisation chain A_i <=> A_{i+1},
    bimolecular exchange A_i + A_j <=> A_k + A_l, a few three-body and falloff steps.  Every species has the same
    composition (so any such reaction is balanced) and an N2-like NASA-7 fit with its own formation enthalpy."""
    sp = [f"A{i}" for i in range(n_active)]
    out = ["units: {length: cm, quantity: mol, activation-energy: cal/mol}", "phases:", "- name: gas", "  thermo: ideal-gas",
           "  elements: [N]", f"  species: [{', '.join(sp + ['N2'])}]", "species:"]
    for i, s in enumerate(sp + ["N2"]):
        h = -1020.9 + (0.0 if s == "N2" else 400.0 * ((i * 7) % 11 - 5))
        out += [f"- name: {s}", "  composition: {N: 2}",
                "  thermo: {model: NASA7, temperature-ranges: [300.0, 1000.0, 5000.0], data: " + N2_LIKE.format(h=h, h2=h + 98.0) + "}"]
    out.append("reactions:")
    for i in range(n_active - 1):
        out += [f"- equation: A{i} <=> A{i + 1}", f"  rate-constant: {{A: {3.0e6 * (1 + i % 5):.3e}, b: 0.5, Ea: {8000.0 + 900.0 * (i % 7)}}}"]
    for i in range(0, n_active - 3, 2):
        out += [f"- equation: A{i} + A{i + 3} <=> A{i + 1} + A{i + 2}", f"  rate-constant: {{A: {2.0e12 * (1 + i % 3):.3e}, b: 0.0, Ea: {3000.0 + 500.0 * (i % 5)}}}"]
    for i in range(0, n_active - 1, 7):
        out += [f"- equation: 2 A{i} + M <=> A{i + 1} + A{i} + M", "  type: three-body", f"  rate-constant: {{A: 4.0e15, b: -0.5, Ea: {1000.0 + 100.0 * i}}}",
                f"  efficiencies: {{A{i}: 2.5, N2: 0.7}}"]
    for i in range(1, n_active - 2, 9):
        out += [f"- equation: A{i} (+M) <=> A{i + 2} (+M)", "  type: falloff", "  low-P-rate-constant: {A: 6.0e+14, b: 0.0, Ea: 9000.0}",
                "  high-P-rate-constant: {A: 2.0e+08, b: 0.4, Ea: 12000.0}", "  Troe: {A: 0.6, T3: 120.0, T1: 1800.0, T2: 5200.0}"]
    return "\n".join(out) + "\n"


@pytest.fixture(scope="module")
def emu(built):
    lib = ctypes.CDLL(built.EMU)
    lib.emu_create.restype = ctypes.c_void_p
    lib.emu_create.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    return lib


def random_states(m, n, seed):
    rng = np.random.default_rng(seed)
    Y = rng.uniform(0.0, 1.0, (n, m.N)) ** 3
    Y /= Y.sum(1, keepdims=True)
    T = rng.uniform(900.0, 2400.0, n)
    p = rng.uniform(0.5, 8.0, n) * 101325.0
    return T, p, Y


def emu_wdot_error(emu, blob, m, seed=1, n=24):
    h = ctypes.c_void_p(emu.emu_create(blob, len(blob)))
    assert h.value, "table builder rejected the mechanism"
    info = (ctypes.c_int * 15)()
    emu.emu_info(h, info)
    T, p, Y = random_states(m, n, seed)
    rho = K.density(m, T, p, Y)
    ref = (K.wdot(m, K.RateConstants(m, T), rho[:, None] * Y / m.W[None]) * m.W[None]).T
    wr = np.zeros((m.N, n))
    Yt = np.ascontiguousarray(Y.T)
    assert emu.emu_wdot(h, n, T.ctypes.data_as(P), p.ctypes.data_as(P), Yt.ctypes.data_as(P), n, wr.ctypes.data_as(P), n) == 0
    emu.emu_destroy(h)
    return float((np.abs(wr - ref) / (np.abs(ref).max(0, keepdims=True) + 1e-300)).max()), int(info[2])


@pytest.mark.skipif(not HAVE_REFERENCE, reason="reference examples only exist in the build container")
@pytest.mark.parametrize("name", list(CASE_LOCAL))
def test_case_local_mechanisms_compile_and_evaluate(built, emu, tmp_path, name):
    src = os.path.join("/root/reference", CASE_LOCAL[name])
    out = tmp_path / "m.blob"
    subprocess.run([built.MECHC, src, str(out)], check=True)
    m = OM.load_mechanism(src)
    blob = out.read_bytes()
    assert blob == OM.serialize(m)                       # C++ front end == PyYAML twin, byte for byte
    err, NP = emu_wdot_error(emu, blob, m)
    assert err < 1e-11, (name, err)


@pytest.mark.skipif(not HAVE_REFERENCE, reason="reference examples only exist in the build container")
def test_real_fluid_case_is_rejected(built, tmp_path):
    src = "/root/reference/examples/dfLowMachFoam/notorch/twoD_supercirticalTGV/2s_CH4_BFER_N2.yaml"
    r = subprocess.run([built.MECHC, src, str(tmp_path / "o.blob")], capture_output=True)
    assert r.returncode != 0 and b"error" in r.stderr     # Peng-Robinson / non-ideal-gas phase: outside the dialect


@pytest.mark.parametrize("n_active,NP", [(38, 40), (45, 48), (60, 64)])
def test_large_kernel_classes_on_a_synthetic_mechanism(built, emu, tmp_path, n_active, NP):
    y = tmp_path / "syn.yaml"
    y.write_text(synthetic_mechanism(n_active))
    out = tmp_path / "syn.blob"
    subprocess.run([built.MECHC, str(y), str(out)], check=True)
    m = OM.load_mechanism(str(y))
    blob = out.read_bytes()
    assert blob == OM.serialize(m) and m.N == n_active + 1
    err, np_used = emu_wdot_error(emu, blob, m)
    assert np_used == NP and err < 1e-11
    # one chemistry sub-step of three cells through the emulated kernel vs the NumPy model of the same integrator run tighter
    from oracle import rosenbrock as RB

    T, p, Y = random_states(m, 3, 7)
    rho = K.density(m, T, p, Y)
    Cref = RB.integrate(m, T, rho, rho[:, None] * Y / m.W[None], 1e-6, 1e-12, 1e-22)
    RRref = ((Cref * m.W[None] / rho[:, None]) - Y).T / 1e-6 * rho[None]
    h = ctypes.c_void_p(emu.emu_create(blob, len(blob)))
    RR, Q, cost = np.zeros((m.N, 3)), np.zeros(3), np.zeros(3)
    st = np.zeros(3, np.int32)
    cnt = (ctypes.c_ulonglong * 4)()
    Yt = np.ascontiguousarray(Y.T)
    rc = emu.emu_solve(h, 3, T.ctypes.data_as(P), p.ctypes.data_as(P), rho.ctypes.data_as(P), None, ctypes.c_double(1e-6),
                       Yt.ctypes.data_as(P), 3, RR.ctypes.data_as(P), 3, Q.ctypes.data_as(P), cost.ctypes.data_as(P),
                       st.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), ctypes.c_double(1e-10), ctypes.c_double(1e-20), 200000, cnt)
    emu.emu_destroy(h)
    assert rc == 0 and (st == 0).all()
    scale = np.abs(RRref).max(0, keepdims=True)
    assert (np.abs(RR - RRref) / scale).max() < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("n_active", [38, 45, 60])
def test_large_kernel_classes_on_the_gpu(built, tmp_path, n_active):
    import dfchem_host

    y = tmp_path / "syn.yaml"
    y.write_text(synthetic_mechanism(n_active))
    h = dfchem_host.Handle(str(y), 0)                     # the C++ YAML front end inside the library
    m = OM.load_mechanism(str(y))
    T, p, Y = random_states(m, 64, 3)
    rho = K.density(m, T, p, Y)
    Yt = np.ascontiguousarray(Y.T)
    ref = (K.wdot(m, K.RateConstants(m, T), rho[:, None] * Y / m.W[None]) * m.W[None]).T
    wr = h.wdot(T, p, Yt)
    assert (np.abs(wr - ref) / np.abs(ref).max(0, keepdims=True)).max() < 1e-11
    h.set_tolerances(1e-10, 1e-20)
    st = np.zeros(64, np.int32)
    RR, Q = h.solve(T, p, rho, 1e-6, Yt, status=st)
    assert (st == 0).all()
    from oracle import rosenbrock as RB

    Cref = RB.integrate(m, T[:4], rho[:4], rho[:4, None] * Y[:4] / m.W[None], 1e-6, 1e-12, 1e-22)
    RRref = ((Cref * m.W[None] / rho[:4, None]) - Y[:4]).T / 1e-6 * rho[None, :4]
    assert (np.abs(RR[:, :4] - RRref) / np.abs(RRref).max(0, keepdims=True)).max() < 1e-6
    assert (np.abs(RR.sum(0)) <= 1e-8 * np.abs(RR).sum(0) + 1e-6).all()       # mass conservation
    h.close()
