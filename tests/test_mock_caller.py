"""The OpenFOAM-side call sequence as compiled C++: tests/mock_caller/solve_cvode_mock.cpp is built with plain g++
against include/dfchem.h (no CUDA / torch / OpenFOAM headers), holds its species fields as N separately allocated arrays
(PtrList<volScalarField>, dfChemistryModel.H:105,121) and runs solve(scalar) / solve(field) / `chemistry off` / the
integrator-failure path and correctThermo exactly as integration/dfChemistryModel_b200.patch wires them
(dfChemistryModel.C:386-408, 1010-1083, 418-543).  On the GPU its outputs must equal the ctypes path bit for bit."""
import os
import struct
import subprocess

import numpy as np
import pytest
from conftest import ROOT, golden


def _write_input(path, g, deltaT, rtol, atol):
    N, n = g["Y"].shape
    with open(path, "wb") as f:
        f.write(struct.pack("<iiddd", N, n, deltaT, rtol, atol))
        for k in ("T", "p", "rho", "dt"):
            f.write(np.ascontiguousarray(g[k], dtype="<f8").tobytes())
        f.write(np.ascontiguousarray(g["Y"], dtype="<f8").tobytes())


def _read_output(path, N, n):
    raw = np.fromfile(path, dtype="<f8")
    assert raw.size == 1 + N * n + 5 * n
    o = 1
    out = {"ret": raw[0]}
    out["RR"] = raw[o:o + N * n].reshape(N, n); o += N * n
    for k in ("Qdot", "cost", "T2", "psi", "rho2"):
        out[k] = raw[o:o + n]; o += n
    return out


def test_mock_caller_builds_against_the_header_and_fails_loudly_without_a_gpu(built, tmp_path):
    import torch

    from dfchem_host import synth

    assert os.path.exists(built.MOCK)
    src = open(os.path.join(ROOT, "tests", "mock_caller", "solve_cvode_mock.cpp")).read()
    assert '#include "dfchem.h"' in src and "cuda" not in src.lower().replace("no cuda", "") and "torch" not in src.replace("torch, OpenFOAM", "")
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    g = golden("h2_li")
    _write_input(tmp_path / "in.bin", g, 1e-6, 1e-9, 1e-15)
    r = subprocess.run([built.MOCK, synth.blob_path("h2_li"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin")],
                       capture_output=True, text=True)
    # no CUDA device -> dfchem_create fails -> the wrapper raises FatalError (exit code 3); never a CPU fallback
    assert r.returncode == 3 and "FOAM FATAL ERROR" in r.stderr and "no CUDA device" in r.stderr, (r.returncode, r.stderr)
    assert not (tmp_path / "out.bin").exists()


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/dfChemistryModel"), reason="reference tree not present")
def test_patch_applies_to_the_reference_tree(tmp_path):
    import shutil

    dst = tmp_path / "src" / "dfChemistryModel"
    dst.mkdir(parents=True)
    for f in ("dfChemistryModel.C", "dfChemistryModel.H", "CMakeLists.txt"):
        shutil.copy(os.path.join("/root/reference/src/dfChemistryModel", f), dst / f)
        os.chmod(dst / f, 0o644)
    patch = os.path.join(ROOT, "integration", "dfChemistryModel_b200.patch")
    r = subprocess.run(["patch", "-p1", "--dry-run", "-i", patch], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    txt = open(patch).read()
    for sym in ("dfchem_create", "dfchem_solve_ptrs", "dfchem_correct_thermo", "dfchem_transport", "dfchem_bind_deltaTChem", "dfchem_destroy"):
        assert sym in txt


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["h2_li", "gri30", "c7_onestep"])
def test_mock_caller_equals_ctypes_path(built, tmp_path, name):
    import dfchem_host
    from dfchem_host import GREAT, synth

    g = golden(name)
    N, n = g["Y"].shape
    rtol, atol, deltaT = 1e-9, 1e-15, 1e-6
    h = dfchem_host.Handle(synth.load_blob(name), 0)
    h.set_tolerances(rtol, atol)
    T, p, rho, Y = (np.ascontiguousarray(g[k]) for k in ("T", "p", "rho", "Y"))
    he = h.h_from_T(T, Y)
    for mode, dt in (("scalar", deltaT), ("field", np.ascontiguousarray(g["dt"]))):
        cost = np.zeros(n)
        RR, Q = h.solve(T, p, rho, dt, Y, cost=cost)
        # same operation order as the C++ caller: Y += (RR * dt) / rho
        Ynew = np.ascontiguousarray(Y + RR * (dt[None] if mode == "field" else deltaT) / rho[None])
        T2, psi, rho2, _, _ = h.correct_thermo(he, p, Ynew, T)
        _write_input(tmp_path / "in.bin", g, deltaT, rtol, atol)
        r = subprocess.run([built.MOCK, synth.blob_path(name), str(tmp_path / "in.bin"), str(tmp_path / f"out_{mode}.bin"), mode],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        o = _read_output(tmp_path / f"out_{mode}.bin", N, n)
        assert o["ret"] == (min(GREAT, 2 * deltaT) if mode == "scalar" else GREAT)      # :393-397 / :894
        assert np.array_equal(o["RR"], RR) and np.array_equal(o["Qdot"], Q) and np.array_equal(o["cost"], cost)
        assert np.array_equal(o["T2"], T2) and np.array_equal(o["psi"], psi) and np.array_equal(o["rho2"], rho2)
    # chemistry off: great (2 deltaT after the min), RR stays zero
    r = subprocess.run([built.MOCK, synth.blob_path(name), str(tmp_path / "in.bin"), str(tmp_path / "out_off.bin"), "off"],
                       capture_output=True, text=True)
    assert r.returncode == 0
    o = _read_output(tmp_path / "out_off.bin", N, n)
    assert o["ret"] == 2 * deltaT and not o["RR"].any()
    # integrator failure (max steps 1) -> FatalError like the reference's uncaught CVODE error
    r = subprocess.run([built.MOCK, synth.blob_path(name), str(tmp_path / "in.bin"), str(tmp_path / "out_fail.bin"), "fail"],
                       capture_output=True, text=True)
    if name != "c7_onestep":       # (the one-step mechanism's golden cells finish in a single step)
        assert r.returncode == 3 and "integrator failed" in r.stderr
    h.close()
