"""Boundary faces of correctThermo on the device (dfchem_correct_thermo_patches_dev; reference: dfChemistryModel.C:546-727,
device form src_gpu/dfThermo.cu:603-657): fixed-T patches go T -> he, the others he -> T, processor patches carry
[neighbour side | internal side] with the internal side copied from the owner cells."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_patch_loop_matches_per_patch_calls(built):
    import torch

    import dfchem_host
    from dfchem_host import synth

    name = "es80"
    h = dfchem_host.Handle(synth.load_blob(name), 0)
    ref = np.load(__import__("os").path.join(__import__("conftest").ROOT, "tests", "golden", "reference_thermo_es80.npz"))
    h.set_transport_fits(ref["visc"], ref["cond"], ref["diff"])
    N = h.N
    n_cells = 5000
    cells = synth.make_batch(name, h.species, h.W, n_cells, seed=3)
    # internal fields (consistent thermo state)
    he_c = h.h_from_T(cells["T"], cells["Y"])
    T_c, psi_c, rho_c, cp_c, _ = h.correct_thermo(he_c, cells["p"], cells["Y"], cells["T"], want_cp=True)
    mu_c, al_c, rD_c = h.transport(T_c, cells["p"], rho_c, cp_c, cells["Y"])
    # patches: generic (700 faces), fixed-T (300), empty (0), processor (400 -> 800 entries), generic (123)
    size = np.array([700, 300, 0, 400, 123], dtype=np.int32)
    ptype = np.array([0, 1, 0, 2, 0], dtype=np.int32)
    span = np.where(ptype == 2, 2 * size, size)
    nb = int(span.sum())
    faces = synth.make_batch(name, h.species, h.W, nb, seed=4)
    bT_true = faces["T"].copy()
    b_he = h.h_from_T(bT_true, faces["Y"])
    bT = bT_true + 40.0                      # the Newton guess on he -> T patches
    off = np.concatenate([[0], np.cumsum(span)])
    bT[off[1]:off[2]] = bT_true[off[1]:off[2]]        # fixed-T patch: T is the datum ...
    b_he[off[1]:off[2]] = -1.0                         # ... and he is overwritten
    rng = np.random.default_rng(0)
    face_cell = rng.integers(0, n_cells, nb).astype(np.int32)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)   # noqa: E731
    d = dict(he=t(b_he), p=t(faces["p"]), Y=t(faces["Y"]), T=t(bT))
    o = {k: torch.full((nb,), float("nan"), dtype=torch.float64, device=dev) for k in ("psi", "rho", "cp", "mu", "alpha")}
    o["rhoD"] = torch.full((N, nb), float("nan"), dtype=torch.float64, device=dev)
    ci = dict(T=t(T_c), he=t(he_c), psi=t(psi_c), rho=t(rho_c), mu=t(mu_c), alpha=t(al_c), rhoD=t(rD_c))
    fc = t(face_cell)
    ip = ctypes.POINTER(ctypes.c_int)
    h.check(h.lib.dfchem_correct_thermo_patches_dev(
        h.h, len(size), size.ctypes.data_as(ip), ptype.ctypes.data_as(ip), nb, d["he"].data_ptr(), d["p"].data_ptr(), d["Y"].data_ptr(),
        d["T"].data_ptr(), o["psi"].data_ptr(), o["rho"].data_ptr(), o["cp"].data_ptr(), o["mu"].data_ptr(), o["alpha"].data_ptr(),
        o["rhoD"].data_ptr(), 0, n_cells, fc.data_ptr(), ci["T"].data_ptr(), ci["he"].data_ptr(), ci["psi"].data_ptr(),
        ci["rho"].data_ptr(), ci["mu"].data_ptr(), ci["alpha"].data_ptr(), ci["rhoD"].data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    got = {k: v.cpu().numpy() for k, v in {**o, "T": d["T"], "he": d["he"]}.items()}
    for i in range(len(size)):
        a, sz = off[i], size[i]
        if sz == 0:
            continue
        sl = slice(a, a + sz)
        Ys = np.ascontiguousarray(faces["Y"][:, sl])
        if ptype[i] == 1:
            he_i = h.h_from_T(bT_true[sl], Ys)
            assert np.array_equal(got["he"][sl], he_i)
        else:
            he_i = b_he[sl]
        Ti, psi_i, rho_i, cp_i, _ = h.correct_thermo(np.ascontiguousarray(he_i), np.ascontiguousarray(faces["p"][sl]), Ys,
                                                     np.ascontiguousarray(bT[sl]), want_cp=True)
        assert np.abs(Ti - bT_true[sl]).max() < 1e-6
        assert np.array_equal(got["T"][sl], Ti) and np.array_equal(got["psi"][sl], psi_i) and np.array_equal(got["rho"][sl], rho_i)
        mu_i, al_i, rD_i = h.transport(Ti, np.ascontiguousarray(faces["p"][sl]), rho_i, cp_i, Ys)
        assert np.array_equal(got["mu"][sl], mu_i) and np.array_equal(got["alpha"][sl], al_i) and np.array_equal(got["rhoD"][:, sl], rD_i)
        if ptype[i] == 2:      # internal side = owner cell values
            s2 = slice(a + sz, a + 2 * sz)
            own = face_cell[sl]
            assert np.array_equal(got["T"][s2], T_c[own]) and np.array_equal(got["he"][s2], he_c[own])
            assert np.array_equal(got["psi"][s2], psi_c[own]) and np.array_equal(got["rho"][s2], rho_c[own])
            assert np.array_equal(got["mu"][s2], mu_c[own]) and np.array_equal(got["alpha"][s2], al_c[own])
            assert np.array_equal(got["rhoD"][:, s2], rD_c[:, own])
    h.close()
