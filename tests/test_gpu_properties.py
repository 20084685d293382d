"""Full-size, size-independent properties of the CUDA path (no oracle at these sizes), the
host-mirror class `DfChemistryModel`, the chunked host pipeline and the device-pointer entry."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api(built):
    import dfchem_host
    from dfchem_host import synth

    return dfchem_host, synth


def comp_matrix(name):
    from dfchem_host import synth
    from oracle import mech as OM

    m = OM.deserialize(synth.load_blob(name))
    return m.comp, m.W


@pytest.mark.parametrize("name,n", [("gri30", 1 << 18), ("es80", 1 << 20), ("c7_onestep", 1 << 18)])
def test_conservation_and_determinism_at_scale(api, name, n):
    """BASELINE.json config sizes (CH4: a quarter of the 4M batch to keep the test short):
    mass conservation sum_k RR_k = 0, element conservation sum_k a_ek RR_k / W_k = 0,
    Qdot = -sum hc_k RR_k, bounded cost, bit-identical reruns."""
    dfchem_host, synth = api
    h = dfchem_host.Handle(synth.load_blob(name), 0)
    h.set_tolerances(1e-6, 1e-10)
    b = synth.make_batch(name, h.species, h.W, n, seed=99)
    cost, st = np.zeros(n), np.zeros(n, np.int32)
    RR, Q = h.solve(b["T"], b["p"], b["rho"], 1e-6, b["Y"], cost=cost, status=st)
    assert (st == 0).all() and np.isfinite(RR).all() and np.isfinite(Q).all()
    # RR is taken against the RAW Y while the integrator advances Y/sum(Y) (dfChemistryModel.C:752-770),
    # so the invariants hold for the integrated part  RRn = RR - (Y/sum Y - Y) rho/dt
    sc = b["rho"] / 1e-6
    RRn = RR - (b["Y"] / b["Y"].sum(0, keepdims=True) - b["Y"]) * sc[None]
    floor = 256 * np.finfo(float).eps * sc               # round-off of O(1) mass fractions seen through rho/dt
    # conservation holds to the linear-solve residual of the Rosenbrock stages, not to round-off
    assert (np.abs(RRn.sum(0)) <= 1e-8 * np.abs(RRn).sum(0) + floor).all()
    comp, W = comp_matrix(name)
    el = (comp / W[:, None]).T @ RRn                      # [n_elem, n]
    assert (np.abs(el) <= 1e-8 * ((np.abs(comp) / W[:, None]).T @ np.abs(RRn)) + floor[None] / W.min()).all()
    assert np.allclose(Q, -(h.hc[:, None] * RR).sum(0), rtol=1e-9, atol=1e-9 * np.abs(Q).max())
    assert cost.min() >= 1 and cost.max() < 5000
    # nothing happens in cold fuel/air cells (no radicals: the unburnt half of the batch): one step
    cold = (b["T"] < 500.0) & ((b["Y"] > 0).sum(0) <= 4)
    assert cold.sum() > 0 and (cost[cold] <= 4).all()      # h0 = 10/||f0/sc||, growth x6: a handful of steps at most
    RR2, Q2 = h.solve(b["T"], b["p"], b["rho"], 1e-6, b["Y"])
    assert np.array_equal(RR, RR2) and np.array_equal(Q, Q2)
    # RR is linear in the caller's density (it only scales the result, dfChemistryModel.C:770)
    # (16384 cells: a batch of an H2-class mechanism below 12288 cells runs the other solve kernel, whose sums differ in the last bits)
    k = 16384
    RR3, _ = h.solve(b["T"][:k], b["p"][:k], 2.0 * b["rho"][:k], 1e-6, np.ascontiguousarray(b["Y"][:, :k]))
    assert np.array_equal(RR3, 2.0 * RR[:, :k])
    h.close()


def test_chunked_host_pipeline_equals_device_entry(api):
    """dfchem_solve (host pointers, 2^18-cell chunks on two streams) == dfchem_solve_dev on the
    same data held in device memory (torch tensors, torch's current stream)."""
    import torch

    dfchem_host, synth = api
    h = dfchem_host.Handle(synth.load_blob("es80"), 0)
    h.set_tolerances(1e-6, 1e-10)
    n = (1 << 18) * 2 + 12345                               # ragged last chunk
    b = synth.make_batch("es80", h.species, h.W, n, seed=3)
    RR, Q = h.solve(b["T"], b["p"], b["rho"], 1e-6, b["Y"])
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(v).to(dev) for k, v in b.items()}
    dRR = torch.empty((h.N, n), dtype=torch.float64, device=dev)
    dQ = torch.empty(n, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    h.check(h.lib.dfchem_solve_dev(h.h, n, t["T"].data_ptr(), t["p"].data_ptr(), t["rho"].data_ptr(), None, 1e-6,
                                   t["Y"].data_ptr(), n, dRR.data_ptr(), n, dQ.data_ptr(), None, None, stream))
    torch.cuda.synchronize()
    assert np.array_equal(dRR.cpu().numpy(), RR) and np.array_equal(dQ.cpu().numpy(), Q)
    assert h.counters()["failed"] == 0
    h.close()


def test_model_mirrors_reference_interface(api):
    """solve(deltaT) return convention (dfChemistryModel.C:386-408, 871, 894), chemistry off,
    correctThermo/updateEnergy round trip, calculateW, and the operator-split 0-D loop of
    df0DFoam (df0DFoam.C:106-111) igniting inside the published window."""
    dfchem_host, synth = api
    from dfchem_host import GREAT, DfChemistryModel

    n = 8
    chem = DfChemistryModel(synth.load_blob("es80"), n, rel_tol=1e-10, abs_tol=1e-24)
    sp = chem.species()
    Y = chem.Y()
    Y[sp.index("H2")] = 0.02555366261652
    Y[sp.index("O2")] = 0.227143668370244
    Y[sp.index("N2")] = 0.7473026689382
    chem.T[:] = 1000.0 + np.arange(n)          # eight slightly different reactors
    chem.p[:] = 101325.0
    chem.updateEnergy()
    chem.correctThermo()
    assert np.abs(chem.T - (1000.0 + np.arange(n))).max() < 1e-8
    dt = 1e-6
    assert chem.solve(dt) == 2 * dt             # min(great, 2*deltaT)
    assert chem.solve(np.full(n, dt)) == GREAT  # field form returns great
    chem.calculateW()
    assert np.abs(sum(chem.wrate(i) for i in range(chem.nSpecies()))).max() < 1e-12 * np.abs(chem.wrate(0)).max() + 1e-30
    off = DfChemistryModel(synth.load_blob("es80"), 2, chemistry=False)
    assert off.solve(dt) == 2 * dt and np.all(off.RR(0) == 0)
    # df0DFoam loop: YEqn (Y += RR dt / rho at frozen h, p) then correctThermo
    Tmax_rate, t_ign, Tprev = 0.0, 0.0, chem.T[0]
    for step in range(400):
        chem.solve(dt)
        for i in range(chem.nSpecies()):
            Y[i] += chem.RR(i) * dt / chem.rho
        chem.correctThermo()
        rate = (chem.T[0] - Tprev) / dt
        if rate > Tmax_rate:
            Tmax_rate, t_ign = rate, (step + 1) * dt
        Tprev = chem.T[0]
    assert 2.2e-4 < t_ign < 2.7e-4, t_ign      # docs/source/qs/0Dcvode.jpg; test/corrtest.cpp:117
    assert chem.T[0] > 2200.0                # oracle loop: 2247.7 K at 400 us (tests/golden/ignition_es80.npz)
    assert abs(Y.sum(0) - 1.0).max() < 1e-9
