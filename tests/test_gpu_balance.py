"""In-library load balancer (dfchem_solve_balanced_dev, csrc/rebalance.cuh) on real GPUs over NCCL:
results must come back bit-identical to the unbalanced call.  The two-rank test needs two GPUs
(gpurun --gpus 2; profiles/r2 keeps its log); with one GPU only the single-rank path runs."""
import ctypes
import os
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _solve(h, lib, torch, b, dev, balanced_with=None, alg=0):
    n, N = b["T"].shape[0], h.N
    d = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in b.items()}
    RR = torch.full((N, n), float("nan"), dtype=torch.float64, device=dev)
    Q, cost = (torch.full((n,), float("nan"), dtype=torch.float64, device=dev) for _ in range(2))
    st = torch.full((n,), -1, dtype=torch.int32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    if balanced_with is None:
        h.check(lib.dfchem_solve_dev(h.h, n, d["T"].data_ptr(), d["p"].data_ptr(), d["rho"].data_ptr(), None, 1e-6,
                                     d["Y"].data_ptr(), n, RR.data_ptr(), n, Q.data_ptr(), cost.data_ptr(), st.data_ptr(), s))
    else:
        h.check(lib.dfchem_solve_balanced_dev(h.h, n, d["T"].data_ptr(), d["p"].data_ptr(), d["rho"].data_ptr(), None, 1e-6,
                                              d["Y"].data_ptr(), n, RR.data_ptr(), n, Q.data_ptr(), balanced_with.data_ptr(),
                                              cost.data_ptr(), st.data_ptr(), alg, s))
    torch.cuda.synchronize()
    return RR, Q, cost, st


def test_single_rank_communicator_is_the_plain_step(built):
    import torch

    import dfchem_host
    from dfchem_host import synth

    lib = dfchem_host.load_library()
    h = dfchem_host.Handle(synth.load_blob("drm19"), 0)
    h.set_tolerances(1e-6, 1e-10)
    idb = ctypes.create_string_buffer(128)
    h.check(lib.dfchem_comm_unique_id(ctypes.cast(idb, ctypes.c_void_p)))
    h.check(lib.dfchem_comm_init(h.h, 1, 0, ctypes.cast(idb, ctypes.c_void_p)))
    dev = torch.device("cuda", 0)
    b = synth.make_batch("drm19", h.species, h.W, 20000, seed=5)
    RR0, Q0, c0, s0 = _solve(h, lib, torch, b, dev)
    RR1, Q1, c1, s1 = _solve(h, lib, torch, b, dev, balanced_with=c0)
    assert torch.equal(RR0, RR1) and torch.equal(Q0, Q1) and torch.equal(c0, c1) and torch.equal(s0, s1)
    out = (ctypes.c_double * 8)()
    h.check(lib.dfchem_balance_stats(h.h, out))
    assert out[0] == 0 and out[1] == 0
    h.close()


def _rank_main(rank, world, idfile, outdir, alg):
    import torch

    import dfchem_host
    from dfchem_host import synth

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    lib = dfchem_host.load_library()
    h = dfchem_host.Handle(synth.load_blob("drm19"), rank)
    h.set_tolerances(1e-6, 1e-10)
    idb = ctypes.create_string_buffer(128)
    if rank == 0:
        h.check(lib.dfchem_comm_unique_id(ctypes.cast(idb, ctypes.c_void_p)))
        with open(idfile + ".tmp", "wb") as f:
            f.write(idb.raw)
        os.replace(idfile + ".tmp", idfile)
    else:
        for _ in range(600):
            if os.path.exists(idfile):
                break
            time.sleep(0.1)
        idb = ctypes.create_string_buffer(open(idfile, "rb").read(), 128)
    h.check(lib.dfchem_comm_init(h.h, world, rank, ctypes.cast(idb, ctypes.c_void_p)))
    # rank 0 holds the hot (expensive) cells, rank 1 cold unburnt ones; different sizes on purpose
    n = 60000 if rank == 0 else 45000
    T_range = (1800.0, 2500.0) if rank == 0 else (300.0, 700.0)
    b = synth.make_batch("drm19", h.species, h.W, n, seed=11 + rank, T_range=T_range)
    RR0, Q0, c0, s0 = _solve(h, lib, torch, b, dev)
    RR1, Q1, c1, s1 = _solve(h, lib, torch, b, dev, balanced_with=c0, alg=alg)
    st = (ctypes.c_double * 8)()
    h.check(lib.dfchem_balance_stats(h.h, st))
    same = bool(torch.equal(RR0, RR1) and torch.equal(Q0, Q1) and torch.equal(c0, c1) and torch.equal(s0, s1))
    np.save(os.path.join(outdir, f"r{rank}.npy"), np.array([float(same), st[0], st[1], st[2], float(c0.sum().item()), st[7]]))
    h.close()


@pytest.mark.parametrize("alg", [0, 1])
def test_two_rank_nccl_rebalance_is_bit_identical(built, tmp_path, alg):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp

    world = 2
    mp.spawn(_rank_main, args=(world, str(tmp_path / "nccl_id"), str(tmp_path), alg), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "r0.npy"), np.load(tmp_path / "r1.npy")
    assert r0[0] == 1.0 and r1[0] == 1.0                  # RR, Qdot, cost, status bit-identical to the unbalanced call
    assert r0[1] > 0 and r0[2] == 0                       # the hot rank sent cells and received none
    assert r1[2] == r0[1] and r1[1] == 0                  # ... and the cold rank received exactly those
    assert r0[4] > 1.5 * r1[4]                            # the loads were unequal to begin with
    print("two-rank rebalance: cells moved", int(r0[1]), "bytes", int(r0[3] + r1[3]), "loads", r0[4], r1[4], "ms", r0[5], r1[5])
