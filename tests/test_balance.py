"""Cost-weighted rebalancing (reference: loadBalancing/LoadBalancer.C:29-232,
LoadBalancerBase.H:143-255) on CPU: plan arithmetic, and the all-to-all-v exchange at world_size 2 on
gloo with a stand-in solver (the exchange is independent of what solves the cells)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_counts_follow_times_to_problem_counts(built):
    from dfchem_host import balance

    cost = np.array([9.0, 7.0, 5.0, 3.0, 1.0, 1.0])           # already sorted, expensive first
    # 10 units to rank A, 5 to rank B: prefix sums 9, 16 -> two cells reach >= 10; next 5 -> one cell
    assert balance.counts_for_units(cost, np.array([10.0, 0.0, 5.0])).tolist() == [2, 0, 1]
    # never send everything: at least one cell stays
    assert balance.counts_for_units(cost, np.array([1e9])).tolist() == [5]
    assert balance.counts_for_units(np.array([]), np.array([3.0])).tolist() == [0]


def test_plan_moves_load_towards_the_mean(built):
    from dfchem_host import balance

    loads = np.array([100.0, 20.0, 60.0, 20.0])
    send = balance.plan_matrix(loads)
    assert send.shape == (4, 4) and (np.diag(send) == 0).all() and (send >= 0).all()
    after = loads - send.sum(1) + send.sum(0)
    assert np.allclose(after, loads.mean())
    # operations below 1 % of the mean are dropped (LoadBalancer.C:170-178)
    assert (balance.plan_matrix(np.array([50.2, 49.8])) == 0).all()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _stand_in_solve(block):
    """[K, n] -> [3, n]: any per-cell function of the record will do (element-wise, so that the
    result does not depend on the memory layout of the block)."""
    return torch.stack([block[3] * block[4] + block[0], 2.0 * block[0] - block[1], block[2] ** 2])


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dfchem_host import balance

        g = torch.Generator().manual_seed(100 + rank)
        n = 300 + 50 * rank
        fields = torch.rand((5, n), generator=g, dtype=torch.float64)
        # rank 0 carries the expensive cells
        cost = torch.rand(n, generator=g, dtype=torch.float64) * (20.0 if rank == 0 else 1.0) + 1.0
        rb = balance.Rebalancer()
        out, plan = balance.solve_balanced(rb, _stand_in_solve, fields, cost, 3)
        ref = _stand_in_solve(fields)
        ok = bool(torch.equal(out, ref))
        moved = plan.n_send if rank == 0 else plan.n_recv
        load_after = float(cost.sum()) - float(cost[plan.order[:plan.n_send]].sum())
        got = torch.tensor([load_after], dtype=torch.float64)
        extra = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
        sent_load = torch.tensor([float(cost[plan.order[:plan.n_send]].sum())], dtype=torch.float64)
        dist.all_gather(extra, sent_load)
        np.save(os.path.join(tmp, f"r{rank}.npy"),
                np.array([float(ok), float(moved), float(got), float(plan.loads[0]), float(plan.loads[1]),
                          float(extra[0]), float(plan.n_send), float(plan.n_recv)]))
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_returns_results_to_the_owner(built, tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0 = np.load(tmp_path / "r0.npy")
    r1 = np.load(tmp_path / "r1.npy")
    assert r0[0] == 1.0 and r1[0] == 1.0                     # every rank got exactly its own cells' results back
    assert r0[1] > 0 and r1[1] == r0[1]                      # rank 0 sent cells, rank 1 received the same number
    assert r1[6] == 0 and r0[7] == 0                         # one-directional: the light rank sends nothing
    l0, l1, sent = r0[3], r0[4], r0[5]
    mean = 0.5 * (l0 + l1)
    assert l0 > l1
    # after the exchange both ranks sit within one (expensive) cell of the mean
    assert abs((l0 - sent) - mean) <= 21.0 and abs((l1 + sent) - mean) <= 21.0
    assert (l0 - sent) < l0 and (l1 + sent) > l1
