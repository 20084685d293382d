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
    """dfchem_balance_counts = timesToProblemCounts (LoadBalancer.C:97-122): leading cells in index order until the
    summed cost reaches the share; at least one cell stays."""
    from dfchem_host import balance

    cost = np.array([9.0, 7.0, 5.0, 3.0, 1.0, 1.0])
    # two ranks, rank 0 heavy: loads (26, 6) -> mean 16 -> rank 0 gives 10 units: cell 0 (9 <= 10); cell 1 would
    # bring the sum to 16 > 10 and stays (count_while(sum <= time) of the reference)
    assert balance.counts_row(np.array([26.0, 6.0]), 0, cost).tolist() == [0, 1]
    assert balance.counts_row(np.array([42.0, 6.0]), 0, cost).tolist() == [0, 2]              # 18 units: 9 + 7 <= 18 < 21
    assert balance.counts_row(np.array([26.0, 6.0]), 1, np.ones(6)).tolist() == [0, 0]       # the receiver sends nothing
    # three ranks (26, 16, 3): mean 15 -> rank 0 gives 11 to rank 2 (lightest), rank 1 gives its 1 unit to rank 2 as well
    assert balance.counts_row(np.array([26.0, 16.0, 3.0]), 0, cost).tolist() == [0, 0, 1]
    # never send everything
    assert balance.counts_row(np.array([1e9, 0.0]), 0, cost).sum() == 5
    # a single cell cannot be given away
    assert balance.counts_row(np.array([10.0, 0.0]), 0, np.array([10.0])).sum() == 0


def _reference_pairing(loads, head_tail):
    """LoadBalancer::getOperations / getOperationsRedezVous restated literally (LoadBalancer.C:124-232)."""
    n = len(loads)
    mean = float(np.mean(loads))
    order = sorted(range(n), key=lambda i: loads[i])
    v = [float(loads[i]) for i in order]
    send = np.zeros((n, n))
    if head_tail:
        for i in range(n // 2):
            hi, lo = n - 1 - i, i
            x = (v[hi] - v[lo]) / 2
            if x > 0.01 * mean:
                send[order[hi], order[lo]] += x
            v[hi] -= x
            v[lo] += x
        return send
    hi, lo = n - 1, 0
    while hi != lo:
        x = min(v[hi] - mean, mean - v[lo])
        if x > 0.01 * mean:
            send[order[hi], order[lo]] += x
        v[hi] -= x
        v[lo] += x
        if abs(v[hi] - mean) < 1e-15 * max(mean, 1.0) or v[hi] - mean <= 1e-12 * mean:
            hi -= 1
        else:
            lo += 1
    return send


def test_plan_matches_the_reference_pairing_rules(built):
    from dfchem_host import balance

    rng = np.random.default_rng(3)
    for n in (2, 3, 4, 8):
        for _ in range(20):
            loads = rng.uniform(1.0, 100.0, n)
            for alg in (balance.ALL_AVERAGE, balance.HEAD_TAIL):
                send = balance.plan_matrix(loads, alg)
                assert np.allclose(send, _reference_pairing(loads, alg == balance.HEAD_TAIL), rtol=1e-12, atol=1e-9), (n, alg, loads)
                assert (np.diag(send) == 0).all() and (send >= 0).all()
                # a rank is sender or receiver, never both (runtime_assert in the reference)
                assert not ((send.sum(1) > 0) & (send.sum(0) > 0)).any()


def test_plan_moves_load_towards_the_mean(built):
    from dfchem_host import balance

    loads = np.array([100.0, 20.0, 60.0, 20.0])
    send = balance.plan_matrix(loads)
    after = loads - send.sum(1) + send.sum(0)
    assert np.allclose(after, loads.mean())
    ht = balance.plan_matrix(loads, balance.HEAD_TAIL)
    after_ht = loads - ht.sum(1) + ht.sum(0)
    assert after_ht.max() - after_ht.min() < loads.max() - loads.min() and np.isclose(after_ht.sum(), loads.sum())
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
        load_after = float(cost.sum()) - float(cost[:plan.n_send].sum())
        got = torch.tensor([load_after], dtype=torch.float64)
        extra = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
        sent_load = torch.tensor([float(cost[:plan.n_send].sum())], dtype=torch.float64)
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
