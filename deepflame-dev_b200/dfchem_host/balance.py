"""Cost-weighted cell migration between the GPUs of one box: the reference's chemistry load
balancer restated as an all-to-all-v over `torch.distributed` (NCCL/NVLink on the GPU box, gloo in
the CPU tests).

Reference (read-only), src/dfChemistryModel/loadBalancing/:
  LoadBalancerBase::updateState   LoadBalancer.C:29-56     all-gather of per-rank loads (sum of the
                                                            previous step's per-cell cpuTime)
  LoadBalancer::getOperations     LoadBalancer.C:124-180   allAverage pairing  -> dfchem_balance_plan
  timesToProblemCounts            LoadBalancer.C:97-122    load units -> number of leading cells
  LoadBalancerBase::balance       LoadBalancerBase.H:143-189  send the chosen cells' ChemistryProblems
  getRemaining                    LoadBalancerBase.H:205-225  the cells that stay
  LoadBalancerBase::unbalance     LoadBalancerBase.H:227-255  solutions return to the owner

Differences (B200-first): records are packed species-major `[K, n]` blocks (no AoS
ChemistryProblem serialisation); the migrated cells are the most expensive ones (cells are ordered
by decreasing cost before the prefix is cut; the reference cuts a prefix in mesh order); cost is the
integrator's step-attempt count of the previous flow step instead of wall seconds.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist

from . import load_library


def plan_matrix(loads: np.ndarray) -> np.ndarray:
    """send[from, to] in load units (dfchem_balance_plan = allAverage rule of getOperations)."""
    lib = load_library()
    loads = np.ascontiguousarray(loads, dtype=np.float64)
    n = loads.shape[0]
    send = np.zeros((n, n))
    dp = ctypes.POINTER(ctypes.c_double)
    rc = lib.dfchem_balance_plan(loads.ctypes.data_as(dp), n, send.ctypes.data_as(dp))
    if rc != 0:
        raise RuntimeError(f"dfchem_balance_plan failed: {rc}")
    return send


def counts_for_units(sorted_cost: np.ndarray, units: np.ndarray) -> np.ndarray:
    """timesToProblemCounts (LoadBalancer.C:97-122): walk the (sorted) cost list once; destination
    j receives cells until their summed cost reaches units[j].  At least one cell always stays."""
    counts = np.zeros(len(units), dtype=np.int64)
    if len(sorted_cost) == 0:
        return counts
    csum = np.concatenate([[0.0], np.cumsum(sorted_cost)])
    start = 0
    limit = len(sorted_cost) - 1
    for j, u in enumerate(units):
        if u <= 0.0 or start >= limit:
            continue
        end = int(np.searchsorted(csum, csum[start] + u, side="left"))
        end = min(max(end, start), limit)
        counts[j] = end - start
        start = end
    return counts


@dataclass
class Plan:
    order: torch.Tensor        # local cell indices by decreasing cost
    send_counts: list          # cells this rank sends to each rank
    recv_counts: list          # guest cells this rank receives from each rank
    n_send: int
    n_recv: int
    loads: np.ndarray          # all ranks' loads (the all-gather of updateState)


class Rebalancer:
    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def make_plan(self, cost: torch.Tensor) -> Plan:
        """cost: [nC] per-cell cost of the previous step on this rank (any device)."""
        dev = cost.device
        load = cost.sum(dtype=torch.float64).reshape(1)
        if self.world > 1:
            allv = [torch.zeros_like(load) for _ in range(self.world)]
            dist.all_gather(allv, load, group=self.group)
            loads = torch.cat(allv).cpu().numpy()
        else:
            loads = load.cpu().numpy()
        send = plan_matrix(loads)
        order = torch.argsort(cost, descending=True, stable=True)
        units = send[self.rank]
        if units.any():
            sorted_cost = cost[order].double().cpu().numpy()
            counts = counts_for_units(sorted_cost, units)
        else:
            counts = np.zeros(self.world, dtype=np.int64)
        sc = torch.tensor(counts, dtype=torch.int64, device=dev)
        rc = torch.zeros_like(sc)
        if self.world > 1:
            dist.all_to_all_single(rc, sc, group=self.group)
        recv = [int(x) for x in rc.cpu().tolist()]
        return Plan(order, [int(x) for x in counts], recv, int(counts.sum()), int(sum(recv)), loads)

    def _a2a(self, src: torch.Tensor, out_cells: int, in_splits, out_splits) -> torch.Tensor:
        """src: [n_in, K] cell-major rows -> [out_cells, K]."""
        out = torch.empty((out_cells, src.shape[1]), dtype=src.dtype, device=src.device)
        if self.world > 1:
            dist.all_to_all_single(out, src.contiguous(), output_split_sizes=out_splits,
                                   input_split_sizes=in_splits, group=self.group)
        return out

    def balance(self, plan: Plan, fields: torch.Tensor):
        """fields: [K, nC] species-major block (T, p, rho, dt, Y...).  Returns (own [K, n_own],
        guests [K, n_recv]); own = the cells that stay (getRemaining), in `plan.order[n_send:]` order."""
        sent = fields[:, plan.order[:plan.n_send]].t().contiguous()
        guests = self._a2a(sent, plan.n_recv, plan.send_counts, plan.recv_counts)
        own = fields[:, plan.order[plan.n_send:]]
        return own, guests.t().contiguous()

    def unbalance(self, plan: Plan, own_res: torch.Tensor, guest_res: torch.Tensor, out: torch.Tensor):
        """own_res [K2, n_own], guest_res [K2, n_recv] -> out [K2, nC] in the caller's cell order."""
        back = self._a2a(guest_res.t().contiguous(), plan.n_send, plan.recv_counts, plan.send_counts)
        out[:, plan.order[plan.n_send:]] = own_res
        if plan.n_send:
            out[:, plan.order[:plan.n_send]] = back.t()
        return out


def solve_balanced(rb: Rebalancer, solve_fn, fields: torch.Tensor, prev_cost: torch.Tensor, n_out: int):
    """One balanced chemistry sub-step.  solve_fn(block [K, n]) -> [n_out, n] (RR rows, Qdot, cost...).
    Guests are solved in the same launch as the resident cells (one concatenated block), so the
    exchange is the only extra work; results return to the owning rank in its own cell order."""
    plan = rb.make_plan(prev_cost)
    own, guests = rb.balance(plan, fields)
    n_own = own.shape[1]
    block = torch.cat([own, guests], dim=1) if plan.n_recv else own.contiguous()
    res = solve_fn(block)
    out = torch.empty((n_out, fields.shape[1]), dtype=res.dtype, device=res.device)
    rb.unbalance(plan, res[:, :n_own], res[:, n_own:], out)
    return out, plan
