"""Cost-weighted cell migration between the GPUs of one box.

The product path is INSIDE the library: `dfchem_solve_balanced_dev` (csrc/rebalance.cuh) plans on the device and
moves cells with grouped ncclSend/ncclRecv; `init_library_comm` below only hands the NCCL id around.  This module also
keeps a small mirror of the same protocol over `torch.distributed`, driven by the library's own host planner
(`dfchem_balance_plan2`, `dfchem_balance_counts`: the code the device planner runs), so that the N > 1 host logic --
who sends how many cells to whom, that results return to their owner -- is testable at world_size 2 on gloo without a GPU.

Reference (read-only), src/dfChemistryModel/loadBalancing/:
  LoadBalancerBase::updateState   LoadBalancer.C:29-56     all-gather of per-rank loads (sum of the
                                                            previous step's per-cell cpuTime)
  LoadBalancer::getOperations     LoadBalancer.C:124-180   allAverage pairing   (algorithm 0)
  getOperationsRedezVous          LoadBalancer.C:183-232   headTail pairing     (algorithm 1)
  timesToProblemCounts            LoadBalancer.C:97-122    load units -> number of LEADING cells (index order)
  LoadBalancerBase::balance       LoadBalancerBase.H:143-189  send the chosen cells' ChemistryProblems
  getRemaining                    LoadBalancerBase.H:205-225  the cells that stay
  LoadBalancerBase::unbalance     LoadBalancerBase.H:227-255  solutions return to the owner
Cost is the integrator's step-attempt count of the previous flow step instead of wall seconds.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist

from . import load_library

ALL_AVERAGE, HEAD_TAIL = 0, 1
_dp = ctypes.POINTER(ctypes.c_double)


def plan_matrix(loads: np.ndarray, algorithm: int = ALL_AVERAGE) -> np.ndarray:
    """send[from, to] in load units (dfchem_balance_plan2)."""
    lib = load_library()
    loads = np.ascontiguousarray(loads, dtype=np.float64)
    n = loads.shape[0]
    send = np.zeros((n, n))
    rc = lib.dfchem_balance_plan2(loads.ctypes.data_as(_dp), n, int(algorithm), send.ctypes.data_as(_dp))
    if rc != 0:
        raise RuntimeError(f"dfchem_balance_plan2 failed: {rc}")
    return send


def counts_row(loads: np.ndarray, me: int, cost: np.ndarray, algorithm: int = ALL_AVERAGE) -> np.ndarray:
    """Cells rank `me` sends to every rank: leading cells in index order until their summed cost reaches the share
    (dfchem_balance_counts = timesToProblemCounts; at least one cell stays)."""
    lib = load_library()
    loads = np.ascontiguousarray(loads, dtype=np.float64)
    cost = np.ascontiguousarray(cost, dtype=np.float64)
    out = np.zeros(loads.shape[0], dtype=np.int32)
    rc = lib.dfchem_balance_counts(loads.ctypes.data_as(_dp), loads.shape[0], int(me), int(algorithm),
                                   cost.ctypes.data_as(_dp), cost.shape[0], out.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
    if rc != 0:
        raise RuntimeError(f"dfchem_balance_counts failed: {rc}")
    return out


def init_library_comm(handle, device=None):
    """Create the library's NCCL communicator across the ranks of the default torch.distributed group: rank 0 draws
    the id (dfchem_comm_unique_id), torch broadcasts its 128 bytes, every rank calls dfchem_comm_init."""
    lib = load_library()
    world, rank = dist.get_world_size(), dist.get_rank()
    buf = (ctypes.c_ubyte * 128)()
    if rank == 0:
        handle.check(lib.dfchem_comm_unique_id(ctypes.cast(buf, ctypes.c_void_p)))
    t = torch.tensor(list(buf), dtype=torch.uint8)
    if device is not None:
        t = t.to(device)
    dist.broadcast(t, 0)
    raw = bytes(t.cpu().tolist())
    idb = ctypes.create_string_buffer(raw, 128)
    handle.check(lib.dfchem_comm_init(handle.h, world, rank, ctypes.cast(idb, ctypes.c_void_p)))


@dataclass
class Plan:
    send_counts: list          # cells this rank sends to each rank (leading cells, consecutive ranges)
    recv_counts: list          # guest cells this rank receives from each rank
    n_send: int
    n_recv: int
    loads: np.ndarray          # all ranks' loads (the all-gather of updateState)


class Rebalancer:
    """torch.distributed mirror of the in-library protocol (CPU tests)."""

    def __init__(self, group=None, algorithm: int = ALL_AVERAGE):
        self.group = group
        self.algorithm = algorithm
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def make_plan(self, cost: torch.Tensor) -> Plan:
        load = cost.sum(dtype=torch.float64).reshape(1)
        allv = [torch.zeros_like(load) for _ in range(self.world)]
        if self.world > 1:
            dist.all_gather(allv, load, group=self.group)
        else:
            allv = [load]
        loads = torch.cat(allv).cpu().numpy()
        row = torch.from_numpy(counts_row(loads, self.rank, cost.double().cpu().numpy(), self.algorithm).astype(np.int64))
        rows = [torch.zeros_like(row) for _ in range(self.world)]
        if self.world > 1:
            dist.all_gather(rows, row, group=self.group)      # the n x n count matrix, as in the library
        else:
            rows = [row]
        counts = torch.stack(rows).numpy()
        send = [int(x) for x in counts[self.rank]]
        recv = [int(x) for x in counts[:, self.rank]]
        return Plan(send, recv, sum(send), sum(recv), loads)

    def _a2a(self, src: torch.Tensor, out_cells: int, in_splits, out_splits) -> torch.Tensor:
        out = torch.empty((out_cells, src.shape[1]), dtype=src.dtype, device=src.device)
        if self.world > 1:
            dist.all_to_all_single(out, src.contiguous(), output_split_sizes=out_splits,
                                   input_split_sizes=in_splits, group=self.group)
        return out

    def balance(self, plan: Plan, fields: torch.Tensor):
        """fields [K, nC] -> (own [K, nC - n_send] = the cells that stay (getRemaining), guests [K, n_recv])."""
        sent = fields[:, :plan.n_send].t().contiguous()
        guests = self._a2a(sent, plan.n_recv, plan.send_counts, plan.recv_counts)
        return fields[:, plan.n_send:], guests.t().contiguous()

    def unbalance(self, plan: Plan, own_res: torch.Tensor, guest_res: torch.Tensor, out: torch.Tensor):
        back = self._a2a(guest_res.t().contiguous(), plan.n_send, plan.recv_counts, plan.send_counts)
        out[:, plan.n_send:] = own_res
        if plan.n_send:
            out[:, :plan.n_send] = back.t()
        return out


def solve_balanced(rb: Rebalancer, solve_fn, fields: torch.Tensor, prev_cost: torch.Tensor, n_out: int):
    """One balanced sub-step with a caller-supplied solver: guests first, then the residents (the library's order)."""
    plan = rb.make_plan(prev_cost)
    own, guests = rb.balance(plan, fields)
    guest_res = solve_fn(guests) if plan.n_recv else torch.empty((n_out, 0), dtype=fields.dtype)
    own_res = solve_fn(own.contiguous())
    out = torch.empty((n_out, fields.shape[1]), dtype=own_res.dtype, device=own_res.device)
    rb.unbalance(plan, own_res, guest_res, out)
    return out, plan
