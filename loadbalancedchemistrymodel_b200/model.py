"""Host driver mirroring ``LoadBalancedChemistryModel::solve<DeltaTType>`` for one rank (= one GPU).

Reference: src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C
  :453-509  first step: everything is solved locally (``firstTime_``)
  :512-519  later steps: refresh the cell list, re-plan every ``updateIter + 1`` steps
  :521-570  ship the FIRST cells of the list to each planned receiver (``cellsToSend``), keep the tail
  :572-611  receive, solve own tail first, then the received cells
  :613-673  return results to their owners in the same order; :676-701 unpack, sum the cost
and src/pointToPointBuffer/pointToPointBuffer.C:53-168 (p2p exchange with swapped sizes on the way back).

What is different on the B200: cells are columns of structure-of-arrays device tensors (the reference's
``baseDataContainer`` records, src/dataContainer/baseDataContainer/baseDataContainer.H:85-117, become rows of a
packed ``[nSp+5, count]`` message), the per-cell cost is the sub-step count the kernel returns (not a wall
clock), ``cellsToSend`` is an exact device prefix scan, and the transport is ``torch.distributed`` point to
point (NCCL over NVLink; gloo in the CPU tests). The arithmetic is behind ``cell_solver`` — in production the
CUDA library (``capi.Chemistry``); there is no CPU fallback here.
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np
import torch
import torch.distributed as dist

from . import capi

GREAT = 1e15
N_FWD_EXTRA = 5   # T, p, rho, deltaT, deltaTChem after the nSp mass fractions
N_RET_EXTRA = 4   # deltaTChem, T, cost, status after the nSp reaction rates


class CudaCellSolver:
    """``solveCellList`` on the device through the C ABI (lbchem_solve_batch)."""

    def __init__(self, chem: "capi.Chemistry"):
        self.chem = chem
        self.launches = 0

    def __call__(self, Y, T, p, rho, deltaT, deltaTChem):
        nSp, n = Y.shape
        RR = torch.empty_like(Y)
        Tout = torch.empty_like(T)
        cost = torch.empty_like(T)
        status = torch.empty(n, dtype=torch.int32, device=Y.device)
        if n:
            assert Y.is_cuda and Y.is_contiguous(), "the CUDA cell solver needs contiguous device tensors"
            stream = torch.cuda.current_stream(Y.device).cuda_stream
            self.chem.solve_device(n, n, Y, T, p, rho, deltaT, deltaTChem, RR, Tout, cost, status, stream=stream)
            self.launches += 1
        return RR, Tout, cost, status


def cells_to_send_device(excl_cost: torch.Tensor, target: float, start: int) -> int:
    """``cellsToSend`` (reference .C:143-167) on an exclusive prefix sum ``excl_cost[i] = sum(cost[:i])``.

    Exact (bit-identical to the sequential loop) when the costs are integers stored as doubles, which is what
    the kernel writes; for arbitrary floating-point costs use ``capi.cells_to_send`` on the host.
    """
    n = excl_cost.shape[0] - 1
    if start >= n:
        return n
    goal = excl_cost[start] + target
    i = int(torch.searchsorted(excl_cost[:n], goal, right=False).item())
    return max(i, start)


class LoadBalancedChemistryModel:
    """One rank of the load-balanced chemistry model. Fields live on ``device`` as float64 tensors."""

    def __init__(self, n_species: int, cell_solver: Callable, device, group=None, update_iter: int = 0,
                 min_fraction: float = 0.02, deltaTChemMax: float = GREAT, host_cells_to_send: bool = False):
        self.nSp = int(n_species)
        self.cell_solver = cell_solver
        self.device = torch.device(device)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.max_iter_update = int(update_iter)          # LoadBalancedCoeffs{updateIter}, reference .C:41-45
        self.min_fraction = float(min_fraction)          # LoadBalancedCoeffs{minFractionOfCellsToSend}
        self.deltaTChemMax = float(deltaTChemMax)
        self.host_cells_to_send = host_cells_to_send
        self.first_time = True                           # reference .H:104
        self.iter = self.max_iter_update                 # reference .C:58: plan on the first balanced step
        self.total_cost = 0.0                            # totalCpuTime_
        self.cell_cost: Optional[torch.Tensor] = None    # cellDataList_[i].cpuTime() of the last step
        self.send_plan: list = []                        # [(toProc, fraction)]  sendAndReceiveData_.first()
        self.recv_plan: list = []                        # [fromProc]            sendAndReceiveData_.second()
        self.RR = self.deltaTChem = self.T_out = self.status = None
        self.cells_on_ranks = None
        self.last = {}                                   # counters of the last step (sent/received cells, bytes)

    # ---- collectives (the reference's gatherList/scatterList pairs) -------------------------------------
    def _all_gather(self, vec: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return vec.reshape(1, -1)
        flat = vec.contiguous().reshape(-1)
        out = torch.empty(self.world * flat.numel(), dtype=vec.dtype, device=vec.device)
        dist.all_gather_into_tensor(out, flat, group=self.group)
        return out.reshape((self.world,) + tuple(vec.shape))

    def _update_processor_balancing(self, n_local: int):
        """getSortedCPUTimesOnProcessor + updateProcessorBalancing (reference .C:170-347): every rank derives the
        whole plan from the all-gathered cost vector with the bit-exact C function."""
        mine = torch.tensor([self.total_cost, float(n_local)], dtype=torch.float64, device=self.device)
        g = self._all_gather(mine).cpu().numpy()
        cost = np.ascontiguousarray(g[:, 0])
        cells = np.ascontiguousarray(g[:, 1]).astype(np.int32)
        send, recv = capi.balance_plan(cost, cells, self.min_fraction)
        self.send_plan, self.recv_plan = send[self.rank], recv[self.rank]
        self.all_send_plan = send
        self.gathered_cost = cost

    # ---- one flow step -------------------------------------------------------------------------------
    def _finish(self, n, RR, dtc, Tout, cost, status):
        """unpack + updateTotalCpuTime (reference .C:676-701, :350-362)."""
        self.RR, self.T_out, self.status = RR, Tout, status
        dt_min = float(dtc.min().item()) if n else GREAT
        self.deltaTChem = torch.clamp(dtc, max=self.deltaTChemMax)
        self.cell_cost = cost
        self.total_cost = float(cost.sum().item())
        return min(dt_min, GREAT)

    def solve(self, Y, T, p, rho, deltaT, deltaTChem=None):
        """Integrate every local cell over ``deltaT`` (scalar, or a per-cell tensor for LTS). Returns the minimum
        chemical time step (reference .C:453-704). Results: ``self.RR [nSp, n]``, ``self.deltaTChem [n]``,
        ``self.T_out [n]``, ``self.cell_cost [n]``, ``self.status [n]`` in the caller's cell order."""
        nSp, n = Y.shape
        assert nSp == self.nSp
        dev = Y.device
        if deltaTChem is None:
            deltaTChem = self.deltaTChem
        dT = deltaT if torch.is_tensor(deltaT) else torch.full((n,), float(deltaT), dtype=torch.float64, device=dev)
        self.last = {"sent_cells": 0, "recv_cells": 0, "sent_bytes": 0, "recv_bytes": 0, "planned": False}

        if self.first_time or self.world == 1:
            dtc = deltaTChem.clone()
            RR, Tout, cost, status = self.cell_solver(Y.contiguous(), T, p, rho, dT, dtc)
            self.first_time = False
            return self._finish(n, RR, dtc, Tout, cost, status)

        if self.iter >= self.max_iter_update:            # reference .C:515-519 (iter_++ >= maxIterUpdate_)
            self._update_processor_balancing(n)
            self.iter = 0
            self.last["planned"] = True
        else:
            self.iter += 1

        # -- which cells go where: cellsToSend on last step's per-cell cost (reference .C:528-562)
        bounds = [0]
        if self.send_plan:
            if self.host_cells_to_send or not self.cell_cost.is_cuda:
                cc = self.cell_cost.cpu().numpy()
                for _, frac in self.send_plan:
                    bounds.append(capi.cells_to_send(cc, self.total_cost * frac, bounds[-1]))
            else:
                excl = torch.zeros(n + 1, dtype=torch.float64, device=dev)
                torch.cumsum(self.cell_cost, 0, out=excl[1:])
                for _, frac in self.send_plan:
                    bounds.append(cells_to_send_device(excl, self.total_cost * frac, bounds[-1]))
        start = bounds[-1]

        # -- size exchange: the reference sends 8 bytes to each partner (pointToPointBuffer.C:194-267); here one
        #    all-gather of the per-destination counts gives every rank its receive sizes
        counts = torch.zeros(self.world, dtype=torch.int64, device=dev)
        for k, (to, _) in enumerate(self.send_plan):
            counts[to] = bounds[k + 1] - bounds[k]
        all_counts = self._all_gather(counts).cpu().numpy()          # [from, to]
        recv_sizes = [int(all_counts[src, self.rank]) for src in self.recv_plan]

        # -- forward trip: pack contiguous cell ranges into [nSp+5, count] messages
        fwd_out, ops = [], []
        for k, (to, _) in enumerate(self.send_plan):
            s, e = bounds[k], bounds[k + 1]
            msg = torch.empty((nSp + N_FWD_EXTRA, e - s), dtype=torch.float64, device=dev)
            msg[:nSp] = Y[:, s:e]
            msg[nSp], msg[nSp + 1], msg[nSp + 2] = T[s:e], p[s:e], rho[s:e]
            msg[nSp + 3], msg[nSp + 4] = dT[s:e], deltaTChem[s:e]
            fwd_out.append(msg)
            if e > s:
                ops.append(dist.P2POp(dist.isend, msg, self._peer(to), group=self.group))
        fwd_in = [torch.empty((nSp + N_FWD_EXTRA, m), dtype=torch.float64, device=dev) for m in recv_sizes]
        for src, buf in zip(self.recv_plan, fwd_in):
            if buf.shape[1]:
                ops.append(dist.P2POp(dist.irecv, buf, self._peer(src), group=self.group))
        reqs = dist.batch_isend_irecv(ops) if ops else []

        # -- own tail first (reference .C:608), overlapping the transfer
        dtc_loc = deltaTChem[start:].clone()
        RRl, Tl, cl, sl = self.cell_solver(Y[:, start:].contiguous(), T[start:].contiguous(), p[start:].contiguous(),
                                           rho[start:].contiguous(), dT[start:].contiguous(), dtc_loc)
        for r in reqs:
            r.wait()

        # -- received cells, all sources in one batch (reference .C:611), then the return trip in the same order
        ops, ret_out = [], []
        if fwd_in:
            allin = torch.cat(fwd_in, dim=1) if len(fwd_in) > 1 else fwd_in[0]
            m = allin.shape[1]
            dtc_r = allin[nSp + 4].clone()
            RRr, Tr, cr, sr = self.cell_solver(allin[:nSp].contiguous(), allin[nSp].contiguous(), allin[nSp + 1].contiguous(),
                                               allin[nSp + 2].contiguous(), allin[nSp + 3].contiguous(), dtc_r)
            ret = torch.empty((nSp + N_RET_EXTRA, m), dtype=torch.float64, device=dev)
            ret[:nSp] = RRr
            ret[nSp], ret[nSp + 1], ret[nSp + 2], ret[nSp + 3] = dtc_r, Tr, cr, sr.to(torch.float64)
            off = 0
            for src, sz in zip(self.recv_plan, recv_sizes):
                part = ret[:, off:off + sz].contiguous()
                off += sz
                ret_out.append(part)
                if sz:
                    ops.append(dist.P2POp(dist.isend, part, self._peer(src), group=self.group))
        ret_in = []
        for k, (to, _) in enumerate(self.send_plan):
            buf = torch.empty((nSp + N_RET_EXTRA, bounds[k + 1] - bounds[k]), dtype=torch.float64, device=dev)
            ret_in.append(buf)
            if buf.shape[1]:
                ops.append(dist.P2POp(dist.irecv, buf, self._peer(to), group=self.group))
        for r in (dist.batch_isend_irecv(ops) if ops else []):
            r.wait()

        # -- unpack into the caller's cell order (reference .C:642-699)
        RR = torch.empty((nSp, n), dtype=torch.float64, device=dev)
        dtc = torch.empty(n, dtype=torch.float64, device=dev)
        Tout = torch.empty(n, dtype=torch.float64, device=dev)
        cost = torch.empty(n, dtype=torch.float64, device=dev)
        status = torch.empty(n, dtype=torch.int32, device=dev)
        for k, buf in enumerate(ret_in):
            s, e = bounds[k], bounds[k + 1]
            RR[:, s:e] = buf[:nSp]
            dtc[s:e], Tout[s:e], cost[s:e] = buf[nSp], buf[nSp + 1], buf[nSp + 2]
            status[s:e] = buf[nSp + 3].to(torch.int32)
        RR[:, start:], dtc[start:], Tout[start:], cost[start:], status[start:] = RRl, dtc_loc, Tl, cl, sl
        self.last.update(sent_cells=start, recv_cells=int(sum(recv_sizes)),
                         sent_bytes=8 * start * (nSp + N_FWD_EXTRA), recv_bytes=8 * int(sum(recv_sizes)) * (nSp + N_FWD_EXTRA),
                         bounds=bounds, recv_sizes=recv_sizes)
        return self._finish(n, RR, dtc, Tout, cost, status)

    def _peer(self, r: int) -> int:
        return dist.get_global_rank(self.group, r) if self.group is not None else r
