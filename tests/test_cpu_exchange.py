"""N>1 host logic on CPU (gloo): plan -> cellsToSend -> forward exchange -> solve -> return -> order-preserving unpack.

Covers what the reference's own tests leave out (SURVEY section 4 "gaps": only its firstTime_ branch ever runs):
redistribution must not change results, messages return in the same order and size
(reference tests/src/pointToPointBuffer-Test.C:73-290 checks that for raw label lists), updateIter > 0.
The cell "solver" here is a deterministic stand-in (the CUDA library needs a GPU). The step logic under test is the
product's C++ driver, csrc/lbchem_lb_core.hpp -- the same source liblbchem.so runs on the device -- compiled for host
memory (liblbchem_lbhost.so) with gloo behind its transport callbacks (model.HostStepModel).
"""
import os
import socket
import traceback

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from loadbalancedchemistrymodel_b200 import capi
from loadbalancedchemistrymodel_b200.model import HostStepModel

NSP = 5


def fake_solver(Y, T, p, rho, dT, dtc):
    """Cell-local and deterministic: any redistribution must reproduce it bit for bit."""
    w = torch.arange(1, Y.shape[0] + 1, dtype=torch.float64).reshape(-1, 1)
    RR = Y * T * 1e-3 * w + rho * dT * 1e3
    Tout = T + p * 1e-5 + dtc * 1e6
    cost = 1.0 + torch.floor(T / 300.0)          # hot cells cost more; integers like the kernel's sub-step counts
    dtc.mul_(0.5)
    return RR, Tout, cost, torch.zeros(Y.shape[1], dtype=torch.int32)


def make_state(rank, n, seed=0):
    g = torch.Generator().manual_seed(1234 + 17 * rank + seed)
    Y = torch.rand((NSP, n), generator=g, dtype=torch.float64)
    # rank 0 is hot (expensive), the others cold: a flame front sitting in one sub-domain
    T = (1500.0 + 900.0 * torch.rand(n, generator=g, dtype=torch.float64)) if rank == 0 else \
        (300.0 + 50.0 * torch.rand(n, generator=g, dtype=torch.float64))
    p = torch.full((n,), 1e5, dtype=torch.float64)
    rho = 1.0 + torch.rand(n, generator=g, dtype=torch.float64)
    dtc = torch.full((n,), 1e-7, dtype=torch.float64)
    return Y, T, p, rho, dtc


def _worker(rank, world, port, n, update_iter, tdac, lts, q):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        nloc = n + 13 * rank                      # ragged shard sizes
        Y, T, p, rho, dtc0 = make_state(rank, nloc)
        dT = (1e-6 * (1.0 + torch.arange(nloc, dtype=torch.float64) / nloc)) if lts else 1e-6
        model = HostStepModel(NSP, fake_solver, update_iter=update_iter, tdac=tdac)
        single = HostStepModel(NSP, fake_solver, balance=False, tdac=tdac)   # never exchanges: the non-load-balanced model
        out = []
        dtc_lb, dtc_ref = dtc0, dtc0
        for step in range(4):
            m1 = model.solve(Y, T, p, rho, dT, dtc_lb)
            m2 = single.solve(Y, T, p, rho, dT, dtc_ref)
            ok = (torch.equal(model.RR, single.RR) and torch.equal(model.T_out, single.T_out)
                  and torch.equal(model.deltaTChem, single.deltaTChem) and torch.equal(model.cell_cost, single.cell_cost)
                  and torch.equal(model.status, single.status) and m1 == m2)
            out.append(dict(step=step, ok=bool(ok), last={k: v for k, v in model.last.items()},
                            send=list(model.send_plan), recv=list(model.recv_plan), total=model.total_cost))
            if not tdac:                          # LoadBalancedTDAC never writes deltaTChem back (reference TDAC .C:968-971)
                dtc_lb, dtc_ref = model.deltaTChem, single.deltaTChem
            else:
                ok = ok and torch.equal(model.deltaTChem, dtc0)
                out[-1]["ok"] = bool(ok)
        q.put((rank, out, None))
        dist.barrier()
        dist.destroy_process_group()
    except Exception:
        q.put((rank, None, traceback.format_exc()))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _run(world, n=400, update_iter=0, tdac=False, lts=False):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, update_iter, tdac, lts, q)) for r in range(world)]
    for p_ in procs:
        p_.start()
    res = {}
    for _ in range(world):
        rank, out, err = q.get(timeout=180)
        assert err is None, err
        res[rank] = out
    for p_ in procs:
        p_.join(timeout=60)
    return res


@pytest.mark.parametrize("world,tdac,lts", [(2, False, False), (2, True, True), (4, False, True), (4, True, False)])
def test_redistribution_preserves_results_and_order(world, tdac, lts):
    res = _run(world, tdac=tdac, lts=lts)
    for rank, out in res.items():
        assert all(o["ok"] for o in out), (rank, out)
        assert out[0]["last"]["sent_cells"] == 0 and not out[0]["last"]["planned"]      # firstTime_: all local
    # rank 0 holds the expensive cells: from step 1 on it ships cells, and what it ships is what the others receive
    for step in (1, 2, 3):
        sent = sum(res[r][step]["last"]["sent_cells"] for r in res)
        recv = sum(res[r][step]["last"]["recv_cells"] for r in res)
        assert sent == recv and res[0][step]["last"]["sent_cells"] > 0
        assert res[0][step]["last"]["planned"]                                         # updateIter 0: re-plan every step
    # the plan evens out the load: cost solved by rank 0 after balancing is below its own total
    assert res[0][1]["send"] and all(not res[r][1]["send"] for r in res if r != 0)


def test_update_iter_replans_every_third_step():
    res = _run(2, update_iter=2)
    planned = [o["last"]["planned"] for o in res[0]]
    assert planned == [False, True, False, False]       # reference .C:515-519 with iter_ starting at maxIterUpdate_
    assert all(o["ok"] for o in res[0]) and all(o["ok"] for o in res[1])


def test_single_rank_without_process_group():
    """world 1: every step is the local branch, deltaTChem is carried, the total cost is the sum of the cell costs."""
    Y, T, p, rho, dtc0 = make_state(0, 97)
    m = HostStepModel(NSP, fake_solver)
    m1 = m.solve(Y, T, p, rho, 1e-6, dtc0)
    assert m.last["sent_cells"] == 0 and not m.last["exchanged"]
    assert m.total_cost == float(m.cell_cost.sum()) and m1 == float(m.deltaTChem.min())
    m.solve(Y, T, p, rho, 1e-6)
    assert torch.equal(m.deltaTChem, dtc0 * 0.25)
    e = m.solve(Y[:, :0], T[:0], p[:0], rho[:0], 1e-6, dtc0[:0])       # a rank without cells (reference .C:455: GREAT)
    assert e == 1e15 and m.RR.shape == (NSP, 0)
