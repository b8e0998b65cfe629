"""N>1 host logic on CPU (gloo): plan -> cellsToSend -> forward exchange -> solve -> return -> order-preserving unpack.

Covers what the reference's own tests leave out (SURVEY section 4 "gaps": only its firstTime_ branch ever runs):
redistribution must not change results, messages return in the same order and size
(reference tests/src/pointToPointBuffer-Test.C:73-290 checks that for raw label lists), updateIter > 0.
The cell "solver" here is a deterministic stand-in (the CUDA library needs a GPU); the driver under test is
the product's loadbalancedchemistrymodel_b200/model.py.
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
from loadbalancedchemistrymodel_b200.model import LoadBalancedChemistryModel, cells_to_send_device

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


def _worker(rank, world, port, n, update_iter, host_c2s, lts, q):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        nloc = n + 13 * rank                      # ragged shard sizes
        Y, T, p, rho, dtc0 = make_state(rank, nloc)
        dT = (1e-6 * (1.0 + torch.arange(nloc, dtype=torch.float64) / nloc)) if lts else 1e-6
        model = LoadBalancedChemistryModel(NSP, fake_solver, "cpu", update_iter=update_iter, host_cells_to_send=host_c2s)
        single = LoadBalancedChemistryModel(NSP, fake_solver, "cpu")
        single.world = 1                          # never exchanges: the reference's non-load-balanced model
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
            dtc_lb, dtc_ref = model.deltaTChem, single.deltaTChem
        q.put((rank, out, None))
        dist.barrier()
        dist.destroy_process_group()
    except Exception:
        q.put((rank, None, traceback.format_exc()))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _run(world, n=400, update_iter=0, host_c2s=False, lts=False):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, update_iter, host_c2s, lts, q)) for r in range(world)]
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


@pytest.mark.parametrize("world,host_c2s,lts", [(2, False, False), (2, True, True), (4, False, True)])
def test_redistribution_preserves_results_and_order(world, host_c2s, lts):
    res = _run(world, host_c2s=host_c2s, lts=lts)
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


def test_cells_to_send_device_scan_matches_sequential_loop():
    rng = np.random.default_rng(5)
    for _ in range(200):
        n = int(rng.integers(0, 60))
        cost = rng.integers(0, 30, n).astype(np.float64)
        excl = torch.zeros(n + 1, dtype=torch.float64)
        if n:
            excl[1:] = torch.cumsum(torch.from_numpy(cost), 0)
        start = int(rng.integers(0, n + 2))
        target = float(rng.integers(0, 200)) * float(rng.choice([1.0, 0.37]))
        want = capi.cells_to_send(cost, target, min(start, n))
        assert cells_to_send_device(excl, target, min(start, n)) == want
