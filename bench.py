#!/usr/bin/env python
"""bench.py — cell-chemistry solves/s for one flow step (BASELINE.json metric).

Workload at every N: BASELINE config 3, "GRI-Mech 3.0 (53 species, 325 reactions), 1M synthetic cells,
uniform cost" per GPU (weak scaling: cells are independent, shards carry no data-path collective; the only
exchange is the reference's all-gather of one cost scalar per rank feeding the balancing plan).
A step = one LoadBalancedChemistryModel::solve(deltaT): every cell integrated over deltaT = 1e-6 s from
deltaTChem = 1e-7 s (reference tests/src/standardChemistryModel-Test.C:84).

  value  device-resident: inputs already in HBM, CUDA events around K steps, max over ranks
  e2e    through lbchem_solve_batch_host with pinned HOST buffers (H2D + integrate + D2H inside the timing)
  --impl reference : the CPU oracle (restatement of the reference's solveCell path; the reference itself
                     needs OpenFOAM and cannot be built here) on all host cores, bounded sample per step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

_emit = print

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "cell-chemistry solves/s (GRI-3.0, 1 flow step)"
CPU_SCALE = {"h2_air": 20, "ch4_mixing": 200}      # cells of the CPU sample relative to GRI: the small mechanisms are that much cheaper per cell
METRIC_OF = {"h2_air": "cell-chemistry solves/s (H2/O2, 1 flow step)", "ch4_mixing": "cell-chemistry solves/s (1-step CH4, 1 flow step)"}
UNIT = "cells/s"
SOLVERS = {"seulex": ("seulex", 1e-12, 1e-1), "Rosenbrock34": ("Rosenbrock34", 1e-12, 1e-2),
           "EulerImplicit": ("EulerImplicit", 1e-12, 1e-2)}     # EulerImplicitCoeffs{cTauChem 1; equilibriumRateLimiter off}
MECH_OF = {"h2_air": "H2/O2 sub-mechanism of GRI-3.0 (10 species, 29 reactions)", "ch4_mixing": "1-step CH4 (5 species, 1 reaction)"}


PIVOTING = 1      # lbchem_options.pivoting of every handle this run creates (--pivoting)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cells", type=int, default=1_000_000, help="cells per GPU")
    ap.add_argument("--solver", default="seulex", choices=list(SOLVERS))
    ap.add_argument("--workload", default="gri_uniform", choices=["gri_uniform", "gri_skewed", "gri_dac", "gri_tiled", "h2_air", "ch4_mixing"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="cells in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-imbalanced", action="store_true", help="skip the load-balancing record (north_star imbalanced batch)")
    ap.add_argument("--no-extra-configs", action="store_true", help="N = 8: skip the BASELINE config 4 / 5 records")
    ap.add_argument("--no-pageable", action="store_true")
    ap.add_argument("--imbalanced-cells", type=int, default=1_000_000, help="cells in TOTAL of the imbalanced batch")
    ap.add_argument("--imbalanced-steps", type=int, default=2)
    ap.add_argument("--dac", action="store_true", help="switch the DAC reduction on for any GRI workload (as --workload gri_dac does for the tiled fixture)")
    ap.add_argument("--pivoting", type=int, default=1, choices=[0, 1, 2],
                    help="lbchem_options.pivoting: 1 = diagonal pivots with the growth test and the reference's rule as fallback (default), 0 = the reference's rule on every matrix")
    ap.add_argument("--carry-dtchem", action="store_true",
                    help="carry deltaTChem from step to step like the model does between flow steps (default for gri_skewed)")
    return ap.parse_args()


def make_workload(name, n, rank, world=1):
    from loadbalancedchemistrymodel_b200 import workloads as W
    if name == "gri_skewed":
        return W.gri_skewed(n, rank=rank, world=world)
    if name in ("gri_dac", "gri_tiled"):
        return W.gri_tiled(n, rank=rank)
    if name == "h2_air":
        return W.h2_air(n, rank=rank)
    if name == "ch4_mixing":
        return W.ch4_mixing_line(n, rank=rank)
    return W.gri_uniform(n, rank=rank)


DAC_TOL, DAC_INIT = float(os.environ.get("LBCHEM_DAC_TOL", "1e-4")), ("CO", "CH4", "HO2")   # reference tests/Cases/Case-chemistry/constant/chemistryProperties:46,52-57


def dac_initial_set(mech):
    return [list(mech.species).index(s) for s in DAC_INIT]


WORKLOAD_TEXT = {
    "gri_uniform": "BASELINE config 3; fixture cells T>1500 K resampled + noise, seed 1003",
    "gri_skewed": "BASELINE config 4; hot cells (T>1500 K) contiguous at the start of the global index space, cold cells T<350 K, seed 1004",
    "gri_tiled": "BASELINE config 5 input without the reduction: the unfiltered 4000-cell fixture tiled + noise, seed 1005",
    "h2_air": "BASELINE config 2; H2/air, phi ~ U(0.5, 2), T ~ U(900, 2500) K, p = 1e5, radical seed, seed 1002",
    "ch4_mixing": "BASELINE config 1 input; CH4 / air mixing line, z ~ U(0, 1), T ~ U(293, 2000) K, seed 1001",
    "gri_dac": "BASELINE config 5; the unfiltered 4000-cell fixture tiled + noise, seed 1005; reduction{active on; method DAC; "
               "tolerance 1e-4; initialSet CO CH4 HO2}, tabulation off",
}


def config_dict(args, n_gpus, extra=None):
    s, atol, rtol = SOLVERS[args.solver]
    cfg = {
        "workload": f"{args.workload}: {MECH_OF.get(args.workload, 'GRI-Mech 3.0 (53 species, 325 reactions)')}, {args.cells} synthetic cells per GPU "
                    f"({WORKLOAD_TEXT[args.workload]})",
        "mechanism": MECH_OF.get(args.workload, "GRI-3.0"), "cells_per_gpu": args.cells, "solver": s, "absTol": atol, "relTol": rtol,
        "deltaT": 1e-6, "deltaTChem0": 1e-7, "parallelism": (f"cells sharded over {n_gpus} GPUs; every step through lbchem_lb_solve (C++ driver: all-gather of the cost, bit-exact plan, device cellsToSend, NCCL p2p of field-major row slices, one two-segment launch)" if n_gpus > 1 else "1 GPU, one persistent kernel per step"),
        "l2": "inputs (464 B/cell x cells) larger than L2; no flush needed",
        "pivoting": {1: "diagonal pivots, growth test, reference rule as fallback (lbchem_options.pivoting = 1, default)",
                     0: "the reference's partial pivoting on every matrix (lbchem_options.pivoting = 0)",
                     2: "test hook: diagonal path always rejected"}[args.pivoting],
    }
    if args.carry_dtchem or args.workload == "gri_skewed":
        cfg["deltaTChem0"] = "1e-07 on the first (warm-up) step, then carried from step to step as the model does"
    if extra:
        cfg.update(extra)
    return cfg


def cpu_oracle_rate(args, n_sample, nthreads, steps=1, warmup=0):
    """Oracle throughput on a bounded sample of the same workload, all host threads."""
    from loadbalancedchemistrymodel_b200 import capi
    from oracle import oracle as O
    s, atol, rtol = SOLVERS[args.solver]
    mech, Y, T, p, rho = make_workload(args.workload, n_sample, rank=0)
    om = O.OracleModel(mech, capi.default_options(solver=s, abs_tol=atol, rel_tol=rtol))
    if args.workload == "gri_dac" or getattr(args, "dac", False):
        om.set_reduction(DAC_TOL, dac_initial_set(mech))
    for _ in range(warmup):
        om.solve_batch(Y, T, p, rho, 1e-6, 1e-7, nthreads=nthreads)
    t0 = time.perf_counter()
    for _ in range(steps):
        r = om.solve_batch(Y, T, p, rho, 1e-6, 1e-7, nthreads=nthreads)
    dt = (time.perf_counter() - t0) / steps
    return n_sample / dt, dt, r["counters"]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_sample = args.cpu_sample or min(args.cells, 1500 * cores * CPU_SCALE.get(args.workload, 1))
    rate, dt, _ = cpu_oracle_rate(args, n_sample, cores, steps=args.steps, warmup=args.warmup)
    sample = f"{n_sample} cells of the workload per step, {cores} threads, static contiguous partition"
    line = {
        "impl": "reference", "metric": METRIC_OF.get(args.workload, METRIC), "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": config_dict(args, args.gpus),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU oracle = restatement of the reference's solveCell + OpenFOAM ODE stack (reference needs OpenFOAM; unbuildable here)",
    }
    _emit(json.dumps(line))


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for ln in self.f.read().splitlines():
            c = [x.strip() for x in ln.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return out


def _events():
    import torch
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


class Ranks:
    """torch.distributed plumbing of the bench: barrier, max / sum over ranks (device-timed numbers are max over ranks)."""

    def __init__(self, world, dev):
        self.world, self.dev = world, dev

    def barrier(self):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(self, x, op="max"):
        import torch
        import torch.distributed as dist
        t = torch.tensor([float(x)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())


IMB = {"solver": "seulex", "absTol": 1e-12, "relTol": 1e-3, "deltaT": 1e-4}


def cpu_plan_baseline(ranks, cells_per_rank):
    """The reference's own path on the host cores for the imbalanced batch: `ranks` MPI-like sub-domains (one oracle thread
    each, static contiguous partition) integrate a bounded sample of the same workload, per-cell wall clock recorded as
    the reference does (LoadBalancedChemistryModel.C:431-448); then the reference's plan (updateProcessorBalancing +
    cellsToSend on those per-cell times) is applied and the makespan of the re-distributed work is evaluated from the
    measured per-cell times. Returns cells/s without and with the plan."""
    from loadbalancedchemistrymodel_b200 import capi
    from loadbalancedchemistrymodel_b200 import workloads as W
    from oracle import oracle as O
    parts = [W.gri_skewed(cells_per_rank, rank=r, world=ranks) for r in range(ranks)]
    mech = parts[0][0]
    Y = np.ascontiguousarray(np.concatenate([q[1] for q in parts], axis=1))
    T, p, rho = (np.concatenate([q[k] for q in parts]) for k in (2, 3, 4))
    om = O.OracleModel(mech, capi.default_options(solver=IMB["solver"], abs_tol=IMB["absTol"], rel_tol=IMB["relTol"]))
    r1 = om.solve_batch(Y, T, p, rho, IMB["deltaT"], 1e-7, nthreads=ranks)               # firstTime_ step
    t0 = time.perf_counter()
    r2 = om.solve_batch(Y, T, p, rho, IMB["deltaT"], r1["deltaTChem"], nthreads=ranks)   # carried sub-step guess, timed
    wall = time.perf_counter() - t0
    ct = r2["cpu_time"].reshape(ranks, cells_per_rank)
    load = ct.sum(1)
    send, recv = capi.balance_plan(load, np.full(ranks, cells_per_rank, np.int32), 0.02)
    after = load.copy()
    for r in range(ranks):
        start = 0
        for to, frac in send[r]:
            end = capi.cells_to_send(ct[r], load[r] * frac, start)
            moved = ct[r, start:end].sum()
            after[r] -= moved; after[to] += moved
            start = end
    n = ranks * cells_per_rank
    return {"ranks": ranks, "sample_cells": n, "wall_s_unbalanced": wall, "cells_per_s_unbalanced": n / load.max(),
            "cells_per_s_with_reference_plan": n / after.max(), "rank_load_s": [float(x) for x in load],
            "rank_load_s_after_plan": [float(x) for x in after],
            "how": "oracle threads as MPI ranks; plan applied to the measured per-cell wall times (makespan = slowest rank)"}


def run_imbalanced(R, rank, local, total_cells, dac=False, steps=2, label="north_star", cpu_plan=False, economy=False):
    """The load-balancing experiment: `total_cells` in TOTAL (strong scaling), the hot 10 % contiguous at the start of the
    global index space (rank 0 holds them), deltaTChem carried from step to step, integrated once without and once
    with the redistribution (same lbchem_lb entry point, balance off / on). seulex 1e-12 / 1e-3 over deltaT = 1e-4 is the
    calibration at which the hot 10 % hold 90 % of the sub-steps (oracle counters: 82.4 vs 1.0 attempts)."""
    import torch
    from loadbalancedchemistrymodel_b200 import capi
    from loadbalancedchemistrymodel_b200 import workloads as W
    from loadbalancedchemistrymodel_b200.model import LoadBalancedChemistryModel

    world, dev = R.world, R.dev
    n = total_cells // world
    if dac:
        mech, Y, T, p, rho = W.gri_tiled(n, rank=rank)
        solver, atol, rtol, dT = "seulex", 1e-12, 1e-1, W.DELTA_T
    else:
        mech, Y, T, p, rho = W.gri_skewed(n, rank=rank, world=world)
        solver, atol, rtol, dT = IMB["solver"], IMB["absTol"], IMB["relTol"], IMB["deltaT"]
    nSp = mech.n_species
    opts = capi.default_options(solver=solver, abs_tol=atol, rel_tol=rtol, cost_model=1, pivoting=PIVOTING)
    chem = capi.Chemistry(mech, opts, device=local)
    if dac:
        chem.set_reduction(DAC_TOL, dac_initial_set(mech))
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    hY, hT, hp, hrho = pin(Y), pin(T), pin(p), pin(rho)
    dY, dT_, dp, drho = (t.to(dev) for t in (hY, hT, hp, hrho))
    dtc0 = torch.full((n,), W.DELTA_T_CHEM0, dtype=torch.float64, device=dev)
    rec = {"label": label, "total_cells": n * world, "cells_per_gpu": n, "solver": solver, "absTol": atol, "relTol": rtol, "deltaT": dT,
           "tdac": bool(dac), "cost_model": "work units 9 nLU + nSolve + 3 nRHS + 12 nJac (lbchem_options.cost_model = 1)",
           "deltaTChem": "1e-07 on step 0 (the firstTime_ local step), then carried"}
    results = {}
    shared = LoadBalancedChemistryModel(chem, tdac=dac, balance=False) if economy else None
    carried = dtc0
    for mode in ("unbalanced", "balanced"):
        # economy (the 4 M-cell config): ONE model -- step 0 local, step 1 timed with the exchange off, then the exchange is
        # switched on: one planning step, then the timed balanced steps; the expensive local step 0 is not repeated
        model = shared if economy else LoadBalancedChemistryModel(chem, tdac=dac, balance=(mode == "balanced"))
        if economy and mode == "balanced":
            model.set_balance(True)
        out = {"RR": torch.empty((nSp, n), dtype=torch.float64, device=dev), "T": torch.empty(n, dtype=torch.float64, device=dev),
               "cost": torch.empty(n, dtype=torch.float64, device=dev), "status": torch.empty(n, dtype=torch.int32, device=dev)}
        dtc = carried if economy else dtc0
        nwarm = 2 if not economy else 1                    # step 0 is local in both; the balanced model plans on step 1
        times, moved, nbytes, kms = [], 0, 0, []
        for it in range(nwarm + steps):
            R.barrier()
            e0, e1 = _events()
            e0.record()
            model.solve(dY, dT_, dp, drho, dT, dtc, out=out)
            e1.record()
            R.barrier()
            if not dac:
                dtc = model.deltaTChem                     # carried (LoadBalancedTDAC never writes it back, quirk D.2)
            if it >= nwarm:
                times.append(R.reduce(e0.elapsed_time(e1)))
                moved, nbytes = model.last["sent_cells"], model.last["sent_bytes"]
                kms.append(chem.stats()["kernel_ms"])
        st = chem.stats()
        results[mode] = dict(ms=float(np.mean(times)), RR=out["RR"].clone(), dtc=model.deltaTChem.clone(), cost=out["cost"].clone(),
                             moved=moved, nbytes=nbytes, total_cost=model.total_cost, kernel_ms=float(np.mean(kms)), failed=int(st["n_failed"]),
                             plan=list(model.send_plan))
        if mode == "balanced":
            # end to end through the load-balanced entry point: pinned HOST arrays, H2D + step + D2H inside the timing
            ho = {"RR": pin(np.empty((nSp, n))).numpy(), "T": pin(np.empty(n)).numpy(), "cost": pin(np.empty(n)).numpy(),
                  "status": pin(np.empty(n, np.int32)).numpy(), "deltaTChem": pin(np.empty(n)).numpy()}
            hd = model.deltaTChem.cpu().numpy() if not dac else np.full(n, W.DELTA_T_CHEM0)
            model.solve_host(hY.numpy(), hT.numpy(), hp.numpy(), hrho.numpy(), dT, hd, ho)
            R.barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                model.solve_host(hY.numpy(), hT.numpy(), hp.numpy(), hrho.numpy(), dT, hd, ho)
                _ = float(ho["cost"].sum())
            R.barrier()
            results["e2e_s"] = R.reduce((time.perf_counter() - t0) / steps)
        carried = dtc
        if not economy or mode == "balanced":
            model.close()
    ub, bl = results["unbalanced"], results["balanced"]
    same = bool(torch.equal(ub["RR"], bl["RR"]) and torch.equal(ub["dtc"], bl["dtc"]) and torch.equal(ub["cost"], bl["cost"]))
    same = (R.reduce(0.0 if same else 1.0) == 0.0) if not economy else None     # economy: the two results are different flow steps
    loads = None
    if world > 1:
        import torch.distributed as dist
        mine = torch.tensor([ub["total_cost"]], dtype=torch.float64, device=dev)
        allc = torch.empty(world, dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(allc, mine)
        loads = allc.cpu().numpy()
    else:
        loads = np.array([ub["total_cost"]])
    moved = R.reduce(bl["moved"], "sum"); nbytes = R.reduce(bl["nbytes"], "sum")
    ideal_ms = ub["ms"] * loads.mean() / loads.max()      # ms per cost unit (slowest rank, unbalanced) x the even share
    # measured split on a sample: attempts of hot and cold cells with the sub-step guess carried
    split = None
    if rank == 0 and not dac:
        ns = 2000
        mech2, Ys, Ts, ps, rhos = W.gri_skewed(2 * ns, hot_fraction=0.5, seed=77)
        c2 = capi.Chemistry(mech2, capi.default_options(solver=solver, abs_tol=atol, rel_tol=rtol, pivoting=PIVOTING), device=local)
        r1 = c2.solve_host(Ys, Ts, ps, rhos, dT, W.DELTA_T_CHEM0)
        r2 = c2.solve_host(Ys, Ts, ps, rhos, dT, r1["deltaTChem"])
        c2.close()
        h, c = float(r2["cost"][:ns].mean()), float(r2["cost"][ns:].mean())
        split = {"attempts_hot_cell": h, "attempts_cold_cell": c, "hot_10pct_share_of_attempts": 0.1 * h / (0.1 * h + 0.9 * c)}
    if cpu_plan and rank == 0 and not dac:
        cores = os.cpu_count() or 1
        rec["cpu_reference_plan"] = cpu_plan_baseline(cores, 96)
    rec.update({
        "ms_per_step_unbalanced": ub["ms"], "ms_per_step_balanced": bl["ms"], "speedup_from_balancing": ub["ms"] / bl["ms"],
        "cells_per_s_balanced": n * world / (bl["ms"] * 1e-3), "cells_per_s_unbalanced": n * world / (ub["ms"] * 1e-3),
        "cells_moved": int(moved), "nvlink_bytes_per_step": int(nbytes),
        "rank_cost_unbalanced": [float(x) for x in loads], "ideal_ms": ideal_ms, "efficiency_vs_ideal": ideal_ms / bl["ms"],
        "ideal_definition": "ms_per_step_unbalanced x mean(rank cost) / max(rank cost): the slowest rank's time per cost unit x an even share",
        "bit_identical_to_unbalanced": same, "failed_cells": int(R.reduce(ub["failed"] + bl["failed"], "sum")),
        "rank0_kernel_ms_balanced": bl["kernel_ms"], "rank0_kernel_ms_unbalanced": ub["kernel_ms"],
        "rank0_send_plan": [[int(t), float(f)] for t, f in bl["plan"]],
        "e2e": {"value": n * world / results["e2e_s"], "unit": UNIT, "ms_per_step": results["e2e_s"] * 1e3,
                "entry": "lbchem_lb_solve_host (pinned host arrays; H2D, plan, NVLink exchange, integrate, return, D2H)",
                "h2d_bytes_per_step": (nSp + 4) * n * 8, "d2h_bytes_per_step": (nSp + 3) * n * 8 + n * 4},
        "measured_split": split,
    })
    chem.close()
    return rec


def run_b200(args):
    import torch
    import torch.distributed as dist
    from loadbalancedchemistrymodel_b200 import capi
    from loadbalancedchemistrymodel_b200 import workloads as W

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    R = Ranks(world, dev)

    s, atol, rtol = SOLVERS[args.solver]
    n = args.cells
    mech, Y, T, p, rho = make_workload(args.workload, n, rank, world)
    nSp = mech.n_species
    chem = capi.Chemistry(mech, capi.default_options(solver=s, abs_tol=atol, rel_tol=rtol, pivoting=args.pivoting), device=local)
    if args.workload == "gri_dac" or args.dac:
        chem.set_reduction(DAC_TOL, dac_initial_set(mech))

    def pinned(a):
        return torch.from_numpy(np.ascontiguousarray(a)).pin_memory()

    hY, hT, hp, hrho = pinned(Y), pinned(T), pinned(p), pinned(rho)
    dY, dT_, dp, drho = (t.to(dev, non_blocking=True) for t in (hY, hT, hp, hrho))
    dtc0 = torch.full((n,), W.DELTA_T_CHEM0, dtype=torch.float64, device=dev)
    dtc = torch.empty_like(dtc0)
    dout = {"RR": torch.empty((nSp, n), dtype=torch.float64, device=dev), "T": torch.empty(n, dtype=torch.float64, device=dev),
            "cost": torch.empty(n, dtype=torch.float64, device=dev), "status": torch.empty(n, dtype=torch.int32, device=dev),
            "deltaTChem": dtc}
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream

    from loadbalancedchemistrymodel_b200.model import LoadBalancedChemistryModel
    # N > 1: every step goes through the library's load-balanced entry point (lbchem_lb_solve: C++ driver + NCCL)
    model = LoadBalancedChemistryModel(chem, tdac=(args.workload == "gri_dac")) if world > 1 else None
    moved = {"sent": 0, "recv": 0, "bytes": 0}
    carry = args.carry_dtchem or args.workload == "gri_skewed"   # reference .C:130,398,490: the sub-step guess persists
    state = {"first": True, "launches": 0}

    def step_device():
        # one LoadBalancedChemistryModel::solve(deltaT). N = 1: one kernel over the shard. N > 1: lbchem_lb_solve --
        # all-gather of the cost, bit-exact plan, cellsToSend on the device, NCCL p2p forward trip, own tail + received
        # cells in one launch, return trip, cost total (reference .C:453-704); the first call is its firstTime_ branch.
        if not carry or state["first"]:
            dtc.copy_(dtc0)
        state["first"] = False
        if model is None:
            chem.solve_device(n, n, dY, dT_, dp, drho, W.DELTA_T, dtc, dout["RR"], dout["T"], dout["cost"], dout["status"], stream=stream)
            state["launches"] += 1
        else:
            model.solve(dY, dT_, dp, drho, W.DELTA_T, dtc, out=dout)
            state["launches"] += 3                      # integration kernel + the two cost-vector kernels of the step
            moved["sent"] = model.last["sent_cells"]; moved["recv"] = model.last["recv_cells"]; moved["bytes"] = model.last["sent_bytes"]

    for _ in range(max(args.warmup, 3)):
        step_device()
    R.barrier()
    launches0 = state["launches"]
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = _events()
    R.barrier()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    R.barrier()
    ms = R.reduce(e0.elapsed_time(e1))
    gpu_launches = state["launches"] - launches0
    st = chem.stats()                      # counters + CUDA-event duration of the last integration kernel
    k_ms = float(st["kernel_ms"])
    ms_per_step = ms / args.steps
    value = n * world / (ms_per_step * 1e-3)
    n_failed = int(st["n_failed"])
    ref_RR = dout["RR"].cpu().numpy() if not carry else None

    # ---- end to end: pinned host buffers through the C ABI (H2D, integrate, D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        out = {"RR": pinned(np.empty((nSp, n))).numpy(), "T": pinned(np.empty(n)).numpy(), "cost": pinned(np.empty(n)).numpy(),
               "status": pinned(np.empty(n, np.int32)).numpy(), "deltaTChem": pinned(np.empty(n)).numpy()}
        args_h = (hY.numpy(), hT.numpy(), hp.numpy(), hrho.numpy(), W.DELTA_T, W.DELTA_T_CHEM0)

        def step_host():
            if model is None:
                chem.solve_host(*args_h, out=out)
            else:
                model.solve_host(*args_h, out)
        for _ in range(2):
            step_host()
        R.barrier()
        t0 = time.perf_counter()
        ke = args.steps
        for _ in range(ke):
            step_host()
            _ = float(out["cost"].sum())            # the step's result read on the host
        R.barrier()
        el = R.reduce((time.perf_counter() - t0) / ke)
        h2d = (nSp + 4) * n * 8
        d2h = (nSp + 3) * n * 8 + n * 4
        e2e = {"value": n * world / el, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": el * 1e3,
               "entry": "lbchem_solve_batch_host (3-stream chunked pipeline)" if model is None else "lbchem_lb_solve_host (load-balanced entry point)",
               "timing": "host wall clock around the blocking call, max over ranks"}
        if ref_RR is not None:
            # the chunks of the host path overlap on the device: their results must be those of the one-launch device path
            e2e["rr_bit_identical_to_device_path"] = bool(np.array_equal(out["RR"], ref_RR))
        if model is None and not args.no_pageable:
            pg = {k: np.empty_like(v) for k, v in out.items()}
            args_p = (np.array(Y), np.array(T), np.array(p), np.array(rho), W.DELTA_T, W.DELTA_T_CHEM0)
            chem.solve_host(*args_p, out=pg)
            t0 = time.perf_counter()
            for _ in range(2):
                chem.solve_host(*args_p, out=pg)
            e2e["pageable_value"] = n / ((time.perf_counter() - t0) / 2)
    clocks = sampler.stop() if sampler else None

    # ---- the load-balancing records (every N): the north_star batch, and at N = 8 BASELINE configs 4 and 5
    imbalanced, extra = None, {}
    if not args.no_imbalanced and args.workload == "gri_uniform":
        if model is not None:
            model.close(); model = None
        imbalanced = run_imbalanced(R, rank, local, args.imbalanced_cells, steps=args.imbalanced_steps,
                                    cpu_plan=(world == 1 and not args.no_cpu_baseline))
        if world == 8 and not args.no_extra_configs:
            extra["config4_gri_skewed_4M"] = run_imbalanced(R, rank, local, 4_000_000, steps=1, label="BASELINE config 4", economy=True)
            extra["config5_gri_dac_1M"] = run_imbalanced(R, rank, local, 1_000_000, dac=True, steps=3, label="BASELINE config 5")

    if rank == 0:
        # ---- roofline of the dominant (only) kernel: FP64 pipe; HBM figure beside it
        fm = W.flop_model(mech)
        flops = W.algorithmic_flops(fm, st)
        peak = capi.fp64_peak(local, 300.0)
        ach = flops / (k_ms * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        cell_bytes = (2 * nSp + 8) * 8                                  # (nSp + 5) doubles in, (nSp + 3) out: 912 B for GRI
        hbm_ach = cell_bytes * st["n_cells"] / (k_ms * 1e-3) / 1e9
        traffic, traffic_src = None, None   # dram bytes per launch from the committed ncu --set full capture of this kernel
        try:
            tr = json.loads((ROOT / "profiles" / "r02_traffic.json").read_text())
            key = f"{args.workload}:{args.solver}:{n}"
            if key in tr:
                traffic, traffic_src = tr[key]["dram_bytes"], tr[key].get("source")
        except Exception:
            pass
        roofline = {
            "bound": "fp64", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
            "traffic_source": traffic_src,
            "kernel": (f"integrate_v3_kernel<{min(w for w in (8, 16, 32, 64) if w >= nSp + 2)} threads per cell, solver> "
                       "(one launch per shard; lbchem_kernels_v3.cu)" if s != "EulerImplicit"
                       else "integrate_kernel<8 threads per cell> (general kernel, sub-warp groups; lbchem_kernels.cu)"), "kernel_ms": k_ms,
            "peak_source": "DFMA micro-benchmark (lbchem_fp64_peak) run in this process after the timed region; MEASURED_PEAKS.json has no FP64 entry",
            "flops_per_cell": flops / max(1, st["n_cells"]), "flop_model": "SURVEY 8(d): nLU*2/3 n^3 + nSolve*2 n^2 + nRHS*F_rhs + nJac*F_jac, transcendentals x20, reference-algorithm counts"
                          + (" at the FULL system size n = 55 (upper bound: the reduced systems are smaller)" if args.workload == "gri_dac" else ""),
            "hbm": {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                    "bytes_per_cell": cell_bytes, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
        }
        cpu = None
        if not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            n_sample = args.cpu_sample or min(n, 600 * cores * CPU_SCALE.get(args.workload, 1))
            rate, dtc_s, _ = cpu_oracle_rate(args, n_sample, cores, steps=1, warmup=1)
            cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{n_sample} cells of the same workload, second of two passes, {cores} threads ({dtc_s:.1f} s)"}
        line = {
            "metric": METRIC_OF.get(args.workload, METRIC), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config_dict(args, world), "clocks": clocks, "e2e": e2e,
            "gpu_launches": gpu_launches, "roofline": roofline, "cpu_baseline": cpu,
            "stats_last_step": {k: int(v) for k, v in st.items() if k.startswith("n_")}, "failed_cells": n_failed,
            "balance_last_step": ({"rank0_sent_cells": moved["sent"], "rank0_recv_cells": moved["recv"],
                                   "note": "uniform cost: the plan is empty by construction; see `imbalanced` for the exchange"} if world > 1 else None),
            "imbalanced": imbalanced,
        }
        line.update(extra)
        _emit(json.dumps(line))
    if model is not None:
        model.close()
    chem.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    global PIVOTING
    args = parse()
    PIVOTING = args.pivoting
    # Libraries (NCCL's version banner, torchrun notices) may write to stdout; the contract is ONE JSON line there.
    # Everything but the final line goes to stderr.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    global _emit
    def _emit(line):
        sys.stdout.flush()
        os.write(real_stdout, (line + "\n").encode())
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
