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
UNIT = "cells/s"
SOLVERS = {"seulex": ("seulex", 1e-12, 1e-1), "Rosenbrock34": ("Rosenbrock34", 1e-12, 1e-2)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cells", type=int, default=1_000_000, help="cells per GPU")
    ap.add_argument("--solver", default="seulex", choices=list(SOLVERS))
    ap.add_argument("--workload", default="gri_uniform", choices=["gri_uniform", "gri_skewed", "gri_dac", "gri_tiled"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="cells in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--carry-dtchem", action="store_true",
                    help="carry deltaTChem from step to step like the model does between flow steps (default for gri_skewed)")
    return ap.parse_args()


def make_workload(name, n, rank, world=1):
    from loadbalancedchemistrymodel_b200 import workloads as W
    if name == "gri_skewed":
        return W.gri_skewed(n, rank=rank, world=world)
    if name in ("gri_dac", "gri_tiled"):
        return W.gri_tiled(n, rank=rank)
    return W.gri_uniform(n, rank=rank)


DAC_TOL, DAC_INIT = float(os.environ.get("LBCHEM_DAC_TOL", "1e-4")), ("CO", "CH4", "HO2")   # reference tests/Cases/Case-chemistry/constant/chemistryProperties:46,52-57


def dac_initial_set(mech):
    return [list(mech.species).index(s) for s in DAC_INIT]


WORKLOAD_TEXT = {
    "gri_uniform": "BASELINE config 3; fixture cells T>1500 K resampled + noise, seed 1003",
    "gri_skewed": "BASELINE config 4; hot cells (T>1500 K) contiguous at the start of the global index space, cold cells T<350 K, seed 1004",
    "gri_tiled": "BASELINE config 5 input without the reduction: the unfiltered 4000-cell fixture tiled + noise, seed 1005",
    "gri_dac": "BASELINE config 5; the unfiltered 4000-cell fixture tiled + noise, seed 1005; reduction{active on; method DAC; "
               "tolerance 1e-4; initialSet CO CH4 HO2}, tabulation off",
}


def config_dict(args, n_gpus, extra=None):
    s, atol, rtol = SOLVERS[args.solver]
    cfg = {
        "workload": f"{args.workload}: GRI-Mech 3.0 (53 species, 325 reactions), {args.cells} synthetic cells per GPU "
                    f"({WORKLOAD_TEXT[args.workload]})",
        "mechanism": "GRI-3.0", "cells_per_gpu": args.cells, "solver": s, "absTol": atol, "relTol": rtol,
        "deltaT": 1e-6, "deltaTChem0": 1e-7, "parallelism": (f"cells sharded over {n_gpus} GPUs; per step: all-gather of the cost, bit-exact plan, NCCL p2p forward/return of planned cells (model.py)" if n_gpus > 1 else "1 GPU, one persistent kernel per step"),
        "l2": "inputs (464 B/cell x cells) larger than L2; no flush needed",
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
    if args.workload == "gri_dac":
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
    n_sample = args.cpu_sample or min(args.cells, 1500 * cores)
    rate, dt, _ = cpu_oracle_rate(args, n_sample, cores, steps=args.steps, warmup=args.warmup)
    sample = f"{n_sample} cells of the workload per step, {cores} threads, static contiguous partition"
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
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

    s, atol, rtol = SOLVERS[args.solver]
    n = args.cells
    mech, Y, T, p, rho = make_workload(args.workload, n, rank, world)
    nSp = mech.n_species
    chem = capi.Chemistry(mech, capi.default_options(solver=s, abs_tol=atol, rel_tol=rtol), device=local)
    if args.workload == "gri_dac":
        chem.set_reduction(DAC_TOL, dac_initial_set(mech))

    def pinned(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t

    hY, hT, hp, hrho = pinned(Y), pinned(T), pinned(p), pinned(rho)
    dY, dT_, dp, drho = (t.to(dev, non_blocking=True) for t in (hY, hT, hp, hrho))
    dtc0 = torch.full((n,), W.DELTA_T_CHEM0, dtype=torch.float64, device=dev)
    dtc = torch.empty_like(dtc0)
    dRR = torch.empty((nSp, n), dtype=torch.float64, device=dev)
    dTout = torch.empty(n, dtype=torch.float64, device=dev)
    dcost = torch.empty(n, dtype=torch.float64, device=dev)
    dstatus = torch.empty(n, dtype=torch.int32, device=dev)
    gathered = torch.zeros(world, dtype=torch.float64, device=dev)
    cells_per_rank = np.full(world, n, np.int32)
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream

    from loadbalancedchemistrymodel_b200.model import CudaCellSolver, LoadBalancedChemistryModel
    cell_solver = CudaCellSolver(chem)
    model = LoadBalancedChemistryModel(nSp, cell_solver, dev) if world > 1 else None
    moved = {"sent": 0, "recv": 0}
    carry = args.carry_dtchem or args.workload == "gri_skewed"   # reference .C:130,398,490: the sub-step guess persists
    state = {"first": True}

    def step_device():
        # one LoadBalancedChemistryModel::solve(deltaT). N = 1: one kernel over the shard. N > 1: the host driver
        # (model.py) -- all-gather of the cost, bit-exact plan, cellsToSend scan, NCCL p2p forward trip, own tail then
        # received cells, return trip, unpack (reference .C:453-704); the first call is its firstTime_ branch.
        if model is None:
            if not carry or state["first"]:
                dtc.copy_(dtc0)
            state["first"] = False
            chem.solve_device(n, n, dY, dT_, dp, drho, W.DELTA_T, dtc, dRR, dTout, dcost, dstatus, stream=stream)
            cell_solver.launches += 1
        else:
            model.solve(dY, dT_, dp, drho, W.DELTA_T, None if (carry and model.deltaTChem is not None) else dtc0)
            moved["sent"] = model.last.get("sent_cells", 0); moved["recv"] = model.last.get("recv_cells", 0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    launches0 = cell_solver.launches
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    gpu_launches = cell_solver.launches - launches0
    st = chem.stats()                      # counters + CUDA-event duration of the last integration kernel
    kernel_ms.append(st["kernel_ms"])
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    ms_per_step = ms / args.steps
    value = n * world / (ms_per_step * 1e-3)
    n_failed = int(st["n_failed"])

    # ---- end to end: pinned host buffers through the C ABI (H2D, integrate, D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        out = {"RR": pinned(np.empty((nSp, n))).numpy(), "T": pinned(np.empty(n)).numpy(), "cost": pinned(np.empty(n)).numpy(),
               "status": pinned(np.empty(n, np.int32)).numpy(), "deltaTChem": pinned(np.empty(n)).numpy()}
        args_h = (hY.numpy(), hT.numpy(), hp.numpy(), hrho.numpy(), W.DELTA_T, W.DELTA_T_CHEM0)
        for _ in range(2):
            chem.solve_host(*args_h, out=out)
        barrier()
        t0 = time.perf_counter()
        ke = args.steps
        for _ in range(ke):
            r = chem.solve_host(*args_h, out=out)
            _ = float(r["cost"].sum())            # the step's result read on the host
        barrier()
        el = (time.perf_counter() - t0) / ke
        t = torch.tensor([el], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        el = float(t.item())
        h2d = (nSp + 4) * n * 8
        d2h = (nSp + 3) * n * 8 + n * 4
        e2e = {"value": n * world / el, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": el * 1e3, "timing": "host wall clock around lbchem_solve_batch_host (blocking), max over ranks"}
    clocks = sampler.stop() if sampler else None

    if rank == 0:
        # ---- roofline of the dominant (only) kernel: FP64 pipe; HBM figure beside it
        model = W.flop_model(mech)
        flops = W.algorithmic_flops(model, st)
        k_ms = float(np.mean(kernel_ms))
        peak = capi.fp64_peak(local, 300.0)
        ach = flops / (k_ms * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        hbm_ach = 912.0 * n / (k_ms * 1e-3) / 1e9
        traffic = None          # dram bytes per launch from the committed ncu --set full capture of this same configuration
        try:
            tr = json.loads((ROOT / "profiles" / "r01_traffic.json").read_text())
            key = f"{args.solver}:{n}"
            if key in tr:
                traffic = tr[key]["dram_bytes"]
        except Exception:
            pass
        roofline = {
            "bound": "fp64", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
            "kernel": "v3::integrate_v3_kernel<solver> (one launch per shard; lbchem_kernels_v3.cu)", "kernel_ms": k_ms,
            "peak_source": "DFMA micro-benchmark (lbchem_fp64_peak) run in this process after the timed region; MEASURED_PEAKS.json has no FP64 entry",
            "flops_per_cell": flops / n, "flop_model": "SURVEY 8(d): nLU*2/3 n^3 + nSolve*2 n^2 + nRHS*F_rhs + nJac*F_jac, transcendentals x20, reference-algorithm counts"
                          + (" at the FULL system size n = 55 (upper bound: the reduced systems are smaller)" if args.workload == "gri_dac" else ""),
            "hbm": {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                    "bytes_per_cell": 912, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
        }
        cpu = None
        if not args.no_cpu_baseline and world >= 1:
            cores = os.cpu_count() or 1
            n_sample = args.cpu_sample or min(n, 600 * cores)
            rate, dtc_s, _ = cpu_oracle_rate(args, n_sample, cores)
            cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{n_sample} cells of the same workload, one pass, {cores} threads ({dtc_s:.1f} s)"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config_dict(args, world), "clocks": clocks, "e2e": e2e,
            "gpu_launches": gpu_launches, "roofline": roofline, "cpu_baseline": cpu,
            "stats_last_step": {k: int(v) for k, v in st.items() if k.startswith("n_")}, "failed_cells": n_failed,
            "balance_last_step": ({"rank0_sent_cells": moved["sent"], "rank0_recv_cells": moved["recv"]} if world > 1 else None),
        }
        _emit(json.dumps(line))
    chem.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
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
