"""Tail-of-launch effect of cost-ordered dispatch (lbchem_options.sort_by_cost) on a SHUFFLED skewed batch.

The north_star batch (10 % hot cells holding 90 % of the sub-steps; seulex 1e-12/1e-3 over deltaT = 1e-4, deltaTChem carried)
with the hot cells scattered at random positions instead of lying first: with the natural queue order the last cells
handed out may be 80-attempt cells that keep a few SMs busy while the rest idle; ordered by last step's cost they start first.
Run on the GPU box: python scripts/sort_effect.py [cells]"""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from loadbalancedchemistrymodel_b200 import capi, workloads as W

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
dev = torch.device("cuda:0")
res = {}
for layout in ("hot_first", "shuffled"):
    mech, Y, T, p, rho = W.gri_skewed(n, hot_first=(layout == "hot_first"))
    tY, tT, tp, trho = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (Y, T, p, rho))
    for sort in (0, 1):
        chem = capi.Chemistry(mech, capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-3, sort_by_cost=sort, cost_model=1))
        dtc = torch.full((n,), 1e-7, dtype=torch.float64, device=dev)
        RR = torch.empty((mech.n_species, n), dtype=torch.float64, device=dev)
        To = torch.empty(n, dtype=torch.float64, device=dev); co = torch.zeros(n, dtype=torch.float64, device=dev)
        so = torch.empty(n, dtype=torch.int32, device=dev)
        ms = []
        for step in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            chem.solve_device(n, n, tY, tT, tp, trho, 1e-4, dtc, RR, To, co, so)
            e1.record(); torch.cuda.synchronize()
            if step >= 2:
                ms.append(e0.elapsed_time(e1))
        res[(layout, sort)] = (float(np.mean(ms)), RR.cpu().numpy())
        print(f"{layout:10s} sort_by_cost={sort}: {np.mean(ms):8.2f} ms per step ({n} cells, steps 3-5, launches {chem.stats()['kernel_launches']})")
        chem.close()
    assert np.array_equal(res[(layout, 0)][1], res[(layout, 1)][1]), "dispatch order changed a result"
print("results identical with and without cost-ordered dispatch")
