"""Small fixed workload for ncu: N hot GRI cells, one solve on device-resident... (host path, 1 chunk)."""
import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from loadbalancedchemistrymodel_b200 import capi, workloads as W

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
solver = sys.argv[2] if len(sys.argv) > 2 else "seulex"
rtol = {"seulex": 1e-1, "Rosenbrock34": 1e-2}[solver]
mech, Y, T, p, rho = W.gri_uniform(n)
chem = capi.Chemistry(mech, capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rtol))
import os
os.environ["LBCHEM_CHUNK"] = str(n)
for _ in range(2):
    r = chem.solve_host(Y, T, p, rho, W.DELTA_T, W.DELTA_T_CHEM0)
st = chem.stats()
print(solver, n, "cells", st, "cells/s (kernel)", n / (st["kernel_ms"] * 1e-3), "failed", int((r["status"] > 1).sum()))
