"""Diagnostic: worst parity offenders GPU vs oracle on the flame fixture (run on the GPU box)."""
import sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from loadbalancedchemistrymodel_b200 import capi
from loadbalancedchemistrymodel_b200.foam_io import Mechanism
from oracle import oracle as O

solver, rtol = (sys.argv[1], float(sys.argv[2])) if len(sys.argv) > 2 else ("Rosenbrock34", 1e-2)
mech = Mechanism.load(ROOT / "tests/golden/gri30.npz")
z = np.load(ROOT / "tests/golden/case_chemistry_0p2.npz")
T, p, Y = z["T"], z["p"], np.ascontiguousarray(z["Y"])
rho = O.density(mech, Y, T, p)
opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rtol)
chem = capi.Chemistry(mech, opts)
t0 = time.time(); got = chem.solve_host(Y, T, p, rho, 1e-6, 1e-7); t1 = time.time()
print("gpu host-path wall", t1 - t0, chem.stats())
t0 = time.time(); ref = O.OracleModel(mech, opts).solve_batch(Y, T, p, rho, 1e-6, 1e-7, nthreads=8); t1 = time.time()
print("oracle wall (8 thr)", t1 - t0)
dt = 1e-6
scale = np.maximum(np.abs(ref["RR"]), 1e-9 * rho * np.maximum(Y, 1e-30) / dt)
eR = np.abs(got["RR"] - ref["RR"]) / scale
print("cost equal", np.array_equal(got["cost"], ref["attempts"]), "status eq", np.array_equal(got["status"], ref["status"]))
idx = np.dstack(np.unravel_index(np.argsort(-eR, axis=None)[:15], eR.shape))[0]
for i, c in idx:
    print(f"{mech.species[i]:8s} cell {c:5d} T {T[c]:8.2f} Y {Y[i,c]:.3e} RRref {ref['RR'][i,c]: .6e} RRgpu {got['RR'][i,c]: .6e} "
          f"eR {eR[i,c]:.2e} dY/Y {abs(ref['RR'][i,c])*dt/rho[c]/max(Y[i,c],1e-300):.2e} att {ref['attempts'][c]:.0f}")
relRR = np.abs(got["RR"] - ref["RR"]) / np.maximum(np.abs(ref["RR"]), 1e-300)
for thr in (1e-6, 1e-8, 1e-10):
    print("frac entries with pure rel RR err >", thr, (relRR > thr).mean())
Yg = Y + got["RR"] * dt / rho; Yr = Y + ref["RR"] * dt / rho
eY = np.abs(Yg - Yr) / np.maximum(np.abs(Yr), 1e-300)
i, c = np.unravel_index(np.argmax(np.where(np.abs(Yr) > 1e-20, eY, 0)), eY.shape)
print("worst Y:", mech.species[i], c, Yr[i, c], eY[i, c])
print("dtChem max rel", (np.abs(got["deltaTChem"] - ref["deltaTChem"]) / ref["deltaTChem"]).max())
d = np.nonzero(got["cost"] != ref["attempts"])[0]
print("cells with different attempt counts:", len(d))
for c in d[:20]:
    print(f" cell {c} T {T[c]:.2f} att gpu {got['cost'][c]:.0f} ref {ref['attempts'][c]:.0f} dtChem gpu {got['deltaTChem'][c]:.6e} ref {ref['deltaTChem'][c]:.6e} Tout diff {got['T'][c]-ref['T'][c]:.3e}")
e = np.abs(got["deltaTChem"] - ref["deltaTChem"]) / ref["deltaTChem"]
b = np.nonzero(e > 1e-6)[0]
print("cells with dtChem rel diff > 1e-6:", len(b), b[:20], e[b[:20]])
