"""Diagnosis of test_seulex_tight_tolerance...: where does the GPU-vs-oracle RR difference sit (same / different step sequence)?"""
import numpy as np
from loadbalancedchemistrymodel_b200 import capi, workloads as W
from oracle import oracle as O

gri = W.load_gri(); T0, p0, Y0 = W.load_flame()
sel = np.argsort(-T0)[::29][:96]
T, p, Y = T0[sel], p0[sel], np.ascontiguousarray(Y0[:, sel]); rho = O.density(gri, Y, T, p)
dt = 1e-5
for piv in (0, 1):
    opts = capi.default_options(solver="seulex", abs_tol=1e-14, rel_tol=1e-8, pivoting=piv)
    chem = capi.Chemistry(gri, opts); got = chem.solve_host(Y, T, p, rho, dt, 1e-7); chem.close()
    ref = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, dt, 1e-7, nthreads=8)
    Wm = gri.W[:, None]
    eR = np.abs(got["RR"] - ref["RR"]) / (np.abs(ref["RR"]) + 1e-14 * Wm / dt)
    same = got["cost"] == ref["attempts"]
    i, c = np.unravel_index(np.argmax(eR), eR.shape)
    c0 = rho[c] * Y[i, c] / gri.W[i]
    print(f"pivoting {piv}: eR max {eR.max():.3e} (same-sequence cells {eR[:, same].max():.3e}, others {eR[:, ~same].max() if (~same).any() else 0:.3e}), "
          f"other cells {int((~same).sum())}/96; worst: species {gri.species[i]} cell {c} same={bool(same[c])} attempts gpu {got['cost'][c]} oracle {ref['attempts'][c]} "
          f"RR gpu {got['RR'][i, c]:.6e} oracle {ref['RR'][i, c]:.6e} c0 {c0:.3e} |dc|/c0 {abs(ref['RR'][i, c]) * dt / gri.W[i] / max(c0, 1e-300):.2e}")
    # the oracle against itself with deltaT moved by one ulp: how far do two valid runs at this tolerance differ?
    ref2 = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, np.nextafter(dt, 1.0), 1e-7, nthreads=8)
    eO = np.abs(ref2["RR"] - ref["RR"]) / (np.abs(ref["RR"]) + 1e-14 * Wm / dt)
    print(f"   oracle vs oracle (deltaT + 1 ulp): eR max {eO.max():.3e}, cells with another attempt count {int((ref2['attempts'] != ref['attempts']).sum())}")
