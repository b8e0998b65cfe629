"""Regenerate tests/golden/dac_regression.npz: the CPU oracle's DAC-reduced outputs on 64 cells of the flame fixture
(`reduction{active on; method DAC; tolerance 1e-4; initialSet CO CH4 HO2}`, reference
tests/Cases/Case-chemistry/constant/chemistryProperties:35-66, seulex 1e-12 / 1e-1).

Drift guard for oracle/lbchem_oracle.c (dac_reduce, tdac_*), NOT a pin against OpenFOAM (parity unpinned).
Run in the build container: python tests/golden/make_dac_regression.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from loadbalancedchemistrymodel_b200 import capi  # noqa: E402
from loadbalancedchemistrymodel_b200.foam_io import Mechanism  # noqa: E402
from oracle import oracle as O  # noqa: E402

TOL, INIT = 1e-4, ("CO", "CH4", "HO2")


def main():
    mech = Mechanism.load(ROOT / "tests/golden/gri30.npz")
    z = np.load(ROOT / "tests/golden/case_chemistry_0p2.npz")
    sel = np.argsort(-z["T"])[::62][:64]
    T, p, Y = z["T"][sel], z["p"][sel], np.ascontiguousarray(z["Y"][:, sel])
    rho = O.density(mech, Y, T, p)
    init = [list(mech.species).index(s) for s in INIT]
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    om = O.OracleModel(mech, opts)
    om.set_reduction(TOL, init)
    r = om.solve_batch(Y, T, p, rho, 1e-6, 1e-7)
    active = np.stack([om.dac_reduce(rho[k] * Y[:, k] / mech.W, T[k], TOL, init)[0] for k in range(sel.size)], 1)
    np.savez_compressed(ROOT / "tests/golden/dac_regression.npz", cells=sel, RR=r["RR"], deltaTChem=r["deltaTChem"],
                        attempts=r["attempts"], n_active=r["n_active"], active=active)
    print("wrote", sel.size, "cells; NsDAC", r["n_active"].min(), r["n_active"].max())


if __name__ == "__main__":
    main()
