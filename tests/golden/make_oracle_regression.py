"""Regenerate tests/golden/oracle_regression.npz: the CPU oracle's own outputs on 32 cells of the flame fixture.

This is a drift guard for oracle/lbchem_oracle.c, NOT a pin against OpenFOAM (parity unpinned).
Run in the build container: python tests/golden/make_oracle_regression.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from loadbalancedchemistrymodel_b200 import capi  # noqa: E402
from loadbalancedchemistrymodel_b200.foam_io import Mechanism  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    mech = Mechanism.load(ROOT / "tests/golden/gri30.npz")
    z = np.load(ROOT / "tests/golden/case_chemistry_0p2.npz")
    sel = np.argsort(-z["T"])[::125][:32]
    T, p, Y = z["T"][sel], z["p"][sel], np.ascontiguousarray(z["Y"][:, sel])
    rho = O.density(mech, Y, T, p)
    out = {"cells": sel}
    for solver, rel_tol in (("Rosenbrock34", 1e-2), ("seulex", 1e-1)):
        opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol)
        r = O.OracleModel(mech, opts).solve_batch(Y, T, p, rho, 1e-6, 1e-7)
        for k in ("RR", "T", "deltaTChem", "attempts"):
            out[f"{solver}_{k}"] = r[k]
    np.savez_compressed(ROOT / "tests/golden/oracle_regression.npz", **out)
    print("wrote", len(sel), "cells")


if __name__ == "__main__":
    main()
