"""Regenerate the committed fixtures from the reference checkout (run in the build container only).

Inputs (read-only, never present on the GPU box):
  /root/reference/tests/Cases/Case-chemistry/constant/{reactionsGRI,thermo.compressibleGasGRI}
  /root/reference/tests/Cases/Case-chemistry/0.2/{T,p,<53 species>,Qdot}
  /root/reference/tutorial/counterFlowFlame2D/constant/{reactions,thermo.compressibleGas}
Outputs (committed): tests/golden/gri30.npz, ch4_1step.npz, case_chemistry_0p2.npz
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from loadbalancedchemistrymodel_b200.foam_io import read_case_state, read_mechanism, read_scalar_field  # noqa: E402

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent


def main():
    case = REF / "tests/Cases/Case-chemistry"
    gri = read_mechanism(case / "constant/reactionsGRI", case / "constant/thermo.compressibleGasGRI")
    gri.save(OUT / "gri30.npz")
    tut = REF / "tutorial/counterFlowFlame2D/constant"
    ch4 = read_mechanism(tut / "reactions", tut / "thermo.compressibleGas")
    ch4.save(OUT / "ch4_1step.npz")
    T, p, Y = read_case_state(case / "0.2", gri.species)
    Qdot = read_scalar_field(case / "0.2/Qdot", T.shape[0])
    np.savez_compressed(OUT / "case_chemistry_0p2.npz", T=T, p=p, Y=Y, Qdot=Qdot)
    print("gri", gri.n_species, gri.n_reactions, "ch4", ch4.n_species, ch4.n_reactions, "cells", T.shape[0])


if __name__ == "__main__":
    main()
