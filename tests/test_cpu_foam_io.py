"""The OpenFOAM dictionary / field readers against the reference's own case files.

`tests/golden/*.npz` were produced by `tests/golden/make_fixtures.py` from `/root/reference`; here the reference files are
parsed AGAIN at test time (build container only: the GPU box has no reference checkout, the test skips there) and compared
with the committed fixtures bit for bit, so that neither the parser nor the fixtures can drift unnoticed."""
from pathlib import Path

import numpy as np
import pytest

from loadbalancedchemistrymodel_b200.foam_io import Mechanism, read_case_state, read_mechanism, read_scalar_field

ROOT = Path(__file__).resolve().parents[1]
REF = Path("/root/reference")
pytestmark = pytest.mark.skipif(not (REF / "tests/Cases/Case-chemistry").exists(), reason="reference checkout not present")


def _same_mechanism(a: Mechanism, b: Mechanism):
    assert list(a.species) == list(b.species)
    for name in ("W", "Tlow", "Thigh", "Tcommon", "janaf_high", "janaf_low", "rtype", "reversible", "arr", "arr0", "troe",
                 "eff_index", "eff", "lhs_ptr", "lhs_species", "lhs_nu", "lhs_exp", "rhs_ptr", "rhs_species", "rhs_nu", "rhs_exp"):
        x, y = np.asarray(getattr(a, name)), np.asarray(getattr(b, name))
        assert x.shape == y.shape and np.array_equal(x, y), name


def test_gri_mechanism_files_parse_to_the_committed_fixture():
    case = REF / "tests/Cases/Case-chemistry/constant"
    mech = read_mechanism(case / "reactionsGRI", case / "thermo.compressibleGasGRI")
    assert mech.n_species == 53 and mech.n_reactions == 325
    _same_mechanism(mech, Mechanism.load(ROOT / "tests/golden/gri30.npz"))
    # reaction classes of SURVEY section 8: 268 + 16 Arrhenius, 12 third-body, 26 Troe, 3 Lindemann
    rt, rev = np.asarray(mech.rtype), np.asarray(mech.reversible)
    assert ((rt == 0) & (rev == 1)).sum() == 268 and ((rt == 0) & (rev == 0)).sum() == 16
    assert (rt == 1).sum() == 12 and (rt == 3).sum() == 26 and (rt == 2).sum() == 3
    assert int(np.asarray(mech.lhs_ptr)[-1] + np.asarray(mech.rhs_ptr)[-1]) == 1262


def test_tutorial_mechanism_parses_to_the_committed_fixture():
    tut = REF / "tutorial/counterFlowFlame2D/constant"
    mech = read_mechanism(tut / "reactions", tut / "thermo.compressibleGas")
    assert mech.n_species == 5 and mech.n_reactions == 1
    _same_mechanism(mech, Mechanism.load(ROOT / "tests/golden/ch4_1step.npz"))


def test_binary_fields_parse_to_the_committed_fixture():
    case = REF / "tests/Cases/Case-chemistry"
    gri = Mechanism.load(ROOT / "tests/golden/gri30.npz")
    T, p, Y = read_case_state(case / "0.2", list(gri.species))
    z = np.load(ROOT / "tests/golden/case_chemistry_0p2.npz")
    assert T.shape == (4000,) and np.array_equal(T, z["T"]) and np.array_equal(np.broadcast_to(p, (4000,)), np.broadcast_to(z["p"], (4000,)))
    assert np.array_equal(Y, z["Y"])
    assert np.array_equal(read_scalar_field(case / "0.2/Qdot", 4000), z["Qdot"])
    # what the survey recorded about the state (Appendix C)
    assert abs(T.min() - 293.0) < 0.1 and abs(T.max() - 1875.54) < 0.01 and ((Y.sum(0) - 1.0).__abs__() < 2e-6).all()
