"""CPU checks of the oracle's DAC path (LoadBalancedTDAC solveCell with `reduction{active on; method DAC}`,
reference src/chemistryModel/LoadBalancedTDACChemistryModel/LoadBalancedTDACChemistryModel.C:360-435).

The reference has no test with DAC switched on (gap iii), and OpenFOAM's DAC.C is not on this
machine, so the restatement is held by: an independent pure-Python transliteration of the graph search on the same
reaction rates, the limits of the tolerance (0 -> nothing removed that the search can reach, > 1 -> only the initial
set), conservation, convergence towards the full mechanism as the tolerance shrinks, the reference's T/p aliasing
quirk, and a committed regression vector.
"""
from collections import deque
from pathlib import Path

import numpy as np
import pytest

from loadbalancedchemistrymodel_b200 import capi
from oracle import oracle as O

ROOT = Path(__file__).resolve().parents[1]
DT, DTCHEM = 1e-6, 1e-7
TOL, INIT = 1e-4, ("CO", "CH4", "HO2")     # tests/Cases/Case-chemistry/constant/chemistryProperties:46,52-57


def _state(mech, flame, sel):
    T, p, Y = flame["T"][sel], flame["p"][sel], np.ascontiguousarray(flame["Y"][:, sel])
    return Y, T, p, O.density(mech, Y, T, p)


def _init(mech):
    return [list(mech.species).index(s) for s in INIT]


def py_dac(mech, omega, tol, init):
    """DAC graph search written independently of the C code: dictionaries keyed by (A, B), insertion-ordered."""
    nSp = mech.n_species
    num = [dict() for _ in range(nSp)]           # A -> {B: signed numerator}, first-appearance order of B
    PA, CA = np.zeros(nSp), np.zeros(nSp)
    for r in range(mech.n_reactions):
        net = {}
        for k in range(mech.lhs_ptr[r], mech.lhs_ptr[r + 1]):
            s = int(mech.lhs_species[k]); net[s] = net.get(s, 0.0) - mech.lhs_nu[k] * omega[r]
        for k in range(mech.rhs_ptr[r], mech.rhs_ptr[r + 1]):
            s = int(mech.rhs_species[k]); net[s] = net.get(s, 0.0) + mech.rhs_nu[k] * omega[r]
        for A, wA in net.items():
            for B in net:
                if B != A:
                    num[A][B] = num[A].get(B, 0.0) + wA
            if wA > 0:
                PA[A] += wA
            else:
                CA[A] -= wA
    R = np.zeros(nSp); active = np.zeros(nSp, bool); q = deque()
    for s in init:
        if not active[s]:
            active[s] = True; R[s] = 1.0; q.append(s)
    while q:
        u = q.popleft()
        den = max(PA[u], CA[u])
        if den > 1e-300:
            for B, n_ in num[u].items():
                rt = R[u] * min(abs(n_) / den, 1.0)
                if rt >= tol:
                    R[B] = max(R[B], rt)
                    if not active[B]:
                        active[B] = True; q.append(B)
    return active


def _omega(mech, om, c, T):
    kf, kr = om.rates(c, T)
    w = np.zeros(mech.n_reactions)
    cc = np.maximum(c, 0.0)
    for r in range(mech.n_reactions):
        f = kf[r]; b = kr[r]
        for k in range(mech.lhs_ptr[r], mech.lhs_ptr[r + 1]):
            f *= cc[mech.lhs_species[k]] ** mech.lhs_exp[k]
        for k in range(mech.rhs_ptr[r], mech.rhs_ptr[r + 1]):
            b *= cc[mech.rhs_species[k]] ** mech.rhs_exp[k]
        w[r] = f - b
    return w


def test_dac_search_matches_python_transliteration(gri, flame):
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    om = O.OracleModel(gri, opts)
    init = _init(gri)
    sel = np.argsort(-flame["T"])[::50][:80]
    Y, T, p, rho = _state(gri, flame, sel)
    differing = 0
    for k in range(sel.size):
        c = rho[k] * Y[:, k] / gri.W
        act, dis = om.dac_reduce(c, T[k], TOL, init)
        want = py_dac(gri, _omega(gri, om, c, T[k]), TOL, init)
        # the two evaluate omega in different multiplication orders: a path value within round-off of the tolerance may flip
        differing += int((act != want).any())
        # a reaction is disabled exactly when it touches a removed species
        for r in range(gri.n_reactions):
            sp = list(gri.lhs_species[gri.lhs_ptr[r]:gri.lhs_ptr[r + 1]]) + list(gri.rhs_species[gri.rhs_ptr[r]:gri.rhs_ptr[r + 1]])
            assert bool(dis[r]) == (not act[sp].all())
        assert act[init].all()
    assert differing <= 1, differing


def test_dac_tolerance_limits(gri, flame):
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    om = O.OracleModel(gri, opts)
    init = _init(gri)
    k = int(np.argmax(flame["T"]))
    Y, T, p, rho = _state(gri, flame, np.array([k]))
    c = rho[0] * Y[:, 0] / gri.W
    act_hi, dis_hi = om.dac_reduce(c, T[0], 2.0, init)          # r_AB <= 1 < tolerance: only the initial set survives
    assert act_hi.sum() == len(init) and act_hi[init].all()
    act_lo, _ = om.dac_reduce(c, T[0], 1e-300, init)            # everything the graph reaches is kept
    act_mid, _ = om.dac_reduce(c, T[0], TOL, init)
    assert act_lo.sum() >= act_mid.sum() >= act_hi.sum()
    assert (act_lo | ~act_mid).all() and act_lo.sum() >= gri.n_species - 3   # AR and the C3 species may be unreachable
    # nested: a looser tolerance never keeps a species a tighter one removed
    prev = act_hi
    for tol in (1e-1, 1e-2, 1e-3, 1e-4, 1e-6):
        a, _ = om.dac_reduce(c, T[0], tol, init)
        assert (a | ~prev).all(), tol
        prev = a


def test_dac_integration_properties(gri, flame):
    """Removed species do not react, mass is conserved, the temperature comes back unchanged (reference aliasing quirk,
    LoadBalancedTDACChemistryModel.C:406-409), and the result converges to the full mechanism as the tolerance shrinks."""
    sel = np.argsort(-flame["T"])[::25][:128]
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    init = _init(gri)
    full = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    scale = np.abs(full["RR"]).max(0) + 1e-30
    errs = []
    for tol in (1e-2, 1e-4, 1e-8):
        om = O.OracleModel(gri, opts)
        om.set_reduction(tol, init)
        r = om.solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
        assert (r["status"] == 0).all()
        assert np.array_equal(r["T"], T)
        # molWeights are rounded to ~6 digits (1e-5 relative); cold cells move ~1e-10 kg/m3/s in total, i.e. absTol noise
        assert (np.abs(r["RR"].sum(0)) <= 1e-4 * np.abs(r["RR"]).sum(0) + 1e-8).all()
        for k in range(0, sel.size, 16):
            act, _ = om.dac_reduce(rho[k] * Y[:, k] / gri.W, T[k], tol, init)
            assert r["n_active"][k] == act.sum()
            assert (r["RR"][~act, k] == 0).all()
        errs.append(np.median(np.abs(r["RR"] - full["RR"]).max(0) / scale))
    assert errs[0] > errs[1] > errs[2], errs
    assert errs[2] < 1e-6 and errs[1] < 1e-3, errs
    # the full model is untouched by a reduction that was set and cleared
    om.set_reduction(None)
    again = om.solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    assert np.array_equal(again["RR"], full["RR"]) and np.array_equal(again["T"], full["T"])


def test_dac_regression_vector(gri, flame):
    """Oracle outputs committed with the script that made them (tests/golden/make_dac_regression.py)."""
    z = np.load(ROOT / "tests/golden/dac_regression.npz")
    sel = z["cells"]
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    om = O.OracleModel(gri, opts)
    om.set_reduction(TOL, _init(gri))
    r = om.solve_batch(Y, T, p, rho, DT, DTCHEM)
    assert np.array_equal(r["n_active"], z["n_active"])
    assert np.array_equal(r["attempts"], z["attempts"])
    assert np.allclose(r["RR"], z["RR"], rtol=1e-9, atol=1e-12 * np.abs(z["RR"]).max())
    assert np.allclose(r["deltaTChem"], z["deltaTChem"], rtol=1e-9)
    for k in range(0, sel.size, 8):
        act, _ = om.dac_reduce(rho[k] * Y[:, k] / gri.W, T[k], TOL, _init(gri))
        assert np.array_equal(act, z["active"][:, k])
