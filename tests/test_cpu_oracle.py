"""What constrains the CPU oracle, given that parity with OpenFOAM is UNPINNED.

The reference ships no golden vectors for this path (tests/src/standardChemistryModel-Test.C:90-102 compares two
OpenFOAM-backed runs), so the oracle is held by physics and by independent numerics instead:
conservation laws, an independent stiff integrator (SciPy Radau) on the same right-hand side, the fixture's own
Qdot field, invariance under partitioning (the reference's LB == non-LB criterion), linear algebra against
LAPACK, the analytic Jacobian against finite differences, and a committed regression vector of its own outputs.
"""
import re
from pathlib import Path

import numpy as np
import pytest

from loadbalancedchemistrymodel_b200 import capi
from oracle import oracle as O

ROOT = Path(__file__).resolve().parents[1]
DT, DTCHEM = 1e-6, 1e-7


def composition(species):
    """Element counts from the species names of the GRI mechanism (C, H, O, N, Ar)."""
    elems = ["C", "H", "O", "N", "Ar"]
    A = np.zeros((len(species), len(elems)))
    for i, s in enumerate(species):
        name = s.replace("(S)", "")
        if name == "AR":
            A[i, 4] = 1
            continue
        for el, cnt in re.findall(r"([CHON])(\d*)", name):
            A[i, elems.index(el)] += int(cnt) if cnt else 1
    return elems, A


def hot_cells(flame, k, stride=1):
    return np.argsort(-flame["T"])[::stride][:k]


def state(mech, flame, sel):
    T, p, Y = flame["T"][sel], flame["p"][sel], np.ascontiguousarray(flame["Y"][:, sel])
    return Y, T, p, O.density(mech, Y, T, p)


def test_species_names_give_the_molecular_weights(gri):
    elems, A = composition(gri.species)
    aw = np.array([12.011, 1.00794, 15.9994, 14.0067, 39.948])
    assert np.allclose(A @ aw, gri.W, rtol=2e-4), "formula parser disagrees with the thermo file's molWeight"


@pytest.mark.parametrize("solver,rel_tol", [("Rosenbrock34", 1e-2), ("seulex", 1e-1)])
def test_mass_and_element_conservation(gri, flame, solver, rel_tol):
    sel = hot_cells(flame, 160, stride=8)
    Y, T, p, rho = state(gri, flame, sel)
    opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol)
    r = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    assert (r["status"] == 0).all()
    RR = r["RR"]
    scale = np.abs(RR).sum(0) + 1e-30
    # the thermo file's molWeight entries are rounded to ~6 digits, so reactions balance mass to ~1e-5 only;
    # element balances (in kmol) are free of W and must close to round-off
    assert (np.abs(RR.sum(0)) / scale).max() < 1e-4, "sum_i RR_i must vanish (mass conservation)"
    _, A = composition(gri.species)
    el = (RR / gri.W[:, None]).T @ A                      # kmol of each element per m3 per s
    # every reaction conserves elements exactly; the only leak is the reference's clamp c = max(c, 0) after each
    # ODE solve (LoadBalancedChemistryModel.C:400-402), which acts on species at the absTol (1e-12 kmol/m3) level
    assert np.abs(el).max() * DT < 1e-11, "element drift over the step must stay at the absTol level"
    ctot = (rho * Y / gri.W[:, None]).sum(0)
    assert (np.abs(el).max(1) * DT / ctot).max() < 1e-9
    assert np.isfinite(r["T"]).all() and (r["deltaTChem"] > 0).all()


def test_partitioning_does_not_change_results(gri, flame):
    """The reference's acceptance criterion (LB == non-LB to 1e-6): cells are independent, so any partition of
    the batch over workers must give bitwise identical results."""
    sel = hot_cells(flame, 64, stride=20)
    Y, T, p, rho = state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    a = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=1)
    b = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=7)
    perm = np.random.default_rng(0).permutation(sel.size)
    c = O.OracleModel(gri, opts).solve_batch(Y[:, perm], T[perm], p[perm], rho[perm], DT, DTCHEM, nthreads=3)
    for k in ("RR", "T", "deltaTChem", "attempts"):
        assert np.array_equal(a[k], b[k]), k
        assert np.array_equal(a[k][..., perm], c[k]), k


def test_integrators_converge_to_scipy_radau(gri, flame):
    """Same RHS (oracle derivatives), independent integrator: the oracle's Rosenbrock34 and seulex approach the
    Radau IIA solution as relTol tightens."""
    from scipy.integrate import solve_ivp
    sel = hot_cells(flame, 3, stride=150)
    Y, T, p, rho = state(gri, flame, sel)
    base = capi.default_options(solver="Rosenbrock34", abs_tol=1e-14, rel_tol=1e-8)
    om = O.OracleModel(gri, base)
    for c in range(sel.size):
        y0 = np.concatenate([rho[c] * Y[:, c] / gri.W, [T[c], p[c]]])
        sol = solve_ivp(lambda t, y: om.derivatives(np.maximum(y, 0) * (np.arange(y.size) < gri.n_species) + y * (np.arange(y.size) >= gri.n_species)),
                        (0.0, DT), y0, method="Radau", rtol=1e-10, atol=1e-16)
        assert sol.success
        yr = sol.y[:, -1]
        errs = {}
        for solver in ("Rosenbrock34", "seulex"):
            for rt in (1e-2, 1e-4, 1e-7):
                o = capi.default_options(solver=solver, abs_tol=1e-14, rel_tol=rt)
                r = O.OracleModel(gri, o).solve_batch(Y[:, c:c + 1], T[c:c + 1], p[c:c + 1], rho[c:c + 1], DT, DTCHEM)
                cend = rho[c] * (Y[:, c] + r["RR"][:, 0] * DT / rho[c]) / gri.W
                big = yr[:-2] > 1e-9 * yr[:-2].max()
                e = max(np.abs(cend[big] - yr[:-2][big]).max() / yr[:-2].max(), abs(r["T"][0] - yr[-2]) / yr[-2])
                errs[(solver, rt)] = e
            assert errs[(solver, 1e-7)] < 1e-6, errs      # the temperature row of OpenFOAM's Jacobian is zero: reduced order
            assert errs[(solver, 1e-7)] <= errs[(solver, 1e-2)] + 1e-12, errs
            assert errs[(solver, 1e-2)] < 1e-3, errs


def test_heat_release_matches_the_fixture_qdot_field(gri, flame):
    """Qdot = -sum_i hf_i RR_i from the oracle's RR against the Qdot field the reference's case stores
    (produced by a previous step with ISAT tol 1e-4: sign / location / magnitude only)."""
    n = flame["T"].shape[0]
    sel = np.arange(0, n, 5)
    Y, T, p, rho = state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    r = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    Tstd = 298.15
    a = np.where((Tstd < gri.Tcommon)[:, None], gri.janaf_low, gri.janaf_high)
    hf = capi.RR_DEFAULT * (((((a[:, 4] / 5 * Tstd + a[:, 3] / 4) * Tstd + a[:, 2] / 3) * Tstd + a[:, 1] / 2) * Tstd + a[:, 0]) * Tstd + a[:, 5]) / gri.W
    Q = -(hf[:, None] * r["RR"]).sum(0)
    Qref = flame["Qdot"][sel]
    strong = np.abs(Qref) > 1e6
    assert strong.sum() > 50
    # The stored field came from ISAT retrievals (tolerance 1e-4 on a 1e-6 s step: the tabulation error is of the
    # order of the state change itself), so only sign, ranking and order of magnitude can be compared.
    from scipy.stats import spearmanr
    assert (np.sign(Q[strong]) == np.sign(Qref[strong])).mean() > 0.98
    assert spearmanr(Q[strong], Qref[strong]).statistic > 0.8
    ratio = np.abs(Q[strong]).sum() / np.abs(Qref[strong]).sum()
    assert 0.05 < ratio < 20.0, ratio
    cold = flame["T"][sel] < 400.0
    assert np.abs(Q[cold]).max() < 1e3, "no heat release in the cold inflow"


def test_lu_matches_lapack(gri, flame):
    rng = np.random.default_rng(3)
    for n in (1, 2, 7, 12, 55):
        A = rng.standard_normal((n, n)) + n * np.eye(n) * rng.choice([0.0, 1.0])
        b = rng.standard_normal(n)
        st, x = O.lu_solve(A, b)
        assert st == 0
        assert np.allclose(x, np.linalg.solve(A, b), rtol=1e-9, atol=1e-12)
    st, _ = O.lu_solve(np.zeros((3, 3)), np.ones(3))
    assert st == capi.CELL_SINGULAR


def only_reactions(m, keep):
    """Copy of mechanism ``m`` restricted to the reactions in ``keep`` (indices)."""
    import dataclasses

    def side(ptr, *arrs):
        newp, outs = [0], [[] for _ in arrs]
        for r in keep:
            for k in range(ptr[r], ptr[r + 1]):
                for o, a in zip(outs, arrs):
                    o.append(a[k])
            newp.append(len(outs[0]))
        return [np.array(newp, np.int32)] + [np.array(o, a.dtype) for o, a in zip(outs, arrs)]
    lp, ls, ln, le = side(m.lhs_ptr, m.lhs_species, m.lhs_nu, m.lhs_exp)
    rp, rs, rn, re_ = side(m.rhs_ptr, m.rhs_species, m.rhs_nu, m.rhs_exp)
    return dataclasses.replace(
        m, rtype=m.rtype[keep].copy(), reversible=m.reversible[keep].copy(), arr=m.arr[keep].copy(), arr0=m.arr0[keep].copy(),
        troe=m.troe[keep].copy(), eff_index=m.eff_index[keep].copy(), lhs_ptr=lp, lhs_species=ls, lhs_nu=ln, lhs_exp=le,
        rhs_ptr=rp, rhs_species=rs, rhs_nu=rn, rhs_exp=re_, names=[m.names[r] for r in keep])


def test_analytic_jacobian_against_finite_differences(gri, flame):
    """OpenFOAM's jacobian() differentiates the mass-action products only (kf, kr and the third-body
    concentration M are held constant), so the comparison with finite differences is exact only for the
    plain Arrhenius reactions: 284 of the 325."""
    plain = only_reactions(gri, np.where(gri.rtype == 0)[0])
    assert plain.n_reactions == 284
    sel = hot_cells(flame, 2, stride=300)
    Y, T, p, rho = state(gri, flame, sel)
    om = O.OracleModel(plain, capi.default_options(solver="Rosenbrock34"))
    nSp = gri.n_species
    for c in range(sel.size):
        y = np.concatenate([rho[c] * Y[:, c] / gri.W, [T[c], p[c]]])
        dfdx, J = om.jacobian(y)
        f0 = om.derivatives(y)
        assert np.allclose(dfdx[:nSp], f0[:nSp], rtol=1e-12, atol=0), "OpenFOAM quirk: dfdx slot carries dcdt (SURVEY A.4)"
        assert np.all(J[nSp:, :] == 0.0), "T and p rows of the Jacobian are zero in OpenFOAM v2306"
        cols = np.argsort(-y[:nSp])[:12]
        for j in cols:
            h = 1e-6 * y[j]
            yp, ym = y.copy(), y.copy()
            yp[j] += h; ym[j] -= h
            fd = (om.derivatives(yp) - om.derivatives(ym))[:nSp] / (2 * h)
            assert np.allclose(J[:nSp, j], fd, rtol=2e-5, atol=1e-7 * np.abs(fd).max()), (c, j)
        # temperature column: OpenFOAM itself uses a central difference with delta = 1e-3 K
        yp, ym = y.copy(), y.copy()
        yp[nSp] += 1e-3; ym[nSp] -= 1e-3
        fd = (om.derivatives(yp) - om.derivatives(ym))[:nSp] / 2e-3
        assert np.allclose(J[:nSp, nSp], fd, rtol=1e-6, atol=1e-9 * np.abs(fd).max())


def test_oracle_regression_vector(gri, flame):
    """Outputs of the oracle for 32 fixture cells, committed with the script that made them
    (tests/golden/make_oracle_regression.py). Guards the oracle against drift; it does NOT pin it to OpenFOAM."""
    z = np.load(ROOT / "tests/golden/oracle_regression.npz")
    sel = z["cells"]
    Y, T, p, rho = state(gri, flame, sel)
    for solver, rel_tol in (("Rosenbrock34", 1e-2), ("seulex", 1e-1)):
        opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol)
        r = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=4)
        assert np.array_equal(r["attempts"], z[f"{solver}_attempts"])
        assert np.allclose(r["T"], z[f"{solver}_T"], rtol=1e-12, atol=0)
        assert np.allclose(r["deltaTChem"], z[f"{solver}_deltaTChem"], rtol=1e-9, atol=0)
        ref = z[f"{solver}_RR"]
        assert np.allclose(r["RR"], ref, rtol=1e-8, atol=1e-12 * np.abs(ref).max())


def test_last_substep_coin_toss_belongs_to_the_algorithm(gri):
    """ODESolver::solve ends with dx = xEnd - x; whether x + dx then equals xEnd to the last bit decides if one more
    ~1e-22 s sub-step is taken. That is round-off of the reference algorithm, not of an implementation: the
    oracle ITSELF changes the attempt count of about 1 % of hot cells when deltaT, T or deltaTChem move by one ulp, the
    state (RR) stays put and the returned sub-step guess jumps by the same factors (~0.26 / ~3.8) the GPU parity test
    sees in its 'cells with a different attempt count' (tests/test_gpu_parity.py::_compare bounds them with this evidence)."""
    from loadbalancedchemistrymodel_b200 import capi, workloads as W
    mech, Y, T, p, rho = W.gri_uniform(1000)
    om = O.OracleModel(mech, capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1))
    r0 = om.solve_batch(Y, T, p, rho, 1e-6, 1e-7, nthreads=8)
    flipped, ratios = 0, []
    for args in ((Y, T, p, rho, np.nextafter(1e-6, 1), 1e-7), (Y, np.nextafter(T, 1e9), p, rho, 1e-6, 1e-7),
                 (Y, T, p, rho, 1e-6, np.nextafter(1e-7, 1))):
        r1 = om.solve_batch(*args, nthreads=8)
        d = r1["attempts"] != r0["attempts"]
        flipped += int(d.sum())
        ratios += list(r1["deltaTChem"][d] / r0["deltaTChem"][d])
        assert (np.abs(r1["attempts"] - r0["attempts"])[d] == 1).all()
        eRR = np.abs(r1["RR"] - r0["RR"]) / (np.abs(r0["RR"]) + 1e-12 * mech.W[:, None] / 1e-6)
        assert eRR.max() < 1e-8                                    # same state in every cell, flipped or not
        same = ~d
        assert (np.abs(r1["deltaTChem"][same] / r0["deltaTChem"][same] - 1)).max() < 1e-9
    ratios = np.array(ratios)
    print("cells flipped by one-ulp input changes:", flipped, "of", 3 * 1000, "deltaTChem ratios", np.round(np.sort(ratios), 3))
    assert flipped >= 10 and ratios.min() > 0.1 and ratios.max() < 10.0
