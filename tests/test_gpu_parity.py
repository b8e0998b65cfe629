"""GPU parity: the CUDA path through the C ABI against the CPU oracle on the same inputs.

Mirrors the reference's own acceptance test (tests/src/standardChemistryModel-Test.C:84-100: one
solve(1e-6) on the GRI flame case, RR compared per species per cell at rel 1e-6).
Tolerances (BASELINE.json north_star): species mass fractions and temperature after the flow step at
max relative error 1e-6; RR within the same bound relative to the cell's state change scale.
"""
import numpy as np
import pytest

from loadbalancedchemistrymodel_b200 import capi
from oracle import oracle as O

pytestmark = pytest.mark.gpu

DT = 1e-6        # tests/src/standardChemistryModel-Test.C:84
DTCHEM = 1e-7    # initialChemicalTimeStep, tests/Cases/Case-chemistry/constant/chemistryProperties:26
RTOL = 1e-6


def _state(mech, flame, sel):
    T, p, Y = flame["T"][sel], flame["p"][sel], np.ascontiguousarray(flame["Y"][:, sel])
    return Y, T, p, O.density(mech, Y, T, p)


def _compare(mech, Y, rho, got, ref, dt, tag, abs_tol=1e-12):
    """rel 1e-6 on Y, T and RR. The absolute floor is 1e-6 x the integrator's own absTol (a concentration,
    kmol/m3) expressed in each quantity's units: species far below absTol (Y ~ 1e-26 radicals in cold cells)
    are round-off of both implementations, not results."""
    assert (got["status"] == ref["status"]).all(), f"{tag}: status differs"
    W = mech.W[:, None]
    Yg = Y + got["RR"] * dt / rho
    Yr = Y + ref["RR"] * dt / rho
    eY = np.abs(Yg - Yr) / (np.abs(Yr) + abs_tol * W / rho)
    eT = np.abs(got["T"] - ref["T"]) / ref["T"]
    eR = np.abs(got["RR"] - ref["RR"]) / (np.abs(ref["RR"]) + abs_tol * W / dt)
    same = got["cost"] == ref["attempts"]
    eD = np.abs(got["deltaTChem"] - ref["deltaTChem"])[same] / ref["deltaTChem"][same]
    print(f"{tag}: max rel err Y {eY.max():.2e} T {eT.max():.2e} RR {eR.max():.2e} dtChem {eD.max():.2e} "
          f"cells with a different attempt count {int((~same).sum())}/{same.size}")
    assert eY.max() < RTOL and eT.max() < RTOL and eR.max() < RTOL and eD.max() < 1e-5
    # The last sub-step is dx = xEnd - x; whether x + dx == xEnd exactly is a round-off coin toss of the
    # reference algorithm itself (one more ~1e-22 s step, other deltaTChem, same state): allow 1 % of cells.
    assert (~same).mean() <= 0.01
    return eY.max(), eT.max(), eR.max()


def test_probes_match_oracle(gri, flame):
    """derivatives / jacobian / rate constants / LU of the device code against the oracle, state by state."""
    sel = np.argsort(-flame["T"])[::40][:64]
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="Rosenbrock34", abs_tol=1e-12, rel_tol=1e-2)
    chem = capi.Chemistry(gri, opts)
    om = O.OracleModel(gri, opts)
    n = gri.n_species + 2
    y = np.vstack([rho * Y / gri.W[:, None], T[None], p[None]]).T.copy()   # [items, n]
    kfkr = chem.probe(capi.PROBE_RATES, y)
    d = chem.probe(capi.PROBE_DERIV, y)
    j = chem.probe(capi.PROBE_JAC, y)
    for k in range(y.shape[0]):
        kf, kr = om.rates(y[k, :-2], y[k, -2])
        assert np.allclose(kfkr[k, :gri.n_reactions], kf, rtol=1e-12, atol=0), k
        assert np.allclose(kfkr[k, gri.n_reactions:], kr, rtol=1e-12, atol=0), k
        dr = om.derivatives(y[k])
        assert np.allclose(d[k], dr, rtol=1e-9, atol=1e-12 * np.abs(dr).max()), k
        dfdx, J = om.jacobian(y[k])
        assert np.allclose(j[k, :n], dfdx, rtol=1e-9, atol=1e-12 * np.abs(dfdx).max()), k
        Jg = j[k, n:].reshape(n, n)
        # analytic block tight; finite-difference temperature column carries 1e-16/1e-3 noise
        assert np.allclose(Jg[:, :-2], J[:, :-2], rtol=1e-9, atol=1e-12 * np.abs(J).max()), k
        assert np.allclose(Jg[:, -2], J[:, -2], rtol=1e-6, atol=1e-9 * np.abs(J[:, -2]).max()), k
    # LU solve of I/(gamma dx) - J
    items = []
    for k in range(8):
        _, J = om.jacobian(y[k])
        A = -J + np.eye(n) / (0.5 * 1e-7)
        b = om.derivatives(y[k])
        items.append(np.concatenate([A.ravel(), b]))
    x = chem.probe(capi.PROBE_LUSOLVE, np.array(items))
    for k in range(8):
        A = items[k][: n * n].reshape(n, n); b = items[k][n * n:]
        st, xr = O.lu_solve(A, b)
        assert st == 0
        assert np.allclose(x[k], xr, rtol=1e-9, atol=1e-13 * np.abs(xr).max()), k
    chem.close()


@pytest.mark.parametrize("kernel", ["v3", "v2", "general"])
@pytest.mark.parametrize("solver,rel_tol", [("Rosenbrock34", 1e-2), ("seulex", 1e-1)])
def test_gri_flame_fixture_parity(gri, flame, solver, rel_tol, kernel, monkeypatch):
    """All 4000 cells of the reference's flame fixture, both tolerance sets of BASELINE config 3, through the
    compact-code kernel (v3, default), the register-LU kernel (LBCHEM_KERNEL=v2) and the general kernel
    (threads_per_cell = 64)."""
    tpc = 64 if kernel == "general" else 0
    if kernel == "v2":
        monkeypatch.setenv("LBCHEM_KERNEL", "v2")
    else:
        monkeypatch.delenv("LBCHEM_KERNEL", raising=False)
    sel = np.arange(flame["T"].shape[0])
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol, threads_per_cell=tpc)
    chem = capi.Chemistry(gri, opts)
    got = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
    st = chem.stats()
    ref = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    cnt = ref["counters"]
    print(solver, "gpu stats", st, "oracle", cnt)
    _compare(gri, Y, rho, got, ref, DT, f"gri/{solver}")
    for k in ("n_steps", "n_attempts", "n_lu", "n_solve", "n_rhs", "n_jac"):
        assert abs(st[k] - cnt[k]) <= 0.002 * cnt[k], k
    chem.close()
