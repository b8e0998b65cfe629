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


def _compare(mech, Y, rho, got, ref, dt, tag, abs_tol=1e-12, max_other_count=0.01, dtchem_tol=1e-5, rr_noise_ulps=0):
    """rel 1e-6 on Y, T and RR. The absolute floor is 1e-6 x the integrator's own absTol (a concentration,
    kmol/m3) expressed in each quantity's units: species far below absTol (Y ~ 1e-26 radicals in cold cells)
    are round-off of both implementations, not results."""
    assert (got["status"] == ref["status"]).all(), f"{tag}: status differs"
    W = mech.W[:, None]
    Yg = Y + got["RR"] * dt / rho
    Yr = Y + ref["RR"] * dt / rho
    eY = np.abs(Yg - Yr) / (np.abs(Yr) + abs_tol * W / rho)
    eT = np.abs(got["T"] - ref["T"]) / ref["T"]
    # RR = (c - c0) W / deltaT carries the representation noise of c itself: `rr_noise_ulps` ulps of c0 (accumulated over the
    # sub-steps of a long integration) are not a difference of results, only what goes beyond them is held to 1e-6
    noise = rr_noise_ulps * np.finfo(float).eps * (rho * Y / W) * W / np.asarray(dt)
    eR = np.maximum(np.abs(got["RR"] - ref["RR"]) - noise, 0.0) / (np.abs(ref["RR"]) + abs_tol * W / np.asarray(dt))
    same = got["cost"] == ref["attempts"]
    eD = np.abs(got["deltaTChem"] - ref["deltaTChem"])[same] / ref["deltaTChem"][same]
    other = ~same
    ratio = got["deltaTChem"][other] / ref["deltaTChem"][other]
    dA = (got["cost"] - ref["attempts"])[other]
    print(f"{tag}: max rel err Y {eY.max():.2e} T {eT.max():.2e} RR {eR.max():.2e} dtChem {eD.max():.2e} "
          f"cells with a different attempt count {int(other.sum())}/{same.size}"
          + (f" (attempt difference {np.unique(dA)}, deltaTChem ratio there {ratio.min():.3f} .. {ratio.max():.3f})" if other.any() else ""))
    assert eY.max() < RTOL and eT.max() < RTOL and eR.max() < RTOL and eD.max() < dtchem_tol
    # The last sub-step is dx = xEnd - x; whether x + dx lands on xEnd exactly is a round-off coin toss of the reference
    # algorithm itself: the oracle flips the same way in 1.1-1.5 % of these cells when deltaT, T or deltaTChem move by ONE
    # ulp (tests/test_cpu_oracle.py::test_last_substep_coin_toss_belongs_to_the_algorithm). A flipped cell takes one
    # ~1e-22 s sub-step more or less: same state (Y, T, RR are checked above for ALL cells), and the returned sub-step
    # guess is the proposal of the other branch of ODESolver::solve, a factor 0.26 or 3.8 away in the oracle's
    # own flips. Bounded here, not masked:
    assert (~same).mean() <= max_other_count
    if other.any():
        assert (np.abs(dA) == 1).all() and ratio.min() > 0.1 and ratio.max() < 10.0
    return eY.max(), eT.max(), eR.max()


@pytest.mark.parametrize("kernel", ["v3", "general"])
def test_probes_match_oracle(gri, flame, kernel):
    """derivatives / jacobian / rate constants / LU of the device code against the oracle, state by state: the device
    functions of the PRODUCTION compact-code kernel (v3: prepare, reaction_sweep, derivatives, jacobian, lu_factor,
    lu_solve) and, as a second column, those of the general kernel (threads_per_cell = 64)."""
    sel = np.argsort(-flame["T"])[::40][:64]
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="Rosenbrock34", abs_tol=1e-12, rel_tol=1e-2, threads_per_cell=0 if kernel == "v3" else 64)
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
    dg = 1.0 / (0.5 * 1e-7)
    if kernel == "general":
        items = []
        for k in range(8):
            _, J = om.jacobian(y[k])
            A = -J + np.eye(n) * dg
            b = om.derivatives(y[k])
            items.append(np.concatenate([A.ravel(), b]))
        x = chem.probe(capi.PROBE_LUSOLVE, np.array(items))
        for k in range(8):
            A = items[k][: n * n].reshape(n, n); b = items[k][n * n:]
            st, xr = O.lu_solve(A, b)
            assert st == 0
            assert np.allclose(x[k], xr, rtol=1e-9, atol=1e-13 * np.abs(xr).max()), k
    else:
        # the compact-code kernel factorises what its own jacobian() left in the sparse store: probe the chain
        # jacobian -> lu_factor -> lu_solve against LUDecompose / LUBacksubstitute of the oracle on dg I - J_oracle
        items, want = [], []
        for k in range(16):
            _, J = om.jacobian(y[k])
            b = om.derivatives(y[k])
            items.append(np.concatenate([y[k], [dg], b]))
            st, xr = O.lu_solve(-J + np.eye(n) * dg, b)
            assert st == 0
            want.append(xr)
        x = chem.probe(capi.PROBE_JAC_LUSOLVE, np.array(items))
        for k in range(16):
            assert np.allclose(x[k], want[k], rtol=1e-8, atol=1e-12 * np.abs(want[k]).max()), k
    chem.close()


@pytest.mark.parametrize("kernel", ["v3", "general"])
@pytest.mark.parametrize("solver,rel_tol", [("Rosenbrock34", 1e-2), ("seulex", 1e-1)])
def test_gri_flame_fixture_parity(gri, flame, solver, rel_tol, kernel, monkeypatch):
    """All 4000 cells of the reference's flame fixture, both tolerance sets of BASELINE config 3, through the
    compact-code kernel (v3, default) and the general kernel (threads_per_cell = 64)."""
    tpc = 64 if kernel == "general" else 0
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


@pytest.mark.parametrize("solver,rel_tol", [("Rosenbrock34", 1e-2), ("seulex", 1e-1)])
def test_pivoting_rules(gri, flame, solver, rel_tol):
    """lbchem_options.pivoting. The parity tests above run the default (1: diagonal pivots, growth test, the reference's
    rule as fallback). Here: (a) rule 0 -- LUDecompose's own pivot sequence on every matrix -- against the oracle on all
    4000 cells at the same 1e-6; (b) the fallback of rule 1 (hook 2 rejects every diagonal factorisation) reproduces
    rule 0 bit for bit; (c) rule 1 stays within 1e-6 of rule 0 in every output field, DAC-reduced systems included."""
    sel = np.arange(flame["T"].shape[0])
    Y, T, p, rho = _state(gri, flame, sel)
    out = {}
    for piv in (0, 1, 2):
        opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol, pivoting=piv)
        chem = capi.Chemistry(gri, opts)
        out[piv] = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
        out[piv, "lu"] = chem.stats()["n_lu"]
        chem.close()
    ref = O.OracleModel(gri, capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol)).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    _compare(gri, Y, rho, out[0], ref, DT, f"gri/{solver}/pivoting=0")
    for k in ("RR", "T", "deltaTChem", "cost", "status"):
        assert np.array_equal(out[0][k], out[2][k]), f"fallback differs from rule 0 in {k}"
    assert out[0, "lu"] == out[2, "lu"]                       # a rejected diagonal attempt is not an algorithmic LU
    a, b = out[0], out[1]
    W = gri.W[:, None]
    eR = (np.abs(a["RR"] - b["RR"]) / (np.abs(a["RR"]) + 1e-12 * W / DT)).max()
    eT = (np.abs(a["T"] - b["T"]) / a["T"]).max()
    print(f"{solver}: pivoting 1 vs 0: RR {eR:.2e} T {eT:.2e}, cells with another attempt count {int((a['cost'] != b['cost']).sum())}")
    assert eR < RTOL and eT < RTOL and (a["status"] == b["status"]).all()
    assert (a["cost"] != b["cost"]).mean() <= 0.01
    # DAC-reduced systems (compacted rows and columns): rule 1 against rule 0
    init = [list(gri.species).index(s) for s in ("CO", "CH4", "HO2")]
    dac = {}
    for piv in (0, 1):
        chem = capi.Chemistry(gri, capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol, pivoting=piv))
        chem.set_reduction(1e-4, init)
        dac[piv] = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
        chem.close()
    eR = (np.abs(dac[0]["RR"] - dac[1]["RR"]) / (np.abs(dac[0]["RR"]) + 1e-12 * W / DT)).max()
    print(f"{solver}: DAC, pivoting 1 vs 0: RR {eR:.2e}")
    assert eR < RTOL and (dac[0]["status"] == dac[1]["status"]).all()


@pytest.mark.parametrize("workload", ["h2_air", "ch4_mixing_line"])
def test_pivoting_rules_on_sub_warp_groups(workload):
    """The same three rules where several cells share one warp (16 threads per cell for H2/O2, 8 for the 1-step CH4
    mechanism): rule 0 against the oracle, the fallback of rule 1 bit-identical to rule 0, rule 1 within 1e-6 of rule 0."""
    from loadbalancedchemistrymodel_b200 import workloads as W
    mech, Y, T, p, rho = getattr(W, workload)(2000)
    out = {}
    for piv in (0, 1, 2):
        opts = capi.default_options(solver="Rosenbrock34", abs_tol=1e-12, rel_tol=1e-2, pivoting=piv)
        chem = capi.Chemistry(mech, opts)
        out[piv] = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
        chem.close()
    ref = O.OracleModel(mech, capi.default_options(solver="Rosenbrock34", abs_tol=1e-12, rel_tol=1e-2)).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    _compare(mech, Y, rho, out[0], ref, DT, f"{workload}/pivoting=0")
    for k in ("RR", "T", "deltaTChem", "cost", "status"):
        assert np.array_equal(out[0][k], out[2][k]), f"fallback differs from rule 0 in {k}"
    W_ = mech.W[:, None]
    eR = (np.abs(out[0]["RR"] - out[1]["RR"]) / (np.abs(out[0]["RR"]) + 1e-12 * W_ / DT)).max()
    print(f"{workload}: pivoting 1 vs 0: RR {eR:.2e}")
    assert eR < RTOL and (out[0]["status"] == out[1]["status"]).all()


# ---- the other BASELINE configurations (parity-test cases, not bench lines) -------------------------
def _solve_both(mech, Y, T, p, rho, opts, dt=DT, dtchem=DTCHEM):
    chem = capi.Chemistry(mech, opts)
    got = chem.solve_host(Y, T, p, rho, dt, dtchem)
    st = chem.stats()
    chem.close()
    ref = O.OracleModel(mech, opts).solve_batch(Y, T, p, rho, dt, dtchem, nthreads=8)
    return got, ref, st


def test_h2_air_parity():
    """BASELINE config 2: H2/O2 sub-mechanism of GRI-3.0 (10 species, 29 reactions), Rosenbrock34 1e-12/1e-2."""
    from loadbalancedchemistrymodel_b200 import workloads as W
    mech, Y, T, p, rho = W.h2_air(3000)
    assert mech.n_species == 10 and mech.n_reactions == 29
    opts = capi.default_options(solver="Rosenbrock34", abs_tol=1e-12, rel_tol=1e-2)
    got, ref, st = _solve_both(mech, Y, T, p, rho, opts)
    _compare(mech, Y, rho, got, ref, DT, "h2/Rosenbrock34")
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    got, ref, st = _solve_both(mech, Y, T, p, rho, opts)
    _compare(mech, Y, rho, got, ref, DT, "h2/seulex")


@pytest.mark.parametrize("solver", ["EulerImplicit", "Rosenbrock34"])
def test_ch4_one_step_parity(solver):
    """BASELINE config 1: the tutorial's 1-step CH4 mechanism (5 species), EulerImplicit(cTauChem 1, limiter off) as
    the tutorial selects (tutorial/counterFlowFlame2D/constant/chemistryProperties:19-31) and its unused odeCoeffs."""
    from loadbalancedchemistrymodel_b200 import workloads as W
    mech, Y, T, p, rho = W.ch4_mixing_line(4000)
    opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=1e-2, cTauChem=1.0, eq_rate_limiter=0)
    got, ref, st = _solve_both(mech, Y, T, p, rho, opts)
    same = got["cost"] == ref["attempts"]
    print(solver, "cells with a different sub-step count", int((~same).sum()))
    W_ = mech.W[:, None]
    Yg = Y + got["RR"] * DT / rho
    Yr = Y + ref["RR"] * DT / rho
    assert (got["status"] == ref["status"]).all() and same.all()
    assert (np.abs(Yg - Yr) / (np.abs(Yr) + 1e-12 * W_ / rho)).max() < RTOL
    assert (np.abs(got["T"] - ref["T"]) / ref["T"]).max() < RTOL
    # RR = (c - c0) W / deltaT: where nothing reacts (cold cells) c - c0 is the round-off of c itself, eps*|c0|*W/deltaT in
    # RR units, which no implementation of the reference formula can beat -- the ORACLE's own RR moves by 6.9e-6 of the 1e-6
    # floor when rho changes by one ulp (shown below). The bound is therefore 1e-6 relative (north_star) plus that
    # representation noise, nothing wider.
    eps = np.finfo(float).eps
    c0 = rho * Y / W_
    noise = 8 * eps * c0 * W_ / DT
    scale = np.abs(ref["RR"]).max(0, keepdims=True)
    excess = np.abs(got["RR"] - ref["RR"]) - noise
    eRR = (np.maximum(excess, 0.0) / (scale + 1e-12 * W_ / DT)).max()
    print(solver, "max RR error beyond the representation noise, relative to the cell's rate scale", eRR)
    assert eRR < RTOL
    eD = np.abs(got["deltaTChem"] - ref["deltaTChem"]) / ref["deltaTChem"]
    if solver != "EulerImplicit":
        assert eD.max() < 1e-5
        return
    # EulerImplicit returns deltaTChem = cTauChem * tMin with tMin ~ c_O2 / (2 pf c_CH4): in hot lean cells the algorithm
    # burns CH4 down to ~1e-15 kmol/m3, where that quotient amplifies the last bits of the 5x5 solve. How much is a
    # property of the reference algorithm: measure it on the oracle itself (inputs moved by ONE ulp) and hold the GPU to it.
    om = O.OracleModel(mech, opts)
    self_d, self_rr = 0.0, 0.0
    for args in ((Y, T, p, np.nextafter(rho, 1e9)), (np.nextafter(Y, 2.0), T, p, rho), (Y, np.nextafter(T, 1e9), p, rho)):
        r1 = om.solve_batch(*args, DT, DTCHEM, nthreads=8)
        self_d = max(self_d, (np.abs(r1["deltaTChem"] - ref["deltaTChem"]) / ref["deltaTChem"]).max())
        self_rr = max(self_rr, (np.abs(r1["RR"] - ref["RR"]) / (scale + 1e-12 * W_ / DT)).max())
    print(f"EulerImplicit: oracle moves under one-ulp input changes by deltaTChem {self_d:.2e}, RR {self_rr:.2e} of the floor; "
          f"GPU vs oracle deltaTChem {eD.max():.2e}")
    assert self_rr > 1e-6                                  # the oracle itself leaves the naive 1e-6 band in cold cells
    assert eD.max() < 10 * self_d and np.median(eD) < 1e-12 and (eD > 1e-9).mean() < 0.03
    one = capi.Chemistry(mech, opts)
    g1 = one.solve_host(Y, T, p, rho, 1e-9, DTCHEM)
    one.close()
    r1 = O.OracleModel(mech, opts).solve_batch(Y, T, p, rho, 1e-9, DTCHEM, nthreads=8)
    assert (np.abs(g1["deltaTChem"] - r1["deltaTChem"]) / r1["deltaTChem"]).max() < 1e-10


def test_edge_cases(gri, flame):
    """Empty batch, Treact gate (LoadBalancedChemistryModel.C:371,413-419), per-cell deltaT (LTS), ragged batch sizes."""
    sel = np.argsort(-flame["T"])[::13][:301]          # odd size: last CTA has idle groups
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="Rosenbrock34", abs_tol=1e-12, rel_tol=1e-2, Treact=1000.0)
    chem = capi.Chemistry(gri, opts)
    e = chem.solve_host(Y[:, :0], T[:0], p[:0], rho[:0], DT, DTCHEM)
    assert e["RR"].shape == (gri.n_species, 0)
    got = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
    ref = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    cold = T <= 1000.0
    assert cold.any() and (~cold).any()
    assert (got["status"][cold] == capi.CELL_SKIPPED).all() and (got["RR"][:, cold] == 0).all()
    assert (got["deltaTChem"][cold] == DTCHEM).all()
    _compare(gri, Y, rho, got, ref, DT, "gri/Treact")
    # local time stepping: a different flow step per cell
    dts = DT * (0.25 + 1.5 * np.random.default_rng(3).random(sel.size))
    got = chem.solve_host(Y, T, p, rho, dts, DTCHEM)
    ref = O.OracleModel(gri, opts).solve_batch(Y, T, p, rho, dts, DTCHEM, nthreads=8)
    _compare(gri, Y, rho, got, ref, dts, "gri/LTS")
    chem.close()


def test_driver_on_one_gpu_equals_direct_call(gri, flame):
    """lbchem_lb_solve with one rank (the library's load-balanced entry point over a 1-rank NCCL communicator): same
    fields as the plain batch call, the second step starts from the first step's deltaTChem (reference .C:130,398,490),
    and the LoadBalancedTDAC driver leaves deltaTChem untouched (reference TDAC .C:968-971)."""
    import torch
    from loadbalancedchemistrymodel_b200.model import LoadBalancedChemistryModel
    sel = np.argsort(-flame["T"])[:512]
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    chem = capi.Chemistry(gri, opts, device=0)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    tY, tT, tp, trho = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (Y, T, p, rho))
    model = LoadBalancedChemistryModel(chem)
    dtc0 = torch.full((sel.size,), DTCHEM, dtype=torch.float64, device=dev)
    m1 = model.solve(tY, tT, tp, trho, DT, dtc0)
    direct = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
    assert np.array_equal(model.RR.cpu().numpy(), direct["RR"])
    assert np.array_equal(model.deltaTChem.cpu().numpy(), direct["deltaTChem"])
    assert m1 == direct["deltaTChem"].min()
    assert model.total_cost == direct["cost"].sum()
    m2 = model.solve(tY, tT, tp, trho, DT)                      # carries deltaTChem
    direct2 = chem.solve_host(Y, T, p, rho, DT, direct["deltaTChem"])
    assert np.array_equal(model.RR.cpu().numpy(), direct2["RR"])
    # host-array entry point of the same step
    ho = {"RR": np.empty_like(Y), "T": np.empty_like(T), "cost": np.empty_like(T), "status": np.empty(T.size, np.int32),
          "deltaTChem": np.empty_like(T)}
    m3 = model.solve_host(Y, T, p, rho, DT, direct["deltaTChem"], ho)
    assert np.array_equal(ho["RR"], direct2["RR"]) and np.array_equal(ho["deltaTChem"], direct2["deltaTChem"]) and m3 == m2
    model.close()
    tdac = LoadBalancedChemistryModel(chem, tdac=True)
    m4 = tdac.solve(tY, tT, tp, trho, DT, dtc0)
    assert np.array_equal(tdac.RR.cpu().numpy(), direct["RR"]) and torch.equal(tdac.deltaTChem, dtc0) and m4 == DTCHEM
    tdac.close()
    chem.close()


# ---- LoadBalancedTDAC with DAC reduction (BASELINE config 5) ------------------------
DAC_TOL = 1e-4                      # tests/Cases/Case-chemistry/constant/chemistryProperties:46
DAC_INIT = ("CO", "CH4", "HO2")     # initialSet, chemistryProperties:52-57


@pytest.mark.parametrize("solver,rel_tol", [("seulex", 1e-1), ("Rosenbrock34", 1e-2)])
def test_gri_flame_fixture_dac_parity(gri, flame, solver, rel_tol):
    """All 4000 fixture cells with `reduction{active on; method DAC; tolerance 1e-4; initialSet CO CH4 HO2}`:
    the kernel integrates the reduced system embedded in the full index space (removed species' rows and columns
    skipped), the oracle integrates the genuinely reduced NsDAC_+2 system as TDACChemistryModel does. Same active
    sets cell by cell, RR / deltaTChem within 1e-6, temperature returned unchanged (reference T/p aliasing)."""
    sel = np.arange(flame["T"].shape[0])
    Y, T, p, rho = _state(gri, flame, sel)
    init = [list(gri.species).index(s) for s in DAC_INIT]
    opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol)
    chem = capi.Chemistry(gri, opts)
    chem.set_reduction(DAC_TOL, init)
    got = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
    ns_gpu = chem.active_species(T.size)
    st = chem.stats()
    om = O.OracleModel(gri, opts)
    om.set_reduction(DAC_TOL, init)
    ref = om.solve_batch(Y, T, p, rho, DT, DTCHEM, nthreads=8)
    print(solver, "NsDAC min/mean/max", ns_gpu.min(), ns_gpu.mean(), ns_gpu.max(), "gpu stats", st, "oracle", ref["counters"])
    # a species whose path value sits within round-off of the tolerance may flip: none expected on this fixture
    assert (ns_gpu == ref["n_active"]).all(), np.nonzero(ns_gpu != ref["n_active"])[0][:10]
    assert ns_gpu.min() >= len(init) and ns_gpu.max() < gri.n_species
    assert np.array_equal(got["T"], T) and np.array_equal(ref["T"], T)
    _compare(gri, Y, rho, got, ref, DT, f"gri-dac/{solver}")
    # switching the reduction off again gives the full-mechanism result
    chem.set_reduction(None)
    full = chem.solve_host(Y[:, :256], T[:256], p[:256], rho[:256], DT, DTCHEM)
    ref_full = O.OracleModel(gri, opts).solve_batch(Y[:, :256], T[:256], p[:256], rho[:256], DT, DTCHEM, nthreads=8)
    _compare(gri, Y[:, :256], rho[:256], full, ref_full, DT, f"gri-dac-off/{solver}")
    chem.close()


def test_qdot_kernel(gri, flame):
    """lbchem_qdot against -sum_i hf_i RR_i in numpy (hf from the low janaf range at Tstd, janafThermo::hc), on the RR of
    a fixture solve and on a 2 M-cell random block where its HBM throughput is printed (466 B per cell algorithmic)."""
    import torch
    sel = np.arange(0, flame["T"].shape[0], 4)
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    chem = capi.Chemistry(gri, opts)
    got = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
    Ts = 298.15
    a = gri.janaf_low
    hf = opts.gas_constant * (((((a[:, 4] / 5 * Ts + a[:, 3] / 4) * Ts + a[:, 2] / 3) * Ts + a[:, 1] / 2) * Ts + a[:, 0]) * Ts + a[:, 5]) / gri.W
    want = -(hf[:, None] * got["RR"]).sum(0)
    dev = torch.device("cuda:0")
    RR = torch.from_numpy(got["RR"]).to(dev)
    Q = torch.empty(sel.size, dtype=torch.float64, device=dev)
    chem.qdot_device(sel.size, sel.size, RR, Q)
    torch.cuda.synchronize()
    q = Q.cpu().numpy()
    assert np.allclose(q, want, rtol=1e-12, atol=1e-9 * np.abs(want).max())
    # no heat release in the cold inflow, strong release in the flame (the fixture's own Qdot field peaks at 1.1e8 W/m3)
    assert np.abs(q[T < 400.0]).max() < 1e3 and np.abs(q).max() > 1e7
    # ragged leading dimension + empty batch
    n, ldy = 1000, 1536
    RRp = torch.randn((gri.n_species, ldy), dtype=torch.float64, device=dev)
    Qp = torch.zeros(n, dtype=torch.float64, device=dev)
    chem.qdot_device(n, ldy, RRp, Qp)
    chem.qdot_device(0, ldy, RRp, Qp)
    torch.cuda.synchronize()
    terms = hf[:, None] * RRp[:, :n].cpu().numpy()
    assert (np.abs(Qp.cpu().numpy() + terms.sum(0)) <= 1e-13 * np.abs(terms).sum(0)).all()   # random signs: sums cancel
    # throughput on a block larger than L2
    n = 2_000_000
    RRb = torch.randn((gri.n_species, n), dtype=torch.float64, device=dev)
    Qb = torch.empty(n, dtype=torch.float64, device=dev)
    for _ in range(3):
        chem.qdot_device(n, n, RRb, Qb)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        chem.qdot_device(n, n, RRb, Qb)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"qdot: {n} cells in {ms:.3f} ms = {(gri.n_species + 1) * 8 * n / ms / 1e6:.0f} GB/s")
    chem.close()


@pytest.mark.parametrize("dac", [False, True])
def test_full_size_properties(gri, dac):
    """BASELINE config 3 / 5 at their full size (1 M cells per GPU) through size-independent properties: the batch is
    1000 distinct cells of the workload tiled 1000 times in a shuffled order, so (i) every copy of a cell must come back
    bit-identical whatever CTA, group or queue position integrated it, (ii) the distinct cells match the oracle at 1e-6,
    (iii) mass is conserved cell by cell, (iv) nothing fails and the counters are 1000 x those of the distinct cells."""
    from loadbalancedchemistrymodel_b200 import workloads as W
    nd, reps = 1000, 1000
    mech, Yd, Td, pd_, rhod = (W.gri_tiled if dac else W.gri_uniform)(nd)
    rng = np.random.default_rng(11)
    idx = np.tile(np.arange(nd), reps)
    rng.shuffle(idx)
    Y, T, p, rho = np.ascontiguousarray(Yd[:, idx]), Td[idx], pd_[idx], rhod[idx]
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    chem = capi.Chemistry(mech, opts)
    om = O.OracleModel(mech, opts)
    if dac:
        init = [list(mech.species).index(s) for s in DAC_INIT]
        chem.set_reduction(DAC_TOL, init); om.set_reduction(DAC_TOL, init)
    got = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
    st = chem.stats()
    assert st["n_cells"] == nd * reps and st["n_failed"] == 0 and (got["status"] == 0).all()
    first = np.full(nd, -1)
    first[idx[::-1]] = np.arange(idx.size)[::-1]          # first position of every distinct cell
    for k in ("RR", "T", "deltaTChem", "cost"):
        a = got[k]
        rep = a[..., first][..., idx]                      # what every position should hold
        assert np.array_equal(a, rep), f"{k}: copies of the same cell differ"
    one = {k: np.ascontiguousarray(got[k][..., first]) for k in ("RR", "T", "deltaTChem", "cost", "status")}
    ref = om.solve_batch(Yd, Td, pd_, rhod, DT, DTCHEM, nthreads=8)
    # hot cells only (config 3): 14 of 1000 land on the other side of the last-sub-step coin toss described in _compare
    _compare(mech, Yd, rhod, one, ref, DT, f"full-size/{'dac' if dac else 'uniform'}", max_other_count=0.02)
    assert st["n_attempts"] == reps * int(one["cost"].sum())
    RR = got["RR"]
    assert (np.abs(RR.sum(0)) <= 1e-4 * np.abs(RR).sum(0) + 1e-8).all()
    if dac:
        ns = chem.active_species(idx.size)
        assert np.array_equal(ns, ns[first][idx]) and np.array_equal(ns[first], ref["n_active"])
    chem.close()


def test_tc_kernel(gri, flame):
    """lbchem_tc (StandardChemistryModel::tc) against the oracle on all fixture cells, plus a ragged leading dimension."""
    import torch
    dev = torch.device("cuda:0")
    sel = np.arange(flame["T"].shape[0])
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    chem = capi.Chemistry(gri, opts)
    want = O.OracleModel(gri, opts).tc(Y, T, rho)
    n, ldy = sel.size, sel.size + 96
    Yd = torch.zeros((gri.n_species, ldy), dtype=torch.float64, device=dev)
    Yd[:, :n] = torch.from_numpy(Y).to(dev)
    Td, pd_, rd = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (T, p, rho))
    out = torch.full((n,), -1.0, dtype=torch.float64, device=dev)
    chem.tc_device(n, ldy, Yd, Td, pd_, rd, out)
    chem.tc_device(0, ldy, Yd, Td, pd_, rd, out)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    err = np.abs(got - want) / want
    print(f"tc: min {got.min():.3e} s max {got.max():.3e} s, max rel err vs oracle {err.max():.2e}")
    assert (got > 0).all() and err.max() < 1e-9
    chem.close()


# ---- option switches, host pipeline, cost order, device cellsToSend, the load-balanced step on real kernels ----------
@pytest.mark.parametrize("solver,rel_tol", [("seulex", 1e-1), ("Rosenbrock34", 1e-2)])
@pytest.mark.parametrize("switch", [dict(dfdx_is_dcdt=0), dict(kc_floor=1e-12), dict(gas_constant=8314.47),
                                    dict(seulex_errold_max=1)])
def test_doubtful_openfoam_details_are_switches_on_both_sides(gri, flame, solver, rel_tol, switch):
    """Every OpenFOAM detail SURVEY Appendix A marks doubtful (dfdx = dcdt quirk, Kc floor, gas constant, seulex errOld)
    is an option of the kernel AND of the oracle: parity with the non-default branch, on every 8th fixture cell."""
    if "seulex_errold_max" in switch and solver != "seulex":
        pytest.skip("seulex only")
    sel = np.arange(0, flame["T"].shape[0], 8)
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol, **switch)
    if "gas_constant" in switch:
        rho = rho * capi.RR_DEFAULT / 8314.47
    got, ref, st = _solve_both(gri, Y, T, p, rho, opts)
    _compare(gri, Y, rho, got, ref, DT, f"switch {switch}/{solver}", max_other_count=0.02)
    base, _, _ = _solve_both(gri, Y, T, p, rho, capi.default_options(solver=solver, abs_tol=1e-12, rel_tol=rel_tol))
    changed = not np.array_equal(base["RR"], got["RR"])
    print("switch", switch, "changes the result:", changed)
    if "gas_constant" in switch or ("dfdx_is_dcdt" in switch and solver == "Rosenbrock34"):
        assert changed                      # the switch reaches the kernel (seulex never reads dfdx; the Kc floor and errOld
                                            # branches are rarely active on this fixture)


def test_seulex_tight_tolerance_uses_the_full_extrapolation_table(gri, flame):
    """relTol 1e-8 starts seulex at kTarg = 5 and lets it climb: the kernel keeps the reference's full 11-row table
    (no order cap), parity with the oracle at the same tolerance."""
    sel = np.argsort(-flame["T"])[::29][:96]
    Y, T, p, rho = _state(gri, flame, sel)
    opts = capi.default_options(solver="seulex", abs_tol=1e-14, rel_tol=1e-8)
    got, ref, st = _solve_both(gri, Y, T, p, rho, opts, dt=1e-5)
    assert (got["status"] == 0).all()
    # many more sub-steps per cell at this tolerance: more last-sub-step coin tosses (see _compare), same states
    # (the proposed next sub-step is err^(-1/(k+1)) of an error estimate that is itself a difference of nearly equal
    # extrapolation columns at this tolerance: its last digits are round-off, hence 1e-4 on deltaTChem here)
    # ~600 sub-steps per cell: N2, which changes by 1e-9 of its concentration over the step, comes back 8 ulps of c apart
    # from the oracle (1.2e-6 of its RR; Y agrees to 6e-11): 32 ulps of c0 are allowed as representation noise of RR
    _compare(gri, Y, rho, got, ref, 1e-5, "seulex 1e-8", abs_tol=1e-14, max_other_count=0.15, dtchem_tol=1e-4, rr_noise_ulps=32)


def test_host_pipeline_with_pinned_buffers_equals_device_path(gri, flame, monkeypatch):
    """lbchem_solve_batch_host with PINNED caller arrays and many chunks: the kernels of consecutive chunks overlap on the
    device (three streams), each with its own work queue and its own scratch; the result must be bit-identical to one
    launch over device-resident arrays."""
    import torch
    monkeypatch.setenv("LBCHEM_CHUNK", "1024")
    rep = 12
    sel = np.tile(np.argsort(-flame["T"])[:1500], rep)
    Y, T, p, rho = _state(gri, flame, sel)
    n = sel.size
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1)
    chem = capi.Chemistry(gri, opts)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    out = {"RR": pin(np.empty((gri.n_species, n))), "T": pin(np.empty(n)), "cost": pin(np.empty(n)),
           "status": pin(np.empty(n, np.int32)), "deltaTChem": pin(np.empty(n))}
    for _ in range(3):
        got = chem.solve_host(pin(Y), pin(T), pin(p), pin(rho), DT, DTCHEM, out=out)
        dev = torch.device("cuda:0")
        tY, tT, tp, trho = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (Y, T, p, rho))
        dtc = torch.full((n,), DTCHEM, dtype=torch.float64, device=dev)
        RR = torch.empty((gri.n_species, n), dtype=torch.float64, device=dev)
        To = torch.empty(n, dtype=torch.float64, device=dev); co = torch.empty(n, dtype=torch.float64, device=dev)
        so = torch.empty(n, dtype=torch.int32, device=dev)
        chem.solve_device(n, n, tY, tT, tp, trho, DT, dtc, RR, To, co, so)
        torch.cuda.synchronize()
        assert np.array_equal(got["RR"], RR.cpu().numpy()) and np.array_equal(got["deltaTChem"], dtc.cpu().numpy())
        assert np.array_equal(got["T"], To.cpu().numpy()) and np.array_equal(got["cost"], co.cpu().numpy())
    chem.close()


def test_cost_ordered_dispatch_gives_the_same_results(gri, flame):
    """sort_by_cost: the work queue hands the cells out in descending order of the previous step's cost. Any order gives
    bit-identical results; the order itself is a permutation sorted by cost."""
    import torch
    rng = np.random.default_rng(5)
    sel = rng.permutation(flame["T"].shape[0])[:3000]
    Y, T, p, rho = _state(gri, flame, sel)
    n = sel.size
    dev = torch.device("cuda:0")
    tY, tT, tp, trho = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (Y, T, p, rho))
    res = []
    for sort in (0, 1):
        chem = capi.Chemistry(gri, capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1, sort_by_cost=sort))
        dtc = torch.full((n,), DTCHEM, dtype=torch.float64, device=dev)
        RR = torch.empty((gri.n_species, n), dtype=torch.float64, device=dev)
        To = torch.empty(n, dtype=torch.float64, device=dev); co = torch.zeros(n, dtype=torch.float64, device=dev)
        so = torch.empty(n, dtype=torch.int32, device=dev)
        for step in range(2):                            # the second step is dispatched by the first step's cost
            chem.solve_device(n, n, tY, tT, tp, trho, DT, dtc, RR, To, co, so)
        torch.cuda.synchronize()
        res.append((RR.cpu().numpy(), dtc.cpu().numpy(), co.cpu().numpy()))
        chem.close()
    for a, b in zip(*res):
        assert np.array_equal(a, b)


def test_cells_to_send_on_the_device_is_bit_exact_at_full_size(gri):
    """lbchem_cells_to_send_device (what lbchem_lb_solve runs) against the sequential host loop on 1 M integer costs with
    fractional targets, chained bounds, targets of zero, unreachable targets, ragged sizes."""
    import torch
    chem = capi.Chemistry(gri, capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1))
    rng = np.random.default_rng(9)
    for n in (1, 5, 1023, 1024, 1025, 4000, 1_000_003):
        cost = rng.integers(0, 90, n).astype(np.float64)
        if n > 10:
            cost[: n // 10] += rng.integers(50, 5000, n // 10)     # hot head
        total = cost.sum()
        tc = torch.from_numpy(cost).cuda()
        for targets in ([total * 0.37], [total * 0.1234567, total * 0.25, 0.0, total * 0.031], [total * 2.0, 1.0], [0.5, 0.5, 1e-3],
                        [total / 7.0] * 7):
            want = [0]
            for t in targets:
                want.append(capi.cells_to_send(cost, t, want[-1]))
            assert chem.cells_to_send_device(tc, targets) == want, (n, targets)
    chem.close()


@pytest.mark.parametrize("tdac", [False, True])
def test_two_ranks_on_one_gpu_redistribute_real_cells(gri, flame, tdac):
    """The load-balanced flow step (C++ driver, lbchem_lb_core.hpp) with the REAL CUDA solver: two logical ranks on the one
    test GPU through the in-process loopback transport (device-to-device copies standing in for ncclSend / ncclRecv).
    Rank 0 holds the flame (hot, expensive cells), rank 1 cold inflow cells: from the second step on rank 0 ships cells, and
    every field must be bit-identical to the non-balanced solve of the same cells (the reference's own acceptance
    criterion LB == non-LB, tests/src/standardChemistryModel-Test.C:90-102, here to the last bit)."""
    import threading
    import torch
    from loadbalancedchemistrymodel_b200.model import LoadBalancedChemistryModel
    order = np.argsort(-flame["T"])
    sels = [order[:2500], order[-700:]]          # rank 0: 2500 flame cells, rank 1: 700 cold inflow cells
    dev = torch.device("cuda:0")
    opts = capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1, cost_model=1)
    init = [list(gri.species).index(s) for s in DAC_INIT]
    chems = [capi.Chemistry(gri, opts, device=0) for _ in range(2)]
    single = capi.Chemistry(gri, opts, device=0)
    if tdac:
        for c in chems + [single]:
            c.set_reduction(DAC_TOL, init)
    models = LoadBalancedChemistryModel.make_loopback(chems, tdac=tdac)
    states = []
    for sel in sels:
        Y, T, p, rho = _state(gri, flame, sel)
        states.append(tuple(torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (Y, T, p, rho)) + ((Y, T, p, rho),))
    dtc = [torch.full((sel.size,), DTCHEM, dtype=torch.float64, device=dev) for sel in sels]
    ref_dtc = [np.full(sel.size, DTCHEM) for sel in sels]
    errors = []
    for step in range(3):
        mins = [None, None]

        def work(r):
            try:
                torch.cuda.set_device(0)
                with torch.cuda.stream(torch.cuda.Stream(dev)):
                    tY, tT, tp, trho, _ = states[r]
                    mins[r] = models[r].solve(tY, tT, tp, trho, DT, dtc[r])
                    torch.cuda.current_stream().synchronize()
            except Exception as e:          # pragma: no cover
                errors.append(e)
        th = [threading.Thread(target=work, args=(r,)) for r in range(2)]
        for t in th:
            t.start()
        for t in th:
            t.join(timeout=600)
        assert not errors, errors
        sent = [m.last["sent_cells"] for m in models]; recv = [m.last["recv_cells"] for m in models]
        print(f"step {step}: sent {sent} recv {recv} plan {models[0].send_plan} total cost {[m.total_cost for m in models]}")
        if step == 0:
            assert sent == [0, 0] and not models[0].last["exchanged"]           # firstTime_: all local
        else:
            assert sent[0] > 0 and sent[0] == recv[1] and sent[1] == 0 and recv[0] == 0
            assert models[0].send_plan and models[0].send_plan[0][0] == 1 and models[1].recv_plan == [0]
        for r in range(2):
            _, _, _, _, (Y, T, p, rho) = states[r]
            want = single.solve_host(Y, T, p, rho, DT, ref_dtc[r])
            m = models[r]
            assert np.array_equal(m.RR.cpu().numpy(), want["RR"]), (step, r)
            assert np.array_equal(m.T_out.cpu().numpy(), want["T"]) and np.array_equal(m.cell_cost.cpu().numpy(), want["cost"])
            assert np.array_equal(m.status.cpu().numpy(), want["status"])
            if tdac:      # LoadBalancedTDAC never copies deltaTChem back (reference TDAC .C:968-971)
                assert np.array_equal(m.deltaTChem.cpu().numpy(), ref_dtc[r]) and mins[r] == DTCHEM
            else:
                assert np.array_equal(m.deltaTChem.cpu().numpy(), want["deltaTChem"]) and mins[r] == want["deltaTChem"].min()
                ref_dtc[r] = want["deltaTChem"]
                dtc[r] = m.deltaTChem
            assert m.total_cost == want["cost"].sum()
    for m in models:
        m.close()
    for c in chems + [single]:
        c.close()


def test_miss_list_integrates_only_the_listed_cells(gri, flame):
    """lbchem_solve_batch_indexed: the device half of ISAT-aware batching (reference
    LoadBalancedTDACChemistryModel.C:683-740 builds the miss list on the host). A random 30 % of the fixture cells are
    'misses': they come back exactly as in the full solve, every other cell's outputs stay untouched."""
    import torch
    dev = torch.device("cuda:0")
    sel = np.arange(flame["T"].shape[0])
    Y, T, p, rho = _state(gri, flame, sel)
    n = sel.size
    chem = capi.Chemistry(gri, capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1))
    full = chem.solve_host(Y, T, p, rho, DT, DTCHEM)
    rng = np.random.default_rng(21)
    miss = np.sort(rng.choice(n, int(0.3 * n), replace=False)).astype(np.int32)
    rng.shuffle(miss)                                                   # any order
    tY, tT, tp, trho = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (Y, T, p, rho))
    dtc = torch.full((n,), DTCHEM, dtype=torch.float64, device=dev)
    RR = torch.full((gri.n_species, n), -7.0, dtype=torch.float64, device=dev)
    To = torch.full((n,), -7.0, dtype=torch.float64, device=dev); co = torch.full((n,), -7.0, dtype=torch.float64, device=dev)
    so = torch.full((n,), -7, dtype=torch.int32, device=dev)
    chem.solve_indexed(torch.from_numpy(miss).to(dev), n, tY, tT, tp, trho, DT, dtc, RR, To, co, so)
    torch.cuda.synchronize()
    st = chem.stats()
    assert st["n_cells"] == miss.size
    hit = np.ones(n, bool); hit[miss] = False
    RRh, dtch, Toh, coh, soh = RR.cpu().numpy(), dtc.cpu().numpy(), To.cpu().numpy(), co.cpu().numpy(), so.cpu().numpy()
    assert np.array_equal(RRh[:, miss], full["RR"][:, miss]) and np.array_equal(dtch[miss], full["deltaTChem"][miss])
    assert np.array_equal(Toh[miss], full["T"][miss]) and np.array_equal(coh[miss], full["cost"][miss]) and (soh[miss] == 0).all()
    assert (RRh[:, hit] == -7.0).all() and (dtch[hit] == DTCHEM).all() and (Toh[hit] == -7.0).all() and (soh[hit] == -7).all()
    chem.solve_indexed(torch.zeros(0, dtype=torch.int32, device=dev), n, tY, tT, tp, trho, DT, dtc, RR, To, co, so)   # empty list
    chem.close()
