"""Synthetic cell-state batches for the BASELINE.json configs.

All generators return field-major arrays like OpenFOAM's ``Y_[i][celli]``: ``Y [nSp, n]``, ``T [n]``,
``p [n]``, ``rho [n]`` with ``rho = p*Wbar/(R*T)``. They read only the committed fixtures under
``tests/golden`` (made from the reference's own case files by ``tests/golden/make_fixtures.py``).
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

from .capi import RR_DEFAULT
from .foam_io import Mechanism

GOLDEN = Path(__file__).resolve().parents[1] / "tests" / "golden"
H2_SPECIES = ["H2", "H", "O", "O2", "OH", "H2O", "HO2", "H2O2", "N2", "AR"]
DELTA_T = 1e-6          # flow time step of both reference cases (controlDict deltaT 1e-6)
DELTA_T_CHEM0 = 1e-7    # initialChemicalTimeStep of both reference cases


def density(mech: Mechanism, Y, T, p, R=RR_DEFAULT):
    Wbar = 1.0 / (np.asarray(Y) / mech.W[:, None]).sum(0)
    return np.asarray(p) * Wbar / (R * np.asarray(T))


def load_gri() -> Mechanism:
    return Mechanism.load(GOLDEN / "gri30.npz")


def load_ch4() -> Mechanism:
    return Mechanism.load(GOLDEN / "ch4_1step.npz")


def load_h2() -> Mechanism:
    return load_gri().subset(H2_SPECIES)


def load_flame():
    z = np.load(GOLDEN / "case_chemistry_0p2.npz")
    return z["T"], z["p"], np.ascontiguousarray(z["Y"])


def _perturb(rng, Y, T, sigma_y=1e-3, sigma_t=1.0):
    Y = Y * np.exp(rng.normal(0.0, sigma_y, Y.shape))
    Y /= Y.sum(0, keepdims=True)
    return np.ascontiguousarray(Y), T + rng.normal(0.0, sigma_t, T.shape)


def gri_uniform(n: int, seed: int = 1003, rank: int = 0):
    """Config 3: resample fixture cells with T > 1500 K, 0.1 % log-normal noise on Y, 1 K noise on T."""
    mech = load_gri()
    T0, p0, Y0 = load_flame()
    hot = np.nonzero(T0 > 1500.0)[0]
    rng = np.random.default_rng([seed, rank])
    pick = hot[rng.integers(0, hot.size, n)]
    Y, T = _perturb(rng, Y0[:, pick], T0[pick])
    p = np.full(n, 1e5)
    return mech, Y, T, p, density(mech, Y, T, p)


def gri_skewed(n: int, hot_fraction: float = 0.1, seed: int = 1004, rank: int = 0, hot_first: bool = True, world: int = 1):
    """Config 4: `hot_fraction` of the cells drawn as in config 3, the rest from fixture cells with T < 350 K
    (1-2 sub-steps); hot cells contiguous at the start of the GLOBAL index space (the flame front sits in few ranks):
    with `world` ranks of n cells each, rank r holds clamp(hot_fraction*world - r, 0, 1) of its cells hot, so at
    world = 8 rank 0 is 80 % hot and ranks 1..7 are cold."""
    hot_fraction = min(1.0, max(0.0, hot_fraction * world - rank))
    mech = load_gri()
    T0, p0, Y0 = load_flame()
    hot = np.nonzero(T0 > 1500.0)[0]
    cold = np.nonzero(T0 < 350.0)[0]
    rng = np.random.default_rng([seed, rank])
    nh = int(round(n * hot_fraction))
    pick = np.concatenate([hot[rng.integers(0, hot.size, nh)], cold[rng.integers(0, cold.size, n - nh)]])
    if not hot_first:
        pick = pick[rng.permutation(n)]
    Y, T = _perturb(rng, Y0[:, pick], T0[pick])
    p = np.full(n, 1e5)
    return mech, Y, T, p, density(mech, Y, T, p)


def gri_tiled(n: int, seed: int = 1005, rank: int = 0):
    """Config 5 input: the unfiltered 4000-cell fixture tiled to n cells with the same noise."""
    mech = load_gri()
    T0, p0, Y0 = load_flame()
    rng = np.random.default_rng([seed, rank])
    pick = np.arange(n) % T0.size
    Y, T = _perturb(rng, Y0[:, pick], T0[pick])
    p = np.full(n, 1e5)
    return mech, Y, T, p, density(mech, Y, T, p)


def h2_air(n: int, seed: int = 1002, rank: int = 0):
    """Config 2: H2/air, phi ~ U(0.5, 2), T ~ U(900, 2500) K, p = 1e5, radical seed Y_H = 1e-6 U."""
    mech = load_h2()
    rng = np.random.default_rng([seed, rank])
    sp = mech.species
    phi = rng.uniform(0.5, 2.0, n)
    X = np.zeros((len(sp), n))
    X[sp.index("H2")] = 2.0 * phi
    X[sp.index("O2")] = 1.0
    X[sp.index("N2")] = 3.76
    Y = X * mech.W[:, None]
    Y /= Y.sum(0, keepdims=True)
    Y[sp.index("H")] = 1e-6 * rng.uniform(0.0, 1.0, n)
    Y /= Y.sum(0, keepdims=True)
    T = rng.uniform(900.0, 2500.0, n)
    p = np.full(n, 1e5)
    return mech, np.ascontiguousarray(Y), T, p, density(mech, Y, T, p)


def ch4_mixing_line(n: int, seed: int = 1001, rank: int = 0):
    """Config 1: fuel (CH4 = 1) / air (O2 .23, N2 .77) mixing line, z ~ U(0,1), T ~ U(293, 2000) K."""
    mech = load_ch4()
    rng = np.random.default_rng([seed, rank])
    sp = mech.species
    z = rng.uniform(0.0, 1.0, n)
    Y = np.zeros((len(sp), n))
    Y[sp.index("CH4")] = z
    Y[sp.index("O2")] = 0.23 * (1 - z)
    Y[sp.index("N2")] = 0.77 * (1 - z)
    T = rng.uniform(293.0, 2000.0, n)
    p = np.full(n, 1e5)
    return mech, Y, T, p, density(mech, Y, T, p)


# ------------------------------------------------------------------------------------------------
# algorithmic flop model: counts what the REFERENCE algorithm performs per call,
# add/mul/fma-half/div = 1 flop, every exp/log/log10/pow = 1 "special" (weighted 20 in the headline).
# ------------------------------------------------------------------------------------------------
def flop_model(mech: Mechanism, special_weight: float = 20.0) -> dict:
    nSp, nR = mech.n_species, mech.n_reactions
    n = nSp + 2
    f_k = 0.0   # flops of one kf/kr evaluation over all reactions
    s_k = 0.0   # specials of the same
    f_c = 0.0   # concentration products + net rate + scatter into dcdt
    for r in range(nR):
        nl = int(mech.lhs_ptr[r + 1] - mech.lhs_ptr[r]); nr = int(mech.rhs_ptr[r + 1] - mech.rhs_ptr[r])

        def arr(a):
            return (2 + (1 if a[2] != 0 else 0), (1 if a[1] != 0 else 0) + (1 if a[2] != 0 else 0))
        fa, sa = arr(mech.arr[r])
        f_k += fa; s_k += sa
        kind = int(mech.rtype[r])
        if kind >= 1:
            f_k += 2 * nSp + 1                       # M = sum eff_i c_i
        if kind >= 2:
            f0, s0 = arr(mech.arr0[r]); f_k += f0 + 6; s_k += s0
        if kind == 3:
            f_k += 25; s_k += 6                      # 3 exp, 2 log10, 1 pow
        if mech.reversible[r]:
            dn = mech.rhs_nu[mech.rhs_ptr[r]:mech.rhs_ptr[r + 1]].sum() - mech.lhs_nu[mech.lhs_ptr[r]:mech.lhs_ptr[r + 1]].sum()
            f_k += 11 + 10 + 6 + 2                   # Ha poly, S poly, G, arg; max + div
            s_k += 2 + (1 if dn != 0 else 0)         # log T (S), exp, pow(Pstd/RT, dn)
        f_c += (nl + nr) + 3 + 2 * (nl + nr)
    f_thermo = 2 * nSp + nSp * (8 + 3) + nSp * (10 + 3) + 6
    F_rhs = f_k + f_c + f_thermo
    S_rhs = s_k
    f_d = 0.0
    for r in range(nR):
        nl = int(mech.lhs_ptr[r + 1] - mech.lhs_ptr[r]); nr = int(mech.rhs_ptr[r + 1] - mech.rhs_ptr[r])
        f_d += nl * nl + nr * nr + 2 * (nl + nr) * (nl + nr)
    # jacobian = omega(T) + kf/kr again for the analytic block + D terms/scatter + omega(T+d) + omega(T-d) + FD
    F_jac = 3 * (f_k + f_c) + f_k + f_d + 3 * nSp
    S_jac = 4 * s_k
    return dict(n=n, F_lu=2.0 / 3.0 * n ** 3, F_solve=2.0 * n * n, F_rhs=F_rhs, S_rhs=S_rhs, F_jac=F_jac, S_jac=S_jac,
                special_weight=special_weight)


def algorithmic_flops(model: dict, stats: dict) -> float:
    w = model["special_weight"]
    return (stats["n_lu"] * model["F_lu"] + stats["n_solve"] * model["F_solve"]
            + stats["n_rhs"] * (model["F_rhs"] + w * model["S_rhs"])
            + stats["n_jac"] * (model["F_jac"] + w * model["S_jac"]))
