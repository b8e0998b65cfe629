"""ctypes wrapper around oracle/liblbchem_oracle.so — CPU ORACLE, TEST INFRASTRUCTURE ONLY.

May be imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs (see the header of lbchem_oracle.c). PARITY UNPINNED against OpenFOAM.
"""
from __future__ import annotations

import ctypes as C
import subprocess
import sys
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(_HERE.parent))
from loadbalancedchemistrymodel_b200.capi import (  # noqa: E402
    MechArrays, MechDesc, Options, _as_dp, _as_ip, _dp, _ip, default_options,
)

LIB = _HERE / "liblbchem_oracle.so"


def build(force: bool = False) -> Path:
    src = _HERE / "lbchem_oracle.c"
    if force or not LIB.exists() or LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(["make", "-C", str(_HERE), "-B" if force else "-s"])
    return LIB


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(LIB))
        L.oracle_mech_create.argtypes = [C.POINTER(MechDesc)]
        L.oracle_mech_create.restype = C.c_void_p
        L.oracle_mech_destroy.argtypes = [C.c_void_p]
        L.oracle_solve_batch.argtypes = [
            C.c_void_p, C.POINTER(Options), C.c_int, C.c_int64, C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double,
            _dp, _dp, _dp, _dp, _dp, _ip, C.POINTER(C.c_int64), C.c_int,
        ]
        L.oracle_derivatives.argtypes = [C.c_void_p, C.POINTER(Options), _dp, _dp]
        L.oracle_jacobian.argtypes = [C.c_void_p, C.POINTER(Options), _dp, _dp, _dp]
        L.oracle_rates.argtypes = [C.c_void_p, C.POINTER(Options), _dp, C.c_double, _dp, _dp]
        L.oracle_lu_solve.argtypes = [C.c_int, _dp, _dp]
        L.oracle_balance_plan.argtypes = [C.c_int32, _dp, _ip, C.c_double, C.c_int32, _ip, _ip, _dp, _ip, _ip]
        L.oracle_cells_to_send.argtypes = [_dp, C.c_int32, C.c_double, C.c_int32]
        L.oracle_cells_to_send.restype = C.c_int32
        L.oracle_tc.argtypes = [C.c_void_p, C.POINTER(Options), C.c_int64, C.c_int64, _dp, _dp, _dp, _dp]
        L.oracle_tc.restype = None
        L.oracle_set_reduction.argtypes = [C.c_int, C.c_double, C.c_int, _ip, _ip]
        L.oracle_set_reduction.restype = None
        L.oracle_dac_reduce.argtypes = [C.c_void_p, C.POINTER(Options), C.c_double, C.c_int, _ip, _dp, C.c_double, _ip, _ip]
        L.oracle_dac_reduce.restype = C.c_int
        _lib = L
    return _lib


class OracleModel:
    """CPU restatement of solveCellList over a batch; same array conventions as the C ABI."""

    def __init__(self, mech, options: Options | None = None, seulex_errold_max: bool = False):
        self.arrays = MechArrays(mech)
        self.mech = mech
        self.h = lib().oracle_mech_create(C.byref(self.arrays.desc))
        self.options = options or default_options()
        self.seulex_errold_max = int(seulex_errold_max)
        self.counters = None
        self.reduction = None          # (tolerance, [initial-set species indices]) = reduction{active on; method DAC; ...}

    def set_reduction(self, tolerance=None, initial_set=None):
        """DAC as LoadBalancedTDAC runs it (reference LoadBalancedTDACChemistryModel.C:385-393); None switches it off."""
        self.reduction = None if tolerance is None else (float(tolerance), [int(i) for i in initial_set])

    def dac_reduce(self, c, T, tolerance, initial_set):
        """Active-species flags [nSp] and disabled-reaction flags [nR] DAC gives for one state."""
        c = np.ascontiguousarray(c, dtype=np.float64)
        init = np.ascontiguousarray(initial_set, dtype=np.int32)
        act = np.zeros(self.mech.n_species, np.int32); dis = np.zeros(self.mech.n_reactions, np.int32)
        lib().oracle_dac_reduce(self.h, C.byref(self.options), float(tolerance), len(init), _as_ip(init), _as_dp(c),
                                float(T), _as_ip(act), _as_ip(dis))
        return act.astype(bool), dis.astype(bool)

    def __del__(self):
        try:
            if self.h:
                lib().oracle_mech_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def solve_batch(self, Y, T, p, rho, deltaT, deltaTChem, nthreads: int = 1):
        """Y [nSp,n]; returns dict(RR [nSp,n], T, deltaTChem, cpu_time, attempts, status, counters)."""
        Y = np.ascontiguousarray(Y, dtype=np.float64)
        nSp, n = Y.shape
        T = np.ascontiguousarray(T, dtype=np.float64)
        p = np.ascontiguousarray(np.broadcast_to(p, (n,)), dtype=np.float64)
        rho = np.ascontiguousarray(rho, dtype=np.float64)
        dtc = np.array(np.broadcast_to(deltaTChem, (n,)), dtype=np.float64)
        if np.ndim(deltaT) == 0:
            dT_arr, dT_u = None, float(deltaT)
        else:
            dT_arr, dT_u = np.ascontiguousarray(deltaT, dtype=np.float64), 0.0
        RR = np.zeros_like(Y); Tout = np.zeros(n); cpu = np.zeros(n); att = np.zeros(n)
        status = np.zeros(n, np.int32)
        cnt = (C.c_int64 * 6)()
        n_active = np.full(n, nSp, np.int32)
        if self.reduction is not None:
            init = np.ascontiguousarray(self.reduction[1], dtype=np.int32)
            lib().oracle_set_reduction(1, self.reduction[0], len(init), _as_ip(init), _as_ip(n_active))
        try:
            rc = self._call(n, Y, T, p, rho, dT_arr, dT_u, dtc, RR, Tout, cpu, att, status, cnt, nthreads)
        finally:
            if self.reduction is not None:
                lib().oracle_set_reduction(0, 0.0, 0, None, None)
        assert rc == 0
        names = ["n_steps", "n_attempts", "n_jac", "n_lu", "n_solve", "n_rhs"]
        self.counters = {k: int(v) for k, v in zip(names, cnt)}
        return dict(RR=RR, T=Tout, deltaTChem=dtc, cpu_time=cpu, attempts=att, status=status,
                    counters=self.counters, n_active=n_active)

    def _call(self, n, Y, T, p, rho, dT_arr, dT_u, dtc, RR, Tout, cpu, att, status, cnt, nthreads):
        return lib().oracle_solve_batch(
            self.h, C.byref(self.options), self.seulex_errold_max, n, n, _as_dp(Y), _as_dp(T), _as_dp(p),
            _as_dp(rho), _as_dp(dT_arr), dT_u, _as_dp(dtc), _as_dp(RR), _as_dp(Tout), _as_dp(cpu), _as_dp(att),
            _as_ip(status), cnt, int(nthreads),
        )

    def tc(self, Y, T, rho):
        """Chemical time scale of every cell (StandardChemistryModel::tc)."""
        Y = np.ascontiguousarray(Y, dtype=np.float64); T = np.ascontiguousarray(T, dtype=np.float64)
        rho = np.ascontiguousarray(rho, dtype=np.float64)
        out = np.zeros(T.size)
        lib().oracle_tc(self.h, C.byref(self.options), T.size, T.size, _as_dp(Y), _as_dp(T), _as_dp(rho), _as_dp(out))
        return out

    def derivatives(self, y):
        y = np.ascontiguousarray(y, dtype=np.float64)
        out = np.zeros_like(y)
        lib().oracle_derivatives(self.h, C.byref(self.options), _as_dp(y), _as_dp(out))
        return out

    def jacobian(self, y):
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = y.shape[0]
        dfdx = np.zeros(n); dfdy = np.zeros((n, n))
        lib().oracle_jacobian(self.h, C.byref(self.options), _as_dp(y), _as_dp(dfdx), _as_dp(dfdy))
        return dfdx, dfdy

    def rates(self, c, T):
        c = np.ascontiguousarray(c, dtype=np.float64)
        kf = np.zeros(self.mech.n_reactions); kr = np.zeros(self.mech.n_reactions)
        lib().oracle_rates(self.h, C.byref(self.options), _as_dp(c), float(T), _as_dp(kf), _as_dp(kr))
        return kf, kr


def lu_solve(A, b):
    A = np.array(A, dtype=np.float64, order="C"); b = np.array(b, dtype=np.float64)
    st = lib().oracle_lu_solve(A.shape[0], _as_dp(A), _as_dp(b))
    return st, b


def balance_plan(cost, cells=None, min_fraction=0.02, tdac=False):
    cost = np.ascontiguousarray(cost, dtype=np.float64)
    P = cost.shape[0]
    cells = np.zeros(P, np.int32) if cells is None else np.ascontiguousarray(cells, dtype=np.int32)
    sp = np.zeros(P + 1, np.int32); rp = np.zeros(P + 1, np.int32)
    st = np.zeros(P * P, np.int32); sf = np.zeros(P * P); rf = np.zeros(P * P, np.int32)
    lib().oracle_balance_plan(P, _as_dp(cost), _as_ip(cells), float(min_fraction), int(bool(tdac)),
                              _as_ip(sp), _as_ip(st), _as_dp(sf), _as_ip(rp), _as_ip(rf))
    send = [[(int(st[k]), float(sf[k])) for k in range(sp[r], sp[r + 1])] for r in range(P)]
    recv = [[int(rf[k]) for k in range(rp[r], rp[r + 1])] for r in range(P)]
    return send, recv


def cells_to_send(cell_cost, target, start=0) -> int:
    cell_cost = np.ascontiguousarray(cell_cost, dtype=np.float64)
    return int(lib().oracle_cells_to_send(_as_dp(cell_cost), cell_cost.shape[0], float(target), int(start)))


def density(mech, Y, T, p, RR=None):
    """rho = p*Wbar/(RR*T), Wbar = 1/sum(Y_i/W_i)."""
    from loadbalancedchemistrymodel_b200.capi import RR_DEFAULT
    RR = RR or RR_DEFAULT
    Wbar = 1.0 / (np.asarray(Y) / mech.W[:, None]).sum(0)
    return np.asarray(p) * Wbar / (RR * np.asarray(T))
