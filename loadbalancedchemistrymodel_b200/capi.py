"""ctypes binding of the C ABI in ``include/lbchem.h`` (the product library ``liblbchem.so``).

The library is hand-written CUDA for sm_100a plus a C++ host layer; there is no CPU fallback: if the
shared object is missing or a CUDA device is absent, calls fail loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

from .foam_io import Mechanism

_HERE = Path(__file__).resolve().parent
# LBCHEM_LIB: another build of the same library (A/B runs of kernel variants, scripts/gpu_var.sh); never a different implementation
LIB_PATH = Path(os.environ["LBCHEM_LIB"]) if os.environ.get("LBCHEM_LIB") else _HERE / "csrc" / "liblbchem.so"

SOLVER_NONE, SOLVER_EULER_IMPLICIT, SOLVER_ROSENBROCK34, SOLVER_SEULEX = 0, 1, 2, 3
SOLVERS = {
    "noChemistrySolver": SOLVER_NONE,
    "EulerImplicit": SOLVER_EULER_IMPLICIT,
    "Rosenbrock34": SOLVER_ROSENBROCK34,
    "seulex": SOLVER_SEULEX,
}
CELL_OK, CELL_SKIPPED, CELL_MAX_STEPS, CELL_UNDERFLOW, CELL_SINGULAR, CELL_NAN = range(6)

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


class MechDesc(C.Structure):
    _fields_ = [
        ("n_species", C.c_int32), ("n_reactions", C.c_int32), ("n_eff", C.c_int32), ("reserved", C.c_int32),
        ("W", _dp), ("Tlow", _dp), ("Thigh", _dp), ("Tcommon", _dp), ("janaf_high", _dp), ("janaf_low", _dp),
        ("rtype", _ip), ("reversible", _ip), ("arr", _dp), ("arr0", _dp), ("troe", _dp),
        ("eff_index", _ip), ("eff", _dp),
        ("lhs_ptr", _ip), ("lhs_species", _ip), ("lhs_nu", _dp), ("lhs_exp", _dp),
        ("rhs_ptr", _ip), ("rhs_species", _ip), ("rhs_nu", _dp), ("rhs_exp", _dp),
    ]


class Options(C.Structure):
    _fields_ = [
        ("solver", C.c_int32), ("max_steps", C.c_int32), ("abs_tol", C.c_double), ("rel_tol", C.c_double),
        ("Treact", C.c_double), ("deltaTChemMax", C.c_double), ("cTauChem", C.c_double),
        ("eq_rate_limiter", C.c_int32), ("dfdx_is_dcdt", C.c_int32), ("gas_constant", C.c_double),
        ("kc_floor", C.c_double), ("threads_per_cell", C.c_int32), ("sort_by_cost", C.c_int32),
        ("seulex_errold_max", C.c_int32), ("cost_model", C.c_int32),
        ("pivoting", C.c_int32), ("reserved", C.c_int32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("n_cells", C.c_int64), ("n_steps", C.c_int64), ("n_attempts", C.c_int64), ("n_jac", C.c_int64),
        ("n_lu", C.c_int64), ("n_solve", C.c_int64), ("n_rhs", C.c_int64), ("n_failed", C.c_int64),
        ("kernel_ms", C.c_double), ("kernel_launches", C.c_int32), ("reserved", C.c_int32),
    ]


class LbOptions(C.Structure):
    """lbchem_lb_options: LoadBalancedCoeffs{updateIter; minFractionOfCellsToSend} + the TDAC driver switch."""
    _fields_ = [("update_iter", C.c_int32), ("tdac", C.c_int32), ("min_fraction", C.c_double), ("balance", C.c_int32),
                ("reserved", C.c_int32)]


class LbInfo(C.Structure):
    _fields_ = [("rank", C.c_int32), ("nranks", C.c_int32), ("planned", C.c_int32), ("exchanged", C.c_int32),
                ("reallocated", C.c_int32), ("n_send_partners", C.c_int32), ("n_recv_partners", C.c_int32),
                ("reserved", C.c_int32), ("n_cells", C.c_int64), ("sent_cells", C.c_int64), ("recv_cells", C.c_int64),
                ("sent_bytes", C.c_int64), ("recv_bytes", C.c_int64), ("total_cost", C.c_double)]


class P2POp(C.Structure):
    _fields_ = [("is_send", C.c_int32), ("peer", C.c_int32), ("ptr", C.c_void_p), ("bytes", C.c_int64)]


ALL_GATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, _dp, _dp, C.c_int32)
EXCHANGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.POINTER(P2POp))


class Transport(C.Structure):
    _fields_ = [("ctx", C.c_void_p), ("all_gather", ALL_GATHER_FN), ("exchange", EXCHANGE_FN)]


# OpenFOAM v2306 thermodynamic::RR = 1e3 * NA * k (etc/controlDict DimensionedConstants, SI set)
RR_DEFAULT = 1e3 * (6.0221417930e23 * 1.3806504e-23)


def default_options(**kw) -> Options:
    """OpenFOAM defaults (ODESolver: absTol SMALL, relTol 1e-4, maxSteps 10000; Treact 0; deltaTChemMax GREAT)."""
    o = Options(
        solver=SOLVER_ROSENBROCK34, max_steps=10000, abs_tol=1e-15, rel_tol=1e-4, Treact=0.0,
        deltaTChemMax=1e15, cTauChem=1.0, eq_rate_limiter=0, dfdx_is_dcdt=1, gas_constant=RR_DEFAULT,
        kc_floor=1e-6, threads_per_cell=0, sort_by_cost=0, seulex_errold_max=0, cost_model=0, pivoting=1,
    )
    for k, v in kw.items():
        if k == "solver" and isinstance(v, str):
            v = SOLVERS[v]
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


class MechArrays:
    """Keeps contiguous numpy copies alive and exposes them as a ``lbchem_mech_desc``."""

    def __init__(self, mech: Mechanism):
        self.mech = mech
        self._keep = {}
        d = MechDesc()
        d.n_species, d.n_reactions, d.n_eff = mech.n_species, mech.n_reactions, int(mech.eff.shape[0])
        for name, typ in MechDesc._fields_[4:]:
            arr = getattr(mech, name)
            if typ is _dp:
                a = np.ascontiguousarray(arr, dtype=np.float64)
                if a.size == 0:
                    a = np.zeros(1)
                ptr = a.ctypes.data_as(_dp)
            else:
                a = np.ascontiguousarray(arr, dtype=np.int32)
                if a.size == 0:
                    a = np.zeros(1, np.int32)
                ptr = a.ctypes.data_as(_ip)
            self._keep[name] = a
            setattr(d, name, ptr)
        self.desc = d


def _as_dp(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _as_ip(a):
    return None if a is None else a.ctypes.data_as(_ip)


_lib = None


def load_library(path: Path | None = None) -> C.CDLL:
    """dlopen liblbchem.so and declare prototypes. Raises if the CUDA library has not been built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise RuntimeError(
            f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the chemistry path)"
        )
    lib = C.CDLL(str(p))
    H = C.c_void_p
    lib.lbchem_version.restype = C.c_char_p
    lib.lbchem_default_options.argtypes = [C.POINTER(Options)]
    lib.lbchem_create.argtypes = [C.POINTER(MechDesc), C.c_int, C.POINTER(H)]
    lib.lbchem_destroy.argtypes = [H]
    lib.lbchem_set_options.argtypes = [H, C.POINTER(Options)]
    lib.lbchem_get_options.argtypes = [H, C.POINTER(Options)]
    lib.lbchem_last_error.argtypes = [H]
    lib.lbchem_last_error.restype = C.c_char_p
    sig = [H, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.lbchem_solve_batch.argtypes = sig + [C.c_void_p]
    lib.lbchem_solve_batch_host.argtypes = sig
    lib.lbchem_solve_batch_indexed.argtypes = [H, C.c_int64, C.c_void_p, C.c_int64] + [C.c_void_p] * 5 + [C.c_double] + [C.c_void_p] * 6
    lib.lbchem_get_stats.argtypes = [H, C.POINTER(Stats)]
    lib.lbchem_set_reduction.argtypes = [H, C.c_int32, C.c_double, C.c_int32, _ip]
    lib.lbchem_get_active_species.argtypes = [H, C.c_int64, _ip]
    lib.lbchem_qdot.argtypes = [H, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.lbchem_tc.argtypes = [H, C.c_int64, C.c_int64] + [C.c_void_p] * 6
    lib.lbchem_probe.argtypes = [H, C.c_int32, C.c_int32, _dp, C.c_int32, _dp, C.c_int32]
    lib.lbchem_fp64_peak.argtypes = [C.c_int, C.c_double, _dp, _dp]
    lib.lbchem_balance_plan.argtypes = [C.c_int32, _dp, _ip, C.c_double, C.c_int32, _ip, _ip, _dp, _ip, _ip]
    lib.lbchem_cells_to_send.argtypes = [_dp, C.c_int32, C.c_double, C.c_int32]
    lib.lbchem_cells_to_send.restype = C.c_int32
    lib.lbchem_cells_to_send_device.argtypes = [H, C.c_void_p, C.c_int64, _dp, C.c_int32, C.POINTER(C.c_int64), C.c_void_p]
    lib.lbchem_cells_to_send_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32]
    declare_lb(lib)
    L = C.c_void_p
    lib.lbchem_lb_nccl_unique_id.argtypes = [C.c_void_p]
    lib.lbchem_lb_create_nccl.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(LbOptions), C.POINTER(L)]
    lib.lbchem_lb_create_transport.argtypes = [H, C.POINTER(Transport), C.c_int32, C.c_int32, C.POINTER(LbOptions), C.POINTER(L)]
    lib.lbchem_lb_create_loopback.argtypes = [C.c_int32, C.POINTER(H), C.POINTER(LbOptions), C.POINTER(L)]
    lib.lbchem_lb_solve_host.argtypes = [L, C.c_int64, C.c_int64] + [C.c_void_p] * 5 + [C.c_double] + [C.c_void_p] * 5 + [_dp]
    if path is None:
        _lib = lib
    return lib


def declare_lb(lib):
    """Prototypes of the lbchem_lb_* entry points shared by liblbchem.so and the CPU test build liblbchem_lbhost.so."""
    L = C.c_void_p
    lib.lbchem_lb_default_options.argtypes = [C.POINTER(LbOptions)]
    lib.lbchem_lb_set_options.argtypes = [L, C.POINTER(LbOptions)]
    lib.lbchem_lb_destroy.argtypes = [L]
    lib.lbchem_lb_last_error.argtypes = [L]
    lib.lbchem_lb_last_error.restype = C.c_char_p
    lib.lbchem_lb_solve.argtypes = [L, C.c_int64, C.c_int64] + [C.c_void_p] * 5 + [C.c_double] + [C.c_void_p] * 5 + [_dp, C.c_void_p]
    lib.lbchem_lb_get_info.argtypes = [L, C.POINTER(LbInfo)]
    lib.lbchem_lb_get_plan.argtypes = [L, _ip, _ip, _dp, C.POINTER(C.c_int64), _ip, _ip, C.POINTER(C.c_int64)]


LBHOST_PATH = _HERE / "csrc" / "liblbchem_lbhost.so"
HOST_SOLVER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _ip)
_lbhost = None


def load_lbhost() -> C.CDLL:
    """The CPU TEST build of the load-balanced step (same C++ step logic, host memory, stand-in cell solver)."""
    global _lbhost
    if _lbhost is None:
        if not LBHOST_PATH.exists():
            raise RuntimeError(f"{LBHOST_PATH} not found: run __graft_entry__.build()")
        lib = C.CDLL(str(LBHOST_PATH))
        declare_lb(lib)
        lib.lbchem_lbhost_create.argtypes = [C.c_int32, HOST_SOLVER_FN, C.c_void_p, C.POINTER(Transport), C.c_int32, C.c_int32,
                                             C.POINTER(LbOptions), C.POINTER(C.c_void_p)]
        _lbhost = lib
    return _lbhost


def lb_options(update_iter=0, min_fraction=0.02, tdac=False, balance=True) -> LbOptions:
    return LbOptions(update_iter=int(update_iter), tdac=int(bool(tdac)), min_fraction=float(min_fraction), balance=int(bool(balance)), reserved=0)


EXPORTED_SYMBOLS = (
    "lbchem_version lbchem_default_options lbchem_create lbchem_destroy lbchem_set_options lbchem_get_options "
    "lbchem_last_error lbchem_solve_batch lbchem_solve_batch_host lbchem_solve_batch_indexed lbchem_get_stats lbchem_balance_plan "
    "lbchem_cells_to_send lbchem_cells_to_send_device lbchem_cells_to_send_noalloc lbchem_probe lbchem_fp64_peak lbchem_set_reduction lbchem_get_active_species lbchem_qdot lbchem_tc "
    "lbchem_lb_default_options lbchem_lb_nccl_unique_id lbchem_lb_create_nccl lbchem_lb_create_transport lbchem_lb_create_loopback "
    "lbchem_lb_set_options lbchem_lb_destroy lbchem_lb_last_error lbchem_lb_solve lbchem_lb_solve_host lbchem_lb_get_info lbchem_lb_get_plan"
).split()


NOALLOC_MAX_CELLS = 4194304     # LBCHEM_NOALLOC_MAX_CELLS
NOALLOC_MAX_TARGETS = 1024      # LBCHEM_NOALLOC_MAX_TARGETS


def cells_to_send_noalloc(cost_ptr, n, targets_ptr, nT, bounds_ptr, stream=0, lib=None):
    """``lbchem_cells_to_send_noalloc``: raw DEVICE pointers in and out, nothing allocated on either side of the boundary
    (this wrapper only builds ctypes scalars and a 256-byte host message buffer). Returns (rc, message)."""
    lib = lib or load_library()
    err = C.create_string_buffer(256)
    rc = lib.lbchem_cells_to_send_noalloc(C.c_void_p(cost_ptr), int(n), C.c_void_p(targets_ptr), int(nT), C.c_void_p(bounds_ptr),
                                          C.c_void_p(stream), err, 256)
    return int(rc), err.value.decode()


def balance_plan(cost, cells=None, min_fraction=0.02, tdac=False, lib=None):
    """``updateProcessorBalancing`` for every rank. Returns (send, recv): send[r] = [(to, frac), ...]."""
    lib = lib or load_library()
    cost = np.ascontiguousarray(cost, dtype=np.float64)
    P = cost.shape[0]
    cells = np.zeros(P, np.int32) if cells is None else np.ascontiguousarray(cells, dtype=np.int32)
    sp = np.zeros(P + 1, np.int32); rp = np.zeros(P + 1, np.int32)
    st = np.zeros(P * P, np.int32); sf = np.zeros(P * P); rf = np.zeros(P * P, np.int32)
    rc = lib.lbchem_balance_plan(P, _as_dp(cost), _as_ip(cells), float(min_fraction), int(bool(tdac)),
                                 _as_ip(sp), _as_ip(st), _as_dp(sf), _as_ip(rp), _as_ip(rf))
    if rc != 0:
        raise RuntimeError(f"lbchem_balance_plan failed: {rc}")
    send = [[(int(st[k]), float(sf[k])) for k in range(sp[r], sp[r + 1])] for r in range(P)]
    recv = [[int(rf[k]) for k in range(rp[r], rp[r + 1])] for r in range(P)]
    return send, recv


def cells_to_send(cell_cost, target, start=0, lib=None) -> int:
    lib = lib or load_library()
    cell_cost = np.ascontiguousarray(cell_cost, dtype=np.float64)
    return int(lib.lbchem_cells_to_send(_as_dp(cell_cost), cell_cost.shape[0], float(target), int(start)))


PROBE_DERIV, PROBE_JAC, PROBE_LUSOLVE, PROBE_RATES, PROBE_JAC_LUSOLVE = 0, 1, 2, 3, 4


class Chemistry:
    """Owner of one ``lbchem_handle``: a mechanism bound to one CUDA device (one per rank / host thread).

    Mirrors what the reference model object holds after construction
    (``LoadBalancedChemistryModel.C:32-59``): the mechanism, the solver options and scratch.
    """

    def __init__(self, mech: Mechanism, options: Options | None = None, device: int = -1, lib=None):
        self.lib = lib or load_library()
        self.mech = mech
        self.arrays = MechArrays(mech)
        self.h = C.c_void_p()
        rc = self.lib.lbchem_create(C.byref(self.arrays.desc), int(device), C.byref(self.h))
        if rc != 0:
            msg = self.lib.lbchem_last_error(self.h).decode() if self.h else "no CUDA device"
            if self.h:
                self.lib.lbchem_destroy(self.h)
                self.h = C.c_void_p()
            raise RuntimeError(f"lbchem_create failed ({rc}): {msg}")
        self.options = options or default_options()
        self.set_options(self.options)

    def close(self):
        if getattr(self, "h", None):
            self.lib.lbchem_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed ({rc}): {self.lib.lbchem_last_error(self.h).decode()}")

    def set_options(self, options: Options):
        self._check(self.lib.lbchem_set_options(self.h, C.byref(options)), "lbchem_set_options")
        self.options = options

    def set_reduction(self, tolerance=None, initial_set=None):
        """``reduction{active on; method DAC; tolerance; initialSet}`` of LoadBalancedTDAC; ``tolerance=None`` switches it off."""
        if tolerance is None:
            self._check(self.lib.lbchem_set_reduction(self.h, 0, 0.0, 0, None), "lbchem_set_reduction")
            return
        init = np.ascontiguousarray(initial_set, dtype=np.int32)
        self._check(self.lib.lbchem_set_reduction(self.h, 1, float(tolerance), int(init.size), _as_ip(init)),
                    "lbchem_set_reduction")

    def active_species(self, n: int) -> np.ndarray:
        """NsDAC_ of the first n cells of the last reduced batch."""
        out = np.zeros(int(n), np.int32)
        self._check(self.lib.lbchem_get_active_species(self.h, int(n), _as_ip(out)), "lbchem_get_active_species")
        return out

    def stats(self) -> dict:
        st = Stats()
        self._check(self.lib.lbchem_get_stats(self.h, C.byref(st)), "lbchem_get_stats")
        return {k: getattr(st, k) for k, _ in Stats._fields_ if k != "reserved"}

    # ---- host buffers (numpy) -------------------------------------------------------------
    def solve_host(self, Y, T, p, rho, deltaT, deltaTChem, out=None):
        """Y [nSp, n] field-major float64 numpy arrays; returns dict(RR, T, deltaTChem, cost, status)."""
        Y = np.ascontiguousarray(Y, dtype=np.float64)
        nSp, n = Y.shape
        T = np.ascontiguousarray(T, dtype=np.float64)
        p = np.ascontiguousarray(np.broadcast_to(p, (n,)), dtype=np.float64)
        rho = np.ascontiguousarray(rho, dtype=np.float64)
        if out is not None and "deltaTChem" in out:
            dtc = out["deltaTChem"]; dtc[...] = deltaTChem
        else:
            dtc = np.array(np.broadcast_to(deltaTChem, (n,)), dtype=np.float64)
        if np.ndim(deltaT) == 0:
            dT_arr, dT_u = None, float(deltaT)
        else:
            dT_arr, dT_u = np.ascontiguousarray(deltaT, dtype=np.float64), 0.0
        o = out or {}
        RR = o.get("RR") if o.get("RR") is not None else np.empty_like(Y)
        Tout = o.get("T") if o.get("T") is not None else np.empty(n)
        cost = o.get("cost") if o.get("cost") is not None else np.empty(n)
        status = o.get("status") if o.get("status") is not None else np.empty(n, np.int32)
        vp = lambda a: None if a is None else C.c_void_p(a.ctypes.data)
        rc = self.lib.lbchem_solve_batch_host(self.h, n, n, vp(Y), vp(T), vp(p), vp(rho), vp(dT_arr), dT_u,
                                              vp(dtc), vp(RR), vp(Tout), vp(cost), vp(status))
        self._check(rc, "lbchem_solve_batch_host")
        return dict(RR=RR, T=Tout, deltaTChem=dtc, cost=cost, status=status)

    # ---- device buffers (anything with data_ptr(): torch CUDA tensors) ----------------------
    def solve_device(self, n, ldy, Y, T, p, rho, deltaT, deltaTChem, RR, Tout=None, cost=None, status=None, stream=0):
        ptr = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        if isinstance(deltaT, (int, float)):
            dT_arr, dT_u = None, float(deltaT)
        else:
            dT_arr, dT_u = deltaT, 0.0
        rc = self.lib.lbchem_solve_batch(self.h, int(n), int(ldy), ptr(Y), ptr(T), ptr(p), ptr(rho), ptr(dT_arr), dT_u,
                                         ptr(deltaTChem), ptr(RR), ptr(Tout), ptr(cost), ptr(status),
                                         C.c_void_p(int(stream)))
        self._check(rc, "lbchem_solve_batch")

    def solve_indexed(self, index, ldy, Y, T, p, rho, deltaT, deltaTChem, RR, Tout=None, cost=None, status=None, stream=0):
        """Integrate only the cells listed in ``index`` (int32 CUDA tensor): the ISAT miss list (lbchem_solve_batch_indexed)."""
        ptr = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        if isinstance(deltaT, (int, float)):
            dT_arr, dT_u = None, float(deltaT)
        else:
            dT_arr, dT_u = deltaT, 0.0
        rc = self.lib.lbchem_solve_batch_indexed(self.h, int(index.shape[0]), ptr(index), int(ldy), ptr(Y), ptr(T), ptr(p), ptr(rho),
                                                 ptr(dT_arr), dT_u, ptr(deltaTChem), ptr(RR), ptr(Tout), ptr(cost), ptr(status),
                                                 C.c_void_p(int(stream)))
        self._check(rc, "lbchem_solve_batch_indexed")

    def qdot_device(self, n, ldy, RR, Qdot, stream=0):
        """Qdot [W/m3] from the RR block of the last solve (torch CUDA tensors or anything with data_ptr())."""
        rc = self.lib.lbchem_qdot(self.h, int(n), int(ldy), C.c_void_p(RR.data_ptr()), C.c_void_p(Qdot.data_ptr()),
                                  C.c_void_p(int(stream)))
        self._check(rc, "lbchem_qdot")

    def tc_device(self, n, ldy, Y, T, p, rho, tc, stream=0):
        """Chemical time scale [s] of every cell (torch CUDA tensors or anything with data_ptr())."""
        ptr = lambda a: C.c_void_p(a.data_ptr())
        rc = self.lib.lbchem_tc(self.h, int(n), int(ldy), ptr(Y), ptr(T), ptr(p), ptr(rho), ptr(tc), C.c_void_p(int(stream)))
        self._check(rc, "lbchem_tc")

    def cells_to_send_device(self, cost, targets, stream=0):
        """cellsToSend on a device cost vector (torch CUDA tensor) for a sequence of targets; returns the bounds list."""
        t = np.ascontiguousarray(targets, dtype=np.float64)
        b = np.zeros(t.size + 1, np.int64)
        rc = self.lib.lbchem_cells_to_send_device(self.h, C.c_void_p(cost.data_ptr()), int(cost.shape[0]), _as_dp(t), int(t.size),
                                                  b.ctypes.data_as(C.POINTER(C.c_int64)), C.c_void_p(int(stream)))
        self._check(rc, "lbchem_cells_to_send_device")
        return [int(x) for x in b]

    def probe(self, kind, x):
        """x [nItems, inStride] -> out [nItems, outStride] (see include/lbchem.h: lbchem_probe)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = self.mech.n_species + 2
        nR = self.mech.n_reactions
        out_w = {PROBE_DERIV: n, PROBE_JAC: n + n * n, PROBE_LUSOLVE: n, PROBE_RATES: 2 * nR, PROBE_JAC_LUSOLVE: n}[kind]
        out = np.zeros((x.shape[0], out_w))
        self._check(self.lib.lbchem_probe(self.h, kind, x.shape[0], _as_dp(x), x.shape[1], _as_dp(out), out_w),
                    "lbchem_probe")
        return out


def fp64_peak(device: int = -1, ms_target: float = 200.0, lib=None) -> float:
    """Measured DFMA rate of the device in TFLOP/s (lbchem_fp64_peak)."""
    lib = lib or load_library()
    t = C.c_double(); s = C.c_double()
    rc = lib.lbchem_fp64_peak(int(device), float(ms_target), C.byref(t), C.byref(s))
    if rc != 0:
        raise RuntimeError(f"lbchem_fp64_peak failed: {rc}")
    return t.value
