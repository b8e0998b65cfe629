"""``LoadBalancedChemistryModel`` / ``LoadBalancedTDAC`` for one rank (= one GPU): a thin binding of ``lbchem_lb_*``.

The flow step itself -- ``LoadBalancedChemistryModel::solve<DeltaTType>``, reference
src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C:453-704 (first step local, plan, cellsToSend,
forward trip, own tail then received cells, return trip, unpack) -- lives in the C++ library
(``csrc/lbchem_lb_core.hpp``, ``csrc/lbchem_lb_cuda.cu``): device-resident cost vector, bit-exact host plan, NCCL
``ncclSend``/``ncclRecv`` of field-major row slices over NVLink, one two-segment launch of the integration kernel.
This module only owns tensors and hands pointers over; ``torch.distributed`` is used once, to broadcast the NCCL
unique id at construction. There is no CPU fallback.

``HostStepModel`` binds the CPU TEST build of the very same step logic (``liblbchem_lbhost.so``: host memory, a stand-in
cell solver, gloo behind the transport callbacks) so that world_size > 1 can be exercised without a GPU
(tests/test_cpu_exchange.py). It contains no chemistry.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Optional

import numpy as np
import torch
import torch.distributed as dist

from . import capi

GREAT = 1e15


def _plan_of(lib, lb, world):
    ns, nr = C.c_int32(), C.c_int32()
    st = np.zeros(world, np.int32); sa = np.zeros(world); bd = np.zeros(world + 1, np.int64)
    rf = np.zeros(world, np.int32); rs = np.zeros(world, np.int64)
    rc = lib.lbchem_lb_get_plan(lb, C.byref(ns), st.ctypes.data_as(capi._ip), sa.ctypes.data_as(capi._dp),
                                bd.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(nr), rf.ctypes.data_as(capi._ip),
                                rs.ctypes.data_as(C.POINTER(C.c_int64)))
    if rc != 0:
        raise RuntimeError("lbchem_lb_get_plan failed")
    return dict(send=[(int(st[k]), float(sa[k])) for k in range(ns.value)], bounds=[int(b) for b in bd[:ns.value + 1]],
                recv=[int(rf[k]) for k in range(nr.value)], recv_sizes=[int(rs[k]) for k in range(nr.value)])


class _StepBase:
    lib = None
    lb = None
    world = 1

    def _info(self):
        info = capi.LbInfo()
        self.lib.lbchem_lb_get_info(self.lb, C.byref(info))
        return info

    def _after(self):
        info = self._info()
        self.total_cost = float(info.total_cost)
        plan = _plan_of(self.lib, self.lb, self.world)
        self.send_plan, self.recv_plan = plan["send"], plan["recv"]
        self.last = dict(sent_cells=int(info.sent_cells), recv_cells=int(info.recv_cells), sent_bytes=int(info.sent_bytes),
                         recv_bytes=int(info.recv_bytes), planned=bool(info.planned), exchanged=bool(info.exchanged),
                         reallocated=bool(info.reallocated), bounds=plan["bounds"], recv_sizes=plan["recv_sizes"])

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed ({rc}): {self.lib.lbchem_lb_last_error(self.lb).decode()}")

    def set_balance(self, on: bool):
        """Switch the exchange off (every step takes the local branch: the reference's non-balanced model) or back on."""
        self.options.balance = int(bool(on))
        self._check(self.lib.lbchem_lb_set_options(self.lb, C.byref(self.options)), "lbchem_lb_set_options")

    def close(self):
        if self.lb:
            self.lib.lbchem_lb_destroy(self.lb)
            self.lb = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class LoadBalancedChemistryModel(_StepBase):
    """One rank of the load-balanced chemistry model on its GPU. Fields are float64 CUDA tensors, field-major.

    ``transport``: "nccl" (default when torch.distributed is initialised with world > 1: one NCCL communicator owned by
    the library, unique id broadcast through ``group``), "single" (one rank), or an ``lbchem_lb`` handle made elsewhere
    (``make_loopback``: several ranks inside one process, device-to-device copies).
    """

    def __init__(self, chem: "capi.Chemistry", group=None, update_iter: int = 0, min_fraction: float = 0.02,
                 tdac: bool = False, balance: bool = True, transport="auto", _lb=None, _rank=0, _world=1):
        self.chem = chem
        self.lib = chem.lib
        self.nSp = chem.mech.n_species
        self.options = capi.lb_options(update_iter, min_fraction, tdac, balance)
        self.group = group
        self.first_time = True
        self.RR = self.deltaTChem = self.T_out = self.status = self.cell_cost = None
        self.total_cost = 0.0
        self.send_plan, self.recv_plan, self.last = [], [], {}
        if _lb is not None:
            self.lb, self.rank, self.world = _lb, _rank, _world
            return
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if transport == "auto":
            transport = "nccl"
        dev = torch.device("cuda", torch.cuda.current_device())
        idbuf = torch.zeros(128, dtype=torch.uint8, device=dev)
        if self.rank == 0:
            host = (C.c_ubyte * 128)()
            rc = self.lib.lbchem_lb_nccl_unique_id(C.cast(host, C.c_void_p))
            if rc != 0:
                raise RuntimeError(f"lbchem_lb_nccl_unique_id failed ({rc})")
            idbuf.copy_(torch.frombuffer(bytearray(host), dtype=torch.uint8))
        if self.world > 1:
            dist.broadcast(idbuf, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        idhost = (C.c_ubyte * 128).from_buffer_copy(bytes(idbuf.cpu().numpy().tobytes()))
        lb = C.c_void_p()
        rc = self.lib.lbchem_lb_create_nccl(chem.h, C.cast(idhost, C.c_void_p), self.rank, self.world, C.byref(self.options), C.byref(lb))
        if rc != 0:
            raise RuntimeError(f"lbchem_lb_create_nccl failed ({rc}): {self.lib.lbchem_last_error(chem.h).decode()}")
        self.lb = lb

    @staticmethod
    def make_loopback(chems, update_iter=0, min_fraction=0.02, tdac=False, balance=True):
        """``len(chems)`` ranks inside this process (call ``solve`` of every rank from its own thread, concurrently)."""
        lib = chems[0].lib
        n = len(chems)
        opts = capi.lb_options(update_iter, min_fraction, tdac, balance)
        hs = (C.c_void_p * n)(*[c.h for c in chems])
        out = (C.c_void_p * n)()
        rc = lib.lbchem_lb_create_loopback(n, hs, C.byref(opts), out)
        if rc != 0:
            raise RuntimeError(f"lbchem_lb_create_loopback failed ({rc}): {lib.lbchem_last_error(chems[0].h).decode()}")
        return [LoadBalancedChemistryModel(chems[r], update_iter=update_iter, min_fraction=min_fraction, tdac=tdac, balance=balance,
                                           _lb=C.c_void_p(out[r]), _rank=r, _world=n) for r in range(n)]

    def solve(self, Y, T, p, rho, deltaT, deltaTChem=None, out=None):
        """Integrate every local cell over ``deltaT`` (scalar, or a per-cell tensor for LTS). Returns the minimum chemical
        time step (reference .C:453-704). Results: ``self.RR [nSp, n]``, ``self.deltaTChem [n]``, ``self.T_out [n]``,
        ``self.cell_cost [n]``, ``self.status [n]`` in the caller's cell order. ``out``: optional dict of preallocated
        tensors (RR, deltaTChem, T, cost, status) to write into."""
        nSp, n = Y.shape
        assert nSp == self.nSp and Y.is_cuda and Y.stride(1) == 1
        dev = Y.device
        if deltaTChem is None:
            deltaTChem = self.deltaTChem
        o = out or {}
        dtc = o.get("deltaTChem")
        if dtc is None:
            dtc = deltaTChem.clone()
        elif dtc.data_ptr() != deltaTChem.data_ptr():
            dtc.copy_(deltaTChem)
        RR = o["RR"] if "RR" in o else torch.empty((nSp, n), dtype=torch.float64, device=dev)
        Tout = o["T"] if "T" in o else torch.empty(n, dtype=torch.float64, device=dev)
        cost = o["cost"] if "cost" in o else torch.empty(n, dtype=torch.float64, device=dev)
        status = o["status"] if "status" in o else torch.empty(n, dtype=torch.int32, device=dev)
        if torch.is_tensor(deltaT):
            dT_ptr, dT_u = C.c_void_p(deltaT.data_ptr()), 0.0
        else:
            dT_ptr, dT_u = None, float(deltaT)
        ptr = lambda t: C.c_void_p(t.data_ptr())
        mn = C.c_double()
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = self.lib.lbchem_lb_solve(self.lb, n, Y.stride(0) if n else 0, ptr(Y), ptr(T), ptr(p), ptr(rho), dT_ptr, dT_u,
                                      ptr(dtc), ptr(RR), ptr(Tout), ptr(cost), ptr(status), C.byref(mn), C.c_void_p(stream))
        self._check(rc, "lbchem_lb_solve")
        self.first_time = False
        self.RR, self.T_out, self.status, self.cell_cost = RR, Tout, status, cost
        self.deltaTChem = dtc
        self._after()
        return mn.value

    def solve_host(self, Y, T, p, rho, deltaT, deltaTChem, out):
        """The same step with HOST numpy arrays (pinned or pageable) through ``lbchem_lb_solve_host``: H2D, step, D2H.
        ``out``: dict of numpy arrays RR, T, cost, status, deltaTChem (deltaTChem is filled with the input first)."""
        nSp, n = Y.shape
        out["deltaTChem"][...] = deltaTChem
        vp = lambda a: C.c_void_p(a.ctypes.data)
        if np.ndim(deltaT) == 0:
            dT_ptr, dT_u = None, float(deltaT)
        else:
            dT_ptr, dT_u = vp(deltaT), 0.0
        mn = C.c_double()
        rc = self.lib.lbchem_lb_solve_host(self.lb, n, n, vp(Y), vp(T), vp(p), vp(rho), dT_ptr, dT_u, vp(out["deltaTChem"]),
                                           vp(out["RR"]), vp(out["T"]), vp(out["cost"]), vp(out["status"]), C.byref(mn))
        self._check(rc, "lbchem_lb_solve_host")
        self._after()
        return mn.value


# ------------------------------------------------------------------------------------------------------------------
# CPU test build: the same step logic over host memory (tests only)
# ------------------------------------------------------------------------------------------------------------------
class HostStepModel(_StepBase):
    """``lbchem_lb_core.hpp`` compiled for the host (``liblbchem_lbhost.so``). ``cell_solver(Y, T, p, rho, dT, dtc)`` is a
    deterministic stand-in for solveCellList on CPU torch tensors returning ``(RR, Tout, cost, status)`` and updating
    ``dtc`` in place; the transport is torch.distributed (gloo) behind the library's transport callbacks."""

    def __init__(self, n_species: int, cell_solver: Callable, group=None, update_iter: int = 0, min_fraction: float = 0.02,
                 tdac: bool = False, balance: bool = True):
        self.lib = capi.load_lbhost()
        self.nSp = int(n_species)
        self.cell_solver = cell_solver
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.options = capi.lb_options(update_iter, min_fraction, tdac, balance)
        self.RR = self.deltaTChem = self.T_out = self.status = self.cell_cost = None
        self.total_cost = 0.0
        self.send_plan, self.recv_plan, self.last = [], [], {}
        self.callback_error = None

        def view1(ptr, n, ctype=C.c_double):
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(n,))

        def view2(ptr, rows, n, ldy):                    # field-major block with leading dimension ldy
            flat = view1(ptr, (rows - 1) * ldy + n)
            return np.lib.stride_tricks.as_strided(flat, shape=(rows, n), strides=(ldy * 8, 8))

        def solver_cb(ctx, n, ldy, Y, T, p, rho, dT, dT_u, dtc, RR, Tout, cost, status):
            try:
                n = int(n); ldy = int(ldy)
                Yv = torch.from_numpy(np.array(view2(Y, self.nSp, n, ldy)))
                vec = lambda q: torch.from_numpy(view1(q, n))
                dTt = vec(dT).clone() if dT else torch.full((n,), float(dT_u), dtype=torch.float64)
                dtct = vec(dtc)                           # in/out: the stand-in solver updates it in place
                r = self.cell_solver(Yv, vec(T).clone(), vec(p).clone(), vec(rho).clone(), dTt, dtct)
                view2(RR, self.nSp, n, ldy)[...] = r[0].numpy()
                view1(Tout, n)[...] = r[1].numpy()
                view1(cost, n)[...] = r[2].numpy()
                view1(status, n, C.c_int32)[...] = r[3].numpy()
                return 0
            except Exception as e:                      # pragma: no cover - surfaced through the return code
                self.callback_error = e
                return 1

        def all_gather_cb(ctx, send, recv, count):
            try:
                mine = torch.from_numpy(np.ctypeslib.as_array(send, shape=(count,)).copy())
                parts = [torch.empty(count, dtype=torch.float64) for _ in range(self.world)]
                dist.all_gather(parts, mine, group=self.group)
                np.ctypeslib.as_array(recv, shape=(count * self.world,))[...] = torch.cat(parts).numpy()
                return 0
            except Exception as e:                      # pragma: no cover
                self.callback_error = e
                return 1

        def exchange_cb(ctx, nops, ops):
            try:
                reqs, keep = [], []
                for k in range(nops):
                    o = ops[k]
                    buf = torch.from_numpy(np.ctypeslib.as_array(C.cast(o.ptr, C.POINTER(C.c_uint8)), shape=(int(o.bytes),)))
                    keep.append(buf)
                    peer = dist.get_global_rank(self.group, o.peer) if self.group is not None else o.peer
                    reqs.append(dist.P2POp(dist.isend if o.is_send else dist.irecv, buf, peer, group=self.group))
                for r in (dist.batch_isend_irecv(reqs) if reqs else []):
                    r.wait()
                return 0
            except Exception as e:                      # pragma: no cover
                self.callback_error = e
                return 1

        self._cbs = (capi.HOST_SOLVER_FN(solver_cb), capi.ALL_GATHER_FN(all_gather_cb), capi.EXCHANGE_FN(exchange_cb))
        self._transport = capi.Transport(ctx=None, all_gather=self._cbs[1], exchange=self._cbs[2])
        lb = C.c_void_p()
        rc = self.lib.lbchem_lbhost_create(self.nSp, self._cbs[0], None, C.byref(self._transport), self.rank, self.world,
                                           C.byref(self.options), C.byref(lb))
        if rc != 0:
            raise RuntimeError(f"lbchem_lbhost_create failed ({rc})")
        self.lb = lb

    def solve(self, Y, T, p, rho, deltaT, deltaTChem=None):
        nSp, n = Y.shape
        if deltaTChem is None:
            deltaTChem = self.deltaTChem
        Y = Y.contiguous(); T = T.contiguous(); p = p.contiguous(); rho = rho.contiguous()
        dtc = deltaTChem.clone()
        RR = torch.empty((nSp, n), dtype=torch.float64)
        Tout = torch.empty(n, dtype=torch.float64); cost = torch.empty(n, dtype=torch.float64)
        status = torch.empty(n, dtype=torch.int32)
        if torch.is_tensor(deltaT):
            deltaT = deltaT.contiguous()
            dT_ptr, dT_u = C.c_void_p(deltaT.data_ptr()), 0.0
        else:
            dT_ptr, dT_u = None, float(deltaT)
        ptr = lambda t: C.c_void_p(t.data_ptr())
        mn = C.c_double()
        rc = self.lib.lbchem_lb_solve(self.lb, n, n, ptr(Y), ptr(T), ptr(p), ptr(rho), dT_ptr, dT_u, ptr(dtc), ptr(RR), ptr(Tout),
                                      ptr(cost), ptr(status), C.byref(mn), None)
        if rc != 0 and self.callback_error is not None:
            raise self.callback_error
        self._check(rc, "lbchem_lb_solve (host test build)")
        self.RR, self.T_out, self.status, self.cell_cost, self.deltaTChem = RR, Tout, status, cost, dtc
        self._after()
        return mn.value
