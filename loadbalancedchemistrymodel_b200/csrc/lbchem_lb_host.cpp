// lbchem_lb_host.cpp -- TEST BUILD of the load-balanced flow step: the step logic of lbchem_lb_core.hpp (the very code
// liblbchem.so runs on the device) over plain host memory, with a caller-supplied stand-in for solveCellList and a
// caller-supplied transport (gloo in tests/test_cpu_exchange.py). It contains no chemistry and is not part of the
// product library: it exists so that plan -> cellsToSend -> forward trip -> own tail, then received -> return trip ->
// unpack can be exercised with world_size > 1 on a machine without a GPU. Built as liblbchem_lbhost.so.
#include <cstdlib>

#include "lbchem_lb_capi.hpp"

extern "C" {
// stand-in for solveCellList on one field-major block (host pointers); returns 0 on success
typedef int (*lbchem_host_solver)(void* ctx, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p,
                                  const double* rho, const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR,
                                  double* Tout, double* cost, int32_t* status);
}

namespace {

struct HostBackend : lbcore::Backend {
    int nSp; lbchem_host_solver fn; void* ctx;
    HostBackend(int n, lbchem_host_solver f, void* c) : nSp(n), fn(f), ctx(c) {}
    int n_species() const override { return nSp; }
    void* alloc(size_t bytes) override { return std::malloc(bytes ? bytes : 8); }
    void release(void* p) override { std::free(p); }
    int copy(void* dst, const void* src, size_t bytes, void*) override { std::memmove(dst, src, bytes); return 0; }
    int sync(void*) override { return 0; }
    int call(const lbcore::Block& b, long long s, double dtU) {
        if (b.n - s <= 0) return 0;
        const int rc = fn(ctx, b.n - s, b.ldy, b.Y + s, b.T + s, b.p + s, b.rho + s, b.deltaT ? b.deltaT + s : nullptr, dtU,
                          b.deltaTChem + s, b.RR + s, b.Tout + s, b.cost + s, b.status + s);
        if (rc) err = "cell solver callback failed";
        return rc ? LBCHEM_E_INVALID : 0;
    }
    int solve(const lbcore::Block& a, long long aStart, const lbcore::Block* b, double dtU, void*) override {
        if (call(a, aStart, dtU)) return LBCHEM_E_INVALID;              // own tail first (reference .C:608)
        return b ? call(*b, 0, dtU) : 0;                               // then the received cells (.C:611)
    }
    int reduce(const double* cost, const double* dtc, long long n, double* total, double* dtMin, void*) override {
        double s = 0.0, m = 1e15;
        for (long long i = 0; i < n; ++i) { s += cost[i]; m = dtc[i] < m ? dtc[i] : m; }   // updateTotalCpuTime, .C:350-362
        *total = s; *dtMin = m;
        return 0;
    }
    int bounds(const double* cost, long long n, const double* targets, int nT, long long* out, void*) override {
        out[0] = 0;
        for (int k = 0; k < nT; ++k) out[k + 1] = lbchem_cells_to_send(cost, (int32_t)n, targets[k], (int32_t)out[k]);
        return 0;
    }
};

}  // namespace

extern "C" {

int lbchem_lbhost_create(int32_t n_species, lbchem_host_solver solver, void* solver_ctx, const lbchem_transport* t, int32_t rank,
                         int32_t nranks, const lbchem_lb_options* o, lbchem_lb* out) {
    if (!solver || !out || n_species < 1 || nranks < 1 || rank < 0 || rank >= nranks) return LBCHEM_E_INVALID;
    if (nranks > 1 && (!t || !t->all_gather || !t->exchange)) return LBCHEM_E_INVALID;
    lbchem_lb_options opt;
    if (o) opt = *o; else lbchem_lb_default_options(&opt);
    lbchem_transport none{};
    std::unique_ptr<lbcore::Transport> tp(new lbcore::CallbackTransport(t ? *t : none, rank, nranks));
    lbchem_lb lb = new lbchem_lb_t();
    lb->drv.reset(new lbcore::Driver(std::unique_ptr<lbcore::Backend>(new HostBackend(n_species, solver, solver_ctx)), std::move(tp), opt));
    *out = lb;
    return LBCHEM_OK;
}

int lbchem_lb_destroy(lbchem_lb lb) {
    if (!lb) return LBCHEM_E_INVALID;
    delete lb;
    return LBCHEM_OK;
}

int lbchem_lb_solve(lbchem_lb lb, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
                    const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR, double* Tout, double* cost,
                    int32_t* status, double* deltaTMin, void* stream) {
    if (!lb) return LBCHEM_E_INVALID;
    lbcore::Block b;
    b.n = n; b.ldy = ldy; b.Y = Y; b.T = T; b.p = p; b.rho = rho; b.deltaT = deltaT; b.deltaTChem = deltaTChem; b.RR = RR;
    b.Tout = Tout; b.cost = cost; b.status = status;
    return lb->drv->solve(b, deltaT_uniform, deltaTMin, stream);
}

}  // extern "C"
