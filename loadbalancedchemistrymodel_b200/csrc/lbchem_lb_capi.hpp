// lbchem_lb_capi.hpp -- the backend-independent half of the lbchem_lb_* C ABI (include/lbchem.h). Included by exactly one
// translation unit per library: lbchem_lb_cuda.cu (liblbchem.so) and lbchem_lb_host.cpp (the CPU test build).
#pragma once
#include "lbchem_lb_core.hpp"

extern "C" {

void lbchem_lb_default_options(lbchem_lb_options* o) {
    if (!o) return;
    std::memset(o, 0, sizeof(*o));
    o->update_iter = 0;          // LoadBalancedCoeffs defaults, reference LoadBalancedChemistryModel.C:41-45
    o->min_fraction = 0.02;
    o->tdac = 0;
    o->balance = 1;
}

int lbchem_lb_set_options(lbchem_lb lb, const lbchem_lb_options* o) {
    if (!lb || !o || o->update_iter < 0 || !(o->min_fraction >= 0)) return LBCHEM_E_INVALID;
    lb->drv->set_options(*o);
    return LBCHEM_OK;
}

const char* lbchem_lb_last_error(lbchem_lb lb) {
    if (!lb) return "null lbchem_lb";
    return lb->err.empty() ? lb->drv->error().c_str() : lb->err.c_str();
}

int lbchem_lb_get_info(lbchem_lb lb, lbchem_lb_info* out) {
    if (!lb || !out) return LBCHEM_E_INVALID;
    *out = lb->drv->info();
    return LBCHEM_OK;
}

int lbchem_lb_get_plan(lbchem_lb lb, int32_t* nsend, int32_t* send_to, double* send_amount, int64_t* bounds,
                       int32_t* nrecv, int32_t* recv_from, int64_t* recv_sizes) {
    if (!lb) return LBCHEM_E_INVALID;
    const lbcore::Driver& d = *lb->drv;
    const size_t ns = d.send_to().size(), nr = d.recv_from().size();
    if (nsend) *nsend = (int32_t)ns;
    if (nrecv) *nrecv = (int32_t)nr;
    for (size_t k = 0; k < ns; ++k) {
        if (send_to) send_to[k] = d.send_to()[k];
        if (send_amount) send_amount[k] = d.send_amount()[k];
    }
    if (bounds) for (size_t k = 0; k <= ns; ++k) bounds[k] = k < d.bounds().size() ? d.bounds()[k] : 0;
    for (size_t k = 0; k < nr; ++k) {
        if (recv_from) recv_from[k] = d.recv_from()[k];
        if (recv_sizes) recv_sizes[k] = k < d.recv_sizes().size() ? d.recv_sizes()[k] : 0;
    }
    return LBCHEM_OK;
}

}  // extern "C"
