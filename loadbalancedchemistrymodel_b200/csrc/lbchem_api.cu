// lbchem_api.cu — C ABI of include/lbchem.h: mechanism flattening, option handling, batch solve on
// device or host buffers, probes. No CPU fallback: every compute entry point needs a CUDA device.
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "lbchem_internal.hpp"

int lb_fail(lbchem_handle h, int code, const std::string& msg) {
    if (h) h->err = msg;
    return code;
}

namespace {

inline int fail(lbchem_handle h, int code, const std::string& msg) { return lb_fail(h, code, msg); }

template <class T>
int upload(lbchem_handle h, const std::vector<T>& v, const T** out) {
    void* d = nullptr;
    const size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    CK(cudaMalloc(&d, bytes));
    h->allocs.push_back(d);
    if (!v.empty()) CK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    *out = static_cast<const T*>(d);
    return LBCHEM_OK;
}

#define UP(vec, field)                                              \
    do {                                                            \
        int rc_ = upload(h, vec, &h->M.field);                      \
        if (rc_ != LBCHEM_OK) return rc_;                           \
    } while (0)

// species-major SoA: coefficient q of species i at [q*nSp + i]
std::vector<double> soa(const double* a, int nSp, const int* idx, const double* div, int nq) {
    std::vector<double> v((size_t)nq * nSp);
    for (int i = 0; i < nSp; ++i)
        for (int q = 0; q < nq; ++q) v[(size_t)q * nSp + i] = a[7 * i + idx[q]] / div[q];
    return v;
}

int build_mechanism(lbchem_handle h, const lbchem_mech_desc* m) {
    const int nSp = m->n_species, nR = m->n_reactions, nEff = m->n_eff;
    if (nSp < 1 || nR < 0 || nEff < 0) return fail(h, LBCHEM_E_INVALID, "bad mechanism sizes");
    DevMech& M = h->M;
    M.nSp = nSp; M.nR = nR; M.nEqn = nSp + 2; M.nEff = nEff;
    M.lda = (M.nEqn & 1) ? M.nEqn : M.nEqn + 1;

    std::vector<double> W(m->W, m->W + nSp), Tc(m->Tcommon, m->Tcommon + nSp);
    double tlo = m->Tlow[0], thi = m->Thigh[0];
    for (int i = 1; i < nSp; ++i) { tlo = std::max(tlo, m->Tlow[i]); thi = std::min(thi, m->Thigh[i]); }
    M.TlowMix = tlo; M.ThighMix = thi;
    {
        const int cpI[5] = {0, 1, 2, 3, 4}; const double cpD[5] = {1, 1, 1, 1, 1};
        const int haI[6] = {4, 3, 2, 1, 0, 5}; const double haD[6] = {5, 4, 3, 2, 1, 1};
        const int sI[6] = {4, 3, 2, 1, 0, 6};  const double sD[6] = {4, 3, 2, 1, 1, 1};
        // cp Horner reads a[4*st]..a[0]: stored in natural order a0..a4
        UP(W, W); UP(Tc, Tc);
        UP(soa(m->janaf_low, nSp, cpI, cpD, 5), cpLo); UP(soa(m->janaf_high, nSp, cpI, cpD, 5), cpHi);
        UP(soa(m->janaf_low, nSp, haI, haD, 6), haLo); UP(soa(m->janaf_high, nSp, haI, haD, 6), haHi);
        UP(soa(m->janaf_low, nSp, sI, sD, 6), sLo);    UP(soa(m->janaf_high, nSp, sI, sD, 6), sHi);
    }

    std::vector<int> rtype(m->rtype, m->rtype + nR), rev(m->reversible, m->reversible + nR), effIdx(m->eff_index, m->eff_index + nR);
    std::vector<double> A(nR), beta(nR), Ta(nR), A0(nR), beta0(nR), Ta0(nR), troe((size_t)4 * nR), rTc(nR), dn(nR);
    std::vector<int> nl(nR), nr(nR), lsp((size_t)LB_MAXSIDE * nR, 0), rsp((size_t)LB_MAXSIDE * nR, 0), dbase(nR + 1, 0);
    std::vector<double> lnu((size_t)LB_MAXSIDE * nR, 0.0), rnu(lnu), lex(lnu), rex(lnu);
    int fracExp = 0;
    for (int r = 0; r < nR; ++r) {
        A[r] = m->arr[3 * r]; beta[r] = m->arr[3 * r + 1]; Ta[r] = m->arr[3 * r + 2];
        A0[r] = m->arr0[3 * r]; beta0[r] = m->arr0[3 * r + 1]; Ta0[r] = m->arr0[3 * r + 2];
        for (int q = 0; q < 4; ++q) troe[(size_t)q * nR + r] = m->troe[4 * r + q];
        if (rtype[r] != LBCHEM_R_ARRHENIUS && (effIdx[r] < 0 || effIdx[r] >= nEff))
            return fail(h, LBCHEM_E_INVALID, "reaction needs third-body efficiencies but eff_index is out of range");
        nl[r] = m->lhs_ptr[r + 1] - m->lhs_ptr[r];
        nr[r] = m->rhs_ptr[r + 1] - m->rhs_ptr[r];
        if (nl[r] < 1 || nr[r] < 1 || nl[r] > LB_MAXSIDE || nr[r] > LB_MAXSIDE)
            return fail(h, LBCHEM_E_UNSUPPORTED, "reaction with 0 or more than 4 species on one side");
        double d = 0;
        for (int s = 0; s < nl[r]; ++s) {
            const int k = m->lhs_ptr[r] + s;
            if (m->lhs_species[k] < 0 || m->lhs_species[k] >= nSp) return fail(h, LBCHEM_E_INVALID, "species index out of range");
            lsp[(size_t)s * nR + r] = m->lhs_species[k]; lnu[(size_t)s * nR + r] = m->lhs_nu[k]; lex[(size_t)s * nR + r] = m->lhs_exp[k];
            d -= m->lhs_nu[k];
            const double e = m->lhs_exp[k];
            if (!(e == 1.0 || e == 2.0 || e == 3.0)) fracExp = 1;
        }
        for (int s = 0; s < nr[r]; ++s) {
            const int k = m->rhs_ptr[r] + s;
            if (m->rhs_species[k] < 0 || m->rhs_species[k] >= nSp) return fail(h, LBCHEM_E_INVALID, "species index out of range");
            rsp[(size_t)s * nR + r] = m->rhs_species[k]; rnu[(size_t)s * nR + r] = m->rhs_nu[k]; rex[(size_t)s * nR + r] = m->rhs_exp[k];
            d += m->rhs_nu[k];
            const double e = m->rhs_exp[k];
            if (!(e == 1.0 || e == 2.0 || e == 3.0)) fracExp = 1;
        }
        // dn accumulated like the oracle: rhs first, then lhs
        double dn2 = 0;
        for (int s = 0; s < nr[r]; ++s) dn2 += m->rhs_nu[m->rhs_ptr[r] + s];
        for (int s = 0; s < nl[r]; ++s) dn2 -= m->lhs_nu[m->lhs_ptr[r] + s];
        (void)d;
        dn[r] = dn2;
        rTc[r] = m->Tcommon[m->lhs_species[m->lhs_ptr[r]]];
        dbase[r + 1] = dbase[r] + nl[r] + nr[r];
    }
    M.fracExp = fracExp;
    M.nTerms = dbase[nR];
    UP(rtype, rtype); UP(rev, rev); UP(effIdx, effIdx);
    UP(A, A); UP(beta, beta); UP(Ta, Ta); UP(A0, A0); UP(beta0, beta0); UP(Ta0, Ta0);
    UP(troe, troe); UP(rTc, rTc); UP(dn, dn);
    UP(std::vector<double>(m->eff, m->eff + (size_t)nEff * nSp), eff);
    UP(nl, nl); UP(nr, nr); UP(lsp, lsp); UP(rsp, rsp); UP(lnu, lnu); UP(rnu, rnu); UP(lex, lex); UP(rex, rex);
    UP(dbase, dbase);

    // omega gather lists (species-major, reaction order, lhs before rhs)
    std::vector<int> spPtr(nSp + 1, 0), spReac; std::vector<double> spCoef;
    {
        std::vector<std::vector<std::pair<int, double>>> L(nSp);
        for (int r = 0; r < nR; ++r) {
            for (int s = 0; s < nl[r]; ++s) L[lsp[(size_t)s * nR + r]].push_back({r, -lnu[(size_t)s * nR + r]});
            for (int s = 0; s < nr[r]; ++s) L[rsp[(size_t)s * nR + r]].push_back({r, rnu[(size_t)s * nR + r]});
        }
        for (int i = 0; i < nSp; ++i) {
            for (auto& e : L[i]) { spReac.push_back(e.first); spCoef.push_back(e.second); }
            spPtr[i + 1] = (int)spReac.size();
        }
    }
    UP(spPtr, spPtr); UP(spReac, spReac); UP(spCoef, spCoef);

    h->H.nSp = nSp; h->H.nR = nR; h->H.nEff = nEff; h->H.nl = nl; h->H.nr = nr; h->H.lsp = lsp; h->H.rsp = rsp;
    h->H.lnu = lnu; h->H.rnu = rnu; h->H.dbase = dbase; h->H.nTerms = M.nTerms;
    return LBCHEM_OK;
}

// Jacobian gather lists: entries sorted by (row, col), terms in the reference's accumulation order,
// entries dealt to the tpc threads in contiguous runs balanced by term count.
int build_jacobian_lists(lbchem_handle h, int tpc) {
    const HostMech& H = h->H;
    const int nR = H.nR;
    std::map<std::pair<int, int>, std::vector<std::pair<int, double>>> E;
    for (int r = 0; r < nR; ++r) {
        const int nl = H.nl[r], nr = H.nr[r];
        for (int j = 0; j < nl; ++j) {
            const int col = H.lsp[(size_t)j * nR + r], term = H.dbase[r] + j;
            for (int i = 0; i < nl; ++i) E[{H.lsp[(size_t)i * nR + r], col}].push_back({term, -H.lnu[(size_t)i * nR + r]});
            for (int i = 0; i < nr; ++i) E[{H.rsp[(size_t)i * nR + r], col}].push_back({term, H.rnu[(size_t)i * nR + r]});
        }
        for (int j = 0; j < nr; ++j) {
            const int col = H.rsp[(size_t)j * nR + r], term = H.dbase[r] + nl + j;
            for (int i = 0; i < nl; ++i) E[{H.lsp[(size_t)i * nR + r], col}].push_back({term, H.lnu[(size_t)i * nR + r]});
            for (int i = 0; i < nr; ++i) E[{H.rsp[(size_t)i * nR + r], col}].push_back({term, -H.rnu[(size_t)i * nR + r]});
        }
    }
    std::vector<int> entRow, entCol, entPtr{0}, termIdx; std::vector<double> termCoef;
    for (auto& kv : E) {
        entRow.push_back(kv.first.first); entCol.push_back(kv.first.second);
        for (auto& t : kv.second) { termIdx.push_back(t.first); termCoef.push_back(t.second); }
        entPtr.push_back((int)termIdx.size());
    }
    const int nEnt = (int)entRow.size();
    std::vector<int> thrPtr(tpc + 1, nEnt);
    {
        // cost of an entry = its terms + 2 (store and loop overhead)
        const double total = (double)termIdx.size() + 2.0 * nEnt;
        double acc = 0; int t = 0; thrPtr[0] = 0;
        for (int e = 0; e < nEnt; ++e) {
            while (t + 1 < tpc && acc >= total * (t + 1) / tpc) thrPtr[++t] = e;
            acc += (entPtr[e + 1] - entPtr[e]) + 2.0;
        }
        while (t + 1 <= tpc) thrPtr[++t] = nEnt;
    }
    h->M.nEntries = nEnt;
    UP(thrPtr, thrPtr); UP(entRow, entRow); UP(entCol, entCol); UP(entPtr, entPtr); UP(termIdx, termIdx); UP(termCoef, termCoef);
    return LBCHEM_OK;
}

int configure_v2(lbchem_handle h) {
    h->useV2 = false; h->useV3 = false;
    const lbchem_options& o = h->opts;
    if (!h->v2Built || o.threads_per_cell != 0 || getenv("LBCHEM_GENERAL_KERNEL")) return LBCHEM_OK;
    if (o.solver != LBCHEM_SOLVER_SEULEX && o.solver != LBCHEM_SOLVER_ROSENBROCK34) return LBCHEM_OK;
    // small mechanisms (H2/O2: 12 equations, 1-step CH4: 7) run the same compact-code kernel with 16 / 8 threads per cell
    // (LBCHEM_SMALL_KERNEL=general selects the general kernel with sub-warp groups instead, for A/B runs)
    {
        const char* sk = getenv("LBCHEM_SMALL_KERNEL");
        if (sk && !strcmp(sk, "general") && h->M.nEqn <= 16) return LBCHEM_OK;
    }
    if (lb_v3_supported(h->V2)) {
        const int ws = lb_v3_workspace(h->V2, o.solver, o.rel_tol, o.abs_tol);
        const size_t per3 = (size_t)ws * sizeof(double);
        const size_t fixed = (size_t)lb_v3_ctx_bytes() + h->V2.hotBytes;
        if (fixed + per3 <= h->smemMax) {
            int groups = (int)std::min<size_t>(lb_v3_max_groups(h->V2.tpc), (h->smemMax - fixed) / per3);
            if (const char* env = getenv("LBCHEM_GROUPS")) if (atoi(env) > 0) groups = std::min(groups, atoi(env));
            h->v3Groups = groups; h->v3Ws = ws;
            h->scratchStride = (h->V2.nEnt + 14 * h->V2.nEqn + h->V2.nSp + 1) & ~1;   // sparse Jacobian, seulex table, owner-only vectors, dT column
            const size_t need = (size_t)h->nSM * groups * h->scratchStride * sizeof(double);
            if (need > h->scratchCap) {
                if (h->dScratch) cudaFree(h->dScratch);
                h->dScratch = nullptr; h->scratchCap = 0;
                CK(cudaMalloc(&h->dScratch, need));
                h->scratchCap = need;
            }
            h->v3Smem = fixed + per3 * groups;
            CK(lb_v3_configure(h->V2.tpc, o.solver, h->v3Smem));
            if (getenv("LBCHEM_VERBOSE") && h->V2.tpc == 64) {
                int r = 0, mt = 0, ss = 0;
                lb_v3_kernel_info(o.solver, &r, &mt, &ss);
                fprintf(stderr, "[lbchem] v3 kernel attributes: %d registers, maxThreadsPerBlock %d, static smem %d B\n", r, mt, ss);
            }
            if (getenv("LBCHEM_VERBOSE"))
                fprintf(stderr, "[lbchem] v3 kernel: %d threads per cell, %d cells per CTA, %d B workspace per cell, %d B mechanism in smem, %zu B dynamic smem\n",
                        h->V2.tpc, groups, (int)per3, h->V2.hotBytes, h->v3Smem);
            h->useV3 = true;
            return LBCHEM_OK;
        }
    }
    return LBCHEM_OK;
}

int configure(lbchem_handle h) {
    {
        int rc = configure_v2(h);
        if (rc != LBCHEM_OK) return rc;
    }
    int tpc = h->opts.threads_per_cell;
    if (tpc == 0) {
        tpc = h->M.nEqn <= 8 ? 8 : (h->M.nEqn <= 16 ? 16 : (h->M.nEqn <= 32 ? 32 : 64));
        if (const char* e = getenv("LBCHEM_TPC")) if (atoi(e) > 0) tpc = atoi(e);
    }
    if (tpc != 8 && tpc != 16 && tpc != 32 && tpc != 64) return fail(h, LBCHEM_E_INVALID, "threads_per_cell must be 0, 8, 16, 32 or 64");
    if (h->configuredSolver == h->opts.solver && h->configuredTpc == tpc) return LBCHEM_OK;
    if (h->configuredTpc != tpc) {
        int rc = build_jacobian_lists(h, tpc);
        if (rc != LBCHEM_OK) return rc;
    }
    h->tpc = tpc;
    h->wsDoubles = lb_work_doubles(h->M, h->opts.solver);
    const size_t per = h->wsDoubles * sizeof(double);
    if (per > h->smemMax) return fail(h, LBCHEM_E_UNSUPPORTED, "mechanism too large for one cell's shared-memory workspace");
    int groups = (int)std::min<size_t>(tpc < 32 ? 16 : 8, h->smemMax / per);
    const char* env = getenv("LBCHEM_GROUPS");
    if (env && atoi(env) > 0) groups = std::min(groups, atoi(env));
    h->groups = groups;
    h->smemBytes = per * groups;
    int ctas = 0;
    CK(lb_configure(tpc, groups, h->smemBytes, &ctas));
    if (ctas < 1) return fail(h, LBCHEM_E_CUDA, "integration kernel does not fit on an SM");
    h->ctasPerSM = ctas;
    h->configuredSolver = h->opts.solver; h->configuredTpc = tpc;
    return LBCHEM_OK;
}

void set_devopts(lbchem_handle h) {
    const lbchem_options& o = h->opts;
    DevOpts& d = h->O;
    d.solver = o.solver; d.maxSteps = o.max_steps; d.absTol = o.abs_tol; d.relTol = o.rel_tol; d.Treact = o.Treact;
    d.deltaTChemMax = o.deltaTChemMax; d.cTauChem = o.cTauChem; d.eqRateLimiter = o.eq_rate_limiter;
    d.dfdxIsDcdt = o.dfdx_is_dcdt; d.RR = o.gas_constant; d.kcFloor = o.kc_floor;
    d.seulexErrOldMax = o.seulex_errold_max; d.pivoting = o.pivoting;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

const char* lbchem_version(void) { return "lbchem-b200 0.1 (sm_100a)"; }

void lbchem_default_options(lbchem_options* o) {
    if (!o) return;
    memset(o, 0, sizeof(*o));
    o->solver = LBCHEM_SOLVER_ROSENBROCK34;
    o->max_steps = 10000;
    o->abs_tol = 1e-15;          // ODESolver: absTol default SMALL
    o->rel_tol = 1e-4;
    o->Treact = 0.0;
    o->deltaTChemMax = 1e15;     // GREAT
    o->cTauChem = 1.0;
    o->eq_rate_limiter = 0;
    o->dfdx_is_dcdt = 1;
    o->gas_constant = 1e3 * (6.0221417930e23 * 1.3806504e-23);
    o->kc_floor = 1e-6;
    o->threads_per_cell = 0;
    o->sort_by_cost = 0;
    o->seulex_errold_max = 0;
    o->cost_model = 0;
    o->pivoting = 1;
}

int lbchem_create(const lbchem_mech_desc* mech, int device, lbchem_handle* out) {
    if (!mech || !out) return LBCHEM_E_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return LBCHEM_E_CUDA;   // no CPU fallback
    lbchem_handle h = new lbchem_handle_t();
    *out = h;   // returned even on failure so that lbchem_last_error can be read; caller destroys it
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) return fail(h, LBCHEM_E_CUDA, "cudaGetDevice failed"); }
    CK(cudaSetDevice(device));
    h->device = device;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    h->nSM = prop.multiProcessorCount;
    h->smemMax = prop.sharedMemPerBlockOptin;
    lbchem_default_options(&h->opts);
    int rc = build_mechanism(h, mech);
    if (rc != LBCHEM_OK) return rc;
    {
        std::string e;
        h->hfOverR.resize(mech->n_species); h->Wsp.assign(mech->W, mech->W + mech->n_species);
        for (int i = 0; i < mech->n_species; ++i) {
            // janafThermo::hc(): the low-range enthalpy polynomial at Tstd = 298.15 K
            const double* a = mech->janaf_low + 7 * i; const double Ts = 298.15;
            h->hfOverR[i] = ((((a[4] / 5.0 * Ts + a[3] / 4.0) * Ts + a[2] / 3.0) * Ts + a[1] / 2.0) * Ts + a[0]) * Ts + a[5];
        }
        {
            const int32_t seed = 0;
            if (lb_dac_program(mech, 1e-4, 1, &seed, h->dacProg, &h->dacPairs) != 0) h->dacProg.clear();
        }
        const int r2 = lb_v2_build(mech, &h->V2, h->allocs, e, &h->v3PosOf);
        if (r2 < 0) return fail(h, r2, e);
        h->v2Built = (r2 == 0);
    }
    CK(cudaMalloc(&h->dQueue, sizeof(unsigned long long)));
    CK(cudaMalloc(&h->dStats, 8 * sizeof(unsigned long long)));
    CK(cudaMemset(h->dStats, 0, 8 * sizeof(unsigned long long)));
    CK(cudaEventCreate(&h->ev0));
    CK(cudaEventCreate(&h->ev1));
    set_devopts(h);
    return configure(h);
}

int lbchem_destroy(lbchem_handle h) {
    if (!h) return LBCHEM_E_INVALID;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    for (void* p : h->allocs) cudaFree(p);
    if (h->dQueue) cudaFree(h->dQueue);
    if (h->dStats) cudaFree(h->dStats);
    if (h->dScratch) cudaFree(h->dScratch);
    if (h->dDac) cudaFree(h->dDac);
    if (h->dHf) cudaFree(h->dHf);
    if (h->dNActive) cudaFree(h->dNActive);
    if (h->dOrder) cudaFree(h->dOrder);
    if (h->dProbeIn) cudaFree(h->dProbeIn);
    if (h->dProbeOut) cudaFree(h->dProbeOut);
    if (h->dHist) cudaFree(h->dHist);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    for (Slot& s : h->slots) {
        void* ptrs[] = {s.Y, s.RR, s.T, s.p, s.rho, s.dT, s.dtc, s.Tout, s.cost, s.status, s.queue, s.scratch};
        for (void* p : ptrs) if (p) cudaFree(p);
        if (s.k0) cudaEventDestroy(s.k0);
        if (s.k1) cudaEventDestroy(s.k1);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    delete h;
    return LBCHEM_OK;
}

int lbchem_set_options(lbchem_handle h, const lbchem_options* o) {
    if (!h || !o) return LBCHEM_E_INVALID;
    if (o->solver < LBCHEM_SOLVER_NONE || o->solver > LBCHEM_SOLVER_SEULEX) return fail(h, LBCHEM_E_INVALID, "unknown solver");
    if (!(o->abs_tol >= 0) || !(o->rel_tol > 0) || o->max_steps < 1 || !(o->gas_constant > 0))
        return fail(h, LBCHEM_E_INVALID, "bad tolerances / maxSteps / gas constant");
    CK(cudaSetDevice(h->device));
    h->opts = *o;
    set_devopts(h);
    return configure(h);
}

int lbchem_get_options(lbchem_handle h, lbchem_options* o) {
    if (!h || !o) return LBCHEM_E_INVALID;
    *o = h->opts;
    return LBCHEM_OK;
}

const char* lbchem_last_error(lbchem_handle h) { return h ? h->err.c_str() : "null handle"; }

int lbchem_set_reduction(lbchem_handle h, int32_t active, double tolerance, int32_t n_init, const int32_t* init_species) {
    if (!h) return LBCHEM_E_INVALID;
    CK(cudaSetDevice(h->device));
    if (!active) { h->dacOn = false; h->V2.dac = nullptr; return configure(h); }     // the workspace layout depends on it
    if (h->dacProg.empty() || !h->v2Built || h->V2.tpc != 64)
        return fail(h, LBCHEM_E_UNSUPPORTED, "DAC reduction needs the packed mechanism of the compact-code kernel (nSp <= 62, integer stoichiometry)");
    if (!(tolerance > 0.0) || n_init < 1 || n_init > 8 || !init_species)
        return fail(h, LBCHEM_E_INVALID, "reduction: tolerance must be > 0 and the initial set must hold 1..8 species");
    for (int i = 0; i < n_init; ++i)
        if (init_species[i] < 0 || init_species[i] >= h->M.nSp) return fail(h, LBCHEM_E_INVALID, "reduction: initial-set species out of range");
    {
        V2Mech probe = h->V2;
        lb_v3_workspace(probe, LBCHEM_SOLVER_ROSENBROCK34, 1e-2, 1e-12);
        const int cap = std::max(h->V2.nSp * (h->V2.nSp + 1), 2 * h->V2.nR);
        if (h->dacPairs + (h->dacPairs + 7) / 8 > cap) return fail(h, LBCHEM_E_UNSUPPORTED, "reduction: species-pair table does not fit the per-cell workspace");
        // the enabled-reaction list may live in the tail of the matrix region when nothing else reaches it: the largest
        // reduced matrix (nSp - 1 species), KF / KR + the Jacobian's D terms, the pair numerators of the reduction pass
        const int nSp = h->V2.nSp, lda = (nSp + 1) | 1, region = std::max(nSp * lda, 2 * h->V2.nR + h->V2.nD);
        const int tail = region - (1 + (h->V2.nR + 3) / 4);
        const int redMat = (nSp - 1) * (nSp | 1);
        h->V2.rlInMatrix = (redMat <= tail && 2 * h->V2.nR + h->V2.nD <= tail && h->dacPairs + (h->dacPairs + 7) / 8 <= tail) ? 1 : 0;
        if (getenv("LBCHEM_RL_SEPARATE")) h->V2.rlInMatrix = 0;
    }
    h->dacProg[7] = n_init;
    for (int i = 0; i < 8; ++i) h->dacProg[8 + i] = i < n_init ? init_species[i] : 0;
    memcpy(&h->dacProg[16], &tolerance, 8);
    if (!h->dDac) CK(cudaMalloc(&h->dDac, h->dacProg.size() * sizeof(int)));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h->dDac, h->dacProg.data(), h->dacProg.size() * sizeof(int), cudaMemcpyHostToDevice));
    h->V2.dac = h->dDac;
    h->dacOn = true;
    return configure(h);        // the reduced kernel's workspace holds the reaction masks and the enabled-reaction list
}

int lbchem_qdot(lbchem_handle h, int64_t n, int64_t ldy, const double* RR, double* Qdot, void* stream) {
    if (!h || n < 0 || ldy < n || (n > 0 && (!RR || !Qdot))) return h ? fail(h, LBCHEM_E_INVALID, "bad Qdot arguments") : LBCHEM_E_INVALID;
    CK(cudaSetDevice(h->device));
    const int nSp = h->M.nSp;
    if (!h->dHf) CK(cudaMalloc(&h->dHf, sizeof(double) * nSp));
    if (h->hfGas != h->opts.gas_constant) {
        std::vector<double> hf(nSp);
        for (int i = 0; i < nSp; ++i) hf[i] = h->opts.gas_constant / h->Wsp[i] * h->hfOverR[i];
        CK(cudaMemcpy(h->dHf, hf.data(), sizeof(double) * nSp, cudaMemcpyHostToDevice));
        h->hfGas = h->opts.gas_constant;
    }
    if (n > 0) CK(lb_launch_qdot(n, ldy, nSp, h->dHf, RR, Qdot, static_cast<cudaStream_t>(stream)));
    return LBCHEM_OK;
}

int lbchem_tc(lbchem_handle h, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
              double* tc, void* stream) {
    if (!h) return LBCHEM_E_INVALID;
    if (n < 0 || ldy < n || (n > 0 && (!Y || !T || !p || !rho || !tc))) return fail(h, LBCHEM_E_INVALID, "bad tc arguments");
    CK(cudaSetDevice(h->device));
    if (!h->useV3)
        return fail(h, LBCHEM_E_UNSUPPORTED, "tc() runs on the compact-code kernel's rate sweep (solver seulex or Rosenbrock34, nEqn <= 64)");
    if (n == 0) return LBCHEM_OK;
    BatchArgs B{};
    B.a.n = n; B.a.ldy = ldy; B.a.Y = Y; B.a.T = T; B.a.p = p; B.a.rho = rho;
    B.scratch = h->dScratch; B.scratchStride = h->scratchStride;
    long long grid = h->nSM;
    const long long need = (n + h->v3Groups - 1) / h->v3Groups;
    if (grid > need) grid = need;
    CK(lb_v3_tc(h->V2.tpc, (int)grid, h->v3Groups, h->v3Smem, h->v3Ws, h->V2, h->O, B, tc, static_cast<cudaStream_t>(stream)));
    return LBCHEM_OK;
}

int lbchem_get_active_species(lbchem_handle h, int64_t n, int32_t* n_active) {
    if (!h || !n_active || n < 0) return LBCHEM_E_INVALID;
    CK(cudaSetDevice(h->device));
    if (!h->dacOn || !h->dNActive || n > h->lastN || n > h->nActiveCap)
        return fail(h, LBCHEM_E_INVALID, "no reduced batch of that size has been solved");
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(n_active, h->dNActive, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost));
    return LBCHEM_OK;
}

static int reserve_nactive(lbchem_handle h, long long n) {
    if (!h->dacOn || n <= h->nActiveCap) return LBCHEM_OK;
    if (h->dNActive) cudaFree(h->dNActive);
    h->dNActive = nullptr; h->nActiveCap = 0;
    CK(cudaMalloc(&h->dNActive, (size_t)n * sizeof(int)));
    h->nActiveCap = n;
    return LBCHEM_OK;
}

}  // extern "C"

int lb_launch_batch(lbchem_handle h, const CellSeg* a, const CellSeg* b, double deltaTUniform, const int* order,
                    unsigned long long* queue, double* scratch, cudaStream_t s) {
    BatchArgs B{};
    B.a = *a;
    if (b) B.b = *b;
    const long long n = B.a.n + B.b.n;
    B.deltaTUniform = deltaTUniform; B.order = order;
    B.queue = queue; B.stats = h->dStats; B.scratch = scratch; B.scratchStride = h->scratchStride;
    B.costModel = h->opts.cost_model;
    if (n <= 0) return LBCHEM_OK;
    CK(cudaMemsetAsync(queue, 0, sizeof(unsigned long long), s));
    if (h->dacOn && !h->useV3)
        return lb_fail(h, LBCHEM_E_UNSUPPORTED, "DAC reduction runs in the compact-code kernel only (solver seulex or Rosenbrock34, nEqn <= 64)");
    if (h->useV3) {
        long long grid3 = h->nSM;
        const long long need3 = (n + h->v3Groups - 1) / h->v3Groups;
        if (grid3 > need3) grid3 = need3;
        CK(lb_v3_launch(h->V2.tpc, h->opts.solver, (int)grid3, h->v3Groups, h->v3Smem, h->v3Ws, h->V2, h->O, B, s));
        h->launches++;
        return LBCHEM_OK;
    }
    // general kernel: one segment per launch, no dispatch order
    for (int seg = 0; seg < 2; ++seg) {
        BatchArgs B1 = B;
        if (seg == 1) B1.a = B.b;
        B1.b = CellSeg{}; B1.order = nullptr;
        if (B1.a.n <= 0) continue;
        if (seg == 1) CK(cudaMemsetAsync(queue, 0, sizeof(unsigned long long), s));
        long long grid = (long long)h->nSM * h->ctasPerSM;
        const long long need = (B1.a.n + h->groups - 1) / h->groups;
        if (grid > need) grid = need;
        CK(lb_launch_integrate(h->tpc, (int)grid, h->groups, h->smemBytes, (int)h->wsDoubles, h->M, h->O, B1, s));
        h->launches++;
    }
    return LBCHEM_OK;
}

// sort_by_cost: counting sort of the queue positions by descending last cost (device), into the handle's order buffer
int lb_cost_order(lbchem_handle h, const double* lastCost, long long n, cudaStream_t s, const int** orderOut) {
    *orderOut = nullptr;
    if (!h->useV3 || !lastCost || n < 2 || n > 0x7fffffffLL) return LBCHEM_OK;
    if (n > h->orderCap) {
        if (h->dOrder) cudaFree(h->dOrder);
    if (h->dProbeIn) cudaFree(h->dProbeIn);
    if (h->dProbeOut) cudaFree(h->dProbeOut);
        h->dOrder = nullptr; h->orderCap = 0;
        CK(cudaMalloc(&h->dOrder, (size_t)n * sizeof(int)));
        h->orderCap = n;
    }
    if (!h->dHist) CK(cudaMalloc(&h->dHist, 4096 * sizeof(int)));
    CK(lb_launch_cost_order(lastCost, n, h->opts.cost_model ? 0.125 : 1.0, h->dHist, h->dOrder, s));
    h->launches += 3;
    *orderOut = h->dOrder;
    return LBCHEM_OK;
}

extern "C" {

static CellSeg make_seg(long long n, long long ldy, const double* Y, const double* T, const double* p, const double* rho,
                        const double* deltaT, double* deltaTChem, double* RR, double* Tout, double* cost, int32_t* status, int* nActive) {
    CellSeg a{};
    a.n = n; a.ldy = ldy; a.Y = Y; a.T = T; a.p = p; a.rho = rho; a.deltaT = deltaT; a.deltaTChem = deltaTChem; a.RR = RR;
    a.Tout = Tout; a.cost = cost; a.status = status; a.nActive = nActive;
    return a;
}

int lbchem_solve_batch(lbchem_handle h, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p,
                       const double* rho, const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR,
                       double* Tout, double* cost, int32_t* status, void* stream) {
    if (!h) return LBCHEM_E_INVALID;
    if (n < 0 || ldy < n || (n > 0 && (!Y || !T || !p || !rho || !deltaTChem || !RR)) || (!deltaT && !(deltaT_uniform > 0)))
        return fail(h, LBCHEM_E_INVALID, "bad batch arguments");
    CK(cudaSetDevice(h->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    h->launches = 0; h->hostPath = false; h->lastN = n;
    { int rc = reserve_nactive(h, n); if (rc != LBCHEM_OK) return rc; }
    h->nActiveOff = 0;
    CK(cudaMemsetAsync(h->dStats, 0, 8 * sizeof(unsigned long long), s));
    // sort_by_cost: `cost` holds the previous step's per-cell cost on entry (any finite values; zeros on a first step)
    const int* order = nullptr;
    if (h->opts.sort_by_cost && cost && n > 1) { int rc = lb_cost_order(h, cost, n, s, &order); if (rc != LBCHEM_OK) return rc; }
    CK(cudaEventRecord(h->ev0, s));
    if (n > 0) {
        const CellSeg a = make_seg(n, ldy, Y, T, p, rho, deltaT, deltaTChem, RR, Tout, cost, status,
                                   (h->dacOn && h->dNActive && n <= h->nActiveCap) ? h->dNActive : nullptr);
        int rc = lb_launch_batch(h, &a, nullptr, deltaT_uniform, order, h->dQueue, h->dScratch, s);
        if (rc != LBCHEM_OK) return rc;
    }
    CK(cudaEventRecord(h->ev1, s));
    h->evValid = true;
    return LBCHEM_OK;
}

int lbchem_solve_batch_indexed(lbchem_handle h, int64_t n_index, const int32_t* index, int64_t ldy, const double* Y, const double* T,
                               const double* p, const double* rho, const double* deltaT, double deltaT_uniform, double* deltaTChem,
                               double* RR, double* Tout, double* cost, int32_t* status, void* stream) {
    if (!h) return LBCHEM_E_INVALID;
    if (n_index < 0 || (n_index > 0 && (!index || !Y || !T || !p || !rho || !deltaTChem || !RR)) || (!deltaT && !(deltaT_uniform > 0)))
        return fail(h, LBCHEM_E_INVALID, "bad batch arguments");
    CK(cudaSetDevice(h->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    h->launches = 0; h->hostPath = false; h->lastN = 0;
    CK(cudaMemsetAsync(h->dStats, 0, 8 * sizeof(unsigned long long), s));
    CK(cudaEventRecord(h->ev0, s));
    if (n_index > 0) {
        CellSeg a = make_seg(n_index, ldy, Y, T, p, rho, deltaT, deltaTChem, RR, Tout, cost, status, nullptr);
        a.index = index;
        int rc = lb_launch_batch(h, &a, nullptr, deltaT_uniform, nullptr, h->dQueue, h->dScratch, s);
        if (rc != LBCHEM_OK) return rc;
    }
    CK(cudaEventRecord(h->ev1, s));
    h->evValid = true;
    return LBCHEM_OK;
}

static int slot_reserve(lbchem_handle h, Slot& s, long long cap, int nSp) {
    if (!s.stream) {
        CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        CK(cudaEventCreate(&s.k0));
        CK(cudaEventCreate(&s.k1));
        CK(cudaMalloc(&s.queue, sizeof(unsigned long long)));
    }
    // every slot owns its work queue AND its kernel scratch: the kernels of consecutive chunks overlap on the device
    if (h->useV3 && s.scratchCap < h->scratchCap) {
        if (s.scratch) cudaFree(s.scratch);
        s.scratch = nullptr; s.scratchCap = 0;
        CK(cudaMalloc(&s.scratch, h->scratchCap));
        s.scratchCap = h->scratchCap;
    }
    if (cap <= s.cap) return LBCHEM_OK;
    void** ptrs[] = {(void**)&s.Y, (void**)&s.RR, (void**)&s.T, (void**)&s.p, (void**)&s.rho, (void**)&s.dT,
                     (void**)&s.dtc, (void**)&s.Tout, (void**)&s.cost, (void**)&s.status};
    for (void** p : ptrs) if (*p) { cudaFree(*p); *p = nullptr; }
    const size_t c = (size_t)cap;
    CK(cudaMalloc(&s.Y, c * nSp * 8)); CK(cudaMalloc(&s.RR, c * nSp * 8));
    CK(cudaMalloc(&s.T, c * 8)); CK(cudaMalloc(&s.p, c * 8)); CK(cudaMalloc(&s.rho, c * 8)); CK(cudaMalloc(&s.dT, c * 8));
    CK(cudaMalloc(&s.dtc, c * 8)); CK(cudaMalloc(&s.Tout, c * 8)); CK(cudaMalloc(&s.cost, c * 8)); CK(cudaMalloc(&s.status, c * 4));
    s.cap = cap;
    return LBCHEM_OK;
}

// Host-pointer entry point: cells are cut into chunks that cycle through kStreams streams, so the
// H2D copy of chunk k+1 and the D2H copy of chunk k-1 overlap the integration of chunk k (when the caller's
// buffers are pinned; copies from pageable memory block the host and the chunks run one after the other).
int lbchem_solve_batch_host(lbchem_handle h, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p,
                            const double* rho, const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR,
                            double* Tout, double* cost, int32_t* status) {
    if (!h) return LBCHEM_E_INVALID;
    if (n < 0 || ldy < n || (n > 0 && (!Y || !T || !p || !rho || !deltaTChem || !RR)) || (!deltaT && !(deltaT_uniform > 0)))
        return fail(h, LBCHEM_E_INVALID, "bad batch arguments");
    CK(cudaSetDevice(h->device));
    const int nSp = h->M.nSp;
    long long chunk = 32768;
    if (const char* e = getenv("LBCHEM_CHUNK")) if (atoll(e) > 0) chunk = atoll(e);
    if (chunk > n) chunk = n > 0 ? n : 1;
    h->launches = 0; h->hostPath = true; h->lastN = n; h->evValid = false;
    { int rc = reserve_nactive(h, n); if (rc != LBCHEM_OK) return rc; }
    for (Slot& s : h->slots) s.used = false;
    CK(cudaMemset(h->dStats, 0, 8 * sizeof(unsigned long long)));
    int k = 0;
    for (long long c0 = 0; c0 < n; c0 += chunk, ++k) {
        const long long m = std::min(chunk, (long long)n - c0);
        Slot& s = h->slots[k % kStreams];
        int rc = slot_reserve(h, s, chunk, nSp);
        if (rc != LBCHEM_OK) return rc;
        cudaStream_t st = s.stream;
        const size_t w = (size_t)m * 8;
        CK(cudaMemcpy2DAsync(s.Y, (size_t)s.cap * 8, Y + c0, (size_t)ldy * 8, w, nSp, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(s.T, T + c0, w, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(s.p, p + c0, w, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(s.rho, rho + c0, w, cudaMemcpyHostToDevice, st));
        if (deltaT) CK(cudaMemcpyAsync(s.dT, deltaT + c0, w, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(s.dtc, deltaTChem + c0, w, cudaMemcpyHostToDevice, st));
        if (!s.used) CK(cudaEventRecord(s.k0, st));
        const CellSeg a = make_seg(m, s.cap, s.Y, s.T, s.p, s.rho, deltaT ? s.dT : nullptr, s.dtc, s.RR, s.Tout, s.cost, s.status,
                                   (h->dacOn && h->dNActive && c0 + m <= h->nActiveCap) ? h->dNActive + c0 : nullptr);
        rc = lb_launch_batch(h, &a, nullptr, deltaT_uniform, nullptr, s.queue, s.scratch ? s.scratch : h->dScratch, st);
        if (rc != LBCHEM_OK) return rc;
        CK(cudaEventRecord(s.k1, st));
        s.used = true;
        CK(cudaMemcpy2DAsync(RR + c0, (size_t)ldy * 8, s.RR, (size_t)s.cap * 8, w, nSp, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(deltaTChem + c0, s.dtc, w, cudaMemcpyDeviceToHost, st));
        if (Tout) CK(cudaMemcpyAsync(Tout + c0, s.Tout, w, cudaMemcpyDeviceToHost, st));
        if (cost) CK(cudaMemcpyAsync(cost + c0, s.cost, w, cudaMemcpyDeviceToHost, st));
        if (status) CK(cudaMemcpyAsync(status + c0, s.status, (size_t)m * 4, cudaMemcpyDeviceToHost, st));
    }
    for (Slot& s : h->slots) if (s.used) CK(cudaStreamSynchronize(s.stream));
    return LBCHEM_OK;
}

int lbchem_get_stats(lbchem_handle h, lbchem_stats* out) {
    if (!h || !out) return LBCHEM_E_INVALID;
    CK(cudaSetDevice(h->device));
    memset(out, 0, sizeof(*out));
    float ms = 0.f;
    if (h->hostPath) {
        // first-start to last-stop per stream is not meaningful across streams; report the busiest stream's span
        for (Slot& s : h->slots) if (s.used) {
            CK(cudaStreamSynchronize(s.stream));
            float t = 0.f; CK(cudaEventElapsedTime(&t, s.k0, s.k1)); ms = std::max(ms, t);
        }
    } else if (h->evValid) {
        CK(cudaEventSynchronize(h->ev1));
        CK(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    }
    unsigned long long st[8];
    CK(cudaMemcpy(st, h->dStats, sizeof(st), cudaMemcpyDeviceToHost));
    out->n_steps = st[0]; out->n_attempts = st[1]; out->n_jac = st[2]; out->n_lu = st[3]; out->n_solve = st[4];
    out->n_rhs = st[5]; out->n_failed = st[6]; out->n_cells = st[7];
    out->kernel_ms = ms; out->kernel_launches = h->launches;
    return LBCHEM_OK;
}

// Test hook: evaluate derivatives / jacobian / LU-solve / rate constants for nItems states with the device
// functions of the integrator. HOST pointers. Strides are in doubles (see probe_kernel for layouts).
int lbchem_probe(lbchem_handle h, int32_t kind, int32_t nItems, const double* in, int32_t inStride, double* out, int32_t outStride) {
    if (!h || !in || !out || nItems < 1 || kind < 0 || kind > LBCHEM_PROBE_JAC_LUSOLVE) return LBCHEM_E_INVALID;
    CK(cudaSetDevice(h->device));
    const bool v3 = h->useV3 && kind != LBCHEM_PROBE_LUSOLVE;
    if (kind == LBCHEM_PROBE_JAC_LUSOLVE && !v3)
        return fail(h, LBCHEM_E_UNSUPPORTED, "LBCHEM_PROBE_JAC_LUSOLVE probes the compact-code kernel (solver seulex or Rosenbrock34)");
    // staging buffers are kept with the handle (test hook, but no allocation per call)
    const size_t needIn = (size_t)nItems * inStride * 8, needOut = (size_t)nItems * outStride * 8;
    if (needIn > h->probeInCap) { if (h->dProbeIn) cudaFree(h->dProbeIn); h->dProbeIn = nullptr; h->probeInCap = 0; CK(cudaMalloc(&h->dProbeIn, needIn)); h->probeInCap = needIn; }
    if (needOut > h->probeOutCap) { if (h->dProbeOut) cudaFree(h->dProbeOut); h->dProbeOut = nullptr; h->probeOutCap = 0; CK(cudaMalloc(&h->dProbeOut, needOut)); h->probeOutCap = needOut; }
    double *dIn = h->dProbeIn, *dOut = h->dProbeOut;
    CK(cudaMemcpy(dIn, in, needIn, cudaMemcpyHostToDevice));
    CK(cudaMemset(dOut, 0, needOut));
    if (v3) {
        BatchArgs B{};
        B.scratch = h->dScratch; B.scratchStride = h->scratchStride;
        long long grid = h->nSM;
        const long long need = (nItems + h->v3Groups - 1) / h->v3Groups;
        if (grid > need) grid = need;
        CK(lb_v3_probe(h->V2.tpc, (int)grid, h->v3Groups, h->v3Smem, h->v3Ws, h->V2, h->O, B, kind, nItems, dIn, inStride, dOut, outStride, 0));
    } else {
        CK(lb_launch_probe(h->tpc, h->groups, h->smemBytes, (int)h->wsDoubles, h->M, h->O, kind, nItems, dIn, dOut, inStride, outStride, 0));
    }
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(out, dOut, needOut, cudaMemcpyDeviceToHost));
    if (v3 && kind == LBCHEM_PROBE_RATES) {
        // packed order -> mechanism order
        const int nR = h->M.nR;
        std::vector<double> tmp(2 * (size_t)nR);
        for (int it = 0; it < nItems; ++it) {
            double* o = out + (size_t)it * outStride;
            std::copy(o, o + 2 * nR, tmp.begin());
            for (int r = 0; r < nR; ++r) { o[r] = tmp[h->v3PosOf[r]]; o[nR + r] = tmp[nR + h->v3PosOf[r]]; }
        }
    }
    return LBCHEM_OK;
}

}  // extern "C"
