/*
 * lbchem_oracle.c — CPU ORACLE for the per-cell chemistry integration path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build,
 * load or call this file. The product path (loadbalancedchemistrymodel_b200/csrc) never links it and
 * has no CPU fallback.
 *
 * PARITY UNPINNED: the arithmetic of this path lives in OpenFOAM v2306 (libODE, libchemistryModel,
 * libspecie), which is neither vendored in the reference nor installed here, and the reference ships no
 * golden vectors for it. This file is a literal single-threaded-per-cell
 * restatement of:
 *   reference  src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C:365-450
 *              (solveCell, solveCellList) and :143-167 (cellsToSend), :170-347 (balancing plan)
 *   OpenFOAM   chemistrySolver/ode/ode.C, ODE/ODESolvers/{ODESolver,adaptiveSolver,Rosenbrock34,seulex},
 *              chemistrySolver/EulerImplicit/EulerImplicit.C,
 *              chemistryModel/StandardChemistryModel/StandardChemistryModel.C (omega, derivatives, jacobian),
 *              specie/reaction/reactionRate/{Arrhenius,thirdBodyArrhenius,FallOff,Troe,Lindemann,thirdBodyEfficiencies},
 *              specie/thermo/{thermo/thermoI.H (K, Kc), janaf/janafThermoI.H}, matrices/scalarMatrices (LU)
 * as described in SURVEY.md Appendix A. Doubtful constants are parameters (gas constant, Kc floor,
 * dfdx quirk, seulex errOld form).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/lbchem.h"

#define O_SMALL 1e-15
#define O_VSMALL 1e-300
#define O_GREAT 1e15
#define O_VGREAT 1e300
#define O_PSTD 1e5

#define MAXSIDE 8

typedef struct {
    int64_t n_steps, n_attempts, n_jac, n_lu, n_solve, n_rhs;
} ocounters;

/* Oracle-side mechanism: desc + per-reaction combined janaf coefficients (Reaction::setThermo) */
typedef struct oracle_mech {
    lbchem_mech_desc d;   /* deep copies */
    int nSp, nR, nEqn;
    double* rhi;          /* [nR][7]  sum_rhs nu*a_hi - sum_lhs nu*a_hi */
    double* rlo;          /* [nR][7] */
    double* rTc;          /* [nR] Tcommon of the first lhs species (janafThermo::operator== keeps jt1.Tcommon_) */
    double* rdn;          /* [nR] sum nu_rhs - sum nu_lhs */
} oracle_mech;

typedef struct {
    lbchem_options o;
    int seulex_errold_max;  /* 0: errOld=min(4err,1) (OpenFOAM text as recalled); 1: NR3 max(4err,1) */
    /* reduction{} dictionary of LoadBalancedTDAC (reference tests/Cases/Case-chemistry/constant/chemistryProperties:35-66) */
    int dac_active; double dac_tol; int dac_ninit; int dac_init[64];
} oracle_opts;

static void* dup_mem(const void* p, size_t bytes) {
    void* q = malloc(bytes ? bytes : 1);
    if (p && bytes) memcpy(q, p, bytes);
    return q;
}

oracle_mech* oracle_mech_create(const lbchem_mech_desc* m) {
    oracle_mech* M = (oracle_mech*)calloc(1, sizeof(oracle_mech));
    int nSp = m->n_species, nR = m->n_reactions;
    M->nSp = nSp; M->nR = nR; M->nEqn = nSp + 2;
    M->d = *m;
    M->d.W = dup_mem(m->W, 8 * nSp);
    M->d.Tlow = dup_mem(m->Tlow, 8 * nSp);
    M->d.Thigh = dup_mem(m->Thigh, 8 * nSp);
    M->d.Tcommon = dup_mem(m->Tcommon, 8 * nSp);
    M->d.janaf_high = dup_mem(m->janaf_high, 56 * nSp);
    M->d.janaf_low = dup_mem(m->janaf_low, 56 * nSp);
    M->d.rtype = dup_mem(m->rtype, 4 * nR);
    M->d.reversible = dup_mem(m->reversible, 4 * nR);
    M->d.arr = dup_mem(m->arr, 24 * nR);
    M->d.arr0 = dup_mem(m->arr0, 24 * nR);
    M->d.troe = dup_mem(m->troe, 32 * nR);
    M->d.eff_index = dup_mem(m->eff_index, 4 * nR);
    M->d.eff = dup_mem(m->eff, 8 * (size_t)m->n_eff * nSp);
    M->d.lhs_ptr = dup_mem(m->lhs_ptr, 4 * (nR + 1));
    M->d.rhs_ptr = dup_mem(m->rhs_ptr, 4 * (nR + 1));
    int nl = m->lhs_ptr[nR], nr = m->rhs_ptr[nR];
    M->d.lhs_species = dup_mem(m->lhs_species, 4 * nl);
    M->d.lhs_nu = dup_mem(m->lhs_nu, 8 * nl);
    M->d.lhs_exp = dup_mem(m->lhs_exp, 8 * nl);
    M->d.rhs_species = dup_mem(m->rhs_species, 4 * nr);
    M->d.rhs_nu = dup_mem(m->rhs_nu, 8 * nr);
    M->d.rhs_exp = dup_mem(m->rhs_exp, 8 * nr);
    M->rhi = (double*)calloc(7 * nR, 8);
    M->rlo = (double*)calloc(7 * nR, 8);
    M->rTc = (double*)calloc(nR, 8);
    M->rdn = (double*)calloc(nR, 8);
    for (int r = 0; r < nR; ++r) {
        double dn = 0;
        for (int k = m->rhs_ptr[r]; k < m->rhs_ptr[r + 1]; ++k) {
            int s = m->rhs_species[k];
            dn += m->rhs_nu[k];
            for (int q = 0; q < 7; ++q) {
                M->rhi[7 * r + q] += m->rhs_nu[k] * m->janaf_high[7 * s + q];
                M->rlo[7 * r + q] += m->rhs_nu[k] * m->janaf_low[7 * s + q];
            }
        }
        for (int k = m->lhs_ptr[r]; k < m->lhs_ptr[r + 1]; ++k) {
            int s = m->lhs_species[k];
            dn -= m->lhs_nu[k];
            for (int q = 0; q < 7; ++q) {
                M->rhi[7 * r + q] -= m->lhs_nu[k] * m->janaf_high[7 * s + q];
                M->rlo[7 * r + q] -= m->lhs_nu[k] * m->janaf_low[7 * s + q];
            }
        }
        M->rdn[r] = dn;
        M->rTc[r] = m->Tcommon[m->lhs_species[m->lhs_ptr[r]]];
    }
    return M;
}

void oracle_mech_destroy(oracle_mech* M) {
    if (!M) return;
    free((void*)M->d.W); free((void*)M->d.Tlow); free((void*)M->d.Thigh); free((void*)M->d.Tcommon);
    free((void*)M->d.janaf_high); free((void*)M->d.janaf_low); free((void*)M->d.rtype);
    free((void*)M->d.reversible); free((void*)M->d.arr); free((void*)M->d.arr0); free((void*)M->d.troe);
    free((void*)M->d.eff_index); free((void*)M->d.eff); free((void*)M->d.lhs_ptr); free((void*)M->d.rhs_ptr);
    free((void*)M->d.lhs_species); free((void*)M->d.lhs_nu); free((void*)M->d.lhs_exp);
    free((void*)M->d.rhs_species); free((void*)M->d.rhs_nu); free((void*)M->d.rhs_exp);
    free(M->rhi); free(M->rlo); free(M->rTc); free(M->rdn);
    free(M);
}

/* ------------------------------------------------------------------------------------------------
 * per-thread workspace (the reference model's mutable scratch: c_, dcdt_, ODE solver members)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    const oracle_mech* M;
    const oracle_opts* O;
    int n;                 /* nEqn */
    double *c_, *dcdt_;    /* StandardChemistryModel::c_, dcdt_ */
    double *cTp;           /* ode::cTp_ */
    /* adaptiveSolver / Rosenbrock34 */
    double *dydx0, *yTemp, *k1, *k2, *k3, *k4, *err, *dydx, *dfdx;
    double *dfdy, *a;      /* n*n */
    int* piv;
    double* vv;
    /* seulex */
    double *y0, *scale, *ySeq, *dy, *yT, *table; /* table [kMaxx][n] */
    double theta; int kTarg;
    double dxOpt[12], temp[12], cpu[12], coeff[12][12]; int nSeq[12];
    double jacRedo;
    /* EulerImplicit */
    double *RRm, *src;
    ocounters cnt;
    int fail;              /* LBCHEM_CELL_* of the current cell */
    /* TDACChemistryModel members used while a DAC-reduced mechanism is being integrated */
    int red;               /* 1 while nSpecie_ == NsDAC_ (between reduceMechanism and the reset at TDAC solveCell :431-434) */
    int NsDAC;
    int *s2c, *c2s;        /* simplifiedToCompleteIndex_, completeToSimplifiedIndex_ */
    unsigned char* rdis;   /* reactionsDisabled_ */
    double *completeC, *simplC;
    /* DAC::reduceMechanism scratch */
    double *rABNum, *PA, *CA, *Rvalue; int *rABPos, *rABOther, *NbInit, *queue; unsigned char* act;
} ows;

static ows* ows_create(const oracle_mech* M, const oracle_opts* O) {
    ows* w = (ows*)calloc(1, sizeof(ows));
    int n = M->nEqn;
    w->M = M; w->O = O; w->n = n;
    double** v[] = {&w->c_, &w->dcdt_, &w->cTp, &w->dydx0, &w->yTemp, &w->k1, &w->k2, &w->k3, &w->k4,
                    &w->err, &w->dydx, &w->dfdx, &w->vv, &w->y0, &w->scale, &w->ySeq, &w->dy, &w->yT, &w->src};
    for (size_t i = 0; i < sizeof(v) / sizeof(v[0]); ++i) *v[i] = (double*)calloc(n, 8);
    w->dfdy = (double*)calloc((size_t)n * n, 8);
    w->a = (double*)calloc((size_t)n * n, 8);
    w->RRm = (double*)calloc((size_t)n * n, 8);
    w->table = (double*)calloc((size_t)12 * n, 8);
    w->piv = (int*)calloc(n, sizeof(int));
    {
        const int nSp = M->nSp, nR = M->nR;
        w->s2c = (int*)calloc(nSp, sizeof(int)); w->c2s = (int*)calloc(nSp, sizeof(int));
        w->rdis = (unsigned char*)calloc(nR ? nR : 1, 1);
        w->completeC = (double*)calloc(nSp, 8); w->simplC = (double*)calloc(n, 8);
        w->rABNum = (double*)calloc((size_t)nSp * nSp, 8);
        w->PA = (double*)calloc(nSp, 8); w->CA = (double*)calloc(nSp, 8); w->Rvalue = (double*)calloc(nSp, 8);
        w->rABPos = (int*)calloc((size_t)nSp * nSp, sizeof(int)); w->rABOther = (int*)calloc((size_t)nSp * nSp, sizeof(int));
        w->NbInit = (int*)calloc(nSp, sizeof(int)); w->queue = (int*)calloc(nSp + 1, sizeof(int));
        w->act = (unsigned char*)calloc(nSp, 1);
    }
    /* seulex constructor */
    const double cpuFunc = 1, cpuJac = 5, cpuLU = 1, cpuSolve = 1;
    w->nSeq[0] = 2; w->nSeq[1] = 3;
    for (int i = 2; i < 12; ++i) w->nSeq[i] = 2 * w->nSeq[i - 2];
    w->cpu[0] = cpuJac + cpuLU + w->nSeq[0] * (cpuFunc + cpuSolve);
    for (int k = 0; k < 11; ++k) w->cpu[k + 1] = w->cpu[k] + (w->nSeq[k + 1] - 1) * (cpuFunc + cpuSolve) + cpuLU;
    for (int k = 0; k < 12; ++k)
        for (int l = 0; l < k; ++l) {
            double ratio = (double)w->nSeq[k] / w->nSeq[l];
            w->coeff[k][l] = 1 / (ratio - 1);
        }
    w->jacRedo = fmin(1e-4, O->o.rel_tol);
    w->theta = 2 * w->jacRedo;
    w->kTarg = 1;
    return w;
}

static void ows_destroy(ows* w) {
    free(w->c_); free(w->dcdt_); free(w->cTp); free(w->dydx0); free(w->yTemp); free(w->k1); free(w->k2);
    free(w->k3); free(w->k4); free(w->err); free(w->dydx); free(w->dfdx); free(w->vv); free(w->y0);
    free(w->scale); free(w->ySeq); free(w->dy); free(w->yT); free(w->src); free(w->dfdy); free(w->a);
    free(w->RRm); free(w->table); free(w->piv);
    free(w->s2c); free(w->c2s); free(w->rdis); free(w->completeC); free(w->simplC); free(w->rABNum); free(w->PA);
    free(w->CA); free(w->Rvalue); free(w->rABPos); free(w->rABOther); free(w->NbInit); free(w->queue); free(w->act);
    free(w);
}

/* ------------------------------------------------------------------------------------------------
 * janaf thermo (janafThermoI.H): coefficients are the cp/R form; molar values = poly * RR
 * ---------------------------------------------------------------------------------------------- */
static inline double janaf_cp(const double* a, double T) {
    return (((a[4] * T + a[3]) * T + a[2]) * T + a[1]) * T + a[0];
}
static inline double janaf_ha(const double* a, double T) {
    return ((((a[4] / 5.0 * T + a[3] / 4.0) * T + a[2] / 3.0) * T + a[1] / 2.0) * T + a[0]) * T + a[5];
}
static inline double janaf_s(const double* a, double T) {
    return (((a[4] / 4.0 * T + a[3] / 3.0) * T + a[2] / 2.0) * T + a[1]) * T + a[0] * log(T) + a[6];
}
static inline const double* sp_coeffs(const oracle_mech* M, int i, double T) {
    return (T < M->d.Tcommon[i]) ? M->d.janaf_low + 7 * i : M->d.janaf_high + 7 * i;
}

/* species::thermo::Kc for the reaction's combined thermo (thermoI.H K/Kp/Kc) */
static double reaction_Kc(const oracle_mech* M, const oracle_opts* O, int r, double T) {
    const double RR = O->o.gas_constant;
    const double* a = (T < M->rTc[r]) ? M->rlo + 7 * r : M->rhi + 7 * r;
    /* Y*Gstd = (Ha - T*S) molar of the net reaction; a[] dimensionless => multiply by RR */
    double G = RR * (janaf_ha(a, T) - T * janaf_s(a, T));
    double arg = -G / (RR * T);
    double Kp = (arg < 600) ? exp(arg) : O_VGREAT;
    double nm = M->rdn[r];
    if (nm == 0.0) return Kp;
    return Kp * pow(O_PSTD / (RR * T), nm);
}

static inline double arrhenius(const double* k, double T) {
    double ak = k[0];
    if (fabs(k[1]) > O_VSMALL) ak *= pow(T, k[1]);
    if (fabs(k[2]) > O_VSMALL) ak *= exp(-k[2] / T);
    return ak;
}

static double third_body_M(const oracle_mech* M, int r, const double* c) {
    const double* e = M->d.eff + (size_t)M->d.eff_index[r] * M->nSp;
    double m = 0;
    for (int i = 0; i < M->nSp; ++i) m += e[i] * c[i];
    return m;
}

static double reaction_kf(const oracle_mech* M, int r, double T, const double* c) {
    const lbchem_mech_desc* d = &M->d;
    switch (d->rtype[r]) {
        case LBCHEM_R_ARRHENIUS:
            return arrhenius(d->arr + 3 * r, T);
        case LBCHEM_R_THIRD_BODY:
            return third_body_M(M, r, c) * arrhenius(d->arr + 3 * r, T);
        default: {
            double k0 = arrhenius(d->arr0 + 3 * r, T);
            double kInf = arrhenius(d->arr + 3 * r, T);
            double Pr = k0 * third_body_M(M, r, c) / kInf;
            double F = 1.0;
            if (d->rtype[r] == LBCHEM_R_FALLOFF_TROE) {
                const double* t = d->troe + 4 * r; /* alpha, Tsss, Ts, Tss */
                double logFcent = log10(fmax((1 - t[0]) * exp(-T / t[1]) + t[0] * exp(-T / t[2]) + exp(-t[3] / T), O_SMALL));
                double cc = -0.4 - 0.67 * logFcent;
                const double dd = 0.14;
                double nn = 0.75 - 1.27 * logFcent;
                double logPr = log10(fmax(Pr, O_SMALL));
                double q = (logPr + cc) / (nn - dd * (logPr + cc));
                F = pow(10.0, logFcent / (1.0 + q * q));
            }
            return kInf * (Pr / (1 + Pr)) * F;
        }
    }
}

static inline double reaction_kr(const oracle_mech* M, const oracle_opts* O, int r, double kf, double T) {
    if (!M->d.reversible[r]) return 0.0;
    return kf / fmax(reaction_Kc(M, O, r, T), O->o.kc_floor);
}

/* StandardChemistryModel::omega(R, c, T, p, pf, cf, lRef, pr, cr, rRef) */
static double omega_reaction(const oracle_mech* M, const oracle_opts* O, int r, const double* c, double T,
                             double* pf, double* cf, int* lRef, double* pr, double* cr, int* rRef) {
    const lbchem_mech_desc* d = &M->d;
    const double kf = reaction_kf(M, r, T, c);
    const double kr = reaction_kr(M, O, r, kf, T);
    const int l0 = d->lhs_ptr[r], Nl = d->lhs_ptr[r + 1] - l0;
    const int r0 = d->rhs_ptr[r], Nr = d->rhs_ptr[r + 1] - r0;

    int slRef = 0;
    *lRef = d->lhs_species[l0];
    *pf = kf;
    for (int s = 1; s < Nl; ++s) {
        int si = d->lhs_species[l0 + s];
        if (c[si] < c[*lRef]) {
            double ex = d->lhs_exp[l0 + slRef];
            *pf *= pow(fmax(c[*lRef], 0.0), ex);
            *lRef = si; slRef = s;
        } else {
            double ex = d->lhs_exp[l0 + s];
            *pf *= pow(fmax(c[si], 0.0), ex);
        }
    }
    *cf = fmax(c[*lRef], 0.0);
    {
        double ex = d->lhs_exp[l0 + slRef];
        if (ex < 1.0) {
            if (*cf > O_SMALL) *pf *= pow(*cf, ex - 1.0); else *pf = 0.0;
        } else {
            *pf *= pow(*cf, ex - 1.0);
        }
    }
    int srRef = 0;
    *rRef = d->rhs_species[r0];
    *pr = kr;
    for (int s = 1; s < Nr; ++s) {
        int si = d->rhs_species[r0 + s];
        if (c[si] < c[*rRef]) {
            double ex = d->rhs_exp[r0 + srRef];
            *pr *= pow(fmax(c[*rRef], 0.0), ex);
            *rRef = si; srRef = s;
        } else {
            double ex = d->rhs_exp[r0 + s];
            *pr *= pow(fmax(c[si], 0.0), ex);
        }
    }
    *cr = fmax(c[*rRef], 0.0);
    {
        double ex = d->rhs_exp[r0 + srRef];
        if (ex < 1.0) {
            if (*cr > O_SMALL) *pr *= pow(*cr, ex - 1.0); else *pr = 0.0;
        } else {
            *pr *= pow(*cr, ex - 1.0);
        }
    }
    return (*pf) * (*cf) - (*pr) * (*cr);
}

/* StandardChemistryModel::omega(c, T, p, dcdt): zeroes all nEqn entries, fills species entries */
static void omega_all(ows* w, const double* c, double T, double* dcdt) {
    const oracle_mech* M = w->M;
    const lbchem_mech_desc* d = &M->d;
    for (int i = 0; i < w->n; ++i) dcdt[i] = 0;
    for (int r = 0; r < M->nR; ++r) {
        double pf, cf, pr, cr; int lRef, rRef;
        double om = omega_reaction(M, w->O, r, c, T, &pf, &cf, &lRef, &pr, &cr, &rRef);
        for (int k = d->lhs_ptr[r]; k < d->lhs_ptr[r + 1]; ++k) dcdt[d->lhs_species[k]] -= d->lhs_nu[k] * om;
        for (int k = d->rhs_ptr[r]; k < d->rhs_ptr[r + 1]; ++k) dcdt[d->rhs_species[k]] += d->rhs_nu[k] * om;
    }
}

static void tdac_derivatives(ows* w, const double* y, double* dcdt);
static void tdac_jacobian(ows* w, const double* y, double* dfdx, double* dfdc);

/* StandardChemistryModel::derivatives(t, c, dcdt) */
static void derivatives(ows* w, const double* y, double* dcdt) {
    if (w->red) { tdac_derivatives(w, y, dcdt); return; }
    const oracle_mech* M = w->M;
    const int nSp = M->nSp;
    const double RR = w->O->o.gas_constant;
    const double T = y[nSp];
    w->cnt.n_rhs++;
    for (int i = 0; i < nSp; ++i) w->c_[i] = fmax(y[i], 0.0);
    omega_all(w, w->c_, T, dcdt);
    double rho = 0, cSum = 0;
    for (int i = 0; i < nSp; ++i) { cSum += w->c_[i]; rho += M->d.W[i] * w->c_[i]; }
    (void)cSum;
    double cp = 0;
    for (int i = 0; i < nSp; ++i) cp += w->c_[i] * (RR * janaf_cp(sp_coeffs(M, i, T), T));
    cp /= rho;
    double dT = 0;
    for (int i = 0; i < nSp; ++i) dT += (RR * janaf_ha(sp_coeffs(M, i, T), T)) * dcdt[i];
    dT /= rho * cp;
    dcdt[nSp] = -dT;
    dcdt[nSp + 1] = 0;
}

/* StandardChemistryModel::jacobian(t, c, dcdt, dfdc); dcdt is the ODE solver's dfdx slot (quirk A.4) */
static void jacobian(ows* w, const double* y, double* dfdx, double* dfdc) {
    if (w->red) { tdac_jacobian(w, y, dfdx, dfdc); return; }
    const oracle_mech* M = w->M;
    const lbchem_mech_desc* d = &M->d;
    const int nSp = M->nSp, n = w->n;
    const double T = y[nSp];
    double* c = w->c_;
    w->cnt.n_jac++;
    for (int i = 0; i < nSp; ++i) c[i] = fmax(y[i], 0.0);
    memset(dfdc, 0, sizeof(double) * n * n);
    omega_all(w, c, T, dfdx);
    if (!w->O->o.dfdx_is_dcdt) for (int i = 0; i < n; ++i) dfdx[i] = 0;

    for (int r = 0; r < M->nR; ++r) {
        const double kf0 = reaction_kf(M, r, T, c);
        const double kr0 = reaction_kr(M, w->O, r, kf0, T);
        const int l0 = d->lhs_ptr[r], l1 = d->lhs_ptr[r + 1];
        const int r0 = d->rhs_ptr[r], r1 = d->rhs_ptr[r + 1];
        for (int j = l0; j < l1; ++j) {
            const int sj = d->lhs_species[j];
            double kf = kf0;
            for (int i = l0; i < l1; ++i) {
                const int si = d->lhs_species[i];
                const double el = d->lhs_exp[i];
                if (i == j) {
                    if (el < 1.0) {
                        if (c[si] > O_SMALL) kf *= el * pow(c[si] + O_VSMALL, el - 1.0); else kf = 0.0;
                    } else {
                        kf *= el * pow(c[si], el - 1.0);
                    }
                } else {
                    kf *= pow(c[si], el);
                }
            }
            for (int i = l0; i < l1; ++i) dfdc[d->lhs_species[i] * n + sj] -= d->lhs_nu[i] * kf;
            for (int i = r0; i < r1; ++i) dfdc[d->rhs_species[i] * n + sj] += d->rhs_nu[i] * kf;
        }
        for (int j = r0; j < r1; ++j) {
            const int sj = d->rhs_species[j];
            double kr = kr0;
            for (int i = r0; i < r1; ++i) {
                const int si = d->rhs_species[i];
                const double er = d->rhs_exp[i];
                if (i == j) {
                    if (er < 1.0) {
                        if (c[si] > O_SMALL) kr *= er * pow(c[si] + O_VSMALL, er - 1.0); else kr = 0.0;
                    } else {
                        kr *= er * pow(c[si], er - 1.0);
                    }
                } else {
                    kr *= pow(c[si], er);
                }
            }
            for (int i = l0; i < l1; ++i) dfdc[d->lhs_species[i] * n + sj] += d->lhs_nu[i] * kr;
            for (int i = r0; i < r1; ++i) dfdc[d->rhs_species[i] * n + sj] -= d->rhs_nu[i] * kr;
        }
    }
    const double delta = 1.0e-3;
    omega_all(w, c, T + delta, w->dcdt_);
    for (int i = 0; i < nSp; ++i) dfdc[i * n + nSp] = w->dcdt_[i];
    omega_all(w, c, T - delta, w->dcdt_);
    for (int i = 0; i < nSp; ++i) dfdc[i * n + nSp] = 0.5 * (dfdc[i * n + nSp] - w->dcdt_[i]) / delta;
    dfdc[nSp * n + nSp] = 0;
    dfdc[(nSp + 1) * n + nSp] = 0;
}


/* ------------------------------------------------------------------------------------------------
 * TDACChemistryModel::omega / derivatives / jacobian while a DAC-reduced mechanism is active
 * (OpenFOAM v2306 chemistryModel/TDACChemistryModel/TDACChemistryModel.C, SURVEY A.12 [from memory]):
 * the ODE vector is simplifiedC_ (NsDAC_ species, T, p); disabled reactions are skipped; rates, third-body
 * sums, rho and cp use the COMPLETE concentration vector c_ (= completeC_ with the active entries replaced).
 * ---------------------------------------------------------------------------------------------- */
static void tdac_load_c(ows* w, const double* y) {
    const int nSp = w->M->nSp;
    for (int i = 0; i < nSp; ++i) w->c_[i] = w->completeC[i];
    for (int i = 0; i < w->NsDAC; ++i) w->c_[w->s2c[i]] = fmax(y[i], 0.0);
}

static void tdac_omega_all(ows* w, const double* c, double T, double* dcdt) {
    const oracle_mech* M = w->M;
    const lbchem_mech_desc* d = &M->d;
    for (int i = 0; i < w->n; ++i) dcdt[i] = 0;
    for (int r = 0; r < M->nR; ++r) {
        if (w->rdis[r]) continue;
        double pf, cf, pr, cr; int lRef, rRef;
        double om = omega_reaction(M, w->O, r, c, T, &pf, &cf, &lRef, &pr, &cr, &rRef);
        for (int k = d->lhs_ptr[r]; k < d->lhs_ptr[r + 1]; ++k) dcdt[w->c2s[d->lhs_species[k]]] -= d->lhs_nu[k] * om;
        for (int k = d->rhs_ptr[r]; k < d->rhs_ptr[r + 1]; ++k) dcdt[w->c2s[d->rhs_species[k]]] += d->rhs_nu[k] * om;
    }
}

static void tdac_derivatives(ows* w, const double* y, double* dcdt) {
    const oracle_mech* M = w->M;
    const int nSp = M->nSp, Ns = w->NsDAC;
    const double RR = w->O->o.gas_constant;
    const double T = y[Ns];
    w->cnt.n_rhs++;
    tdac_load_c(w, y);
    tdac_omega_all(w, w->c_, T, dcdt);
    double rho = 0;
    for (int i = 0; i < nSp; ++i) rho += M->d.W[i] * w->c_[i];
    double cp = 0;
    for (int i = 0; i < nSp; ++i) cp += w->c_[i] * (RR * janaf_cp(sp_coeffs(M, i, T), T));
    cp /= rho;
    double dT = 0;
    for (int i = 0; i < Ns; ++i) { const int si = w->s2c[i]; dT += (RR * janaf_ha(sp_coeffs(M, si, T), T)) * dcdt[i]; }
    dT /= rho * cp;
    dcdt[Ns] = -dT;
    dcdt[Ns + 1] = 0;
}

static void tdac_jacobian(ows* w, const double* y, double* dfdx, double* dfdc) {
    const oracle_mech* M = w->M;
    const lbchem_mech_desc* d = &M->d;
    const int Ns = w->NsDAC, n = w->n;
    const int* c2s = w->c2s;
    const double T = y[Ns];
    double* c = w->c_;
    w->cnt.n_jac++;
    tdac_load_c(w, y);
    memset(dfdc, 0, sizeof(double) * n * n);
    /* the dfdx slot: same option as the standard model (irrelevant for seulex, which never reads it) */
    tdac_omega_all(w, c, T, dfdx);
    if (!w->O->o.dfdx_is_dcdt) for (int i = 0; i < n; ++i) dfdx[i] = 0;
    for (int r = 0; r < M->nR; ++r) {
        if (w->rdis[r]) continue;
        const double kf0 = reaction_kf(M, r, T, c);
        const double kr0 = reaction_kr(M, w->O, r, kf0, T);
        const int l0 = d->lhs_ptr[r], l1 = d->lhs_ptr[r + 1];
        const int r0 = d->rhs_ptr[r], r1 = d->rhs_ptr[r + 1];
        for (int j = l0; j < l1; ++j) {
            const int sj = c2s[d->lhs_species[j]];
            double kf = kf0;
            for (int i = l0; i < l1; ++i) {
                const int si = d->lhs_species[i];
                const double el = d->lhs_exp[i];
                if (i == j) {
                    if (el < 1.0) { if (c[si] > O_SMALL) kf *= el * pow(c[si] + O_VSMALL, el - 1.0); else kf = 0.0; }
                    else kf *= el * pow(c[si], el - 1.0);
                } else kf *= pow(c[si], el);
            }
            for (int i = l0; i < l1; ++i) dfdc[c2s[d->lhs_species[i]] * n + sj] -= d->lhs_nu[i] * kf;
            for (int i = r0; i < r1; ++i) dfdc[c2s[d->rhs_species[i]] * n + sj] += d->rhs_nu[i] * kf;
        }
        for (int j = r0; j < r1; ++j) {
            const int sj = c2s[d->rhs_species[j]];
            double kr = kr0;
            for (int i = r0; i < r1; ++i) {
                const int si = d->rhs_species[i];
                const double er = d->rhs_exp[i];
                if (i == j) {
                    if (er < 1.0) { if (c[si] > O_SMALL) kr *= er * pow(c[si] + O_VSMALL, er - 1.0); else kr = 0.0; }
                    else kr *= er * pow(c[si], er - 1.0);
                } else kr *= pow(c[si], er);
            }
            for (int i = l0; i < l1; ++i) dfdc[c2s[d->lhs_species[i]] * n + sj] += d->lhs_nu[i] * kr;
            for (int i = r0; i < r1; ++i) dfdc[c2s[d->rhs_species[i]] * n + sj] -= d->rhs_nu[i] * kr;
        }
    }
    const double delta = 1.0e-3;
    tdac_omega_all(w, c, T + delta, w->dcdt_);
    for (int i = 0; i < Ns; ++i) dfdc[i * n + Ns] = w->dcdt_[i];
    tdac_omega_all(w, c, T - delta, w->dcdt_);
    for (int i = 0; i < Ns; ++i) dfdc[i * n + Ns] = 0.5 * (dfdc[i * n + Ns] - w->dcdt_[i]) / delta;
    dfdc[Ns * n + Ns] = 0;
    dfdc[(Ns + 1) * n + Ns] = 0;
}

/* ------------------------------------------------------------------------------------------------
 * LUDecompose / LUBacksubstitute (scalarMatrices.C; Numerical-Recipes ludcmp / lubksb)
 * ---------------------------------------------------------------------------------------------- */
static int lu_decompose(ows* w, double* A, int m, int* piv) {
    double* vv = w->vv;
    w->cnt.n_lu++;
    for (int i = 0; i < m; ++i) {
        double largest = 0.0, temp;
        const double* Ai = A + (size_t)i * m;
        for (int j = 0; j < m; ++j)
            if ((temp = fabs(Ai[j])) > largest) largest = temp;
        if (largest == 0.0) return LBCHEM_CELL_SINGULAR;
        vv[i] = 1.0 / largest;
    }
    for (int j = 0; j < m; ++j) {
        double* Aj = A + (size_t)j * m;
        for (int i = 0; i < j; ++i) {
            double* Ai = A + (size_t)i * m;
            double sum = Ai[j];
            for (int k = 0; k < i; ++k) sum -= Ai[k] * A[(size_t)k * m + j];
            Ai[j] = sum;
        }
        int iMax = 0;
        double largest = 0.0;
        for (int i = j; i < m; ++i) {
            double* Ai = A + (size_t)i * m;
            double sum = Ai[j];
            for (int k = 0; k < j; ++k) sum -= Ai[k] * A[(size_t)k * m + j];
            Ai[j] = sum;
            double temp;
            if ((temp = vv[i] * fabs(sum)) >= largest) { largest = temp; iMax = i; }
        }
        piv[j] = iMax;
        if (j != iMax) {
            double* Am = A + (size_t)iMax * m;
            for (int k = 0; k < m; ++k) { double t = Aj[k]; Aj[k] = Am[k]; Am[k] = t; }
            vv[iMax] = vv[j];
        }
        if (Aj[j] == 0.0) Aj[j] = O_SMALL;
        if (j != m - 1) {
            double rDiag = 1.0 / Aj[j];
            for (int i = j + 1; i < m; ++i) A[(size_t)i * m + j] *= rDiag;
        }
    }
    return 0;
}

static void lu_backsubst(ows* w, const double* A, int m, const int* piv, double* b) {
    int ii = 0;
    w->cnt.n_solve++;
    for (int i = 0; i < m; ++i) {
        int ip = piv[i];
        double sum = b[ip];
        b[ip] = b[i];
        const double* Ai = A + (size_t)i * m;
        if (ii != 0) {
            for (int j = ii - 1; j < i; ++j) sum -= Ai[j] * b[j];
        } else if (sum != 0.0) {
            ii = i + 1;
        }
        b[i] = sum;
    }
    for (int i = m - 1; i >= 0; --i) {
        double sum = b[i];
        const double* Ai = A + (size_t)i * m;
        for (int j = i + 1; j < m; ++j) sum -= Ai[j] * b[j];
        b[i] = sum / Ai[i];
    }
}

/* ------------------------------------------------------------------------------------------------
 * ODESolver::normalizeError, Rosenbrock34::solve, adaptiveSolver::solve
 * ---------------------------------------------------------------------------------------------- */
static double normalize_error(const ows* w, const double* y0, const double* y, const double* err) {
    double maxErr = 0.0;
    for (int i = 0; i < w->n; ++i) {
        double tol = w->O->o.abs_tol + w->O->o.rel_tol * fmax(fabs(y0[i]), fabs(y[i]));
        maxErr = fmax(maxErr, fabs(err[i]) / tol);
    }
    return maxErr;
}

static double rosenbrock34_step(ows* w, const double* y0, const double* dydx0, double dx, double* y) {
    /* constants by Shampine (Rosenbrock34.C) */
    static const double a21 = 2, a31 = 48.0 / 25.0, a32 = 6.0 / 25.0, c21 = -8, c31 = 372.0 / 25.0,
                        c32 = 12.0 / 5.0, c41 = -112.0 / 125.0, c42 = -54.0 / 125.0, c43 = -2.0 / 5.0,
                        b1 = 19.0 / 9.0, b2 = 1.0 / 2.0, b3 = 25.0 / 108.0, b4 = 125.0 / 108.0,
                        e1 = 34.0 / 108.0, e2 = 7.0 / 36.0, e3 = 0, e4 = 125.0 / 108.0, gamma = 1.0 / 2.0,
                        c1X = 1.0 / 2.0, c2X = -3.0 / 2.0, c3X = 121.0 / 50.0, c4X = 29.0 / 250.0;
    const int n = w->n;
    double *k1 = w->k1, *k2 = w->k2, *k3 = w->k3, *k4 = w->k4, *dfdx = w->dfdx, *dydx = w->dydx, *a = w->a;
    jacobian(w, y0, dfdx, w->dfdy);
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) a[i * n + j] = -w->dfdy[i * n + j];
        a[i * n + i] += 1.0 / (gamma * dx);
    }
    int st = lu_decompose(w, a, n, w->piv);
    if (st) { w->fail = st; return 0.0; }
    for (int i = 0; i < n; ++i) k1[i] = dydx0[i] + dx * c1X * dfdx[i];
    lu_backsubst(w, a, n, w->piv, k1);
    for (int i = 0; i < n; ++i) y[i] = y0[i] + a21 * k1[i];
    derivatives(w, y, dydx);
    for (int i = 0; i < n; ++i) k2[i] = dydx[i] + dx * c2X * dfdx[i] + c21 * k1[i] / dx;
    lu_backsubst(w, a, n, w->piv, k2);
    for (int i = 0; i < n; ++i) y[i] = y0[i] + a31 * k1[i] + a32 * k2[i];
    derivatives(w, y, dydx);
    for (int i = 0; i < n; ++i) k3[i] = dydx[i] + dx * c3X * dfdx[i] + (c31 * k1[i] + c32 * k2[i]) / dx;
    lu_backsubst(w, a, n, w->piv, k3);
    for (int i = 0; i < n; ++i) k4[i] = dydx[i] + dx * c4X * dfdx[i] + (c41 * k1[i] + c42 * k2[i] + c43 * k3[i]) / dx;
    lu_backsubst(w, a, n, w->piv, k4);
    for (int i = 0; i < n; ++i) {
        y[i] = y0[i] + b1 * k1[i] + b2 * k2[i] + b3 * k3[i] + b4 * k4[i];
        w->err[i] = e1 * k1[i] + e2 * k2[i] + e3 * k3[i] + e4 * k4[i];
    }
    return normalize_error(w, y0, y, w->err);
}

/* adaptiveSolver::solve(odes, x, y, dxTry) */
static void adaptive_solve(ows* w, double* x, double* y, double* dxTry) {
    const double safeScale = 0.9, alphaInc = 0.2, alphaDec = 0.25, minScale = 0.2, maxScale = 10;
    double dx = *dxTry, err = 0;
    derivatives(w, y, w->dydx0);
    do {
        w->cnt.n_attempts++;
        err = rosenbrock34_step(w, y, w->dydx0, dx, w->yTemp);
        if (w->fail) return;
        if (!(err == err)) { w->fail = LBCHEM_CELL_NAN; return; }
        if (err > 1) {
            double scale = fmax(safeScale * pow(err, -alphaDec), minScale);
            dx *= scale;
            if (dx < O_VSMALL) { w->fail = LBCHEM_CELL_UNDERFLOW; return; }
        }
    } while (err > 1);
    *x += dx;
    memcpy(y, w->yTemp, sizeof(double) * w->n);
    if (err > pow(maxScale / safeScale, -1.0 / alphaInc))
        *dxTry = fmin(fmax(safeScale * pow(err, -alphaInc), minScale), maxScale) * dx;
    else
        *dxTry = safeScale * maxScale * dx;
}

/* ------------------------------------------------------------------------------------------------
 * seulex (seulex.C; NR3 StepperSie with OpenFOAM's stepState)
 * ---------------------------------------------------------------------------------------------- */
typedef struct { int forward, first, last, reject, prevReject; double dxTry, dxDid; } step_state;

static int seul(ows* w, const double* y0, double dxTot, int k, double* y, const double* scale) {
    const int n = w->n;
    int nSteps = w->nSeq[k];
    double dx = dxTot / nSteps;
    double *a = w->a, *dy = w->dy, *yT = w->yT, *dydx = w->dydx;
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) a[i * n + j] = -w->dfdy[i * n + j];
        a[i * n + i] += 1 / dx;
    }
    int st = lu_decompose(w, a, n, w->piv);
    if (st) { w->fail = st; return 0; }
    derivatives(w, y0, dy);
    lu_backsubst(w, a, n, w->piv, dy);
    memcpy(yT, y0, sizeof(double) * n);
    for (int nn = 1; nn < nSteps; ++nn) {
        for (int i = 0; i < n; ++i) yT[i] += dy[i];
        if (nn == 1 && k <= 1) {
            double dy1 = 0;
            for (int i = 0; i < n; ++i) { double q = dy[i] / scale[i]; dy1 += q * q; }
            dy1 = sqrt(dy1);
            derivatives(w, yT, dydx);
            for (int i = 0; i < n; ++i) dy[i] = dydx[i] - dy[i] / dx;
            lu_backsubst(w, a, n, w->piv, dy);
            const double denom = fmin(1.0, dy1 + O_SMALL);
            double dy2 = 0;
            for (int i = 0; i < n; ++i) {
                if (fabs(dy[i]) > scale[i] * denom) { w->theta = 1; return 0; }
                double q = dy[i] / scale[i];
                dy2 += q * q;
            }
            dy2 = sqrt(dy2);
            w->theta = dy2 / denom;
            if (w->theta > 1) return 0;
        }
        derivatives(w, yT, dy);
        lu_backsubst(w, a, n, w->piv, dy);
    }
    for (int i = 0; i < n; ++i) y[i] = yT[i] + dy[i];
    return 1;
}

static void seulex_extrapolate(ows* w, int k, double* y) {
    const int n = w->n;
    double* tb = w->table;
    for (int j = k - 1; j > 0; --j)
        for (int i = 0; i < n; ++i)
            tb[(j - 1) * n + i] = tb[j * n + i] + w->coeff[k][j] * (tb[j * n + i] - tb[(j - 1) * n + i]);
    for (int i = 0; i < n; ++i) y[i] = tb[i] + w->coeff[k][0] * (tb[i] - y[i]);
}

static void seulex_solve(ows* w, double* x, double* y, step_state* step) {
    const int n = w->n;
    const int kMaxx = 11;
    const double stepFactor1 = 0.6, stepFactor2 = 0.93, stepFactor3 = 0.1, stepFactor4 = 4, stepFactor5 = 0.5,
                 kFactor1 = 0.7, kFactor2 = 0.9;
    const double absTol = w->O->o.abs_tol, relTol = w->O->o.rel_tol;
    double *temp = w->temp, *dxOpt = w->dxOpt, *cpu = w->cpu, *scale = w->scale, *y0 = w->y0, *ySeq = w->ySeq;
    const int* nSeq = w->nSeq;
    temp[0] = O_GREAT;
    double dx = step->dxTry;
    memcpy(y0, y, sizeof(double) * n);
    dxOpt[0] = fabs(0.1 * dx);
    if (step->first || step->prevReject) w->theta = 2.0 * w->jacRedo;
    if (step->first) {
        double logTol = -log10(relTol + absTol) * 0.6 + 0.5;
        int kt = (int)logTol;
        if (kt > kMaxx - 1) kt = kMaxx - 1;
        if (kt < 1) kt = 1;
        w->kTarg = kt;
    }
    for (int i = 0; i < n; ++i) scale[i] = absTol + relTol * fabs(y[i]);
    int jacUpdated = 0;
    if (w->theta > w->jacRedo) { jacobian(w, y, w->dfdx, w->dfdy); jacUpdated = 1; }
    int k = 0;
    double dxNew = fabs(dx);
    int firstk = 1;
    while (firstk || step->reject) {
        dx = step->forward ? dxNew : -dxNew;
        firstk = 0;
        step->reject = 0;
        w->cnt.n_attempts++;
        if (fabs(dx) < O_VSMALL || w->cnt.n_attempts > 100000) { w->fail = LBCHEM_CELL_UNDERFLOW; return; }
        double errOld = 0;
        for (k = 0; k <= w->kTarg + 1; ++k) {
            int success = seul(w, y0, dx, k, ySeq, scale);
            if (w->fail) return;
            if (!success) { step->reject = 1; dxNew = fabs(dx) * stepFactor5; break; }
            if (k == 0) memcpy(y, ySeq, sizeof(double) * n);
            else memcpy(w->table + (size_t)(k - 1) * n, ySeq, sizeof(double) * n);
            if (k != 0) {
                seulex_extrapolate(w, k, y);
                double err = 0;
                for (int i = 0; i < n; ++i) {
                    scale[i] = absTol + relTol * fabs(y0[i]);
                    double q = (y[i] - w->table[i]) / scale[i];
                    err += q * q;
                }
                err = sqrt(err / n);
                if (!(err == err)) { w->fail = LBCHEM_CELL_NAN; return; }
                if (err > 1 / O_SMALL || (k > 1 && err >= errOld)) {
                    step->reject = 1; dxNew = fabs(dx) * stepFactor5; break;
                }
                errOld = w->O->seulex_errold_max ? fmax(4 * err, 1.0) : fmin(4 * err, 1.0);
                double expo = 1.0 / (k + 1);
                double facmin = pow(stepFactor3, expo);
                double fac;
                if (err == 0) fac = 1 / facmin;
                else {
                    fac = stepFactor2 / pow(err / stepFactor1, expo);
                    fac = fmax(facmin / stepFactor4, fmin(1 / facmin, fac));
                }
                dxOpt[k] = fabs(dx * fac);
                temp[k] = cpu[k] / dxOpt[k];
                if ((step->first || step->last) && err <= 1) break;
                if (k == w->kTarg - 1 && !step->prevReject && !step->first && !step->last) {
                    if (err <= 1) break;
                    else if (err > nSeq[w->kTarg] * nSeq[w->kTarg + 1] * 4) {
                        step->reject = 1;
                        w->kTarg = k;
                        if (w->kTarg > 1 && temp[k - 1] < kFactor1 * temp[k]) w->kTarg--;
                        dxNew = dxOpt[w->kTarg];
                        break;
                    }
                }
                if (k == w->kTarg) {
                    if (err <= 1) break;
                    else if (err > nSeq[k + 1] * 2) {
                        step->reject = 1;
                        if (w->kTarg > 1 && temp[k - 1] < kFactor1 * temp[k]) w->kTarg--;
                        dxNew = dxOpt[w->kTarg];
                        break;
                    }
                }
                if (k == w->kTarg + 1) {
                    if (err > 1) {
                        step->reject = 1;
                        if (w->kTarg > 1 && temp[w->kTarg - 1] < kFactor1 * temp[w->kTarg]) w->kTarg--;
                        dxNew = dxOpt[w->kTarg];
                    }
                    break;
                }
            }
        }
        if (step->reject) {
            step->prevReject = 1;
            if (!jacUpdated) {
                w->theta = 2.0 * w->jacRedo;
                if (w->theta > w->jacRedo && !jacUpdated) { jacobian(w, y, w->dfdx, w->dfdy); jacUpdated = 1; }
            }
        }
    }
    jacUpdated = 0;
    step->dxDid = dx;
    *x += dx;
    int kopt;
    if (k == 1) kopt = 2;
    else if (k <= w->kTarg) {
        kopt = k;
        if (temp[k - 1] < kFactor1 * temp[k]) kopt = k - 1;
        else if (temp[k] < kFactor2 * temp[k - 1]) kopt = (k + 1 < kMaxx - 1) ? k + 1 : kMaxx - 1;
    } else {
        kopt = k - 1;
        if (k > 2 && temp[k - 2] < kFactor1 * temp[k - 1]) kopt = k - 2;
        if (temp[k] < kFactor2 * temp[kopt]) kopt = (k < kMaxx - 1) ? k : kMaxx - 1;
    }
    if (step->prevReject) {
        w->kTarg = (kopt < k) ? kopt : k;
        dxNew = fmin(fabs(dx), dxOpt[w->kTarg]);
        step->prevReject = 0;
    } else {
        if (kopt <= k) dxNew = dxOpt[kopt];
        else {
            if (k < w->kTarg && temp[k] < kFactor2 * temp[k - 1]) dxNew = dxOpt[k] * cpu[kopt + 1] / cpu[k];
            else dxNew = dxOpt[k] * cpu[kopt] / cpu[k];
        }
        w->kTarg = kopt;
    }
    step->dxTry = step->forward ? dxNew : -dxNew;
}

/* ODESolver::solve(xStart, xEnd, y, dxTry) */
static void ode_solve(ows* w, double xStart, double xEnd, double* y, double* dxTry) {
    step_state step = {1, 1, 0, 0, 0, *dxTry, 0};
    double x = xStart;
    const int solver = w->O->o.solver;
    for (int nStep = 0; nStep < w->O->o.max_steps; ++nStep) {
        double dxTry0 = step.dxTry;
        step.reject = 0;
        if ((x + step.dxTry - xEnd) * (x + step.dxTry - xStart) > 0) {
            step.last = 1;
            step.dxTry = xEnd - x;
        }
        if (solver == LBCHEM_SOLVER_SEULEX) {
            seulex_solve(w, &x, y, &step);
        } else {
            double x0 = x;
            adaptive_solve(w, &x, y, &step.dxTry);
            step.dxDid = x - x0;
        }
        if (w->fail) return;
        w->cnt.n_steps++;
        if ((x - xEnd) * (xEnd - xStart) >= 0) {
            if (nStep > 0 && step.last) step.dxTry = dxTry0;
            *dxTry = step.dxTry;
            return;
        }
        step.first = 0;
        if (step.reject) step.prevReject = 1;
    }
    w->fail = LBCHEM_CELL_MAX_STEPS;
}

/* ------------------------------------------------------------------------------------------------
 * chemistrySolver::solve(c, T, p, deltaT, subDeltaT): ode<> and EulerImplicit<>
 * ---------------------------------------------------------------------------------------------- */
static void ode_chem_solve(ows* w, double* c, double* T, double* p, double* deltaT, double* subDeltaT) {
    const int nSp = w->n - 2;   /* this->nSpecie(): NsDAC_ while a reduced mechanism is active (ode::solve resizes cTp_) */
    for (int i = 0; i < nSp; ++i) w->cTp[i] = c[i];
    w->cTp[nSp] = *T;
    w->cTp[nSp + 1] = *p;
    ode_solve(w, 0, *deltaT, w->cTp, subDeltaT);
    for (int i = 0; i < nSp; ++i) c[i] = fmax(0.0, w->cTp[i]);
    *T = w->cTp[nSp];
    *p = w->cTp[nSp + 1];
}

/* mixture janaf: mass-weighted sum over species, all switching at the first species' Tcommon
 * (janafThermo::operator+= keeps the left operand's Tcommon_). Returned as sum_i c_i * molar value. */
static double mix_Ha(const ows* w, const double* c, double T) {
    const oracle_mech* M = w->M;
    const int lowSet = T < M->d.Tcommon[0];
    double h = 0;
    for (int i = 0; i < M->nSp; ++i)
        h += c[i] * janaf_ha((lowSet ? M->d.janaf_low : M->d.janaf_high) + 7 * i, T);
    return h * w->O->o.gas_constant;
}
static double mix_Cp(const ows* w, const double* c, double T) {
    const oracle_mech* M = w->M;
    const int lowSet = T < M->d.Tcommon[0];
    double s = 0;
    for (int i = 0; i < M->nSp; ++i)
        s += c[i] * janaf_cp((lowSet ? M->d.janaf_low : M->d.janaf_high) + 7 * i, T);
    return s * w->O->o.gas_constant;
}

/* species::thermo::THa -> T(f, p, T0, Ha, Cp, limit): Newton, tol 1e-4*T0, maxIter 100 */
static double mix_THa(ows* w, const double* c, double ha, double T0) {
    const oracle_mech* M = w->M;
    double Tlo = M->d.Tlow[0], Thi = M->d.Thigh[0];
    for (int i = 1; i < M->nSp; ++i) { Tlo = fmax(Tlo, M->d.Tlow[i]); Thi = fmin(Thi, M->d.Thigh[i]); }
    double Test = T0, Tnew = T0, Ttol = T0 * 1e-4;
    int iter = 0;
    do {
        Test = Tnew;
        Tnew = Test - (mix_Ha(w, c, Test) - ha) / mix_Cp(w, c, Test);
        if (Tnew < Tlo || Tnew > Thi) Tnew = fmin(fmax(Tnew, Tlo), Thi);
        if (iter++ > 100) { w->fail = LBCHEM_CELL_MAX_STEPS; break; }
    } while (fabs(Tnew - Test) > Ttol);
    return Tnew;
}

static void euler_implicit_solve(ows* w, double* c, double* T, double* p, double* deltaT, double* subDeltaT) {
    const oracle_mech* M = w->M;
    const lbchem_mech_desc* d = &M->d;
    const int nSp = M->nSp;
    double* RR = w->RRm;  /* nSp x nSp */
    (void)p;
    w->cnt.n_attempts++; w->cnt.n_steps++;
    for (int i = 0; i < nSp; ++i) c[i] = fmax(0.0, c[i]);
    double cTot = 0;
    for (int i = 0; i < nSp; ++i) cTot += c[i];
    const double ha = mix_Ha(w, c, *T);
    const double deltaTEst = fmin(*deltaT, *subDeltaT);
    memset(RR, 0, sizeof(double) * nSp * nSp);
    for (int r = 0; r < M->nR; ++r) {
        double pf, cf, pr, cr; int lRef, rRef;
        double om = omega_reaction(M, w->O, r, c, *T, &pf, &cf, &lRef, &pr, &cr, &rRef);
        double corr = 1;
        if (w->O->o.eq_rate_limiter) corr = (om < 0) ? 1 / (1 + pr * deltaTEst) : 1 / (1 + pf * deltaTEst);
        for (int k = d->lhs_ptr[r]; k < d->lhs_ptr[r + 1]; ++k) {
            int si = d->lhs_species[k]; double sl = d->lhs_nu[k];
            RR[si * nSp + rRef] -= sl * pr * corr;
            RR[si * nSp + lRef] += sl * pf * corr;
        }
        for (int k = d->rhs_ptr[r]; k < d->rhs_ptr[r + 1]; ++k) {
            int si = d->rhs_species[k]; double sr = d->rhs_nu[k];
            RR[si * nSp + lRef] -= sr * pf * corr;
            RR[si * nSp + rRef] += sr * pr * corr;
        }
    }
    w->cnt.n_rhs++;
    double tMin = O_GREAT;
    for (int i = 0; i < nSp; ++i) {
        double dd = 0;
        for (int j = 0; j < nSp; ++j) dd -= RR[i * nSp + j] * c[j];
        if (dd < -O_SMALL) tMin = fmin(tMin, -(c[i] + O_SMALL) / dd);
        else {
            dd = fmax(dd, O_SMALL);
            const double cm = fmax(cTot - c[i], 1e-5);
            tMin = fmin(tMin, cm / dd);
        }
    }
    *subDeltaT = w->O->o.cTauChem * tMin;
    *deltaT = fmin(*deltaT, *subDeltaT);
    for (int i = 0; i < nSp; ++i) { RR[i * nSp + i] += 1 / (*deltaT); w->src[i] = c[i] / (*deltaT); }
    int st = lu_decompose(w, RR, nSp, w->piv);
    if (st) { w->fail = st; return; }
    lu_backsubst(w, RR, nSp, w->piv, w->src);
    for (int i = 0; i < nSp; ++i) c[i] = fmax(0.0, w->src[i]);
    *T = mix_THa(w, c, ha, *T);
}

/* ------------------------------------------------------------------------------------------------
 * LoadBalancedChemistryModel::solveCell (reference LoadBalancedChemistryModel.C:365-420)
 * ---------------------------------------------------------------------------------------------- */
static void solve_cell(ows* w, const double* Y, int64_t ldy, int64_t cell, double* T, double* p, double rho,
                       double deltaT, double* deltaTChem, double* RR, int32_t* status) {
    const oracle_mech* M = w->M;
    const int nSp = M->nSp;
    const int solver = w->O->o.solver;
    w->fail = 0;
    if (*T > w->O->o.Treact) {
        double c0[nSp], c[nSp];
        for (int i = 0; i < nSp; ++i) { c[i] = rho * Y[i * ldy + cell] / M->d.W[i]; c0[i] = c[i]; }
        double timeLeft = deltaT;
        int guard = 0;
        while (timeLeft > O_SMALL) {
            double dt = timeLeft;
            if (solver == LBCHEM_SOLVER_EULER_IMPLICIT) euler_implicit_solve(w, c, T, p, &dt, deltaTChem);
            else if (solver == LBCHEM_SOLVER_NONE) { /* noChemistrySolver::solve does nothing */ }
            else ode_chem_solve(w, c, T, p, &dt, deltaTChem);
            if (w->fail) break;
            timeLeft -= dt;
            if (++guard > 10000000) { w->fail = LBCHEM_CELL_MAX_STEPS; break; }
        }
        *deltaTChem = fmin(*deltaTChem, w->O->o.deltaTChemMax);
        for (int i = 0; i < nSp; ++i) RR[i * ldy + cell] = (c[i] - c0[i]) * M->d.W[i] / deltaT;
        *status = w->fail ? w->fail : LBCHEM_CELL_OK;
    } else {
        for (int i = 0; i < nSp; ++i) RR[i * ldy + cell] = 0;
        *status = LBCHEM_CELL_SKIPPED;
    }
}


/* ------------------------------------------------------------------------------------------------
 * chemistryReductionMethods::DAC::reduceMechanism(c, T, p), automaticSIS off
 * (OpenFOAM v2306 chemistryModel/TDACChemistryModel/reduction/DAC/DAC.C, SURVEY A.12 [from memory]; called at
 * reference LoadBalancedTDACChemistryModel.C:388). Per reaction: omega_i; per species A of the reaction the net
 * nu_A * omega_i (a species on both sides is counted once, summed); P_A / C_A = sums of the positive / negative
 * contributions; numerators of r_AB (signed sums over the reactions that hold both A and B) are stored per A in
 * first-appearance order of B. Breadth-first search from the initial set with R(seed) = 1; an edge is followed
 * when R_u * min(|num_AB| / max(P_A, C_A), 1) >= tolerance, R_B = max(R_B, R_u r_AB); a species is queued once.
 * Reactions touching an inactive species are disabled; the simplified ordering is ascending complete index.
 * ---------------------------------------------------------------------------------------------- */
static void dac_reduce(ows* w, const double* c, double T) {
    const oracle_mech* M = w->M;
    const lbchem_mech_desc* d = &M->d;
    const int nSp = M->nSp, nR = M->nR;
    for (int i = 0; i < nSp; ++i) { w->completeC[i] = c[i]; w->PA[i] = 0; w->CA[i] = 0; w->NbInit[i] = 0; w->Rvalue[i] = 0; w->act[i] = 0; }
    for (int i = 0; i < nSp * nSp; ++i) { w->rABPos[i] = -1; w->rABOther[i] = -1; w->rABNum[i] = 0; }
    for (int r = 0; r < nR; ++r) {
        double pf, cf, pr, cr; int lRef, rRef;
        const double omegai = omega_reaction(M, w->O, r, c, T, &pf, &cf, &lRef, &pr, &cr, &rRef);
        double wA[2 * MAXSIDE]; int wAID[2 * MAXSIDE]; int nw = 0;
        for (int k = d->lhs_ptr[r]; k < d->lhs_ptr[r + 1]; ++k) {
            const int ss = d->lhs_species[k]; const double sl = -d->lhs_nu[k];
            int found = 0;
            for (int id = 0; id < nw; ++id) if (wAID[id] == ss) { wA[id] += sl * omegai; found = 1; break; }
            if (!found) { wA[nw] = sl * omegai; wAID[nw++] = ss; }
        }
        for (int k = d->rhs_ptr[r]; k < d->rhs_ptr[r + 1]; ++k) {
            const int ss = d->rhs_species[k]; const double sl = d->rhs_nu[k];
            int found = 0;
            for (int id = 0; id < nw; ++id) if (wAID[id] == ss) { wA[id] += sl * omegai; found = 1; break; }
            if (!found) { wA[nw] = sl * omegai; wAID[nw++] = ss; }
        }
        for (int id = 0; id < nw; ++id) {
            const int A = wAID[id];
            /* the other species of this reaction, each once, in lhs-then-rhs order (the FIFO of the reference) */
            for (int side = 0; side < 2; ++side) {
                const int k0 = side ? d->rhs_ptr[r] : d->lhs_ptr[r], k1 = side ? d->rhs_ptr[r + 1] : d->lhs_ptr[r + 1];
                const int32_t* sp = side ? d->rhs_species : d->lhs_species;
                for (int k = k0; k < k1; ++k) {
                    const int B = sp[k];
                    if (B == A) continue;
                    int seen = 0;   /* deltaBi: count B once per (reaction, A) */
                    for (int s2 = 0; s2 < 2 && !seen; ++s2) {
                        const int q0 = s2 ? d->rhs_ptr[r] : d->lhs_ptr[r];
                        const int q1 = (s2 == side) ? k : (s2 ? d->rhs_ptr[r + 1] : d->lhs_ptr[r + 1]);
                        if (s2 > side) break;
                        const int32_t* sq = s2 ? d->rhs_species : d->lhs_species;
                        for (int q = q0; q < q1; ++q) if (sq[q] == B) { seen = 1; break; }
                    }
                    if (seen) continue;
                    if (w->rABPos[A * nSp + B] == -1) {
                        const int pos = w->NbInit[A]++;
                        w->rABPos[A * nSp + B] = pos;
                        w->rABNum[A * nSp + pos] = wA[id];
                        w->rABOther[A * nSp + pos] = B;
                    } else {
                        w->rABNum[A * nSp + w->rABPos[A * nSp + B]] += wA[id];
                    }
                }
            }
            if (wA[id] > 0.0) w->PA[A] += wA[id]; else w->CA[A] += -wA[id];
        }
    }
    /* search from the initial set */
    int head = 0, tail = 0, nActive = 0;
    for (int i = 0; i < w->O->dac_ninit; ++i) {
        const int q = w->O->dac_init[i];
        if (q < 0 || q >= nSp || w->act[q]) continue;
        w->act[q] = 1; nActive++; w->queue[tail++] = q; w->Rvalue[q] = 1.0;
    }
    const double tol = w->O->dac_tol;
    while (head < tail) {
        const int u = w->queue[head++];
        const double Den = fmax(w->PA[u], w->CA[u]);
        if (Den > O_VSMALL) {
            for (int v = 0; v < w->NbInit[u]; ++v) {
                const int other = w->rABOther[u * nSp + v];
                double rAB = fabs(w->rABNum[u * nSp + v]) / Den;
                if (rAB > 1) rAB = 1;
                const double Rt = w->Rvalue[u] * rAB;
                if (Rt >= tol) {
                    if (Rt > w->Rvalue[other]) w->Rvalue[other] = Rt;
                    if (!w->act[other]) { w->act[other] = 1; nActive++; w->queue[tail++] = other; }
                }
            }
        }
    }
    for (int r = 0; r < nR; ++r) {
        int dis = 0;
        for (int k = d->lhs_ptr[r]; k < d->lhs_ptr[r + 1] && !dis; ++k) if (!w->act[d->lhs_species[k]]) dis = 1;
        for (int k = d->rhs_ptr[r]; k < d->rhs_ptr[r + 1] && !dis; ++k) if (!w->act[d->rhs_species[k]]) dis = 1;
        w->rdis[r] = (unsigned char)dis;
    }
    int j = 0;
    for (int i = 0; i < nSp; ++i) {
        if (w->act[i]) { w->s2c[j] = i; w->simplC[j] = c[i]; w->c2s[i] = j++; } else w->c2s[i] = -1;
    }
    w->NsDAC = nActive;
}

/* LoadBalancedTDACChemistryModel::solveCell with mechRed()->active() (reference LoadBalancedTDACChemistryModel.C:360-435).
 * The reduced call passes cData.T() for BOTH T and p (:406-409): p is dead for perfectGas, and on return the aliased
 * reference is first set to the final temperature and then overwritten by the p slot (= the initial T), so the cell
 * temperature comes back unchanged. Reproduced here by passing the same pointer twice. */
static void solve_cell_tdac(ows* w, const double* Y, int64_t ldy, int64_t cell, double* T, double* p, double rho,
                            double deltaT, double* deltaTChem, double* RR, int32_t* status, int32_t* nActive) {
    const oracle_mech* M = w->M;
    const int nSp = M->nSp;
    const int nFull = w->n;
    (void)p;
    w->fail = 0;
    double c0[nSp], c[nSp];
    for (int i = 0; i < nSp; ++i) { c[i] = rho * Y[i * ldy + cell] / M->d.W[i]; c0[i] = c[i]; }
    dac_reduce(w, c, *T);
    if (nActive) *nActive = w->NsDAC;
    w->red = 1; w->n = w->NsDAC + 2;
    double timeLeft = deltaT;
    int guard = 0;
    while (timeLeft > O_SMALL) {
        double dt = timeLeft;
        for (int i = 0; i < nSp; ++i) w->completeC[i] = c[i];
        ode_chem_solve(w, w->simplC, T, T, &dt, deltaTChem);
        for (int i = 0; i < w->NsDAC; ++i) c[w->s2c[i]] = w->simplC[i];
        if (w->fail) break;
        timeLeft -= dt;
        if (++guard > 10000000) { w->fail = LBCHEM_CELL_MAX_STEPS; break; }
    }
    w->red = 0; w->n = nFull;
    /* the TDAC variant has no deltaTChemMax clip in solveCell; RR as in its solve<DeltaT> (:905-915) */
    for (int i = 0; i < nSp; ++i) RR[i * ldy + cell] = (c[i] - c0[i]) * M->d.W[i] / deltaT;
    *status = w->fail ? w->fail : LBCHEM_CELL_OK;
}

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

/*
 * solveCellList over a batch (reference :423-450): per-cell wall-clock seconds go to cpu_time[]
 * (the reference's cost model); attempts[] gets the attempted-sub-step count (the GPU cost model).
 * counters[6] = sums of {n_steps, n_attempts, n_jac, n_lu, n_solve, n_rhs}. nthreads static-partitions
 * the cells into contiguous ranges (one range per thread = one MPI rank's sub-domain).
 */
typedef struct {
    const oracle_mech* M; const oracle_opts* O;
    int64_t lo, hi, ldy;
    const double *Y, *T, *p, *rho, *deltaT; double deltaT_uniform;
    double *deltaTChem, *RR, *Tout, *cpu_time, *attempts; int32_t* status;
    int32_t* nActive;      /* DAC: NsDAC_ per cell */
    int64_t cnt[6];
} batch_job;

static void* batch_worker(void* arg) {
    batch_job* j = (batch_job*)arg;
    ows* w = ows_create(j->M, j->O);
    for (int64_t cidx = j->lo; cidx < j->hi; ++cidx) {
        memset(&w->cnt, 0, sizeof(w->cnt));
        double Tc = j->T[cidx], pc = j->p[cidx];
        int32_t st = 0;
        double t0 = now_s();
        if (j->O->dac_active)
            solve_cell_tdac(w, j->Y, j->ldy, cidx, &Tc, &pc, j->rho[cidx], j->deltaT ? j->deltaT[cidx] : j->deltaT_uniform,
                            &j->deltaTChem[cidx], j->RR, &st, j->nActive ? &j->nActive[cidx] : NULL);
        else
        solve_cell(w, j->Y, j->ldy, cidx, &Tc, &pc, j->rho[cidx], j->deltaT ? j->deltaT[cidx] : j->deltaT_uniform,
                   &j->deltaTChem[cidx], j->RR, &st);
        double t1 = now_s();
        if (j->Tout) j->Tout[cidx] = Tc;
        if (j->cpu_time) j->cpu_time[cidx] = t1 - t0;
        if (j->attempts) j->attempts[cidx] = (double)w->cnt.n_attempts;
        if (j->status) j->status[cidx] = st;
        j->cnt[0] += w->cnt.n_steps; j->cnt[1] += w->cnt.n_attempts; j->cnt[2] += w->cnt.n_jac;
        j->cnt[3] += w->cnt.n_lu; j->cnt[4] += w->cnt.n_solve; j->cnt[5] += w->cnt.n_rhs;
    }
    ows_destroy(w);
    return NULL;
}

/* reduction{} for the batches that follow (process-wide; the tests set it and reset it around a call) */
static struct { int active; double tol; int ninit; int init[64]; int32_t* nActive; } g_dac;
void oracle_set_reduction(int active, double tol, int n_init, const int32_t* init_species, int32_t* n_active_out) {
    g_dac.active = active; g_dac.tol = tol; g_dac.ninit = n_init > 64 ? 64 : n_init;
    for (int i = 0; i < g_dac.ninit; ++i) g_dac.init[i] = init_species[i];
    g_dac.nActive = n_active_out;
}

/* probe: the active-species flags DAC gives for one state (c[nSp] in kmol/m3, T) */
int oracle_dac_reduce(const oracle_mech* M, const lbchem_options* opts, double tol, int n_init, const int32_t* init_species,
                      const double* c, double T, int32_t* active_out, int32_t* disabled_out) {
    oracle_opts O; memset(&O, 0, sizeof(O)); O.o = *opts;
    O.dac_active = 1; O.dac_tol = tol; O.dac_ninit = n_init > 64 ? 64 : n_init;
    for (int i = 0; i < O.dac_ninit; ++i) O.dac_init[i] = init_species[i];
    ows* w = ows_create(M, &O);
    dac_reduce(w, c, T);
    for (int i = 0; i < M->nSp; ++i) active_out[i] = w->act[i];
    if (disabled_out) for (int r = 0; r < M->nR; ++r) disabled_out[r] = w->rdis[r];
    const int ns = w->NsDAC;
    ows_destroy(w);
    return ns;
}

int oracle_solve_batch(const oracle_mech* M, const lbchem_options* opts, int seulex_errold_max, int64_t n,
                       int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
                       const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR,
                       double* Tout, double* cpu_time, double* attempts, int32_t* status, int64_t* counters,
                       int nthreads) {
    oracle_opts O;
    memset(&O, 0, sizeof(O));
    O.o = *opts;
    O.seulex_errold_max = seulex_errold_max || opts->seulex_errold_max;
    O.dac_active = g_dac.active; O.dac_tol = g_dac.tol; O.dac_ninit = g_dac.ninit;
    memcpy(O.dac_init, g_dac.init, sizeof(O.dac_init));
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    batch_job jobs[256];
    pthread_t th[256];
    for (int t = 0; t < nthreads; ++t) {
        batch_job* j = &jobs[t];
        memset(j, 0, sizeof(*j));
        j->M = M; j->O = &O; j->ldy = ldy;
        j->lo = n * t / nthreads; j->hi = n * (t + 1) / nthreads;
        j->Y = Y; j->T = T; j->p = p; j->rho = rho; j->deltaT = deltaT; j->deltaT_uniform = deltaT_uniform;
        j->deltaTChem = deltaTChem; j->RR = RR; j->Tout = Tout; j->cpu_time = cpu_time; j->attempts = attempts;
        j->status = status; j->nActive = g_dac.active ? g_dac.nActive : NULL;
        if (t > 0) pthread_create(&th[t], NULL, batch_worker, j);
    }
    batch_worker(&jobs[0]);
    for (int t = 1; t < nthreads; ++t) pthread_join(th[t], NULL);
    if (counters) {
        for (int q = 0; q < 6; ++q) counters[q] = 0;
        for (int t = 0; t < nthreads; ++t) for (int q = 0; q < 6; ++q) counters[q] += jobs[t].cnt[q];
    }
    return 0;
}

/* StandardChemistryModel::tc(): tc = nReaction * cSum / (SMALL + sum_r sum_rhs nu * pf * cf), c = rho Y / W unclamped */
void oracle_tc(const oracle_mech* M, const lbchem_options* opts, int64_t n, int64_t ldy, const double* Y, const double* T,
               const double* rho, double* tc) {
    oracle_opts O; memset(&O, 0, sizeof(O)); O.o = *opts;
    const lbchem_mech_desc* d = &M->d;
    double c[M->nSp];
    for (int64_t cell = 0; cell < n; ++cell) {
        double cSum = 0, acc = O_SMALL;
        for (int i = 0; i < M->nSp; ++i) { c[i] = rho[cell] * Y[i * ldy + cell] / d->W[i]; cSum += c[i]; }
        for (int r = 0; r < M->nR; ++r) {
            double pf, cf, pr, cr; int lRef, rRef;
            omega_reaction(M, &O, r, c, T[cell], &pf, &cf, &lRef, &pr, &cr, &rRef);
            for (int k = d->rhs_ptr[r]; k < d->rhs_ptr[r + 1]; ++k) acc += d->rhs_nu[k] * pf * cf;
        }
        tc[cell] = M->nR * cSum / acc;
    }
}

/* probes used by the tests: RHS and Jacobian of one state (y = c[nSp], T, p) */
void oracle_derivatives(const oracle_mech* M, const lbchem_options* opts, const double* y, double* dydt) {
    oracle_opts O; memset(&O, 0, sizeof(O)); O.o = *opts;
    ows* w = ows_create(M, &O);
    derivatives(w, y, dydt);
    ows_destroy(w);
}
void oracle_jacobian(const oracle_mech* M, const lbchem_options* opts, const double* y, double* dfdx, double* dfdy) {
    oracle_opts O; memset(&O, 0, sizeof(O)); O.o = *opts;
    ows* w = ows_create(M, &O);
    jacobian(w, y, dfdx, dfdy);
    ows_destroy(w);
}
void oracle_rates(const oracle_mech* M, const lbchem_options* opts, const double* c, double T, double* kf, double* kr) {
    oracle_opts O; memset(&O, 0, sizeof(O)); O.o = *opts;
    for (int r = 0; r < M->nR; ++r) { kf[r] = reaction_kf(M, r, T, c); kr[r] = reaction_kr(M, &O, r, kf[r], T); }
}
int oracle_lu_solve(int m, double* A, double* b) {
    oracle_opts O; memset(&O, 0, sizeof(O)); O.o.rel_tol = 1e-4;
    oracle_mech M; memset(&M, 0, sizeof(M)); M.nEqn = m;
    ows* w = ows_create(&M, &O);
    int st = lu_decompose(w, A, m, w->piv);
    if (!st) lu_backsubst(w, A, m, w->piv, b);
    ows_destroy(w);
    return st;
}

/* ------------------------------------------------------------------------------------------------
 * Balancing plan: getSortedCPUTimesOnProcessor + updateProcessorBalancing
 * (reference LoadBalancedChemistryModel.C:311-347, :170-308; TDAC variant
 *  LoadBalancedTDACChemistryModel.C:149-287) and cellsToSend (:143-167)
 * ---------------------------------------------------------------------------------------------- */
typedef struct { double t; int proc; int cells; } sorted_entry;

/* std::greater<std::pair<scalar, Pair<label>>>: lexicographic on (t, proc, cells) */
static int entry_greater(const sorted_entry* a, const sorted_entry* b) {
    if (a->t > b->t) return 1;
    if (b->t > a->t) return 0;
    if (a->proc > b->proc) return 1;
    if (b->proc > a->proc) return 0;
    return a->cells > b->cells;
}

int oracle_balance_plan(int32_t P, const double* cost, const int32_t* cells, double minFrac, int32_t tdac,
                        int32_t* send_ptr, int32_t* send_to, double* send_frac,
                        int32_t* recv_ptr, int32_t* recv_from) {
    sorted_entry* S = (sorted_entry*)malloc(sizeof(sorted_entry) * P);
    for (int i = 0; i < P; ++i) { S[i].t = cost[i]; S[i].proc = i; S[i].cells = cells ? cells[i] : 0; }
    /* stable insertion sort, descending */
    for (int i = 1; i < P; ++i) {
        sorted_entry e = S[i];
        int j = i - 1;
        while (j >= 0 && entry_greater(&e, &S[j])) { S[j + 1] = S[j]; --j; }
        S[j + 1] = e;
    }
    double avg = 0;
    for (int i = 0; i < P; ++i) avg += S[i].t;
    avg /= (double)P;
    /* per-rank push-order lists, capacity P each */
    int* ns = (int*)calloc(P, sizeof(int));
    int* nr = (int*)calloc(P, sizeof(int));
    int* sto = (int*)malloc(sizeof(int) * P * P);
    double* sfr = (double*)malloc(sizeof(double) * P * P);
    int* rfr = (int*)malloc(sizeof(int) * P * P);
    for (int i = 0; i < P; ++i) {
        const int procI = S[i].proc;
        const double tI = S[i].t;
        double over = tI - avg;
        if (tdac && over < 0) continue;
        for (int k = P - 1; k > 0; --k) {
            const int procK = S[k].proc;
            const double cap = avg - S[k].t;
            if (cap <= 0) continue;
            const double newCap = cap - over;
            if ((fmin(cap, over) / tI) < minFrac) continue;
            if (newCap > 0) {
                S[k].t += over; S[i].t -= over;
                rfr[procK * P + nr[procK]++] = procI;
                sto[procI * P + ns[procI]] = procK;
                sfr[procI * P + ns[procI]++] = tdac ? over : over / tI;
                over = 0;
                break;
            } else {
                S[k].t += cap; S[i].t -= cap; over -= cap;
                rfr[procK * P + nr[procK]++] = procI;
                sto[procI * P + ns[procI]] = procK;
                sfr[procI * P + ns[procI]++] = tdac ? cap : cap / tI;
            }
        }
    }
    int a = 0, b = 0;
    for (int r = 0; r < P; ++r) {
        send_ptr[r] = a; recv_ptr[r] = b;
        for (int q = 0; q < ns[r]; ++q) { send_to[a] = sto[r * P + q]; send_frac[a] = sfr[r * P + q]; ++a; }
        for (int q = 0; q < nr[r]; ++q) recv_from[b++] = rfr[r * P + q];
    }
    send_ptr[P] = a; recv_ptr[P] = b;
    free(S); free(ns); free(nr); free(sto); free(sfr); free(rfr);
    return 0;
}

int32_t oracle_cells_to_send(const double* cell_cost, int32_t n, double target, int32_t start) {
    int32_t end = n;
    double cpuTime = 0;
    for (int32_t i = start; i < n; ++i) {
        if (cpuTime >= target) { end = i; break; }
        cpuTime += cell_cost[i];
    }
    return end;
}
