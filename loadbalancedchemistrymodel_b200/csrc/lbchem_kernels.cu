// lbchem_kernels.cu — hand-written sm_100a FP64 kernels for the per-cell stiff chemistry integration.
//
// One thread group (TPC = 32 or 64 threads) integrates one cell; a CTA holds several groups, each
// pulling cells from a global atomic work queue (persistent kernel), so cells needing many sub-steps
// do not stall their neighbours. All per-cell state lives in shared memory.
//
// Restates (reference call site LoadBalancedChemistryModel.C:392-399):
//   solveCell -> ode::solve -> ODESolver::solve -> adaptiveSolver/Rosenbrock34 | seulex,
//   EulerImplicit::solve, StandardChemistryModel::{omega,derivatives,jacobian}, reaction rates,
//   species::thermo::Kc, LUDecompose/LUBacksubstitute.
// No tensor cores: the per-cell LU is at most ~55x55.
#include <cuda_runtime.h>
#include <math.h>
#include "lbchem_device.cuh"
#include "../../include/lbchem.h"

// ------------------------------------------------------------------------------------------------
// thread group = the TPC threads working on one cell
// ------------------------------------------------------------------------------------------------
template <int TPC>
struct Grp {
    int tid;        // thread index inside the group
    int gid;        // group index inside the CTA
    double* red;    // 2 slots x 2 warps of reduction scratch (shared)
    int tog;
    unsigned mask;  // lanes of this group inside its warp (TPC < 32: several cells share one warp, each an independent
                    // sub-warp group that only ever synchronises and shuffles inside its own lanes)

    static constexpr int kHalf = TPC < 32 ? TPC / 2 : 16;
    __device__ __forceinline__ void sync() const {
        if (TPC <= 32) __syncwarp(mask);
        else asm volatile("bar.sync %0, %1;" ::"r"(gid + 1), "r"(TPC) : "memory");
    }
    __device__ __forceinline__ double sum(double v) {
#pragma unroll
        for (int o = kHalf; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
        if (TPC <= 32) return v;
        double* r = red + 4 * tog; tog ^= 1;
        if ((tid & 31) == 0) r[tid >> 5] = v;
        sync();
        double s = r[0];
#pragma unroll
        for (int w = 1; w < TPC / 32; ++w) s += r[w];
        return s;
    }
    __device__ __forceinline__ double max(double v) {
#pragma unroll
        for (int o = kHalf; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(mask, v, o));
        if (TPC <= 32) return v;
        double* r = red + 4 * tog; tog ^= 1;
        if ((tid & 31) == 0) r[tid >> 5] = v;
        sync();
        double s = r[0];
#pragma unroll
        for (int w = 1; w < TPC / 32; ++w) s = fmax(s, r[w]);
        return s;
    }
    // NaN-propagating max for error norms: returns NaN if any lane holds NaN
    __device__ __forceinline__ double maxnan(double v) {
        double bad = (v != v) ? 1.0 : 0.0;
        double m = max(v != v ? 0.0 : v);
        double b = max(bad);
        return b > 0.0 ? __longlong_as_double(0x7ff8000000000000LL) : m;
    }
    __device__ __forceinline__ int any(int p) {
        int a = __any_sync(mask, p);
        if (TPC <= 32) return a;
        double* r = red + 4 * tog; tog ^= 1;
        if ((tid & 31) == 0) r[tid >> 5] = a ? 1.0 : 0.0;
        sync();
        double s = r[0];
#pragma unroll
        for (int w = 1; w < TPC / 32; ++w) s += r[w];
        return s > 0.0;
    }
    // lexicographic arg-max of (key, idx); ties on key go to the larger idx (LUDecompose's ">=" scan)
    __device__ __forceinline__ int argmax(unsigned long long key, int idx) {
#pragma unroll
        for (int o = kHalf; o > 0; o >>= 1) {
            unsigned long long k2 = __shfl_xor_sync(mask, key, o);
            int i2 = __shfl_xor_sync(mask, idx, o);
            if (k2 > key || (k2 == key && i2 > idx)) { key = k2; idx = i2; }
        }
        if (TPC <= 32) return idx;
        double* r = red + 4 * tog; tog ^= 1;
        if ((tid & 31) == 0) {
            r[tid >> 5] = __longlong_as_double((long long)key);
            reinterpret_cast<int*>(r + 2)[tid >> 5] = idx;
        }
        sync();
        unsigned long long bk = (unsigned long long)__double_as_longlong(r[0]);
        int bi = reinterpret_cast<int*>(r + 2)[0];
#pragma unroll
        for (int w = 1; w < TPC / 32; ++w) {
            unsigned long long k2 = (unsigned long long)__double_as_longlong(r[w]);
            int i2 = reinterpret_cast<int*>(r + 2)[w];
            if (k2 > bk || (k2 == bk && i2 > bi)) { bk = k2; bi = i2; }
        }
        return bi;
    }
};

// per-cell shared-memory workspace (pointers into the group's slice of dynamic shared memory)
struct Work {
    double *Y0, *YN, *D0, *DX, *K1, *K2, *K3, *K4, *DY;   // nEqn each
    double *CC;        // nSp  clamped concentrations
    double *GL, *GH;   // nSp  (h - T s)/R for the low / high janaf set
    double *ME;        // nEff third-body concentrations
    double *VV, *RD;   // nEqn LU implicit-scaling factors / reciprocal pivots
    double *R2;        // max(nTerms + nR, 2 nR): omega_r and D-terms, then omega(T+d), omega(T-d)
    double *A;         // nEqn x lda   Jacobian, then LU factors
    double *A2;        // second matrix (seulex: LU copy; EulerImplicit unused) or null
    double *TB;        // seulex extrapolation table (kMaxx x nEqn) or null
    int* perm;         // nEqn
    // counters (registers would do; kept here so all threads agree)
    int nSteps, nAttempts, nJac, nLU, nSolve, nRhs;
};

__host__ __device__ inline size_t work_doubles(const DevMech& M, int solver) {
    size_t n = M.nEqn;
    size_t r2 = (size_t)M.nTerms + M.nR;
    if (r2 < 2 * (size_t)M.nR) r2 = 2 * (size_t)M.nR;
    size_t d = 9 * n + 3 * (size_t)M.nSp + M.nEff + 2 * n + r2 + n * M.lda;
    if (solver == LBCHEM_SOLVER_SEULEX) d += n * M.lda + 12 * n + 5 * n;   // A2, TB, seulex vectors
    d += 8;                                                               // reduction scratch
    d += (n + 1) / 2 + 1;                                                 // perm (ints)
    return d;
}

// ------------------------------------------------------------------------------------------------
// janaf thermo (janafThermoI.H), Horner forms identical to the oracle
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double janaf_cp(const double* a, int st, double T) {
    return (((a[4 * st] * T + a[3 * st]) * T + a[2 * st]) * T + a[1 * st]) * T + a[0];
}
__device__ __forceinline__ double janaf_ha(const double* q, int st, double T) {
    // q: a4/5, a3/4, a2/3, a1/2, a0, a5
    return ((((q[0] * T + q[1 * st]) * T + q[2 * st]) * T + q[3 * st]) * T + q[4 * st]) * T + q[5 * st];
}
__device__ __forceinline__ double janaf_s(const double* q, int st, double T, double lnT) {
    // q: a4/4, a3/3, a2/2, a1, a0, a6
    return (((q[0] * T + q[1 * st]) * T + q[2 * st]) * T + q[3 * st]) * T + q[4 * st] * lnT + q[5 * st];
}

__device__ __forceinline__ double ipow(double c, double e) {
    if (e == 1.0) return c;
    if (e == 2.0) return c * c;
    if (e == 3.0) return c * c * c;
    if (e == 0.0) return 1.0;
    return pow(c, e);
}

// Arrhenius: A T^beta exp(-Ta/T), evaluated as A exp(beta ln T - Ta/T) (one exp; DESIGN.md "rate constants")
__device__ __forceinline__ double arrhenius(double A, double beta, double Ta, double lnT, double invT) {
    const bool hb = fabs(beta) > LB_VSMALL, ht = fabs(Ta) > LB_VSMALL;
    if (!hb && !ht) return A;
    double arg = 0.0;
    if (hb) arg = beta * lnT;
    if (ht) arg -= Ta * invT;
    return A * exp(arg);
}

// species Gibbs functions g_i = h_i/R - T s_i/R for both janaf sets (inputs of Kc)
template <int TPC>
__device__ __forceinline__ void species_gibbs(const Grp<TPC>& g, const DevMech& M, const Work& w, double T, double lnT) {
    const int nSp = M.nSp;
    for (int i = g.tid; i < nSp; i += TPC) {
        w.GL[i] = janaf_ha(M.haLo + i, nSp, T) - T * janaf_s(M.sLo + i, nSp, T, lnT);
        w.GH[i] = janaf_ha(M.haHi + i, nSp, T) - T * janaf_s(M.sHi + i, nSp, T, lnT);
    }
}

// third-body concentrations M_e = sum_i eff[e][i] c_i (thirdBodyEfficiencies::M), sequential like the reference
template <int TPC>
__device__ __forceinline__ void third_body(const Grp<TPC>& g, const DevMech& M, const Work& w) {
    for (int e = g.tid; e < M.nEff; e += TPC) {
        const double* ef = M.eff + (size_t)e * M.nSp;
        double m = 0.0;
        for (int i = 0; i < M.nSp; ++i) m = fma(ef[i], w.CC[i], m);
        w.ME[e] = m;
    }
}

// forward / reverse rate constants of reaction r (Reaction::kf, kr; A.6). GL/GH and ME must be current.
__device__ __forceinline__ void rate_constants(const DevMech& M, const DevOpts& O, const Work& w, int r, double T,
                                               double lnT, double invT, double& kf, double& kr) {
    const int nR = M.nR;
    const int ty = M.rtype[r];
    double k = arrhenius(M.A[r], M.beta[r], M.Ta[r], lnT, invT);
    if (ty == LBCHEM_R_THIRD_BODY) {
        k = w.ME[M.effIdx[r]] * k;
    } else if (ty >= LBCHEM_R_FALLOFF_LINDEMANN) {
        const double k0 = arrhenius(M.A0[r], M.beta0[r], M.Ta0[r], lnT, invT);
        const double kInf = k;
        const double Pr = k0 * w.ME[M.effIdx[r]] / kInf;
        double F = 1.0;
        if (ty == LBCHEM_R_FALLOFF_TROE) {
            const double al = M.troe[r], Tsss = M.troe[nR + r], Ts = M.troe[2 * nR + r], Tss = M.troe[3 * nR + r];
            const double logFcent = log10(fmax((1 - al) * exp(-T / Tsss) + al * exp(-T / Ts) + exp(-Tss / T), LB_SMALL));
            const double cc = -0.4 - 0.67 * logFcent;
            const double nn = 0.75 - 1.27 * logFcent;
            const double logPr = log10(fmax(Pr, LB_SMALL));
            const double q = (logPr + cc) / (nn - 0.14 * (logPr + cc));
            F = pow(10.0, logFcent / (1.0 + q * q));
        }
        k = kInf * (Pr / (1 + Pr)) * F;
    }
    kf = k;
    kr = 0.0;
    if (M.rev[r]) {
        const double* G = (T < M.rTc[r]) ? w.GL : w.GH;
        double gs = 0.0;
        const int nl = M.nl[r], nr = M.nr[r];
        for (int s = 0; s < nr; ++s) gs = fma(M.rnu[s * nR + r], G[M.rsp[s * nR + r]], gs);
        for (int s = 0; s < nl; ++s) gs = fma(-M.lnu[s * nR + r], G[M.lsp[s * nR + r]], gs);
        const double arg = -gs * invT;
        double Kc = (arg < 600.0) ? exp(arg) : LB_VGREAT;
        const double dn = M.dn[r];
        if (dn != 0.0) {
            const double x = LB_PSTD / (O.RR * T);
            if (dn == 1.0) Kc *= x;
            else if (dn == -1.0) Kc *= 1.0 / x;
            else Kc *= pow(x, dn);
        }
        kr = k / fmax(Kc, O.kcFloor);
    }
}

// ------------------------------------------------------------------------------------------------
// StandardChemistryModel::omega for one reaction, literal form with the reference-species logic
// (A.7). Used by EulerImplicit and by mechanisms with fractional exponents.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double omega_faithful(const DevMech& M, const double* c, int r, double kf, double kr,
                                                 double& pf, double& cf, int& lRef, double& pr, double& cr, int& rRef) {
    const int nR = M.nR;
    const int nl = M.nl[r], nr = M.nr[r];
    int sl = 0;
    lRef = M.lsp[r];
    pf = kf;
    for (int s = 1; s < nl; ++s) {
        const int si = M.lsp[s * nR + r];
        if (c[si] < c[lRef]) {
            pf *= pow(fmax(c[lRef], 0.0), M.lex[sl * nR + r]);
            lRef = si; sl = s;
        } else {
            pf *= pow(fmax(c[si], 0.0), M.lex[s * nR + r]);
        }
    }
    cf = fmax(c[lRef], 0.0);
    {
        const double ex = M.lex[sl * nR + r];
        if (ex < 1.0) { if (cf > LB_SMALL) pf *= pow(cf, ex - 1.0); else pf = 0.0; }
        else pf *= pow(cf, ex - 1.0);
    }
    int sr = 0;
    rRef = M.rsp[r];
    pr = kr;
    for (int s = 1; s < nr; ++s) {
        const int si = M.rsp[s * nR + r];
        if (c[si] < c[rRef]) {
            pr *= pow(fmax(c[rRef], 0.0), M.rex[sr * nR + r]);
            rRef = si; sr = s;
        } else {
            pr *= pow(fmax(c[si], 0.0), M.rex[s * nR + r]);
        }
    }
    cr = fmax(c[rRef], 0.0);
    {
        const double ex = M.rex[sr * nR + r];
        if (ex < 1.0) { if (cr > LB_SMALL) pr *= pow(cr, ex - 1.0); else pr = 0.0; }
        else pr *= pow(cr, ex - 1.0);
    }
    return pf * cf - pr * cr;
}

// net rate of reaction r from rate constants and clamped concentrations
__device__ __forceinline__ double omega_r(const DevMech& M, const double* c, int r, double kf, double kr) {
    const int nR = M.nR;
    if (M.fracExp) {
        double pf, cf, pr, cr; int lRef, rRef;
        return omega_faithful(M, c, r, kf, kr, pf, cf, lRef, pr, cr, rRef);
    }
    const int nl = M.nl[r], nr = M.nr[r];
    double f = kf, b = kr;
    for (int s = 0; s < nl; ++s) f *= ipow(c[M.lsp[s * nR + r]], M.lex[s * nR + r]);
    for (int s = 0; s < nr; ++s) b *= ipow(c[M.rsp[s * nR + r]], M.rex[s * nR + r]);
    return f - b;
}

// d(omega_r)/d(c_j) factors of StandardChemistryModel::jacobian (A.9): one value per lhs slot (times kf)
// and per rhs slot (times kr), written to D[0 .. nl+nr)
__device__ __forceinline__ void dterms_r(const DevMech& M, const double* c, int r, double kf0, double kr0, double* D) {
    const int nR = M.nR;
    const int nl = M.nl[r], nr = M.nr[r];
    for (int j = 0; j < nl; ++j) {
        double kf = kf0;
        for (int i = 0; i < nl; ++i) {
            const double ci = c[M.lsp[i * nR + r]];
            const double el = M.lex[i * nR + r];
            if (i == j) {
                if (el < 1.0) { if (ci > LB_SMALL) kf *= el * pow(ci + LB_VSMALL, el - 1.0); else kf = 0.0; }
                else kf *= el * ipow(ci, el - 1.0);
            } else {
                kf *= ipow(ci, el);
            }
        }
        D[j] = kf;
    }
    for (int j = 0; j < nr; ++j) {
        double kr = kr0;
        for (int i = 0; i < nr; ++i) {
            const double ci = c[M.rsp[i * nR + r]];
            const double er = M.rex[i * nR + r];
            if (i == j) {
                if (er < 1.0) { if (ci > LB_SMALL) kr *= er * pow(ci + LB_VSMALL, er - 1.0); else kr = 0.0; }
                else kr *= er * ipow(ci, er - 1.0);
            } else {
                kr *= ipow(ci, er);
            }
        }
        D[nl + j] = kr;
    }
}

// species gather: dcdt_i = sum coef * om[reac] in reaction order (the reference's accumulation order)
__device__ __forceinline__ double gather_species(const DevMech& M, const double* om, int i) {
    double acc = 0.0;
    const int e = M.spPtr[i + 1];
    for (int k = M.spPtr[i]; k < e; ++k) acc = fma(M.spCoef[k], om[M.spReac[k]], acc);
    return acc;
}

// clamp concentrations, refresh T-dependent species data and third-body sums; ends with a group sync
template <int TPC>
__device__ __forceinline__ void prepare_state(Grp<TPC>& g, const DevMech& M, Work& w, const double* y, double T) {
    for (int i = g.tid; i < M.nSp; i += TPC) w.CC[i] = fmax(y[i], 0.0);
    species_gibbs(g, M, w, T, log(T));
    g.sync();
    third_body(g, M, w);
    g.sync();
}

// StandardChemistryModel::derivatives(t, y, dydt) (A.8). y, dydt: shared vectors of nEqn.
template <int TPC>
__device__ void derivatives(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, const double* y, double* dydt) {
    const int nSp = M.nSp;
    g.sync();   // y is final; previous readers of CC / R2 / dydt are done
    const double T = y[nSp];
    const double lnT = log(T), invT = 1.0 / T;
    prepare_state(g, M, w, y, T);
    double* om = w.R2;
    for (int r = g.tid; r < M.nR; r += TPC) {
        double kf, kr;
        rate_constants(M, O, w, r, T, lnT, invT, kf, kr);
        om[r] = omega_r(M, w.CC, r, kf, kr);
    }
    g.sync();
    double rho = 0.0, cp = 0.0, dT = 0.0;
    for (int i = g.tid; i < nSp; i += TPC) {
        const double d = gather_species(M, om, i);
        dydt[i] = d;
        const double ci = w.CC[i];
        const bool lo = T < M.Tc[i];
        rho = fma(M.W[i], ci, rho);
        cp = fma(ci, O.RR * janaf_cp((lo ? M.cpLo : M.cpHi) + i, nSp, T), cp);
        dT = fma(O.RR * janaf_ha((lo ? M.haLo : M.haHi) + i, nSp, T), d, dT);
    }
    rho = g.sum(rho); cp = g.sum(cp); dT = g.sum(dT);
    cp /= rho;
    dT /= rho * cp;
    if (g.tid == 0) { dydt[nSp] = -dT; dydt[nSp + 1] = 0.0; }
    if (g.tid == 0) w.nRhs++;
    g.sync();
}

// StandardChemistryModel::jacobian(t, y, dfdx, dfdy) (A.9): analytic in c, central difference in T.
// dfdx receives the species rates (quirk A.4) or zeros.
template <int TPC>
__device__ void jacobian(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, const double* y, double* dfdx, double* J) {
    const int nSp = M.nSp, n = M.nEqn, lda = M.lda, nR = M.nR;
    g.sync();
    const double T = y[nSp];
    for (int k = g.tid; k < n * lda; k += TPC) J[k] = 0.0;
    prepare_state(g, M, w, y, T);
    {
        const double lnT = log(T), invT = 1.0 / T;
        double* om = w.R2;
        double* D = w.R2 + nR;
        for (int r = g.tid; r < nR; r += TPC) {
            double kf, kr;
            rate_constants(M, O, w, r, T, lnT, invT, kf, kr);
            om[r] = omega_r(M, w.CC, r, kf, kr);
            dterms_r(M, w.CC, r, kf, kr, D + M.dbase[r]);
        }
    }
    g.sync();
    for (int i = g.tid; i < n; i += TPC) dfdx[i] = (i < nSp && O.dfdxIsDcdt) ? gather_species(M, w.R2, i) : 0.0;
    {
        const double* D = w.R2 + nR;
        const int e1 = M.thrPtr[g.tid + 1];
        for (int e = M.thrPtr[g.tid]; e < e1; ++e) {
            double acc = 0.0;
            const int t1 = M.entPtr[e + 1];
            for (int t = M.entPtr[e]; t < t1; ++t) acc = fma(M.termCoef[t], D[M.termIdx[t]], acc);
            J[M.entRow[e] * lda + M.entCol[e]] = acc;
        }
    }
    g.sync();
    const double delta = 1.0e-3;
    for (int pass = 0; pass < 2; ++pass) {
        const double Tp = pass == 0 ? T + delta : T - delta;
        const double lnT = log(Tp), invT = 1.0 / Tp;
        species_gibbs(g, M, w, Tp, lnT);
        g.sync();
        double* om = w.R2 + pass * nR;
        for (int r = g.tid; r < nR; r += TPC) {
            double kf, kr;
            rate_constants(M, O, w, r, Tp, lnT, invT, kf, kr);
            om[r] = omega_r(M, w.CC, r, kf, kr);
        }
        g.sync();
    }
    for (int i = g.tid; i < nSp; i += TPC) {
        const double dp = gather_species(M, w.R2, i);
        const double dm = gather_species(M, w.R2 + nR, i);
        J[i * lda + nSp] = 0.5 * (dp - dm) / delta;
    }
    if (g.tid == 0) w.nJac++;
    g.sync();
}

// ------------------------------------------------------------------------------------------------
// LUDecompose / LUBacksubstitute (A.10). Right-looking elimination performs, element by element,
// the same sequence of fused multiply-subtracts as the reference's Crout loops, with the same
// implicit-scaling pivot rule (ties -> last row). Rows are physically swapped in shared memory.
// ------------------------------------------------------------------------------------------------
template <int TPC>
__device__ int lu_decompose(Grp<TPC>& g, const DevMech& M, Work& w, double* A, int n) {
    const int lda = M.lda;
    int sing = 0;
    g.sync();
    for (int i = g.tid; i < n; i += TPC) {
        double largest = 0.0;
        for (int j = 0; j < n; ++j) largest = fmax(largest, fabs(A[i * lda + j]));
        if (largest == 0.0) sing = 1;
        w.VV[i] = 1.0 / largest;
        w.perm[i] = i;
    }
    if (g.any(sing)) return LBCHEM_CELL_SINGULAR;
    g.sync();
    for (int j = 0; j < n; ++j) {
        unsigned long long bk = 0ull; int bi = -1;
        for (int i = g.tid; i < n; i += TPC) {
            if (i >= j) {
                const unsigned long long key = (unsigned long long)__double_as_longlong(w.VV[i] * fabs(A[i * lda + j]));
                if (key > bk || (key == bk && i > bi)) { bk = key; bi = i; }
            }
        }
        const int iMax = g.argmax(bk, bi);
        if (iMax != j) {
            for (int k = g.tid; k < n; k += TPC) {
                const double t = A[j * lda + k]; A[j * lda + k] = A[iMax * lda + k]; A[iMax * lda + k] = t;
            }
            if (g.tid == 0) {
                w.VV[iMax] = w.VV[j];
                const int t = w.perm[j]; w.perm[j] = w.perm[iMax]; w.perm[iMax] = t;
            }
        }
        g.sync();
        double piv = A[j * lda + j];
        if (piv == 0.0) piv = LB_SMALL;
        const double rd = 1.0 / piv;
        // zero pivot -> SMALL: late readers see 0 or SMALL and derive the same piv (benign race)
        if (g.tid == 0) { if (A[j * lda + j] == 0.0) A[j * lda + j] = piv; w.RD[j] = rd; }
        for (int i = g.tid; i < n; i += TPC) {
            if (i > j) {
                const double l = A[i * lda + j] * rd;
                A[i * lda + j] = l;
                double* Ai = A + i * lda;
                const double* Aj = A + j * lda;
#pragma unroll 4
                for (int k = j + 1; k < n; ++k) Ai[k] = fma(-l, Aj[k], Ai[k]);
            }
        }
        g.sync();
    }
    if (g.tid == 0) w.nLU++;
    return 0;
}

// b <- A^-1 b with the factors above. Warp 0 of the group holds two rows per lane and broadcasts
// finished unknowns by shuffle (n <= 64); larger systems use the shared-memory column sweep.
template <int TPC>
__device__ void lu_solve(Grp<TPC>& g, const DevMech& M, Work& w, const double* A, int n, double* b) {
    const int lda = M.lda;
    g.sync();
    if (TPC >= 32 && n <= 64) {
        if (g.tid < 32) {
            const int lane = g.tid, i0 = lane, i1 = lane + 32;
            double b0 = i0 < n ? b[w.perm[i0]] : 0.0;
            double b1 = i1 < n ? b[w.perm[i1]] : 0.0;
            const double* A0 = A + (i0 < n ? i0 : 0) * lda;
            const double* A1 = A + (i1 < n ? i1 : 0) * lda;
            for (int j = 0; j < n - 1; ++j) {
                const double bj = __shfl_sync(0xffffffffu, j < 32 ? b0 : b1, j & 31);
                if (i0 > j && i0 < n) b0 = fma(-A0[j], bj, b0);
                if (i1 > j && i1 < n) b1 = fma(-A1[j], bj, b1);
            }
            for (int j = n - 1; j >= 0; --j) {
                const double v = (j < 32 ? b0 : b1) * w.RD[j];
                const double xj = __shfl_sync(0xffffffffu, v, j & 31);
                if (lane == (j & 31)) { if (j < 32) b0 = xj; else b1 = xj; }
                if (i0 < j) b0 = fma(-A0[j], xj, b0);
                if (i1 < j) b1 = fma(-A1[j], xj, b1);
            }
            __syncwarp();
            if (i0 < n) b[i0] = b0;
            if (i1 < n) b[i1] = b1;
        }
    } else {
        double* t = w.VV;   // scaling factors are dead after the factorisation
        for (int i = g.tid; i < n; i += TPC) t[i] = b[w.perm[i]];
        g.sync();
        for (int i = g.tid; i < n; i += TPC) b[i] = t[i];
        for (int j = 0; j < n - 1; ++j) {
            g.sync();
            const double bj = b[j];
            for (int i = g.tid; i < n; i += TPC) if (i > j) b[i] = fma(-A[i * lda + j], bj, b[i]);
        }
        for (int j = n - 1; j >= 0; --j) {
            g.sync();
            const double xj = b[j] * w.RD[j];
            g.sync();
            if (g.tid == 0) b[j] = xj;
            for (int i = g.tid; i < n; i += TPC) if (i < j) b[i] = fma(-A[i * lda + j], xj, b[i]);
        }
    }
    if (g.tid == 0) w.nSolve++;
    g.sync();
}

// ------------------------------------------------------------------------------------------------
// ODESolver::normalizeError, Rosenbrock34::solve (A.4), adaptiveSolver::solve (A.3)
// ------------------------------------------------------------------------------------------------
struct Cell {          // uniform per-cell scalars (identical in every thread of the group)
    int fail;
    double theta;      // seulex
    int kTarg;         // seulex
};

// one Rosenbrock34 attempt from Y0 with step dx; result in YN; returns the normalised error
template <int TPC>
__device__ double rosenbrock34_step(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, Cell& cs, double dx) {
    const double a21 = 2, a31 = 48.0 / 25.0, a32 = 6.0 / 25.0, c21 = -8, c31 = 372.0 / 25.0, c32 = 12.0 / 5.0,
                 c41 = -112.0 / 125.0, c42 = -54.0 / 125.0, c43 = -2.0 / 5.0, b1 = 19.0 / 9.0, b2 = 1.0 / 2.0,
                 b3 = 25.0 / 108.0, b4 = 125.0 / 108.0, e1 = 34.0 / 108.0, e2 = 7.0 / 36.0, e3 = 0, e4 = 125.0 / 108.0,
                 gamma = 1.0 / 2.0, c1X = 1.0 / 2.0, c2X = -3.0 / 2.0, c3X = 121.0 / 50.0, c4X = 29.0 / 250.0;
    const int n = M.nEqn, lda = M.lda;
    double *Y0 = w.Y0, *Y = w.YN, *D0 = w.D0, *DX = w.DX, *K1 = w.K1, *K2 = w.K2, *K3 = w.K3, *K4 = w.K4, *DY = w.DY;
    jacobian(g, M, O, w, Y0, DX, w.A);
    {
        const double dg = 1.0 / (gamma * dx);
        for (int k = g.tid; k < n * lda; k += TPC) w.A[k] = -w.A[k];
        g.sync();
        for (int i = g.tid; i < n; i += TPC) w.A[i * lda + i] += dg;
    }
    const int st = lu_decompose(g, M, w, w.A, n);
    if (st) { cs.fail = st; return 0.0; }
    for (int i = g.tid; i < n; i += TPC) K1[i] = D0[i] + dx * c1X * DX[i];
    lu_solve(g, M, w, w.A, n, K1);
    for (int i = g.tid; i < n; i += TPC) Y[i] = Y0[i] + a21 * K1[i];
    derivatives(g, M, O, w, Y, DY);
    for (int i = g.tid; i < n; i += TPC) K2[i] = DY[i] + dx * c2X * DX[i] + c21 * K1[i] / dx;
    lu_solve(g, M, w, w.A, n, K2);
    for (int i = g.tid; i < n; i += TPC) Y[i] = Y0[i] + a31 * K1[i] + a32 * K2[i];
    derivatives(g, M, O, w, Y, DY);
    for (int i = g.tid; i < n; i += TPC) K3[i] = DY[i] + dx * c3X * DX[i] + (c31 * K1[i] + c32 * K2[i]) / dx;
    lu_solve(g, M, w, w.A, n, K3);
    for (int i = g.tid; i < n; i += TPC) K4[i] = DY[i] + dx * c4X * DX[i] + (c41 * K1[i] + c42 * K2[i] + c43 * K3[i]) / dx;
    lu_solve(g, M, w, w.A, n, K4);
    double em = 0.0;
    for (int i = g.tid; i < n; i += TPC) {
        const double yi = Y0[i] + b1 * K1[i] + b2 * K2[i] + b3 * K3[i] + b4 * K4[i];
        const double er = e1 * K1[i] + e2 * K2[i] + e3 * K3[i] + e4 * K4[i];
        Y[i] = yi;
        const double tol = O.absTol + O.relTol * fmax(fabs(Y0[i]), fabs(yi));
        const double q = fabs(er) / tol;
        em = (q != q || em != em) ? __longlong_as_double(0x7ff8000000000000LL) : fmax(em, q);
    }
    return g.maxnan(em);
}

// adaptiveSolver::solve: state Y0 at x advanced by one accepted step; dxTry updated
template <int TPC>
__device__ void adaptive_solve(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, Cell& cs, double& x, double& dxTry) {
    const double safeScale = 0.9, alphaInc = 0.2, alphaDec = 0.25, minScale = 0.2, maxScale = 10;
    const int n = M.nEqn;
    double dx = dxTry, err = 0.0;
    derivatives(g, M, O, w, w.Y0, w.D0);
    do {
        if (g.tid == 0) w.nAttempts++;
        err = rosenbrock34_step(g, M, O, w, cs, dx);
        if (cs.fail) return;
        if (!(err == err)) { cs.fail = LBCHEM_CELL_NAN; return; }
        if (err > 1) {
            const double scale = fmax(safeScale * pow(err, -alphaDec), minScale);
            dx *= scale;
            if (dx < LB_VSMALL) { cs.fail = LBCHEM_CELL_UNDERFLOW; return; }
        }
    } while (err > 1);
    x += dx;
    g.sync();
    for (int i = g.tid; i < n; i += TPC) w.Y0[i] = w.YN[i];
    g.sync();
    if (err > pow(maxScale / safeScale, -1.0 / alphaInc))
        dxTry = fmin(fmax(safeScale * pow(err, -alphaInc), minScale), maxScale) * dx;
    else
        dxTry = safeScale * maxScale * dx;
}

// ------------------------------------------------------------------------------------------------
// seulex (A.5): extrapolated linearly-implicit Euler, NR3 StepperSie control with OpenFOAM's stepState
// ------------------------------------------------------------------------------------------------
struct StepState { int forward, first, last, reject, prevReject; double dxTry, dxDid; };

__constant__ int c_nSeq[12] = {2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96};
// cpu[0] = cpuJac + cpuLU + nSeq[0]*(cpuFunc + cpuSolve); cpu[k+1] = cpu[k] + (nSeq[k+1]-1)*2 + 1
__constant__ double c_cpu[12] = {10, 15, 22, 33, 48, 71, 102, 149, 212, 307, 434, 625};

template <int TPC>
__device__ bool seul(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, Cell& cs, const double* y0, double dxTot,
                     int k, double* yOut, const double* scale) {
    const int n = M.nEqn, lda = M.lda;
    const int nSteps = c_nSeq[k];
    const double dx = dxTot / nSteps;
    double *dy = w.K4, *yT = w.YN, *dydx = w.DY;
    g.sync();
    for (int q = g.tid; q < n * lda; q += TPC) w.A2[q] = -w.A[q];
    g.sync();
    for (int i = g.tid; i < n; i += TPC) w.A2[i * lda + i] += 1 / dx;
    const int st = lu_decompose(g, M, w, w.A2, n);
    if (st) { cs.fail = st; return false; }
    derivatives(g, M, O, w, y0, dy);
    lu_solve(g, M, w, w.A2, n, dy);
    for (int i = g.tid; i < n; i += TPC) yT[i] = y0[i];
    for (int nn = 1; nn < nSteps; ++nn) {
        for (int i = g.tid; i < n; i += TPC) yT[i] += dy[i];
        if (nn == 1 && k <= 1) {
            double s1 = 0.0;
            for (int i = g.tid; i < n; i += TPC) { const double q = dy[i] / scale[i]; s1 = fma(q, q, s1); }
            const double dy1 = sqrt(g.sum(s1));
            derivatives(g, M, O, w, yT, dydx);
            for (int i = g.tid; i < n; i += TPC) dy[i] = dydx[i] - dy[i] / dx;
            lu_solve(g, M, w, w.A2, n, dy);
            const double denom = fmin(1.0, dy1 + LB_SMALL);
            double s2 = 0.0; int big = 0;
            for (int i = g.tid; i < n; i += TPC) {
                if (fabs(dy[i]) > scale[i] * denom) big = 1;
                const double q = dy[i] / scale[i];
                s2 = fma(q, q, s2);
            }
            if (g.any(big)) { cs.theta = 1; return false; }
            const double dy2 = sqrt(g.sum(s2));
            cs.theta = dy2 / denom;
            if (cs.theta > 1) return false;
        }
        derivatives(g, M, O, w, yT, dy);
        lu_solve(g, M, w, w.A2, n, dy);
    }
    for (int i = g.tid; i < n; i += TPC) yOut[i] = yT[i] + dy[i];
    g.sync();
    return true;
}

__device__ __forceinline__ double seulex_coeff(int k, int l) {
    const double ratio = (double)c_nSeq[k] / c_nSeq[l];
    return 1 / (ratio - 1);
}

// seulex::solve(x, y, step); y is w.Y0
template <int TPC>
__device__ void seulex_solve(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, Cell& cs, double& x, StepState& step,
                             int& attempts) {
    const int n = M.nEqn;
    const int kMaxx = 11;
    const double stepFactor1 = 0.6, stepFactor2 = 0.93, stepFactor3 = 0.1, stepFactor4 = 4, stepFactor5 = 0.5,
                 kFactor1 = 0.7, kFactor2 = 0.9;
    const double absTol = O.absTol, relTol = O.relTol;
    const double jacRedo = fmin(1e-4, relTol);
    double *y = w.Y0, *y0 = w.K1, *scale = w.K2, *ySeq = w.K3, *tb = w.TB;
    double temp[12], dxOpt[12];
    temp[0] = LB_GREAT;
    double dx = step.dxTry;
    g.sync();
    for (int i = g.tid; i < n; i += TPC) { y0[i] = y[i]; scale[i] = absTol + relTol * fabs(y[i]); }
    dxOpt[0] = fabs(0.1 * dx);
    if (step.first || step.prevReject) cs.theta = 2.0 * jacRedo;
    if (step.first) {
        const double logTol = -log10(relTol + absTol) * 0.6 + 0.5;
        int kt = (int)logTol;
        if (kt > kMaxx - 1) kt = kMaxx - 1;
        if (kt < 1) kt = 1;
        cs.kTarg = kt;
    }
    bool jacUpdated = false;
    if (cs.theta > jacRedo) { jacobian(g, M, O, w, y, w.DX, w.A); jacUpdated = true; }
    int k = 0;
    double dxNew = fabs(dx);
    bool firstk = true;
    while (firstk || step.reject) {
        dx = step.forward ? dxNew : -dxNew;
        firstk = false;
        step.reject = 0;
        attempts++;
        if (g.tid == 0) w.nAttempts++;
        if (fabs(dx) < LB_VSMALL || attempts > 100000) { cs.fail = LBCHEM_CELL_UNDERFLOW; return; }
        double errOld = 0;
        for (k = 0; k <= cs.kTarg + 1; ++k) {
            const bool success = seul(g, M, O, w, cs, y0, dx, k, ySeq, scale);
            if (cs.fail) return;
            if (!success) { step.reject = 1; dxNew = fabs(dx) * stepFactor5; break; }
            if (k == 0) { for (int i = g.tid; i < n; i += TPC) y[i] = ySeq[i]; }
            else { for (int i = g.tid; i < n; i += TPC) tb[(k - 1) * n + i] = ySeq[i]; }
            if (k != 0) {
                // extrapolate(k, table, y): each thread owns its components, no cross-thread traffic
                double s = 0.0;
                for (int i = g.tid; i < n; i += TPC) {
                    for (int j = k - 1; j > 0; --j)
                        tb[(j - 1) * n + i] = tb[j * n + i] + seulex_coeff(k, j) * (tb[j * n + i] - tb[(j - 1) * n + i]);
                    const double yi = tb[i] + seulex_coeff(k, 0) * (tb[i] - y[i]);
                    y[i] = yi;
                    const double sc = absTol + relTol * fabs(y0[i]);
                    scale[i] = sc;
                    const double q = (yi - tb[i]) / sc;
                    s = fma(q, q, s);
                }
                const double err = sqrt(g.sum(s) / n);
                if (!(err == err)) { cs.fail = LBCHEM_CELL_NAN; return; }
                if (err > 1 / LB_SMALL || (k > 1 && err >= errOld)) { step.reject = 1; dxNew = fabs(dx) * stepFactor5; break; }
                errOld = O.seulexErrOldMax ? fmax(4 * err, 1.0) : fmin(4 * err, 1.0);
                const double expo = 1.0 / (k + 1);
                const double facmin = pow(stepFactor3, expo);
                double fac;
                if (err == 0) fac = 1 / facmin;
                else {
                    fac = stepFactor2 / pow(err / stepFactor1, expo);
                    fac = fmax(facmin / stepFactor4, fmin(1 / facmin, fac));
                }
                dxOpt[k] = fabs(dx * fac);
                temp[k] = c_cpu[k] / dxOpt[k];
                if ((step.first || step.last) && err <= 1) break;
                if (k == cs.kTarg - 1 && !step.prevReject && !step.first && !step.last) {
                    if (err <= 1) break;
                    else if (err > c_nSeq[cs.kTarg] * c_nSeq[cs.kTarg + 1] * 4) {
                        step.reject = 1;
                        cs.kTarg = k;
                        if (cs.kTarg > 1 && temp[k - 1] < kFactor1 * temp[k]) cs.kTarg--;
                        dxNew = dxOpt[cs.kTarg];
                        break;
                    }
                }
                if (k == cs.kTarg) {
                    if (err <= 1) break;
                    else if (err > c_nSeq[k + 1] * 2) {
                        step.reject = 1;
                        if (cs.kTarg > 1 && temp[k - 1] < kFactor1 * temp[k]) cs.kTarg--;
                        dxNew = dxOpt[cs.kTarg];
                        break;
                    }
                }
                if (k == cs.kTarg + 1) {
                    if (err > 1) {
                        step.reject = 1;
                        if (cs.kTarg > 1 && temp[cs.kTarg - 1] < kFactor1 * temp[cs.kTarg]) cs.kTarg--;
                        dxNew = dxOpt[cs.kTarg];
                    }
                    break;
                }
            }
        }
        if (step.reject) {
            step.prevReject = 1;
            if (!jacUpdated) {
                cs.theta = 2.0 * jacRedo;
                if (cs.theta > jacRedo && !jacUpdated) { jacobian(g, M, O, w, y, w.DX, w.A); jacUpdated = true; }
            }
        }
    }
    step.dxDid = dx;
    x += dx;
    int kopt;
    if (k == 1) kopt = 2;
    else if (k <= cs.kTarg) {
        kopt = k;
        if (temp[k - 1] < kFactor1 * temp[k]) kopt = k - 1;
        else if (temp[k] < kFactor2 * temp[k - 1]) kopt = (k + 1 < kMaxx - 1) ? k + 1 : kMaxx - 1;
    } else {
        kopt = k - 1;
        if (k > 2 && temp[k - 2] < kFactor1 * temp[k - 1]) kopt = k - 2;
        if (temp[k] < kFactor2 * temp[kopt]) kopt = (k < kMaxx - 1) ? k : kMaxx - 1;
    }
    if (step.prevReject) {
        cs.kTarg = (kopt < k) ? kopt : k;
        dxNew = fmin(fabs(dx), dxOpt[cs.kTarg]);
        step.prevReject = 0;
    } else {
        if (kopt <= k) dxNew = dxOpt[kopt];
        else {
            if (k < cs.kTarg && temp[k] < kFactor2 * temp[k - 1]) dxNew = dxOpt[k] * c_cpu[kopt + 1] / c_cpu[k];
            else dxNew = dxOpt[k] * c_cpu[kopt] / c_cpu[k];
        }
        cs.kTarg = kopt;
    }
    step.dxTry = step.forward ? dxNew : -dxNew;
}

// ------------------------------------------------------------------------------------------------
// EulerImplicit::solve (A.11). c lives in Y0[0..nSp), T in Y0[nSp]; matrix in w.A (nSp x nSp).
// ------------------------------------------------------------------------------------------------
template <int TPC>
__device__ double mix_sum(Grp<TPC>& g, const DevMech& M, const DevOpts& O, const double* c, double T, bool wantCp) {
    const int nSp = M.nSp;
    const bool lo = T < M.Tc[0];   // janafThermo::operator+= keeps the first species' Tcommon
    double s = 0.0;
    for (int i = g.tid; i < nSp; i += TPC)
        s = fma(c[i], wantCp ? janaf_cp((lo ? M.cpLo : M.cpHi) + i, nSp, T) : janaf_ha((lo ? M.haLo : M.haHi) + i, nSp, T), s);
    return g.sum(s) * O.RR;
}

template <int TPC>
__device__ void euler_implicit_solve(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, Cell& cs, double& deltaT,
                                     double& subDeltaT) {
    const int nSp = M.nSp, nR = M.nR, lda = M.lda;
    double* c = w.Y0;
    double* RRm = w.A;
    double* src = w.K1;
    if (g.tid == 0) { w.nAttempts++; w.nSteps++; w.nRhs++; }
    g.sync();
    const double T = c[nSp];
    double ct = 0.0;
    for (int i = g.tid; i < nSp; i += TPC) { const double v = fmax(0.0, c[i]); c[i] = v; w.CC[i] = v; ct += v; }
    const double cTot = g.sum(ct);
    g.sync();
    const double ha = mix_sum(g, M, O, c, T, false);
    const double deltaTEst = fmin(deltaT, subDeltaT);
    for (int q = g.tid; q < nSp * lda; q += TPC) RRm[q] = 0.0;
    species_gibbs(g, M, w, T, log(T));
    g.sync();
    third_body(g, M, w);
    g.sync();
    {
        const double lnT = log(T), invT = 1.0 / T;
        double* PF = w.R2; double* PR = w.R2 + nR; int* REF = reinterpret_cast<int*>(w.R2 + 2 * nR);
        for (int r = g.tid; r < nR; r += TPC) {
            double kf, kr, pf, cf, pr, cr; int lRef, rRef;
            rate_constants(M, O, w, r, T, lnT, invT, kf, kr);
            const double om = omega_faithful(M, w.CC, r, kf, kr, pf, cf, lRef, pr, cr, rRef);
            double corr = 1.0;
            if (O.eqRateLimiter) corr = (om < 0) ? 1 / (1 + pr * deltaTEst) : 1 / (1 + pf * deltaTEst);
            PF[r] = pf * corr; PR[r] = pr * corr; REF[2 * r] = lRef; REF[2 * r + 1] = rRef;
        }
        g.sync();
        for (int i = g.tid; i < nSp; i += TPC) {
            double* Ri = RRm + i * lda;
            const int e = M.spPtr[i + 1];
            for (int q = M.spPtr[i]; q < e; ++q) {
                const int r = M.spReac[q];
                const double cf = M.spCoef[q];      // -nu on the lhs, +nu on the rhs
                Ri[REF[2 * r + 1]] += cf * PR[r];
                Ri[REF[2 * r]] -= cf * PF[r];
            }
        }
    }
    g.sync();
    double tMin = LB_GREAT;
    for (int i = g.tid; i < nSp; i += TPC) {
        double d = 0.0;
        for (int j = 0; j < nSp; ++j) d = fma(-RRm[i * lda + j], c[j], d);
        if (d < -LB_SMALL) tMin = fmin(tMin, -(c[i] + LB_SMALL) / d);
        else {
            d = fmax(d, LB_SMALL);
            tMin = fmin(tMin, fmax(cTot - c[i], 1e-5) / d);
        }
    }
    tMin = -g.max(-tMin);
    subDeltaT = O.cTauChem * tMin;
    deltaT = fmin(deltaT, subDeltaT);
    for (int i = g.tid; i < nSp; i += TPC) { RRm[i * lda + i] += 1 / deltaT; src[i] = c[i] / deltaT; }
    const int st = lu_decompose(g, M, w, RRm, nSp);
    if (st) { cs.fail = st; return; }
    lu_solve(g, M, w, RRm, nSp, src);
    for (int i = g.tid; i < nSp; i += TPC) c[i] = fmax(0.0, src[i]);
    g.sync();
    // species::thermo::THa -> Newton on Ha with Cp, tol 1e-4*T0, at most 100 iterations
    double Test = T, Tnew = T;
    const double Ttol = T * 1e-4;
    int iter = 0;
    do {
        Test = Tnew;
        const double h = mix_sum(g, M, O, c, Test, false);
        const double cp = mix_sum(g, M, O, c, Test, true);
        Tnew = Test - (h - ha) / cp;
        if (Tnew < M.TlowMix || Tnew > M.ThighMix) Tnew = fmin(fmax(Tnew, M.TlowMix), M.ThighMix);
        if (iter++ > 100) { cs.fail = LBCHEM_CELL_MAX_STEPS; break; }
    } while (fabs(Tnew - Test) > Ttol);
    g.sync();
    if (g.tid == 0) c[nSp] = Tnew;
    g.sync();
}

// ------------------------------------------------------------------------------------------------
// ODESolver::solve(xStart, xEnd, y, dxTry) (A.2) and ode<>::solve (A.1); y is w.Y0
// ------------------------------------------------------------------------------------------------
template <int TPC>
__device__ void ode_solve(Grp<TPC>& g, const DevMech& M, const DevOpts& O, Work& w, Cell& cs, double xStart, double xEnd,
                          double& dxTry) {
    StepState step = {1, 1, 0, 0, 0, dxTry, 0.0};
    double x = xStart;
    int attempts = 0;
    for (int nStep = 0; nStep < O.maxSteps; ++nStep) {
        const double dxTry0 = step.dxTry;
        step.reject = 0;
        if ((x + step.dxTry - xEnd) * (x + step.dxTry - xStart) > 0) {
            step.last = 1;
            step.dxTry = xEnd - x;
        }
        if (O.solver == LBCHEM_SOLVER_SEULEX) {
            seulex_solve(g, M, O, w, cs, x, step, attempts);
        } else {
            const double x0 = x;
            adaptive_solve(g, M, O, w, cs, x, step.dxTry);
            step.dxDid = x - x0;
        }
        if (cs.fail) return;
        if (g.tid == 0) w.nSteps++;
        if ((x - xEnd) * (xEnd - xStart) >= 0) {
            if (nStep > 0 && step.last) step.dxTry = dxTry0;
            dxTry = step.dxTry;
            return;
        }
        step.first = 0;
        if (step.reject) step.prevReject = 1;
    }
    cs.fail = LBCHEM_CELL_MAX_STEPS;
}

// ------------------------------------------------------------------------------------------------
// workspace carving, solveCell (reference LoadBalancedChemistryModel.C:365-420), kernels
// ------------------------------------------------------------------------------------------------
template <int TPC>
__device__ __forceinline__ void carve(Work& w, Grp<TPC>& g, double* base, const DevMech& M, int solver) {
    const int n = M.nEqn;
    size_t r2 = (size_t)M.nTerms + M.nR;
    if (r2 < 2 * (size_t)M.nR) r2 = 2 * (size_t)M.nR;
    double* p = base;
    g.red = p; p += 8;
    w.Y0 = p; p += n; w.YN = p; p += n; w.D0 = p; p += n; w.DX = p; p += n;
    w.K1 = p; p += n; w.K2 = p; p += n; w.K3 = p; p += n; w.K4 = p; p += n; w.DY = p; p += n;
    w.CC = p; p += M.nSp; w.GL = p; p += M.nSp; w.GH = p; p += M.nSp;
    w.ME = p; p += M.nEff;
    w.VV = p; p += n; w.RD = p; p += n;
    w.R2 = p; p += r2;
    w.A = p; p += (size_t)n * M.lda;
    w.A2 = nullptr; w.TB = nullptr;
    if (solver == LBCHEM_SOLVER_SEULEX) { w.A2 = p; p += (size_t)n * M.lda; w.TB = p; p += 12 * n; }
    w.perm = reinterpret_cast<int*>(p);
    w.nSteps = w.nAttempts = w.nJac = w.nLU = w.nSolve = w.nRhs = 0;
}

template <int TPC>
__global__ void __launch_bounds__(TPC * (TPC < 32 ? 16 : 8)) integrate_kernel(DevMech M, DevOpts O, BatchArgs B, int groups, int wsDoubles) {
    extern __shared__ double smem[];
    Grp<TPC> g;
    g.tid = threadIdx.x % TPC;
    g.gid = threadIdx.x / TPC;
    g.tog = 0;
    g.mask = TPC >= 32 ? 0xffffffffu : (((1u << (TPC & 31)) - 1u) << ((threadIdx.x & 31) & ~(TPC - 1)));
    Work w;
    carve(w, g, smem + (size_t)g.gid * wsDoubles, M, O.solver);
    const int nSp = M.nSp;
    unsigned long long st_steps = 0, st_att = 0, st_jac = 0, st_lu = 0, st_solve = 0, st_rhs = 0, st_fail = 0, st_cells = 0;

    for (;;) {
        unsigned long long q = 0;
        if (g.tid == 0) q = atomicAdd(B.queue, 1ull);
        q = __shfl_sync(g.mask, q, TPC >= 32 ? 0 : ((threadIdx.x & 31) & ~(TPC - 1)));
        if (TPC > 32) {
            // broadcast across the group's warps through shared memory
            g.sync();
            if (g.tid == 0) reinterpret_cast<unsigned long long*>(g.red)[7] = q;
            g.sync();
            q = reinterpret_cast<unsigned long long*>(g.red)[7];
        }
        if (q >= (unsigned long long)B.a.n) break;
        long long cell = B.order ? (long long)B.order[q] : (long long)q;
        if (B.a.index) cell = B.a.index[cell];
        const double T0 = B.a.T[cell];
        const double p0 = B.a.p[cell];
        const double rho = B.a.rho[cell];
        const double deltaT = B.a.deltaT ? B.a.deltaT[cell] : B.deltaTUniform;
        double dtChem = B.a.deltaTChem[cell];
        w.nSteps = w.nAttempts = w.nJac = w.nLU = w.nSolve = w.nRhs = 0;
        Cell cs; cs.fail = 0; cs.theta = 0.0; cs.kTarg = 1;
        int status;
        double Tend = T0;
        if (T0 > O.Treact) {
            g.sync();
            for (int i = g.tid; i < nSp; i += TPC) w.Y0[i] = rho * B.a.Y[(size_t)i * B.a.ldy + cell] / M.W[i];
            if (g.tid == 0) { w.Y0[nSp] = T0; w.Y0[nSp + 1] = p0; }
            g.sync();
            double timeLeft = deltaT;
            int guard = 0;
            while (timeLeft > LB_SMALL) {
                double dt = timeLeft;
                if (O.solver == LBCHEM_SOLVER_EULER_IMPLICIT) {
                    euler_implicit_solve(g, M, O, w, cs, dt, dtChem);
                } else if (O.solver == LBCHEM_SOLVER_NONE) {
                    // noChemistrySolver::solve does nothing
                } else {
                    // ode<>::solve: integrate cTp over [0, dt], then clip c >= 0
                    ode_solve(g, M, O, w, cs, 0.0, dt, dtChem);
                    g.sync();
                    for (int i = g.tid; i < nSp; i += TPC) w.Y0[i] = fmax(0.0, w.Y0[i]);
                    g.sync();
                }
                if (cs.fail) break;
                timeLeft -= dt;
                if (++guard > 10000000) { cs.fail = LBCHEM_CELL_MAX_STEPS; break; }
            }
            g.sync();
            dtChem = fmin(dtChem, O.deltaTChemMax);
            for (int i = g.tid; i < nSp; i += TPC) {
                const double c0 = rho * B.a.Y[(size_t)i * B.a.ldy + cell] / M.W[i];
                B.a.RR[(size_t)i * B.a.ldy + cell] = (w.Y0[i] - c0) * M.W[i] / deltaT;
            }
            Tend = w.Y0[nSp];
            status = cs.fail ? cs.fail : LBCHEM_CELL_OK;
        } else {
            for (int i = g.tid; i < nSp; i += TPC) B.a.RR[(size_t)i * B.a.ldy + cell] = 0.0;
            status = LBCHEM_CELL_SKIPPED;
        }
        if (g.tid == 0) {
            B.a.deltaTChem[cell] = dtChem;
            if (B.a.Tout) B.a.Tout[cell] = Tend;
            if (B.a.cost) B.a.cost[cell] = (double)w.nAttempts;
            if (B.a.status) B.a.status[cell] = status;
            st_steps += w.nSteps; st_att += w.nAttempts; st_jac += w.nJac; st_lu += w.nLU;
            st_solve += w.nSolve; st_rhs += w.nRhs; st_fail += (status >= LBCHEM_CELL_MAX_STEPS); st_cells += 1;
        }
    }
    if (g.tid == 0 && st_cells) {
        atomicAdd(B.stats + 0, st_steps); atomicAdd(B.stats + 1, st_att); atomicAdd(B.stats + 2, st_jac);
        atomicAdd(B.stats + 3, st_lu); atomicAdd(B.stats + 4, st_solve); atomicAdd(B.stats + 5, st_rhs);
        atomicAdd(B.stats + 6, st_fail); atomicAdd(B.stats + 7, st_cells);
    }
}

// Probe kernel: one group per item evaluates derivatives / jacobian / LU-solve / rate constants with the
// same device functions the integrator uses (test hook behind lbchem_probe).
//   DERIV   in: y[nEqn]                 out: dydt[nEqn]
//   JAC     in: y[nEqn]                 out: dfdx[nEqn] then J[nEqn*nEqn] row-major
//   LUSOLVE in: A[nEqn*nEqn], b[nEqn]   out: x[nEqn]
//   RATES   in: y[nEqn]                 out: kf[nR], kr[nR]
template <int TPC>
__global__ void __launch_bounds__(TPC * (TPC < 32 ? 16 : 8)) probe_kernel(DevMech M, DevOpts O, int kind, int nItems, const double* in, double* out,
                                                         int inStride, int outStride, int wsDoubles) {
    extern __shared__ double smem[];
    Grp<TPC> g;
    g.tid = threadIdx.x % TPC; g.gid = threadIdx.x / TPC; g.tog = 0;
    Work w;
    carve(w, g, smem + (size_t)g.gid * wsDoubles, M, O.solver);
    const int item = blockIdx.x * (blockDim.x / TPC) + g.gid;
    if (item >= nItems) return;
    const int n = M.nEqn, lda = M.lda;
    const double* x = in + (size_t)item * inStride;
    double* o = out + (size_t)item * outStride;
    if (kind == LB_PROBE_LUSOLVE) {
        for (int q = g.tid; q < n * n; q += TPC) w.A[(q / n) * lda + (q % n)] = x[q];
        for (int i = g.tid; i < n; i += TPC) w.K1[i] = x[n * n + i];
        g.sync();
        const int st = lu_decompose(g, M, w, w.A, n);
        if (!st) lu_solve(g, M, w, w.A, n, w.K1);
        for (int i = g.tid; i < n; i += TPC) o[i] = st ? __longlong_as_double(0x7ff8000000000000LL) : w.K1[i];
        return;
    }
    for (int i = g.tid; i < n; i += TPC) w.Y0[i] = x[i];
    g.sync();
    if (kind == LB_PROBE_DERIV) {
        derivatives(g, M, O, w, w.Y0, w.DY);
        for (int i = g.tid; i < n; i += TPC) o[i] = w.DY[i];
    } else if (kind == LB_PROBE_JAC) {
        jacobian(g, M, O, w, w.Y0, w.DX, w.A);
        for (int i = g.tid; i < n; i += TPC) o[i] = w.DX[i];
        for (int q = g.tid; q < n * n; q += TPC) o[n + q] = w.A[(q / n) * lda + (q % n)];
    } else if (kind == LB_PROBE_RATES) {
        const double T = w.Y0[M.nSp];
        prepare_state(g, M, w, w.Y0, T);
        const double lnT = log(T), invT = 1.0 / T;
        for (int r = g.tid; r < M.nR; r += TPC) {
            double kf, kr;
            rate_constants(M, O, w, r, T, lnT, invT, kf, kr);
            o[r] = kf; o[M.nR + r] = kr;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// launchers (called from lbchem_api.cu)
// ------------------------------------------------------------------------------------------------
size_t lb_work_doubles(const DevMech& M, int solver) {
    size_t d = work_doubles(M, solver);
    return (d + 1) & ~(size_t)1;   // keep every group's slice 16-byte aligned
}

template <int TPC>
static cudaError_t configure(int groups, size_t smemBytes, int* ctasPerSM) {
    cudaError_t e = cudaFuncSetAttribute(integrate_kernel<TPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(probe_kernel<TPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctasPerSM, integrate_kernel<TPC>, groups * TPC, smemBytes);
}

cudaError_t lb_configure(int tpc, int groups, size_t smemBytes, int* ctasPerSM) {
    switch (tpc) {
        case 8: return configure<8>(groups, smemBytes, ctasPerSM);
        case 16: return configure<16>(groups, smemBytes, ctasPerSM);
        case 32: return configure<32>(groups, smemBytes, ctasPerSM);
        default: return configure<64>(groups, smemBytes, ctasPerSM);
    }
}

cudaError_t lb_launch_integrate(int tpc, int grid, int groups, size_t smemBytes, int wsDoubles, const DevMech& M,
                                const DevOpts& O, const BatchArgs& B, cudaStream_t s) {
    switch (tpc) {
        case 8: integrate_kernel<8><<<grid, groups * 8, smemBytes, s>>>(M, O, B, groups, wsDoubles); break;
        case 16: integrate_kernel<16><<<grid, groups * 16, smemBytes, s>>>(M, O, B, groups, wsDoubles); break;
        case 32: integrate_kernel<32><<<grid, groups * 32, smemBytes, s>>>(M, O, B, groups, wsDoubles); break;
        default: integrate_kernel<64><<<grid, groups * 64, smemBytes, s>>>(M, O, B, groups, wsDoubles);
    }
    return cudaGetLastError();
}

cudaError_t lb_launch_probe(int tpc, int groups, size_t smemBytes, int wsDoubles, const DevMech& M, const DevOpts& O,
                            int kind, int nItems, const double* in, double* out, int inStride, int outStride, cudaStream_t s) {
    const int grid = (nItems + groups - 1) / groups;
    switch (tpc) {
        case 8: probe_kernel<8><<<grid, groups * 8, smemBytes, s>>>(M, O, kind, nItems, in, out, inStride, outStride, wsDoubles); break;
        case 16: probe_kernel<16><<<grid, groups * 16, smemBytes, s>>>(M, O, kind, nItems, in, out, inStride, outStride, wsDoubles); break;
        case 32: probe_kernel<32><<<grid, groups * 32, smemBytes, s>>>(M, O, kind, nItems, in, out, inStride, outStride, wsDoubles); break;
        default: probe_kernel<64><<<grid, groups * 64, smemBytes, s>>>(M, O, kind, nItems, in, out, inStride, outStride, wsDoubles);
    }
    return cudaGetLastError();
}
