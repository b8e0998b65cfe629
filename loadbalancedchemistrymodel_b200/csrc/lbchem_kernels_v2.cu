// lbchem_kernels_v2.cu — fast path of the per-cell chemistry integration for sm_100a.
//
// 64 threads (2 warps) integrate one cell; a persistent CTA holds up to 6 cells and stages the packed
// mechanism into shared memory once. Differences from the general kernel (lbchem_kernels.cu):
//   * the dense (I/(g dx) - J) matrix is factorised in REGISTERS: thread t owns row t (N doubles), partial
//     pivoting is implicit (rows never move; the reference's pivot order incl. its tie rule is tracked through
//     per-row positions), the pivot row is broadcast through shared memory once per column;
//   * the Jacobian is kept SPARSE in shared memory (its structural non-zeros + the dense temperature column),
//     so seulex needs no second dense matrix and Rosenbrock34 re-uses it after a rejected step;
//   * triangular solves run in column blocks of 8: one barrier per block instead of one per column;
//   * rate constants: reactions sorted into homogeneous rounds (exp / no exp / third body / fall-off).
// Same arithmetic as the general kernel where order matters (LU elimination order, forward substitution,
// Jacobian and omega accumulation order); see DESIGN.md for the places where rounding differs.
#include <cuda_runtime.h>
#include <math.h>
#include "lbchem_device.cuh"
#include "lbchem_v2.cuh"
#include "../../include/lbchem.h"

namespace v2 {

struct SMech {            // typed views into the staged blob
    int nSp, nR, nEff;
    const double *W, *Tc, *cpLo, *cpHi, *haLo, *haHi, *sLo, *sHi;
    const double *A, *beta, *Ta, *fall, *tcTab, *effVal;
    const unsigned* info; const unsigned long long *lpk, *rpk;
    const int *effPtr, *effSp, *spPtr, *rowPtr;
    const unsigned* spList; const unsigned char* entCol;
};

__device__ __forceinline__ void bind(SMech& S, const V2Mech& M, const unsigned char* b) {
    S.nSp = M.nSp; S.nR = M.nR; S.nEff = M.nEff;
#define P(T, off) reinterpret_cast<const T*>(b + M.off)
    S.W = P(double, oW); S.Tc = P(double, oTc); S.cpLo = P(double, oCpLo); S.cpHi = P(double, oCpHi);
    S.haLo = P(double, oHaLo); S.haHi = P(double, oHaHi); S.sLo = P(double, oSLo); S.sHi = P(double, oSHi);
    S.A = P(double, oA); S.beta = P(double, oBeta); S.Ta = P(double, oTa); S.fall = P(double, oFall);
    S.tcTab = P(double, oTcTab); S.effVal = P(double, oEffVal);
    S.info = P(unsigned, oInfo); S.lpk = P(unsigned long long, oLpk); S.rpk = P(unsigned long long, oRpk);
    S.effPtr = P(int, oEffPtr); S.effSp = P(int, oEffSp); S.spPtr = P(int, oSpPtr); S.rowPtr = P(int, oRowPtr);
    S.spList = P(unsigned, oSpList); S.entCol = P(unsigned char, oEntCol);
#undef P
}

struct Grp {
    int tid, gid, lane, warp;
    double* red;      // 16 doubles of exchange scratch
    int tog;
    __device__ __forceinline__ void sync() const { asm volatile("bar.sync %0, %1;" ::"r"(gid + 1), "r"(V2_TPC) : "memory"); }
    // three sums at once (one barrier)
    __device__ __forceinline__ void sum3(double& a, double& b, double& c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); c += __shfl_xor_sync(0xffffffffu, c, o);
        }
        double* r = red + 8 * tog; tog ^= 1;
        if (lane == 0) { r[warp * 3] = a; r[warp * 3 + 1] = b; r[warp * 3 + 2] = c; }
        sync();
        a = r[0] + r[3]; b = r[1] + r[4]; c = r[2] + r[5];
    }
    __device__ __forceinline__ double sum(double a) { double b = 0, c = 0; sum3(a, b, c); return a; }
    // max with NaN propagation (b carries the NaN flag)
    __device__ __forceinline__ double maxnan(double v) {
        double bad = (v != v) ? 1.0 : 0.0;
        v = (v != v) ? 0.0 : v;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o)); bad = fmax(bad, __shfl_xor_sync(0xffffffffu, bad, o));
        }
        double* r = red + 8 * tog; tog ^= 1;
        if (lane == 0) { r[warp * 2] = v; r[warp * 2 + 1] = bad; }
        sync();
        v = fmax(r[0], r[2]); bad = fmax(r[1], r[3]);
        return bad > 0.0 ? __longlong_as_double(0x7ff8000000000000LL) : v;
    }
    __device__ __forceinline__ int any(int p) {
        const int a = __any_sync(0xffffffffu, p);
        double* r = red + 8 * tog; tog ^= 1;
        if (lane == 0) r[warp] = a ? 1.0 : 0.0;
        sync();
        return (r[0] + r[1]) > 0.0;
    }
};

template <int N>
struct Work {
    double* V[10];      // N each: general vectors
    double* TB;         // tbRows x N
    double *CC, *GL, *GH, *ME, *OM, *JS, *JT, *RD, *BSH, *XS;
    double* AUX;        // LU aux: LB, UB, pivot-row buffers; KF/KR while the Jacobian is assembled
    double* XCH;        // 16 doubles: pivot exchange slots
    double *LB, *UB, *ROWB, *KF, *KR;
    int nSteps, nAttempts, nJac, nLU, nSolve, nRhs;
};

template <int N>
__device__ __forceinline__ void carve(Work<N>& w, Grp& g, double* base, const V2Mech& M) {
    double* p = base;
    g.red = p; p += 16;
    for (int i = 0; i < 10; ++i) { w.V[i] = p; p += N; }
    w.TB = p; p += M.tbRows * N;
    w.CC = p; p += M.nSp + 1; w.GL = p; p += M.nSp; w.GH = p; p += M.nSp; w.ME = p; p += M.nEff;
    w.OM = p; p += M.nR; w.JS = p; p += M.nEnt; w.JT = p; p += M.nSp;
    w.RD = p; p += N; w.BSH = p; p += N; w.XS = p; p += 16;
    w.AUX = p;
    constexpr int nBlk = (N + V2_BLK - 1) / V2_BLK;
    w.LB = w.AUX; w.UB = w.AUX + nBlk * V2_LSTRIDE; w.ROWB = w.UB + nBlk * V2_USTRIDE;
    w.KF = w.AUX; w.KR = w.AUX + M.nR;
    w.XCH = base + M.wsDoubles - 16;
    w.nSteps = w.nAttempts = w.nJac = w.nLU = w.nSolve = w.nRhs = 0;
}

// ---- janaf polynomials (Horner forms identical to the oracle) -----------------------------------
__device__ __forceinline__ double jcp(const double* a, int st, double T) {
    return (((a[4 * st] * T + a[3 * st]) * T + a[2 * st]) * T + a[1 * st]) * T + a[0];
}
__device__ __forceinline__ double jha(const double* q, int st, double T) {
    return ((((q[0] * T + q[1 * st]) * T + q[2 * st]) * T + q[3 * st]) * T + q[4 * st]) * T + q[5 * st];
}
__device__ __forceinline__ double js(const double* q, int st, double T, double lnT) {
    return (((q[0] * T + q[1 * st]) * T + q[2 * st]) * T + q[3 * st]) * T + q[4 * st] * lnT + q[5 * st];
}

__device__ __forceinline__ double cpow(double c, unsigned e) {   // e in 1..3
    return e == 1 ? c : (e == 2 ? c * c : c * c * c);
}

// product of clamped concentrations of one reaction side
__device__ __forceinline__ double side_product(const double* c, unsigned long long pk, unsigned n) {
    double f = 1.0;
    for (unsigned s = 0; s < n; ++s) { const unsigned sl = V2_SLOT(pk, s); f *= cpow(c[V2_SP(sl)], V2_EX(sl)); }
    return f;
}

// forward / reverse rate constants of the reaction at position p (A.6); GL/GH/ME must be current
__device__ __forceinline__ void rate_constants(const SMech& S, const DevOpts& O, const double* GL, const double* GH,
                                               const double* ME, int p, double T, double lnT, double invT, double& kf, double& kr) {
    const unsigned w = S.info[p];
    double k = S.A[p];
    if (V2_HASEXP(w)) k *= exp(fma(S.beta[p], lnT, -S.Ta[p] * invT));
    const unsigned ty = V2_RTYPE(w);
    if (ty == LBCHEM_R_THIRD_BODY) {
        k = ME[V2_EFF(w)] * k;
    } else if (ty >= LBCHEM_R_FALLOFF_LINDEMANN) {
        const double* f = S.fall + 8 * V2_FALL(w);
        double k0 = f[0];
        if (fabs(f[1]) > LB_VSMALL || fabs(f[2]) > LB_VSMALL) k0 *= exp(fma(f[1], lnT, -f[2] * invT));
        const double kInf = k;
        const double Pr = k0 * ME[V2_EFF(w)] / kInf;
        double F = 1.0;
        if (ty == LBCHEM_R_FALLOFF_TROE) {
            const double al = f[3], Tsss = f[4], Ts = f[5], Tss = f[6];
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
    if (V2_REV(w)) {
        const double* G = (T < S.tcTab[V2_TCI(w)]) ? GL : GH;
        const unsigned long long lp = S.lpk[p], rp = S.rpk[p];
        double gs = 0.0;
        for (unsigned s = 0; s < V2_NR(w); ++s) { const unsigned sl = V2_SLOT(rp, s); gs = fma((double)V2_NU(sl), G[V2_SP(sl)], gs); }
        for (unsigned s = 0; s < V2_NL(w); ++s) { const unsigned sl = V2_SLOT(lp, s); gs = fma(-(double)V2_NU(sl), G[V2_SP(sl)], gs); }
        const double arg = -gs * invT;
        double Kc = (arg < 600.0) ? exp(arg) : LB_VGREAT;
        const int dn = (int)V2_DNC(w) - 3;
        if (dn != 0) {
            const double x = LB_PSTD / (O.RR * T);
            if (dn == 1) Kc *= x;
            else if (dn == -1) Kc *= 1.0 / x;
            else Kc *= pow(x, (double)dn);
        }
        kr = k / fmax(Kc, O.kcFloor);
    }
}

// clamp c, refresh Gibbs functions and third-body sums for temperature T; ends with a group sync.
// y: state vector (c, T, p). If `clamp` is false CC is taken as current (finite-difference passes).
template <int N>
__device__ __forceinline__ void prepare(Grp& g, const SMech& S, Work<N>& w, const double* y, double T, double lnT, bool clamp) {
    const int nSp = S.nSp, t = g.tid;
    double cs = 0.0;
    if (t < nSp) {
        if (clamp) { const double c = fmax(y[t], 0.0); w.CC[t] = c; cs = c; }
        w.GL[t] = jha(S.haLo + t, nSp, T) - T * js(S.sLo + t, nSp, T, lnT);
        w.GH[t] = jha(S.haHi + t, nSp, T) - T * js(S.sHi + t, nSp, T, lnT);
    }
    if (clamp) {
        if (t == nSp) w.CC[nSp] = 1.0;
        const double cTot = g.sum(cs);              // barrier inside: CC, GL, GH visible afterwards
        if (t < S.nEff) {
            double m = cTot;
            for (int k = S.effPtr[t]; k < S.effPtr[t + 1]; ++k) m = fma(S.effVal[k], w.CC[S.effSp[k]], m);
            w.ME[t] = m;
        }
    }
    g.sync();
}

// omega of every reaction at (CC, T) into om[pos]; optionally keeps kf, kr
template <int N, bool KEEP>
__device__ __forceinline__ void reaction_sweep(const Grp& g, const SMech& S, const DevOpts& O, Work<N>& w, double T, double lnT,
                                               double invT, double* om) {
    for (int p = g.tid; p < S.nR; p += V2_TPC) {
        double kf, kr;
        rate_constants(S, O, w.GL, w.GH, w.ME, p, T, lnT, invT, kf, kr);
        const unsigned iw = S.info[p];
        const double f = kf * side_product(w.CC, S.lpk[p], V2_NL(iw));
        const double b = kr * side_product(w.CC, S.rpk[p], V2_NR(iw));
        om[p] = f - b;
        if (KEEP) { w.KF[p] = kf; w.KR[p] = kr; }
    }
}

__device__ __forceinline__ double gather_species(const SMech& S, const double* om, int i) {
    double acc = 0.0;
    const int e = S.spPtr[i + 1];
    for (int k = S.spPtr[i]; k < e; ++k) {
        const unsigned u = S.spList[k];
        acc = fma((double)(int)(signed char)(u >> 16), om[u & 0xffffu], acc);
    }
    return acc;
}

// StandardChemistryModel::derivatives (A.8)
template <int N>
__device__ __forceinline__ void derivatives(Grp& g, const SMech& S, const DevOpts& O, Work<N>& w, const double* y, double* dydt) {
    const int nSp = S.nSp, t = g.tid;
    g.sync();
    const double T = y[nSp];
    const double lnT = log(T), invT = 1.0 / T;
    prepare(g, S, w, y, T, lnT, true);
    reaction_sweep<N, false>(g, S, O, w, T, lnT, invT, w.OM);
    g.sync();
    double rho = 0.0, cp = 0.0, dT = 0.0;
    if (t < nSp) {
        const double d = gather_species(S, w.OM, t);
        dydt[t] = d;
        const double ci = w.CC[t];
        const bool lo = T < S.Tc[t];
        rho = S.W[t] * ci;
        cp = ci * (O.RR * jcp((lo ? S.cpLo : S.cpHi) + t, nSp, T));
        dT = (O.RR * jha((lo ? S.haLo : S.haHi) + t, nSp, T)) * d;
    }
    g.sum3(rho, cp, dT);
    cp /= rho;
    dT /= rho * cp;
    if (t == 0) { dydt[nSp] = -dT; dydt[nSp + 1] = 0.0; w.nRhs++; }
    g.sync();
}

// StandardChemistryModel::jacobian (A.9) into the sparse store: JS[entry] (species block, structural
// non-zeros in (row, col) order), JT[i] = d(dc_i/dt)/dT by central difference; dfdx = species rates (A.4 quirk)
template <int N>
__device__ __forceinline__ void jacobian(Grp& g, const V2Mech& M, const SMech& S, const DevOpts& O, Work<N>& w, const double* y, double* dfdx) {
    const int nSp = S.nSp, t = g.tid;
    g.sync();
    const double T = y[nSp];
    {
        const double lnT = log(T), invT = 1.0 / T;
        prepare(g, S, w, y, T, lnT, true);
        reaction_sweep<N, true>(g, S, O, w, T, lnT, invT, w.OM);
    }
    g.sync();
    if (t < N) dfdx[t] = (t < nSp && O.dfdxIsDcdt) ? gather_species(S, w.OM, t) : 0.0;
    {
        // gather program: acc += coef * k * c1 * c2 * c3, entry closed by the `last` flag
        const int nT = M.jthrN[t];
        int e = M.jthrEnt[t];
        double acc = 0.0;
        const unsigned long long* prog = M.jterm + t;
        unsigned long long wd = nT > 0 ? prog[0] : 0ull;
        for (int k = 0; k < nT; ++k) {
            const unsigned long long nx = (k + 1 < nT) ? prog[(size_t)(k + 1) * V2_TPC] : 0ull;   // prefetch
            const unsigned kid = V2_T_K(wd);
            const double kk = (kid & 1u) ? w.KR[kid >> 1] : w.KF[kid >> 1];
            const double d = ((kk * w.CC[V2_T_S1(wd)]) * w.CC[V2_T_S2(wd)]) * w.CC[V2_T_S3(wd)];
            acc = fma((double)V2_T_COEF(wd), d, acc);
            if (V2_T_LAST(wd)) { w.JS[e++] = acc; acc = 0.0; }
            wd = nx;
        }
    }
    g.sync();
    const double delta = 1.0e-3;
    double dp = 0.0;
    for (int pass = 0; pass < 2; ++pass) {
        const double Tp = pass == 0 ? T + delta : T - delta;
        const double lnT = log(Tp), invT = 1.0 / Tp;
        prepare(g, S, w, y, Tp, lnT, false);
        reaction_sweep<N, false>(g, S, O, w, Tp, lnT, invT, w.OM);
        g.sync();
        if (t < nSp) {
            const double d = gather_species(S, w.OM, t);
            if (pass == 0) dp = d; else w.JT[t] = 0.5 * (dp - d) / delta;
        }
        g.sync();
    }
    if (t == 0) w.nJac++;
}

// ------------------------------------------------------------------------------------------------
// dense matrix in registers: thread t holds row t of  dg*I - J
// ------------------------------------------------------------------------------------------------
template <int N>
__device__ __forceinline__ void build_matrix(const Grp& g, const SMech& S, const Work<N>& w, double (&a)[N], double dg) {
    constexpr int nSp = N - 2;
    const int t = g.tid;
    int e = 0, eEnd = 0;
    if (t < nSp) { e = S.rowPtr[t]; eEnd = S.rowPtr[t + 1]; }
    int nextCol = e < eEnd ? (int)S.entCol[e] : 255;
    const double jt = t < nSp ? -w.JT[t] : 0.0;
#pragma unroll
    for (int k = 0; k < N; ++k) {
        double v = 0.0;
        if (k < nSp) {
            if (nextCol == k) {
                v = -w.JS[e];
                ++e;
                nextCol = e < eEnd ? (int)S.entCol[e] : 255;
            }
        } else if (k == nSp) {
            v = jt;
        }
        if (k == t) v += dg;
        a[k] = v;
    }
}

// LUDecompose (A.10) on the register-resident rows. Row t is eliminated in place; rows never move: `pos`
// is the index the row would have in the reference's physically swapped matrix (drives the tie rule),
// `step` the column at which this row became the pivot row. Returns 0 or LBCHEM_CELL_SINGULAR.
template <int N>
__device__ __forceinline__ int lu_factor(Grp& g, Work<N>& w, double (&a)[N], int& step) {
    const int t = g.tid;
    bool done = (t >= N);
    int pos = t;
    step = -1;
    double big = 0.0;
#pragma unroll
    for (int k = 0; k < N; ++k) big = fmax(big, fabs(a[k]));
    if (g.any(!done && big == 0.0)) return LBCHEM_CELL_SINGULAR;
    const double vv = 1.0 / big;
    double myrd = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) {
        const double piv = a[j];
        const unsigned long long key = done ? 0ull : (unsigned long long)__double_as_longlong(vv * fabs(piv)) + 1ull;
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
        const bool c1 = hi == mh;
        const unsigned ml = __reduce_max_sync(0xffffffffu, c1 ? lo : 0u);
        const bool c2 = c1 && lo == ml;
        const unsigned pp = c2 ? (unsigned)pos + 1u : 0u;
        const unsigned mp = __reduce_max_sync(0xffffffffu, pp);
        const bool win = c2 && pp == mp;
        const int par = j & 1;
        double* rowb = w.ROWB + (g.warp * 2 + par) * (N + 2);
        double* slot = w.XCH + (g.warp * 2 + par) * 4;
        if (win) {
            slot[0] = __longlong_as_double((long long)key);
            reinterpret_cast<int*>(slot + 1)[0] = pos;
            reinterpret_cast<int*>(slot + 1)[1] = t;
            const double pv = (piv == 0.0) ? LB_SMALL : piv;
            rowb[N] = 1.0 / pv;
#pragma unroll
            for (int k = 0; k < N; ++k) if (k > j) rowb[k] = a[k];
        }
        g.sync();
        const double* s0 = w.XCH + (0 * 2 + par) * 4;
        const double* s1 = w.XCH + (1 * 2 + par) * 4;
        const unsigned long long k0 = (unsigned long long)__double_as_longlong(s0[0]);
        const unsigned long long k1 = (unsigned long long)__double_as_longlong(s1[0]);
        const int p0 = reinterpret_cast<const int*>(s0 + 1)[0], p1 = reinterpret_cast<const int*>(s1 + 1)[0];
        const int gw = (k1 > k0 || (k1 == k0 && p1 > p0)) ? 1 : 0;
        const int ppos = gw ? p1 : p0;
        const int ptid = gw ? reinterpret_cast<const int*>(s1 + 1)[1] : reinterpret_cast<const int*>(s0 + 1)[1];
        const double* row = w.ROWB + (gw * 2 + par) * (N + 2);
        const double rd = row[N];
        if (t == ptid) {
            done = true; step = j; myrd = rd;
            if (piv == 0.0) a[j] = LB_SMALL;
        } else if (!done) {
            if (pos == j) pos = ppos;
            const double l = a[j] * rd;
            a[j] = l;
#pragma unroll
            for (int k = 0; k < N; ++k) if (k > j) a[k] = fma(-l, row[k], a[k]);
        }
    }
    // publish what the blocked triangular solves need: reciprocal pivots and the 8x8 diagonal blocks
    if (t < N) {
        w.RD[step] = myrd;
        const int mb = step >> 3, mk = step & 7;
        constexpr int nBlk = (N + V2_BLK - 1) / V2_BLK;
#pragma unroll
        for (int b = 0; b < nBlk; ++b) {
            if (b == mb) {
#pragma unroll
                for (int q = 0; q < V2_BLK; ++q) {
                    if (b * V2_BLK + q < N) {
                        if (q < mk) w.LB[b * V2_LSTRIDE + mk * (mk - 1) / 2 + q] = a[b * V2_BLK + q];
                        else w.UB[b * V2_USTRIDE + (8 * mk - mk * (mk - 1) / 2) + (q - mk)] = a[b * V2_BLK + q];
                    }
                }
            }
        }
    }
    if (t == 0) w.nLU++;
    g.sync();
    return 0;
}

// x <- A^-1 x with the register-resident factors (x: shared vector of N, natural order in and out)
template <int N>
__device__ __forceinline__ void lu_solve(Grp& g, Work<N>& w, const double (&a)[N], int step, double* x) {
    constexpr int nBlk = (N + V2_BLK - 1) / V2_BLK;
    const int t = g.tid;
    const int mb = step >> 3;           // -1 >> 3 == -1 for the idle threads t >= N
    g.sync();
    double bt = t < N ? x[t] : 0.0;
    double* xs = w.XS + g.warp * 8;
    // forward substitution, unit lower triangle
#pragma unroll
    for (int b = 0; b < nBlk; ++b) {
        if (mb == b) w.BSH[step] = bt;
        g.sync();
        if (g.lane == 0) {
            double yb[V2_BLK];
#pragma unroll
            for (int k = 0; k < V2_BLK; ++k) {
                if (b * V2_BLK + k < N) {
                    double v = w.BSH[b * V2_BLK + k];
#pragma unroll
                    for (int q = 0; q < V2_BLK; ++q) if (q < k) v = fma(-w.LB[b * V2_LSTRIDE + k * (k - 1) / 2 + q], yb[q], v);
                    yb[k] = v;
                    xs[k] = v;
                }
            }
        }
        __syncwarp();
        if (mb > b) {
#pragma unroll
            for (int q = 0; q < V2_BLK; ++q)
                if (b * V2_BLK + q < N) bt = fma(-a[b * V2_BLK + q], xs[q], bt);
        } else if (mb == b) {
            bt = xs[step & 7];
        }
    }
    // backward substitution, upper triangle, multiply by the reciprocal pivots
#pragma unroll
    for (int b = nBlk - 1; b >= 0; --b) {
        if (mb == b) w.BSH[step] = bt;
        g.sync();
        if (g.lane == 0) {
            double xb[V2_BLK];
#pragma unroll
            for (int k = V2_BLK - 1; k >= 0; --k) {
                if (b * V2_BLK + k < N) {
                    double v = w.BSH[b * V2_BLK + k];
#pragma unroll
                    for (int q = 0; q < V2_BLK; ++q)
                        if (q > k && b * V2_BLK + q < N) v = fma(-w.UB[b * V2_USTRIDE + (8 * k - k * (k - 1) / 2) + (q - k)], xb[q], v);
                    v *= w.RD[b * V2_BLK + k];
                    xb[k] = v;
                    xs[k] = v;
                    if (g.warp == 0) x[b * V2_BLK + k] = v;
                }
            }
        }
        __syncwarp();
        if (mb >= 0 && mb < b) {
#pragma unroll
            for (int q = 0; q < V2_BLK; ++q)
                if (b * V2_BLK + q < N) bt = fma(-a[b * V2_BLK + q], xs[q], bt);
        }
    }
    if (t == 0) w.nSolve++;
    g.sync();
}

// ------------------------------------------------------------------------------------------------
// integrators
// ------------------------------------------------------------------------------------------------
struct Cell { int fail; double theta; int kTarg; };
struct StepState { int forward, first, last, reject, prevReject; double dxTry, dxDid; };

__constant__ int c_nSeq[12] = {2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96};
__constant__ double c_cpu[12] = {10, 15, 22, 33, 48, 71, 102, 149, 212, 307, 434, 625};

// one Rosenbrock34 attempt (A.4) from Y=V0 with step dx; result in V1. The sparse Jacobian (and dfdx in V3)
// of the step start are re-used after a rejected attempt: the reference recomputes identical values.
template <int N>
__device__ __forceinline__ double rosenbrock34_step(Grp& g, const SMech& S, const DevOpts& O, Work<N>& w, Cell& cs, double dx) {
    const double a21 = 2, a31 = 48.0 / 25.0, a32 = 6.0 / 25.0, c21 = -8, c31 = 372.0 / 25.0, c32 = 12.0 / 5.0,
                 c41 = -112.0 / 125.0, c42 = -54.0 / 125.0, c43 = -2.0 / 5.0, b1 = 19.0 / 9.0, b2 = 1.0 / 2.0,
                 b3 = 25.0 / 108.0, b4 = 125.0 / 108.0, e1 = 34.0 / 108.0, e2 = 7.0 / 36.0, e3 = 0, e4 = 125.0 / 108.0,
                 gamma = 1.0 / 2.0, c1X = 1.0 / 2.0, c2X = -3.0 / 2.0, c3X = 121.0 / 50.0, c4X = 29.0 / 250.0;
    const int t = g.tid;
    double *Y0 = w.V[0], *Y = w.V[1], *D0 = w.V[2], *DX = w.V[3], *K1 = w.V[4], *K2 = w.V[5], *K3 = w.V[6], *K4 = w.V[7], *DY = w.V[8];
    double a[N];
    int step;
    g.sync();
    build_matrix<N>(g, S, w, a, 1.0 / (gamma * dx));
    const int st = lu_factor<N>(g, w, a, step);
    if (st) { cs.fail = st; return 0.0; }
    if (t == 0) w.nJac++;    // algorithmic count: the reference evaluates the Jacobian in every attempt
#pragma unroll 1
    for (int s = 0; s < 4; ++s) {
        if (s == 1 || s == 2) {
            if (t < N) Y[t] = (s == 1) ? Y0[t] + a21 * K1[t] : Y0[t] + a31 * K1[t] + a32 * K2[t];
            derivatives<N>(g, S, O, w, Y, DY);
        }
        double* K = w.V[4] + s * N;   // V[4..7] are contiguous
        if (t < N) {
            double r;
            if (s == 0) r = D0[t] + dx * c1X * DX[t];
            else if (s == 1) r = DY[t] + dx * c2X * DX[t] + c21 * K1[t] / dx;
            else if (s == 2) r = DY[t] + dx * c3X * DX[t] + (c31 * K1[t] + c32 * K2[t]) / dx;
            else r = DY[t] + dx * c4X * DX[t] + (c41 * K1[t] + c42 * K2[t] + c43 * K3[t]) / dx;
            K[t] = r;
        }
        lu_solve<N>(g, w, a, step, K);
    }
    double em = 0.0;
    if (t < N) {
        const double yi = Y0[t] + b1 * K1[t] + b2 * K2[t] + b3 * K3[t] + b4 * K4[t];
        const double er = e1 * K1[t] + e2 * K2[t] + e3 * K3[t] + e4 * K4[t];
        Y[t] = yi;
        em = fabs(er) / (O.absTol + O.relTol * fmax(fabs(Y0[t]), fabs(yi)));
    }
    return g.maxnan(em);
}

// adaptiveSolver::solve (A.3)
template <int N>
__device__ __forceinline__ void adaptive_solve(Grp& g, const V2Mech& M, const SMech& S, const DevOpts& O, Work<N>& w, Cell& cs,
                                               double& x, double& dxTry) {
    const double safeScale = 0.9, alphaInc = 0.2, alphaDec = 0.25, minScale = 0.2, maxScale = 10;
    const int t = g.tid;
    double dx = dxTry, err = 0.0;
    // dydx0 and the Jacobian share the rate evaluation at (y0, T0): dydx0[i] == dfdx[i] for species
    jacobian<N>(g, M, S, O, w, w.V[0], w.V[3]);
    if (t == 0) w.nJac--;                 // counted per attempt below
    derivatives<N>(g, S, O, w, w.V[0], w.V[2]);
    if (!O.dfdxIsDcdt) { /* DX already zero */ }
    do {
        if (t == 0) w.nAttempts++;
        err = rosenbrock34_step<N>(g, S, O, w, cs, dx);
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
    if (t < N) w.V[0][t] = w.V[1][t];
    g.sync();
    if (err > pow(maxScale / safeScale, -1.0 / alphaInc))
        dxTry = fmin(fmax(safeScale * pow(err, -alphaInc), minScale), maxScale) * dx;
    else
        dxTry = safeScale * maxScale * dx;
}

// seulex::seul (A.5): one extrapolation column; y0 = V4, scale = V5, result in V6
template <int N>
__device__ __forceinline__ bool seul(Grp& g, const SMech& S, const DevOpts& O, Work<N>& w, Cell& cs, double dxTot, int k) {
    const int t = g.tid;
    const int nSteps = c_nSeq[k];
    const double dx = dxTot / nSteps;
    double *y0 = w.V[4], *scale = w.V[5], *yOut = w.V[6], *dy = w.V[7], *yT = w.V[1], *f = w.V[8], *dyc = w.V[9];
    double a[N];
    int step;
    g.sync();
    build_matrix<N>(g, S, w, a, 1 / dx);
    const int st = lu_factor<N>(g, w, a, step);
    if (st) { cs.fail = st; return false; }
    bool ok = true;
#pragma unroll 1
    for (int it = 0; it < nSteps; ++it) {
        if (it == 1) { if (t < N) yT[t] = y0[t] + dy[t]; }
        else if (it > 1) { if (t < N) yT[t] += dy[t]; }
        derivatives<N>(g, S, O, w, it == 0 ? y0 : yT, f);
        if (it == 1 && k <= 1) {
            double s1 = 0.0;
            if (t < N) { const double q = dy[t] / scale[t]; s1 = q * q; dyc[t] = f[t] - dy[t] / dx; }
            const double dy1 = sqrt(g.sum(s1));
            lu_solve<N>(g, w, a, step, dyc);
            const double denom = fmin(1.0, dy1 + LB_SMALL);
            double s2 = 0.0; int big = 0;
            if (t < N) {
                if (fabs(dyc[t]) > scale[t] * denom) big = 1;
                const double q = dyc[t] / scale[t];
                s2 = q * q;
            }
            if (g.any(big)) { cs.theta = 1; ok = false; break; }
            const double dy2 = sqrt(g.sum(s2));
            cs.theta = dy2 / denom;
            if (cs.theta > 1) { ok = false; break; }
            if (t == 0) w.nRhs++;      // the reference evaluates derivatives(yT) a second time (same value)
        }
        if (t < N) dy[t] = f[t];
        lu_solve<N>(g, w, a, step, dy);
    }
    if (ok) { if (t < N) yOut[t] = yT[t] + dy[t]; }
    g.sync();
    return ok;
}

__device__ __forceinline__ double seulex_coeff(int k, int l) {
    const double ratio = (double)c_nSeq[k] / c_nSeq[l];
    return 1 / (ratio - 1);
}

// seulex::solve (A.5); y = V0
template <int N>
__device__ __forceinline__ void seulex_solve(Grp& g, const V2Mech& M, const SMech& S, const DevOpts& O, Work<N>& w, Cell& cs,
                                             double& x, StepState& step, int& attempts) {
    const int t = g.tid;
    const int kMaxx = 11;
    const double stepFactor1 = 0.6, stepFactor2 = 0.93, stepFactor3 = 0.1, stepFactor4 = 4, stepFactor5 = 0.5,
                 kFactor1 = 0.7, kFactor2 = 0.9;
    const double absTol = O.absTol, relTol = O.relTol;
    const double jacRedo = fmin(1e-4, relTol);
    double *y = w.V[0], *y0 = w.V[4], *scale = w.V[5], *ySeq = w.V[6], *tb = w.TB;
    double temp[12], dxOpt[12];
    temp[0] = LB_GREAT;
    double dx = step.dxTry;
    g.sync();
    if (t < N) { y0[t] = y[t]; scale[t] = absTol + relTol * fabs(y[t]); }
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
    if (cs.theta > jacRedo) { jacobian<N>(g, M, S, O, w, y, w.V[3]); jacUpdated = true; }
    int k = 0;
    double dxNew = fabs(dx);
    bool firstk = true;
    while (firstk || step.reject) {
        dx = step.forward ? dxNew : -dxNew;
        firstk = false;
        step.reject = 0;
        attempts++;
        if (t == 0) w.nAttempts++;
        if (fabs(dx) < LB_VSMALL || attempts > 100000) { cs.fail = LBCHEM_CELL_UNDERFLOW; return; }
        if (cs.kTarg + 1 > M.tbRows) { cs.fail = LBCHEM_CELL_MAX_STEPS; return; }   // extrapolation table too small for this order
        double errOld = 0;
#pragma unroll 1
        for (k = 0; k <= cs.kTarg + 1; ++k) {
            const bool success = seul<N>(g, S, O, w, cs, dx, k);
            if (cs.fail) return;
            if (!success) { step.reject = 1; dxNew = fabs(dx) * stepFactor5; break; }
            double s = 0.0;
            if (t < N) {
                if (k == 0) y[t] = ySeq[t];
                else {
                    tb[(k - 1) * N + t] = ySeq[t];
                    for (int j = k - 1; j > 0; --j)
                        tb[(j - 1) * N + t] = tb[j * N + t] + seulex_coeff(k, j) * (tb[j * N + t] - tb[(j - 1) * N + t]);
                    const double yi = tb[t] + seulex_coeff(k, 0) * (tb[t] - y[t]);
                    y[t] = yi;
                    const double sc = absTol + relTol * fabs(y0[t]);
                    scale[t] = sc;
                    const double q = (yi - tb[t]) / sc;
                    s = q * q;
                }
            }
            if (k != 0) {
                const double err = sqrt(g.sum(s) / N);
                if (!(err == err)) { cs.fail = LBCHEM_CELL_NAN; return; }
                if (err > 1 / LB_SMALL || (k > 1 && err >= errOld)) { step.reject = 1; dxNew = fabs(dx) * stepFactor5; break; }
                errOld = fmin(4 * err, 1.0);
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
                if (cs.theta > jacRedo && !jacUpdated) { jacobian<N>(g, M, S, O, w, y, w.V[3]); jacUpdated = true; }
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

// ODESolver::solve (A.2); y = V0
template <int N>
__device__ __forceinline__ void ode_solve(Grp& g, const V2Mech& M, const SMech& S, const DevOpts& O, Work<N>& w, Cell& cs,
                                          double xStart, double xEnd, double& dxTry) {
    StepState step = {1, 1, 0, 0, 0, dxTry, 0.0};
    double x = xStart;
    int attempts = 0;
#pragma unroll 1
    for (int nStep = 0; nStep < O.maxSteps; ++nStep) {
        const double dxTry0 = step.dxTry;
        step.reject = 0;
        if ((x + step.dxTry - xEnd) * (x + step.dxTry - xStart) > 0) {
            step.last = 1;
            step.dxTry = xEnd - x;
        }
        if (O.solver == LBCHEM_SOLVER_SEULEX) {
            seulex_solve<N>(g, M, S, O, w, cs, x, step, attempts);
        } else {
            const double x0 = x;
            adaptive_solve<N>(g, M, S, O, w, cs, x, step.dxTry);
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

template <int N, int SOLVER>
__global__ void __launch_bounds__(V2_TPC * 6, 1) integrate_v2_kernel(V2Mech M, DevOpts O, BatchArgs B, int groups) {
    extern __shared__ double smem[];
    // stage the packed mechanism once per CTA
    {
        const unsigned long long* src = reinterpret_cast<const unsigned long long*>(M.blob);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(smem);
        for (int i = threadIdx.x; i < M.blobBytes / 8; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    SMech S;
    bind(S, M, reinterpret_cast<const unsigned char*>(smem));
    Grp g;
    g.tid = threadIdx.x % V2_TPC; g.gid = threadIdx.x / V2_TPC; g.lane = threadIdx.x & 31; g.warp = g.tid >> 5; g.tog = 0;
    Work<N> w;
    carve<N>(w, g, smem + M.blobBytes / 8 + (size_t)g.gid * M.wsDoubles, M);
    O.solver = SOLVER;
    constexpr int nSp = N - 2;
    const int t = g.tid;
    unsigned long long st_steps = 0, st_att = 0, st_jac = 0, st_lu = 0, st_solve = 0, st_rhs = 0, st_fail = 0, st_cells = 0;

    for (;;) {
        g.sync();
        if (t == 0) reinterpret_cast<unsigned long long*>(w.XS)[0] = atomicAdd(B.queue, 1ull);
        g.sync();
        const unsigned long long q = reinterpret_cast<unsigned long long*>(w.XS)[0];
        if (q >= (unsigned long long)B.n) break;
        const long long cell = B.order ? (long long)B.order[q] : (long long)q;
        const double T0 = B.T[cell];
        const double p0 = B.p[cell];
        const double rho = B.rho[cell];
        const double deltaT = B.deltaT ? B.deltaT[cell] : B.deltaTUniform;
        double dtChem = B.deltaTChem[cell];
        w.nSteps = w.nAttempts = w.nJac = w.nLU = w.nSolve = w.nRhs = 0;
        Cell cs; cs.fail = 0; cs.theta = 0.0; cs.kTarg = 1;
        int status;
        double Tend = T0;
        if (T0 > O.Treact) {
            g.sync();
            if (t < nSp) w.V[0][t] = rho * B.Y[(size_t)t * B.ldy + cell] / S.W[t];
            if (t == nSp) w.V[0][nSp] = T0;
            if (t == nSp + 1) w.V[0][nSp + 1] = p0;
            g.sync();
            double timeLeft = deltaT;
            while (timeLeft > LB_SMALL) {
                double dt = timeLeft;
                ode_solve<N>(g, M, S, O, w, cs, 0.0, dt, dtChem);
                g.sync();
                if (t < nSp) w.V[0][t] = fmax(0.0, w.V[0][t]);
                g.sync();
                if (cs.fail) break;
                timeLeft -= dt;
            }
            dtChem = fmin(dtChem, O.deltaTChemMax);
            if (t < nSp) {
                const double c0 = rho * B.Y[(size_t)t * B.ldy + cell] / S.W[t];
                B.RR[(size_t)t * B.ldy + cell] = (w.V[0][t] - c0) * S.W[t] / deltaT;
            }
            Tend = w.V[0][nSp];
            status = cs.fail ? cs.fail : LBCHEM_CELL_OK;
        } else {
            if (t < nSp) B.RR[(size_t)t * B.ldy + cell] = 0.0;
            status = LBCHEM_CELL_SKIPPED;
        }
        if (t == 0) {
            B.deltaTChem[cell] = dtChem;
            if (B.Tout) B.Tout[cell] = Tend;
            if (B.cost) B.cost[cell] = (double)w.nAttempts;
            if (B.status) B.status[cell] = status;
            st_steps += w.nSteps; st_att += w.nAttempts; st_jac += w.nJac; st_lu += w.nLU;
            st_solve += w.nSolve; st_rhs += w.nRhs; st_fail += (status >= LBCHEM_CELL_MAX_STEPS); st_cells += 1;
        }
    }
    if (t == 0 && st_cells) {
        atomicAdd(B.stats + 0, st_steps); atomicAdd(B.stats + 1, st_att); atomicAdd(B.stats + 2, st_jac);
        atomicAdd(B.stats + 3, st_lu); atomicAdd(B.stats + 4, st_solve); atomicAdd(B.stats + 5, st_rhs);
        atomicAdd(B.stats + 6, st_fail); atomicAdd(B.stats + 7, st_cells);
    }
}

}  // namespace v2

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
template <int N, int SOLVER>
static cudaError_t v2_configure_one(size_t smemBytes) {
    return cudaFuncSetAttribute(v2::integrate_v2_kernel<N, SOLVER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
}

bool lb_v2_supported_size(int nEqn) { return nEqn == 55; }

cudaError_t lb_v2_configure(int nEqn, int solver, size_t smemBytes) {
    if (nEqn != 55) return cudaErrorInvalidValue;
    return solver == LBCHEM_SOLVER_SEULEX ? v2_configure_one<55, LBCHEM_SOLVER_SEULEX>(smemBytes)
                                          : v2_configure_one<55, LBCHEM_SOLVER_ROSENBROCK34>(smemBytes);
}

cudaError_t lb_v2_launch(int nEqn, int solver, int grid, int groups, size_t smemBytes, const V2Mech& M, const DevOpts& O,
                         const BatchArgs& B, cudaStream_t s) {
    if (nEqn != 55) return cudaErrorInvalidValue;
    if (solver == LBCHEM_SOLVER_SEULEX)
        v2::integrate_v2_kernel<55, LBCHEM_SOLVER_SEULEX><<<grid, groups * V2_TPC, smemBytes, s>>>(M, O, B, groups);
    else
        v2::integrate_v2_kernel<55, LBCHEM_SOLVER_ROSENBROCK34><<<grid, groups * V2_TPC, smemBytes, s>>>(M, O, B, groups);
    return cudaGetLastError();
}
