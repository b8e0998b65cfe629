// lbchem_kernels_v3.cu — compact-code generation of the per-cell chemistry integration for sm_100a.
//
// Why: the v2 kernel (register-resident, fully unrolled LU) is 29 k SASS instructions; ncu shows 71 % of its
// warp samples stalled on instruction fetch (profiles/r01_ncu_v2_seulex.txt). This kernel keeps every hot
// phase a SHORT LOOP that stays resident in the 32 KB instruction cache:
//   * the (I/(g dx) - J) matrix lives in shared memory (species block + temperature column, 53 x 54 for GRI;
//     the T and p rows of OpenFOAM's Jacobian are zero, so those two equations decouple exactly),
//     LU with implicit row pivoting: rows never move, one named barrier per column;
//   * triangular solves are column loops over the same storage;
//   * rate evaluation is split into passes with ONE exp() call site: (1) arguments of every Arrhenius and
//     equilibrium exponential, (2) a flat loop exponentiating 2*nR values, (3) assembly of kf, kr and omega,
//     fall-off reactions in their own loop.
// Arithmetic (operation order inside each formula, elimination order, pivot tie rule) is that of v2, which
// is parity-green against the oracle; see DESIGN.md.
// 64 threads integrate one cell; a persistent CTA stages the packed mechanism (lbchem_v2.cuh) once.
#include <cuda_runtime.h>
#include <math.h>
#include "lbchem_device.cuh"
#include "lbchem_v2.cuh"
#include "../../include/lbchem.h"

namespace v3 {


struct SMech {            // typed views into the staged blob
    int nSp, nR, nEff;
    const double *W, *Tc, *cpLo, *cpHi, *haLo, *haHi, *sLo, *sHi;
    const double *A, *beta, *Ta, *fall, *tcTab, *effVal;
    const unsigned* info; const unsigned long long *lpk, *rpk;
    const int *effPtr, *effSp, *spPtr, *rowPtr, *gaTab;
    const unsigned short* spList; const unsigned char* entCol;
};

__device__ __forceinline__ void bind(SMech& S, const V2Mech& M, const unsigned char* b) {
    S.nSp = M.nSp; S.nR = M.nR; S.nEff = M.nEff;
#define P(T, off) reinterpret_cast<const T*>(b + M.off)
    S.W = P(double, oW); S.Tc = P(double, oTc); S.cpLo = P(double, oCpLo); S.cpHi = P(double, oCpHi);
    S.haLo = P(double, oHaLo); S.haHi = P(double, oHaHi); S.sLo = P(double, oSLo); S.sHi = P(double, oSHi);
    S.A = P(double, oA); S.beta = P(double, oBeta); S.Ta = P(double, oTa); S.fall = P(double, oFall);
    S.tcTab = P(double, oTcTab); S.effVal = P(double, oEffVal);
    S.info = P(unsigned, oInfo); S.lpk = P(unsigned long long, oLpk); S.rpk = P(unsigned long long, oRpk);
    S.effPtr = P(int, oEffPtr); S.effSp = P(int, oEffSp); S.spPtr = P(int, oSpPtr); S.rowPtr = P(int, oRowPtr); S.gaTab = P(int, oGaTab);
    S.spList = P(unsigned short, oSpList); S.entCol = P(unsigned char, oEntCol);
#undef P
}

// The thermodynamic polynomials (one batch of independent coalesced loads per prepare() / finish_derivatives()), the
// Arrhenius beta / Ta (prefetched a round ahead), the fall-off parameters and the matrix sparsity lists stay in global
// memory (L2-resident): the shared-memory copy of the mechanism shrinks from 28 KB to 13 KB for GRI, which is what lets a
// seventh cell fit per SM.
__device__ __forceinline__ void bind_cold(SMech& S, const V2Mech& M) {
    if (M.hotBytes >= M.blobBytes) return;
#define PC(T, off) reinterpret_cast<const T*>(M.blob + M.off)
    S.cpLo = PC(double, oCpLo); S.cpHi = PC(double, oCpHi); S.rowPtr = PC(int, oRowPtr); S.entCol = PC(unsigned char, oEntCol);
    S.sLo = PC(double, oSLo); S.sHi = PC(double, oSHi); S.haLo = PC(double, oHaLo); S.haHi = PC(double, oHaHi);
    S.beta = PC(double, oBeta); S.Ta = PC(double, oTa); S.fall = PC(double, oFall);
#undef PC
}

// Per-CTA context in shared memory: mechanism views, options and the layout of a group's workspace
// (offsets in doubles from the group's base). Every routine takes (cx, base).
struct Ctx {
    SMech S;
    DevOpts O;
    int nEqn, nEnt, nFall, tbRows, lda, wsDoubles;
    int oV, oCC, oGL, oGH, oME, oOM, oBS, oSTEP, oXCH, oXS, oCNT, oA, oMSK, oSX, oRL;
    const unsigned* jterm; const unsigned long long* dterm; int nD; const int* jthrEnt; const int* jthrN;
    const int* dac;                               // DAC program, null when reduction is off
    double* scratch; int scratchStride, groups;   // per-group global (L2-resident) scratch: sparse Jacobian, seulex table
};
enum { CNT_STEPS = 0, CNT_ATTEMPTS, CNT_JAC, CNT_LU, CNT_SOLVE, CNT_RHS, CNT_N, CNT_DIAG = 6, CNT_NODE = 7 };   // CNT_DIAG: the factors in the workspace come from the diagonal-pivot path; CNT_NODE: equations of the ODE system (NsDAC_+2 when reduced)

__host__ __device__ __forceinline__ int ctx_layout(Ctx& c, const V2Mech& M) {
    const int N = M.nEqn;
    int p = 32;                               // reduction scratch
    c.oXCH = p; p += 16; c.oXS = p; p += 2; c.oCNT = p; p += 4;
    c.oMSK = p; c.oRL = p;
    if (M.dac) {                              // DAC only (the reduced instantiation of the kernel):
        p += (M.nR + 63) / 64 + 1;            //   disabled-reaction bits (by packed position), then the active-species word
        //   int count + 16-bit list of the packed positions of the ENABLED reactions: its own block, or (rlInMatrix) the
        //   tail of the matrix region -- a reduced matrix never reaches it, and when NO species was removed there is no
        //   list (count -1 = every reaction, identity order). That is what lets a seventh reduced cell fit per SM.
        c.oRL = p; if (!M.rlInMatrix) p += 1 + (M.nR + 3) / 4;
    }
    // vectors that other threads read (inputs of derivatives(), right-hand sides of lu_solve()) live here; the ones only
    // their owner thread touches (seulex: f(y0), y0, scale; Rosenbrock34: y0, dydx0, dfdx) sit in the group's global
    // scratch, and seul()'s column result travels in a register
    c.oV = p; p += 5 * N;
    c.oCC = p; p += M.nSp + 1; c.oGL = p; p += M.nSp; c.oGH = p; p += M.nSp; c.oME = p; p += M.nEff;
    c.oOM = p; p += M.nR;
    c.oBS = p; p += N;
    c.oSX = p; if (M.tbRows > 0) p += 24;     // seulex: temp[12], dxOpt[12] (group-uniform step-size bookkeeping)
    c.oSTEP = p; p += N + 1;                  // int STEP[N] (pivot column of each row), then int SPOF[N] (DAC: species of a compact column)
    c.lda = (M.nSp + 1) | 1;                  // odd (GRI: 55): one column read across the 16 rows of a half-warp touches 16 distinct 8-byte banks
    c.oA = p; p += (M.nSp * c.lda > 2 * M.nR + M.nD) ? M.nSp * c.lda : 2 * M.nR + M.nD;   // KF/KR and the Jacobian's D terms alias the matrix
    if (M.dac && M.rlInMatrix) c.oRL = p - (1 + (M.nR + 3) / 4);
    return (p + 1) & ~1;
}
#define WS(name) (base + cx->o##name)
// global scratch of this group: [nEnt] sparse Jacobian JS, then [tbRows x nEqn] seulex extrapolation table
#define GS_JS (cx->scratch + (size_t)(blockIdx.x * cx->groups + threadIdx.x / TPC) * cx->scratchStride)
#define GS_TB (GS_JS + cx->nEnt)
#define GS_GV(i) (GS_JS + cx->nEnt + 11 * cx->nEqn + (i) * cx->nEqn)     /* owner-only vectors, i = 0..2 */
#define GS_JT (GS_JS + cx->nEnt + 14 * cx->nEqn)                         /* d(dc/dt)/dT column of the Jacobian */
#define VEC(i) (base + cx->oV + (i) * cx->nEqn)

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

__device__ __forceinline__ double cpow(double c, unsigned e) {   // e in 1..3; 1 in nearly every slot: two untaken branches
    double r = c;
    if (e > 1) { r = c * c; if (e > 2) r = r * c; }
    return r;
}

__device__ __forceinline__ double side_product(const double* c, unsigned long long pk, unsigned n) {
    double f = 1.0;
#pragma unroll 1
    for (unsigned s = 0; s < n; ++s) { const unsigned sl = V2_SLOT(pk, s); f *= cpow(c[V2_SP(sl)], V2_EX(sl)); }
    return f;
}


struct Cell { int fail; double theta; int kTarg; int f0valid; };
struct StepState { int forward, first, last, reject, prevReject; double dxTry, dxDid; };

__constant__ int c_nSeq[12] = {2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96};
__constant__ double c_cpu[12] = {10, 15, 22, 33, 48, 71, 102, 149, 212, 307, 434, 625};

// smem layout: [Ctx (padded to 512 B)] [mechanism blob] [groups x workspace]
constexpr int CTX_BYTES = 512;
static_assert(sizeof(Ctx) <= CTX_BYTES, "Ctx grew: raise CTX_BYTES");

// threads per CTA: seven 64-thread groups = 14 warps, i.e. four warps on two of the SM's sub-partitions: 16384 / (4 x 32) =
// 128 registers per thread at most. The sub-warp widths (8 / 16 / 32 threads per cell) run 512 threads = 4 warps on every
// sub-partition at the same 128 registers: 64 / 32 / 16 cells per CTA (measured: H2/O2 3.63 -> 4.28 M cells/s, one-step
// CH4 8.94 -> 10.4 M cells/s against 384 threads at 168 registers; the few spills stay in L1)
template <int TPC> struct Width { static constexpr int kThreads = TPC == 64 ? 448 : 512; static constexpr int kMaxGroups = kThreads / TPC; static constexpr int kRegs = 128; };

// ================================================================================================
// Everything below is written once for a group of TPC threads integrating one cell: TPC = 64 (two warps, named
// barriers; GRI-class mechanisms up to 62 species), 32, 16 or 8 (sub-warp groups: several cells share one warp, each
// group synchronises and shuffles only inside its own lanes; H2/O2 with 12 equations runs 24 cells per SM on 16 lanes
// each, the 1-step CH4 mechanism 48 cells on 8 lanes each). Thread t of a group owns species row t, so TPC >= nEqn.
// ================================================================================================
template <int TPC>
struct K {

// Group-level helpers. Threads derive their coordinates from threadIdx; nothing lives in local memory, so the
// heavy routines below can be real (non-inlined) functions with exactly one copy in the instruction stream.
struct Grp {
    int tid, gid, lane, warp;
    unsigned mask;    // lanes of this group inside its warp (all of them when TPC >= 32)
    double* red;      // 32 doubles of exchange scratch at the start of the group's workspace: 4 slots of 8
    static constexpr int kHalf = TPC < 32 ? TPC / 2 : 16;
    static constexpr bool kTwoWarps = TPC > 32;
    __device__ __forceinline__ Grp(double* base) {
        tid = threadIdx.x % TPC; gid = threadIdx.x / TPC; lane = threadIdx.x & 31; warp = tid >> 5; red = base;
        mask = TPC >= 32 ? 0xffffffffu : (((1u << (TPC & 31)) - 1u) << (lane & ~(TPC - 1)));
    }
    // group barrier (also orders the group's shared and global memory accesses): a named barrier for two warps, a
    // (sub-)warp barrier otherwise
    __device__ __forceinline__ void sync() const {
        if (kTwoWarps) asm volatile("bar.sync %0, %1;" ::"r"(gid + 1), "n"(TPC) : "memory");
        else __syncwarp(mask);
    }
    // Reductions write slot `s` (0..3). Two reductions with no barrier in between must use different slots. Every
    // reduction ends with a group barrier (callers rely on it to publish what they wrote before the call).
    __device__ __forceinline__ void sum3(double& a, double& b, double& c, int s) const {
#pragma unroll
        for (int o = kHalf; o > 0; o >>= 1) {
            a += __shfl_xor_sync(mask, a, o); b += __shfl_xor_sync(mask, b, o); c += __shfl_xor_sync(mask, c, o);
        }
        if (!kTwoWarps) { sync(); return; }
        double* r = red + 8 * s;
        if (lane == 0) { r[warp * 3] = a; r[warp * 3 + 1] = b; r[warp * 3 + 2] = c; }
        sync();
        a = r[0] + r[3]; b = r[1] + r[4]; c = r[2] + r[5];
    }
    __device__ __forceinline__ double sum(double a, int s) const {
#pragma unroll
        for (int o = kHalf; o > 0; o >>= 1) a += __shfl_xor_sync(mask, a, o);
        if (!kTwoWarps) { sync(); return a; }
        double* r = red + 8 * s;
        if (lane == 0) r[warp] = a;
        sync();
        return r[0] + r[1];
    }
    __device__ __forceinline__ double maxnan(double v, int s) const {   // max with NaN propagation
        double bad = (v != v) ? 1.0 : 0.0;
        v = (v != v) ? 0.0 : v;
#pragma unroll
        for (int o = kHalf; o > 0; o >>= 1) {
            v = fmax(v, __shfl_xor_sync(mask, v, o)); bad = fmax(bad, __shfl_xor_sync(mask, bad, o));
        }
        if (kTwoWarps) {
            double* r = red + 8 * s;
            if (lane == 0) { r[warp * 2] = v; r[warp * 2 + 1] = bad; }
            sync();
            v = fmax(r[0], r[2]); bad = fmax(r[1], r[3]);
        } else sync();
        return bad > 0.0 ? __longlong_as_double(0x7ff8000000000000LL) : v;
    }
    __device__ __forceinline__ int any(int p, int s) const {
        const int a = __any_sync(mask, p);
        if (!kTwoWarps) { sync(); return a; }
        double* r = red + 8 * s;
        if (lane == 0) r[warp] = a ? 1.0 : 0.0;
        sync();
        return (r[0] + r[1]) > 0.0;
    }
};

// clamp c, refresh Gibbs functions and third-body sums for temperature T; ends with a group sync.
// y: state vector (c, T, p). If `clamp` is false CC is taken as current (finite-difference passes).
static __device__ __forceinline__ void prepare(const Grp& g, const Ctx* cx, double* base, const double* y, double T, double lnT, bool clamp) {
    const SMech& S = cx->S;
    const int nSp = S.nSp, t = g.tid;
    double* CC = WS(CC);
    double cs = 0.0;
    if (t < nSp) {
        if (clamp) { const double c = fmax(y[t], 0.0); CC[t] = c; cs = c; }
        // 24 independent coalesced loads from the L2-resident polynomials: all issued before the first use
        const double gl = jha(S.haLo + t, nSp, T) - T * js(S.sLo + t, nSp, T, lnT);
        const double gh = jha(S.haHi + t, nSp, T) - T * js(S.sHi + t, nSp, T, lnT);
        WS(GL)[t] = gl; WS(GH)[t] = gh;
    }
    if (clamp) {
        if (t == nSp) CC[nSp] = 1.0;
        const double cTot = g.sum(cs, 0);           // barrier inside: CC, GL, GH visible afterwards
        if (t < S.nEff) {
            double m = cTot;
#pragma unroll 1
            for (int k = S.effPtr[t]; k < S.effPtr[t + 1]; ++k) m = fma(S.effVal[k], CC[S.effSp[k]], m);
            WS(ME)[t] = m;
        }
    }
    g.sync();
}

// omega of every reaction at (CC, T) into OM[pos]; keep != 0 also stores kf, kr (Jacobian assembly).
// Pass structure (one exp call site): exponentials -> fall-off blending -> kf, kr, mass action.
template <bool DAC>
static __device__ __noinline__ void reaction_sweep(const Ctx* cx, double* base, int keep, double T, double lnT, double invT) {
    const Grp g(base);
    const SMech& S = cx->S;
    const int nR = S.nR, nFall = cx->nFall;
    double* OM = WS(OM);
    const double *GL = WS(GL), *GH = WS(GH), *ME = WS(ME), *CC = WS(CC);
    // DAC: reactions touching a removed species are skipped (TDACChemistryModel::omega, reactionsDisabled_)
    const unsigned long long* MSK = reinterpret_cast<const unsigned long long*>(WS(MSK));
    constexpr bool dac = DAC;
#define RX_OFF(p) (dac && ((MSK[(p) >> 6] >> ((p) & 63)) & 1ull))
    // DAC: the sweep walks the compacted list of enabled reactions (half of GRI is typically switched off: three rounds
    // of 64 instead of six); OM of a disabled reaction was zeroed once by dac_reduce and is never touched again
    const int nList = dac ? reinterpret_cast<const int*>(WS(XS))[0] : -1;        // -1: every reaction, identity order (the count sits in the
                                                                                  // queue-ticket slot, free once the cell is fetched: the list itself may share the matrix tail)
    const int nLoop = nList < 0 ? nR : nList;
    const unsigned short* RL = reinterpret_cast<const unsigned short*>(WS(RL) + 1);
#define RX_AT(q) ((dac && nList >= 0) ? (int)RL[q] : (q))
    // (1) OM[p] = exp(beta ln T - Ta/T) (1 when the reaction has no such factor)
    // beta and Ta stream from global memory (L2): the next round's pair is in flight while this round's exp() runs; the
    // fall-off parameters of pass (2) are fetched here as well, ahead of their use
    double ff[7] = {0, 0, 0, 0, 0, 0, 0};
    if (nFall > 0 && g.tid < nFall) {
        const double* f = S.fall + 8 * V2_FALL(S.info[nR - nFall + g.tid]);
#pragma unroll
        for (int i = 0; i < 7; ++i) ff[i] = f[i];
    }
    {
        int q = g.tid, p = 0; unsigned iw = 0; double be = 0.0, ta = 0.0;
        if (q < nLoop) { p = RX_AT(q); iw = S.info[p]; if (V2_HASEXP(iw)) { be = S.beta[p]; ta = S.Ta[p]; } }
#pragma unroll 1
        while (q < nLoop) {
            const int qn = q + TPC;
            int pn = 0; unsigned iwn = 0; double ben = 0.0, tan_ = 0.0;
            if (qn < nLoop) { pn = RX_AT(qn); iwn = S.info[pn]; if (V2_HASEXP(iwn)) { ben = S.beta[pn]; tan_ = S.Ta[pn]; } }
            OM[p] = V2_HASEXP(iw) ? exp(fma(be, lnT, -ta * invT)) : 1.0;
            q = qn; p = pn; iw = iwn; be = ben; ta = tan_;
        }
    }
    g.sync();
    // (2) fall-off reactions (sorted last): OM[p] <- blended rate constant
    if (nFall > 0) {
#pragma unroll 1
        for (int q = g.tid; q < nFall; q += TPC) {
            const int p = nR - nFall + q;
            if (RX_OFF(p)) continue;
            const unsigned iw = S.info[p];
            double fl[7];
            if (q == g.tid) {
#pragma unroll
                for (int i = 0; i < 7; ++i) fl[i] = ff[i];
            } else {
                const double* fg = S.fall + 8 * V2_FALL(iw);
#pragma unroll
                for (int i = 0; i < 7; ++i) fl[i] = fg[i];
            }
            const double* f = fl;
            double k0 = f[0];
            if (fabs(f[1]) > LB_VSMALL || fabs(f[2]) > LB_VSMALL) k0 *= exp(fma(f[1], lnT, -f[2] * invT));
            const double kInf = S.A[p] * OM[p];
            const double Pr = k0 * ME[V2_EFF(iw)] / kInf;
            double F = 1.0;
            if (V2_RTYPE(iw) == LBCHEM_R_FALLOFF_TROE) {
                const double al = f[3], Tsss = f[4], Ts = f[5], Tss = f[6];
                const double logFcent = log10(fmax((1 - al) * exp(-T / Tsss) + al * exp(-T / Ts) + exp(-Tss / T), LB_SMALL));
                const double cc = -0.4 - 0.67 * logFcent;
                const double nn = 0.75 - 1.27 * logFcent;
                const double logPr = log10(fmax(Pr, LB_SMALL));
                const double q2 = (logPr + cc) / (nn - 0.14 * (logPr + cc));
                F = exp(2.302585092994046 * (logFcent / (1.0 + q2 * q2)));      // 10^y
            }
            OM[p] = kInf * (Pr / (1 + Pr)) * F;
        }
        g.sync();
    }
    // (3) kf, kr, omega
    const double x = LB_PSTD / (cx->O.RR * T);
    const double xinv = 1.0 / x;                  // the factor of every reaction that loses a mole: one division per sweep
    const double kcFloor = cx->O.kcFloor;
    double* KF = WS(A); double* KR = KF + nR;
#pragma unroll 1
    for (int q = g.tid; q < nLoop; q += TPC) {
        const int p = RX_AT(q);
        const unsigned iw = S.info[p];
        const unsigned ty = V2_RTYPE(iw);
        double k = OM[p];
        if (ty < LBCHEM_R_FALLOFF_LINDEMANN) {
            k = S.A[p] * k;
            if (ty == LBCHEM_R_THIRD_BODY) k = ME[V2_EFF(iw)] * k;
        }
        double kr = 0.0, pl = 1.0, pr = 1.0;
        const unsigned long long lp = S.lpk[p], rp = S.rpk[p];
        if (V2_REV(iw)) {
            // one walk over the slots of each side gives the Gibbs sum (right side first, as the reference adds them) and
            // the concentration product
            const double* G = (T < S.tcTab[V2_TCI(iw)]) ? GL : GH;
            double gs = 0.0;
#pragma unroll 1
            for (unsigned s = 0; s < V2_NR(iw); ++s) {
                const unsigned sl = V2_SLOT(rp, s), sp = V2_SP(sl);
                gs = fma((double)V2_NU(sl), G[sp], gs); pr *= cpow(CC[sp], V2_EX(sl));
            }
#pragma unroll 1
            for (unsigned s = 0; s < V2_NL(iw); ++s) {
                const unsigned sl = V2_SLOT(lp, s), sp = V2_SP(sl);
                gs = fma(-(double)V2_NU(sl), G[sp], gs); pl *= cpow(CC[sp], V2_EX(sl));
            }
            const double arg = -gs * invT;
            double Kc = (arg < 600.0) ? exp(arg) : LB_VGREAT;
            const int dn = (int)V2_DNC(iw) - 3;
            if (dn != 0) {
                if (dn == 1) Kc *= x;
                else if (dn == -1) Kc *= xinv;
                else Kc *= pow(x, (double)dn);
            }
            kr = k / fmax(Kc, kcFloor);
        } else pl = side_product(CC, lp, V2_NL(iw));
        const double f = k * pl;
        const double b = kr * pr;
        OM[p] = f - b;
        if (keep) { KF[p] = k; KR[p] = kr; }
    }
#undef RX_OFF
#undef RX_AT
    g.sync();
}

// dc_i/dt = sum over the reactions of species i of nu * omega: four interleaved partial sums (terms k, k+1, k+2, k+3 of the
// reference's order, combined as (a0 + a1) + (a2 + a3)) over 16-bit list entries read four at a time (lists padded with nu = 0).
// The species sums of a whole group, dealt to all its threads (host: gaTab): thread t walks its part of a list, the threads
// past the species rows hand their partial sums to the owners through the exchange scratch, one group barrier. Returns
// dc_t/dt in thread t < nSp (own part + helper parts in list order), 0 elsewhere. Call with every thread of the group.
static __device__ __forceinline__ double gather_all(const Grp& g, const Ctx* cx, double* base) {
    const SMech& S = cx->S;
    const double* om = WS(OM);
    const int t = g.tid, nSp = S.nSp;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const int e = S.gaTab[64 + t];
    const unsigned long long* L = reinterpret_cast<const unsigned long long*>(S.spList);
#pragma unroll 1
    for (int k = S.gaTab[t]; k < e; k += 4) {
        const unsigned long long w = L[k >> 2];
        const int u0 = (int)(short)(w & 0xffffu), u1 = (int)(short)((w >> 16) & 0xffffu), u2 = (int)(short)((w >> 32) & 0xffffu),
                  u3 = (int)(short)(w >> 48);
        a0 = fma((double)(u0 >> 12), om[u0 & 0xfff], a0);
        a1 = fma((double)(u1 >> 12), om[u1 & 0xfff], a1);
        a2 = fma((double)(u2 >> 12), om[u2 & 0xfff], a2);
        a3 = fma((double)(u3 >> 12), om[u3 & 0xfff], a3);
    }
    double v = (a0 + a1) + (a2 + a3);
    if (S.gaTab[191] == 0) return t < nSp ? v : 0.0;          // short lists (small mechanisms): nothing was split
    double* H = g.red + 16;                       // slots 2 and 3 of the exchange scratch
    if (t >= nSp && t < nSp + 16) H[t - nSp] = v;
    g.sync();
    if (t < nSp) {
        const int h = S.gaTab[128 + t];
#pragma unroll 1
        for (int i = 0; i < (h >> 8); ++i) v += H[(h & 255) + i];
    } else v = 0.0;
    return v;
}

// tail of StandardChemistryModel::derivatives: dT/dt = -sum h_i w_i / (rho cp) from the species rates `d` (thread t owns
// species t), CC and T current; writes dydt and counts one derivatives() evaluation
static __device__ __forceinline__ void finish_derivatives(const Grp& g, const Ctx* cx, double* base, double T, double d, double* dydt) {
    const SMech& S = cx->S;
    const int nSp = S.nSp, t = g.tid;
    const double RR = cx->O.RR;
    double rho = 0.0, cp = 0.0, dT = 0.0;
    if (t < nSp) {
        dydt[t] = d;
        const double ci = WS(CC)[t];
        const bool lo = T < S.Tc[t];
        rho = S.W[t] * ci;
        cp = ci * (RR * jcp((lo ? S.cpLo : S.cpHi) + t, nSp, T));
        dT = (RR * jha((lo ? S.haLo : S.haHi) + t, nSp, T)) * d;
    }
    g.sum3(rho, cp, dT, 1);
    cp /= rho;
    dT /= rho * cp;
    if (t == 0) { dydt[nSp] = -dT; dydt[nSp + 1] = 0.0; reinterpret_cast<int*>(WS(CNT))[CNT_RHS]++; }
    g.sync();
}

// StandardChemistryModel::derivatives (A.8): y, dydt are vectors of the group's workspace
template <bool DAC>
static __device__ __noinline__ void derivatives(const Ctx* cx, double* base, const double* y, double* dydt) {
    const Grp g(base);
    const SMech& S = cx->S;
    const int nSp = S.nSp, t = g.tid;
    g.sync();
    const double T = y[nSp];
    const double lnT = log(T), invT = 1.0 / T;
    prepare(g, cx, base, y, T, lnT, true);
    reaction_sweep<DAC>(cx, base, 0, T, lnT, invT);
    finish_derivatives(g, cx, base, T, gather_all(g, cx, base), dydt);
}

// StandardChemistryModel::jacobian (A.9) into the sparse store: JS[entry] (species block, structural
// non-zeros in (row, col) order), JT[i] = d(dc_i/dt)/dT by central difference; dfdx = species rates (A.4 quirk)
template <bool DAC>
static __device__ __noinline__ void jacobian(const Ctx* cx, double* base, const double* y, double* dfdx, double* dydt0) {
    const Grp g(base);
    const SMech& S = cx->S;
    const int nSp = S.nSp, t = g.tid, N = cx->nEqn;
    g.sync();
    const double T = y[nSp];
    const double delta = 1.0e-3;
    double dp = 0.0;
#pragma unroll 1
    for (int pass = 0; pass < 3; ++pass) {
        const double Tp = pass == 0 ? T : (pass == 1 ? T + delta : T - delta);
        const double lnT = log(Tp), invT = 1.0 / Tp;
        prepare(g, cx, base, y, Tp, lnT, pass == 0);
        reaction_sweep<DAC>(cx, base, pass == 0, Tp, lnT, invT);
        if (pass == 0) {
            // species rates at (y, T): the dfdx slot (A.4 quirk) and, on request, the full derivatives() at the same state
            // (adaptiveSolver's dydx0: the reference evaluates it separately from the same rates)
            const double d0 = ((dfdx && cx->O.dfdxIsDcdt) || dydt0) ? gather_all(g, cx, base) : 0.0;
            if (dfdx && t < N) dfdx[t] = (t < nSp && cx->O.dfdxIsDcdt) ? d0 : 0.0;
            if (dydt0) finish_derivatives(g, cx, base, T, d0, dydt0);
            // (1) D terms: k * product of the other concentrations, one per (reaction, side, species), into the matrix
            //     region behind KF/KR; each is shared by every Jacobian row its reaction touches
            const double* KFR = WS(A); const double* CC = WS(CC); double* JS = GS_JS;
            const int nR = S.nR, nD = cx->nD;
            double* D = WS(A) + 2 * nR;
            {
                const unsigned long long* dp = cx->dterm;
                const unsigned long long* MSKj = reinterpret_cast<const unsigned long long*>(WS(MSK));
#pragma unroll 4
                for (int k = t; k < nD; k += TPC) {
                    const unsigned long long wd = dp[k];
                    const unsigned kid = V2_T_K(wd);
                    const unsigned pos = kid >> 1;
                    double d = 0.0;
                    if (!DAC || !((MSKj[pos >> 6] >> (pos & 63)) & 1ull)) {       // a disabled reaction contributes exact zeros
                        const double kk = KFR[(kid & 1u) * nR + pos];
                        d = ((kk * CC[V2_T_S1(wd)]) * CC[V2_T_S2(wd)]) * CC[V2_T_S3(wd)];
                    }
                    D[k] = d;
                }
            }
            g.sync();
            // (2) entry gather program: acc += coef * D[term], entry closed by the `last` flag. The program streams from
            //     global memory (L2): eight words are fetched ahead so that one round trip is paid per eight terms
            const int nT = cx->jthrN[t];
            double* jsp = JS + cx->jthrEnt[t];
            double acc = 0.0;
            const unsigned* prog = cx->jterm + t;
            constexpr int PF = 8;
            unsigned cur[PF];
#pragma unroll
            for (int q = 0; q < PF; ++q) cur[q] = prog[(size_t)q * TPC];          // padded by PF rows on the host
#pragma unroll 1
            for (int k0 = 0; k0 < nT; k0 += PF) {
                unsigned nxt[PF];
#pragma unroll
                for (int q = 0; q < PF; ++q) nxt[q] = prog[(size_t)(k0 + PF + q) * TPC];
#pragma unroll
                for (int q = 0; q < PF; ++q) {
                    // words past a thread's last term are zero (host padding): they add 0 * D[0] to a sum that is never stored
                    const unsigned wd = cur[q];
                    acc = fma((double)V2_J_COEF(wd), D[V2_J_D(wd)], acc);
                    if (V2_J_LAST(wd)) { *jsp++ = acc; acc = 0.0; }
                }
#pragma unroll
                for (int q = 0; q < PF; ++q) cur[q] = nxt[q];
            }
        } else {
            const double d = gather_all(g, cx, base);
            if (t < nSp) { if (pass == 1) dp = d; else GS_JT[t] = 0.5 * (dp - d) / delta; }
        }
        g.sync();
    }
    if (t == 0) reinterpret_cast<int*>(WS(CNT))[CNT_JAC]++;
}

// ------------------------------------------------------------------------------------------------
// dense linear algebra on the shared-memory matrix: thread t owns the row of species t of  dg*I - J
// (columns: species block + temperature column). The T and p rows are (0 .. dg .. 0): handled in closed form.
//
// DAC (LoadBalancedTDAC, reduction active): the kept species are COMPACTED -- species t sits in row / column
// cm(t) = number of kept species below t, the matrix is NsDAC x (NsDAC + 1) with its own odd stride, and every loop runs
// over NsDAC columns: this is TDACChemistryModel's reduced system (simplifiedToCompleteIndex_ is ascending, so the
// compact order is the reference's order, which its pivot tie rule needs). Threads of removed species idle here; their
// equations are decoupled (dc/dt = 0, diagonal dg).
// ------------------------------------------------------------------------------------------------
template <bool DAC>
struct Compact {
    unsigned long long act; int Ns, ldc;
    __device__ __forceinline__ Compact(const Ctx* cx, const double* base) {
        const int nSp = cx->S.nSp;
        if (DAC) {
            act = reinterpret_cast<const unsigned long long*>(base + cx->oMSK)[(cx->S.nR + 63) >> 6];
            Ns = __popcll(act);
            ldc = (Ns + 1) | 1;
        } else { act = ~0ull; Ns = nSp; ldc = cx->lda; }
    }
    __device__ __forceinline__ bool kept(int sp) const { return !DAC || ((act >> sp) & 1ull); }
    __device__ __forceinline__ int index(int sp) const { return DAC ? __popcll(act & ((1ull << sp) - 1ull)) : sp; }
};

// row t of dg I - J from the sparse store into the shared matrix (compacted when DAC); returns max |row| when asked for
template <bool DAC, bool WANT_BIG>
static __device__ __forceinline__ double lu_build(const Ctx* cx, double* base, const Compact<DAC>& cp, double* row, int* SPOF, bool mine, int cm, double dg) {
    const SMech& S = cx->S;
    const int t = threadIdx.x % TPC, Ns = cp.Ns;
    double big = fabs(dg);
    if (mine) {
#pragma unroll 4
        for (int k = 0; k <= Ns; ++k) row[k] = 0.0;
        const double* JS = GS_JS;
        const int e1 = S.rowPtr[t + 1];
        // JS and the column list stream from global memory (L2): four independent loads per round trip
        int e = S.rowPtr[t];
        if (!DAC) {
#pragma unroll 1
            for (; e + 3 < e1; e += 4) {
                const int c0 = S.entCol[e], c1 = S.entCol[e + 1], c2 = S.entCol[e + 2], c3 = S.entCol[e + 3];
                const double v0 = JS[e], v1 = JS[e + 1], v2 = JS[e + 2], v3 = JS[e + 3];
                row[c0] = -v0; row[c1] = -v1; row[c2] = -v2; row[c3] = -v3;
            }
#pragma unroll 1
            for (; e < e1; ++e) row[S.entCol[e]] = -JS[e];
        } else {
#pragma unroll 1
            for (; e + 3 < e1; e += 4) {
                const int c0 = S.entCol[e], c1 = S.entCol[e + 1], c2 = S.entCol[e + 2], c3 = S.entCol[e + 3];
                const double v0 = JS[e], v1 = JS[e + 1], v2 = JS[e + 2], v3 = JS[e + 3];
                if (cp.kept(c0)) row[cp.index(c0)] = -v0;
                if (cp.kept(c1)) row[cp.index(c1)] = -v1;
                if (cp.kept(c2)) row[cp.index(c2)] = -v2;
                if (cp.kept(c3)) row[cp.index(c3)] = -v3;
            }
#pragma unroll 1
            for (; e < e1; ++e) { const int c = S.entCol[e]; if (cp.kept(c)) row[cp.index(c)] = -JS[e]; }
            SPOF[cm] = t;
        }
        row[Ns] = -GS_JT[t];
        row[cm] += dg;
        if (WANT_BIG) {
            big = 0.0;
#pragma unroll 1
            for (int k = 0; k <= Ns; ++k) big = fmax(big, fabs(row[k]));
        }
    }
    return big;
}

// Factorisation with DIAGONAL pivots (lbchem_options.pivoting = 1, the default): no vote, no exchange. Returns true when
// it is accepted -- every multiplier below 2^10; the reference's choice is the diagonal in 96 % of the columns of the GRI
// flame workloads and the largest multiplier met there is 2.0 -- else the caller factorises the matrix again with the
// reference's rule. pivoting = 2 rejects always (test hook of that fallback).
template <bool DAC>
static __device__ __noinline__ bool lu_factor_diag(const Ctx* cx, double* base, double dg) {
    const Grp g(base);
    const int nSp = cx->S.nSp, t = g.tid, N = cx->nEqn;
    const Compact<DAC> cp(cx, base);
    const int Ns = cp.Ns, ldc = cp.ldc;
    const bool mine = t < nSp && cp.kept(t);
    const int cm = mine ? cp.index(t) : 0;
    double* A = WS(A);
    int* STEP = reinterpret_cast<int*>(WS(STEP));
    int* SPOF = STEP + N + (N & 1);               // DAC: species of compact column j
    double* row = A + cm * ldc;
    g.sync();
    lu_build<DAC, false>(cx, base, cp, row, SPOF, mine, cm, dg);
    // ---- diagonal pivots. Row j is final after step j - 1; its owner leaves 1 / pivot in the diagonal slot (what
    // lu_solve needs as well) before the barrier of step j. The reciprocal is computed by EVERY lane on a harmless
    // operand (1.0 except in the owner of row j + 1), right after the first element of the row update, so that its
    // dependent chain runs under the loads and FMAs of the rest of the row instead of serialising one warp.
    if (mine && cm == 0) { const double d = row[0]; row[0] = 1.0 / ((d == 0.0) ? LB_SMALL : d); }
    if (t < N) STEP[t] = mine ? cm : -1;
    if (t == 0) WS(XS)[1] = 1.0 / dg;
    bool grown = false;
#pragma unroll 1
    for (int j = 0; j < Ns; ++j) {
        g.sync();
        if (mine && cm > j) {
            const double* __restrict__ prow = A + j * ldc;
            double* __restrict__ mine_ = row;
            const double l = mine_[j] * prow[j];
            int k = j + 1;
            const double r1 = fma(-l, prow[k], mine_[k]);
            const bool nxt = cm == k;
            const double den = nxt ? ((r1 == 0.0) ? LB_SMALL : r1) : 1.0;
            double rc;
            mine_[j] = l; grown |= !(fabs(l) <= 1024.0);
            ++k;
            if (k + 7 <= Ns) {
                // first batch of the row, peeled: the Newton chain of the reciprocal (the one 1.0 / den compiles to, without
                // its range check: a pivot outside 2^+-1022 gives inf / NaN multipliers, which the growth test below rejects)
                // sits in the same basic block and is scheduled under the batch's loads and FMAs
                const double p0 = prow[k], p1 = prow[k + 1], p2 = prow[k + 2], p3 = prow[k + 3];
                const double p4 = prow[k + 4], p5 = prow[k + 5], p6 = prow[k + 6], p7 = prow[k + 7];
                const double r0 = mine_[k], q1 = mine_[k + 1], r2 = mine_[k + 2], r3 = mine_[k + 3];
                const double r4 = mine_[k + 4], r5 = mine_[k + 5], r6 = mine_[k + 6], r7 = mine_[k + 7];
                double y;
                asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(den));
                double e = fma(-den, y, 1.0);
                e = fma(e, e, e); y = fma(y, e, y);
                e = fma(-den, y, 1.0); rc = fma(y, e, y);
                mine_[k] = fma(-l, p0, r0); mine_[k + 1] = fma(-l, p1, q1);
                mine_[k + 2] = fma(-l, p2, r2); mine_[k + 3] = fma(-l, p3, r3);
                mine_[k + 4] = fma(-l, p4, r4); mine_[k + 5] = fma(-l, p5, r5);
                mine_[k + 6] = fma(-l, p6, r6); mine_[k + 7] = fma(-l, p7, r7);
                k += 8;
            } else rc = 1.0 / den;
#pragma unroll 1
            for (; k + 7 <= Ns; k += 8) {
                const double p0 = prow[k], p1 = prow[k + 1], p2 = prow[k + 2], p3 = prow[k + 3];
                const double p4 = prow[k + 4], p5 = prow[k + 5], p6 = prow[k + 6], p7 = prow[k + 7];
                const double r0 = mine_[k], q1 = mine_[k + 1], r2 = mine_[k + 2], r3 = mine_[k + 3];
                const double r4 = mine_[k + 4], r5 = mine_[k + 5], r6 = mine_[k + 6], r7 = mine_[k + 7];
                mine_[k] = fma(-l, p0, r0); mine_[k + 1] = fma(-l, p1, q1);
                mine_[k + 2] = fma(-l, p2, r2); mine_[k + 3] = fma(-l, p3, r3);
                mine_[k + 4] = fma(-l, p4, r4); mine_[k + 5] = fma(-l, p5, r5);
                mine_[k + 6] = fma(-l, p6, r6); mine_[k + 7] = fma(-l, p7, r7);
            }
#pragma unroll 1
            for (; k + 3 <= Ns; k += 4) {
                const double p0 = prow[k], p1 = prow[k + 1], p2 = prow[k + 2], p3 = prow[k + 3];
                const double r0 = mine_[k], q1 = mine_[k + 1], r2 = mine_[k + 2], r3 = mine_[k + 3];
                mine_[k] = fma(-l, p0, r0); mine_[k + 1] = fma(-l, p1, q1);
                mine_[k + 2] = fma(-l, p2, r2); mine_[k + 3] = fma(-l, p3, r3);
            }
#pragma unroll 1
            for (; k <= Ns; ++k) mine_[k] = fma(-l, prow[k], mine_[k]);
            mine_[j + 1] = nxt ? rc : r1;
        }
    }
    // accepted when no multiplier grew (NaN fails the comparison as well); barrier inside
    if (g.any(cx->O.pivoting == 2 || grown, 2)) return false;
    if (t == 0) { reinterpret_cast<int*>(WS(CNT))[CNT_LU]++; reinterpret_cast<int*>(WS(CNT))[CNT_DIAG] = 1; }
    g.sync();
    return true;
}

// Build the matrix from the sparse Jacobian and factorise it: LUDecompose (A.10) with implicit row scaling and
// partial pivoting. Rows never move: `pos` is the index the row would have in the reference's physically swapped
// matrix (drives its tie rule: the LAST maximal row wins), STEP[t] the column at which row t became the pivot row.
// With lbchem_options.pivoting != 0 the diagonal-pivot factorisation above is tried first.
template <bool DAC>
static __device__ __noinline__ int lu_factor(const Ctx* cx, double* base, double dg) {
    if (cx->O.pivoting && lu_factor_diag<DAC>(cx, base, dg)) return 0;
    const Grp g(base);
    const int nSp = cx->S.nSp, t = g.tid, N = cx->nEqn;
    const Compact<DAC> cp(cx, base);
    const int Ns = cp.Ns, ldc = cp.ldc;
    const bool mine = t < nSp && cp.kept(t);
    const int cm = mine ? cp.index(t) : 0;
    double* A = WS(A);
    int* STEP = reinterpret_cast<int*>(WS(STEP));
    int* SPOF = STEP + N + (N & 1);               // DAC: species of compact column j
    double* XCH = WS(XCH);
    double* row = A + cm * ldc;
    g.sync();
    const double big = lu_build<DAC, true>(cx, base, cp, row, SPOF, mine, cm, dg);
    if (g.any(t < N && big == 0.0, 2)) return LBCHEM_CELL_SINGULAR;
    const double vv = 1.0 / big;
    bool done = !mine;
    int pos = cm;
    if (done && t < N) STEP[t] = -1;              // T, p and removed species: diagonal rows, solved in closed form
    if (t == 0) WS(XS)[1] = 1.0 / dg;
    int par = 1;      // exchange buffer, toggled per pivot column
#pragma unroll 1
    for (int j = 0; j < Ns; ++j) {
        par ^= 1;
        const double piv = done ? 0.0 : row[j];
        const unsigned long long key = done ? 0ull : (unsigned long long)__double_as_longlong(vv * fabs(piv)) + 1ull;
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        // one reduction on the top 32 bits of the key decides the warp's candidate whenever that maximum is unique
        // (the common case); ties in the top half fall through to the exact 64-bit + position comparison
        const unsigned mh = __reduce_max_sync(g.mask, hi);
        const bool c1 = hi == mh;
        bool win = c1;
        if (__popc(__ballot_sync(g.mask, c1)) != 1) {
            const unsigned ml = __reduce_max_sync(g.mask, c1 ? lo : 0u);
            const bool c2 = c1 && lo == ml;
            const unsigned pp = c2 ? (unsigned)pos + 1u : 0u;
            const unsigned mp = __reduce_max_sync(g.mask, pp);
            win = c2 && pp == mp;
        }
        if (win) {
            double* slot = XCH + (g.warp * 2 + par) * 4;
            slot[0] = __longlong_as_double((long long)key);
            reinterpret_cast<int*>(slot + 1)[0] = pos;
            reinterpret_cast<int*>(slot + 1)[1] = cm;
            slot[2] = 1.0 / ((piv == 0.0) ? LB_SMALL : piv);
        }
        g.sync();
        // both warps' candidates are read in one go (no second, dependent round trip through shared memory)
        const double* s0 = XCH + (0 * 2 + par) * 4;
        const double* s1 = XCH + (1 * 2 + par) * 4;
        const unsigned long long k0 = (unsigned long long)__double_as_longlong(s0[0]);
        const int2 i0 = *reinterpret_cast<const int2*>(s0 + 1);
        const double rd0 = s0[2];
        int ppos = i0.x, prw = i0.y; double rd = rd0;
        if (Grp::kTwoWarps) {
            const unsigned long long k1 = (unsigned long long)__double_as_longlong(s1[0]);
            const int2 i1 = *reinterpret_cast<const int2*>(s1 + 1);
            const double rd1 = s1[2];
            if (k1 > k0 || (k1 == k0 && i1.x > i0.x)) { ppos = i1.x; prw = i1.y; rd = rd1; }
        }
        if (!done && cm == prw) {
            done = true; STEP[t] = j;
            row[j] = rd;          // the diagonal slot of the pivot row keeps 1 / pivot (1 / SMALL for a zero pivot): all the solves need
        } else if (!done) {
            if (pos == j) pos = ppos;
            const double l = piv * rd;
            row[j] = l;
            const double* __restrict__ prow = A + prw * ldc;
            double* __restrict__ mine_ = row;
            int k = j + 1;
            // batches of eight, then four: all loads first, so the stores of one batch never wait behind its loads
#pragma unroll 1
            for (; k + 7 <= Ns; k += 8) {
                const double p0 = prow[k], p1 = prow[k + 1], p2 = prow[k + 2], p3 = prow[k + 3];
                const double p4 = prow[k + 4], p5 = prow[k + 5], p6 = prow[k + 6], p7 = prow[k + 7];
                const double r0 = mine_[k], r1 = mine_[k + 1], r2 = mine_[k + 2], r3 = mine_[k + 3];
                const double r4 = mine_[k + 4], r5 = mine_[k + 5], r6 = mine_[k + 6], r7 = mine_[k + 7];
                mine_[k] = fma(-l, p0, r0); mine_[k + 1] = fma(-l, p1, r1);
                mine_[k + 2] = fma(-l, p2, r2); mine_[k + 3] = fma(-l, p3, r3);
                mine_[k + 4] = fma(-l, p4, r4); mine_[k + 5] = fma(-l, p5, r5);
                mine_[k + 6] = fma(-l, p6, r6); mine_[k + 7] = fma(-l, p7, r7);
            }
#pragma unroll 1
            for (; k + 3 <= Ns; k += 4) {
                const double p0 = prow[k], p1 = prow[k + 1], p2 = prow[k + 2], p3 = prow[k + 3];
                const double r0 = mine_[k], r1 = mine_[k + 1], r2 = mine_[k + 2], r3 = mine_[k + 3];
                mine_[k] = fma(-l, p0, r0); mine_[k + 1] = fma(-l, p1, r1);
                mine_[k + 2] = fma(-l, p2, r2); mine_[k + 3] = fma(-l, p3, r3);
            }
#pragma unroll 1
            for (; k <= Ns; ++k) mine_[k] = fma(-l, prow[k], mine_[k]);
        }
    }
    if (t == 0) { reinterpret_cast<int*>(WS(CNT))[CNT_LU]++; reinterpret_cast<int*>(WS(CNT))[CNT_DIAG] = 0; }
    g.sync();
    return 0;
}

// x <- A^-1 x with the factors left by lu_factor (x: shared vector of nEqn, natural order in and out)
//
// Factors from the diagonal-pivot path (CNT_DIAG): row order = column order, so the substitutions stay inside a warp.
// The warp that owns the rows of columns [0, c0) runs them as shuffle steps (the owner lane broadcasts its finished
// value, no shared memory, no barrier), publishes its part once, and the other warp first folds those c0 columns into
// its rows as a run of FMAs and then does its own columns the same way; the backward pass mirrors it. Three group
// barriers per solve instead of one per column; the order of the operations on every row is that of the column loop
// below (ascending columns forward; temperature column, then descending columns backward).
template <bool DAC>
static __device__ __noinline__ void lu_solve(const Ctx* cx, double* base, double* x) {
    const Grp g(base);
    const int nSp = cx->S.nSp, t = g.tid, N = cx->nEqn;
    const Compact<DAC> cp(cx, base);
    const int Ns = cp.Ns;
    const bool sp = t < nSp && cp.kept(t);
    const int cm = sp ? cp.index(t) : 0;
    const double* row = WS(A) + cm * cp.ldc;
    double* BS = WS(BS);
    const int* SPOF = reinterpret_cast<const int*>(WS(STEP)) + N + (N & 1);
    g.sync();
    const int step = t < N ? reinterpret_cast<const int*>(WS(STEP))[t] : -1;
    const double rdg = WS(XS)[1];                   // 1 / dg: the diagonal rows (T, p, removed species)
    double bt = t < N ? x[t] : 0.0;
    if (t < nSp && !sp) x[t] = bt * rdg;            // removed species: decoupled diagonal row
    if (reinterpret_cast<const int*>(WS(CNT))[CNT_DIAG]) {
        constexpr int kW = TPC < 32 ? TPC : 32;                   // lanes of a group inside one warp
        constexpr bool two = Grp::kTwoWarps;
        const int lbase = g.lane & ~(kW - 1);
        // columns [0, c0) belong to the rows of the group's first warp, [c0, Ns) to the second
        const int c0 = two ? (nSp > 32 ? cp.index(32) : Ns) : Ns;
        if (t == nSp + 1) x[t] = bt * rdg;
        if (t == nSp) { const double v = bt * rdg; BS[Ns] = v; x[t] = v; }
        const int cf = sp ? cm : -1;                              // forward: rows below column j take part (cm > j)
        const int cb = sp ? cm : 0x7fffffff;                      // backward: rows above column j (cm < j)
#define LANE_OF(j) (lbase + ((DAC ? SPOF[j] : (j)) & (kW - 1)))
        // One substitution step = broadcast of the owner's value + one FMA (look-ahead loads and a pre-scaled backward chain
        // were measured: no gain, more instructions).
#define FWD_STEPS(j0, j1)                                                                                   \
        {                                                                                                   \
            _Pragma("unroll 1")                                                                             \
            for (int j = (j0); j < (j1); ++j) {                                                             \
                const double xj = __shfl_sync(g.mask, bt, LANE_OF(j));                                      \
                if (cf > j) bt = fma(-row[j], xj, bt);                                                      \
            }                                                                                               \
        }
        // folds the columns [j0, j1) another warp has finished (values in BS) into this row: four partial sums
#define FOLD(j0, j1)                                                                                        \
        {                                                                                                   \
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0; int j = (j0);                                    \
            _Pragma("unroll 1")                                                                             \
            for (; j + 3 < (j1); j += 4) {                                                                  \
                const double a0 = row[j], a1 = row[j + 1], a2 = row[j + 2], a3 = row[j + 3];                \
                const double b0 = BS[j], b1 = BS[j + 1], b2 = BS[j + 2], b3 = BS[j + 3];                    \
                s0 = fma(a0, b0, s0); s1 = fma(a1, b1, s1); s2 = fma(a2, b2, s2); s3 = fma(a3, b3, s3);     \
            }                                                                                               \
            _Pragma("unroll 1")                                                                             \
            for (; j < (j1); ++j) s0 = fma(row[j], BS[j], s0);                                              \
            bt -= (s0 + s1) + (s2 + s3);                                                                    \
        }
        // forward substitution, unit lower triangle
        if (!two || g.warp == 0) {
            FWD_STEPS(0, c0)
            if (two && sp) BS[cm] = bt;
        }
        g.sync();
        if (two && g.warp == 1) {
            if (sp) FOLD(0, c0)
            FWD_STEPS(c0, Ns)
        }
        // backward substitution: the temperature column first (BS[Ns] was published before the barrier above), then the
        // species columns in descending order; the owner of column j broadcasts x_j = b_j / pivot_j
        if (sp) bt = fma(-row[Ns], BS[Ns], bt);
        double s = 0.0;
#define BWD_STEPS(j1, j0)                                                                                   \
        {                                                                                                   \
            _Pragma("unroll 1")                                                                             \
            for (int j = (j1) - 1; j >= (j0); --j) {                                                        \
                const double xj = __shfl_sync(g.mask, bt * row[j], LANE_OF(j));                             \
                if (cb < j) bt = fma(-row[j], xj, bt);                                                      \
                if (cb == j) s = xj;                                                                        \
            }                                                                                               \
        }
        if (!two) { BWD_STEPS(Ns, 0) }
        else {
            if (g.warp == 1) {
                BWD_STEPS(Ns, c0)
                if (sp) BS[cm] = s;
            }
            g.sync();
            if (g.warp == 0) {
                if (sp) { FOLD(c0, Ns) }
                BWD_STEPS(c0, 0)
            }
        }
#undef LANE_OF
#undef FWD_STEPS
#undef BWD_STEPS
#undef FOLD
        if (sp) x[t] = s;
        if (t == 0) reinterpret_cast<int*>(WS(CNT))[CNT_SOLVE]++;
        g.sync();
        return;
    }
    // forward substitution, unit lower triangle (the T and p rows have no sub-diagonal entries)
#pragma unroll 1
    for (int j = 0; j < Ns; ++j) {
        if (step == j) BS[j] = bt;
        g.sync();
        if (sp && step > j) bt = fma(-row[j], BS[j], bt);
    }
    // backward substitution: p and T first (diagonal rows), then the species columns in descending order
    if (t == nSp + 1) x[t] = bt * rdg;
    if (t == nSp) { const double v = bt * rdg; BS[Ns] = v; x[t] = v; }
    g.sync();
    if (sp) bt = fma(-row[Ns], BS[Ns], bt);
#pragma unroll 1
    for (int j = Ns - 1; j >= 0; --j) {
        if (step == j) { const double v = bt * row[j]; BS[j] = v; x[DAC ? SPOF[j] : j] = v; }
        g.sync();
        if (sp && step < j) bt = fma(-row[j], BS[j], bt);
    }
    if (t == 0) reinterpret_cast<int*>(WS(CNT))[CNT_SOLVE]++;
    g.sync();
}


// ------------------------------------------------------------------------------------------------
// DAC::reduceMechanism (automaticSIS off) for the cell whose state is y: fills the disabled-reaction bits and the
// active-species word of MSK and returns NsDAC_. Rates of ALL reactions at (c, T) -> per species A the numerators of
// r_AB (signed sums in reaction order), P_A, C_A -> breadth-first search from the initial set by warp 0 (a species is
// queued once, in the reference's order: neighbours in first-appearance order) with R_B = max(R_B, R_u r_uB).
// Uses the matrix region for the numerators and V1..V4 as scratch: call before the integration starts.
// ------------------------------------------------------------------------------------------------
static __device__ __noinline__ int dac_reduce(const Ctx* cx, double* base, const double* y) {
    const Grp g(base);
    const SMech& S = cx->S;
    const int nSp = S.nSp, nR = S.nR, t = g.tid;
    const int* D = cx->dac;
    unsigned long long* MSK = reinterpret_cast<unsigned long long*>(WS(MSK));
    const int nW = (nR + 63) >> 6;
    g.sync();
    if (t <= nW) MSK[t] = 0ull;                       // every reaction takes part in the rate evaluation
    if (t == 0) reinterpret_cast<int*>(WS(XS))[0] = -1;      // no list: the sweep below visits every reaction
    const double T = y[nSp];
    const double lnT = log(T), invT = 1.0 / T;
    prepare(g, cx, base, y, T, lnT, true);
    reaction_sweep<true>(cx, base, 0, T, lnT, invT);
    const double* OM = WS(OM);
    double* NUM = WS(A);
    unsigned char* NB = reinterpret_cast<unsigned char*>(NUM + D[0]);     // neighbour B of each pair, after the numerators
    double *PA = VEC(1), *CA = VEC(2), *RV = VEC(3);
    int* pairPtr = reinterpret_cast<int*>(VEC(4));                        // nSp + 1 ints; the search queue follows
    constexpr int PF = 8;                                                  // program words fetched ahead (L2 round trips)
    {
        // pair numerators: thread t owns a contiguous run of pairs, balanced by term count
        const unsigned* term = reinterpret_cast<const unsigned*>(D + D[3]);
        const int k0 = (D + D[2])[t], k1 = (D + D[2])[t + 1];
        int e = (D + D[6])[t];
        double acc = 0.0;
#pragma unroll 1
        for (int k = k0; k < k1; k += PF) {
            unsigned w[PF];
#pragma unroll
            for (int q = 0; q < PF; ++q) w[q] = term[k + q];               // padded by PF words on the host
#pragma unroll
            for (int q = 0; q < PF; ++q) {
                if (k + q < k1) {
                    acc = __dadd_rn(acc, __dmul_rn((double)(int)(signed char)(w[q] >> 16), OM[w[q] & 0xffffu]));
                    if ((w[q] >> 24) & 1u) { NUM[e] = acc; NB[e] = (unsigned char)((w[q] >> 25) & 63u); ++e; acc = 0.0; }
                }
            }
        }
    }
    if (t <= nSp) pairPtr[t] = (D + D[1])[t];
    if (t < nSp) {
        const int* pcPtr = D + D[4]; const unsigned* pc = reinterpret_cast<const unsigned*>(D + D[5]);
        double pa = 0.0, ca = 0.0;
        const int q1 = pcPtr[t + 1];
#pragma unroll 1
        for (int k = pcPtr[t]; k < q1; k += PF) {
            unsigned w[PF];
#pragma unroll
            for (int q = 0; q < PF; ++q) w[q] = pc[k + q];
#pragma unroll
            for (int q = 0; q < PF; ++q) {
                if (k + q < q1) {
                    const double wA = __dmul_rn((double)(int)(signed char)(w[q] >> 16), OM[w[q] & 0xffffu]);
                    if (wA > 0.0) pa += wA; else ca += -wA;
                }
            }
        }
        PA[t] = pa; CA[t] = ca; RV[t] = 0.0;
    }
    g.sync();
    // r_AB = min(|num| / max(P_A, C_A), 1) for every pair of species A = t, ahead of the serial search (0 when the
    // denominator is below VSMALL: the reference does not look at the neighbours of such a species)
    if (t < nSp) {
        const double Den = fmax(PA[t], CA[t]);
        const int e1 = pairPtr[t + 1];
#pragma unroll 2
        for (int e = pairPtr[t]; e < e1; ++e) {
            double r = fabs(NUM[e]) / Den;
            if (r > 1.0) r = 1.0;
            NUM[e] = (Den > LB_VSMALL) ? r : 0.0;
        }
    }
    g.sync();
    if (g.warp == 0) {
        unsigned char* Q = reinterpret_cast<unsigned char*>(pairPtr + ((nSp + 2) & ~1));
        const double tol = __hiloint2double(D[17], D[16]);
        unsigned long long act = 0ull;
        int head = 0, tail = 0;
        const int nInit = D[7];
        for (int i = 0; i < nInit; ++i) {
            const int q = D[8 + i];
            if (!((act >> q) & 1ull)) {
                act |= 1ull << q;
                if (g.lane == 0) { Q[tail] = (unsigned char)q; RV[q] = 1.0; }
                ++tail;
            }
        }
        __syncwarp();
#pragma unroll 1
        while (head < tail) {
            const int u = Q[head++];
            {
                const int e0 = pairPtr[u], n = pairPtr[u + 1] - e0;
                const double Ru = RV[u];
#pragma unroll 1
                for (int v0 = 0; v0 < n; v0 += 32) {
                    const int v = v0 + g.lane;
                    bool cond = false; int Bsp = 0; double Rt = 0.0;
                    if (v < n) {
                        Bsp = NB[e0 + v];
                        Rt = Ru * NUM[e0 + v];
                        cond = Rt >= tol;
                    }
                    if (cond && Rt > RV[Bsp]) RV[Bsp] = Rt;
                    const bool nw = cond && !((act >> Bsp) & 1ull);
                    const unsigned m = __ballot_sync(0xffffffffu, nw);
                    if (nw) Q[tail + __popc(m & ((1u << g.lane) - 1u))] = (unsigned char)Bsp;
                    tail += __popc(m);
                    const unsigned lo = __reduce_or_sync(0xffffffffu, (nw && Bsp < 32) ? 1u << Bsp : 0u);
                    const unsigned hi = __reduce_or_sync(0xffffffffu, (nw && Bsp >= 32) ? 1u << (Bsp - 32) : 0u);
                    act |= ((unsigned long long)hi << 32) | lo;
                    __syncwarp();
                }
            }
        }
        if (g.lane == 0) MSK[nW] = act;
    }
    g.sync();
    const unsigned long long act = MSK[nW];
    g.sync();
#pragma unroll 1
    for (int c = 0; c < nW; ++c) {
        const int p = c * 64 + t;
        bool dis = false;
        if (p < nR) {
            const unsigned iw = S.info[p];
            const unsigned long long lp = S.lpk[p], rp = S.rpk[p];
            for (unsigned s = 0; s < V2_NL(iw); ++s) dis = dis || !((act >> V2_SP(V2_SLOT(lp, s))) & 1ull);
            for (unsigned s = 0; s < V2_NR(iw); ++s) dis = dis || !((act >> V2_SP(V2_SLOT(rp, s))) & 1ull);
        }
        const unsigned b = __ballot_sync(0xffffffffu, dis);
        if (g.lane == 0) reinterpret_cast<unsigned*>(MSK)[2 * c + g.warp] = b;
        if (p < nR && dis) WS(OM)[p] = 0.0;           // stays zero: the sweeps only visit enabled reactions
    }
    g.sync();
    // compacted list of the enabled reactions, ascending packed position: thread c lists the reactions of mask word c
    if (__popcll(act) == nSp) {
        if (t == 0) reinterpret_cast<int*>(WS(XS))[0] = -1;  // nothing removed: every reaction stays, no list (the full matrix owns the tail)
    } else if (t < nW) {
        int basep = 0;
        for (int c = 0; c < t; ++c) { const int hiBit = nR - c * 64 < 64 ? nR - c * 64 : 64; basep += hiBit - __popcll(MSK[c]); }
        const int lim = nR - t * 64 < 64 ? nR - t * 64 : 64;
        unsigned long long en = ~MSK[t];
        if (lim < 64) en &= (1ull << lim) - 1ull;
        unsigned short* RLw = reinterpret_cast<unsigned short*>(WS(RL) + 1);
        while (en) { const int bpos = __ffsll((long long)en) - 1; en &= en - 1ull; RLw[basep++] = (unsigned short)(t * 64 + bpos); }
        if (t == nW - 1) reinterpret_cast<int*>(WS(XS))[0] = basep;
    }
    g.sync();
    return __popcll(act);
}

// ------------------------------------------------------------------------------------------------
// integrators (control flow identical to lbchem_kernels_v2.cu; only the heavy routines are calls)
// ------------------------------------------------------------------------------------------------


#define CNT(i) (reinterpret_cast<int*>(WS(CNT))[i])

// one Rosenbrock34 attempt (A.4) from Y0=V0 with step dx; result in V1. The sparse Jacobian (and dfdx in V3)
// of the step start are re-used after a rejected attempt: the reference recomputes identical values.
template <bool DAC>
static __device__ __forceinline__ double rosenbrock34_step(const Grp& g, const Ctx* cx, double* base, Cell& cs, double dx) {
    const double a21 = 2, a31 = 48.0 / 25.0, a32 = 6.0 / 25.0, c21 = -8, c31 = 372.0 / 25.0, c32 = 12.0 / 5.0,
                 c41 = -112.0 / 125.0, c42 = -54.0 / 125.0, c43 = -2.0 / 5.0, b1 = 19.0 / 9.0, b2 = 1.0 / 2.0,
                 b3 = 25.0 / 108.0, b4 = 125.0 / 108.0, e1 = 34.0 / 108.0, e2 = 7.0 / 36.0, e3 = 0, e4 = 125.0 / 108.0,
                 gamma = 1.0 / 2.0, c1X = 1.0 / 2.0, c2X = -3.0 / 2.0, c3X = 121.0 / 50.0, c4X = 29.0 / 250.0;
    const int t = g.tid, N = cx->nEqn;
    double *Y = VEC(0), *K1 = VEC(1), *K2 = VEC(2), *K3 = VEC(3), *K4 = VEC(4);
    double* DY = K4;      // derivatives() output of stages 2-3; stage 4 turns it into k4's right-hand side element by element, in place
    const double *Y0 = GS_GV(0), *D0 = GS_GV(1), *DX = GS_GV(2);     // read by their owner thread only
    const int st = lu_factor<DAC>(cx, base, 1.0 / (gamma * dx));
    if (st) { cs.fail = st; return 0.0; }
    if (t == 0) CNT(CNT_JAC)++;    // algorithmic count: the reference evaluates the Jacobian in every attempt
#pragma unroll 1
    for (int s = 0; s < 4; ++s) {
        if (s == 1 || s == 2) {
            if (t < N) Y[t] = (s == 1) ? Y0[t] + a21 * K1[t] : Y0[t] + a31 * K1[t] + a32 * K2[t];
            derivatives<DAC>(cx, base, Y, DY);
        }
        double* K = VEC(1 + s);
        if (t < N) {
            double r;
            if (s == 0) r = D0[t] + dx * c1X * DX[t];
            else if (s == 1) r = DY[t] + dx * c2X * DX[t] + c21 * K1[t] / dx;
            else if (s == 2) r = DY[t] + dx * c3X * DX[t] + (c31 * K1[t] + c32 * K2[t]) / dx;
            else r = DY[t] + dx * c4X * DX[t] + (c41 * K1[t] + c42 * K2[t] + c43 * K3[t]) / dx;
            K[t] = r;
        }
        lu_solve<DAC>(cx, base, K);
    }
    double em = 0.0;
    if (t < N) {
        const double yi = Y0[t] + b1 * K1[t] + b2 * K2[t] + b3 * K3[t] + b4 * K4[t];
        const double er = e1 * K1[t] + e2 * K2[t] + e3 * K3[t] + e4 * K4[t];
        Y[t] = yi;
        em = fabs(er) / (cx->O.absTol + cx->O.relTol * fmax(fabs(Y0[t]), fabs(yi)));
    }
    return g.maxnan(em, 0);
}

// adaptiveSolver::solve (A.3)
template <bool DAC>
static __device__ __forceinline__ void adaptive_solve(const Grp& g, const Ctx* cx, double* base, Cell& cs, double& x, double& dxTry) {
    const double safeScale = 0.9, alphaInc = 0.2, alphaDec = 0.25, minScale = 0.2, maxScale = 10;
    const int t = g.tid, N = cx->nEqn;
    double dx = dxTry, err = 0.0;
    // dydx0 and the Jacobian are both evaluated at (y0, T0)
    jacobian<DAC>(cx, base, GS_GV(0), GS_GV(2), GS_GV(1));
    if (t == 0) CNT(CNT_JAC)--;                 // counted per attempt in rosenbrock34_step
    do {
        if (t == 0) CNT(CNT_ATTEMPTS)++;
        err = rosenbrock34_step<DAC>(g, cx, base, cs, dx);
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
    if (t < N) GS_GV(0)[t] = VEC(0)[t];
    g.sync();
    if (err > pow(maxScale / safeScale, -1.0 / alphaInc))
        dxTry = fmin(fmax(safeScale * pow(err, -alphaInc), minScale), maxScale) * dx;
    else
        dxTry = safeScale * maxScale * dx;
}

// seulex::seul (A.5): one extrapolation column; y0 = V4, scale = V5, result in V6.
// `phase` keeps derivatives() and lu_solve() at one call site each: phase 1 = the stability check of the
// second sub-step (solves dyc), phase 0 = the regular Euler increment (solves dy).
template <bool DAC>
static __device__ __forceinline__ bool seul(const Grp& g, const Ctx* cx, double* base, Cell& cs, double dxTot, int k, double& yOut) {
    const int t = g.tid, N = cx->nEqn;
    const int nSteps = c_nSeq[k];
    const double dx = dxTot / nSteps;
    double *dy = VEC(3), *yT = VEC(1), *f = VEC(4), *dyc = VEC(2);
    double *f0 = GS_GV(0), *y0 = GS_GV(1), *scale = GS_GV(2);         // each element is read and written by its owner thread only
    const int st = lu_factor<DAC>(cx, base, 1 / dx);
    if (st) { cs.fail = st; return false; }
    bool ok = true;
    int it = 0, phase = 0;
    double dy1 = 0.0;
#pragma unroll 1
    while (it < nSteps) {
        if (phase == 0) {
            if (it == 1) { if (t < N) yT[t] = y0[t] + dy[t]; }
            else if (it > 1) { if (t < N) yT[t] += dy[t]; }
            if (it == 0 && cs.f0valid) {
                // f(y0) is the same for every column and every retry of this step: the reference re-evaluates it,
                // we keep it in V3 (counted as an evaluation all the same)
                if (t < N) f[t] = f0[t];
                if (t == 0) CNT(CNT_RHS)++;
            } else {
                derivatives<DAC>(cx, base, it == 0 ? y0 : yT, f);
                if (it == 0) { if (t < N) f0[t] = f[t]; cs.f0valid = 1; }
            }
            if (it == 1 && k <= 1) phase = 1;
        }
        double* rhs;
        if (phase == 1) {
            double s1 = 0.0;
            if (t < N) { const double q = dy[t] / scale[t]; s1 = q * q; dyc[t] = f[t] - dy[t] / dx; }
            dy1 = sqrt(g.sum(s1, 2));
            rhs = dyc;
        } else {
            if (t < N) dy[t] = f[t];
            rhs = dy;
        }
        lu_solve<DAC>(cx, base, rhs);
        if (phase == 1) {
            const double denom = fmin(1.0, dy1 + LB_SMALL);
            double s2 = 0.0; int big = 0;
            if (t < N) {
                if (fabs(dyc[t]) > scale[t] * denom) big = 1;
                const double q = dyc[t] / scale[t];
                s2 = q * q;
            }
            if (g.any(big, 3)) { cs.theta = 1; ok = false; break; }
            const double dy2 = sqrt(g.sum(s2, 2));
            cs.theta = dy2 / denom;
            if (cs.theta > 1) { ok = false; break; }
            if (t == 0) CNT(CNT_RHS)++;    // the reference evaluates derivatives(yT) a second time (same value)
            phase = 2;                     // next pass: regular increment of the same sub-step, f still current
        } else {
            phase = 0;
            ++it;
        }
    }
    if (ok && t < N) yOut = yT[t] + dy[t];
    g.sync();
    return ok;
}

static __device__ __forceinline__ double seulex_coeff(int k, int l) {
    const double ratio = (double)c_nSeq[k] / c_nSeq[l];
    return 1 / (ratio - 1);
}

// seulex::solve (A.5); y = V0
template <bool DAC>
static __device__ __forceinline__ void seulex_solve(const Grp& g, const Ctx* cx, double* base, Cell& cs, double& x, StepState& step, int& attempts) {
    const int t = g.tid, N = cx->nEqn;
    const DevOpts& O = cx->O;
    const int kMaxx = 11;
    const double stepFactor1 = 0.6, stepFactor2 = 0.93, stepFactor3 = 0.1, stepFactor4 = 4, stepFactor5 = 0.5,
                 kFactor1 = 0.7, kFactor2 = 0.9;
    const double absTol = O.absTol, relTol = O.relTol;
    const double jacRedo = fmin(1e-4, relTol);
    double *y = VEC(0), *y0 = GS_GV(1), *scale = GS_GV(2), *tb = GS_TB;
    // temp / dxOpt are group-uniform: one copy in the workspace (every thread writes the same value; reads follow a group
    // barrier of the column loop), not 48 registers per thread
    double* temp = WS(SX); double* dxOpt = temp + 12;
    double dx = step.dxTry;
    g.sync();
    if (t < N) { y0[t] = y[t]; scale[t] = absTol + relTol * fabs(y[t]); }
    cs.f0valid = 0;
    temp[0] = LB_GREAT; dxOpt[0] = fabs(0.1 * dx);
    if (step.first || step.prevReject) cs.theta = 2.0 * jacRedo;
    if (step.first) {
        const double logTol = -log10(relTol + absTol) * 0.6 + 0.5;
        int kt = (int)logTol;
        if (kt > kMaxx - 1) kt = kMaxx - 1;
        if (kt < 1) kt = 1;
        cs.kTarg = kt;
    }
    bool jacUpdated = false;
    if (cs.theta > jacRedo) {
        // the Jacobian's first rate sweep is at (y0, T0): it also yields f(y0) for the first Euler sub-step of every column
        jacobian<DAC>(cx, base, y, nullptr, GS_GV(0));
        if (t == 0) CNT(CNT_RHS)--;            // counted when seul() consumes it
        cs.f0valid = 1;
        jacUpdated = true;
    }
    int k = 0;
    double dxNew = fabs(dx);
    bool firstk = true;
    while (firstk || step.reject) {
        dx = step.forward ? dxNew : -dxNew;
        firstk = false;
        step.reject = 0;
        attempts++;
        if (t == 0) CNT(CNT_ATTEMPTS)++;
        if (fabs(dx) < LB_VSMALL || attempts > 100000) { cs.fail = LBCHEM_CELL_UNDERFLOW; return; }
        double errOld = 0;
#pragma unroll 1
        for (k = 0; k <= cs.kTarg + 1; ++k) {
            double ySeqT = 0.0;                      // this thread's element of the column result
            const bool success = seul<DAC>(g, cx, base, cs, dx, k, ySeqT);
            if (cs.fail) return;
            if (!success) { step.reject = 1; dxNew = fabs(dx) * stepFactor5; break; }
            double s = 0.0;
            if (t < N) {
                if (k == 0) y[t] = ySeqT;
                else {
                    tb[(k - 1) * N + t] = ySeqT;
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
                const double err = sqrt(g.sum(s, 1) / CNT(CNT_NODE));
                if (!(err == err)) { cs.fail = LBCHEM_CELL_NAN; return; }
                if (err > 1 / LB_SMALL || (k > 1 && err >= errOld)) { step.reject = 1; dxNew = fabs(dx) * stepFactor5; break; }
                errOld = O.seulexErrOldMax ? fmax(4 * err, 1.0) : fmin(4 * err, 1.0);   // doubtful OpenFOAM detail: a switch
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
                if (cs.theta > jacRedo && !jacUpdated) { jacobian<DAC>(cx, base, y, nullptr, nullptr); jacUpdated = true; }
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
template <int SOLVER, bool DAC>
static __device__ __forceinline__ void ode_solve(const Grp& g, const Ctx* cx, double* base, Cell& cs, double xStart, double xEnd, double& dxTry) {
    StepState step = {1, 1, 0, 0, 0, dxTry, 0.0};
    double x = xStart;
    int attempts = 0;
#pragma unroll 1
    for (int nStep = 0; nStep < cx->O.maxSteps; ++nStep) {
        const double dxTry0 = step.dxTry;
        step.reject = 0;
        if ((x + step.dxTry - xEnd) * (x + step.dxTry - xStart) > 0) {
            step.last = 1;
            step.dxTry = xEnd - x;
        }
        if (SOLVER == LBCHEM_SOLVER_SEULEX) {
            seulex_solve<DAC>(g, cx, base, cs, x, step, attempts);
        } else {
            const double x0 = x;
            adaptive_solve<DAC>(g, cx, base, cs, x, step.dxTry);
            step.dxDid = x - x0;
        }
        if (cs.fail) return;
        if (g.tid == 0) CNT(CNT_STEPS)++;
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



template <int SOLVER, bool DAC>
static __device__ __forceinline__ void integrate_body(const V2Mech& M, const DevOpts& O, const BatchArgs& B, int wsDoubles) {
    extern __shared__ double smem[];
    unsigned char* sm = reinterpret_cast<unsigned char*>(smem);
    {
        const unsigned long long* src = reinterpret_cast<const unsigned long long*>(M.blob);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(sm + CTX_BYTES);
        for (int i = threadIdx.x; i < M.hotBytes / 8; i += blockDim.x) dst[i] = src[i];
    }
    Ctx* cxw = reinterpret_cast<Ctx*>(sm);
    if (threadIdx.x == 0) {
        bind(cxw->S, M, sm + CTX_BYTES);
        bind_cold(cxw->S, M);
        cxw->O = O; cxw->O.solver = SOLVER;
        cxw->nEqn = M.nEqn; cxw->nEnt = M.nEnt; cxw->nFall = M.nFall; cxw->tbRows = M.tbRows;
        cxw->wsDoubles = ctx_layout(*cxw, M);
        cxw->jterm = M.jterm; cxw->dterm = M.dterm; cxw->nD = M.nD; cxw->jthrEnt = M.jthrEnt; cxw->jthrN = M.jthrN; cxw->dac = M.dac;
        cxw->scratch = B.scratch; cxw->scratchStride = B.scratchStride; cxw->groups = blockDim.x / TPC;
    }
    __syncthreads();
    const Ctx* cx = cxw;
    double* base = reinterpret_cast<double*>(sm + CTX_BYTES + M.hotBytes) + (size_t)(threadIdx.x / TPC) * wsDoubles;
    const Grp g(base);
    const SMech& S = cx->S;
    const int nSp = S.nSp, t = g.tid;
    unsigned long long st[8] = {0, 0, 0, 0, 0, 0, 0, 0};

    for (;;) {
        g.sync();
        if (t == 0) reinterpret_cast<unsigned long long*>(WS(XS))[0] = atomicAdd(B.queue, 1ull);
        g.sync();
        const unsigned long long q = reinterpret_cast<unsigned long long*>(WS(XS))[0];
        if (q >= (unsigned long long)(B.a.n + B.b.n)) break;
        // queue position -> (segment, cell): own cells first, then the cells received from other ranks
        // (reference LoadBalancedChemistryModel.C:608-611), all through the one queue of this launch
        long long cell = B.order ? (long long)B.order[q] : (long long)q;
        const bool second = cell >= B.a.n;
        if (second) cell -= B.a.n;
#define SEG(f) (second ? B.b.f : B.a.f)
        { const int* ix = SEG(index); if (ix) cell = ix[cell]; }     // ISAT miss list: only the listed cells are integrated
        const double* __restrict__ Yp = SEG(Y);
        double* __restrict__ RRp = SEG(RR);
        const long long ldy = SEG(ldy);
        const double T0 = SEG(T)[cell];
        const double p0 = SEG(p)[cell];
        const double rho = SEG(rho)[cell];
        const double* dTp = SEG(deltaT);
        const double deltaT = dTp ? dTp[cell] : B.deltaTUniform;
        double* dtcp = SEG(deltaTChem);
        double dtChem = dtcp[cell];
        if (t < CNT_N) CNT(t) = 0;
        Cell cs; cs.fail = 0; cs.theta = 0.0; cs.kTarg = 1; cs.f0valid = 0;
        if (t == 0) CNT(CNT_NODE) = cx->nEqn;
        int status;
        double Tend = T0;
        double* y = (SOLVER == LBCHEM_SOLVER_SEULEX) ? VEC(0) : GS_GV(0);      // the state vector the solver advances
        if (T0 > O.Treact) {
            if (t < nSp) y[t] = rho * Yp[(size_t)t * ldy + cell] / S.W[t];
            if (t == nSp) y[nSp] = T0;
            if (t == nSp + 1) y[nSp + 1] = p0;
            g.sync();
            if (DAC) {
                // LoadBalancedTDAC solveCell with mechRed()->active() (reference LoadBalancedTDACChemistryModel.C:385-393)
                const int ns = dac_reduce(cx, base, y);
                if (t == 0) { CNT(CNT_NODE) = ns + 2; int* na = SEG(nActive); if (na) na[cell] = ns; }
            }
            double timeLeft = deltaT;
            while (timeLeft > LB_SMALL) {
                double dt = timeLeft;
                ode_solve<SOLVER, DAC>(g, cx, base, cs, 0.0, dt, dtChem);
                g.sync();
                if (t < nSp) y[t] = fmax(0.0, y[t]);
                g.sync();
                if (cs.fail) break;
                timeLeft -= dt;
            }
            dtChem = fmin(dtChem, O.deltaTChemMax);
            if (t < nSp) {
                const double c0 = rho * Yp[(size_t)t * ldy + cell] / S.W[t];
                RRp[(size_t)t * ldy + cell] = (y[t] - c0) * S.W[t] / deltaT;
            }
            // the reduced call of the reference passes cData.T() for both T and p (:406-409): the aliased reference ends up
            // holding the p slot, i.e. the temperature the cell came in with
            Tend = DAC ? T0 : y[nSp];
            status = cs.fail ? cs.fail : LBCHEM_CELL_OK;
        } else {
            if (t < nSp) RRp[(size_t)t * ldy + cell] = 0.0;
            status = LBCHEM_CELL_SKIPPED;
        }
        if (t == 0) {
            dtcp[cell] = dtChem;
            double* To = SEG(Tout); double* co = SEG(cost); int* so = SEG(status);
            if (To) To[cell] = Tend;
            // per-cell cost (the reference stores wall-clock seconds, LoadBalancedChemistryModel.C:431-448): attempted
            // sub-steps, or integer work units weighted by what each building block costs on this kernel
            if (co) co[cell] = B.costModel ? (double)(9 * CNT(CNT_LU) + CNT(CNT_SOLVE) + 3 * CNT(CNT_RHS) + 12 * CNT(CNT_JAC))
                                           : (double)CNT(CNT_ATTEMPTS);
            if (so) so[cell] = status;
            st[0] += CNT(CNT_STEPS); st[1] += CNT(CNT_ATTEMPTS); st[2] += CNT(CNT_JAC); st[3] += CNT(CNT_LU);
            st[4] += CNT(CNT_SOLVE); st[5] += CNT(CNT_RHS); st[6] += (status >= LBCHEM_CELL_MAX_STEPS); st[7] += 1;
        }
#undef SEG
    }
    if (t == 0 && st[7]) {
#pragma unroll
        for (int i = 0; i < 8; ++i) atomicAdd(B.stats + i, st[i]);
    }
}


// ------------------------------------------------------------------------------------------------
// Chemical time scale of every cell (OpenFOAM StandardChemistryModel::tc()):
//   tc = nReaction * sum_i c_i / (SMALL + sum_r (sum_rhs nu) * kf_r * prod_lhs c^e)
// Same staging, workspace layout and rate sweep as the integration kernel; one cell per 64-thread group.
// ------------------------------------------------------------------------------------------------
static __device__ __forceinline__ void tc_body(const V2Mech& M, const DevOpts& O, const BatchArgs& B, int wsDoubles, double* tcOut) {
    extern __shared__ double smem[];
    unsigned char* sm = reinterpret_cast<unsigned char*>(smem);
    {
        const unsigned long long* src = reinterpret_cast<const unsigned long long*>(M.blob);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(sm + CTX_BYTES);
        for (int i = threadIdx.x; i < M.hotBytes / 8; i += blockDim.x) dst[i] = src[i];
    }
    Ctx* cxw = reinterpret_cast<Ctx*>(sm);
    if (threadIdx.x == 0) {
        bind(cxw->S, M, sm + CTX_BYTES);
        bind_cold(cxw->S, M);
        cxw->O = O;
        cxw->nEqn = M.nEqn; cxw->nEnt = M.nEnt; cxw->nFall = M.nFall; cxw->tbRows = M.tbRows;
        cxw->wsDoubles = ctx_layout(*cxw, M);
        cxw->jterm = M.jterm; cxw->dterm = M.dterm; cxw->nD = M.nD; cxw->jthrEnt = M.jthrEnt; cxw->jthrN = M.jthrN; cxw->dac = nullptr;
        cxw->scratch = B.scratch; cxw->scratchStride = B.scratchStride; cxw->groups = blockDim.x / TPC;
    }
    __syncthreads();
    const Ctx* cx = cxw;
    double* base = reinterpret_cast<double*>(sm + CTX_BYTES + M.hotBytes) + (size_t)(threadIdx.x / TPC) * wsDoubles;
    const Grp g(base);
    const SMech& S = cx->S;
    const int nSp = S.nSp, nR = S.nR, t = g.tid;
    const long long groups = (long long)gridDim.x * cx->groups;
    for (long long cell = (long long)blockIdx.x * cx->groups + threadIdx.x / TPC; cell < B.a.n; cell += groups) {
        g.sync();
        double* y = VEC(0);
        const double T = B.a.T[cell], rho = B.a.rho[cell];
        double c = 0.0;
        if (t < nSp) { c = rho * B.a.Y[(size_t)t * B.a.ldy + cell] / S.W[t]; y[t] = c; }
        g.sync();
        const double cSum = g.sum(c, 3);          // the reference sums the unclamped concentrations
        const double lnT = log(T), invT = 1.0 / T;
        prepare(g, cx, base, y, T, lnT, true);
        reaction_sweep<false>(cx, base, 1, T, lnT, invT);
        const double* KF = WS(A); const double* CC = WS(CC);
        double s = 0.0;
#pragma unroll 1
        for (int p = t; p < nR; p += TPC) {
            const unsigned iw = S.info[p];
            const unsigned long long rp = S.rpk[p];
            double nu = 0.0;
            for (unsigned q = 0; q < V2_NR(iw); ++q) nu += (double)V2_NU(V2_SLOT(rp, q));
            s = fma(nu, KF[p] * side_product(CC, S.lpk[p], V2_NL(iw)), s);
        }
        const double tot = g.sum(s, 3);
        if (t == 0) tcOut[cell] = nR * cSum / (LB_SMALL + tot);
    }
}

// ------------------------------------------------------------------------------------------------
// Test hook (lbchem_probe): the building blocks of THIS kernel -- prepare / reaction_sweep / derivatives / jacobian /
// lu_factor / lu_solve, the very device functions integrate_v3_kernel calls -- on single states, so that a parity failure
// can be localised. One state per 64-thread group. Layouts as in include/lbchem.h (kf / kr come out in packed reaction
// order; the host maps them back).
// ------------------------------------------------------------------------------------------------
static __device__ __forceinline__ void probe_body(const V2Mech& M, const DevOpts& O, const BatchArgs& B, int wsDoubles, int kind, int nItems,
                                                   const double* in, int inStride, double* out, int outStride) {
    extern __shared__ double smem[];
    unsigned char* sm = reinterpret_cast<unsigned char*>(smem);
    {
        const unsigned long long* src = reinterpret_cast<const unsigned long long*>(M.blob);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(sm + CTX_BYTES);
        for (int i = threadIdx.x; i < M.hotBytes / 8; i += blockDim.x) dst[i] = src[i];
    }
    Ctx* cxw = reinterpret_cast<Ctx*>(sm);
    if (threadIdx.x == 0) {
        bind(cxw->S, M, sm + CTX_BYTES);
        bind_cold(cxw->S, M);
        cxw->O = O;
        cxw->nEqn = M.nEqn; cxw->nEnt = M.nEnt; cxw->nFall = M.nFall; cxw->tbRows = M.tbRows;
        cxw->wsDoubles = ctx_layout(*cxw, M);
        cxw->jterm = M.jterm; cxw->dterm = M.dterm; cxw->nD = M.nD; cxw->jthrEnt = M.jthrEnt; cxw->jthrN = M.jthrN; cxw->dac = nullptr;
        cxw->scratch = B.scratch; cxw->scratchStride = B.scratchStride; cxw->groups = blockDim.x / TPC;
    }
    __syncthreads();
    const Ctx* cx = cxw;
    double* base = reinterpret_cast<double*>(sm + CTX_BYTES + M.hotBytes) + (size_t)(threadIdx.x / TPC) * wsDoubles;
    const Grp g(base);
    const SMech& S = cx->S;
    const int nSp = S.nSp, nR = S.nR, t = g.tid, N = cx->nEqn;
    const int groups = gridDim.x * cx->groups;
    for (int item = blockIdx.x * cx->groups + threadIdx.x / TPC; item < nItems; item += groups) {
        const double* x = in + (size_t)item * inStride;
        double* o = out + (size_t)item * outStride;
        g.sync();
        if (t < CNT_N) CNT(t) = 0;
        if (t == 0) CNT(CNT_NODE) = N;
        double* y = VEC(0);
        if (t < N) y[t] = x[t];
        g.sync();
        if (kind == LBCHEM_PROBE_DERIV) {
            derivatives<false>(cx, base, y, VEC(1));
            if (t < N) o[t] = VEC(1)[t];
        } else if (kind == LBCHEM_PROBE_RATES) {
            const double T = y[nSp], lnT = log(T), invT = 1.0 / T;
            prepare(g, cx, base, y, T, lnT, true);
            reaction_sweep<false>(cx, base, 1, T, lnT, invT);
            const double* KF = WS(A);
            for (int p = t; p < 2 * nR; p += TPC) o[p] = KF[p];
        } else {
            jacobian<false>(cx, base, y, VEC(2), nullptr);
            if (kind == LBCHEM_PROBE_JAC) {
                if (t < N) o[t] = VEC(2)[t];
                double* J = o + N;
                for (int k = t; k < N * N; k += TPC) J[k] = 0.0;
                g.sync();
                if (t < nSp) {
                    const double* JS = GS_JS;
                    for (int e = S.rowPtr[t]; e < S.rowPtr[t + 1]; ++e) J[t * N + S.entCol[e]] = JS[e];
                    J[t * N + nSp] = GS_JT[t];
                }
            } else {                                   // LBCHEM_PROBE_JAC_LUSOLVE: x = (dg I - J(y))^-1 b
                const double dg = x[N];
                const int st = lu_factor<false>(cx, base, dg);
                double* b = VEC(1);
                if (t < N) b[t] = x[N + 1 + t];
                lu_solve<false>(cx, base, b);
                if (t < N) o[t] = st ? __longlong_as_double(0x7ff8000000000000LL) : b[t];
            }
        }
        g.sync();
    }
}

};   // struct K<TPC>

template <int TPC, int SOLVER, bool DAC>
__global__ void __maxnreg__(Width<TPC>::kRegs) integrate_v3_kernel(V2Mech M, DevOpts O, BatchArgs B, int wsDoubles) {
    K<TPC>::template integrate_body<SOLVER, DAC>(M, O, B, wsDoubles);
}
template <int TPC>
__global__ void __launch_bounds__(Width<TPC>::kThreads, 1) tc_v3_kernel(V2Mech M, DevOpts O, BatchArgs B, int wsDoubles, double* tcOut) {
    K<TPC>::tc_body(M, O, B, wsDoubles, tcOut);
}
template <int TPC>
__global__ void __launch_bounds__(Width<TPC>::kThreads, 1) probe_v3_kernel(V2Mech M, DevOpts O, BatchArgs B, int wsDoubles, int kind, int nItems,
                                                          const double* in, int inStride, double* out, int outStride) {
    K<TPC>::probe_body(M, O, B, wsDoubles, kind, nItems, in, inStride, out, outStride);
}

}  // namespace v3

// ------------------------------------------------------------------------------------------------
// host side: workspace size (must match v3::ctx_layout), configuration, launch
// ------------------------------------------------------------------------------------------------
int lb_v3_workspace(V2Mech& M, int solver, double relTol, double absTol) {
    // the extrapolation table lives in the global scratch, always sized for the reference's full kMaxx = 11 rows
    (void)relTol; (void)absTol;
    M.tbRows = solver == LBCHEM_SOLVER_SEULEX ? 11 : 0;
    v3::Ctx c;
    return v3::ctx_layout(c, M);
}

bool lb_v3_supported(const V2Mech& M) { return M.nEqn <= 64 && M.nEqn >= 3 && M.tpc >= M.nEqn; }
int lb_v3_ctx_bytes() { return v3::CTX_BYTES; }
int lb_v3_max_groups(int tpc) { return tpc == 64 ? 7 : 512 / tpc; }

namespace {
template <int TPC, int SOLVER, bool DAC>
cudaError_t attr1(size_t smemBytes) {
    return cudaFuncSetAttribute(v3::integrate_v3_kernel<TPC, SOLVER, DAC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
}
template <int TPC>
cudaError_t attr_tpc(int solver, size_t smemBytes) {
    cudaError_t e = solver == LBCHEM_SOLVER_SEULEX ? attr1<TPC, LBCHEM_SOLVER_SEULEX, false>(smemBytes)
                                                   : attr1<TPC, LBCHEM_SOLVER_ROSENBROCK34, false>(smemBytes);
    if (e != cudaSuccess || TPC != 64) return e;
    // the DAC-reduced variant exists for 64-thread groups only (selected per launch by M.dac)
    return solver == LBCHEM_SOLVER_SEULEX ? attr1<64, LBCHEM_SOLVER_SEULEX, true>(smemBytes)
                                          : attr1<64, LBCHEM_SOLVER_ROSENBROCK34, true>(smemBytes);
}
template <int TPC>
cudaError_t launch_tpc(int solver, int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O,
                       const BatchArgs& B, cudaStream_t s) {
    const int block = groups * TPC;
    if (solver == LBCHEM_SOLVER_SEULEX) v3::integrate_v3_kernel<TPC, LBCHEM_SOLVER_SEULEX, false><<<grid, block, smemBytes, s>>>(M, O, B, wsDoubles);
    else v3::integrate_v3_kernel<TPC, LBCHEM_SOLVER_ROSENBROCK34, false><<<grid, block, smemBytes, s>>>(M, O, B, wsDoubles);
    return cudaGetLastError();
}
}  // namespace

// registers per thread, the kernel's own thread limit and its static shared memory (LBCHEM_VERBOSE diagnostics)
cudaError_t lb_v3_kernel_info(int solver, int* regs, int* maxThreads, int* staticSmem) {
    cudaFuncAttributes a{};
    cudaError_t e = solver == LBCHEM_SOLVER_SEULEX ? cudaFuncGetAttributes(&a, v3::integrate_v3_kernel<64, LBCHEM_SOLVER_SEULEX, false>)
                                                   : cudaFuncGetAttributes(&a, v3::integrate_v3_kernel<64, LBCHEM_SOLVER_ROSENBROCK34, false>);
    *regs = a.numRegs; *maxThreads = a.maxThreadsPerBlock; *staticSmem = (int)a.sharedSizeBytes;
    return e;
}

cudaError_t lb_v3_configure(int tpc, int solver, size_t smemBytes) {
    switch (tpc) {
        case 8: return attr_tpc<8>(solver, smemBytes);
        case 16: return attr_tpc<16>(solver, smemBytes);
        case 32: return attr_tpc<32>(solver, smemBytes);
        default: return attr_tpc<64>(solver, smemBytes);
    }
}

cudaError_t lb_v3_launch(int tpc, int solver, int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O,
                         const BatchArgs& B, cudaStream_t s) {
    if (M.dac) {
        if (tpc != 64) return cudaErrorInvalidValue;
        if (solver == LBCHEM_SOLVER_SEULEX) v3::integrate_v3_kernel<64, LBCHEM_SOLVER_SEULEX, true><<<grid, groups * 64, smemBytes, s>>>(M, O, B, wsDoubles);
        else v3::integrate_v3_kernel<64, LBCHEM_SOLVER_ROSENBROCK34, true><<<grid, groups * 64, smemBytes, s>>>(M, O, B, wsDoubles);
        return cudaGetLastError();
    }
    switch (tpc) {
        case 8: return launch_tpc<8>(solver, grid, groups, smemBytes, wsDoubles, M, O, B, s);
        case 16: return launch_tpc<16>(solver, grid, groups, smemBytes, wsDoubles, M, O, B, s);
        case 32: return launch_tpc<32>(solver, grid, groups, smemBytes, wsDoubles, M, O, B, s);
        default: return launch_tpc<64>(solver, grid, groups, smemBytes, wsDoubles, M, O, B, s);
    }
}

namespace {
template <int TPC>
cudaError_t tc_tpc(int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O, const BatchArgs& B, double* tcOut, cudaStream_t s) {
    cudaError_t e = cudaFuncSetAttribute(v3::tc_v3_kernel<TPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess) return e;
    v3::tc_v3_kernel<TPC><<<grid, groups * TPC, smemBytes, s>>>(M, O, B, wsDoubles, tcOut);
    return cudaGetLastError();
}
template <int TPC>
cudaError_t probe_tpc(int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O, const BatchArgs& B,
                      int kind, int nItems, const double* in, int inStride, double* out, int outStride, cudaStream_t s) {
    cudaError_t e = cudaFuncSetAttribute(v3::probe_v3_kernel<TPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess) return e;
    v3::probe_v3_kernel<TPC><<<grid, groups * TPC, smemBytes, s>>>(M, O, B, wsDoubles, kind, nItems, in, inStride, out, outStride);
    return cudaGetLastError();
}
}  // namespace

cudaError_t lb_v3_tc(int tpc, int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O, const BatchArgs& B,
                     double* tcOut, cudaStream_t s) {
    V2Mech Mm = M; Mm.dac = nullptr;
    switch (tpc) {
        case 8: return tc_tpc<8>(grid, groups, smemBytes, wsDoubles, Mm, O, B, tcOut, s);
        case 16: return tc_tpc<16>(grid, groups, smemBytes, wsDoubles, Mm, O, B, tcOut, s);
        case 32: return tc_tpc<32>(grid, groups, smemBytes, wsDoubles, Mm, O, B, tcOut, s);
        default: return tc_tpc<64>(grid, groups, smemBytes, wsDoubles, Mm, O, B, tcOut, s);
    }
}

cudaError_t lb_v3_probe(int tpc, int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O, const BatchArgs& B,
                        int kind, int nItems, const double* in, int inStride, double* out, int outStride, cudaStream_t s) {
    V2Mech Mm = M; Mm.dac = nullptr;
    switch (tpc) {
        case 8: return probe_tpc<8>(grid, groups, smemBytes, wsDoubles, Mm, O, B, kind, nItems, in, inStride, out, outStride, s);
        case 16: return probe_tpc<16>(grid, groups, smemBytes, wsDoubles, Mm, O, B, kind, nItems, in, inStride, out, outStride, s);
        case 32: return probe_tpc<32>(grid, groups, smemBytes, wsDoubles, Mm, O, B, kind, nItems, in, inStride, out, outStride, s);
        default: return probe_tpc<64>(grid, groups, smemBytes, wsDoubles, Mm, O, B, kind, nItems, in, inStride, out, outStride, s);
    }
}
