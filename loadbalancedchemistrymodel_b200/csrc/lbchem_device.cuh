// lbchem_device.cuh — device-side data model shared by the kernels and the host layer.
//
// Path replaced: the per-cell loop solveCellList()/solveCell() of
//   reference src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C:365-450
// plus the OpenFOAM ODE stack it calls through the pure-virtual solve(c,T,p,deltaT,subDeltaT)
// (LoadBalancedChemistryModel.H:235-242; arithmetic per SURVEY.md Appendix A).
#pragma once
#include <stdint.h>

#define LB_SMALL 1e-15
#define LB_VSMALL 1e-300
#define LB_GREAT 1e15
#define LB_VGREAT 1e300
#define LB_PSTD 1e5

#define LB_MAXSIDE 4      // max species per reaction side held in the fixed-width stoichiometry table

// Flattened mechanism in device memory. All arrays are structure-of-arrays so that thread r (reaction)
// or thread i (species) reads consecutive addresses.
struct DevMech {
    int nSp, nR, nEqn, nEff;
    int nTerms;          // number of d(omega_r)/d(c_j) terms = sum_r (nl_r + nr_r)
    int nEntries;        // structurally non-zero entries of the species block of the Jacobian
    int fracExp;         // 1 if any concentration exponent is non-integer or < 1 (faithful slow omega path)
    double TlowMix, ThighMix;  // max_i Tlow_i, min_i Thigh_i (temperature limits of the mixture thermo)
    int lda;             // leading dimension of the per-cell dense matrix (odd: bank-conflict-free row access)
    // species thermo (janaf); coefficient q of species i at [q*nSp + i]
    const double* W;
    const double* Tc;
    const double* cpLo; const double* cpHi;   // 5 x nSp : a0..a4                  (cp/R)
    const double* haLo; const double* haHi;   // 6 x nSp : a4/5,a3/4,a2/3,a1/2,a0,a5 (h/R, Horner order)
    const double* sLo;  const double* sHi;    // 6 x nSp : a4/4,a3/3,a2/2,a1,a0,a6   (s/R, Horner order; a0 multiplies ln T)
    // reactions
    const int* rtype; const int* rev; const int* effIdx;
    const double* A; const double* beta; const double* Ta;          // kInf for fall-off
    const double* A0; const double* beta0; const double* Ta0;       // k0 (fall-off only)
    const double* troe;                                             // 4 x nR : alpha, Tsss, Ts, Tss
    const double* rTc;                                              // Tcommon of the first lhs species (A.6 quirk)
    const double* dn;                                               // sum nu_rhs - sum nu_lhs
    const double* eff;                                              // nEff x nSp
    // fixed-width stoichiometry; slot s of reaction r at [s*nR + r]
    const int* nl; const int* nr;
    const int* lsp; const int* rsp;
    const double* lnu; const double* rnu;
    const double* lex; const double* rex;
    const int* dbase;                                               // first D-term index of reaction r
    // omega gather: species i sums coef * omega[reac] over [spPtr[i], spPtr[i+1]) in reaction order
    const int* spPtr; const int* spReac; const double* spCoef;
    // Jacobian gather: thread t owns entries [thrPtr[t], thrPtr[t+1]); entry e = (row,col) sums
    // coef * D[term] over [entPtr[e], entPtr[e+1]) in the reference's accumulation order
    const int* thrPtr; const int* entRow; const int* entCol; const int* entPtr;
    const int* termIdx; const double* termCoef;
};

struct DevOpts {
    int solver, maxSteps;
    double absTol, relTol, Treact, deltaTChemMax, cTauChem;
    int eqRateLimiter, dfdxIsDcdt;
    double RR, kcFloor;
    int seulexErrOldMax, pivoting;   // pivoting: lbchem_options.pivoting (compact-code kernel)
};

// one contiguous block of cells, field-major (the reference's cell list after pack: baseDataContainer.H:85-117)
struct CellSeg {
    long long n, ldy;
    const double* Y; const double* T; const double* p; const double* rho; const double* deltaT;
    double* deltaTChem; double* RR; double* Tout; double* cost; int* status;
    int* nActive;                // optional: number of active species DAC kept for each cell (NsDAC_)
    const int* index;            // optional miss list: queue position q of this segment integrates cell index[q]
};

// per-batch arguments. A batch is one or two segments handed out through ONE work queue: the load-balanced step
// integrates "own tail, then received cells" (reference LoadBalancedChemistryModel.C:608-611) in a single launch.
struct BatchArgs {
    CellSeg a, b;                // queue positions [0, a.n) -> segment a, [a.n, a.n + b.n) -> segment b
    double deltaTUniform;
    const int* order;            // optional dispatch order over the a.n + b.n queue positions (descending last cost), may be null
    unsigned long long* queue;   // work-queue head
    unsigned long long* stats;   // 8 counters: steps, attempts, jac, lu, solve, rhs, failed, cells
    double* scratch;             // v3 kernel: per-group global scratch (sparse Jacobian + seulex table), L2-resident
    int scratchStride;           // doubles per group
    int costModel;               // 0: cost = attempted sub-steps; 1: integer work units (see lbchem_options.cost_model)
};

// probe kinds (test hooks that run the same device functions as the integrator)
enum { LB_PROBE_DERIV = 0, LB_PROBE_JAC = 1, LB_PROBE_LUSOLVE = 2, LB_PROBE_RATES = 3 };
