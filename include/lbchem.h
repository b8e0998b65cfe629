/*
 * lbchem.h — C ABI of the B200-native per-cell chemistry integration path.
 *
 * This is the drop-in boundary for the hot path of ITV-Stuttgart/loadBalancedChemistryModel:
 * everything below `LoadBalancedChemistryModel::solve<DeltaTType>` that the reference runs
 * per cell on the CPU. Each entry point names the reference interface it replaces
 * (paths relative to the reference checkout):
 *
 *   lbchem_create / lbchem_destroy
 *       model construction: StandardChemistryModel(thermo) + LoadBalancedCoeffs read
 *       src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C:32-59
 *   lbchem_set_options
 *       chemistryProperties: odeCoeffs{solver;absTol;relTol;maxSteps}, EulerImplicitCoeffs,
 *       Treact, maxChemicalTimeStep (tutorial/counterFlowFlame2D/constant/chemistryProperties:17-38)
 *   lbchem_solve_batch / lbchem_solve_batch_host
 *       solveCellList() + solveCell() + the pure-virtual solve(c,T,p,deltaT,subDeltaT)
 *       LoadBalancedChemistryModel.C:365-450, LoadBalancedChemistryModel.H:235-242
 *       (field layout is OpenFOAM's own field-major Y_[i][celli] / RR_[i][celli],
 *        LoadBalancedChemistryModel.C:86-140 and :676-699, so the host layer is zero-copy)
 *   lbchem_balance_plan
 *       updateProcessorBalancing() + getSortedCPUTimesOnProcessor()
 *       LoadBalancedChemistryModel.C:170-308, :311-347
 *   lbchem_cells_to_send
 *       cellsToSend()  LoadBalancedChemistryModel.C:143-167
 *
 * All scalars are IEEE double, all labels int32 (reference: label=32, scalar=64).
 * No entry point aborts; errors come back as negative return codes and per-cell status.
 * A handle is bound to one CUDA device and one host thread at a time (the reference model
 * is not re-entrant either: shared mutable scratch c_, dcdt_).
 */
#ifndef LBCHEM_H
#define LBCHEM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* reaction-rate kinds */
enum {
    LBCHEM_R_ARRHENIUS = 0,         /* ArrheniusReactionRate */
    LBCHEM_R_THIRD_BODY = 1,        /* thirdBodyArrheniusReactionRate */
    LBCHEM_R_FALLOFF_LINDEMANN = 2, /* FallOffReactionRate<Arrhenius, Lindemann> */
    LBCHEM_R_FALLOFF_TROE = 3       /* FallOffReactionRate<Arrhenius, Troe> */
};

/* chemistry solvers (chemistryType{solver ...;}) */
enum {
    LBCHEM_SOLVER_NONE = 0,          /* noChemistrySolver */
    LBCHEM_SOLVER_EULER_IMPLICIT = 1,/* EulerImplicit */
    LBCHEM_SOLVER_ROSENBROCK34 = 2,  /* ode + odeCoeffs{solver Rosenbrock34;} */
    LBCHEM_SOLVER_SEULEX = 3         /* ode + odeCoeffs{solver seulex;} */
};

/* per-cell status codes */
enum {
    LBCHEM_CELL_OK = 0,
    LBCHEM_CELL_SKIPPED = 1,        /* T <= Treact: RR = 0 (LoadBalancedChemistryModel.C:413-419) */
    LBCHEM_CELL_MAX_STEPS = 2,      /* OpenFOAM: FatalError "Integration steps greater than maxSteps" */
    LBCHEM_CELL_UNDERFLOW = 3,      /* OpenFOAM: FatalError "stepsize underflow" */
    LBCHEM_CELL_SINGULAR = 4,       /* OpenFOAM: FatalError "Singular matrix" */
    LBCHEM_CELL_NAN = 5
};

/* return codes */
enum {
    LBCHEM_OK = 0,
    LBCHEM_E_INVALID = -1,
    LBCHEM_E_CUDA = -2,
    LBCHEM_E_UNSUPPORTED = -3,
    LBCHEM_E_NOMEM = -4
};

/* Flattened mechanism (structure of arrays). Pointers are HOST pointers, copied by lbchem_create. */
typedef struct lbchem_mech_desc {
    int32_t n_species;
    int32_t n_reactions;
    int32_t n_eff;                /* rows of the de-duplicated third-body efficiency table */
    int32_t reserved;
    const double* W;              /* [nSp] kg/kmol */
    const double* Tlow;           /* [nSp] */
    const double* Thigh;          /* [nSp] */
    const double* Tcommon;        /* [nSp] */
    const double* janaf_high;     /* [nSp][7] a0..a6 (dimensionless, cp/R form) */
    const double* janaf_low;      /* [nSp][7] */
    const int32_t* rtype;         /* [nR] LBCHEM_R_* */
    const int32_t* reversible;    /* [nR] 0/1 */
    const double* arr;            /* [nR][3] A, beta, Ta (kInf for fall-off) */
    const double* arr0;           /* [nR][3] k0 (fall-off only) */
    const double* troe;           /* [nR][4] alpha, Tsss, Ts, Tss */
    const int32_t* eff_index;     /* [nR] row of eff or -1 */
    const double* eff;            /* [n_eff][nSp] */
    const int32_t* lhs_ptr;       /* [nR+1] */
    const int32_t* lhs_species;
    const double* lhs_nu;         /* stoichiometric coefficient */
    const double* lhs_exp;        /* concentration exponent */
    const int32_t* rhs_ptr;       /* [nR+1] */
    const int32_t* rhs_species;
    const double* rhs_nu;
    const double* rhs_exp;
} lbchem_mech_desc;

/* Solver options. lbchem_default_options() fills OpenFOAM's defaults. */
typedef struct lbchem_options {
    int32_t solver;               /* LBCHEM_SOLVER_* */
    int32_t max_steps;            /* ODESolver maxSteps (default 10000) */
    double abs_tol;               /* odeCoeffs absTol */
    double rel_tol;               /* odeCoeffs relTol */
    double Treact;                /* cells with T <= Treact are skipped (default 0) */
    double deltaTChemMax;         /* maxChemicalTimeStep (default GREAT = 1e15) */
    double cTauChem;              /* EulerImplicitCoeffs */
    int32_t eq_rate_limiter;      /* EulerImplicitCoeffs equilibriumRateLimiter */
    int32_t dfdx_is_dcdt;         /* 1 = OpenFOAM quirk: jacobian() leaves dcdt in the dfdx slot */
    double gas_constant;          /* RR, J/(kmol K); default 8314.47 (OpenFOAM thermodynamicConstants) */
    double kc_floor;              /* kr = kf / max(Kc, kc_floor); default 1e-6 */
    int32_t threads_per_cell;     /* 0 = choose from mechanism size; else 32/64/128 */
    int32_t sort_by_cost;         /* 1 = dispatch cells to the work queue in descending last-cost order */
} lbchem_options;

typedef struct lbchem_handle_t* lbchem_handle;

/* Per-batch counters, filled by lbchem_get_stats after a solve (sums over the cells of the last batch). */
typedef struct lbchem_stats {
    int64_t n_cells;
    int64_t n_steps;      /* accepted sub-steps */
    int64_t n_attempts;   /* attempted sub-steps (accepted + rejected) */
    int64_t n_jac;        /* Jacobian evaluations the algorithm performs */
    int64_t n_lu;         /* LU factorisations */
    int64_t n_solve;      /* back-substitutions */
    int64_t n_rhs;        /* derivatives() evaluations the algorithm performs */
    int64_t n_failed;     /* cells with status >= LBCHEM_CELL_MAX_STEPS */
    double kernel_ms;     /* device time of the integration kernel of the last batch (CUDA events) */
    int32_t kernel_launches; /* kernels launched by the last solve call */
    int32_t reserved;
} lbchem_stats;

const char* lbchem_version(void);
void lbchem_default_options(lbchem_options* opts);

/* device < 0: use the current CUDA device. */
int lbchem_create(const lbchem_mech_desc* mech, int device, lbchem_handle* out);
int lbchem_destroy(lbchem_handle h);
int lbchem_set_options(lbchem_handle h, const lbchem_options* opts);
int lbchem_get_options(lbchem_handle h, lbchem_options* opts);
const char* lbchem_last_error(lbchem_handle h);

/*
 * Mechanism reduction of the LoadBalancedTDAC model: `reduction{active on; method DAC; tolerance; initialSet{...}}`
 * (reference tests/Cases/Case-chemistry/constant/chemistryProperties:35-66; automaticSIS off). When active, every cell
 * of the batches that follow is reduced before it is integrated, as solveCell does with mechRed()->active()
 * (reference src/chemistryModel/LoadBalancedTDACChemistryModel/LoadBalancedTDACChemistryModel.C:385-393,399-415):
 * species are kept by a breadth-first search over the direct-interaction coefficients from the initial set,
 * reactions touching a removed species are switched off, only the kept species (+T, p) are integrated, third-body
 * sums, rho and cp still use every species. Tout returns the temperature the cell came in with (the reference's
 * reduced call aliases T and p, :406-409). Solvers: seulex, Rosenbrock34. tabulation{} (ISAT) is host-side and not
 * part of this library.
 *   init_species [n_init]  indices of the initialSet species (1..8 entries), in dictionary order
 */
int lbchem_set_reduction(lbchem_handle h, int32_t active, double tolerance, int32_t n_init, const int32_t* init_species);
/* NsDAC_ of the first n cells of the last batch (HOST pointer); valid after a solve with reduction active. */
int lbchem_get_active_species(lbchem_handle h, int64_t n, int32_t* n_active);

/*
 * Integrate n cells over their flow time step. DEVICE pointers, stream-ordered on `stream`
 * (a cudaStream_t passed as void*; NULL = legacy default stream).
 *   Y        [nSp][ldy]  mass fractions, field-major (species i, cell c at Y[i*ldy + c]), in
 *   T,p,rho  [n]         in
 *   deltaT   [n]         flow time step per cell, in (LTS); or pass deltaT_uniform > 0 and deltaT = NULL
 *   deltaTChem [n]       chemical sub-step guess, in/out (clipped to deltaTChemMax on return)
 *   RR       [nSp][ldy]  reaction rates kg/m3/s, out
 *   Tout     [n]         temperature after the step, out (may be NULL)
 *   cost     [n]         per-cell cost = attempted sub-steps of this integration, out (may be NULL)
 *   status   [n]         LBCHEM_CELL_*, out (may be NULL)
 * Returns LBCHEM_OK or a negative error; cell-level failures are reported in status only.
 */
int lbchem_solve_batch(lbchem_handle h, int64_t n, int64_t ldy,
                       const double* Y, const double* T, const double* p, const double* rho,
                       const double* deltaT, double deltaT_uniform,
                       double* deltaTChem, double* RR, double* Tout, double* cost,
                       int32_t* status, void* stream);

/* Same call with HOST pointers: stages through the handle's pinned/device buffers (H2D, solve, D2H). */
int lbchem_solve_batch_host(lbchem_handle h, int64_t n, int64_t ldy,
                            const double* Y, const double* T, const double* p, const double* rho,
                            const double* deltaT, double deltaT_uniform,
                            double* deltaTChem, double* RR, double* Tout, double* cost,
                            int32_t* status);

int lbchem_get_stats(lbchem_handle h, lbchem_stats* out);

/*
 * Heat release rate Qdot[c] = -sum_i hf_i RR_i[c]  [W/m3], hf_i = formation enthalpy per mass of species i
 * (OpenFOAM StandardChemistryModel::Qdot(), which the combustion model calls right after solve(); the reference inherits
 * it unchanged, LoadBalancedChemistryModel.H:58-60). DEVICE pointers, stream-ordered; RR as written by
 * lbchem_solve_batch ([nSp][ldy]).
 */
int lbchem_qdot(lbchem_handle h, int64_t n, int64_t ldy, const double* RR, double* Qdot, void* stream);

/*
 * Chemical time scale tc[c] = nReaction * sum_i c_i / (SMALL + sum_r (sum_rhs nu) * forward rate of r)  [s]
 * (OpenFOAM StandardChemistryModel::tc(), used by the EDC / PaSR combustion models next to solve(); inherited unchanged by
 * the reference). DEVICE pointers, stream-ordered; same Y, T, p, rho arrays as lbchem_solve_batch. Needs a handle configured
 * for seulex or Rosenbrock34 (it reuses that kernel's rate sweep).
 */
int lbchem_tc(lbchem_handle h, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
              double* tc, void* stream);

/*
 * Test hook: evaluate one building block of the integrator for nItems states on the device, with the
 * same device functions lbchem_solve_batch uses (so a parity failure can be localised). HOST pointers;
 * strides in doubles. With n = n_species + 2 and y = (c[0..nSp), T, p):
 *   LBCHEM_PROBE_DERIV   in y[n]               out dydt[n]           StandardChemistryModel::derivatives
 *   LBCHEM_PROBE_JAC     in y[n]               out dfdx[n], J[n*n]   StandardChemistryModel::jacobian
 *   LBCHEM_PROBE_LUSOLVE in A[n*n], b[n]       out x[n]              LUDecompose + LUBacksubstitute
 *   LBCHEM_PROBE_RATES   in y[n]               out kf[nR], kr[nR]    Reaction::kf / kr
 */
enum { LBCHEM_PROBE_DERIV = 0, LBCHEM_PROBE_JAC = 1, LBCHEM_PROBE_LUSOLVE = 2, LBCHEM_PROBE_RATES = 3 };
int lbchem_probe(lbchem_handle h, int32_t kind, int32_t n_items, const double* in, int32_t in_stride,
                 double* out, int32_t out_stride);

/*
 * FP64-pipe peak of the device (DFMA micro-benchmark, ~ms_target milliseconds): the roofline denominator of
 * the chemistry kernels (bench.py; MEASURED_PEAKS.json carries no FP64 figure). device < 0: current device.
 */
int lbchem_fp64_peak(int device, double ms_target, double* tflops, double* sm_count);

/*
 * The balancing plan (bit-exact with the reference for the same cost vector).
 *   cost[nranks], cells[nranks]: all-gathered per-rank total cost and cell count.
 * Outputs are the plan of every rank, flattened in push order:
 *   send_ptr[nranks+1], send_to[], send_frac[]   (capacity nranks*nranks each)
 *   recv_ptr[nranks+1], recv_from[]
 * tdac != 0 selects the LoadBalancedTDAC variant (absolute seconds instead of fractions,
 * LoadBalancedTDACChemistryModel.C:149-287): send_frac then holds seconds.
 */
int lbchem_balance_plan(int32_t nranks, const double* cost, const int32_t* cells,
                        double min_fraction, int32_t tdac,
                        int32_t* send_ptr, int32_t* send_to, double* send_frac,
                        int32_t* recv_ptr, int32_t* recv_from);

/* cellsToSend: first index `end` >= start with sum(cost[start..end)) >= target, else n. Host pointers. */
int32_t lbchem_cells_to_send(const double* cell_cost, int32_t n, double target, int32_t start);

#ifdef __cplusplus
}
#endif
#endif /* LBCHEM_H */
