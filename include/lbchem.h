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
    int32_t threads_per_cell;     /* 0 = choose kernel and group width from the mechanism size; 8/16/32/64 = general kernel
                                     with that many threads per cell (8 and 16: several cells share one warp) */
    int32_t sort_by_cost;         /* 1 = lbchem_solve_batch dispatches cells to the work queue in descending order of the
                                     cost[] array as it holds on ENTRY (the previous flow step's per-cell cost) */
    int32_t seulex_errold_max;    /* seulex: errOld = max(4 err, 1) [1, Numerical Recipes] or min(4 err, 1) [0, default]:
                                     doubtful OpenFOAM detail, a switch on both the kernel and the oracle */
    int32_t cost_model;           /* per-cell cost written to cost[]: 0 = attempted sub-steps; 1 = integer work units
                                     9 nLU + nSolve + 3 nRHS + 12 nJac (what those building blocks cost on this kernel) */
    int32_t pivoting;             /* LU of the compact-code kernel (seulex, Rosenbrock34). 0 = the reference's rule on every
                                     matrix: LUDecompose's partial pivoting with implicit row scaling, same pivot sequence and
                                     tie rule. 1 (default) = diagonal pivots first, accepted when every multiplier stays below
                                     2^10, else that matrix is factorised again with rule 0: same solution to round-off (the
                                     reference's choice IS the diagonal in 96 % of the columns of the GRI flame cases; results
                                     differ by ~1e-9 relative, inside the 1e-6 contract), no vote / exchange per pivot column
                                     and warp-local triangular solves. 2 = test hook: diagonal path always rejected (== 0) */
    int32_t reserved;
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
 *   cost     [n]         per-cell cost of this integration (lbchem_options.cost_model), out (may be NULL); with
 *                        sort_by_cost it is also read on entry (the previous step's cost, zeros on a first step)
 *   status   [n]         LBCHEM_CELL_*, out (may be NULL)
 * Returns LBCHEM_OK or a negative error; cell-level failures are reported in status only.
 */
int lbchem_solve_batch(lbchem_handle h, int64_t n, int64_t ldy,
                       const double* Y, const double* T, const double* p, const double* rho,
                       const double* deltaT, double deltaT_uniform,
                       double* deltaTChem, double* RR, double* Tout, double* cost,
                       int32_t* status, void* stream);

/*
 * The same integration for a MISS LIST: only the cells index[0 .. n_index) of the field-major arrays are integrated (gather
 * on load, scatter on store: RR / deltaTChem / Tout / cost / status are written at those cells only, every other entry
 * is left as it is). This is the device half of LoadBalancedTDAC with `tabulation{active on;}`: the host ISAT filter
 * (LoadBalancedTDACChemistryModel.C:683-740) retrieves what it can and lists the misses; no repacking of the fields.
 * DEVICE pointers (index too), stream-ordered; arrays must cover max(index) + 1 cells, ldy as in lbchem_solve_batch.
 */
int lbchem_solve_batch_indexed(lbchem_handle h, int64_t n_index, const int32_t* index, int64_t ldy,
                               const double* Y, const double* T, const double* p, const double* rho,
                               const double* deltaT, double deltaT_uniform,
                               double* deltaTChem, double* RR, double* Tout, double* cost,
                               int32_t* status, void* stream);

/* Same call with HOST pointers: the cells are cut into chunks that cycle through three streams (H2D, integrate, D2H),
 * copied straight from / to the caller's arrays. With pinned (page-locked) caller arrays the copies of one chunk overlap
 * the integration of its neighbours; with pageable arrays every copy blocks the host and the chunks run one after the
 * other. sort_by_cost is ignored here. One call at a time per handle; so is lbchem_solve_batch (one batch in flight per
 * handle: the work queue, the counters and the kernel scratch belong to the handle). */
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
 * Test hook: evaluate one building block of the integrator for nItems states on the device, with the device functions
 * of the kernel the handle is configured to run (the compact-code kernel for seulex / Rosenbrock34 on mechanisms it
 * supports, else the general kernel; threads_per_cell != 0 forces the general kernel), so that a parity failure can be
 * localised. HOST pointers; strides in doubles. With n = n_species + 2 and y = (c[0..nSp), T, p):
 *   LBCHEM_PROBE_DERIV       in y[n]               out dydt[n]           StandardChemistryModel::derivatives
 *   LBCHEM_PROBE_JAC         in y[n]               out dfdx[n], J[n*n]   StandardChemistryModel::jacobian
 *   LBCHEM_PROBE_LUSOLVE     in A[n*n], b[n]       out x[n]              LUDecompose + LUBacksubstitute (general kernel only)
 *   LBCHEM_PROBE_RATES       in y[n]               out kf[nR], kr[nR]    Reaction::kf / kr
 *   LBCHEM_PROBE_JAC_LUSOLVE in y[n], dg, b[n]     out x[n]              x = (dg I - J(y))^-1 b through jacobian +
 *                                                                        LUDecompose + LUBacksubstitute as the ODE solvers
 *                                                                        chain them (compact-code kernel only)
 */
enum { LBCHEM_PROBE_DERIV = 0, LBCHEM_PROBE_JAC = 1, LBCHEM_PROBE_LUSOLVE = 2, LBCHEM_PROBE_RATES = 3, LBCHEM_PROBE_JAC_LUSOLVE = 4 };
int lbchem_probe(lbchem_handle h, int32_t kind, int32_t n_items, const double* in, int32_t in_stride,
                 double* out, int32_t out_stride);

/*
 * cellsToSend (LoadBalancedChemistryModel.C:143-167) on a DEVICE cost vector, nT targets in sequence: bounds[0] = 0,
 * bounds[k+1] = first index >= bounds[k] at which the cost summed from bounds[k] reaches targets[k], else n.
 * targets / bounds are HOST arrays. Bit-identical to lbchem_cells_to_send when the costs are integers (what the
 * kernel writes). This is what lbchem_lb_solve runs; exported for tests.
 */
int lbchem_cells_to_send_device(lbchem_handle h, const double* cell_cost, int64_t n, const double* targets, int32_t nT,
                                int64_t* bounds, void* stream);

/*
 * The same cellsToSend search for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). No counterpart in the reference.
 *   - allocates NOTHING: no cudaMalloc / cudaFree / cudaHostAlloc, no allocating library routine; it only enqueues kernels
 *     (and one cudaMemsetAsync when n == 0) on `stream` of the CURRENT device and returns without synchronising;
 *   - cell_cost[n], targets[nT] (inputs) and bounds[nT + 1] (output) are all DEVICE arrays of the current device;
 *   - its two-level prefix structure (tile sums, exclusive tile prefix: 3 doubles per 1024 cells) lives in one file-scope
 *     device array of the library, sized for LBCHEM_NOALLOC_MAX_CELLS cells (about 96 KiB; every GPU that loads the
 *     library holds it until the process exits). n beyond that, or nT beyond LBCHEM_NOALLOC_MAX_TARGETS, is rejected
 *     with LBCHEM_E_INVALID and a message in err (if given; always NUL-terminated);
 *   - every call on one GPU shares that array: calls that overlap on the GPU (different streams, no ordering) RACE.
 *     Order them on one stream or with events. Nothing in the array has to start at zero: every element a call reads is
 *     written earlier in the same call, so the previous call's contents never show.
 * The first call on each device resolves the array's device address (cudaGetSymbolAddress; not an allocation): make
 * one call per device before starting a graph capture. Results are bit-identical to lbchem_cells_to_send_device.
 */
#define LBCHEM_NOALLOC_MAX_CELLS   4194304   /* 4 Mi cells = 4096 tiles of 1024 */
#define LBCHEM_NOALLOC_MAX_TARGETS 1024
int lbchem_cells_to_send_noalloc(const double* cell_cost, int64_t n, const double* targets, int32_t nT, int64_t* bounds,
                                 void* stream, char* err, int32_t err_len);

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

/* ------------------------------------------------------------------------------------------------------------------
 * The load-balanced flow step: LoadBalancedChemistryModel::solve<DeltaTType>
 *   reference src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C:453-704
 *   (first step local :471-509, plan :515-519 -> :170-347, cellsToSend + ship the first cells :521-570, own tail then
 *    received cells :608-611, results returned in the same order :613-673, unpack + total cost :676-701)
 *   and LoadBalancedTDACChemistryModel::solve<DeltaTType> (LoadBalancedTDACChemistryModel.C:623-979) with `tdac` set,
 *   with src/pointToPointBuffer/pointToPointBuffer.C:53-168 replaced by NCCL point-to-point over NVLink (or a transport
 *   the caller supplies). One lbchem_lb per rank (= per GPU, per host thread), on top of that rank's lbchem_handle.
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct lbchem_lb_t* lbchem_lb;

typedef struct lbchem_lb_options {
    int32_t update_iter;     /* LoadBalancedCoeffs{updateIter}: re-plan every update_iter + 1 steps (.C:41-45, 515-519) */
    int32_t tdac;            /* 1 = LoadBalancedTDAC driver: absolute cost targets (TDAC .C:232,256), literal 2 % threshold
                                (:214), deltaTChem[] is read but never written back (:968-971) */
    double min_fraction;     /* LoadBalancedCoeffs{minFractionOfCellsToSend}, default 0.02 */
    int32_t balance;         /* 0 = never exchange (every step takes the local branch: the non-balanced model) */
    int32_t reserved;
} lbchem_lb_options;

/* what the last lbchem_lb_solve did on this rank */
typedef struct lbchem_lb_info {
    int32_t rank, nranks;
    int32_t planned;          /* the plan was recomputed in this step */
    int32_t exchanged;        /* the step took the exchange branch (0: first step / one rank / balance off) */
    int32_t reallocated;      /* the receive block had to grow in this step */
    int32_t n_send_partners, n_recv_partners;
    int32_t reserved;
    int64_t n_cells;          /* this rank's cells */
    int64_t sent_cells;       /* own cells integrated by other ranks */
    int64_t recv_cells;       /* other ranks' cells integrated here */
    int64_t sent_bytes, recv_bytes;   /* both trips */
    double total_cost;        /* totalCpuTime_ after the step: sum of this rank's per-cell costs */
} lbchem_lb_info;

/* one message of a grouped point-to-point exchange (ptr: device memory for lbchem_lb_create_transport) */
typedef struct lbchem_p2p_op { int32_t is_send; int32_t peer; void* ptr; int64_t bytes; } lbchem_p2p_op;

/* Caller-supplied transport (e.g. MPI from the OpenFOAM side). all_gather: `count` doubles per rank, HOST buffers,
 * blocking. exchange: nops messages; messages between one pair of ranks match in posting order; blocking; called with
 * the step's stream drained, must return when every receive has landed. Both return 0 on success. */
typedef struct lbchem_transport {
    void* ctx;
    int (*all_gather)(void* ctx, const double* send, double* recv, int32_t count);
    int (*exchange)(void* ctx, int32_t nops, const lbchem_p2p_op* ops);
} lbchem_transport;

void lbchem_lb_default_options(lbchem_lb_options* o);
/* 128-byte NCCL unique id, made on one rank and broadcast by the caller (MPI_Bcast / any side channel). */
int lbchem_lb_nccl_unique_id(void* id128);
/* NCCL over NVLink / NVSwitch: one communicator per lbchem_lb (libnccl.so.2 is loaded on first use). */
int lbchem_lb_create_nccl(lbchem_handle h, const void* id128, int32_t rank, int32_t nranks, const lbchem_lb_options* o, lbchem_lb* out);
int lbchem_lb_create_transport(lbchem_handle h, const lbchem_transport* t, int32_t rank, int32_t nranks, const lbchem_lb_options* o, lbchem_lb* out);
/* nranks ranks inside ONE process, one host thread each (lbchem_lb_solve of every rank must be called concurrently):
 * messages are device-to-device copies. handles[r] may sit on the same or on different GPUs. out[nranks]. */
int lbchem_lb_create_loopback(int32_t nranks, const lbchem_handle* handles, const lbchem_lb_options* o, lbchem_lb* out);
int lbchem_lb_set_options(lbchem_lb lb, const lbchem_lb_options* o);
int lbchem_lb_destroy(lbchem_lb lb);
const char* lbchem_lb_last_error(lbchem_lb lb);

/*
 * One flow step over this rank's n cells: the argument list of lbchem_solve_batch (DEVICE pointers, stream-ordered
 * work, but the call itself blocks on two or three small read-backs: the cost total, cellsToSend bounds, message sizes)
 * plus *deltaTMin = min_c deltaTChem[c] after the step (what solve() returns before its 2*deltaT clamp, .C:714-718).
 * Collective over the ranks. Tout, cost, status may be NULL. deltaT: all ranks pass per-cell arrays or all pass NULL.
 * With 2+ ranks the cells [0, sent) of this rank are integrated elsewhere and their results written into RR / deltaTChem /
 * Tout / cost / status by the return trip, exactly as if they had been integrated here.
 */
int lbchem_lb_solve(lbchem_lb lb, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
                    const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR, double* Tout, double* cost,
                    int32_t* status, double* deltaTMin, void* stream);
/* The same step with HOST arrays (H2D, lbchem_lb_solve, D2H): what the OpenFOAM adapter calls with its fields. */
int lbchem_lb_solve_host(lbchem_lb lb, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
                         const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR, double* Tout, double* cost,
                         int32_t* status, double* deltaTMin);
int lbchem_lb_get_info(lbchem_lb lb, lbchem_lb_info* out);
/* The plan this rank holds (sendAndReceiveData_, .H:84-96) and the split of the last step. Capacity nranks each;
 * send_amount = fraction of the rank's total cost (standard) or absolute cost (tdac); bounds[nsend + 1]. */
int lbchem_lb_get_plan(lbchem_lb lb, int32_t* nsend, int32_t* send_to, double* send_amount, int64_t* bounds,
                       int32_t* nrecv, int32_t* recv_from, int64_t* recv_sizes);

#ifdef __cplusplus
}
#endif
#endif /* LBCHEM_H */
