// lbchem_internal.hpp -- what the translation units of liblbchem.so share: the handle, the launch entry of the
// integration kernels, error helpers. Not installed; the public boundary is include/lbchem.h.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>

#include "../../include/lbchem.h"
#include "lbchem_device.cuh"
#include "lbchem_v2.cuh"

int lb_v2_build(const lbchem_mech_desc* m, V2Mech* out, std::vector<void*>& allocs, std::string& err, std::vector<int>* posOfOut);
cudaError_t lb_v3_probe(int tpc, int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O, const BatchArgs& B,
                        int kind, int nItems, const double* in, int inStride, double* out, int outStride, cudaStream_t s);
cudaError_t lb_v3_tc(int tpc, int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O, const BatchArgs& B,
                     double* tcOut, cudaStream_t s);
cudaError_t lb_launch_qdot(long long n, long long ldy, int nSp, const double* hf, const double* RR, double* Qdot, cudaStream_t s);
int lb_dac_program(const lbchem_mech_desc* m, double tol, int nInit, const int32_t* init, std::vector<int>& prog, int* nPairsOut);
int lb_v3_workspace(V2Mech& M, int solver, double relTol, double absTol);
bool lb_v3_supported(const V2Mech& M);
int lb_v3_ctx_bytes();
cudaError_t lb_v3_configure(int tpc, int solver, size_t smemBytes);
int lb_v3_max_groups(int tpc);
cudaError_t lb_v3_kernel_info(int solver, int* regs, int* maxThreads, int* staticSmem);
cudaError_t lb_v3_launch(int tpc, int solver, int grid, int groups, size_t smemBytes, int wsDoubles, const V2Mech& M, const DevOpts& O,
                         const BatchArgs& B, cudaStream_t s);
size_t lb_work_doubles(const DevMech& M, int solver);
cudaError_t lb_configure(int tpc, int groups, size_t smemBytes, int* ctasPerSM);
cudaError_t lb_launch_integrate(int tpc, int grid, int groups, size_t smemBytes, int wsDoubles, const DevMech& M,
                                const DevOpts& O, const BatchArgs& B, cudaStream_t s);
cudaError_t lb_launch_probe(int tpc, int groups, size_t smemBytes, int wsDoubles, const DevMech& M, const DevOpts& O,
                            int kind, int nItems, const double* in, double* out, int inStride, int outStride, cudaStream_t s);

constexpr int kStreams = 3;

struct HostMech {   // host copy of the flattened tables (kept for re-partitioning when threads_per_cell changes)
    int nSp = 0, nR = 0, nEff = 0;
    std::vector<int> nl, nr, lsp, rsp, dbase;
    std::vector<double> lnu, rnu;
    int nTerms = 0;
};

struct Slot {       // per-stream staging for the host-pointer entry point
    cudaStream_t stream = nullptr;
    cudaEvent_t k0 = nullptr, k1 = nullptr;
    double *Y = nullptr, *RR = nullptr, *T = nullptr, *p = nullptr, *rho = nullptr, *dT = nullptr, *dtc = nullptr,
           *Tout = nullptr, *cost = nullptr;
    int* status = nullptr;
    unsigned long long* queue = nullptr;
    double* scratch = nullptr; size_t scratchCap = 0;   // this slot's own kernel scratch (sparse Jacobian + seulex table per group)
    long long cap = 0;
    bool used = false;
};

struct lbchem_handle_t {
    int device = 0;
    int nSM = 0;
    size_t smemMax = 0;
    DevMech M{};
    DevOpts O{};
    lbchem_options opts{};
    HostMech H;
    std::vector<void*> allocs;
    int tpc = 64, groups = 1, ctasPerSM = 1;
    size_t wsDoubles = 0, smemBytes = 0;
    int configuredSolver = -1, configuredTpc = -1;
    V2Mech V2{};                     // packed mechanism of the fast path
    std::vector<int> v3PosOf;        // packed position of reaction r (probe: kf / kr back to mechanism order)
    bool v2Built = false, useV2 = false, useV3 = false;
    int v3Groups = 0, v3Ws = 0; size_t v3Smem = 0;
    double* dScratch = nullptr; int scratchStride = 0; size_t scratchCap = 0;
    // DAC reduction (LoadBalancedTDAC): host copy of the program, device copy, per-cell NsDAC_ of the last batch
    std::vector<int> dacProg; int dacPairs = 0; int* dDac = nullptr; bool dacOn = false;
    int* dNActive = nullptr; long long nActiveCap = 0, nActiveOff = 0;
    int* dOrder = nullptr; long long orderCap = 0; int* dHist = nullptr;   // sort_by_cost dispatch order
    double *dProbeIn = nullptr, *dProbeOut = nullptr; size_t probeInCap = 0, probeOutCap = 0;   // lbchem_probe staging
    // Qdot: dimensionless formation enthalpies ha_low(Tstd)/R per species (host), hf = RR/W * that in J/kg (device)
    std::vector<double> hfOverR, Wsp; double* dHf = nullptr; double hfGas = -1;
    void* dThr = nullptr;            // device copy of thrPtr for the configured tpc
    unsigned long long* dQueue = nullptr;
    unsigned long long* dStats = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool evValid = false, hostPath = false;
    int launches = 0;
    long long lastN = 0;
    Slot slots[kStreams];
    std::string err;
};


// error helpers shared by the host translation units
int lb_fail(lbchem_handle h, int code, const std::string& msg);
#define CK(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess)                                                                        \
            return lb_fail(h, LBCHEM_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));     \
    } while (0)

// One launch of the configured integration kernel over one or two segments (b may be null), stream-ordered on s.
// `queue`: one zero-initialised-by-this-call work-queue word; `scratch`: per-launch global scratch of the v3 kernel
// (handle->dScratch for the device entry points; concurrent launches need their own, see Slot).
int lb_launch_batch(lbchem_handle h, const CellSeg* a, const CellSeg* b, double deltaTUniform, const int* order,
                    unsigned long long* queue, double* scratch, cudaStream_t s);
// Dispatch order for sort_by_cost: positions of [0, n) in descending order of lastCost (counting sort on the device).
int lb_cost_order(lbchem_handle h, const double* lastCost, long long n, cudaStream_t s, const int** orderOut);
cudaError_t lb_launch_cost_order(const double* cost, long long n, double scale, int* hist, int* order, cudaStream_t s);
