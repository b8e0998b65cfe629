// lbchem_lb_cuda.cu -- the load-balanced flow step on the device: CUDA backend of lbchem_lb_core.hpp (two-segment launch
// of the integration kernel, cost total / cellsToSend as small kernels on the resident cost vector), NCCL transport over
// NVLink (grouped ncclSend / ncclRecv of field-major row slices, no packing), and the lbchem_lb_* C ABI.
//   reference: LoadBalancedChemistryModel.C:453-704, pointToPointBuffer.C:53-168 (see lbchem_lb_core.hpp)
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <atomic>
#include <cstdio>

#include "lbchem_internal.hpp"
#include "lbchem_lb_capi.hpp"

namespace {

// ------------------------------------------------------------------------------------------------
// kernels on the per-cell cost vector. Costs are non-negative INTEGERS stored as doubles (sub-step counts or work units),
// so every partial sum is exact in any order and the device answer of cellsToSend is bit-identical to the reference's
// sequential loop (LoadBalancedChemistryModel.C:143-167).
// ------------------------------------------------------------------------------------------------
constexpr int kTile = 1024;       // cells per block of the two-level prefix structure

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// tileSum[b] = sum cost[b*kTile ..), tileMin[b] = min deltaTChem[..]
__global__ void __launch_bounds__(256) lb_tile_reduce(const double* __restrict__ cost, const double* __restrict__ dtc, long long n,
                                                      double* __restrict__ tileSum, double* __restrict__ tileMin) {
    __shared__ double ss[8], sm[8];
    const long long base = (long long)blockIdx.x * kTile;
    double s = 0.0, m = 1e300;
#pragma unroll
    for (int q = 0; q < kTile / 256; ++q) {
        const long long i = base + q * 256 + threadIdx.x;
        if (i < n) { s += cost[i]; m = fmin(m, dtc[i]); }
    }
    s = warp_sum(s); m = warp_min(m);
    if ((threadIdx.x & 31) == 0) { ss[threadIdx.x >> 5] = s; sm[threadIdx.x >> 5] = m; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 1e300;
        for (int w = 0; w < 8; ++w) { a += ss[w]; b = fmin(b, sm[w]); }
        tileSum[blockIdx.x] = a; tileMin[blockIdx.x] = b;
    }
}

// tileExcl[b] = sum of the tiles before b (tileExcl[nb] = total); out[0] = total, out[1] = min
__global__ void __launch_bounds__(1024) lb_tile_scan(const double* __restrict__ tileSum, const double* __restrict__ tileMin, int nb,
                                                     double* __restrict__ tileExcl, double* __restrict__ out) {
    __shared__ double part[1024];
    __shared__ double carry;
    const int t = threadIdx.x;
    if (t == 0) carry = 0.0;
    double m = 1e300;
    __syncthreads();
    for (int b0 = 0; b0 < nb; b0 += 1024) {
        const int b = b0 + t;
        const double v = b < nb ? tileSum[b] : 0.0;
        if (b < nb) m = fmin(m, tileMin[b]);
        part[t] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            const double add = t >= o ? part[t - o] : 0.0;
            __syncthreads();
            part[t] += add;
            __syncthreads();
        }
        if (b < nb) tileExcl[b] = carry + part[t] - v;
        __syncthreads();
        if (t == 1023) carry += part[1023];
        __syncthreads();
    }
    part[t] = m;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) { if (t < o) part[t] = fmin(part[t], part[t + o]); __syncthreads(); }
    if (t == 0) { tileExcl[nb] = carry; out[0] = carry; out[1] = part[0]; }
}

// cellsToSend for nT targets in sequence (one block). E(i) = sum cost[0..i): find the first i >= start with
// E(i) - E(start) >= target, else n. Tile level first (tileExcl), then inside the one tile that holds the crossing.
__global__ void __launch_bounds__(1024) lb_bounds(const double* __restrict__ cost, long long n, const double* __restrict__ tileExcl, int nb,
                                                  const double* __restrict__ targets, int nT, long long* __restrict__ bounds) {
    __shared__ double pre[kTile];
    __shared__ long long hit;
    __shared__ long long sStart;
    __shared__ double sEs;
    const int t = threadIdx.x;
    if (t == 0) { sStart = 0; sEs = 0.0; bounds[0] = 0; }
    __syncthreads();
    for (int k = 0; k < nT; ++k) {
        const long long start = sStart;
        const double Es = sEs, target = targets[k];
        long long end = n; double Eend = 0.0;
        if (!(0.0 < target) || start >= n) {            // acc = 0 already reaches the target (or nothing is left): end = start
            end = start < n ? start : n; Eend = Es;
            if (start >= n) Eend = Es;
        } else {
            // first tile boundary b (as a cell index b*kTile > start) at which the running sum has reached the target
            if (t == 0) hit = (long long)nb + 1;
            __syncthreads();
            const long long b0 = start / kTile + 1;
            for (long long b = b0 + t; b <= nb; b += 1024) {
                if (tileExcl[b] - Es >= target) { atomicMin(&hit, b); break; }   // tileExcl is non-decreasing
            }
            __syncthreads();
            const long long bh = hit;
            __syncthreads();
            if (bh > nb) {                                // never reached: the sequential loop runs off the end
                end = n; Eend = tileExcl[nb];
            } else {
                // the crossing lies in (lo, bh*kTile], lo = max(start, (bh-1)*kTile)
                const long long tb = (bh - 1) * kTile;
                const long long lo = start > tb ? start : tb;
                const double Elo = (lo == start) ? Es : tileExcl[bh - 1];
                const long long i = lo + t;              // pre[t] = E(lo + t + 1) - E(lo)
                pre[t] = (i < n && i < bh * kTile) ? cost[i] : 0.0;
                __syncthreads();
                for (int o = 1; o < kTile; o <<= 1) {
                    const double add = t >= o ? pre[t - o] : 0.0;
                    __syncthreads();
                    pre[t] += add;
                    __syncthreads();
                }
                if (t == 0) hit = n;
                __syncthreads();
                // candidate index lo + t + 1 (<= n): E = Elo + pre[t]
                if (lo + t + 1 <= n && lo + t + 1 <= bh * kTile && (Elo + pre[t]) - Es >= target) atomicMin(&hit, lo + t + 1);
                __syncthreads();
                end = hit;
                Eend = (end - lo - 1 >= 0 && end - lo - 1 < kTile) ? Elo + pre[end - lo - 1] : tileExcl[nb];
                __syncthreads();
            }
        }
        if (t == 0) { bounds[k + 1] = end; sStart = end; sEs = Eend; }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// CUDA backend
// ------------------------------------------------------------------------------------------------
struct CudaBackend : lbcore::Backend {
    lbchem_handle h;
    double *tileSum = nullptr, *tileMin = nullptr, *tileExcl = nullptr, *dOut = nullptr, *dTargets = nullptr;
    long long* dBounds = nullptr;
    double* hPin = nullptr;            // pinned: [0..1] total/min, [2..2+P) targets, then bounds as long long
    long long tileCap = 0; int nbLast = 0; int maxT = 0;
    explicit CudaBackend(lbchem_handle hh, int nranks) : h(hh), maxT(nranks) {}
    ~CudaBackend() override {
        cudaSetDevice(h->device);
        void* p[] = {tileSum, tileMin, tileExcl, dOut, dTargets, dBounds};
        for (void* q : p) if (q) cudaFree(q);
        if (hPin) cudaFreeHost(hPin);
    }
    int ck(cudaError_t e, const char* what) {
        if (e == cudaSuccess) return 0;
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return LBCHEM_E_CUDA;
    }
    int init() {
        cudaSetDevice(h->device);
        if (ck(cudaMalloc(&dOut, 2 * sizeof(double)), "cudaMalloc")) return LBCHEM_E_CUDA;
        if (ck(cudaMalloc(&dTargets, (size_t)(maxT + 1) * sizeof(double)), "cudaMalloc")) return LBCHEM_E_CUDA;
        if (ck(cudaMalloc(&dBounds, (size_t)(maxT + 2) * sizeof(long long)), "cudaMalloc")) return LBCHEM_E_CUDA;
        if (ck(cudaHostAlloc(&hPin, (size_t)(2 * maxT + 8) * sizeof(double), cudaHostAllocDefault), "cudaHostAlloc")) return LBCHEM_E_CUDA;
        return 0;
    }
    int n_species() const override { return h->M.nSp; }
    void* alloc(size_t bytes) override {
        cudaSetDevice(h->device);
        void* p = nullptr;
        return cudaMalloc(&p, bytes ? bytes : 8) == cudaSuccess ? p : nullptr;
    }
    void release(void* p) override { cudaSetDevice(h->device); cudaFree(p); }
    int copy(void* dst, const void* src, size_t bytes, void* stream) override {
        return ck(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, static_cast<cudaStream_t>(stream)), "cudaMemcpyAsync");
    }
    int sync(void* stream) override { return ck(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)), "cudaStreamSynchronize"); }

    static CellSeg seg(const lbcore::Block& b, long long s) {
        CellSeg c{};
        c.n = b.n - s; c.ldy = b.ldy;
        c.Y = b.Y + s; c.T = b.T + s; c.p = b.p + s; c.rho = b.rho + s; c.deltaT = b.deltaT ? b.deltaT + s : nullptr;
        c.deltaTChem = b.deltaTChem + s; c.RR = b.RR + s; c.Tout = b.Tout ? b.Tout + s : nullptr;
        c.cost = b.cost ? b.cost + s : nullptr; c.status = b.status ? b.status + s : nullptr; c.nActive = nullptr;
        return c;
    }
    int solve(const lbcore::Block& a, long long aStart, const lbcore::Block* b, double dtU, void* stream) override {
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (ck(cudaSetDevice(h->device), "cudaSetDevice")) return LBCHEM_E_CUDA;
        h->launches = 0; h->hostPath = false; h->lastN = (a.n - aStart) + (b ? b->n : 0);
        if (ck(cudaMemsetAsync(h->dStats, 0, 8 * sizeof(unsigned long long), s), "cudaMemsetAsync")) return LBCHEM_E_CUDA;
        if (ck(cudaEventRecord(h->ev0, s), "cudaEventRecord")) return LBCHEM_E_CUDA;
        const CellSeg sa = seg(a, aStart);
        CellSeg sb{};
        if (b) sb = seg(*b, 0);
        const int rc = lb_launch_batch(h, &sa, b ? &sb : nullptr, dtU, nullptr, h->dQueue, h->dScratch, s);
        if (rc != LBCHEM_OK) { err = h->err; return rc; }
        if (ck(cudaEventRecord(h->ev1, s), "cudaEventRecord")) return LBCHEM_E_CUDA;
        h->evValid = true;
        return 0;
    }
    int tiles(long long n) {
        const long long nb = (n + kTile - 1) / kTile;
        if (nb + 1 > tileCap) {
            cudaSetDevice(h->device);
            void* p[] = {tileSum, tileMin, tileExcl};
            for (void* q : p) if (q) cudaFree(q);
            tileSum = tileMin = tileExcl = nullptr;
            const long long cap = nb + nb / 4 + 64;
            if (ck(cudaMalloc(&tileSum, cap * 8), "cudaMalloc") || ck(cudaMalloc(&tileMin, cap * 8), "cudaMalloc") ||
                ck(cudaMalloc(&tileExcl, (cap + 1) * 8), "cudaMalloc")) { tileCap = 0; return LBCHEM_E_CUDA; }
            tileCap = cap;
        }
        return 0;
    }
    int reduce(const double* cost, const double* dtc, long long n, double* total, double* dtMin, void* stream) override {
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (ck(cudaSetDevice(h->device), "cudaSetDevice")) return LBCHEM_E_CUDA;
        if (n <= 0) { *total = 0.0; *dtMin = 1e15; nbLast = 0; return ck(cudaStreamSynchronize(s), "cudaStreamSynchronize"); }
        if (tiles(n)) return LBCHEM_E_CUDA;
        const int nb = (int)((n + kTile - 1) / kTile);
        lb_tile_reduce<<<nb, 256, 0, s>>>(cost, dtc, n, tileSum, tileMin);
        lb_tile_scan<<<1, 1024, 0, s>>>(tileSum, tileMin, nb, tileExcl, dOut);
        if (ck(cudaGetLastError(), "lb_tile kernels")) return LBCHEM_E_CUDA;
        if (ck(cudaMemcpyAsync(hPin, dOut, 2 * sizeof(double), cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync")) return LBCHEM_E_CUDA;
        if (ck(cudaStreamSynchronize(s), "cudaStreamSynchronize")) return LBCHEM_E_CUDA;
        *total = hPin[0]; *dtMin = hPin[1]; nbLast = nb;
        h->launches += 2;
        return 0;
    }
    int bounds(const double* cost, long long n, const double* targets, int nT, long long* out, void* stream) override {
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (ck(cudaSetDevice(h->device), "cudaSetDevice")) return LBCHEM_E_CUDA;
        out[0] = 0;
        if (nT <= 0) return 0;
        if (nT > maxT || n <= 0 || nbLast != (int)((n + kTile - 1) / kTile)) { err = "cellsToSend: no cost vector of this size on the device"; return LBCHEM_E_INVALID; }
        double* hT = hPin + 2;
        long long* hB = reinterpret_cast<long long*>(hPin + 2 + maxT + 1);
        for (int k = 0; k < nT; ++k) hT[k] = targets[k];
        if (ck(cudaMemcpyAsync(dTargets, hT, nT * sizeof(double), cudaMemcpyHostToDevice, s), "cudaMemcpyAsync")) return LBCHEM_E_CUDA;
        lb_bounds<<<1, 1024, 0, s>>>(cost, n, tileExcl, nbLast, dTargets, nT, dBounds);
        if (ck(cudaGetLastError(), "lb_bounds")) return LBCHEM_E_CUDA;
        if (ck(cudaMemcpyAsync(hB, dBounds, (nT + 1) * sizeof(long long), cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync")) return LBCHEM_E_CUDA;
        if (ck(cudaStreamSynchronize(s), "cudaStreamSynchronize")) return LBCHEM_E_CUDA;
        for (int k = 0; k <= nT; ++k) out[k] = hB[k];
        return 0;
    }
};

// ------------------------------------------------------------------------------------------------
// NCCL transport. libnccl.so.2 is resolved at run time (a process that already loaded NCCL, e.g. through torch, shares
// that copy), so liblbchem.so itself has no link-time dependency on it.
// ------------------------------------------------------------------------------------------------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
    bool load() {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
        if (!lib) { err = std::string("cannot load libnccl.so.2: ") + dlerror(); return false; }
#define SYM(f) do { *reinterpret_cast<void**>(&f) = dlsym(lib, "nccl" #f); if (!f) { err = "libnccl: missing symbol nccl" #f; lib = nullptr; return false; } } while (0)
        SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(AllGather); SYM(Send); SYM(Recv); SYM(GroupStart); SYM(GroupEnd);
        SYM(GetErrorString);
#undef SYM
        return true;
    }
};
NcclApi& nccl() { static NcclApi a; return a; }

struct NcclTransport : lbcore::Transport {
    ncclComm_t comm = nullptr;
    int device = 0;
    double *dSend = nullptr, *dRecv = nullptr, *hBuf = nullptr; int cap = 0;
    cudaStream_t own = nullptr;
    ~NcclTransport() override {
        cudaSetDevice(device);
        if (comm) nccl().CommDestroy(comm);
        if (dSend) cudaFree(dSend);
        if (dRecv) cudaFree(dRecv);
        if (hBuf) cudaFreeHost(hBuf);
        if (own) cudaStreamDestroy(own);
    }
    int nk(ncclResult_t r, const char* what) {
        if (r == ncclSuccess) return 0;
        err = std::string(what) + ": " + nccl().GetErrorString(r);
        return LBCHEM_E_CUDA;
    }
    int all_gather(const double* send, double* recv, int count) override {
        cudaSetDevice(device);
        if (count > cap) {
            if (dSend) cudaFree(dSend);
            if (dRecv) cudaFree(dRecv);
            if (hBuf) cudaFreeHost(hBuf);
            cap = count + 16;
            if (cudaMalloc(&dSend, cap * 8) != cudaSuccess || cudaMalloc(&dRecv, (size_t)cap * size * 8) != cudaSuccess ||
                cudaHostAlloc(&hBuf, (size_t)cap * (size + 1) * 8, cudaHostAllocDefault) != cudaSuccess) { err = "all_gather staging allocation failed"; cap = 0; return LBCHEM_E_NOMEM; }
        }
        std::memcpy(hBuf, send, count * 8);
        if (cudaMemcpyAsync(dSend, hBuf, count * 8, cudaMemcpyHostToDevice, own) != cudaSuccess) { err = "all_gather H2D failed"; return LBCHEM_E_CUDA; }
        if (nk(nccl().AllGather(dSend, dRecv, count, ncclDouble, comm, own), "ncclAllGather")) return LBCHEM_E_CUDA;
        if (cudaMemcpyAsync(hBuf + cap, dRecv, (size_t)count * size * 8, cudaMemcpyDeviceToHost, own) != cudaSuccess ||
            cudaStreamSynchronize(own) != cudaSuccess) { err = "all_gather D2H failed"; return LBCHEM_E_CUDA; }
        std::memcpy(recv, hBuf + cap, (size_t)count * size * 8);
        return 0;
    }
    bool async() const override { return true; }
    int exchange(const std::vector<lbchem_p2p_op>& ops, void* stream) override {
        if (ops.empty()) return 0;
        cudaSetDevice(device);
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (nk(nccl().GroupStart(), "ncclGroupStart")) return LBCHEM_E_CUDA;
        int rc = 0;
        for (const lbchem_p2p_op& o : ops) {
            if (o.bytes <= 0) continue;
            const ncclResult_t r = o.is_send ? nccl().Send(o.ptr, (size_t)o.bytes, ncclChar, o.peer, comm, s)
                                             : nccl().Recv(o.ptr, (size_t)o.bytes, ncclChar, o.peer, comm, s);
            if (r != ncclSuccess) { rc = nk(r, o.is_send ? "ncclSend" : "ncclRecv"); break; }
        }
        const ncclResult_t e = nccl().GroupEnd();
        if (!rc && e != ncclSuccess) rc = nk(e, "ncclGroupEnd");
        return rc;
    }
};

// staging of the host-pointer entry point
struct HostStage {
    double *Y = nullptr, *RR = nullptr, *T = nullptr, *p = nullptr, *rho = nullptr, *dT = nullptr, *dtc = nullptr, *Tout = nullptr, *cost = nullptr;
    int32_t* status = nullptr;
    long long cap = 0;
    cudaStream_t stream = nullptr;
};

int make_lb(lbchem_handle h, std::unique_ptr<lbcore::Transport> tp, CudaBackend* be, const lbchem_lb_options* o, lbchem_lb* out) {
    lbchem_lb_options opt;
    if (o) opt = *o; else lbchem_lb_default_options(&opt);
    if (opt.update_iter < 0 || !(opt.min_fraction >= 0)) return LBCHEM_E_INVALID;
    lbchem_lb lb = new lbchem_lb_t();
    lb->chem = h;
    lb->drv.reset(new lbcore::Driver(std::unique_ptr<lbcore::Backend>(be), std::move(tp), opt));
    *out = lb;
    return LBCHEM_OK;
}

}  // namespace

extern "C" {

int lbchem_lb_nccl_unique_id(void* id128) {
    if (!id128) return LBCHEM_E_INVALID;
    if (!nccl().load()) return LBCHEM_E_UNSUPPORTED;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (nccl().GetUniqueId(&id) != ncclSuccess) return LBCHEM_E_CUDA;
    std::memcpy(id128, &id, sizeof(id));
    return LBCHEM_OK;
}

int lbchem_lb_create_nccl(lbchem_handle h, const void* id128, int32_t rank, int32_t nranks, const lbchem_lb_options* o, lbchem_lb* out) {
    if (!h || !out || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return LBCHEM_E_INVALID;
    *out = nullptr;
    if (!nccl().load()) return lb_fail(h, LBCHEM_E_UNSUPPORTED, nccl().err);
    CK(cudaSetDevice(h->device));
    std::unique_ptr<NcclTransport> tp(new NcclTransport());
    tp->rank = rank; tp->size = nranks; tp->device = h->device;
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    ncclResult_t r = nccl().CommInitRank(&tp->comm, nranks, id, rank);
    if (r != ncclSuccess) return lb_fail(h, LBCHEM_E_CUDA, std::string("ncclCommInitRank: ") + nccl().GetErrorString(r));
    CK(cudaStreamCreateWithFlags(&tp->own, cudaStreamNonBlocking));
    std::unique_ptr<CudaBackend> be(new CudaBackend(h, nranks));
    if (be->init()) return lb_fail(h, LBCHEM_E_CUDA, be->err);
    return make_lb(h, std::move(tp), be.release(), o, out);
}

int lbchem_lb_create_transport(lbchem_handle h, const lbchem_transport* t, int32_t rank, int32_t nranks, const lbchem_lb_options* o, lbchem_lb* out) {
    if (!h || !out || !t || !t->all_gather || !t->exchange || nranks < 1 || rank < 0 || rank >= nranks) return LBCHEM_E_INVALID;
    *out = nullptr;
    std::unique_ptr<CudaBackend> be(new CudaBackend(h, nranks));
    if (be->init()) return lb_fail(h, LBCHEM_E_CUDA, be->err);
    std::unique_ptr<lbcore::Transport> tp(new lbcore::CallbackTransport(*t, rank, nranks));
    return make_lb(h, std::move(tp), be.release(), o, out);
}

int lbchem_lb_create_loopback(int32_t nranks, const lbchem_handle* handles, const lbchem_lb_options* o, lbchem_lb* out) {
    if (nranks < 1 || !handles || !out) return LBCHEM_E_INVALID;
    for (int r = 0; r < nranks; ++r) { if (!handles[r]) return LBCHEM_E_INVALID; out[r] = nullptr; }
    auto hub = std::make_shared<lbcore::LoopHub>(nranks);
    for (int r = 0; r < nranks; ++r) {
        std::unique_ptr<CudaBackend> be(new CudaBackend(handles[r], nranks));
        if (be->init()) return lb_fail(handles[r], LBCHEM_E_CUDA, be->err);
        std::unique_ptr<lbcore::Transport> tp(new lbcore::LoopTransport(hub, r, be.get()));
        const int rc = make_lb(handles[r], std::move(tp), be.release(), o, &out[r]);
        if (rc != LBCHEM_OK) return rc;
    }
    return LBCHEM_OK;
}

int lbchem_cells_to_send_device(lbchem_handle h, const double* cell_cost, int64_t n, const double* targets, int32_t nT,
                                int64_t* bounds, void* stream) {
    if (!h || !targets || !bounds || nT < 0 || n < 0 || (n > 0 && !cell_cost)) return LBCHEM_E_INVALID;
    bounds[0] = 0;
    if (n == 0) { for (int k = 0; k < nT; ++k) bounds[k + 1] = 0; return LBCHEM_OK; }
    CudaBackend be(h, nT > 1 ? nT : 1);
    if (be.init()) return lb_fail(h, LBCHEM_E_CUDA, be.err);
    // the driver pairs reduce() (tile sums of the cost vector) with bounds(); min(deltaTChem) is not needed here
    double total = 0.0, mn = 0.0;
    if (be.reduce(cell_cost, cell_cost, n, &total, &mn, stream)) return lb_fail(h, LBCHEM_E_CUDA, be.err);
    std::vector<long long> b(nT + 1);
    if (be.bounds(cell_cost, n, targets, nT, b.data(), stream)) return lb_fail(h, LBCHEM_E_CUDA, be.err);
    for (int k = 0; k <= nT; ++k) bounds[k] = b[k];
    return LBCHEM_OK;
}

// ------------------------------------------------------------------------------------------------
// cellsToSend without any allocation (include/lbchem.h: lbchem_cells_to_send_noalloc; no counterpart in the reference).
// Same three kernels as CudaBackend::reduce / bounds; their scratch is the
// file-scope device array below instead of cudaMalloc'ed buffers. In HOST code `g_noalloc_scratch` is only a stand-in
// for the symbol: the device address comes from cudaGetSymbolAddress, once per device, with that device current.
// ------------------------------------------------------------------------------------------------
constexpr int kNoallocTiles = LBCHEM_NOALLOC_MAX_CELLS / kTile;                 // 4096
constexpr int kNoallocDoubles = 3 * kNoallocTiles + 8;                          // tileSum | tileMin | tileExcl[+1] | out[2]
static_assert(LBCHEM_NOALLOC_MAX_CELLS % kTile == 0, "the cell limit is a whole number of tiles");
static __device__ __align__(256) double g_noalloc_scratch[kNoallocDoubles];
constexpr int kNoallocMaxDev = 64;
static std::atomic<double*> g_noalloc_ptr[kNoallocMaxDev];                        // zero-initialised: not resolved yet

static int noalloc_fail(char* err, int32_t len, const char* fmt, long long a, long long b) {
    if (err && len > 0) snprintf(err, (size_t)len, fmt, a, b);
    return LBCHEM_E_INVALID;
}

int lbchem_cells_to_send_noalloc(const double* cell_cost, int64_t n, const double* targets, int32_t nT, int64_t* bounds,
                                 void* stream, char* err, int32_t err_len) {
    if (err && err_len > 0) err[0] = 0;
    if (!targets || !bounds || nT < 0 || n < 0 || (n > 0 && !cell_cost))
        return noalloc_fail(err, err_len, "lbchem_cells_to_send_noalloc: null pointer or negative size (n = %lld, nT = %lld)", (long long)n, nT);
    if (n > LBCHEM_NOALLOC_MAX_CELLS)
        return noalloc_fail(err, err_len, "lbchem_cells_to_send_noalloc: n = %lld exceeds LBCHEM_NOALLOC_MAX_CELLS = %lld "
                            "(the static scratch is sized for that; use lbchem_cells_to_send_device)", (long long)n, LBCHEM_NOALLOC_MAX_CELLS);
    if (nT > LBCHEM_NOALLOC_MAX_TARGETS)
        return noalloc_fail(err, err_len, "lbchem_cells_to_send_noalloc: nT = %lld exceeds LBCHEM_NOALLOC_MAX_TARGETS = %lld", nT, LBCHEM_NOALLOC_MAX_TARGETS);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    auto cuda_fail = [&](cudaError_t e, const char* what) {
        if (err && err_len > 0) snprintf(err, (size_t)err_len, "lbchem_cells_to_send_noalloc: %s: %s", what, cudaGetErrorString(e));
        return LBCHEM_E_CUDA;
    };
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev < 0 || dev >= kNoallocMaxDev) return noalloc_fail(err, err_len, "lbchem_cells_to_send_noalloc: device ordinal %lld not below %lld", dev, kNoallocMaxDev);
    double* base = g_noalloc_ptr[dev].load(std::memory_order_acquire);
    if (!base) {                                  // set-up, once per device (racing first calls resolve the same address)
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&base), g_noalloc_scratch);
        if (e != cudaSuccess || !base) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        g_noalloc_ptr[dev].store(base, std::memory_order_release);
    }
    if (n == 0) {                                 // the sequential loop returns n = 0 for every target
        e = cudaMemsetAsync(bounds, 0, (size_t)(nT + 1) * sizeof(int64_t), s);
        return e == cudaSuccess ? LBCHEM_OK : cuda_fail(e, "cudaMemsetAsync");
    }
    double* tileSum = base;
    double* tileMin = base + kNoallocTiles;
    double* tileExcl = base + 2 * kNoallocTiles;  // nb + 1 entries
    double* out = base + 3 * kNoallocTiles + 1;   // total, min (unused by the caller)
    const int nb = (int)((n + kTile - 1) / kTile);
    // every entry read below (tileSum / tileMin [0, nb), tileExcl [0, nb]) is written by the two kernels before: nothing
    // of the previous call's contents is ever read, so nothing needs clearing
    lb_tile_reduce<<<nb, 256, 0, s>>>(cell_cost, cell_cost, n, tileSum, tileMin);
    lb_tile_scan<<<1, 1024, 0, s>>>(tileSum, tileMin, nb, tileExcl, out);
    static_assert(sizeof(long long) == sizeof(int64_t), "bounds are 64-bit");
    lb_bounds<<<1, 1024, 0, s>>>(cell_cost, n, tileExcl, nb, targets, nT, reinterpret_cast<long long*>(bounds));
    e = cudaGetLastError();
    return e == cudaSuccess ? LBCHEM_OK : cuda_fail(e, "kernel launch");
}

int lbchem_lb_destroy(lbchem_lb lb) {
    if (!lb) return LBCHEM_E_INVALID;
    if (lb->chem) { cudaSetDevice(lb->chem->device); cudaDeviceSynchronize(); }
    if (lb->extra) {
        HostStage* st = static_cast<HostStage*>(lb->extra);
        void* p[] = {st->Y, st->RR, st->T, st->p, st->rho, st->dT, st->dtc, st->Tout, st->cost, st->status};
        for (void* q : p) if (q) cudaFree(q);
        if (st->stream) cudaStreamDestroy(st->stream);
        delete st;
    }
    delete lb;
    return LBCHEM_OK;
}

int lbchem_lb_solve(lbchem_lb lb, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
                    const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR, double* Tout, double* cost,
                    int32_t* status, double* deltaTMin, void* stream) {
    if (!lb) return LBCHEM_E_INVALID;
    lb->err.clear();
    if (n < 0 || ldy < n || (n > 0 && (!Y || !T || !p || !rho || !deltaTChem || !RR)) || (!deltaT && !(deltaT_uniform > 0))) {
        lb->err = "bad batch arguments";
        return LBCHEM_E_INVALID;
    }
    if (cudaSetDevice(lb->chem->device) != cudaSuccess) { lb->err = "cudaSetDevice failed"; return LBCHEM_E_CUDA; }
    lbcore::Block b;
    b.n = n; b.ldy = ldy; b.Y = Y; b.T = T; b.p = p; b.rho = rho; b.deltaT = deltaT; b.deltaTChem = deltaTChem; b.RR = RR;
    b.Tout = Tout; b.cost = cost; b.status = status;
    return lb->drv->solve(b, deltaT_uniform, deltaTMin, stream);
}

int lbchem_lb_solve_host(lbchem_lb lb, int64_t n, int64_t ldy, const double* Y, const double* T, const double* p, const double* rho,
                         const double* deltaT, double deltaT_uniform, double* deltaTChem, double* RR, double* Tout, double* cost,
                         int32_t* status, double* deltaTMin) {
    if (!lb) return LBCHEM_E_INVALID;
    lb->err.clear();
    if (n < 0 || ldy < n || (n > 0 && (!Y || !T || !p || !rho || !deltaTChem || !RR)) || (!deltaT && !(deltaT_uniform > 0))) {
        lb->err = "bad batch arguments";
        return LBCHEM_E_INVALID;
    }
    lbchem_handle h = lb->chem;
#define LCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { lb->err = std::string(#call) + ": " + cudaGetErrorString(e_); return LBCHEM_E_CUDA; } } while (0)
    LCK(cudaSetDevice(h->device));
    if (!lb->extra) lb->extra = new HostStage();
    HostStage& st = *static_cast<HostStage*>(lb->extra);
    const int nSp = h->M.nSp;
    if (!st.stream) LCK(cudaStreamCreateWithFlags(&st.stream, cudaStreamNonBlocking));
    if (n > st.cap) {
        void** ptrs[] = {(void**)&st.Y, (void**)&st.RR, (void**)&st.T, (void**)&st.p, (void**)&st.rho, (void**)&st.dT,
                         (void**)&st.dtc, (void**)&st.Tout, (void**)&st.cost, (void**)&st.status};
        for (void** q : ptrs) if (*q) { cudaFree(*q); *q = nullptr; }
        const size_t c = (size_t)n + (size_t)n / 16 + 64;
        LCK(cudaMalloc(&st.Y, c * nSp * 8)); LCK(cudaMalloc(&st.RR, c * nSp * 8));
        LCK(cudaMalloc(&st.T, c * 8)); LCK(cudaMalloc(&st.p, c * 8)); LCK(cudaMalloc(&st.rho, c * 8)); LCK(cudaMalloc(&st.dT, c * 8));
        LCK(cudaMalloc(&st.dtc, c * 8)); LCK(cudaMalloc(&st.Tout, c * 8)); LCK(cudaMalloc(&st.cost, c * 8)); LCK(cudaMalloc(&st.status, c * 4));
        st.cap = (long long)c;
    }
    cudaStream_t s = st.stream;
    const size_t w = (size_t)n * 8;
    if (n > 0) {
        LCK(cudaMemcpy2DAsync(st.Y, (size_t)st.cap * 8, Y, (size_t)ldy * 8, w, nSp, cudaMemcpyHostToDevice, s));
        LCK(cudaMemcpyAsync(st.T, T, w, cudaMemcpyHostToDevice, s));
        LCK(cudaMemcpyAsync(st.p, p, w, cudaMemcpyHostToDevice, s));
        LCK(cudaMemcpyAsync(st.rho, rho, w, cudaMemcpyHostToDevice, s));
        if (deltaT) LCK(cudaMemcpyAsync(st.dT, deltaT, w, cudaMemcpyHostToDevice, s));
        LCK(cudaMemcpyAsync(st.dtc, deltaTChem, w, cudaMemcpyHostToDevice, s));
    }
    lbcore::Block b;
    b.n = n; b.ldy = st.cap; b.Y = st.Y; b.T = st.T; b.p = st.p; b.rho = st.rho; b.deltaT = deltaT ? st.dT : nullptr;
    b.deltaTChem = st.dtc; b.RR = st.RR; b.Tout = st.Tout; b.cost = st.cost; b.status = st.status;
    const int rc = lb->drv->solve(b, deltaT_uniform, deltaTMin, s);
    if (rc != LBCHEM_OK) return rc;
    if (n > 0) {
        LCK(cudaMemcpy2DAsync(RR, (size_t)ldy * 8, st.RR, (size_t)st.cap * 8, w, nSp, cudaMemcpyDeviceToHost, s));
        LCK(cudaMemcpyAsync(deltaTChem, st.dtc, w, cudaMemcpyDeviceToHost, s));
        if (Tout) LCK(cudaMemcpyAsync(Tout, st.Tout, w, cudaMemcpyDeviceToHost, s));
        if (cost) LCK(cudaMemcpyAsync(cost, st.cost, w, cudaMemcpyDeviceToHost, s));
        if (status) LCK(cudaMemcpyAsync(status, st.status, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    }
    LCK(cudaStreamSynchronize(s));
#undef LCK
    return LBCHEM_OK;
}

}  // extern "C"
