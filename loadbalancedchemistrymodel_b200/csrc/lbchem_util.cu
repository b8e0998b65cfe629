// lbchem_util.cu -- small device utilities next to the integration kernel.
//
// Cost-ordered dispatch (lbchem_options.sort_by_cost): the persistent kernel hands out cells through one atomic work
// queue; with `order` set, queue position q integrates cell order[q]. The order is a counting sort of the cells by the
// cost of their LAST integration, most expensive first, so that the cells that take longest start first and the tail of
// the launch is made of cheap cells (north_star: "a persistent-kernel work queue so cells needing many substeps do not
// stall their neighbours"). Costs are small non-negative integers (sub-step attempts or work units), so 4096 bins with a
// saturating last bin sort them exactly; ties keep no particular order (cells are independent, any order gives the same
// results).
#include <cuda_runtime.h>
#include "lbchem_internal.hpp"

namespace {

constexpr int kBins = 4096;

__device__ __forceinline__ int cost_bin(double c, double scale) {
    c *= scale;
    if (!(c > 0.0)) return 0;               // also NaN
    return c < (double)(kBins - 1) ? (int)c : kBins - 1;
}

__global__ void __launch_bounds__(1024) cost_hist_kernel(const double* __restrict__ cost, long long n, double scale, int* __restrict__ hist) {
    __shared__ int sh[kBins];
    for (int i = threadIdx.x; i < kBins; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        atomicAdd(&sh[cost_bin(cost[i], scale)], 1);
    __syncthreads();
    for (int i = threadIdx.x; i < kBins; i += blockDim.x) if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// hist[b] <- first queue position of bin b when the bins are laid out in DESCENDING cost order
__global__ void __launch_bounds__(1024) cost_scan_kernel(int* __restrict__ hist) {
    __shared__ int part[1024];
    const int t = threadIdx.x;
    // thread t owns the four bins kBins-1-4t .. kBins-4-4t (descending)
    int v[4], s = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[q] = hist[kBins - 1 - (4 * t + q)]; s += v[q]; }
    part[t] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int add = t >= o ? part[t - o] : 0;
        __syncthreads();
        part[t] += add;
        __syncthreads();
    }
    int run = part[t] - s;
#pragma unroll
    for (int q = 0; q < 4; ++q) { hist[kBins - 1 - (4 * t + q)] = run; run += v[q]; }
}

__global__ void __launch_bounds__(1024) cost_scatter_kernel(const double* __restrict__ cost, long long n, double scale, int* __restrict__ cursor,
                                                            int* __restrict__ order) {
    __shared__ int cnt[kBins];
    __shared__ int basep[kBins];
    const long long per = ((n + gridDim.x - 1) / gridDim.x + blockDim.x - 1) / blockDim.x * blockDim.x;   // contiguous tile per block
    const long long lo = (long long)blockIdx.x * per, hi = lo + per < n ? lo + per : n;
    for (int i = threadIdx.x; i < kBins; i += blockDim.x) cnt[i] = 0;
    __syncthreads();
    for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) atomicAdd(&cnt[cost_bin(cost[i], scale)], 1);
    __syncthreads();
    for (int i = threadIdx.x; i < kBins; i += blockDim.x) {
        const int c = cnt[i];
        basep[i] = c ? atomicAdd(&cursor[i], c) : 0;
        cnt[i] = 0;
    }
    __syncthreads();
    for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        const int b = cost_bin(cost[i], scale);
        order[basep[b] + atomicAdd(&cnt[b], 1)] = (int)i;
    }
}

}  // namespace

cudaError_t lb_launch_cost_order(const double* cost, long long n, double scale, int* hist, int* order, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(hist, 0, kBins * sizeof(int), s);
    if (e != cudaSuccess) return e;
    int blocks = (int)((n + 4095) / 4096);
    if (blocks > 592) blocks = 592;
    if (blocks < 1) blocks = 1;
    cost_hist_kernel<<<blocks, 1024, 0, s>>>(cost, n, scale, hist);
    cost_scan_kernel<<<1, 1024, 0, s>>>(hist);
    cost_scatter_kernel<<<blocks, 1024, 0, s>>>(cost, n, scale, hist, order);
    return cudaGetLastError();
}
