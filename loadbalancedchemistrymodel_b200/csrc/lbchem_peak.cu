// lbchem_peak.cu — FP64 pipe peak micro-benchmark (roofline denominator for the chemistry kernels;
// MEASURED_PEAKS.json has no FP64 entry). 8 independent DFMA chains per thread, all SMs busy.
#include <cuda_runtime.h>
#include "../../include/lbchem.h"

__global__ void __launch_bounds__(256) dfma_peak_kernel(double* out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

// Returns the sustained DFMA rate in TFLOP/s (2 flops per FMA) over ~`ms_target` milliseconds.
extern "C" int lbchem_fp64_peak(int device, double ms_target, double* tflops, double* sm_count) {
    if (!tflops) return LBCHEM_E_INVALID;
    if (device >= 0 && cudaSetDevice(device) != cudaSuccess) return LBCHEM_E_CUDA;
    cudaDeviceProp prop; int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return LBCHEM_E_CUDA;
    const int blocks = prop.multiProcessorCount * 8, threads = 256;
    double* out = nullptr;
    if (cudaMalloc(&out, sizeof(double) * blocks * threads) != cudaSuccess) return LBCHEM_E_NOMEM;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int iters = 2000;
    float ms = 0.f;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        dfma_peak_kernel<<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return LBCHEM_E_CUDA; }
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms >= ms_target) break;
        const double scale = ms > 0.05 ? ms_target / ms * 1.1 : 8.0;
        iters = (int)(iters * (scale > 8.0 ? 8.0 : scale)) + 1;
    }
    const double flops = 2.0 * 16 * 8 * (double)iters * blocks * threads;
    *tflops = flops / (ms * 1e-3) / 1e12;
    if (sm_count) *sm_count = prop.multiProcessorCount;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    return LBCHEM_OK;
}

// ------------------------------------------------------------------------------------------------
// Qdot = -sum_i hf_i RR_i  [W/m3]  (OpenFOAM StandardChemistryModel::Qdot, called by the combustion model right after
// solve()). HBM-bound: one thread per cell walks the field-major RR block, consecutive threads read
// consecutive addresses of each species row; (nSp + 1) * 8 bytes per cell.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) qdot_kernel(long long n, long long ldy, int nSp, const double* __restrict__ hf,
                                                   const double* __restrict__ RR, double* __restrict__ Qdot) {
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < n; c += (long long)gridDim.x * blockDim.x) {
        double q0 = 0.0, q1 = 0.0, q2 = 0.0, q3 = 0.0;
        int i = 0;
        // eight independent loads in flight per thread before the first FMA needs one
        for (; i + 7 < nSp; i += 8) {
            double r[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) r[u] = __ldcs(RR + (size_t)(i + u) * ldy + c);
            q0 = fma(-hf[i], r[0], q0); q1 = fma(-hf[i + 1], r[1], q1); q2 = fma(-hf[i + 2], r[2], q2); q3 = fma(-hf[i + 3], r[3], q3);
            q0 = fma(-hf[i + 4], r[4], q0); q1 = fma(-hf[i + 5], r[5], q1); q2 = fma(-hf[i + 6], r[6], q2); q3 = fma(-hf[i + 7], r[7], q3);
        }
        for (; i < nSp; ++i) q0 = fma(-hf[i], __ldcs(RR + (size_t)i * ldy + c), q0);
        Qdot[c] = (q0 + q1) + (q2 + q3);
    }
}

cudaError_t lb_launch_qdot(long long n, long long ldy, int nSp, const double* hf, const double* RR, double* Qdot, cudaStream_t s) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)sms * 8;          // a multiple of the SM count; grid-stride beyond that
    if (blocks > cap) blocks = cap;
    qdot_kernel<<<(int)blocks, 256, 0, s>>>(n, ldy, nSp, hf, RR, Qdot);
    return cudaGetLastError();
}
