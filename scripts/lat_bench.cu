// lat_bench.cu -- dependent-issue latencies the critical-path model of DESIGN.md section 6.3 needs (not part of the product).
// One warp (or one 64-thread group for the named barrier) on one SM, clock64() around a dependent chain of N operations.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/lat_bench scripts/lat_bench.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

#define N 2048

template <int KIND>
__global__ void chain(double* out, long long* cyc, double a, double b, int lanesrc) {
    __shared__ double sm[64];
    sm[threadIdx.x & 63] = a;
    __syncthreads();
    double x = a + threadIdx.x * 1e-9, y = b;
    unsigned u = threadIdx.x * 2654435761u;
    int idx = threadIdx.x & 63;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) {
        if (KIND == 0) x = fma(x, a, b);                                   // DFMA
        if (KIND == 1) x = x * a;                                          // DMUL
        if (KIND == 2) x = x + b;                                          // DADD
        if (KIND == 3) x = __shfl_sync(0xffffffffu, x, lanesrc);           // 64-bit shuffle (two SHFL.IDX)
        if (KIND == 4) { x = fma(__shfl_sync(0xffffffffu, x, lanesrc), a, b); }   // shuffle + DFMA (a substitution step)
        if (KIND == 5) { idx = (int)sm[idx] & 63; }                        // LDS.64 -> F2I -> address (upper bound of an LDS round trip)
        if (KIND == 6) x = 1.0 / x;                                        // full division path
        if (KIND == 7) x = exp(x * 1e-3);                                  // exp
        if (KIND == 8) u = __reduce_max_sync(0xffffffffu, u) + threadIdx.x;   // REDUX
        if (KIND == 9) { asm volatile("bar.sync 1, 64;" ::: "memory"); }   // named barrier, two warps
        if (KIND == 10) { sm[threadIdx.x & 63] = x; asm volatile("bar.sync 1, 64;" ::: "memory"); x = fma(sm[(threadIdx.x + 33) & 63], a, b); }   // STS + bar + LDS + DFMA (old substitution step)
        if (KIND == 11) { float f = __int_as_float(u); f = f * 1.0001f + 1.0f; u = __float_as_int(f); }   // FFMA
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    out[threadIdx.x] = x + y + u + idx;
}

int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 64 * sizeof(double)); cudaMalloc(&cyc, sizeof(long long));
    const char* names[] = {"DFMA", "DMUL", "DADD", "SHFL64", "SHFL64+DFMA", "LDS.64+F2I", "1.0/x", "exp()", "REDUX.MAX", "bar.sync 64",
                           "STS+bar+LDS+DFMA", "FFMA"};
    for (int k = 0; k < 12; ++k) {
        long long h = 0;
        const int threads = (k == 9 || k == 10) ? 64 : 32;
        for (int rep = 0; rep < 2; ++rep) {
            switch (k) {
#define C(K) case K: chain<K><<<1, threads>>>(out, cyc, 0.999999, 1e-9, 5); break;
                C(0) C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8) C(9) C(10) C(11)
            }
            cudaDeviceSynchronize();
        }
        cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        printf("%-18s %7.1f cycles per dependent step\n", names[k], (double)h / N);
    }
    return 0;
}
