// scratch/fp64_peak.cu -- one-off measurement: sustained FP64 rate of a B200 through (a) DFMA, (b) DMMA m8n8k4,
// (c) DMMA m16n8k16, (d) DFMA and DMMA warps side by side.  Decides whether the Gram pass should use the FP64 tensor
// path.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/fp64_peak.bin scratch/fp64_peak.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

template <int CHAINS>
__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double a, double b) {
    double acc[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) acc[k] = threadIdx.x * 1e-3 + k;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) acc[k] = fma(acc[k], a, b);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma16816(double (&d)[4], const double (&a)[8], const double (&b)[4]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0, %1, %2, %3}, "
                 "{%4, %5, %6, %7, %8, %9, %10, %11}, {%12, %13, %14, %15}, {%0, %1, %2, %3};"
                 : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                   "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int CHAINS>
__global__ void __launch_bounds__(256) dmma884_kernel(double* out, int iters, double a, double b) {
    double d0[CHAINS], d1[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) { d0[k] = threadIdx.x * 1e-3 + k; d1[k] = k; }
    const double av = a + threadIdx.x * 1e-9, bv = b;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) dmma884(d0[k], d1[k], av, bv);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += d0[k] + d1[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CHAINS>
__global__ void __launch_bounds__(256) dmma16816_kernel(double* out, int iters, double a, double b) {
    double d[CHAINS][4];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k)
#pragma unroll
        for (int q = 0; q < 4; ++q) d[k][q] = threadIdx.x * 1e-3 + k + q;
    double av[8], bv[4];
#pragma unroll
    for (int q = 0; q < 8; ++q) av[q] = a + q * 1e-9;
#pragma unroll
    for (int q = 0; q < 4; ++q) bv[q] = b + q * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) dmma16816(d[k], av, bv);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += d[k][0] + d[k][1] + d[k][2] + d[k][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// even warps DFMA, odd warps DMMA m8n8k4
__global__ void __launch_bounds__(256) mixed_kernel(double* out, int iters, double a, double b) {
    const int warp = threadIdx.x >> 5;
    double s = 0;
    if (warp & 1) {
        double d0[8], d1[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { d0[k] = threadIdx.x * 1e-3 + k; d1[k] = k; }
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 8; ++k) dmma884(d0[k], d1[k], a, b);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) s += d0[k] + d1[k];
    } else {
        double acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = threadIdx.x * 1e-3 + k;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = fma(acc[k], a, b);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) s += acc[k];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double time_ms(F launch) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(); launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
    const int sms = pr.multiProcessorCount;
    printf("device %s, %d SMs\n", pr.name, sms);
    double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 256));
    const int iters = 20000;
    for (int bps = 1; bps <= 8; bps *= 2) {
        const int grid = sms * bps;
        {
            double ms = time_ms([&] { dfma_kernel<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
            double fl = 2.0 * grid * 256.0 * 8 * iters;
            printf("DFMA       chains 8  ctas/SM %d : %8.3f ms  %7.2f TFLOP/s\n", bps, ms, fl / ms * 1e-9);
        }
        {
            double ms = time_ms([&] { dfma_kernel<16><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
            double fl = 2.0 * grid * 256.0 * 16 * iters;
            printf("DFMA       chains 16 ctas/SM %d : %8.3f ms  %7.2f TFLOP/s\n", bps, ms, fl / ms * 1e-9);
        }
        {
            double ms = time_ms([&] { dmma884_kernel<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
            double fl = 2.0 * grid * 8.0 /*warps*/ * 8 * iters * 256.0;
            printf("DMMA 8x8x4 chains 8  ctas/SM %d : %8.3f ms  %7.2f TFLOP/s\n", bps, ms, fl / ms * 1e-9);
        }
        {
            double ms = time_ms([&] { dmma16816_kernel<4><<<grid, 256>>>(out, iters / 4, 1.0000001, 1e-9); });
            double fl = 2.0 * grid * 8.0 * 4 * (iters / 4) * 2048.0;
            printf("DMMA 16x8x16 chains 4 ctas/SM %d: %8.3f ms  %7.2f TFLOP/s\n", bps, ms, fl / ms * 1e-9);
        }
        {
            double ms = time_ms([&] { mixed_kernel<<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
            double fl = 2.0 * grid * 4.0 * 8 * iters * (256.0 + 32.0);
            printf("mixed (4 DFMA + 4 DMMA warps) ctas/SM %d: %8.3f ms  %7.2f TFLOP/s\n", bps, ms, fl / ms * 1e-9);
        }
    }
    return 0;
}
