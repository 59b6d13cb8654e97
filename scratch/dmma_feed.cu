// scratch/dmma_feed.cu -- one-off: DMMA.8x8x4 rate when the fragments come from shared memory the way pass A feeds
// them (5 LDS.64 + 6 DMMA per 4-point step, 6 accumulator pairs per warp), with and without a predicate on the DMMA.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double lds(unsigned a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
#define PP 132
template <bool PRED, int NA, int NB>
__global__ void __launch_bounds__(512) feed_kernel(double* out, int iters, unsigned mask) {
    extern __shared__ double ct[];      // [72][PP]
    for (int q = threadIdx.x; q < 72 * PP; q += blockDim.x) ct[q] = 1.0 + 1e-9 * q;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned base = (unsigned)__cvta_generic_to_shared(ct) + (((lane >> 2) * PP + (lane & 3)) * 8);
    unsigned ao[NA], bo[NB];
    for (int i = 0; i < NA; ++i) ao[i] = base + ((warp + i) % 9) * 8 * PP * 8;
    for (int j = 0; j < NB; ++j) bo[j] = base + ((warp + 2 + j) % 9) * 8 * PP * 8;
    bool on[NA * NB];
    for (int m = 0; m < NA * NB; ++m) on[m] = (mask >> m) & 1;
    double acc[NA * NB][2];
    for (int m = 0; m < NA * NB; ++m) { acc[m][0] = 0; acc[m][1] = 0; }
    for (int it = 0; it < iters; ++it) {
        for (int s = 0; s < 32; ++s) {
            double a[NA], b[NB];
#pragma unroll
            for (int i = 0; i < NA; ++i) a[i] = lds(ao[i] + s * 32);
#pragma unroll
            for (int j = 0; j < NB; ++j) b[j] = lds(bo[j] + s * 32);
#pragma unroll
            for (int i = 0; i < NA; ++i)
#pragma unroll
                for (int j = 0; j < NB; ++j) {
                    if (PRED) { if (on[i * NB + j]) dmma884(acc[i * NB + j][0], acc[i * NB + j][1], a[i], b[j]); }
                    else dmma884(acc[i * NB + j][0], acc[i * NB + j][1], a[i], b[j]);
                }
        }
    }
    double s = 0;
    for (int m = 0; m < NA * NB; ++m) s += acc[m][0] + acc[m][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F> static double time_ms(F launch) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) { CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms; }
    return best;
}
int main() {
    cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
    const int sms = pr.multiProcessorCount;
    double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 2 * 512));
    const size_t smem = sizeof(double) * 72 * PP;
    CK(cudaFuncSetAttribute(feed_kernel<true, 2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(feed_kernel<false, 2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(feed_kernel<false, 3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(feed_kernel<false, 2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int iters = 200;
    // predicated-off DMMAs: do they still occupy the tensor pipe?
    for (unsigned mask : {0x3fu, 0x07u, 0x01u}) {
        double ms = time_ms([&] { feed_kernel<true, 2, 3><<<sms, 12 * 32, smem>>>(out, iters, mask); });
        printf("2x3 predicated mask 0x%02x (12 warps): %8.3f ms\n", mask, ms);
    }
    for (int ctas = 1; ctas <= 1; ++ctas)
        for (int warps : {4, 8, 10, 12, 16}) {
            const int grid = sms * ctas;
            auto rep = [&](const char* name, double ms, int pairs) {
                printf("%-28s ctas/SM %d warps %2d: %8.3f ms %7.2f TFLOP/s\n", name, ctas, warps, ms,
                       2.0 * 256 * pairs * 32.0 * iters * warps * grid / ms * 1e-9);
            };
            rep("2x3 unpredicated", time_ms([&] { feed_kernel<false, 2, 3><<<grid, warps * 32, smem>>>(out, iters, 0x3f); }), 6);
            rep("2x3 predicated (all on)", time_ms([&] { feed_kernel<true, 2, 3><<<grid, warps * 32, smem>>>(out, iters, 0x3f); }), 6);
            rep("3x3 unpredicated", time_ms([&] { feed_kernel<false, 3, 3><<<grid, warps * 32, smem>>>(out, iters, 0x1ff); }), 9);
            rep("2x4 unpredicated", time_ms([&] { feed_kernel<false, 2, 4><<<grid, warps * 32, smem>>>(out, iters, 0xff); }), 8);
        }
    return 0;
}
