// Development aid: TMEM as per-thread scratch (tcgen05.st / tcgen05.ld, 32x32b) -- data integrity for
// every warp / column slot and the cost of a 28-double round trip per thread.
#include <cstdio>
#include "tmem.cuh"
using namespace ofb;

__global__ void __launch_bounds__(512, 1) probe(double* out, long long* cyc, int iters)
{
    __shared__ uint32_t s_base;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc(&s_base, 512);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t base = s_base;
    const uint32_t a0 = tmem_addr(base, (warp >> 2) * 64);
    double v[28];
#pragma unroll
    for (int k = 0; k < 28; ++k) v[k] = threadIdx.x * 100.0 + k;
#pragma unroll
    for (int c = 0; c < 7; ++c) { double t[4] = {v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]}; tmem_st4(a0 + 8 * c, t); }
    tmem_wait_st();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < 7; ++c) {
            double t[4];
            tmem_ld4(a0 + 8 * c, t);
            tmem_wait_ld();
#pragma unroll
            for (int k = 0; k < 4; ++k) t[k] += 1.0;
            tmem_st4(a0 + 8 * c, t);
        }
        tmem_wait_st();
    }
    const long long t1 = clock64();
    double sum = 0.0;
#pragma unroll
    for (int c = 0; c < 7; ++c) {
        double t[4];
        tmem_ld4(a0 + 8 * c, t);
        tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 4; ++k) sum += t[k] - (threadIdx.x * 100.0 + 4 * c + k + iters);
    }
    out[blockIdx.x * 512 + threadIdx.x] = sum;      // 0 when every value came back intact
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    __syncthreads();
    if (warp == 0) tmem_dealloc(base, 512);
}

int main()
{
    double* o; long long* c;
    cudaMalloc(&o, 148 * 512 * 8); cudaMalloc(&c, 148 * 8);
    const int iters = 1000;
    probe<<<148, 512>>>(o, c, iters);
    cudaError_t e = cudaDeviceSynchronize();
    printf("run: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    static double h[148 * 512]; long long hc[148];
    cudaMemcpy(h, o, sizeof h, cudaMemcpyDeviceToHost); cudaMemcpy(hc, c, sizeof hc, cudaMemcpyDeviceToHost);
    int bad = 0; for (int i = 0; i < 148 * 512; ++i) if (h[i] != 0.0) ++bad;
    printf("bad threads: %d of %d\n", bad, 148 * 512);
    printf("cycles per iteration (16 warps x 28 doubles ld + st, chunked x8 with wait per chunk): %.1f\n", (double)hc[0] / iters);
    printf("  = %.1f bytes/clk/SM each way\n", 512 * 28 * 8.0 / ((double)hc[0] / iters));
    return 0;
}
