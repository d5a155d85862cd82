// Micro-benchmarks that size the fused level kernel: FP64 add/mul/fma issue rate, 64-bit
// shared-memory load rate, shuffle rate, and their overlap.  Prints ops/clk/SM.
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 4096
template <int MODE>
__global__ void __launch_bounds__(1024) k(double* out, long long* cyc, double seed)
{
    __shared__ double sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = seed * i;
    __syncthreads();
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double c = seed * 1.0000001;
    int idx = threadIdx.x;
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < ITER; ++i) {
        if (MODE == 0) { a0 = __dadd_rn(a0, c); a1 = __dadd_rn(a1, c); a2 = __dadd_rn(a2, c); a3 = __dadd_rn(a3, c); a4 = __dadd_rn(a4, c); a5 = __dadd_rn(a5, c); a6 = __dadd_rn(a6, c); a7 = __dadd_rn(a7, c); }
        if (MODE == 1) { a0 = __dmul_rn(a0, c); a1 = __dmul_rn(a1, c); a2 = __dmul_rn(a2, c); a3 = __dmul_rn(a3, c); a4 = __dmul_rn(a4, c); a5 = __dmul_rn(a5, c); a6 = __dmul_rn(a6, c); a7 = __dmul_rn(a7, c); }
        if (MODE == 2) { a0 = __fma_rn(a0, c, c); a1 = __fma_rn(a1, c, c); a2 = __fma_rn(a2, c, c); a3 = __fma_rn(a3, c, c); a4 = __fma_rn(a4, c, c); a5 = __fma_rn(a5, c, c); a6 = __fma_rn(a6, c, c); a7 = __fma_rn(a7, c, c); }
        if (MODE == 3) {  // 8 LDS.64, conflict-free
            a0 += sm[(idx + 0) & 4095]; a1 += sm[(idx + 32) & 4095]; a2 += sm[(idx + 64) & 4095]; a3 += sm[(idx + 96) & 4095];
            a4 += sm[(idx + 128) & 4095]; a5 += sm[(idx + 160) & 4095]; a6 += sm[(idx + 192) & 4095]; a7 += sm[(idx + 224) & 4095];
            idx += 256;
        }
        if (MODE == 4) {  // 8 double shuffles (16 SHFL)
            a0 = __shfl_down_sync(~0u, a0, 1); a1 = __shfl_down_sync(~0u, a1, 1); a2 = __shfl_down_sync(~0u, a2, 2); a3 = __shfl_down_sync(~0u, a3, 3);
            a4 = __shfl_up_sync(~0u, a4, 1); a5 = __shfl_up_sync(~0u, a5, 2); a6 = __shfl_up_sync(~0u, a6, 1); a7 = __shfl_up_sync(~0u, a7, 1);
        }
        if (MODE == 5) {  // 8 DADD + 4 LDS.64 : do they overlap?
            a0 = __dadd_rn(a0, sm[(idx + 0) & 4095]); a1 = __dadd_rn(a1, sm[(idx + 32) & 4095]); a2 = __dadd_rn(a2, sm[(idx + 64) & 4095]); a3 = __dadd_rn(a3, sm[(idx + 96) & 4095]);
            a4 = __dadd_rn(a4, c); a5 = __dadd_rn(a5, c); a6 = __dadd_rn(a6, c); a7 = __dadd_rn(a7, c);
            idx += 128;
        }
        if (MODE == 6) {  // 8 DADD + 4 double shuffles
            a0 = __dadd_rn(a0, __shfl_down_sync(~0u, a4, 1)); a1 = __dadd_rn(a1, __shfl_down_sync(~0u, a5, 1));
            a2 = __dadd_rn(a2, __shfl_down_sync(~0u, a6, 1)); a3 = __dadd_rn(a3, __shfl_down_sync(~0u, a7, 1));
            a4 = __dadd_rn(a4, c); a5 = __dadd_rn(a5, c); a6 = __dadd_rn(a6, c); a7 = __dadd_rn(a7, c);
        }
        if (MODE == 7) {  // 8 LDS.128
            const double2* s2 = (const double2*)sm;
            double2 v0 = s2[(idx + 0) & 2047], v1 = s2[(idx + 32) & 2047], v2 = s2[(idx + 64) & 2047], v3 = s2[(idx + 96) & 2047];
            double2 v4 = s2[(idx + 128) & 2047], v5 = s2[(idx + 160) & 2047], v6 = s2[(idx + 192) & 2047], v7 = s2[(idx + 224) & 2047];
            a0 += v0.x; a1 += v1.y; a2 += v2.x; a3 += v3.y; a4 += v4.x; a5 += v5.y; a6 += v6.x; a7 += v7.y;
            idx += 256;
        }
        if (MODE == 8) {  // 8 x (int->double convert)
            a0 += (double)(idx); a1 += (double)(idx + 1); a2 += (double)(idx + 2); a3 += (double)(idx + 3);
            a4 += (double)(idx + 4); a5 += (double)(idx + 5); a6 += (double)(idx + 6); a7 += (double)(idx + 7);
            idx += 9;
        }
        if (MODE == 9) {  // 8 DDIV
            a0 = __ddiv_rn(a0, c); a1 = __ddiv_rn(a1, c); a2 = __ddiv_rn(a2, c); a3 = __ddiv_rn(a3, c);
            a4 = __ddiv_rn(a4, c); a5 = __ddiv_rn(a5, c); a6 = __ddiv_rn(a6, c); a7 = __ddiv_rn(a7, c);
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, double ops_per_iter, int threads)
{
    double* out; long long* cyc;
    int nb = 148 * (2048 / threads);
    cudaMalloc(&out, sizeof(double) * nb * threads);
    cudaMalloc(&cyc, sizeof(long long) * nb);
    k<MODE><<<nb, threads>>>(out, cyc, 1.0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < 5; ++i) k<MODE><<<nb, threads>>>(out, cyc, 1.0);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
    long long h[4]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    double total = ops_per_iter * ITER * (double)nb * threads;
    double per_clk_sm = ops_per_iter * ITER * 2048.0 / (double)h[0];
    printf("%-28s thr/blk %4d: %.3f ms  %.2f Gop/s  cycles %lld  -> %.1f thread-ops/clk/SM  (eff clk %.0f MHz)\n", name, threads, ms,
           total / ms * 1e-6, h[0], per_clk_sm, h[0] / (ms * 1e3));
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    for (int t : {1024, 256}) {
        if (t == 1024) {
            run<0>("DADD x8", 8, 1024); run<1>("DMUL x8", 8, 1024); run<2>("DFMA x8", 8, 1024);
            run<3>("LDS.64 x8", 8, 1024); run<7>("LDS.128 x8", 8, 1024); run<4>("SHFL.f64 x8", 8, 1024);
            run<5>("DADD x8 + LDS.64 x4", 8, 1024); run<6>("DADD x8 + SHFL.f64 x4", 8, 1024);
            run<8>("I2F.F64 x8 (+DADD x8)", 8, 1024); run<9>("DDIV x8", 8, 1024);
        } else {
            run<0>("DADD x8", 8, 256); run<3>("LDS.64 x8", 8, 256);
        }
    }
    return 0;
}
