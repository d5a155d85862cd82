// Development aid: one bulk-tensor load of a row of doubles, to pin down which tensor-map encodings the
// device accepts.  nvcc -gencode arch=compute_100a,code=sm_100a -I optical-flow_b200/csrc tools/tma_probe.cu
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include "tma.cuh"
using namespace ofb;

__global__ void probe(const __grid_constant__ CUtensorMap tm, double* out, int c0, int c1, int c2, int n, uint32_t bytes)
{
    extern __shared__ __align__(128) double sm[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 512);
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar, bytes);
        tma_load_3d(sm, &tm, c0, c1, c2, bar);
    }
    mbar_wait(bar, 0);
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = sm[i];
}

int main(int argc, char** argv)
{
    const int only = argc > 1 ? atoi(argv[1]) : -1;
    const int w = 100, h = 20, np = 3, pitch = 112;
    const size_t plane = (size_t)pitch * h;
    double* d; double* o;
    cudaMalloc(&d, plane * np * 8); cudaMalloc(&o, 1024 * 8);
    double* hbuf = new double[plane * np];
    for (size_t i = 0; i < plane * np; ++i) hbuf[i] = (double)i;
    cudaMemcpy(d, hbuf, plane * np * 8, cudaMemcpyHostToDevice);
    for (int variant = 0; variant < 3; ++variant) {
        if (only >= 0 && variant != only) continue;
        CUtensorMap tm; memset(&tm, 0, sizeof tm);
        cudaError_t e; int boxw = 66; uint32_t bytes = 66 * 8; int c0 = -5;
        if (variant == 0) e = tma_map_planes_f64(&tm, d, w, h, pitch, plane, np, 66, 1);
        else if (variant == 1) {
            boxw = argc > 2 ? atoi(argv[2]) : 64; c0 = argc > 3 ? atoi(argv[3]) : 4; bytes = boxw * 8;
            e = tma_map_planes_f64(&tm, d, w, h, pitch, plane, np, boxw, 1);
            printf("box %d c0 %d\n", boxw, c0);
        }
        else {
            // the same planes described as 32-bit words
            TmaEncodeTiledFn enc = tma_encoder();
            const cuuint64_t dims[3] = {2 * w, (cuuint64_t)h, (cuuint64_t)np};
            const cuuint64_t strides[2] = {(cuuint64_t)pitch * 8, plane * 8};
            const cuuint32_t box[3] = {132, 1, 1};
            const cuuint32_t es[3] = {1, 1, 1};
            CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            e = r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue; c0 = -10;
        }
        printf("variant %d encode: %s\n", variant, cudaGetErrorString(e));
        if (e != cudaSuccess) continue;
        cudaMemset(o, 0, 1024 * 8);
        probe<<<1, 128, 8192>>>(tm, o, c0, 3, 1, boxw, bytes);
        e = cudaDeviceSynchronize();
        printf("variant %d run: %s\n", variant, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        double res[80]; cudaMemcpy(res, o, boxw * 8, cudaMemcpyDeviceToHost);
        printf("  first: %.0f %.0f %.0f %.0f %.0f %.0f %.0f ... last %.0f (expect row 3 of plane 1: base %zu)\n", res[0], res[1], res[2], res[3],
               res[4], res[5], res[6], res[boxw - 1], plane + 3 * pitch);
    }
    return 0;
}
