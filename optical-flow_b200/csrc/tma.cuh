// Bulk-tensor copies (TMA) and mbarriers for the sm_100a kernels: plain tile movement is handed to
// the copy engine (cp.async.bulk.tensor + mbarrier::complete_tx) instead of per-thread LDG with
// per-thread address arithmetic; out-of-range box elements arrive as 0.
//
// Host side: the tensor map is encoded through the driver entry point (no link against libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ofb {

// ------------------------------------------------------------------------------ host
typedef CUresult (*TmaEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline TmaEncodeTiledFn tma_encoder()
{
    static TmaEncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (TmaEncodeTiledFn)p;
    }();
    return fn;
}

// Batch of row-major double planes (w x h, row pitch / plane stride in doubles, n planes) as a rank-3
// tensor (x, y, plane) with a box of box_w x box_h x 1 elements.  The base must be 16-byte aligned and
// pitch / stride multiples of 2 doubles.
inline cudaError_t tma_map_planes_f64(CUtensorMap* m, const double* base, int w, int h, size_t pitch, size_t plane_stride,
                                      int n_planes, int box_w, int box_h)
{
    TmaEncodeTiledFn enc = tma_encoder();
    if (!enc) return cudaErrorNotSupported;
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (pitch & 1) || (plane_stride & 1) || (box_w & 1)) return cudaErrorInvalidValue;
    if (plane_stride == 0) plane_stride = pitch * (size_t)h;        // a single plane: any valid stride
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n_planes};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch * 8, (cuuint64_t)plane_stride * 8};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

// ------------------------------------------------------------------------------ device
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// makes the initialised barrier visible to the async proxy (the copy engine)
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MBAR_DONE;\n"
        "bra MBAR_WAIT;\n"
        "MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}

// One box of a rank-3 tensor map -> shared memory (128-byte aligned), completion on `bar`.  The box
// must start at a 16-byte aligned element of the innermost dimension (FP64: an even c0; an odd one
// raises "illegal instruction" on sm_100 -- tools/tma_probe.cu); it may start or end outside the
// tensor, the elements outside arrive as 0.
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
        : "memory");
}
#endif

}  // namespace ofb
