// Tensor memory (TMEM, 256 KB per SM: 128 lanes x 512 columns of 32 bits) used as per-thread scratch
// beside shared memory: state that a thread carries from one step of a marching kernel to the next
// lives here instead of in shared memory, so it costs neither shared-memory capacity nor LSU
// bandwidth.  Only the 32x32b shape is used: thread t of warp w reads / writes N consecutive columns
// of lane 32 (w % 4) + t; the four warps that share a lane quarter get disjoint column ranges.
// All calls are warp-collective (.sync.aligned): every lane of the warp must execute them.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ofb {

// Allocates n_cols columns (power of two, 32..512) for the CTA; the TMEM base address is written to
// *slot (shared memory).  One full warp calls this; dealloc is called by the same warp at the end.
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, int n_cols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(slot)),
                 "r"(n_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t base, int n_cols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(n_cols) : "memory");
}
__device__ __forceinline__ void tmem_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Address of column `col` for the calling warp: lane field = first lane of the warp's quarter.
__device__ __forceinline__ uint32_t tmem_addr(uint32_t base, int col) { return base + (((threadIdx.x >> 5) & 3u) << 21) + (uint32_t)col; }

// 4 doubles (8 columns) per thread.  The 64-bit values are split / joined inside the asm block so that
// ptxas can keep them in the register pairs they already live in (no per-word moves).
__device__ __forceinline__ void tmem_st4(uint32_t addr, const double (&v)[4])
{
    asm volatile(
        "{\n"
        ".reg .b32 a<8>;\n"
        "mov.b64 {a0, a1}, %1;\n"
        "mov.b64 {a2, a3}, %2;\n"
        "mov.b64 {a4, a5}, %3;\n"
        "mov.b64 {a6, a7}, %4;\n"
        "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {a0, a1, a2, a3, a4, a5, a6, a7};\n"
        "}\n" ::"r"(addr),
        "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3])
        : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t addr, double (&v)[4])
{
    asm volatile(
        "{\n"
        ".reg .b32 a<8>;\n"
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {a0, a1, a2, a3, a4, a5, a6, a7}, [%4];\n"
        "mov.b64 %0, {a0, a1};\n"
        "mov.b64 %1, {a2, a3};\n"
        "mov.b64 %2, {a4, a5};\n"
        "mov.b64 %3, {a6, a7};\n"
        "}\n"
        : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
        : "r"(addr)
        : "memory");
}

}  // namespace ofb
