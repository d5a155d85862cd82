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

// 4 doubles (8 columns) per thread
__device__ __forceinline__ void tmem_st4(uint32_t addr, const double (&v)[4])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(addr),
                 "r"(__double2loint(v[0])), "r"(__double2hiint(v[0])), "r"(__double2loint(v[1])), "r"(__double2hiint(v[1])),
                 "r"(__double2loint(v[2])), "r"(__double2hiint(v[2])), "r"(__double2loint(v[3])), "r"(__double2hiint(v[3]))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t addr, double (&v)[4])
{
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(addr)
                 : "memory");
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = __hiloint2double((int)r[2 * k + 1], (int)r[2 * k]);
}

}  // namespace ofb
