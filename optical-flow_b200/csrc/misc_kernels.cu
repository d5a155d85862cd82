// Small kernels beside the hot path:
//  * warp_level_kernel   -- materialises the warped level (pyrUp + map + remap), only used for
//                           the reference's in-place write-back quirk (warped.copyTo(pyr[i-1]),
//                           ParallelVideoPyr.cpp:696,701) when of_params.write_back_warped = 1;
//  * gesture_stats_kernel -- the live loop's flow statistics (:845-873) on existing planes;
//  * stats_finalize_kernel -- detected = count > count_thresh (:886).
#include "of_device.cuh"
#include "of_kernels.h"

namespace ofb {

namespace {

__global__ void __launch_bounds__(256) warp_level_kernel(WarpArgs a)
{
    const int x = blockIdx.x * 32 + (threadIdx.x & 31);
    const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= a.w || y >= a.h) return;
    // coarse rows/cols addressed directly in global memory (tile origin 0,0)
    const double upu = pyrup_at(a.cu, (int)a.c_pitch, 0, 0, y, x, a.wc, a.hc);
    const double upv = pyrup_at(a.cv, (int)a.c_pitch, 0, 0, y, x, a.wc, a.hc);
    float m1x, m1y, m2x, m2y;
    map_coords(x, y, upu, upv, m1x, m1y, m2x, m2y);
    const double val = a.sign > 0.0f ? remap_tap4(a.src, a.src_pitch, a.w, a.h, m1x, m1y)
                                     : remap_tap4(a.src, a.src_pitch, a.w, a.h, m2x, m2y);
    a.dst[(size_t)y * a.dst_pitch + x] = val;
}

__global__ void __launch_bounds__(256) gesture_stats_kernel(StatsArgs a)
{
    __shared__ double s_red[3];
    const int pair = blockIdx.z;
    const double* __restrict__ u = a.u + (size_t)pair * a.uv_pair_stride;
    const double* __restrict__ v = a.v + (size_t)pair * a.uv_pair_stride;
    if (threadIdx.x < 3) s_red[threadIdx.x] = 0.0;
    __syncthreads();
    double sx = 0.0, sy = 0.0, sc = 0.0;
    for (int y = blockIdx.y; y < a.h; y += gridDim.y)
        for (int x = blockIdx.x * 256 + threadIdx.x; x < a.w; x += gridDim.x * 256) {
            const double fu = __ldg(u + (size_t)y * a.uv_pitch + x);
            const double fv = __ldg(v + (size_t)y * a.uv_pitch + x);
            const double val = sqrt(dadd(dmul(fu, fu), dmul(fv, fv)));
            if (!(val < a.mag_thresh)) { sx += (double)x; sy += (double)y; sc += 1.0; }
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sc += __shfl_xor_sync(0xffffffffu, sc, o);
    }
    if ((threadIdx.x & 31) == 0 && sc != 0.0) {
        atomicAdd(&s_red[0], sx);
        atomicAdd(&s_red[1], sy);
        atomicAdd(&s_red[2], sc);
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_red[2] != 0.0) {
        atomicAdd(&a.stats[pair].sum_x, s_red[0]);
        atomicAdd(&a.stats[pair].sum_y, s_red[1]);
        atomicAdd(&a.stats[pair].count, s_red[2]);
    }
}

__global__ void stats_finalize_kernel(of_gesture_stats* stats, int n, double count_thresh)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) stats[i].detected = stats[i].count > count_thresh ? 1.0 : 0.0;
}

}  // namespace

cudaError_t launch_warp_level(const WarpArgs& a, cudaStream_t s, unsigned long long* launches)
{
    dim3 grid((a.w + 31) / 32, (a.h + 7) / 8, 1);
    warp_level_kernel<<<grid, 256, 0, s>>>(a);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_gesture_stats(const StatsArgs& a, int n_pairs, cudaStream_t s,
                                 unsigned long long* launches)
{
    int gx = (a.w + 255) / 256;
    if (gx > 8) gx = 8;
    int gy = a.h < 64 ? a.h : 64;
    dim3 grid(gx, gy, n_pairs);
    gesture_stats_kernel<<<grid, 256, 0, s>>>(a);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_stats_finalize(of_gesture_stats* stats, int n_pairs, double count_thresh,
                                  cudaStream_t s, unsigned long long* launches)
{
    stats_finalize_kernel<<<(n_pairs + 127) / 128, 128, 0, s>>>(stats, n_pairs, count_thresh);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

// cv::Mat::convertTo(CV_64FC1, alpha, beta) of 8-bit frames (the harness mains' img.convertTo(img,
// CV_64FC1, 1.0/255), ParallelBasicPyr.cpp:729-730): dst = rn(rn(v * alpha) + beta), v exact in double.
// One thread = 4 consecutive pixels of one row.
namespace {
__global__ void __launch_bounds__(256) u8_to_f64_kernel(const unsigned char* __restrict__ src, size_t src_plane,
                                                         size_t src_step, double* __restrict__ dst, size_t dst_plane,
                                                         size_t dst_pitch, int w, int h, double alpha, double beta)
{
    const int x4 = (blockIdx.x * 64 + (threadIdx.x & 63)) * 4;
    const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (y >= h || x4 >= w) return;
    const unsigned char* s = src + (size_t)blockIdx.z * src_plane + (size_t)y * src_step + x4;
    double* d = dst + (size_t)blockIdx.z * dst_plane + (size_t)y * dst_pitch + x4;
    if (x4 + 4 <= w && ((reinterpret_cast<uintptr_t>(s) & 3) | (reinterpret_cast<uintptr_t>(d) & 15)) == 0) {
        const uint32_t p = __ldg(reinterpret_cast<const uint32_t*>(s));
        double2 a, b;
        a.x = dadd(dmul((double)(p & 255u), alpha), beta);
        a.y = dadd(dmul((double)((p >> 8) & 255u), alpha), beta);
        b.x = dadd(dmul((double)((p >> 16) & 255u), alpha), beta);
        b.y = dadd(dmul((double)(p >> 24), alpha), beta);
        reinterpret_cast<double2*>(d)[0] = a;
        reinterpret_cast<double2*>(d)[1] = b;
    } else {
        for (int i = 0; i < 4 && x4 + i < w; ++i) d[i] = dadd(dmul((double)s[i], alpha), beta);
    }
}
}  // namespace

cudaError_t launch_u8_to_f64(const unsigned char* src, size_t src_plane, size_t src_step, double* dst, size_t dst_plane,
                             size_t dst_pitch, int w, int h, int n_frames, double alpha, double beta, cudaStream_t s,
                             unsigned long long* launches)
{
    dim3 grid((w + 255) / 256, (h + 3) / 4, n_frames);
    u8_to_f64_kernel<<<grid, 256, 0, s>>>(src, src_plane, src_step, dst, dst_plane, dst_pitch, w, h, alpha, beta);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace ofb
