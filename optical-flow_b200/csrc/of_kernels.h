// Host-visible launch interface of the sm_100a kernels (internal to the library; the
// public boundary is include/of_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

#include <atomic>

#include "../../include/of_b200.h"

namespace ofb {

// One pyramid level of a batch of frame pairs.  Pitches and strides are in doubles
// (mask: bytes).  Pair k uses base + k * *_pair_stride.
struct LevelArgs {
    const double* A;          // frame 1 at this level (unwarped pyramid level)
    const double* B;          // frame 2 at this level
    size_t img_pitch;
    size_t a_pair_stride;
    size_t b_pair_stride;
    const double* cu;         // accumulated flow of the next-coarser level, or nullptr
    const double* cv;
    size_t c_pitch;
    size_t c_pair_stride;
    int wc, hc;               // coarse size (w/2, h/2)
    double* u;                // accumulated flow at this level (may be nullptr: stats only)
    double* v;
    size_t uv_pitch;
    size_t uv_pair_stride;
    unsigned char* mask;      // may be nullptr
    size_t mask_pitch;
    size_t mask_pair_stride;
    of_gesture_stats* stats;  // may be nullptr; accumulated with atomics (pre-zeroed)
    double mag_thresh;
    double mag2_thresh;       // smallest s with !(sqrt(s) < mag_thresh): lets kernels skip the sqrt
    int w, h;
    int radius;
    double det_eps;
};

struct PyrDownArgs {
    const double* src;
    size_t src_pitch;
    size_t src_img_stride;
    double* dst;
    size_t dst_pitch;
    size_t dst_img_stride;
    int w, h;                 // source size; destination is ((w+1)/2) x ((h+1)/2)
    int n_images;
};

struct WarpArgs {             // write-back of the warped level (quirk Q1), optional path
    const double* src;
    size_t src_pitch;
    const double* cu;
    const double* cv;
    size_t c_pitch;
    int wc, hc;
    double* dst;
    size_t dst_pitch;
    int w, h;
    float sign;               // +1: frame 1 (x + f/2), -1: frame 2 (x - f/2)
};

struct StatsArgs {
    const double* u;
    const double* v;
    size_t uv_pitch;
    size_t uv_pair_stride;
    int w, h;
    of_gesture_stats* stats;
    double mag_thresh;
};

struct LiveDiffArgs {         // live-loop pre-processing :800-813 (live_pre.cu)
    const unsigned char* cur;
    const unsigned char* prev;
    size_t gray_step;         // bytes
    int w, h;
    float beta;               // (float)(-learning_rate)
    int thresh;               // THRESH_TOZERO threshold
    int iterations;           // dilate / erode iterations (1..4)
    double* out;
    size_t out_pitch;         // doubles
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute, and one process may drive
// several devices from several threads (one context per device): remember the opt-in per device.
struct SmemOptIn {
    std::atomic<unsigned long long> done{0};      // bit d: set on device d (devices >= 64: set every time)
    cudaError_t ensure(const void* func, size_t bytes)
    {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev < 64 && ((done.load(std::memory_order_acquire) >> dev) & 1ull)) return cudaSuccess;
        e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e == cudaSuccess && dev < 64) done.fetch_or(1ull << dev, std::memory_order_release);
        return e;
    }
};

// Number of SMs of the current device (cached per device).
inline int current_sm_count()
{
    static std::atomic<int> cache[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev < 64) {
        const int c = cache[dev].load(std::memory_order_relaxed);
        if (c > 0) return c;
    }
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    if (dev < 64) cache[dev].store(n, std::memory_order_relaxed);
    return n;
}

// Each returns the CUDA error of the launch and adds the number of kernels it launched
// to *launches.
cudaError_t launch_pyr_down(const PyrDownArgs& a, cudaStream_t s, unsigned long long* launches);
cudaError_t launch_lk_level(const LevelArgs& a, int n_pairs, cudaStream_t s,
                            unsigned long long* launches);
cudaError_t launch_warp_level(const WarpArgs& a, cudaStream_t s, unsigned long long* launches);
cudaError_t launch_gesture_stats(const StatsArgs& a, int n_pairs, cudaStream_t s,
                                 unsigned long long* launches);
cudaError_t launch_bgr2gray(const unsigned char* bgr, size_t bgr_step, unsigned char* gray, size_t gray_step, int w,
                            int h, int bits, cudaStream_t s, unsigned long long* launches);
cudaError_t launch_live_diff(const LiveDiffArgs& a, cudaStream_t s, unsigned long long* launches);
// live-loop colour-frame steps :782, :790 (live_frame.cu).  Tables are host-built and device-resident.
void build_cubic_table(int src_n, int dst_n, int* ofs, short* co);
bool build_gauss_kernel_q8(int ksize, double sigma, int* out);
cudaError_t launch_resize_cubic(const unsigned char* src, size_t sstep, int cn, unsigned char* dst, size_t dstep, int dw,
                                int dh, const int* xofs, const short* xco, const int* yofs, const float* ybeta,
                                cudaStream_t s, unsigned long long* launches);
// gray_bits = 0: blurred image out (cn channels); 14 / 15: cn = 3, cvtColor(BGR2GRAY) fused, one channel out
cudaError_t launch_gauss_u8(const unsigned char* src, size_t sstep, int cn, unsigned char* dst, size_t dstep, int w, int h,
                            const int* kernel, int ksize, int gray_bits, cudaStream_t s, unsigned long long* launches);
// frames of 8-bit pixels (dense planes src_plane bytes apart) -> double planes (dst_plane doubles apart)
cudaError_t launch_u8_to_f64(const unsigned char* src, size_t src_plane, size_t src_step, double* dst, size_t dst_plane,
                             size_t dst_pitch, int w, int h, int n_frames, double alpha, double beta, cudaStream_t s,
                             unsigned long long* launches);
cudaError_t launch_stats_finalize(of_gesture_stats* stats, int n_pairs, double count_thresh,
                                  cudaStream_t s, unsigned long long* launches);

}  // namespace ofb
