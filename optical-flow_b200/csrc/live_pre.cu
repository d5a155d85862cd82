// Pre-processing of the reference's live loop between the colour frame and the LK input
// (ParallelVideoPyr.cpp:791-813), the "next" row SURVEY.md 8(f)-2:
//
//   bgr2gray_kernel   cv::cvtColor(CV_BGR2GRAY) :791 -- 8-bit fixed point
//   live_diff_kernel  diff = cur - 0.3 * prev (8-bit, saturating) :800, threshold(TOZERO) :803,
//                     Sobel x + Sobel y -> double :808-811, dilate x n :812, erode x n :813,
//                     fused: one tile of the two gray frames in, one tile of doubles out; the
//                     difference image, the Sobel sum and the dilated image never reach HBM.
//
// HBM traffic of live_diff_kernel per pixel: read 2 bytes (two gray frames, the tile halo comes
// from L2), write 8 = 10 B algorithmic.  resize (:782) and GaussianBlur (:790) stay with the
// capture front end (host OpenCV); they are not on this path.
#include "of_device.cuh"
#include "of_kernels.h"

namespace ofb {

namespace {

// ------------------------------------------------------------------------------ BGR -> gray
// (B*cb + G*cg + R*cr + 2^(bits-1)) >> bits; one thread = 4 pixels (12 bytes in, 4 out).
__global__ void __launch_bounds__(256) bgr2gray_kernel(const unsigned char* __restrict__ bgr, size_t bgr_step,
                                                        unsigned char* __restrict__ gray, size_t gray_step,
                                                        int w, int h, int cb, int cg, int cr, int bits)
{
    const int x4 = (blockIdx.x * 64 + (threadIdx.x & 63)) * 4;
    const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (y >= h || x4 >= w) return;
    const unsigned char* src = bgr + (size_t)y * bgr_step + (size_t)x4 * 3;
    unsigned char* dst = gray + (size_t)y * gray_step + x4;
    const int half = 1 << (bits - 1);
    if (x4 + 4 <= w && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 3) == 0) {
        const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
        const uint32_t a = __ldg(s32), b = __ldg(s32 + 1), c = __ldg(s32 + 2);
        // bytes: a = B0 G0 R0 B1 | b = G1 R1 B2 G2 | c = R2 B3 G3 R3
        const int g0 = (int)(((a & 255) * cb + ((a >> 8) & 255) * cg + ((a >> 16) & 255) * cr + half) >> bits);
        const int g1 = (int)(((a >> 24) * cb + (b & 255) * cg + ((b >> 8) & 255) * cr + half) >> bits);
        const int g2 = (int)((((b >> 16) & 255) * cb + (b >> 24) * cg + (c & 255) * cr + half) >> bits);
        const int g3 = (int)((((c >> 8) & 255) * cb + ((c >> 16) & 255) * cg + (c >> 24) * cr + half) >> bits);
        *reinterpret_cast<uint32_t*>(dst) = (uint32_t)g0 | ((uint32_t)g1 << 8) | ((uint32_t)g2 << 16) | ((uint32_t)g3 << 24);
    } else {
        for (int i = 0; i < 4 && x4 + i < w; ++i)
            dst[i] = (unsigned char)((src[3 * i] * cb + src[3 * i + 1] * cg + src[3 * i + 2] * cr + half) >> bits);
    }
}

// ------------------------------------------------------------------------------ :800-813
constexpr int LT = 32;           // output tile (LT x LT), 256 threads

template <int N>
struct LiveCfg {
    static constexpr int S = LT + 4 * N;          // Sobel-sum tile side (halo 2N: dilate N + erode N)
    static constexpr int DS = S + 2;              // difference-image tile side (+1 for the 3x3 Sobel)
    static constexpr int DP = (DS + 3) & ~3;      // its row pitch in bytes
    static constexpr int M = LT + 2 * N;          // dilated tile side
    static constexpr size_t SMEM = (size_t)(S * S + S * M + M * M) * sizeof(double) + (size_t)DS * DP;
};

__device__ __forceinline__ int reflect101_clamped(int i, int n)
{
    if (n == 1) return 0;
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return min(max(i, 0), n - 1);                 // far outside: value unused, index kept legal
}

// saturate_cast<uchar>(cur * 1.f + prev * beta) with float32 products and sum, round half to
// even (cv::addWeighted on CV_8U), then THRESH_TOZERO.
__device__ __forceinline__ int diff_pixel(int cur, int prev, float beta, int thresh)
{
    const float t = __fadd_rn(__fmul_rn((float)cur, 1.0f), __fmul_rn((float)prev, beta));
    int d = __float2int_rn(t);
    d = min(max(d, 0), 255);
    return d > thresh ? d : 0;
}

template <int N>
__global__ void __launch_bounds__(256) live_diff_kernel(const unsigned char* __restrict__ cur,
                                                         const unsigned char* __restrict__ prev, size_t gstep,
                                                         int w, int h, float beta, int thresh,
                                                         double* __restrict__ out, size_t opitch)
{
    using C = LiveCfg<N>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* s_s = reinterpret_cast<double*>(smem_raw);     // Sobel sum, S x S (-inf outside the image)
    double* s_t = s_s + C::S * C::S;                       // row pass, S rows x M columns
    double* s_d = s_t + C::S * C::M;                       // dilated, M x M (+inf outside the image)
    unsigned char* s_df = reinterpret_cast<unsigned char*>(s_d + C::M * C::M);   // difference image, DS x DP
    const int tx0 = blockIdx.x * LT, ty0 = blockIdx.y * LT;
    const int tid = threadIdx.x;
    const double ninf = __longlong_as_double(0xfff0000000000000ll), pinf = __longlong_as_double(0x7ff0000000000000ll);

    // thresholded difference image with the border the 3x3 Sobel sees (BORDER_REFLECT_101)
    for (int i = tid; i < C::DS * C::DS; i += 256) {
        const int r = i / C::DS, c = i - r * C::DS;
        const int y = reflect101_clamped(ty0 - 2 * N - 1 + r, h), x = reflect101_clamped(tx0 - 2 * N - 1 + c, w);
        const size_t o = (size_t)y * gstep + x;
        s_df[r * C::DP + c] = (unsigned char)diff_pixel(__ldg(cur + o), __ldg(prev + o), beta, thresh);
    }
    __syncthreads();
    // Sobel x + Sobel y (integer valued, exact in double)
    for (int i = tid; i < C::S * C::S; i += 256) {
        const int r = i / C::S, c = i - r * C::S;
        const int y = ty0 - 2 * N + r, x = tx0 - 2 * N + c;
        double v = ninf;
        if ((unsigned)y < (unsigned)h && (unsigned)x < (unsigned)w) {
            const unsigned char* p = s_df + r * C::DP + c;        // (r, c) is the pixel's upper-left neighbour
            const int a00 = p[0], a01 = p[1], a02 = p[2];
            const int a10 = p[C::DP], a12 = p[C::DP + 2];
            const int a20 = p[2 * C::DP], a21 = p[2 * C::DP + 1], a22 = p[2 * C::DP + 2];
            const int sx = (a02 + 2 * a12 + a22) - (a00 + 2 * a10 + a20);
            const int sy = (a20 + 2 * a21 + a22) - (a00 + 2 * a01 + a02);
            v = (double)(sx + sy);
        }
        s_s[i] = v;
    }
    __syncthreads();
    // dilate: max over (2N+1)^2, rows then columns
    for (int i = tid; i < C::S * C::M; i += 256) {
        const int r = i / C::M, c = i - r * C::M;
        const double* p = s_s + r * C::S + c;
        double m = p[0];
#pragma unroll
        for (int k = 1; k <= 2 * N; ++k) m = fmax(m, p[k]);
        s_t[i] = m;
    }
    __syncthreads();
    for (int i = tid; i < C::M * C::M; i += 256) {
        const int r = i / C::M, c = i - r * C::M;
        const int y = ty0 - N + r, x = tx0 - N + c;
        const double* p = s_t + r * C::M + c;
        double m = p[0];
#pragma unroll
        for (int k = 1; k <= 2 * N; ++k) m = fmax(m, p[k * C::M]);
        s_d[i] = ((unsigned)y < (unsigned)h && (unsigned)x < (unsigned)w) ? m : pinf;
    }
    __syncthreads();
    // erode: min over (2N+1)^2, rows then columns (row pass into s_t, M rows x LT columns)
    for (int i = tid; i < C::M * LT; i += 256) {
        const int r = i / LT, c = i - r * LT;
        const double* p = s_d + r * C::M + c;
        double m = p[0];
#pragma unroll
        for (int k = 1; k <= 2 * N; ++k) m = fmin(m, p[k]);
        s_t[i] = m;
    }
    __syncthreads();
    for (int i = tid; i < LT * LT; i += 256) {
        const int r = i / LT, c = i - r * LT;
        const int y = ty0 + r, x = tx0 + c;
        if (y >= h || x >= w) continue;
        const double* p = s_t + r * LT + c;
        double m = p[0];
#pragma unroll
        for (int k = 1; k <= 2 * N; ++k) m = fmin(m, p[k * LT]);
        out[(size_t)y * opitch + x] = m;
    }
}

template <int N>
cudaError_t launch_live_n(const LiveDiffArgs& a, cudaStream_t s, unsigned long long* launches)
{
    using C = LiveCfg<N>;
    static SmemOptIn opt_in;
    if (C::SMEM > 48 * 1024) {
        cudaError_t e = opt_in.ensure((const void*)live_diff_kernel<N>, C::SMEM);
        if (e != cudaSuccess) return e;
    }
    dim3 grid((a.w + LT - 1) / LT, (a.h + LT - 1) / LT, 1);
    live_diff_kernel<N><<<grid, 256, C::SMEM, s>>>(a.cur, a.prev, a.gray_step, a.w, a.h, a.beta, a.thresh, a.out, a.out_pitch);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_bgr2gray(const unsigned char* bgr, size_t bgr_step, unsigned char* gray, size_t gray_step, int w,
                            int h, int bits, cudaStream_t s, unsigned long long* launches)
{
    const int cb = bits == 14 ? 1868 : 3735, cg = bits == 14 ? 9617 : 19235, cr = bits == 14 ? 4899 : 9798;
    dim3 grid((w + 255) / 256, (h + 3) / 4, 1);
    bgr2gray_kernel<<<grid, 256, 0, s>>>(bgr, bgr_step, gray, gray_step, w, h, cb, cg, cr, bits);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_live_diff(const LiveDiffArgs& a, cudaStream_t s, unsigned long long* launches)
{
    switch (a.iterations) {
        case 1: return launch_live_n<1>(a, s, launches);
        case 2: return launch_live_n<2>(a, s, launches);
        case 3: return launch_live_n<3>(a, s, launches);
        case 4: return launch_live_n<4>(a, s, launches);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace ofb
