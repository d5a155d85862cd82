// Gaussian pyramid construction: cv::pyrDown as called by getGaussianPyramid
// (ParallelVideoPyr.cpp:566-575, pyrDown at :571).
//
//   hor[y][x] = s[y][2x]*6 + (s[y][R(2x-1)] + s[y][R(2x+1)])*4 + s[y][R(2x-2)] + s[y][R(2x+2)]
//   dst[y][x] = (hor[2y][x]*6 + (hor[R(2y-1)][x] + hor[R(2y+1)][x])*4
//                + hor[R(2y-2)][x] + hor[R(2y+2)][x]) * (1/256)
//
// R = BORDER_REFLECT_101.  All frames of a batch (both frames of every pair, or every
// frame of a stream) go through one launch: grid.z = image index.
//
// Each CTA produces a TX x TY output tile.  The (2TX+3) x (2TY+3) source patch is read
// once with coalesced loads into shared memory, the horizontal pass is evaluated once per
// (source row, output column) into a second shared plane, and the vertical pass reads that
// plane -- so every source value crosses HBM/L2 once per CTA and `hor` is never recomputed.
#include "of_device.cuh"
#include "of_kernels.h"

namespace ofb {

namespace {

constexpr int PD_TX = 64;   // output tile width
constexpr int PD_TY = 8;    // output tile height
constexpr int PD_NT = 256;
constexpr int PD_SW = 2 * PD_TX + 3;  // source patch width
constexpr int PD_SH = 2 * PD_TY + 3;  // source patch height

__device__ __forceinline__ int r101_any(int i, int n)
{
    if (n == 1) return 0;
    const int p = 2 * (n - 1);
    i %= p;
    if (i < 0) i += p;
    return i >= n ? p - i : i;
}

__global__ void __launch_bounds__(PD_NT) pyr_down_kernel(PyrDownArgs a)
{
    __shared__ double s_src[PD_SH][PD_SW + 1];
    __shared__ double s_hor[PD_SH][PD_TX];

    const int img = blockIdx.z;
    const double* __restrict__ src = a.src + (size_t)img * a.src_img_stride;
    double* __restrict__ dst = a.dst + (size_t)img * a.dst_img_stride;
    const int ow = (a.w + 1) >> 1, oh = (a.h + 1) >> 1;
    const int ox0 = blockIdx.x * PD_TX, oy0 = blockIdx.y * PD_TY;
    const int sx0 = 2 * ox0 - 2, sy0 = 2 * oy0 - 2;

    // Source patch, reflect-101 applied at load time.
    for (int i = threadIdx.x; i < PD_SH * PD_SW; i += PD_NT) {
        const int ry = i / PD_SW, rx = i - ry * PD_SW;
        const int y = r101_any(sy0 + ry, a.h), x = r101_any(sx0 + rx, a.w);
        s_src[ry][rx] = __ldg(src + (size_t)y * a.src_pitch + x);
    }
    __syncthreads();

    // Horizontal pass: one value per (patch row, output column).
    for (int i = threadIdx.x; i < PD_SH * PD_TX; i += PD_NT) {
        const int ry = i / PD_TX, tx = i - ry * PD_TX;
        const double* r = &s_src[ry][2 * tx];   // r[2] is source column 2*(ox0+tx)
        double hv = dmul(r[2], 6.0);
        hv = dadd(hv, dmul(dadd(r[1], r[3]), 4.0));
        hv = dadd(hv, r[0]);
        hv = dadd(hv, r[4]);
        s_hor[ry][tx] = hv;
    }
    __syncthreads();

    // Vertical pass.
    for (int i = threadIdx.x; i < PD_TY * PD_TX; i += PD_NT) {
        const int ty = i / PD_TX, tx = i - ty * PD_TX;
        const int oy = oy0 + ty, oxx = ox0 + tx;
        if (oy >= oh || oxx >= ow) continue;
        const int c = 2 * ty + 2;               // patch row of source row 2*oy
        double o = dmul(s_hor[c][tx], 6.0);
        o = dadd(o, dmul(dadd(s_hor[c - 1][tx], s_hor[c + 1][tx]), 4.0));
        o = dadd(o, s_hor[c - 2][tx]);
        o = dadd(o, s_hor[c + 2][tx]);
        dst[(size_t)oy * a.dst_pitch + oxx] = dmul(o, 1.0 / 256.0);
    }
}

}  // namespace

cudaError_t launch_pyr_down(const PyrDownArgs& a, cudaStream_t s, unsigned long long* launches)
{
    const int ow = (a.w + 1) / 2, oh = (a.h + 1) / 2;
    dim3 grid((ow + PD_TX - 1) / PD_TX, (oh + PD_TY - 1) / PD_TY, a.n_images);
    pyr_down_kernel<<<grid, PD_NT, 0, s>>>(a);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace ofb
