// The two steps of the reference's live loop on the 8-bit colour frame, before cvtColor
// (ParallelVideoPyr.cpp:782, :790):
//
//   resize_cubic_kernel   cv::resize(frame, Size(1080, 1080), 0, 0, INTER_CUBIC)        :782
//   gauss_kernel          cv::GaussianBlur(Size(9, 9), 1.5, 1.5)                        :790
//                         (optionally fused with cvtColor(CV_BGR2GRAY) :791: the blurred colour
//                          frame then never reaches HBM)
//
// Both are OpenCV-defined fixed-point arithmetic on bytes, written out below; the tests pin them (through
// the CPU restatement under oracle/) against cv2 4.13 -- GaussianBlur: also OpenCV 2.4's integer kernel;
// resize: cv2's own code path, IPP off (DESIGN.md states the pin status).
//
//   resize: per destination column / row four source indices (clamped) and four 11-bit coefficients
//           (cv::interpolateCubic, A = -0.75, float32; tables built on the host in the same float32
//           arithmetic and cached per geometry); horizontal pass in int32, vertical pass in float32:
//           acc = S3 b3; acc = S2 b2 + acc; acc = S1 b1 + acc; acc = S0 b0 + acc with b_k = beta_k / 2^22,
//           every product and sum rounded (no FMA), then round half to even and saturate.
//   blur:   integer kernel k[9] = {2, 9, 28, 55, 68, 55, 28, 9, 2} (sum 256) in both passes,
//           BORDER_REFLECT_101, one rounding at the end: (sum + 2^15) >> 16.
//
// Memory-bound byte work: a thread produces one pixel (all channels); loads are L1 / L2 hits
// (every source byte is read by up to 16 / 81 neighbours), stores are coalesced.
#include "of_device.cuh"
#include "of_kernels.h"

#include <math.h>

namespace ofb {

namespace {

// ------------------------------------------------------------------------------ resize
template <int CN>
__global__ void __launch_bounds__(256) resize_cubic_kernel(const unsigned char* __restrict__ src, size_t sstep,
                                                            unsigned char* __restrict__ dst, size_t dstep, int dw, int dh,
                                                            const int* __restrict__ xofs, const short* __restrict__ xco,
                                                            const int* __restrict__ yofs, const float* __restrict__ ybeta)
{
    const int x = blockIdx.x * 32 + (threadIdx.x & 31);
    const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= dw || y >= dh) return;
    int xo[4], xc[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        xo[k] = __ldg(xofs + 4 * x + k) * CN;
        xc[k] = __ldg(xco + 4 * x + k);
    }
    float acc[CN];
#pragma unroll
    for (int r = 3; r >= 0; --r) {                           // rows 3, 2, 1, 0: cv2's accumulation order
        const unsigned char* row = src + (size_t)__ldg(yofs + 4 * y + r) * sstep;
        const float b = __ldg(ybeta + 4 * y + r);
#pragma unroll
        for (int c = 0; c < CN; ++c) {
            const int hsum = (int)__ldg(row + xo[0] + c) * xc[0] + (int)__ldg(row + xo[1] + c) * xc[1] +
                             (int)__ldg(row + xo[2] + c) * xc[2] + (int)__ldg(row + xo[3] + c) * xc[3];
            const float t = __fmul_rn((float)hsum, b);
            acc[c] = r == 3 ? t : __fadd_rn(t, acc[c]);
        }
    }
    unsigned char* d = dst + (size_t)y * dstep + (size_t)x * CN;
#pragma unroll
    for (int c = 0; c < CN; ++c) d[c] = (unsigned char)min(max(__float2int_rn(acc[c]), 0), 255);
}

// ------------------------------------------------------------------------------ blur (+ gray)
constexpr int GT_W = 64, GT_H = 16;      // output tile, 256 threads: 4 pixels each

struct GaussK { int k[31]; int n; };

// GRAY: the epilogue converts the blurred BGR pixel to gray (cvtColor :791) and stores one byte.
template <int CN, bool GRAY>
__global__ void __launch_bounds__(256) gauss_kernel(const unsigned char* __restrict__ src, size_t sstep,
                                                     unsigned char* __restrict__ dst, size_t dstep, int w, int h,
                                                     GaussK gk, int cb, int cg, int cr, int bits)
{
    extern __shared__ __align__(16) unsigned char g_smem[];
    const int r = gk.n >> 1;
    const int tw = GT_W + 2 * r, th = GT_H + 2 * r;
    unsigned char* s_src = g_smem;                                        // th x tw x CN bytes
    unsigned short* s_row = reinterpret_cast<unsigned short*>(g_smem + (((size_t)th * tw * CN + 15) & ~(size_t)15));  // th x GT_W x CN
    const int x0 = blockIdx.x * GT_W, y0 = blockIdx.y * GT_H;
    for (int i = threadIdx.x; i < th * tw; i += 256) {
        const int ty = i / tw, tx = i - ty * tw;
        int sy = y0 - r + ty, sx = x0 - r + tx;
        // BORDER_REFLECT_101, repeated for images narrower than the radius; far outside: unused
        sy = h == 1 ? 0 : (sy < 0 ? -sy : sy);
        sx = w == 1 ? 0 : (sx < 0 ? -sx : sx);
        if (h > 1) { const int p = 2 * (h - 1); sy %= p; if (sy >= h) sy = p - sy; }
        if (w > 1) { const int p = 2 * (w - 1); sx %= p; if (sx >= w) sx = p - sx; }
        const unsigned char* s = src + (size_t)sy * sstep + (size_t)sx * CN;
#pragma unroll
        for (int c = 0; c < CN; ++c) s_src[(size_t)i * CN + c] = __ldg(s + c);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < th * GT_W; i += 256) {
        const int ty = i / GT_W, tx = i - ty * GT_W;
        const unsigned char* p = s_src + ((size_t)ty * tw + tx) * CN;
        int acc[CN];
#pragma unroll
        for (int c = 0; c < CN; ++c) acc[c] = 0;
        for (int k = 0; k < gk.n; ++k)
#pragma unroll
            for (int c = 0; c < CN; ++c) acc[c] += (int)p[k * CN + c] * gk.k[k];
#pragma unroll
        for (int c = 0; c < CN; ++c) s_row[(size_t)i * CN + c] = (unsigned short)acc[c];    // <= 255 * 256
    }
    __syncthreads();
    for (int i = threadIdx.x; i < GT_H * GT_W; i += 256) {
        const int ty = i / GT_W, tx = i - ty * GT_W;
        const int x = x0 + tx, y = y0 + ty;
        if (x >= w || y >= h) continue;
        const unsigned short* p = s_row + ((size_t)ty * GT_W + tx) * CN;
        int acc[CN];
#pragma unroll
        for (int c = 0; c < CN; ++c) acc[c] = 0;
        for (int k = 0; k < gk.n; ++k)
#pragma unroll
            for (int c = 0; c < CN; ++c) acc[c] += (int)p[(size_t)k * GT_W * CN + c] * gk.k[k];
        int v[CN];
#pragma unroll
        for (int c = 0; c < CN; ++c) v[c] = min((acc[c] + (1 << 15)) >> 16, 255);
        if constexpr (GRAY) {
            static_assert(CN == 3, "gray epilogue needs BGR");
            dst[(size_t)y * dstep + x] = (unsigned char)((v[0] * cb + v[1] * cg + v[2] * cr + (1 << (bits - 1))) >> bits);
        } else {
            unsigned char* d = dst + (size_t)y * dstep + (size_t)x * CN;
#pragma unroll
            for (int c = 0; c < CN; ++c) d[c] = (unsigned char)v[c];
        }
    }
}

size_t gauss_smem(int n, int cn)
{
    const int r = n >> 1, tw = GT_W + 2 * r, th = GT_H + 2 * r;
    return (((size_t)th * tw * cn + 15) & ~(size_t)15) + (size_t)th * GT_W * cn * sizeof(unsigned short);
}

}  // namespace

// cv::interpolateCubic (A = -0.75) in float32, OpenCV's operation order; the host compiler must not
// contract these (x86-64 baseline has no FMA; -ffp-contract=off is passed for other targets).
static void cubic_coeffs(float x, float c[4])
{
    const float A = -0.75f;
    c[0] = ((A * (x + 1.0f) - 5.0f * A) * (x + 1.0f) + 8.0f * A) * (x + 1.0f) - 4.0f * A;
    c[1] = ((A + 2.0f) * x - (A + 3.0f)) * x * x + 1.0f;
    c[2] = ((A + 2.0f) * (1.0f - x) - (A + 3.0f)) * (1.0f - x) * (1.0f - x) + 1.0f;
    c[3] = 1.0f - c[0] - c[1] - c[2];
}

void build_cubic_table(int src_n, int dst_n, int* ofs, short* co)
{
    const double scale = 1.0 / ((double)dst_n / src_n);
    for (int d = 0; d < dst_n; ++d) {
        float fx = (float)((d + 0.5) * scale - 0.5);
        const int sx = (int)floorf(fx);
        fx -= (float)sx;
        float c[4];
        cubic_coeffs(fx, c);
        for (int k = 0; k < 4; ++k) {
            int s = sx - 1 + k;
            s = s < 0 ? 0 : (s > src_n - 1 ? src_n - 1 : s);
            ofs[4 * d + k] = s;
            co[4 * d + k] = (short)lrintf(c[k] * 2048.0f);          // cvRound: half to even
        }
    }
}

cudaError_t launch_resize_cubic(const unsigned char* src, size_t sstep, int cn, unsigned char* dst, size_t dstep, int dw,
                                int dh, const int* xofs, const short* xco, const int* yofs, const float* ybeta,
                                cudaStream_t s, unsigned long long* launches)
{
    dim3 grid((dw + 31) / 32, (dh + 7) / 8, 1);
    if (cn == 3) resize_cubic_kernel<3><<<grid, 256, 0, s>>>(src, sstep, dst, dstep, dw, dh, xofs, xco, yofs, ybeta);
    else if (cn == 1) resize_cubic_kernel<1><<<grid, 256, 0, s>>>(src, sstep, dst, dstep, dw, dh, xofs, xco, yofs, ybeta);
    else return cudaErrorInvalidValue;
    if (launches) *launches += 1;
    return cudaGetLastError();
}

// Fixed-point Gaussian kernel of cv::GaussianBlur on CV_8U: exp(-x^2 / (2 sigma^2)) normalised, times 256,
// rounded half to even with the rounding error carried to the next tap; symmetric.
bool build_gauss_kernel_q8(int ksize, double sigma, int* out)
{
    if (ksize < 1 || ksize > 31 || !(ksize & 1) || !(sigma > 0.0)) return false;
    double k[31], sum = 0.0;
    for (int i = 0; i < ksize; ++i) {
        const double x = i - (ksize - 1) / 2.0;
        k[i] = exp(-x * x / (2.0 * sigma * sigma));
        sum += k[i];
    }
    double err = 0.0;
    for (int i = 0; i <= ksize / 2; ++i) {
        const double adj = k[i] / sum * 256.0 + err;
        const double v = nearbyint(adj);
        err = adj - v;
        out[i] = out[ksize - 1 - i] = (int)v;
    }
    return true;
}

cudaError_t launch_gauss_u8(const unsigned char* src, size_t sstep, int cn, unsigned char* dst, size_t dstep, int w, int h,
                            const int* kernel, int ksize, int gray_bits, cudaStream_t s, unsigned long long* launches)
{
    GaussK gk{};
    gk.n = ksize;
    for (int i = 0; i < ksize; ++i) gk.k[i] = kernel[i];
    const size_t smem = gauss_smem(ksize, cn);
    dim3 grid((w + GT_W - 1) / GT_W, (h + GT_H - 1) / GT_H, 1);
    const int cb = gray_bits == 14 ? 1868 : 3735, cg = gray_bits == 14 ? 9617 : 19235, cr = gray_bits == 14 ? 4899 : 9798;
    static SmemOptIn o3g, o3, o1;
    cudaError_t e = cudaSuccess;
    if (cn == 3 && gray_bits) {
        if (smem > 48 * 1024) e = o3g.ensure((const void*)gauss_kernel<3, true>, smem);
        if (e != cudaSuccess) return e;
        gauss_kernel<3, true><<<grid, 256, smem, s>>>(src, sstep, dst, dstep, w, h, gk, cb, cg, cr, gray_bits);
    } else if (cn == 3) {
        if (smem > 48 * 1024) e = o3.ensure((const void*)gauss_kernel<3, false>, smem);
        if (e != cudaSuccess) return e;
        gauss_kernel<3, false><<<grid, 256, smem, s>>>(src, sstep, dst, dstep, w, h, gk, 0, 0, 0, 15);
    } else if (cn == 1 && !gray_bits) {
        if (smem > 48 * 1024) e = o1.ensure((const void*)gauss_kernel<1, false>, smem);
        if (e != cudaSuccess) return e;
        gauss_kernel<1, false><<<grid, 256, smem, s>>>(src, sstep, dst, dstep, w, h, gk, 0, 0, 0, 15);
    } else {
        return cudaErrorInvalidValue;
    }
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace ofb
