// Device-side arithmetic shared by the sm_100a kernels of the pyramidal LK path.
//
// Everything that feeds the 2x2 solve is written with the non-contracting intrinsics
// (__dadd_rn/__dmul_rn/...): the reference materialises every cv::Mat product before it
// subtracts (ParallelVideoPyr.cpp:470-472), so det = rn(rn(Sxx*Syy) - rn(Sxy*Sxy)); an
// FMA-contracted det is non-zero where the reference's is exactly 0 and flips the
// validity mask.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ofb {

__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }

// BORDER_REFLECT_101 for an index known to lie in [-2, n+1] (all this path needs).
__device__ __forceinline__ int r101(int i, int n)
{
    if (n == 1) return 0;
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i < 0 ? 0 : i;          // only reachable for n == 2, i == n+1
}

// Source-row index map of cv::pyrUp's vertical pass: rho(k) = R101(2k, 2H) / 2,
// i.e. rho(-1) = 1 (0 when H == 1), rho(H) = H-1.   (ParallelVideoPyr.cpp:646-647)
__device__ __forceinline__ int pyrup_rho(int k, int hc)
{
    if (k < 0) return hc > 1 ? 1 : 0;
    if (k >= hc) return hc - 1;
    return k;
}

// Horizontal pass of cv::pyrUp for fine column X from one coarse row `s` (n = coarse
// width, s indexed by absolute coarse column minus `c0`).
__device__ __forceinline__ double pyrup_row(const double* __restrict__ s, int c0, int X, int n)
{
    const int j = X >> 1;
    if (n == 1) return dmul(s[0 - c0], 8.0);
    if (X & 1) {
        if (j == n - 1) return dmul(s[n - 1 - c0], 8.0);
        return dmul(dadd(s[j - c0], s[j + 1 - c0]), 4.0);
    }
    if (j == 0) return dadd(dmul(s[0 - c0], 6.0), dmul(s[1 - c0], 2.0));
    if (j == n - 1) return dadd(s[n - 2 - c0], dmul(s[n - 1 - c0], 7.0));
    return dadd(dadd(s[j - 1 - c0], dmul(s[j - c0], 6.0)), s[j + 1 - c0]);
}

// cv::pyrUp value at fine pixel (Y, X) of a coarse plane held in a tile `t` with row
// pitch `tp`, whose element (0,0) is coarse pixel (cy0, cx0).  wc x hc = coarse size.
__device__ __forceinline__ double pyrup_at(const double* __restrict__ t, int tp, int cy0, int cx0,
                                           int Y, int X, int wc, int hc)
{
    const int k = Y >> 1;
    const double* r1 = t + (k - cy0) * tp;
    const double* r2 = t + (pyrup_rho(k + 1, hc) - cy0) * tp;
    const double v1 = pyrup_row(r1, cx0, X, wc);
    const double v2 = pyrup_row(r2, cx0, X, wc);
    if (Y & 1) return dmul(dmul(dadd(v1, v2), 4.0), 1.0 / 64.0);
    const double* r0 = t + (pyrup_rho(k - 1, hc) - cy0) * tp;
    const double v0 = pyrup_row(r0, cx0, X, wc);
    return dmul(dadd(dadd(v0, dmul(v1, 6.0)), v2), 1.0 / 64.0);
}

// OpenCV remap fixed-point coordinate (INTER_BITS = 5): s = rint_half_even(m * 32.f) in
// float, integer part s >> 5 saturated to int16 (remap keeps it in a CV_16SC2 map),
// fraction s & 31.   (cv::remap call sites ParallelVideoPyr.cpp:695,700)
__device__ __forceinline__ void q5(float m, int& ip, int& frac)
{
    float t = __fmul_rn(m, 32.0f);
    if (!(fabsf(t) <= 1073741824.0f)) t = (t < 0.0f) ? -1073741824.0f : 1073741824.0f;
    const int s = __float2int_rn(t);
    int i = s >> 5;
    ip = max(-32768, min(32767, i));
    frac = s & 31;
}

// k / 1024 for an integer 0 <= k <= 1024, exactly, with one FP64 add instead of an
// int->double conversion (conversions issue at 1/4 of the DADD rate on sm_100):
// the double with high word 0x42900000 and low word k is 2^42 + k * 2^-10.
__device__ __forceinline__ double q10_to_double(int k)
{
    return __dsub_rn(__hiloint2double(0x42900000, k), 4398046511104.0);
}

// Bilinear sample of `src` (w x h, row pitch `pitch` doubles) at map coordinate (mx, my):
// float32 weights (exact multiples of 1/1024), double blend left to right,
// BORDER_CONSTANT 0 tested per tap.
__device__ __forceinline__ double remap_tap4(const double* __restrict__ src, size_t pitch,
                                             int w, int h, float mx, float my)
{
    int ix, iy, ax, ay;
    q5(mx, ix, ax);
    q5(my, iy, ay);
    const double w00 = q10_to_double((32 - ay) * (32 - ax));
    const double w01 = q10_to_double((32 - ay) * ax);
    const double w10 = q10_to_double(ay * (32 - ax));
    const double w11 = q10_to_double(ay * ax);
    const bool x0 = (unsigned)ix < (unsigned)w, x1 = (unsigned)(ix + 1) < (unsigned)w;
    const bool y0 = (unsigned)iy < (unsigned)h, y1 = (unsigned)(iy + 1) < (unsigned)h;
    const double* p = src + (ptrdiff_t)iy * (ptrdiff_t)pitch + ix;
    const double s00 = (y0 && x0) ? __ldg(p) : 0.0;
    const double s01 = (y0 && x1) ? __ldg(p + 1) : 0.0;
    const double s10 = (y1 && x0) ? __ldg(p + pitch) : 0.0;
    const double s11 = (y1 && x1) ? __ldg(p + pitch + 1) : 0.0;
    return dadd(dadd(dadd(dmul(s00, w00), dmul(s01, w01)), dmul(s10, w10)), dmul(s11, w11));
}

// Map build (ParallelVideoPyr.cpp:667-684): all float32; sign=+1 frame 1, -1 frame 2.
__device__ __forceinline__ void map_coords(int x, int y, double upu, double upv,
                                           float& m1x, float& m1y, float& m2x, float& m2y)
{
    const float fu = __double2float_rn(upu), fv = __double2float_rn(upv);
    const float hu = __fmul_rn(fu, 0.5f), hv = __fmul_rn(fv, 0.5f);   // == f / 2 exactly
    m1x = __fadd_rn((float)x, hu);
    m1y = __fadd_rn((float)y, hv);
    m2x = __fsub_rn((float)x, hu);
    m2y = __fsub_rn((float)y, hv);
}

}  // namespace ofb
