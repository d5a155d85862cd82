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
// Streaming kernel: a CTA owns 128 output columns and marches down a chunk of output rows; a
// thread owns one output column and keeps the five horizontally filtered source rows of the
// current output row in registers (two new source rows per output row).  A source row is read
// once, as one 16-byte load per thread (columns 2x, 2x+1); the three neighbours come from the
// adjacent lanes by shuffle, warp-edge and image-border lanes fetch theirs with scalar loads.
// HBM traffic = the source once + the destination once (+ 3 extra source rows per chunk).
#include "of_device.cuh"
#include "of_kernels.h"

namespace ofb {

namespace {

#ifndef PD_NT_
#define PD_NT_ 128
#endif
constexpr int PD_NT = PD_NT_;      // threads = output columns per CTA
constexpr int PD_AHEAD = 3;     // output rows whose source loads are in flight

__device__ __forceinline__ int r101_any(int i, int n)
{
    if (n == 1) return 0;
    const int p = 2 * (n - 1);
    i %= p;
    if (i < 0) i += p;
    return i >= n ? p - i : i;
}

struct PdCol {
    int x;            // output column
    int c[5];         // reflected source columns 2x-2 .. 2x+2
    bool valid;       // x < ow
    bool pair;        // columns 2x, 2x+1 both exist and can be read as one 16-byte load
    bool fix_l, fix_r;   // left / right neighbours are not available from the adjacent lanes
};

// The thread's source values of one row: columns 2x, 2x+1 for every thread (the only loads of
// interior lanes) plus, for warp-edge and image-border lanes, the neighbours that no adjacent
// lane can supply.  Issued several rows ahead of their use.
struct PdRow { double2 v; double e0, e1, e4; };

__device__ __forceinline__ PdRow pd_load(const double* __restrict__ row, const PdCol& t)
{
    PdRow r;
    r.v = make_double2(0.0, 0.0);
    r.e0 = r.e1 = r.e4 = 0.0;
    if (t.pair) {
        r.v = __ldg(reinterpret_cast<const double2*>(row + t.c[2]));
    } else if (t.valid) {
        r.v.x = __ldg(row + t.c[2]);
        r.v.y = __ldg(row + t.c[3]);
    }
    if (t.fix_l) { r.e0 = __ldg(row + t.c[0]); r.e1 = __ldg(row + t.c[1]); }
    if (t.fix_r) r.e4 = __ldg(row + t.c[4]);
    return r;
}

// Horizontal pass of one source row for this thread's output column.  Warp-collective.
__device__ __forceinline__ double pd_hrow(const PdCol& t, const PdRow& r)
{
    const double s2 = r.v.x, s3 = r.v.y;
    double s0 = __shfl_up_sync(0xffffffffu, s2, 1);
    double s1 = __shfl_up_sync(0xffffffffu, s3, 1);
    double s4 = __shfl_down_sync(0xffffffffu, s2, 1);
    if (t.fix_l) { s0 = r.e0; s1 = r.e1; }
    if (t.fix_r) s4 = r.e4;
    double hv = dmul(s2, 6.0);
    hv = dadd(hv, dmul(dadd(s1, s3), 4.0));
    hv = dadd(hv, s0);
    return dadd(hv, s4);
}

__global__ void __launch_bounds__(PD_NT) pyr_down_kernel(PyrDownArgs a, int rows_per_cta)
{
    const int img = blockIdx.z;
    const double* __restrict__ src = a.src + (size_t)img * a.src_img_stride;
    double* __restrict__ dst = a.dst + (size_t)img * a.dst_img_stride;
    const int w = a.w, h = a.h;
    const int ow = (w + 1) >> 1, oh = (h + 1) >> 1;
    const int lane = threadIdx.x & 31;

    PdCol t;
    t.x = blockIdx.x * PD_NT + threadIdx.x;
    t.valid = t.x < ow;
#pragma unroll
    for (int k = 0; k < 5; ++k) t.c[k] = t.valid ? r101_any(2 * t.x - 2 + k, w) : 0;
    const bool vec = ((reinterpret_cast<uintptr_t>(src) | (a.src_pitch * 8)) & 15) == 0;
    t.pair = t.valid && vec && 2 * t.x + 1 < w;
    // neighbours by shuffle need an adjacent lane that holds the true (unreflected) columns
    t.fix_l = t.valid && (lane == 0 || t.x == 0);
    t.fix_r = t.valid && (lane == 31 || 2 * t.x + 2 >= w || t.x + 1 >= ow);

    const int oy0 = blockIdx.y * rows_per_cta;
    const int oy1 = min(oy0 + rows_per_cta, oh);
    auto row_ptr = [&](int y) { return src + (size_t)r101_any(y, h) * a.src_pitch; };
    double h0, h1, h2, h3, h4;
    {
        const PdRow r2 = pd_load(row_ptr(2 * oy0 - 2), t), r3 = pd_load(row_ptr(2 * oy0 - 1), t);
        const PdRow r4 = pd_load(row_ptr(2 * oy0), t);
        h2 = pd_hrow(t, r2);
        h3 = pd_hrow(t, r3);
        h4 = pd_hrow(t, r4);
    }
    // software pipeline: the loads of PD_AHEAD output rows (2 source rows each) stay in flight
    PdRow q[PD_AHEAD][2];
#pragma unroll
    for (int k = 0; k < PD_AHEAD; ++k) {
        q[k][0] = pd_load(row_ptr(2 * (oy0 + k) + 1), t);
        q[k][1] = pd_load(row_ptr(2 * (oy0 + k) + 2), t);
    }
    for (int oyb = oy0; oyb < oy1; oyb += PD_AHEAD) {
#pragma unroll
        for (int k = 0; k < PD_AHEAD; ++k) {
            const int oy = oyb + k;
            const PdRow v1 = q[k][0], v2 = q[k][1];
            q[k][0] = pd_load(row_ptr(2 * (oy + PD_AHEAD) + 1), t);      // reflected rows stay in range
            q[k][1] = pd_load(row_ptr(2 * (oy + PD_AHEAD) + 2), t);
            h0 = h2; h1 = h3; h2 = h4;
            h3 = pd_hrow(t, v1);
            h4 = pd_hrow(t, v2);
            double o = dmul(h2, 6.0);
            o = dadd(o, dmul(dadd(h1, h3), 4.0));
            o = dadd(o, h0);
            o = dadd(o, h4);
            if (t.valid && oy < oy1) dst[(size_t)oy * a.dst_pitch + t.x] = dmul(o, 1.0 / 256.0);
        }
    }
}

// ------------------------------------------------------------------------------ fast path
// Even width, 16-byte aligned rows, at least 2 output columns.  Same arithmetic and operation
// order as above, different thread mapping: a lane loads one 16-byte chunk (source columns 2k,
// 2k+1) per source row and nothing else; lanes 1..30 of a warp own output columns, lanes 0 and 31
// only supply their chunk to the neighbour (the chunk is loaded again by the adjacent warp, an L1 /
// L2 hit), so a row costs one load instruction per warp, a thread keeps 4 registers per row in
// flight, and 8 CTAs of 128 threads fit an SM with PDF_AHEAD x 2 rows x 16 B in flight per thread.
#ifndef PDF_AHEAD_
#define PDF_AHEAD_ 4
#endif
constexpr int PDF_AHEAD = PDF_AHEAD_;
constexpr int PDF_WCOLS = 30;                       // output columns per warp
constexpr int PDF_COLS = PDF_WCOLS * (PD_NT / 32);  // per CTA

__device__ __forceinline__ double2 pdf_load(const double* __restrict__ row, int k, bool have)
{
    return have ? __ldg(reinterpret_cast<const double2*>(row) + k) : make_double2(0.0, 0.0);
}

__device__ __forceinline__ double pdf_hrow(double2 v, bool first, bool last)
{
    const double s2 = v.x, s3 = v.y;
    double s0 = __shfl_up_sync(0xffffffffu, s2, 1);
    double s1 = __shfl_up_sync(0xffffffffu, s3, 1);
    double s4 = __shfl_down_sync(0xffffffffu, s2, 1);
    if (last) s4 = s2;                              // column w reflects to w-2 = 2x
    if (first) { s0 = s4; s1 = s3; }                // columns -2, -1 reflect to 2, 1
    double hv = dmul(s2, 6.0);
    hv = dadd(hv, dmul(dadd(s1, s3), 4.0));
    hv = dadd(hv, s0);
    return dadd(hv, s4);
}

__global__ void __launch_bounds__(PD_NT, 1024 / PD_NT) pyr_down_fast_kernel(PyrDownArgs a, int rows_per_cta)
{
    const int img = blockIdx.z;
    const double* __restrict__ src = a.src + (size_t)img * a.src_img_stride;
    double* __restrict__ dst = a.dst + (size_t)img * a.dst_img_stride;
    const int h = a.h;
    const int ow = a.w >> 1, oh = (h + 1) >> 1;
    const int lane = threadIdx.x & 31;
    const int x = blockIdx.x * PDF_COLS + (threadIdx.x >> 5) * PDF_WCOLS + lane - 1;   // chunk = output column
    const bool have = x >= 0 && x < ow;
    const bool owner = have && lane >= 1 && lane <= PDF_WCOLS;
    const bool first = x == 0, last = x == ow - 1;

    const int oy0 = blockIdx.y * rows_per_cta;
    const int oy1 = min(oy0 + rows_per_cta, oh);
    const int per = h > 1 ? 2 * (h - 1) : 1;
    auto row_ptr = [&](int y) {                     // BORDER_REFLECT_101, any y
        int i = y % per;
        if (i < 0) i += per;
        if (i >= h) i = per - i;
        return src + (size_t)i * a.src_pitch;
    };
    double h0, h1, h2, h3, h4;
    {
        const double2 r2 = pdf_load(row_ptr(2 * oy0 - 2), x, have), r3 = pdf_load(row_ptr(2 * oy0 - 1), x, have);
        const double2 r4 = pdf_load(row_ptr(2 * oy0), x, have);
        h2 = pdf_hrow(r2, first, last);
        h3 = pdf_hrow(r3, first, last);
        h4 = pdf_hrow(r4, first, last);
    }
    double2 q[PDF_AHEAD][2];
#pragma unroll
    for (int k = 0; k < PDF_AHEAD; ++k) {
        q[k][0] = pdf_load(row_ptr(2 * (oy0 + k) + 1), x, have);
        q[k][1] = pdf_load(row_ptr(2 * (oy0 + k) + 2), x, have);
    }
    double* out = dst + (size_t)oy0 * a.dst_pitch + x;
    for (int oyb = oy0; oyb < oy1; oyb += PDF_AHEAD) {
#pragma unroll
        for (int k = 0; k < PDF_AHEAD; ++k) {
            const int oy = oyb + k;
            const double2 v1 = q[k][0], v2 = q[k][1];
            q[k][0] = pdf_load(row_ptr(2 * (oy + PDF_AHEAD) + 1), x, have);
            q[k][1] = pdf_load(row_ptr(2 * (oy + PDF_AHEAD) + 2), x, have);
            h0 = h2; h1 = h3; h2 = h4;
            h3 = pdf_hrow(v1, first, last);
            h4 = pdf_hrow(v2, first, last);
            double o = dmul(h2, 6.0);
            o = dadd(o, dmul(dadd(h1, h3), 4.0));
            o = dadd(o, h0);
            o = dadd(o, h4);
            if (owner && oy < oy1) *out = dmul(o, 1.0 / 256.0);
            out += a.dst_pitch;
        }
    }
}

}  // namespace

cudaError_t launch_pyr_down(const PyrDownArgs& a, cudaStream_t s, unsigned long long* launches)
{
    const int ow = (a.w + 1) / 2, oh = (a.h + 1) / 2;
#ifndef PD_NO_FAST
    if ((a.w & 1) == 0 && a.w >= 4 && (((uintptr_t)a.src | (a.src_pitch * 8) | (a.src_img_stride * 8)) & 15) == 0) {
        const int strips = (ow + PDF_COLS - 1) / PDF_COLS;
        // Output rows per CTA: the height that minimises (waves of resident CTAs) x (rows + the ~6 row-times a CTA
        // spends on its three warm-up source rows and on filling its load pipeline).  A fixed 64 rows put the 1080p
        // batch at 1224 CTAs against 1184 resident ones -- two waves, the second one nearly empty.
        const long slots = (long)current_sm_count() * (1024 / PD_NT);
        int rows = 64;
        long best = -1;
        for (int r = 8; r <= 128; r += PDF_AHEAD) {
            const long ctas = (long)strips * ((oh + r - 1) / r) * a.n_images;
            const long cost = ((ctas + slots - 1) / slots) * (r + 6);
            if (best < 0 || cost <= best) { best = cost; rows = r; }
        }
        dim3 grid(strips, (oh + rows - 1) / rows, a.n_images);
        pyr_down_fast_kernel<<<grid, PD_NT, 0, s>>>(a, rows);
        if (launches) *launches += 1;
        return cudaGetLastError();
    }
#endif
    const int strips = (ow + PD_NT - 1) / PD_NT;
    // rows per CTA: long enough to amortise the 3 warm-up rows, short enough to fill the GPU
    int rows = 64;
    while (rows > 16 && (long)strips * ((oh + rows - 1) / rows) * a.n_images < (long)current_sm_count() * 12) rows /= 2;
    dim3 grid(strips, (oh + rows - 1) / rows, a.n_images);
    pyr_down_kernel<<<grid, PD_NT, 0, s>>>(a, rows);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace ofb
