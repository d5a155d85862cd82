// Phase-synchronous fused pyramid-level kernel (throughput path), window radius R = 1..7.
//
// One launch = one pyramid level of every pair of the batch, everything the reference does
// between two levels (ParallelVideoPyr.cpp:610-703) plus the single-level LK (:327-542):
//   pyrUp(coarse flow) -> float32 maps -> remap of both frames -> fx/fy/ft -> products ->
//   (2R+1)^2 window sums -> guarded 2x2 solve -> += upsampled coarse flow.
//
// One CTA (NT threads, one per SM) owns a strip of 128 product columns (OUT_COLS output
// columns plus the window halo) and marches down a chunk of rows in steps of WIN = 2R+1
// rows.  A step is a sequence of block-wide phases; in every phase all threads do the same
// light per-element job, so latency is hidden by thread count instead of by deep per-warp
// pipelines:
//
//   P0  coarse-flow rows the next step needs: one thread hands the rows to the copy engine
//       (cp.async.bulk.tensor, one box of 66 coarse columns per row and plane, zero-filled
//       outside the plane, completion on an mbarrier) while the block runs P2; afterwards
//       the horizontal pass of cv::pyrUp turns them into fine columns                  -> HR ring
//   P1  one warp per fine row, a lane takes columns lane + 32 i: vertical pass of pyrUp,
//       float32 maps, 8 coalesced gathers per pixel, two blends                       -> AB
//       (the upsampled flow of output pixels is parked in u, v for the epilogue)
//   P2  one warp per product row, a lane takes 4 columns: 2x2 gradient stencils of both
//       frames, five products, horizontal window sums in registers + shuffles        -> PROD
//   P3  per (channel, column pair): vertical window sums, van Herk / Gil-Werman; suffix
//       sums of the previous block in shared memory, this block's rows in registers  -> PROD, VS
//   P4  one warp per output row, a lane takes 4 columns: guarded solve, += parked up-flow,
//       stores, mask, statistics
//
// The step loop at the bottom of the file overlaps the latency-bound P1 of the next step with
// P3 and P4 of the current one (see the schedule comment there).
//
// Window sums never subtract: an all-zero window sums to exactly 0, so det == 0 pixels keep
// the reference's "zero divisor -> 0" behaviour (cv::divide, :514,518) bit for bit.  The
// association order of the (2R+1)^2 terms differs from the oracle's row-major order by design
// (no reference order exists for R > 1); tests bound the effect at 1e-3 px.
// Vertical blocks start at absolute rows = -R (mod WIN) whatever the chunking, so results do
// not depend on the batch size.
//
// HBM traffic per level-l pixel: read A, B (gathered, L2-friendly), read 1/4 coarse u, v,
// write u, v = 36 B algorithmic; the parked up-flow adds a write + read of u, v that stays
// in L2 (a step's rows are re-read within microseconds).
#include "lk_common.cuh"
#include "of_kernels.h"
#include "tma.cuh"
#include "tmem.cuh"

#include <string.h>

namespace ofb {

#ifdef PH_PROF
// Development aid: per-warp cycle totals of the first CTA, by region of the step loop.
__device__ unsigned long long g_phase_prof[32 * 8];
#ifdef PH_PROF1
__device__ unsigned long long g_p1_prof[32 * 8];
#define P1_MARK(slot)                                                            \
    do {                                                                         \
        const long long t_now = clock64();                                       \
        p1_acc[slot] += t_now - p1_t;                                            \
        p1_t = t_now;                                                            \
    } while (0)
#endif
#define PH_MARK(slot)                                                            \
    do {                                                                         \
        const long long t_now = clock64();                                       \
        prof_acc[slot] += t_now - prof_t;                                        \
        prof_t = t_now;                                                          \
    } while (0)
#else
#define PH_MARK(slot) do { } while (0)
#endif
#ifndef P1_MARK
#define P1_MARK(slot) do { } while (0)
#endif

namespace {

// The phases of a step and their rare paths (taps outside the image, pyrUp's border columns) are inlined into the
// step loop.  Measured alternatives: every phase a function of its own 1.22 ms per level-0 launch
// against 0.94; the rare paths out of line (step loop 47 KB -> 33 KB of code, about the size of the instruction
// cache) 0.965 against 0.944 -- code size is not what bounds the loop; for R >= 2 a blend written as three
// interpolations (two exact fractions instead of four weights, ~12 integer instructions per pixel less) 0.909
// against 0.920, but it moves the noise that ill-conditioned coarse pixels hand down outside the validity mask from
// 2e-3 to 1.3e-2 px at 4K and was not taken.
#define PH_COLD __forceinline__
#define PH_PHASE __forceinline__

constexpr int PH_NT = 512;       // threads per CTA (16 warps; a warp = one row of 32 quads)
constexpr int PH_NW = PH_NT / 32;
constexpr int PH_PC = 128;       // product columns per strip = 32 lanes x 4 columns
constexpr int AB_HALF = 72;      // warped-frame rows: lo pairs [0,64), hi pairs [72,136)
constexpr int AB_PITCH = 136;
constexpr int HR_RING = 16;      // upsampled coarse-flow rows kept in shared memory: row k lives in slot (k + 2) & 15,
                                 // element (column X, plane p) at (slot * 128 + X) * 2 + p  (u, v interleaved)
constexpr int CR_W = 68;         // raw coarse columns landed per strip: 64 + one neighbour on either side, starting
                                 // one column further left (the copy engine wants a 16-byte aligned box start)
constexpr int CR_PITCH = 80;     // doubles between the rows of the landing buffer: 640 B = 5 x 128 B (TMA alignment)

constexpr int al16(int d) { return (d + 15) / 16 * 16; }

template <int R>
struct PhaseCfg {
    static constexpr int Q = R / 4;
    static constexpr int HL = Q + 1;                  // halo lanes (4 columns each) per side
    static constexpr int OUT_LANES = 32 - 2 * HL;
    static constexpr int OUT_COLS = 4 * OUT_LANES;
    static constexpr int WIN = 2 * R + 1;             // rows per step = vertical block length
    static constexpr int AB_ROWS = WIN + 1;           // + the row above the step
    static constexpr int HR_ROWS = (WIN + 1) / 2 + 3; // coarse rows a step can touch
    // window-sum rows hold the OUT_COLS output columns, pair-split: lo pairs, then hi pairs at PHALF
    static constexpr int PHALF = (2 * OUT_LANES + 7) / 16 * 16 + 8;
    static constexpr int PPITCH = PHALF + 2 * OUT_LANES;
    static constexpr int PROD_D = WIN * 5 * PPITCH;
    static constexpr int AB_D = AB_ROWS * 2 * AB_PITCH;
    static constexpr int HR_D = HR_RING * 2 * PH_PC;
    static constexpr int CR_D = HR_ROWS * 2 * CR_PITCH;   // landing buffer of the coarse rows (raw)
    // offsets in doubles; the landing buffer must be 128-byte aligned
    static constexpr int O_AB = al16(PROD_D);
    static constexpr int O_HR = O_AB + al16(AB_D);
    static constexpr int O_CR = O_HR + HR_D;
    static constexpr int O_BAR = O_CR + CR_D;              // mbarrier (8 bytes), TMEM base address (4 bytes), P3 mbarrier
    // Tensor memory: a P3 thread keeps the 2R raw rows 1 .. 2R of the previous vertical block of its
    // (channel, column pair) there, 2R x 2 doubles = 8R columns; a column slot of 64 per warp
    static constexpr int TM_SLOT = 64;
    static constexpr int TM_COLS = 256;
    static constexpr size_t SMEM_BYTES = (size_t)(O_BAR + 3) * 8;
    static constexpr int VPAIRS = 2 * OUT_LANES;      // column pairs per channel in P3
    static constexpr int NT = PH_NT;
};

// R = 1 (the reference's own 3x3 window, get_Sum9 :114-160): the nine terms are summed in the
// reference's row-major order, which is not separable, so there are no horizontal / vertical
// passes.  A step is WIN = 15 rows (not the window length); raw products of all 128 strip
// columns are kept in a ring of WIN + 2 rows laid out like the warped-frame rows.
#ifndef PH1_WIN
#define PH1_WIN 15
#define PH1_CTAS 1
#endif
template <>
struct PhaseCfg<1> {
    static constexpr int HL = 1;
    static constexpr int OUT_LANES = 30;
    static constexpr int OUT_COLS = 4 * OUT_LANES;
    static constexpr int WIN = PH1_WIN;               // rows per step
    static constexpr int CTAS = PH1_CTAS;             // CTAs per SM
    static constexpr int NT = 32 * (WIN + 1);         // one warp per row of P1 (+ the row above the first step)
    static constexpr int AB_ROWS = WIN + 1;
    static constexpr int HR_ROWS = (WIN + 1) / 2 + 3;
    static constexpr int PR_ROWS = WIN + 2;           // product-row ring
    static constexpr int PROD_D = PR_ROWS * 5 * 136;
    static constexpr int AB_D = AB_ROWS * 2 * 136;
    static constexpr int HR_D = HR_RING * 2 * PH_PC;
    static constexpr int CR_D = HR_ROWS * 2 * CR_PITCH;
    static constexpr int O_AB = al16(PROD_D);
    static constexpr int O_HR = O_AB + al16(AB_D);
    static constexpr int O_CR = O_HR + HR_D;
    static constexpr int O_BAR = O_CR + CR_D;
    static constexpr size_t SMEM_BYTES = (size_t)(O_BAR + 2) * 8;
};

// Shared-memory rows that are read four columns per lane are "pair-split": columns 4j, 4j+1 of
// lane j sit at lo[2j], columns 4j+2, 4j+3 at hi[2j] = row[half + 2j], so that every 128-bit
// access of a warp is contiguous.  half = 8 (mod 16) doubles makes the column-per-lane and
// pair-per-lane accesses of the other phases conflict-free as well.
__device__ __forceinline__ double2 lds2(const double* p) { return *reinterpret_cast<const double2*>(p); }
__device__ __forceinline__ void sts2(double* p, double x, double y) { *reinterpret_cast<double2*>(p) = make_double2(x, y); }
__device__ __forceinline__ void ld4s(const double* p, int half, double (&v)[4])
{
    const double2 a = lds2(p), b = lds2(p + half);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4s(double* p, int half, const double (&v)[4])
{
    sts2(p, v[0], v[1]);
    sts2(p + half, v[2], v[3]);
}

struct PGeo {
    int W, H;
    int c0;           // image column of strip column 0 (multiple of 4, may be negative)
    int xo0;          // first output column
    int y0, y1;       // output rows [y0, y1)
    bool has_coarse;
    bool edge;        // strip reaches outside the image on the left or right
};

// Position of strip column X inside a warped-frame row.
__device__ __forceinline__ int ppos(int X) { return ((X & 2) ? AB_HALF : 0) + 2 * (X >> 2) + (X & 1); }

// ------------------------------------------------------------------------------ P0
// Coarse-flow rows rho(k_from) .. rho(k_from + n_new - 1) (cv::pyrUp's row map, :646-647) of both
// planes -> landing buffer, by the copy engine: one box of CR_W columns starting two coarse columns
// left of the strip (an even column: FP64 boxes must start 16-byte aligned) per row and plane; columns
// outside the plane arrive as 0.  Called by one thread.
__device__ __forceinline__ void up_rows_issue(const CUtensorMap* tmu, const CUtensorMap* tmv, int j0, int hc, int pair,
                                              int k_from, int n_new, double* s_cr, uint64_t* bar)
{
    mbar_arrive_expect_tx(bar, (uint32_t)n_new * (uint32_t)(2 * CR_W * sizeof(double)));
#pragma unroll 1
    for (int r = 0; r < n_new; ++r) {
        const int row = pyrup_rho(k_from + r, hc);
        tma_load_3d(s_cr + (2 * r) * CR_PITCH, tmu, j0, row, pair, bar);
        tma_load_3d(s_cr + (2 * r + 1) * CR_PITCH, tmv, j0, row, pair, bar);
    }
}

// cv::pyrUp's edge formulas for the first / last coarse column (and 0 outside the image), out of line.
static __device__ PH_COLD void up_border(const double* cu, const double* cv, int j0, int j, int wc, double* d)
{
    double eu = 0.0, ou = 0.0, ev = 0.0, ov = 0.0;
    if ((unsigned)j < (unsigned)wc) {
        eu = pyrup_row(cu, j0, 2 * j, wc);
        ou = pyrup_row(cu, j0, 2 * j + 1, wc);
        ev = pyrup_row(cv, j0, 2 * j, wc);
        ov = pyrup_row(cv, j0, 2 * j + 1, wc);
    }
    *reinterpret_cast<double2*>(d) = make_double2(eu, ev);
    *reinterpret_cast<double2*>(d + 2) = make_double2(ou, ov);
}

// Horizontal pass of cv::pyrUp (:646-647) over the landed rows -> HR ring; task = (row, coarse
// column j) -> fine columns 2j, 2j+1 of both planes.  Fine columns outside the image hold 0.
template <int NT>
__device__ PH_PHASE void up_rows_hpass(const PGeo& g, int wc, int k_from, int n_new, const double* s_cr,
                                              double* s_hr)
{
    const int jl = threadIdx.x & 63;
    const int j0 = (g.c0 >> 1) - 2;                     // coarse column of landing column 0
    const int j = j0 + 2 + jl;
    const bool border = (unsigned)(j - 1) >= (unsigned)(wc - 2);      // j <= 0 or j >= wc - 1
    for (int r = threadIdx.x >> 6; r < n_new; r += NT / 64) {
        const double* cu = s_cr + (2 * r) * CR_PITCH + jl + 1;       // cu[0] = column j - 1
        const double* cv = cu + CR_PITCH;
        double* d = s_hr + ((((k_from + r + 2) & (HR_RING - 1)) * PH_PC) + 2 * jl) * 2;
        if (!border) {
            const double um = cu[0], u0 = cu[1], u1 = cu[2];
            const double vm = cv[0], v0 = cv[1], v1 = cv[2];
            sts2(d, dadd(dadd(um, dmul(u0, 6.0)), u1), dadd(dadd(vm, dmul(v0, 6.0)), v1));
            sts2(d + 2, dmul(dadd(u0, u1), 4.0), dmul(dadd(v0, v1), 4.0));
        } else {
            up_border(cu - jl - 1, cv - jl - 1, j0, j, wc, d);
        }
    }
}

// Rare per-tap path of the remap (a tap outside the image, a non-finite flow).
static __device__ PH_COLD double remap_tap4_cold(const double* __restrict__ F, size_t pitch, int W, int H, float mx,
                                                      float my)
{
    return remap_tap4(F, pitch, W, H, mx, my);
}

// ------------------------------------------------------------------------------ P1
// Warped frames A', B' (:667-700) of fine rows vy0 + r - 1, r = r_lo .. WIN, into the AB block
// (slot of r = (sb + r) mod AB_ROWS); the calling warp takes the rows r_first, r_first + r_step,
// ...; one warp per row, a lane takes the columns lane + 32 i so
// that every gather instruction of the warp reads 32 neighbouring pixels.  Row -1 is the
// reflect-101 row above row 0 (image row 1); other rows outside the image are never read.
//
// Map arithmetic (:667-684 and cv::remap's fixed point): the reference forms m = x + f / 2 in float32
// and remap rounds m * 32 to an integer.  Scaling by 32 commutes with the float32 rounding of the
// sum, so s = rint(32 x + 16 f) is the same integer with two instructions less per coordinate
// (32 x is exact; the only inexact cases are |m| below 2^-120, where both give 0).  A non-finite
// or huge flow is clamped to 3e38: the tap then lies outside the image like the reference's.
template <int R>
__device__ PH_PHASE void phase_remap(const LevelArgs& a, const PGeo& g, const double* __restrict__ A,
                                            const double* __restrict__ B, double* __restrict__ U,
                                            double* __restrict__ V, int vy0, int r_first, int r_step, int sb,
                                            const double* s_hr, double* s_ab)
{
    using C = PhaseCfg<R>;
    const int W = g.W, H = g.H;
    const int pitch = (int)a.img_pitch;
    const int lane = threadIdx.x & 31;
    const int x0 = g.c0 + lane;
#ifdef PH_PROF1
    long long p1_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long p1_t = clock64();
#endif
    for (int r = r_first; r <= C::WIN; r += r_step) {
        const int vy = vy0 + r - 1;
        int yy = vy;
        if (vy == -1) yy = H > 1 ? 1 : 0;
        else if ((unsigned)vy >= (unsigned)H) continue;
        int slot = sb + r;
        if (slot >= C::AB_ROWS) slot -= C::AB_ROWS;
        double* dst = s_ab + slot * (2 * AB_PITCH) + ppos(lane);       // column lane + 32 i at dst[16 i]
        P1_MARK(0);
        double va[4] = {0.0, 0.0, 0.0, 0.0}, vb[4] = {0.0, 0.0, 0.0, 0.0};
#ifdef PH_KNOCK_REMAP
        if (false) {
#else
        if (g.has_coarse) {
#endif
            // vertical pass of cv::pyrUp from the rows of P0 (u, v interleaved: one 16-byte load per row)
            const int k = yy >> 1;
            const double* h0 = s_hr + ((((k + 2) & (HR_RING - 1)) * PH_PC) + lane) * 2;   // row k
            const double* hp = s_hr + ((((k + 3) & (HR_RING - 1)) * PH_PC) + lane) * 2;   // row k+1
            double upu[4], upv[4];
            if (yy & 1) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const double2 c = lds2(h0 + 64 * i), p = lds2(hp + 64 * i);
                    // (a * 4) / 64 == a / 16: both scalings are exact
                    upu[i] = dmul(dadd(c.x, p.x), 1.0 / 16.0);
                    upv[i] = dmul(dadd(c.y, p.y), 1.0 / 16.0);
                }
            } else {
                const double* hm = s_hr + ((((k + 1) & (HR_RING - 1)) * PH_PC) + lane) * 2;   // row k-1
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const double2 m = lds2(hm + 64 * i), c = lds2(h0 + 64 * i), p = lds2(hp + 64 * i);
                    upu[i] = dmul(dadd(dadd(m.x, dmul(c.x, 6.0)), p.x), 1.0 / 64.0);
                    upv[i] = dmul(dadd(dadd(m.y, dmul(c.y, 6.0)), p.y), 1.0 / 64.0);
                }
            }
            P1_MARK(1);
            // park the upsampled flow of output pixels for the epilogue (tmpu += upu, :634-637)
            if (vy == yy && vy >= g.y0 && vy < g.y1) {
                const int off = vy * (int)a.uv_pitch + x0;
                double* pu = U + off;
                double* pv = V + off;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    bool ok = true;
                    if (i == 0) ok = lane >= 4 * C::HL;
                    if (i == 3) ok = lane < 4 * C::HL + C::OUT_COLS - 96;
                    if (g.edge) ok = ok && (unsigned)(x0 + 32 * i) < (unsigned)W;
                    if (ok) {
                        pu[32 * i] = upu[i];
                        pv[32 * i] = upv[i];
                    }
                }
            }
            P1_MARK(2);
            float hu[4], hv[4];                              // 32 * (f / 2), float32
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                hu[i] = fminf(__fmul_rn(__double2float_rn(upu[i]), 16.0f), 3.0e38f);
                hv[i] = fminf(__fmul_rn(__double2float_rn(upv[i]), 16.0f), 3.0e38f);
            }
            const float xf = (float)(32 * x0), yf = (float)(32 * yy);
            // frame 1 is sampled at (x + f.x / 2, y + f.y / 2), frame 2 at (x - f.x / 2, y - f.y / 2)
            // both frames' 32 gathers in flight before the first blend (measured: level 0 0.976 -> 0.910 ms)
            int off[2][4], fr[2][4];
            unsigned out = 0;
#pragma unroll
            for (int f = 0; f < 2; ++f)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float xi = __fadd_rn(xf, 1024.0f * (float)i);
                    const int sx = __float2int_rn(f ? __fsub_rn(xi, hu[i]) : __fadd_rn(xi, hu[i]));
                    const int sy = __float2int_rn(f ? __fsub_rn(yf, hv[i]) : __fadd_rn(yf, hv[i]));
                    const int ix = sx >> 5, iy = sy >> 5;
                    const bool interior = (unsigned)ix < (unsigned)(W - 1) && (unsigned)iy < (unsigned)(H - 1);
                    off[f][i] = interior ? iy * pitch + ix : 0;
                    if (!interior) out |= 1u << (4 * f + i);
                    fr[f][i] = (sx & 31) | ((sy & 31) << 8);
                }
            double sv[2][4][4];
#pragma unroll
            for (int f = 0; f < 2; ++f)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const double* pa = (f ? B : A) + (unsigned)off[f][i];
                    sv[f][i][0] = __ldg(pa); sv[f][i][1] = __ldg(pa + 1);
                    sv[f][i][2] = __ldg(pa + pitch); sv[f][i][3] = __ldg(pa + pitch + 1);
                }
#pragma unroll
            for (int f = 0; f < 2; ++f) {
                double* v = f ? vb : va;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int ax = fr[f][i] & 31, ay = fr[f][i] >> 8;
                    v[i] = R == 1 ? blend4_exact(sv[f][i][0], sv[f][i][1], sv[f][i][2], sv[f][i][3], ax, ay)
                                  : blend4(sv[f][i][0], sv[f][i][1], sv[f][i][2], sv[f][i][3], ax, ay);
                }
            }
            if (out) {                                       // rare: a tap outside the image, non-finite flow
#pragma unroll
                for (int f = 0; f < 2; ++f) {
                    const double* __restrict__ F = f ? B : A;
                    double* v = f ? vb : va;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (((out >> (4 * f + i)) & 1u) && (unsigned)(x0 + 32 * i) < (unsigned)W) {
                            const float xr = (float)(x0 + 32 * i), yr = (float)yy;
                            const float hx = __fmul_rn(hu[i], 1.0f / 32.0f), hy = __fmul_rn(hv[i], 1.0f / 32.0f);
                            const float mx = f ? __fsub_rn(xr, hx) : __fadd_rn(xr, hx);
                            const float my = f ? __fsub_rn(yr, hy) : __fadd_rn(yr, hy);
                            v[i] = remap_tap4_cold(F, a.img_pitch, W, H, mx, my);
                        }
                    }
                }
            }
        } else {
            const double* pa = A + (size_t)yy * a.img_pitch + x0;
            const double* pb = B + (size_t)yy * a.img_pitch + x0;
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if ((unsigned)(x0 + 32 * i) < (unsigned)W) { va[i] = __ldg(pa + 32 * i); vb[i] = __ldg(pb + 32 * i); }
        }
        if (g.edge) {                                        // columns outside the image read as 0
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const bool in = (unsigned)(x0 + 32 * i) < (unsigned)W;
                va[i] = in ? va[i] : 0.0;
                vb[i] = in ? vb[i] : 0.0;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            dst[16 * i] = va[i];
            dst[AB_PITCH + 16 * i] = vb[i];
        }
        P1_MARK(7);
    }
#ifdef PH_PROF1
    if (blockIdx.x == 1 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && W >= 1920)
        for (int i = 0; i < 8; ++i) atomicAdd(&g_p1_prof[(threadIdx.x >> 5) * 8 + i], (unsigned long long)p1_acc[i]);
#endif
}

// ------------------------------------------------------------------------------ P2
// One warp per product row vy0 + i (lane = 4 consecutive columns): gradients (2x2 filter2D
// stencils over both frames, accumulation order ((k00 p00 + k01 p01) + k10 p10) + k11 p11,
// :44-84), the five products (:383-405) and their horizontal (2R+1)-column window sums in
// registers + shuffles (additions only).  Rows and columns outside the image contribute 0.
template <int R>
__device__ PH_PHASE void phase_products(const PGeo& g, int vy0, int sb, const double* s_ab, double* s_prod)
{
    using C = PhaseCfg<R>;
    const int W = g.W, H = g.H;
    const int lane = threadIdx.x & 31;
    const bool active = lane >= C::HL && lane < 32 - C::HL;
    const int xb = g.c0 + 4 * lane;
    for (int i = threadIdx.x >> 5; i < C::WIN; i += PH_NW) {
        const int vy = vy0 + i;
        double* d = s_prod + i * (5 * C::PPITCH) + 2 * (lane - C::HL);
        if ((unsigned)vy >= (unsigned)H) {                   // warp-uniform
            if (active) {
                const double z[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
                for (int c = 0; c < 5; ++c) st4s(d + c * C::PPITCH, C::PHALF, z);
            }
            continue;
        }
        int s1 = sb + i + 1, s0 = sb + i;
        if (s1 >= C::AB_ROWS) s1 -= C::AB_ROWS;
        if (s0 >= C::AB_ROWS) s0 -= C::AB_ROWS;
        const double* up = s_ab + s0 * (2 * AB_PITCH) + 2 * lane;
        const double* cr = s_ab + s1 * (2 * AB_PITCH) + 2 * lane;
        double pA[4], pB[4], cA[4], cB[4];
        ld4s(up, AB_HALF, pA);
        ld4s(up + AB_PITCH, AB_HALF, pB);
        ld4s(cr, AB_HALF, cA);
        ld4s(cr + AB_PITCH, AB_HALF, cB);
        // left neighbours (column xb - 1 = hi pair of lane - 1); lane 0 has none: its column 0
        // only exists as the neighbour of column 1 and its products are never used
        double pLA = up[AB_HALF - 1], pLB = up[AB_PITCH + AB_HALF - 1];
        double cLA = cr[AB_HALF - 1], cLB = cr[AB_PITCH + AB_HALF - 1];
        if (xb == 0) {                                       // reflect-101 at the left image border
            pLA = pA[1]; pLB = pB[1]; cLA = cA[1]; cLB = cB[1];
        }
        double fx[4], fy[4], ft[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double a00 = q ? pA[q - 1] : pLA, a01 = pA[q];
            const double a10 = q ? cA[q - 1] : cLA, a11 = cA[q];
            const double b00 = q ? pB[q - 1] : pLB, b01 = pB[q];
            const double b10 = q ? cB[q - 1] : cLB, b11 = cB[q];
            const double fxa = dadd(dsub(dsub(a01, a00), a10), a11);
            const double fxb = dadd(dsub(dsub(b01, b00), b10), b11);
            const double sa = dadd(a00, a01), sbb = dadd(b00, b01);
            const double fya = dadd(dsub(a10, sa), a11);
            const double fyb = dadd(dsub(b10, sbb), b11);
            const double fta = -dadd(dadd(sa, a10), a11);
            const double ftb = dadd(dadd(sbb, b10), b11);
            fx[q] = dadd(fxa, fxb);
            fy[q] = dadd(fya, fyb);
            ft[q] = dadd(fta, ftb);
        }
        if (g.edge) {                                        // columns outside the image contribute nothing
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool in = (unsigned)(xb + q) < (unsigned)W;
                fx[q] = in ? fx[q] : 0.0;
                fy[q] = in ? fy[q] : 0.0;
                ft[q] = in ? ft[q] : 0.0;
            }
        }
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            double x[4], h[4];
#pragma unroll
            for (int q = 0; q < 4; ++q)
                x[q] = dmul(c == 1 || c == 4 ? fy[q] : fx[q], c == 0 ? fx[q] : (c == 1 || c == 2 ? fy[q] : ft[q]));
            hsum4<R>(x, h);
            if (active) st4s(d + c * C::PPITCH, C::PHALF, h);
        }
    }
}

// ------------------------------------------------------------------------------ P3
// Vertical (2R+1)-row window sums of one (channel, column pair), in place.  van Herk /
// Gil-Werman: the window of block row i is the suffix of the previous block from row i+1 plus
// the prefix of this block up to row i.  The previous block's raw rows 1 .. 2R are the thread's
// own state from the last step and live in tensor memory (taddr), not in shared memory: they are
// loaded, turned into the suffix sums in registers, and this block's raw rows are stored back as
// they are read.  Additions only, in the order  row(t+1) + (row(t+2) + ... ).
// Warp-collective (tcgen05.ld / st): threads without a task run it with active = false.
template <int R>
__device__ PH_PHASE void phase_vertical(double* col, uint32_t taddr, bool active)
{
    constexpr int WIN = 2 * R + 1;
    constexpr int ROW = 5 * PhaseCfg<R>::PPITCH;
    double2 sf[2 * R];                                   // sf[t] = rows t+1 .. 2R of the previous block
#pragma unroll
    for (int c = 0; c < R; ++c) {
        double t[4];
        tmem_ld4(taddr + 8 * c, t);
        sf[2 * c] = make_double2(t[0], t[1]);
        sf[2 * c + 1] = make_double2(t[2], t[3]);
    }
    tmem_wait_ld();
#pragma unroll
    for (int t = 2 * R - 2; t >= 0; --t) {
        sf[t].x = dadd(sf[t].x, sf[t + 1].x);
        sf[t].y = dadd(sf[t].y, sf[t + 1].y);
    }
    double2 P = make_double2(0.0, 0.0), keep = P;
#pragma unroll
    for (int i = 0; i < WIN; ++i) {
        const double2 p = lds2(col + i * ROW);            // threads without a task read column pair 0 and drop it
        if (i == 0) P = p;
        else { P.x = dadd(P.x, p.x); P.y = dadd(P.y, p.y); }
        double2 S = P;
        if (i < WIN - 1) {
            S.x = dadd(sf[i].x, P.x);
            S.y = dadd(sf[i].y, P.y);
        }
        if (active) sts2(col + i * ROW, S.x, S.y);
        if (i >= 1) {                                    // raw rows 1 .. 2R -> tensor memory, two rows per store
            if (i & 1) keep = p;
            else {
                const double t[4] = {keep.x, keep.y, p.x, p.y};
                tmem_st4(taddr + 8 * (i / 2 - 1), t);
            }
        }
    }
    tmem_wait_st();
}

// ------------------------------------------------------------------------------ P4
// Guarded 2x2 solve (:407-542), += upsampled coarse flow (:634-637), stores, mask, gesture
// statistics (:845-873).
struct StatAcc { unsigned long long x, y, c; };

// Solve + epilogue of output pixels (y, xb), (y, xb + 1) from their five window sums; the statistics
// of the pair are added to the caller's per-row counters (sx: sum of the columns, sc: count).
template <int R, bool VEC>
__device__ __forceinline__ void solve_pair(const LevelArgs& a, const PGeo& g, double* __restrict__ pu,
                                           double* __restrict__ pv, unsigned char* __restrict__ pm, bool yin, int xb,
                                           double2 sxx, double2 syy, double2 sxy, double2 sxt, double2 syt,
                                           double det_eps, double mag2, unsigned& sx, unsigned& sc)
{
    const int W = g.W;
    double upu[2] = {0.0, 0.0}, upv[2] = {0.0, 0.0};
    if (g.has_coarse) {                                  // parked by P1 (same CTA, barriers between)
        if constexpr (VEC) {
            const double2 u0 = *reinterpret_cast<const double2*>(pu), v0 = *reinterpret_cast<const double2*>(pv);
            upu[0] = u0.x; upu[1] = u0.y; upv[0] = v0.x; upv[1] = v0.y;
        } else {
            upu[0] = pu[0]; upu[1] = pu[1]; upv[0] = pv[0]; upv[1] = pv[1];
        }
    }
    double lu[2], lv[2], nus[2], nvs[2], dets[2];
    unsigned okbits = 0;
    bool fast = true;
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const int x = xb + q;
        const bool interior = yin && x >= R && x <= W - 1 - R;
        const double xx = q ? sxx.y : sxx.x, yy = q ? syy.y : syy.x, xy = q ? sxy.y : sxy.x;
        const double xt = q ? sxt.y : sxt.x, yt = q ? syt.y : syt.x;
        dets[q] = dsub(dmul(xx, yy), dmul(xy, xy));
        nus[q] = dsub(dmul(xy, yt), dmul(yy, xt));
        nvs[q] = dsub(dmul(xt, xy), dmul(xx, yt));
        const bool ok = interior && fabs(dets[q]) > det_eps;
        okbits |= ok ? (1u << q) : 0u;
        fast = (div2_fast(nus[q], nvs[q], dets[q], lu[q], lv[q]) || !ok) && fast;
    }
    if (!fast) {                                         // rare: exponent range outside the fast path
        const Quot4 q4 = div4_slow(nus[0], nvs[0], dets[0], nus[1], nvs[1], dets[1]);
        lu[0] = q4.u0; lv[0] = q4.v0; lu[1] = q4.u1; lv[1] = q4.v1;
    }
    double fu[2], fv[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const bool ok = (okbits >> q) & 1u;
        fu[q] = ok ? lu[q] : 0.0;
        fv[q] = ok ? lv[q] : 0.0;
        if (g.has_coarse) { fu[q] = dadd(fu[q], upu[q]); fv[q] = dadd(fv[q], upv[q]); }
    }
    if constexpr (VEC) {
        *reinterpret_cast<double2*>(pu) = make_double2(fu[0], fu[1]);
        *reinterpret_cast<double2*>(pv) = make_double2(fv[0], fv[1]);
    } else {
        pu[0] = fu[0]; pu[1] = fu[1]; pv[0] = fv[0]; pv[1] = fv[1];
    }
    if (pm) {
        pm[0] = (unsigned char)(okbits & 1u);
        pm[1] = (unsigned char)((okbits >> 1) & 1u);
    }
    if (a.stats) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            // val = sqrt(u^2 + v^2); !(val < mag_thresh)  <=>  !(u^2 + v^2 < mag2)
            const double s2 = dadd(dmul(fu[q], fu[q]), dmul(fv[q], fv[q]));
            if (!(s2 < mag2)) { sx += (unsigned)(xb + q); sc += 1u; }
        }
    }
}

// P4 of one output row y (block row i of the window sums): lane L < OUT_LANES owns the column pairs L
// and L + OUT_LANES of the strip's 2 OUT_LANES pairs, so that the lanes' 16-byte accesses to u, v are
// contiguous; pair p = (quad p / 2, half p % 2) of the pair-split sum rows.
template <int R, bool VEC>
__device__ PH_PHASE void phase_solve_row(const LevelArgs& a, const PGeo& g, double* __restrict__ U,
                                                double* __restrict__ V, unsigned char* __restrict__ M, int y,
                                                const double* s_row, StatAcc& st)
{
    using C = PhaseCfg<R>;
    const int L = threadIdx.x & 31;
    const int xb = g.xo0 + 2 * L;
    if (L >= C::OUT_LANES || xb >= g.W) return;          // W even: a pair exists or it does not
    const double det_eps = a.det_eps, mag2 = a.mag2_thresh;
    const bool yin = y >= R && y <= g.H - 1 - R;
    const int off = y * (int)a.uv_pitch + xb;            // < 2^31 (checked by the launcher)
    unsigned char* pm = M ? M + (size_t)y * a.mask_pitch + xb : nullptr;
    const double* sp = s_row + (L & 1) * C::PHALF + 2 * (L >> 1);
    unsigned sx = 0, sc = 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        if (h == 1 && xb + 2 * C::OUT_LANES >= g.W) break;
        const double* q = sp + h * C::OUT_LANES;
        const int dx = 2 * C::OUT_LANES * h;
        solve_pair<R, VEC>(a, g, U + off + dx, V + off + dx, pm ? pm + dx : nullptr, yin, xb + dx, lds2(q),
                           lds2(q + C::PPITCH), lds2(q + 2 * C::PPITCH), lds2(q + 3 * C::PPITCH), lds2(q + 4 * C::PPITCH),
                           det_eps, mag2, sx, sc);
    }
    st.x += sx;
    st.c += sc;
    st.y += (unsigned long long)(sc * (unsigned)y);      // < 4 * 32767
}

// ------------------------------------------------------------------------------ R = 1: P2, P4
// P2 for R = 1: one warp per product row (lane = 4 columns): gradients and the five raw
// products of all 128 strip columns into the product ring (slot n0 + i).
__device__ PH_PHASE void phase_products1(const PGeo& g, int vy0, int sb, int n0, const double* s_ab,
                                                double* s_prod)
{
    using C = PhaseCfg<1>;
    const int W = g.W, H = g.H;
    const int lane = threadIdx.x & 31;
    const int xb = g.c0 + 4 * lane;
    for (int i = threadIdx.x >> 5; i < C::WIN; i += C::NT / 32) {
        const int vy = vy0 + i;
        int slot = n0 + i;
        if (slot >= C::PR_ROWS) slot -= C::PR_ROWS;
        double* d = s_prod + slot * (5 * AB_PITCH) + 2 * lane;
        if ((unsigned)vy >= (unsigned)H) {                   // warp-uniform
            const double z[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
            for (int c = 0; c < 5; ++c) st4s(d + c * AB_PITCH, AB_HALF, z);
            continue;
        }
        int s1 = sb + i + 1, s0 = sb + i;
        if (s1 >= C::AB_ROWS) s1 -= C::AB_ROWS;
        if (s0 >= C::AB_ROWS) s0 -= C::AB_ROWS;
        const double* up = s_ab + s0 * (2 * AB_PITCH) + 2 * lane;
        const double* cr = s_ab + s1 * (2 * AB_PITCH) + 2 * lane;
        double pA[4], pB[4], cA[4], cB[4];
        ld4s(up, AB_HALF, pA);
        ld4s(up + AB_PITCH, AB_HALF, pB);
        ld4s(cr, AB_HALF, cA);
        ld4s(cr + AB_PITCH, AB_HALF, cB);
        double pLA = up[AB_HALF - 1], pLB = up[AB_PITCH + AB_HALF - 1];
        double cLA = cr[AB_HALF - 1], cLB = cr[AB_PITCH + AB_HALF - 1];
        if (xb == 0) {                                       // reflect-101 at the left image border
            pLA = pA[1]; pLB = pB[1]; cLA = cA[1]; cLB = cB[1];
        }
        double fx[4], fy[4], ft[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double a00 = q ? pA[q - 1] : pLA, a01 = pA[q];
            const double a10 = q ? cA[q - 1] : cLA, a11 = cA[q];
            const double b00 = q ? pB[q - 1] : pLB, b01 = pB[q];
            const double b10 = q ? cB[q - 1] : cLB, b11 = cB[q];
            // filter2D accumulation order ((k00 p00 + k01 p01) + k10 p10) + k11 p11 (:44-84)
            const double fxa = dadd(dsub(dsub(a01, a00), a10), a11);
            const double fxb = dadd(dsub(dsub(b01, b00), b10), b11);
            const double sa = dadd(a00, a01), sbb = dadd(b00, b01);
            const double fya = dadd(dsub(a10, sa), a11);
            const double fyb = dadd(dsub(b10, sbb), b11);
            const double fta = -dadd(dadd(sa, a10), a11);
            const double ftb = dadd(dadd(sbb, b10), b11);
            fx[q] = dadd(fxa, fxb);
            fy[q] = dadd(fya, fyb);
            ft[q] = dadd(fta, ftb);
        }
        if (g.edge) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool in = (unsigned)(xb + q) < (unsigned)W;
                fx[q] = in ? fx[q] : 0.0;
                fy[q] = in ? fy[q] : 0.0;
                ft[q] = in ? ft[q] : 0.0;
            }
        }
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            double x[4];
#pragma unroll
            for (int q = 0; q < 4; ++q)
                x[q] = dmul(c == 1 || c == 4 ? fy[q] : fx[q], c == 0 ? fx[q] : (c == 1 || c == 2 ? fy[q] : ft[q]));
            st4s(d + c * AB_PITCH, AB_HALF, x);
        }
    }
}

// P4 for R = 1: the calling warp owns output row y (window rows in ring slots r0, r1, r2), lane
// L (1 .. 30) the four output columns c0 + 4 L ..; 3x3 sums in get_Sum9's order (:122-157:
// row-major, starting from 0.0), then the solve of both column pairs.
template <bool VEC>
__device__ PH_PHASE void phase_solve1(const LevelArgs& a, const PGeo& g, double* __restrict__ U,
                                             double* __restrict__ V, unsigned char* __restrict__ M, int y,
                                             int r0, int r1, int r2, const double* s_prod, StatAcc& st)
{
    using C = PhaseCfg<1>;
    const int L = threadIdx.x & 31;                          // lane L holds strip columns 4L .. 4L+3
    const int xb = g.c0 + 4 * L;
    const bool active = L >= C::HL && L < 32 - C::HL && xb < g.W;
    double S[5][4];
#pragma unroll
    for (int c = 0; c < 5; ++c) {
        double m[3][6];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double* row = s_prod + (k == 0 ? r0 : (k == 1 ? r1 : r2)) * (5 * AB_PITCH) + c * AB_PITCH + 2 * L;
            const double2 lo = lds2(row), hi = lds2(row + AB_HALF);
            // columns 4L-1 and 4L+4 come from the neighbouring lanes (the halo lanes 0 and 31 only provide)
            m[k][0] = __shfl_up_sync(0xffffffffu, hi.y, 1);
            m[k][1] = lo.x; m[k][2] = lo.y; m[k][3] = hi.x; m[k][4] = hi.y;
            m[k][5] = __shfl_down_sync(0xffffffffu, lo.x, 1);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < 3; ++k)
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    // 0.0 + x == x unless x is -0.0, which a square never is
                    if (k == 0 && j == 0 && c < 2) s = m[0][q];
                    else s = dadd(s, m[k][q + j]);
                }
            S[c][q] = s;
        }
    }
    const double det_eps = a.det_eps, mag2 = a.mag2_thresh;
    if (!active) return;
    const bool yin = y >= 1 && y <= g.H - 2;
    const int off = y * (int)a.uv_pitch + xb;            // < 2^31 (checked by the launcher)
    unsigned char* pm = M ? M + (size_t)y * a.mask_pitch + xb : nullptr;
    unsigned sx = 0, sc = 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        if (xb + 2 * h >= g.W) break;
        solve_pair<1, VEC>(a, g, U + off + 2 * h, V + off + 2 * h, pm ? pm + 2 * h : nullptr, yin, xb + 2 * h,
                           make_double2(S[0][2 * h], S[0][2 * h + 1]), make_double2(S[1][2 * h], S[1][2 * h + 1]),
                           make_double2(S[2][2 * h], S[2][2 * h + 1]), make_double2(S[3][2 * h], S[3][2 * h + 1]),
                           make_double2(S[4][2 * h], S[4][2 * h + 1]), det_eps, mag2, sx, sc);
    }
    st.x += sx;
    st.c += sc;
    st.y += (unsigned long long)(sc * (unsigned)y);
}

// ------------------------------------------------------------------------------ kernel
// Step geometry: step s warps fine rows vy0 + r - 1, r = r_lo .. WIN, and needs the coarse rows
// ka .. ka + nk - 1.
struct Step { int vy0, r_lo, sb, ka, nk; bool any; };

template <int R>
__device__ __forceinline__ Step step_geo(const PGeo& g, int n_steps, int s)
{
    using C = PhaseCfg<R>;
    const int H = g.H;
    Step q;
    q.vy0 = g.y0 - R + s * C::WIN;                 // product row of block row 0
    q.r_lo = s == 0 ? 0 : 1;
    q.sb = (s * C::WIN) % C::AB_ROWS;
    // rows [vlo, vhi]; source rows after border handling: [lo, hi]
    const int vlo = q.vy0 + q.r_lo - 1, vhi = q.vy0 + C::WIN - 1;
    int lo = max(vlo, 0), hi = min(vhi, H - 1);
    if (vlo <= -1 && vhi >= -1) { lo = 0; hi = max(hi, min(1, H - 1)); }
    q.ka = (lo >> 1) - 1;
    q.nk = (hi >> 1) + 2 - q.ka;
    q.any = lo <= hi && s < n_steps;
    return q;
}

__device__ __forceinline__ void stats_flush(const LevelArgs& a, int pair, StatAcc st)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        st.x += __shfl_xor_sync(0xffffffffu, st.x, o);
        st.y += __shfl_xor_sync(0xffffffffu, st.y, o);
        st.c += __shfl_xor_sync(0xffffffffu, st.c, o);
    }
    if ((threadIdx.x & 31) == 0 && st.c != 0) {   // integer-valued sums: exact in any order
        atomicAdd(&a.stats[pair].sum_x, (double)st.x);
        atomicAdd(&a.stats[pair].sum_y, (double)st.y);
        atomicAdd(&a.stats[pair].count, (double)st.c);
    }
}

// HC: the level has a coarser level below it (everything but the coarsest level of a pyramid); a template
// parameter so that each kernel carries one version of the remap and of the epilogue only (code size).
template <int R, bool VEC, bool HC>
__global__ void __launch_bounds__(PH_NT, 1) lk_level_phase_kernel(const __grid_constant__ CUtensorMap tmu,
                                                                  const __grid_constant__ CUtensorMap tmv, LevelArgs a,
                                                                  int th)
{
    using C = PhaseCfg<R>;
    extern __shared__ __align__(128) double sm[];
    double* s_prod = sm;
    double* s_ab = sm + C::O_AB;
    double* s_hr = sm + C::O_HR;
    double* s_cr = sm + C::O_CR;
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(sm + C::O_BAR);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(sm + C::O_BAR + 1);
    uint64_t* s_p3bar = reinterpret_cast<uint64_t*>(sm + C::O_BAR + 2);

    PGeo g;
    g.W = a.w;
    g.H = a.h;
    g.xo0 = blockIdx.x * C::OUT_COLS;
    g.c0 = g.xo0 - 4 * C::HL;
    g.y0 = blockIdx.y * th;
    g.y1 = min(g.y0 + th, a.h);
    g.has_coarse = HC;
    g.edge = g.c0 < 0 || g.c0 + PH_PC > a.w;
    const int pair = blockIdx.z;
    const int warp = threadIdx.x >> 5;

    const double* __restrict__ A = a.A + (size_t)pair * a.a_pair_stride;
    const double* __restrict__ B = a.B + (size_t)pair * a.b_pair_stride;
    double* U = a.u + (size_t)pair * a.uv_pair_stride;
    double* V = a.v + (size_t)pair * a.uv_pair_stride;
    unsigned char* M = a.mask ? a.mask + (size_t)pair * a.mask_pair_stride : nullptr;

    // P3 ownership: thread -> (channel, pair of output columns)
    constexpr int V_WARPS = (5 * C::VPAIRS + 31) / 32;
    static_assert((V_WARPS + 3) / 4 * C::TM_SLOT <= C::TM_COLS && 8 * R <= C::TM_SLOT, "tensor-memory slots");
    const bool v_task = threadIdx.x < 5 * C::VPAIRS;
    int voff;
    {
        const int ch = threadIdx.x / C::VPAIRS, q = threadIdx.x - ch * C::VPAIRS;
        const int half = q / C::OUT_LANES, ol = q - half * C::OUT_LANES;
        voff = v_task ? ch * C::PPITCH + half * C::PHALF + 2 * ol : 0;
    }
    if (threadIdx.x == 0) {
        mbar_init(s_bar, 1);
        mbar_init(s_p3bar, V_WARPS);                     // one arrival per V warp and step
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc(s_tmem, C::TM_COLS);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tm_base = *s_tmem;
    const uint32_t tm_mine = tmem_addr(tm_base, (warp >> 2) * C::TM_SLOT);
    if (warp < V_WARPS) {                                // no previous block yet: its raw rows are 0
        const double z[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int c = 0; c < R; ++c) tmem_st4(tm_mine + 8 * c, z);
        tmem_wait_st();
    }
    StatAcc st = {0, 0, 0};

    const int rows = g.y1 - g.y0;
    const int n_steps = (rows + 2 * R + C::WIN - 1) / C::WIN;

    // Schedule of a step (two block-wide barriers):
    //   beta : thread 0 hands the coarse rows of step s+1 to the copy engine; P2(s); then the horizontal
    //          pyrUp pass over the landed rows
    //   gamma: the first V_WARPS warps ("V") run P3(s) and arrive on a shared-memory barrier; every warp
    //          then warps its row of P1(s+1); finally every warp waits for that barrier (P3 done in all
    //          V warps) and solves its row of P4(s).  The latency-bound gathers thus overlap the
    //          shared-memory-bound vertical sums and the arithmetic of the solve.
    // The loop starts at s = -1 (only P0(0), P1(0)) so that every phase has a single call site:
    // the code of a step has to stay small enough for the instruction cache.
    constexpr int O_WARPS = PH_NW - V_WARPS;
    static_assert(C::WIN - O_WARPS < V_WARPS, "not enough warps for the rows of P1");
    constexpr int N_VR = C::WIN - O_WARPS > 0 ? C::WIN - O_WARPS : 0;     // V warps with a P1 row
    // rows of P1: O warps take WIN, WIN-1, ...; V warps 1, 2, ...; the last V warp has no row in
    // steady state and takes row 0, which only exists in the first step
    const int my_row = warp >= V_WARPS ? C::WIN - (warp - V_WARPS)
                                       : (1 + warp <= C::WIN - O_WARPS ? 1 + warp : (warp == V_WARPS - 1 ? 0 : -1));
    // rows of P4: block row i goes to the warp that warps row i + 1 in P1 (one P1 row and one P4 row per
    // warp and step); PH_P4_MODE 0: the O warps and the V warps without a P1 row share them round robin
#ifndef PH_P4_MODE
#define PH_P4_MODE 1
#endif
    constexpr int P4_N = PH_P4_MODE ? C::WIN : O_WARPS + (V_WARPS - N_VR);
    const int p4_idx = PH_P4_MODE ? (my_row >= 1 ? my_row - 1 : -1)
                                  : (warp >= V_WARPS ? warp - V_WARPS : (warp >= N_VR ? O_WARPS + warp - N_VR : -1));
    Step cur = step_geo<R>(g, n_steps, 0);
    cur.any = false;
    int k_have = -1000;                                // highest coarse row in the ring
    unsigned up_par = 0;                               // phase parity of the landing-buffer barrier
    unsigned p3_par = 0;                               // ... and of the P3 barrier
#ifdef PH_PROF
    long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long prof_t = clock64();
#endif
    for (int s = -1; s < n_steps; ++s) {
        const Step nxt = step_geo<R>(g, n_steps, s + 1);
        PH_MARK(7);
        {
            // coarse rows of the next step that are not in the ring yet (consecutive steps overlap)
            const bool up = nxt.any && g.has_coarse;
            const int k_from = max(nxt.ka, k_have + 1), n_new = max(nxt.ka + nxt.nk - k_from, 0);
            if (up && threadIdx.x == 0)
                up_rows_issue(&tmu, &tmv, (g.c0 >> 1) - 2, a.hc, pair, k_from, n_new, s_cr, s_bar);
            if (s >= 0) phase_products<R>(g, cur.vy0, cur.sb, s_ab, s_prod);
            if (up) {
                mbar_wait(s_bar, up_par);
                up_par ^= 1u;
                up_rows_hpass<PH_NT>(g, a.wc, k_from, n_new, s_cr, s_hr);
                k_have = nxt.ka + nxt.nk - 1;
            }
        }
        PH_MARK(0);
        __syncthreads();
        PH_MARK(1);
        const int m0 = C::WIN * (s - 1) + 1;           // output row (relative to y0) of block row 0
        const int i_lo = max(0, -m0), i_hi = s >= 0 ? min(C::WIN, rows - m0) : 0;
        if (warp < V_WARPS && s >= 0) {
#ifndef PH_KNOCK_P3
            phase_vertical<R>(s_prod + voff, tm_mine, v_task);
#endif
            // "my part of P3 is done": an arrival on a shared-memory barrier, which nobody waits for before P4,
            // so that a V warp goes on to its P1 row at once instead of waiting for the slowest V warp here
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(s_p3bar);
        }
        PH_MARK(2);
        {
            const int r_mine = nxt.any && my_row >= nxt.r_lo ? my_row : C::WIN + 1;
            phase_remap<R>(a, g, A, B, U, V, nxt.vy0, r_mine, C::WIN + 1, nxt.sb, s_hr, s_ab);
        }
        PH_MARK(3);
        if (s >= 0) {
            if (p4_idx >= 0) mbar_wait(s_p3bar, p3_par);     // P3 done in all V warps
            p3_par ^= 1u;
            PH_MARK(4);
#ifndef PH_KNOCK_P4
            if (p4_idx >= 0)
#else
            if (p4_idx >= 1000)
#endif
                for (int i = i_lo + p4_idx; i < i_hi; i += P4_N)
                    phase_solve_row<R, VEC>(a, g, U, V, M, g.y0 + m0 + i, s_prod + i * (5 * C::PPITCH), st);
        }
        PH_MARK(5);
        __syncthreads();
        PH_MARK(6);
        cur = nxt;
    }
#ifdef PH_PROF
    if (blockIdx.x == 1 && blockIdx.y == 0 && blockIdx.z == 0 && (threadIdx.x & 31) == 0 && a.w >= 1920)
        for (int i = 0; i < 8; ++i) atomicAdd(&g_phase_prof[warp * 8 + i], (unsigned long long)prof_acc[i]);
#endif
    if (a.stats) stats_flush(a, pair, st);
    if (warp == 0) tmem_dealloc(tm_base, C::TM_COLS);    // after the loop's last block-wide barrier
}

// ------------------------------------------------------------------------------ kernel, R = 1
// Same strip march with WIN = 15 rows per step.  Schedule of a step (two block-wide barriers):
//   beta : coarse rows of step s+1 -> copy engine; P2(s) raw products -> ring; horizontal pyrUp pass
//   gamma: every warp warps its row of P1(s+1) and solves its output row of P4(s); odd warps
//          solve first, so that the latency-bound gathers of one half overlap the FP64 work
//          of the other.
// Output row n - 2 (relative to y0) is produced when product row n = s WIN + i has arrived.
template <bool VEC, bool HC>
__global__ void __launch_bounds__(PhaseCfg<1>::NT, PhaseCfg<1>::CTAS)
    lk_level_phase1_kernel(const __grid_constant__ CUtensorMap tmu, const __grid_constant__ CUtensorMap tmv, LevelArgs a,
                           int th)
{
    constexpr int R = 1;
    using C = PhaseCfg<1>;
    extern __shared__ __align__(128) double sm[];
    double* s_prod = sm;
    double* s_ab = sm + C::O_AB;
    double* s_hr = sm + C::O_HR;
    double* s_cr = sm + C::O_CR;
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(sm + C::O_BAR);

    PGeo g;
    g.W = a.w;
    g.H = a.h;
    g.xo0 = blockIdx.x * C::OUT_COLS;
    g.c0 = g.xo0 - 4 * C::HL;
    g.y0 = blockIdx.y * th;
    g.y1 = min(g.y0 + th, a.h);
    g.has_coarse = HC;
    g.edge = g.c0 < 0 || g.c0 + PH_PC > a.w;
    const int pair = blockIdx.z;

    const double* __restrict__ A = a.A + (size_t)pair * a.a_pair_stride;
    const double* __restrict__ B = a.B + (size_t)pair * a.b_pair_stride;
    double* U = a.u + (size_t)pair * a.uv_pair_stride;
    double* V = a.v + (size_t)pair * a.uv_pair_stride;
    unsigned char* M = a.mask ? a.mask + (size_t)pair * a.mask_pair_stride : nullptr;
    StatAcc st = {0, 0, 0};
    if (threadIdx.x == 0) {
        mbar_init(s_bar, 1);
        mbar_fence_init();
    }
    __syncthreads();

    const int rows = g.y1 - g.y0;
    const int n_steps = (rows + 2 * R + C::WIN - 1) / C::WIN;
    const int warp = threadIdx.x >> 5;

    // rows of P1: warp w takes row w + 1; the last warp takes row 0, which only exists in the first step
    const int my_row = warp < C::WIN ? warp + 1 : 0;
    Step cur = step_geo<1>(g, n_steps, 0);
    cur.any = false;
    int k_have = -1000;
    unsigned up_par = 0;
    for (int s = -1; s < n_steps; ++s) {
        const Step nxt = step_geo<1>(g, n_steps, s + 1);
        const int n0 = s >= 0 ? (s * C::WIN) % C::PR_ROWS : 0;
        {
            const bool up = nxt.any && g.has_coarse;
            const int k_from = max(nxt.ka, k_have + 1), n_new = max(nxt.ka + nxt.nk - k_from, 0);
            if (up && threadIdx.x == 0)
                up_rows_issue(&tmu, &tmv, (g.c0 >> 1) - 2, a.hc, pair, k_from, n_new, s_cr, s_bar);
            if (s >= 0) phase_products1(g, cur.vy0, cur.sb, n0, s_ab, s_prod);
            if (up) {
                mbar_wait(s_bar, up_par);
                up_par ^= 1u;
                up_rows_hpass<C::NT>(g, a.wc, k_from, n_new, s_cr, s_hr);
                k_have = nxt.ka + nxt.nk - 1;
            }
        }
        __syncthreads();
        const int r_mine = nxt.any && my_row >= nxt.r_lo ? my_row : C::WIN + 1;
        // this warp's output row: product row n = s WIN + warp has arrived -> output row n - 2
        const int n = s * C::WIN + warp;
        const bool solve = s >= 0 && warp < C::WIN && n >= 2 && n - 2 < rows;
        int r2 = n0 + warp;
        if (r2 >= C::PR_ROWS) r2 -= C::PR_ROWS;
        const int r1 = r2 == 0 ? C::PR_ROWS - 1 : r2 - 1;
        const int r0 = r1 == 0 ? C::PR_ROWS - 1 : r1 - 1;
#ifndef PH1_SOLVE_FIRST
#define PH1_SOLVE_FIRST (warp < C::NT / 64)      // two solve-first and two remap-first warps per SM sub-partition
#endif
        if (PH1_SOLVE_FIRST) {
            if (solve) phase_solve1<VEC>(a, g, U, V, M, g.y0 + n - 2, r0, r1, r2, s_prod, st);
            phase_remap<R>(a, g, A, B, U, V, nxt.vy0, r_mine, C::WIN + 1, nxt.sb, s_hr, s_ab);
        } else {
            phase_remap<R>(a, g, A, B, U, V, nxt.vy0, r_mine, C::WIN + 1, nxt.sb, s_hr, s_ab);
            if (solve) phase_solve1<VEC>(a, g, U, V, M, g.y0 + n - 2, r0, r1, r2, s_prod, st);
        }
        __syncthreads();
        cur = nxt;
    }
    if (a.stats) stats_flush(a, pair, st);
}

template <int R, bool VEC, bool HC>
cudaError_t launch_phase_v(const LevelArgs& a, int n_pairs, cudaStream_t s, unsigned long long* launches)
{
    using C = PhaseCfg<R>;
    static SmemOptIn opt_in;
    {
        cudaError_t e;
        if constexpr (R == 1) e = opt_in.ensure((const void*)lk_level_phase1_kernel<VEC, HC>, C::SMEM_BYTES);
        else e = opt_in.ensure((const void*)lk_level_phase_kernel<R, VEC, HC>, C::SMEM_BYTES);
        if (e != cudaSuccess) return e;
    }
    // tensor maps of the coarse flow planes (x, y, pair), one box = CR_W columns of one row
    CUtensorMap tmu, tmv;
    memset(&tmu, 0, sizeof tmu);
    memset(&tmv, 0, sizeof tmv);
    if constexpr (HC) {
        cudaError_t e = tma_map_planes_f64(&tmu, a.cu, a.wc, a.hc, a.c_pitch, a.c_pair_stride, n_pairs, CR_W, 1);
        if (e == cudaSuccess) e = tma_map_planes_f64(&tmv, a.cv, a.wc, a.hc, a.c_pitch, a.c_pair_stride, n_pairs, CR_W, 1);
        if (e != cudaSuccess) return e;
    }
    const int strips = (a.w + C::OUT_COLS - 1) / C::OUT_COLS;
    // Chunk height: a multiple of the vertical block length, so that block boundaries sit at
    // the same absolute rows whatever the chunking (bit-identical results for any batch size).
    // Pick the height that minimises (waves of resident CTAs) x (steps a CTA runs).
    long resident = current_sm_count();
    if constexpr (R == 1) resident *= C::CTAS;
    const int max_m = (a.h + C::WIN - 1) / C::WIN;
    long best_cost = -1;
    int th = C::WIN * max_m;
    for (int m = 1; m <= max_m; ++m) {
        const int t = C::WIN * m;
        const long chunks = (a.h + t - 1) / t;
        const long ctas = (long)strips * chunks * n_pairs;
        const long waves = (ctas + resident - 1) / resident;
        const long steps = (t + 2 * R + C::WIN - 1) / C::WIN;
        const long cost = waves * steps;
        if (best_cost < 0 || cost <= best_cost) { best_cost = cost; th = t; }
    }
    const int chunks = (a.h + th - 1) / th;
    dim3 grid(strips, chunks, n_pairs);
    if constexpr (R == 1) lk_level_phase1_kernel<VEC, HC><<<grid, C::NT, C::SMEM_BYTES, s>>>(tmu, tmv, a, th);
    else lk_level_phase_kernel<R, VEC, HC><<<grid, C::NT, C::SMEM_BYTES, s>>>(tmu, tmv, a, th);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

// VEC: every flow row of every pair is 16-byte aligned (true for the library's own planes and for dense
// caller buffers), so the epilogue uses 16-byte accesses without a second, predicated code path.
template <int R>
cudaError_t launch_phase(const LevelArgs& a, int n_pairs, cudaStream_t s, unsigned long long* launches)
{
    const bool vec = ((reinterpret_cast<uintptr_t>(a.u) | reinterpret_cast<uintptr_t>(a.v) | (a.uv_pitch * 8) |
                       (n_pairs > 1 ? a.uv_pair_stride * 8 : 0)) & 15) == 0;
    if (a.cu) return vec ? launch_phase_v<R, true, true>(a, n_pairs, s, launches) : launch_phase_v<R, false, true>(a, n_pairs, s, launches);
    return vec ? launch_phase_v<R, true, false>(a, n_pairs, s, launches) : launch_phase_v<R, false, false>(a, n_pairs, s, launches);
}

}  // namespace

// Handles R = 1..7 on even-width levels (statistics-only calls come with scratch flow planes from the
// caller, of_api.cu); tiny / odd-sized levels and larger radii go to the tile kernel.  The coarse flow
// planes are read through tensor maps: 16-byte aligned base, even pitch / pair stride (the library's
// own scratch planes always are).
cudaError_t launch_lk_level_phase(const LevelArgs& a, int n_pairs, cudaStream_t s,
                                  unsigned long long* launches, bool* handled)
{
    *handled = false;
    if (!a.u || !a.v) return cudaSuccess;
    if (a.w < 4 || a.h < 2 || (a.w & 1)) return cudaSuccess;
    if (a.cu && (a.wc < 2 || a.w != 2 * a.wc || a.h != 2 * a.hc)) return cudaSuccess;
    if (a.cu && (((reinterpret_cast<uintptr_t>(a.cu) | reinterpret_cast<uintptr_t>(a.cv)) & 15) || (a.c_pitch & 1) ||
                 (a.c_pair_stride & 1) || !tma_encoder()))
        return cudaSuccess;
    if ((unsigned long long)a.img_pitch * (unsigned long long)a.h >= (1ull << 31)) return cudaSuccess;
    if ((unsigned long long)a.uv_pitch * (unsigned long long)a.h >= (1ull << 31)) return cudaSuccess;
    *handled = true;
    switch (a.radius) {
        case 1: return launch_phase<1>(a, n_pairs, s, launches);
        case 2: return launch_phase<2>(a, n_pairs, s, launches);
        case 3: return launch_phase<3>(a, n_pairs, s, launches);
        case 4: return launch_phase<4>(a, n_pairs, s, launches);
        case 5: return launch_phase<5>(a, n_pairs, s, launches);
        case 6: return launch_phase<6>(a, n_pairs, s, launches);
        case 7: return launch_phase<7>(a, n_pairs, s, launches);
        default: break;
    }
    *handled = false;
    return cudaSuccess;
}

}  // namespace ofb

#ifdef PH_PROF
#ifdef PH_PROF1
extern "C" __attribute__((visibility("default"))) int of_b200_debug_p1_prof(unsigned long long* out, int reset)
{
    cudaError_t e = cudaMemcpyFromSymbol(out, ofb::g_p1_prof, sizeof(unsigned long long) * 32 * 8);
    if (e == cudaSuccess && reset) {
        static unsigned long long z[32 * 8];
        e = cudaMemcpyToSymbol(ofb::g_p1_prof, z, sizeof(z));
    }
    return (int)e;
}
#endif
extern "C" __attribute__((visibility("default"))) int of_b200_debug_phase_prof(unsigned long long* out, int reset)
{
    cudaError_t e = cudaMemcpyFromSymbol(out, ofb::g_phase_prof, sizeof(unsigned long long) * 32 * 8);
    if (e == cudaSuccess && reset) {
        static unsigned long long z[32 * 8];
        e = cudaMemcpyToSymbol(ofb::g_phase_prof, z, sizeof(z));
    }
    return (int)e;
}
#endif
