// Strip-marching fused pyramid-level kernel (warp-specialised; fallback of lk_level_phase.cu for calls without a flow output), window radius R = 2..7.
//
// One launch = one pyramid level of every pair of the batch, everything the reference does
// between two levels (ParallelVideoPyr.cpp:610-703) plus the single-level LK (:327-542):
//   pyrUp(coarse flow) -> float32 maps -> remap of both frames -> fx/fy/ft -> products ->
//   (2R+1)^2 window sums -> guarded 2x2 solve -> += upsampled coarse flow.
//
// A CTA (4 warps) owns a strip of OUT_COLS output columns and marches down a chunk of rows:
//
//   producers (warps 0,1; lane = 4 consecutive columns; warp p takes every other row)
//       cv::pyrUp rows of the coarse flow kept in registers while marching, float32 map,
//       32 gathers per lane issued back to back, blend -> warped A', B' row -> AB ring.
//   consumer  (warp 2; lane = 4 consecutive columns)
//       2x2 gradient stencils (previous row in registers), five products, horizontal window
//       sums in registers + shuffles (additions only), vertical window sums by a van Herk /
//       Gil-Werman block scheme over a (2R)-row shared-memory ring private to each lane,
//       guarded solve with one shared reciprocal -> LK flow row -> LK ring.
//   epilogue  (warp 3; lane = 4 consecutive columns)
//       upsampled coarse flow at the output row, += (:634-637), 16-byte coalesced stores of
//       u, v, optional mask and gesture statistics (:845-873).
//
// Shared-memory rows are "pair-split": a lane's columns (0,1) live at lo[2*lane], its columns
// (2,3) at hi[2*lane], so every 128-bit access of a warp is contiguous and conflict-free.
//
// Window sums never subtract: an all-zero window sums to exactly 0, so det == 0 pixels keep
// the reference's "zero divisor -> 0" behaviour (cv::divide, :514,518) bit for bit.  The
// association order of the (2R+1)^2 terms differs from the oracle's row-major order by design
// (no reference order exists for R > 1); tests bound the effect at 1e-3 px.
//
// HBM traffic per level-l pixel: read A, B (gathered, L2-friendly), read 1/4 coarse u, v,
// write u, v  = 36 B algorithmic; everything else stays on chip.
#include "lk_common.cuh"
#include "of_kernels.h"

namespace ofb {

namespace {

constexpr int NB = 4;          // AB ring depth (rows); even: a slot always belongs to one producer
constexpr int NE = 2;          // LK ring depth (rows)
constexpr int AB_SLOT = 256;   // doubles per AB slot: A_lo[64] A_hi[64] B_lo[64] B_hi[64]

template <int R>
struct StripCfg {
    static constexpr int Q = R / 4, RHO = R % 4;
    static constexpr int HL = Q + 1;                  // halo lanes on each side
    static constexpr int OUT_LANES = 32 - 2 * HL;
    static constexpr int OUT_COLS = 4 * OUT_LANES;
    static constexpr int HALF = OUT_COLS / 2;         // offset of the hi pairs inside a pair-split row
    static constexpr int WIN = 2 * R + 1;             // vertical block length
    static constexpr int VSLOTS = 2 * R;              // block rows 1 .. 2R are stored
    static constexpr int VROW = 5 * OUT_COLS;         // doubles per ring slot (5 channels)
    static constexpr size_t V_DOUBLES = (size_t)VSLOTS * VROW;
    static constexpr size_t AB_DOUBLES = (size_t)NB * AB_SLOT;
    static constexpr size_t LK_DOUBLES = (size_t)NE * 2 * OUT_COLS;
    static constexpr size_t SMEM_BYTES = (V_DOUBLES + AB_DOUBLES + LK_DOUBLES) * 8 + (size_t)NE * 32 * 4;
};

// named barriers (id 0 is __syncthreads)
constexpr int BAR_AB_FULL = 1;               // .. +NB-1
constexpr int BAR_AB_EMPTY = 1 + NB;         // .. +NB-1
constexpr int BAR_LK_FULL = 1 + 2 * NB;      // .. +NE-1
constexpr int BAR_LK_EMPTY = 1 + 2 * NB + NE;

__device__ __forceinline__ void bar_sync(int id, int n)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int n)
{
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}

struct Geo {          // per-CTA geometry, identical in all roles
    int W, H;
    int c0;           // image column of (lane 0, i 0); multiple of 4 (may be negative)
    int xo0;          // first output column of the strip
    int y0, y1;       // output rows [y0, y1)
    int vy_first;     // first real product row = max(y0 - R, 0)
    int vy_last;      // last real product row  = min(y1 - 1 + R, H - 1)
    int n_deliv;      // rows the producers deliver = 1 (the row above vy_first) + real rows
    bool has_coarse;
    bool edge;        // strip touches the left or right image border (pyrUp border formulas)
};

__device__ __forceinline__ double2 lds2(const double* p) { return *reinterpret_cast<const double2*>(p); }
__device__ __forceinline__ void sts2(double* p, double x, double y) { *reinterpret_cast<double2*>(p) = make_double2(x, y); }

// pair-split 4-vector: p points at the lane's lo pair, the hi pair is `half` doubles further
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

// ---------------------------------------------------------------- cv::pyrUp while marching
// Horizontal pass of cv::pyrUp (:646-647) for the lane's four fine columns xb .. xb+3 from one
// coarse row: e0 o0 e1 o1.  j0 = xb / 2 (even), n = coarse width >= 2.  Warp-collective.
struct H4 { double v[4]; };
struct RawRow { double s0, s1, se; };       // coarse values j0, j0+1 and (lane 0: j0-1, lane 31: j0+2)

struct UpCtx {
    const double* __restrict__ cu;
    const double* __restrict__ cv;
    size_t cp;
    int wc, hc;
    int j0, xb;
    bool col_ok, edge;
};

// Issues the global loads of one coarse row (no use of the results: they can stay in flight).
__device__ __forceinline__ RawRow raw_load(const double* __restrict__ row, const UpCtx& c)
{
    const int lane = threadIdx.x & 31;
    RawRow r;
    r.s0 = r.s1 = r.se = 0.0;
    if (c.col_ok) {
        if (c.j0 + 1 < c.wc) {
            const double2 t = __ldg(reinterpret_cast<const double2*>(row + c.j0));
            r.s0 = t.x;
            r.s1 = t.y;
        } else {
            r.s0 = __ldg(row + c.j0);
        }
        if (lane == 0 && c.j0 > 0) r.se = __ldg(row + c.j0 - 1);
        if (lane == 31 && c.j0 + 2 < c.wc) r.se = __ldg(row + c.j0 + 2);
    }
    return r;
}

__device__ __forceinline__ H4 hrow_finish(const RawRow& r, const UpCtx& c)
{
    const int lane = threadIdx.x & 31;
    const int n = c.wc, j0 = c.j0;
    const double s0 = r.s0, s1 = r.s1;
    double sl = __shfl_up_sync(0xffffffffu, s1, 1);
    double sr = __shfl_down_sync(0xffffffffu, s0, 1);
    if (lane == 0) sl = r.se;
    if (lane == 31) sr = r.se;
    H4 h;
    h.v[0] = dadd(dadd(sl, dmul(s0, 6.0)), s1);
    h.v[1] = dmul(dadd(s0, s1), 4.0);
    h.v[2] = dadd(dadd(s0, dmul(s1, 6.0)), sr);
    h.v[3] = dmul(dadd(s1, sr), 4.0);
    if (c.edge) {                                 // warp-uniform: only strips touching a border
        if (c.xb == 0) h.v[0] = dadd(dmul(s0, 6.0), dmul(s1, 2.0));
        if (j0 + 1 == n - 1) { h.v[2] = dadd(s0, dmul(s1, 7.0)); h.v[3] = dmul(s1, 8.0); }
        if (j0 == n - 1) { h.v[0] = dadd(sl, dmul(s0, 7.0)); h.v[1] = dmul(s0, 8.0); }
    }
    return h;
}

// Used by the (cold) full reload of the three rows.  (Measured: moving it out of line costs 16 %.)
__device__ __forceinline__ H4 pyrup_hrow4(const double* __restrict__ row, int j0, int xb, int wc, bool col_ok, bool edge)
{
    UpCtx c;
    c.cu = c.cv = nullptr;
    c.cp = 0;
    c.wc = wc;
    c.hc = 0;
    c.j0 = j0;
    c.xb = xb;
    c.col_ok = col_ok;
    c.edge = edge;
    return hrow_finish(raw_load(row, c), c);
}

struct UpRows {       // pyrUp rows rho(k-1), k, rho(k+1) for the lane's 4 columns, u and v
    H4 u0, u1, u2, v0, v1, v2;
    RawRow pu, pv;    // coarse row rho(pk + 1), loads possibly still in flight
    int pk;           // k the prefetch was issued for (-100: none)
};

// Issues the loads of the coarse row the step after `k` will need and leaves them in flight in
// registers, so that their latency overlaps the gathers of the current row.  (Measured: an L1
// prefetch instruction instead is slower than no prefetch at all -- L1 is tiny next to 225 KB of
// shared memory.)
__device__ __forceinline__ void up_prefetch(UpRows& r, const UpCtx& c, int knext)
{
    const size_t o2 = (size_t)pyrup_rho(knext + 1, c.hc) * c.cp;
    r.pu = raw_load(c.cu + o2, c);
    r.pv = raw_load(c.cv + o2, c);
    r.pk = knext;
}

__device__ __forceinline__ void up_advance(UpRows& r, const UpCtx& c, int k, int kprev)
{
    if (k == kprev + 1) {
        r.u0 = r.u1; r.u1 = r.u2;
        r.v0 = r.v1; r.v1 = r.v2;
    } else {
        const size_t o0 = (size_t)pyrup_rho(k - 1, c.hc) * c.cp, o1 = (size_t)k * c.cp;
        r.u0 = pyrup_hrow4(c.cu + o0, c.j0, c.xb, c.wc, c.col_ok, c.edge);
        r.v0 = pyrup_hrow4(c.cv + o0, c.j0, c.xb, c.wc, c.col_ok, c.edge);
        r.u1 = pyrup_hrow4(c.cu + o1, c.j0, c.xb, c.wc, c.col_ok, c.edge);
        r.v1 = pyrup_hrow4(c.cv + o1, c.j0, c.xb, c.wc, c.col_ok, c.edge);
    }
    if (r.pk != k) up_prefetch(r, c, k);
    r.u2 = hrow_finish(r.pu, c);
    r.v2 = hrow_finish(r.pv, c);
}

// Vertical pass of cv::pyrUp.  ((a+b)*4)*(1/64) == (a+b)*(1/16): both scalings are exact.
__device__ __forceinline__ void up_eval(const UpRows& r, int y, double (&upu)[4], double (&upv)[4])
{
    if (y & 1) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            upu[i] = dmul(dadd(r.u1.v[i], r.u2.v[i]), 0.0625);
            upv[i] = dmul(dadd(r.v1.v[i], r.v2.v[i]), 0.0625);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            upu[i] = dmul(dadd(dadd(r.u0.v[i], dmul(r.u1.v[i], 6.0)), r.u2.v[i]), 0.015625);
            upv[i] = dmul(dadd(dadd(r.v0.v[i], dmul(r.v1.v[i], 6.0)), r.v2.v[i]), 0.015625);
        }
    }
}

// ------------------------------------------------------------------------------ producers
// Rare path of the producers: a pixel with a tap outside the image (BORDER_CONSTANT 0 tested
// per tap) or a non-finite map value.  Out of line: it must not widen the hot loop (the hot
// loops of all roles have to fit the 32 KB instruction cache together).
__device__ __noinline__ double remap_cold(const double* __restrict__ F, size_t pitch, int W, int H, float mx, float my)
{
    return remap_tap4(F, pitch, W, H, mx, my);
}

template <int R>
__device__ __forceinline__ void producer_role(const LevelArgs& a, const Geo& g, int pair, double* s_ab, int par)
{
    const int lane = threadIdx.x & 31;
    const int xb = g.c0 + 4 * lane;
    const bool col_ok = xb >= 0 && xb < g.W;          // W is even and xb % 4 == 0: xb, xb+1 valid
    const bool full4 = xb + 3 < g.W;                  // else W % 4 == 2 and only two columns exist
    const int W = g.W, H = g.H;
    const int pitch = (int)a.img_pitch;
    const double* __restrict__ A = a.A + (size_t)pair * a.a_pair_stride;
    const double* __restrict__ B = a.B + (size_t)pair * a.b_pair_stride;
    const bool vec_in = ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) |
                          (a.img_pitch * 8)) & 15) == 0;
    UpCtx uc;
    uc.cu = g.has_coarse ? a.cu + (size_t)pair * a.c_pair_stride : nullptr;
    uc.cv = g.has_coarse ? a.cv + (size_t)pair * a.c_pair_stride : nullptr;
    uc.cp = a.c_pitch;
    uc.wc = a.wc;
    uc.hc = a.hc;
    uc.j0 = xb >> 1;
    uc.xb = xb;
    uc.col_ok = col_ok;
    uc.edge = g.edge;
    UpRows up;
    up.pk = -100;
    int kprev = -100;
    const float xf = (float)xb;

    for (int d = par; d < g.n_deliv; d += 2) {
        const int yy = (d == 0) ? r101(g.vy_first - 1, H) : g.vy_first + d - 1;
        const int slot = d & (NB - 1);
        double* ra = s_ab + slot * AB_SLOT + 2 * lane;
        if (g.has_coarse) {
            const int k = yy >> 1;
            if (k != kprev) up_advance(up, uc, k, kprev);
            kprev = k;
            float hu[4], hv[4];                         // half flow, float32 (map build :667-684)
            unsigned finite = 0;
            {
                double upu[4], upv[4];
                up_eval(up, yy, upu, upv);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    hu[i] = __fmul_rn(__double2float_rn(upu[i]), 0.5f);   // == f / 2 exactly
                    hv[i] = __fmul_rn(__double2float_rn(upv[i]), 0.5f);
                    finite |= (fabsf(hu[i]) + fabsf(hv[i]) < __int_as_float(0x7f800000)) ? (1u << i) : 0u;
                }
            }
            const float yf = (float)yy;
            // One pass per frame (rolled: the loop body is the hot code of this role): frame 1 is
            // sampled at (x + hu, y + hv), frame 2 at (x - hu, y - hv); x - h == x + (-h) exactly.
#pragma unroll 1
            for (int f = 0; f < 2; ++f) {
                const double* __restrict__ F = f ? B : A;
                const float sg = f ? -1.0f : 1.0f;
                Tap t[4];
                unsigned all_in = 0x10000;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float mx = __fadd_rn(xf + (float)i, __fmul_rn(sg, hu[i]));
                    const float my = __fadd_rn(yf, __fmul_rn(sg, hv[i]));
                    t[i] = make_tap(mx, my, W, H, pitch, (finite >> i) & 1u);
                    all_in &= t[i].fr;
                }
                double sv[4][4];                        // 16 gathers per lane, issued back to back
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const double* pa = F + t[i].off;
                    sv[i][0] = __ldg(pa); sv[i][1] = __ldg(pa + 1);
                    sv[i][2] = __ldg(pa + pitch); sv[i][3] = __ldg(pa + pitch + 1);
                }
                if (f == 0) up_prefetch(up, uc, k + 1);   // coarse row of this producer's next step
                double v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = blend4(sv[i][0], sv[i][1], sv[i][2], sv[i][3], t[i].fr);
                if (!all_in && col_ok) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (!(t[i].fr & 0x10000)) {
                            const float mx = __fadd_rn(xf + (float)i, __fmul_rn(sg, hu[i]));
                            const float my = __fadd_rn(yf, __fmul_rn(sg, hv[i]));
                            v[i] = remap_cold(F, a.img_pitch, W, H, mx, my);
                        }
                    }
                }
                if (!col_ok) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[i] = 0.0;
                } else if (!full4) {
                    v[2] = v[3] = 0.0;
                }
                if (f == 0 && d >= NB) bar_sync(BAR_AB_EMPTY + slot, 64);
                st4s(ra + f * 128, 64, v);
            }
        } else {
            double va[4], vb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) va[i] = vb[i] = 0.0;
            if (col_ok) {
                const double* pa = A + (size_t)yy * a.img_pitch + xb;
                const double* pb = B + (size_t)yy * a.img_pitch + xb;
                if (vec_in && full4) {
                    const double2 a0 = __ldg(reinterpret_cast<const double2*>(pa));
                    const double2 a1 = __ldg(reinterpret_cast<const double2*>(pa + 2));
                    const double2 b0 = __ldg(reinterpret_cast<const double2*>(pb));
                    const double2 b1 = __ldg(reinterpret_cast<const double2*>(pb + 2));
                    va[0] = a0.x; va[1] = a0.y; va[2] = a1.x; va[3] = a1.y;
                    vb[0] = b0.x; vb[1] = b0.y; vb[2] = b1.x; vb[3] = b1.y;
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (xb + i < W) { va[i] = __ldg(pa + i); vb[i] = __ldg(pb + i); }
                }
            }
            if (d >= NB) bar_sync(BAR_AB_EMPTY + slot, 64);
            st4s(ra, 64, va);
            st4s(ra + 128, 64, vb);
        }
        bar_arrive(BAR_AB_FULL + slot, 64);
    }
}

// ------------------------------------------------------------------------------ consumer
template <int R>
__device__ __forceinline__ void consumer_role(const LevelArgs& a, const Geo& g, const double* s_ab,
                                              double* s_v, double* s_lk, unsigned* s_mk)
{
    using C = StripCfg<R>;
    const int lane = threadIdx.x & 31;
    const bool active = lane >= C::HL && lane < 32 - C::HL;
    const int xb = g.c0 + 4 * lane;                 // image column of i = 0
    const int ol = active ? lane - C::HL : 0;       // output lane
    const int W = g.W, H = g.H;
    const double* ab_lane = s_ab + 2 * lane;
    double* v_lane = s_v + 2 * ol;
    double* lk_lane = s_lk + 2 * ol;
    const double det_eps = a.det_eps;

    double prevA[4], prevB[4], prevLA, prevLB;
    double P[5][4];
#pragma unroll
    for (int c = 0; c < 5; ++c)
#pragma unroll
        for (int i = 0; i < 4; ++i) P[c][i] = 0.0;

    int dd = 0;   // deliveries consumed so far
    {   // the row above the first product row (reflect-101 at the image top)
        bar_sync(BAR_AB_FULL + 0, 64);
        ld4s(ab_lane, 64, prevA);
        ld4s(ab_lane + 128, 64, prevB);
        prevLA = ab_lane[64 - 1];
        prevLB = ab_lane[192 - 1];
        if (dd + NB < g.n_deliv) bar_arrive(BAR_AB_EMPTY + 0, 64);
        dd = 1;
    }

    const int th = g.y1 - g.y0;
    const int n_virtual = th + 2 * R;
    int ib = 0;
    double* vp = v_lane;                             // ring slot ib
    for (int n = 0; n < n_virtual; ++n) {
        const int vy = g.y0 - R + n;
        double fx[4], fy[4], ft[4];
        if (vy >= 0 && vy < H) {
            const int slot = dd & (NB - 1);
            bar_sync(BAR_AB_FULL + slot, 64);
            const double* ra = ab_lane + slot * AB_SLOT;
            double curA[4], curB[4];
            ld4s(ra, 64, curA);
            ld4s(ra + 128, 64, curB);
            double curLA = ra[64 - 1];
            double curLB = ra[192 - 1];
            if (dd + NB < g.n_deliv) bar_arrive(BAR_AB_EMPTY + slot, 64);
            ++dd;
            if (xb == 0) {                           // reflect-101 at the left image border
                const bool w1 = W > 1;
                curLA = w1 ? curA[1] : curA[0];
                curLB = w1 ? curB[1] : curB[0];
                prevLA = w1 ? prevA[1] : prevA[0];
                prevLB = w1 ? prevB[1] : prevB[0];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const double a00 = i ? prevA[i - 1] : prevLA, a01 = prevA[i];
                const double a10 = i ? curA[i - 1] : curLA, a11 = curA[i];
                const double b00 = i ? prevB[i - 1] : prevLB, b01 = prevB[i];
                const double b10 = i ? curB[i - 1] : curLB, b11 = curB[i];
                // filter2D accumulation order ((k00 p00 + k01 p01) + k10 p10) + k11 p11 (:44-84)
                const double fxa = dadd(dsub(dsub(a01, a00), a10), a11);
                const double fxb = dadd(dsub(dsub(b01, b00), b10), b11);
                const double sa = dadd(a00, a01), sb = dadd(b00, b01);
                const double fya = dadd(dsub(a10, sa), a11);
                const double fyb = dadd(dsub(b10, sb), b11);
                const double fta = -dadd(dadd(sa, a10), a11);
                const double ftb = dadd(dadd(sb, b10), b11);
                fx[i] = dadd(fxa, fxb);
                fy[i] = dadd(fya, fyb);
                ft[i] = dadd(fta, ftb);
            }
            if (g.edge) {                            // columns outside the image contribute nothing
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const bool in = (unsigned)(xb + i) < (unsigned)W;
                    fx[i] = in ? fx[i] : 0.0;
                    fy[i] = in ? fy[i] : 0.0;
                    ft[i] = in ? ft[i] : 0.0;
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) { prevA[i] = curA[i]; prevB[i] = curB[i]; }
            prevLA = curLA;
            prevLB = curLB;
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) fx[i] = fy[i] = ft[i] = 0.0;
        }

        const bool emit = n >= C::WIN - 1;
        const bool last = ib == C::WIN - 1;
        // Horizontal sums of the five products for all lanes, in uniform control flow (the
        // shuffles of the five channels interleave freely).
        double h[5][4];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            double x[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                // fx2, fy2, fxfy, fxft, fyft (:383-405)
                const double p = (c == 1 || c == 4) ? fy[i] : fx[i];
                const double q = c == 0 ? fx[i] : (c <= 2 ? fy[i] : ft[i]);
                x[i] = dmul(p, q);
            }
            hsum4<R>(x, h[c]);
        }
        // Vertical van Herk / Gil-Werman: P = running prefix of the current block (reset to 0
        // when a block ends; 0 + h == h), the ring holds the suffix sums of the previous block,
        // overwritten row by row.  Only loads and stores are predicated -- no divergent branch.
        const bool ld_suf = active && !last;          // block row WIN-1: the window is the block itself
        const bool st_raw = active && ib >= 1;
        double S[5][4];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            double* vrow = vp + c * C::OUT_COLS;                 // slot ib  <-> block row ib+1
            double sp[4] = {0.0, 0.0, 0.0, 0.0};
            if (ld_suf) ld4s(vrow, C::HALF, sp);                 // suffix of previous block from row ib+1
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                P[c][i] = dadd(P[c][i], h[c][i]);
                S[c][i] = dadd(sp[i], P[c][i]);
            }
            if (st_raw) st4s(vrow - C::VROW, C::HALF, h[c]);     // raw row ib -> slot ib-1
        }

        if (last) {
            if (active) {
                // block finished: turn the stored raw rows 1..2R into suffix sums, in place
                // (the five channels advance together: 20 independent chains)
#pragma unroll 1
                for (int c = 0; c < 5; ++c) {                     // rolled: small code, runs once per block
                    double acc[4];
                    double* vr = v_lane + (C::VSLOTS - 1) * C::VROW + c * C::OUT_COLS;
                    ld4s(vr, C::HALF, acc);
#pragma unroll 2
                    for (int t = C::VSLOTS - 2; t >= 0; --t) {
                        vr -= C::VROW;
                        double raw[4];
                        ld4s(vr, C::HALF, raw);
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[i] = dadd(raw[i], acc[i]);
                        st4s(vr, C::HALF, acc);
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < 5; ++c)
#pragma unroll
                for (int i = 0; i < 4; ++i) P[c][i] = 0.0;
        }

        if (emit) {
            const int m = n - (C::WIN - 1);
            const int y = g.y0 + m;
            const int eslot = m & (NE - 1);
            if (m >= NE) bar_sync(BAR_LK_EMPTY + eslot, 64);
            const bool yin = y >= R && y <= H - 1 - R;
            double lu[4], lv[4], nus[4], nvs[4], dets[4];
            unsigned okbits = 0;
            bool fast = true;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int x = xb + i;
                const bool interior = yin && x >= R && x <= W - 1 - R;
                const double sxx = S[0][i], syy = S[1][i], sxy = S[2][i], sxt = S[3][i], syt = S[4][i];
                dets[i] = dsub(dmul(sxx, syy), dmul(sxy, sxy));
                nus[i] = dsub(dmul(sxy, syt), dmul(syy, sxt));
                nvs[i] = dsub(dmul(sxt, sxy), dmul(sxx, syt));
                const bool ok = interior && fabs(dets[i]) > det_eps;
                okbits |= ok ? (1u << (8 * i)) : 0u;
                fast = (div2_fast(nus[i], nvs[i], dets[i], lu[i], lv[i]) || !ok) && fast;
            }
            if (!fast) {                             // rare: exponent range outside the fast path
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if ((okbits >> (8 * i)) & 1u) {
                        lu[i] = ddiv_slow(nus[i], dets[i]);
                        lv[i] = ddiv_slow(nvs[i], dets[i]);
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const bool ok = (okbits >> (8 * i)) & 1u;
                lu[i] = ok ? lu[i] : 0.0;
                lv[i] = ok ? lv[i] : 0.0;
            }
            if (active) {
                double* lk = lk_lane + eslot * 2 * C::OUT_COLS;
                st4s(lk, C::HALF, lu);
                st4s(lk + C::OUT_COLS, C::HALF, lv);
                s_mk[eslot * 32 + lane] = okbits;
            }
            bar_arrive(BAR_LK_FULL + eslot, 64);
        }
        if (last) { ib = 0; vp = v_lane; } else { ++ib; vp += C::VROW; }
    }
}

// ------------------------------------------------------------------------------ epilogue
template <int R>
__device__ __forceinline__ void epilogue_role(const LevelArgs& a, const Geo& g, int pair,
                                              const double* s_lk, const unsigned* s_mk)
{
    using C = StripCfg<R>;
    const int lane = threadIdx.x & 31;
    const bool active = lane >= C::HL && lane < 32 - C::HL;
    const int xb = g.c0 + 4 * lane;
    const int ol = active ? lane - C::HL : 0;
    const int W = g.W;
    const bool col_ok = xb >= 0 && xb < W;
    const bool store = active && col_ok;
    const bool full4 = xb + 3 < W;
    UpCtx uc;
    uc.cu = g.has_coarse ? a.cu + (size_t)pair * a.c_pair_stride : nullptr;
    uc.cv = g.has_coarse ? a.cv + (size_t)pair * a.c_pair_stride : nullptr;
    uc.cp = a.c_pitch;
    uc.wc = a.wc;
    uc.hc = a.hc;
    uc.j0 = xb >> 1;
    uc.xb = xb;
    uc.col_ok = col_ok;
    uc.edge = g.edge;
    double* __restrict__ U = a.u ? a.u + (size_t)pair * a.uv_pair_stride : nullptr;
    double* __restrict__ V = a.v ? a.v + (size_t)pair * a.uv_pair_stride : nullptr;
    unsigned char* __restrict__ M = a.mask ? a.mask + (size_t)pair * a.mask_pair_stride : nullptr;
    const bool vec_out = U && ((reinterpret_cast<uintptr_t>(U) | reinterpret_cast<uintptr_t>(V) |
                                (a.uv_pitch * 8)) & 15) == 0;
    const double* lk_lane = s_lk + 2 * ol;
    const double mag2 = a.mag2_thresh;
    UpRows up;
    up.pk = -100;
    int kprev = -100;
    unsigned long long st_x = 0, st_y = 0, st_c = 0;   // exact integer partial sums of this lane
    const int th = g.y1 - g.y0;
    for (int m = 0; m < th; ++m) {
        const int y = g.y0 + m;
        const int eslot = m & (NE - 1);
        double upu[4], upv[4];
        if (g.has_coarse) {                          // upsampled coarse flow first (global loads)
            const int k = y >> 1;
            if (k != kprev) {
                up_advance(up, uc, k, kprev);
                up_prefetch(up, uc, k + 1);
            }
            kprev = k;
            up_eval(up, y, upu, upv);
        }
        bar_sync(BAR_LK_FULL + eslot, 64);
        double fu[4], fv[4];
        const double* lk = lk_lane + eslot * 2 * C::OUT_COLS;
        ld4s(lk, C::HALF, fu);
        ld4s(lk + C::OUT_COLS, C::HALF, fv);
        const unsigned okbits = s_mk[eslot * 32 + lane];
        if (m + NE < th) bar_arrive(BAR_LK_EMPTY + eslot, 64);
        if (g.has_coarse) {                          // tmpu += upu; tmpv += upv  (:634-637)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                fu[i] = dadd(fu[i], upu[i]);
                fv[i] = dadd(fv[i], upv[i]);
            }
        }
        if (store) {
            if (U) {
                double* pu = U + (size_t)y * a.uv_pitch + xb;
                double* pv = V + (size_t)y * a.uv_pitch + xb;
                if (vec_out && full4) {
                    *reinterpret_cast<double2*>(pu) = make_double2(fu[0], fu[1]);
                    *reinterpret_cast<double2*>(pu + 2) = make_double2(fu[2], fu[3]);
                    *reinterpret_cast<double2*>(pv) = make_double2(fv[0], fv[1]);
                    *reinterpret_cast<double2*>(pv + 2) = make_double2(fv[2], fv[3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (xb + i < W) { pu[i] = fu[i]; pv[i] = fv[i]; }
                }
            }
            if (M) {
                unsigned char* pm = M + (size_t)y * a.mask_pitch + xb;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (xb + i < W) pm[i] = (unsigned char)((okbits >> (8 * i)) & 1u);
            }
            if (a.stats) {
                unsigned rc = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    // val = sqrt(u^2 + v^2); !(val < mag_thresh)  <=>  !(u^2 + v^2 < mag2)   (:850-873)
                    const double s2 = dadd(dmul(fu[i], fu[i]), dmul(fv[i], fv[i]));
                    if (xb + i < W && !(s2 < mag2)) { st_x += (unsigned)(xb + i); rc += 1u; }
                }
                st_c += rc;
                st_y += (unsigned long long)rc * (unsigned)y;
            }
        }
    }
    if (a.stats) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_x += __shfl_xor_sync(0xffffffffu, st_x, o);
            st_y += __shfl_xor_sync(0xffffffffu, st_y, o);
            st_c += __shfl_xor_sync(0xffffffffu, st_c, o);
        }
        if (lane == 0 && st_c != 0) {                // integer-valued sums: exact in any order
            atomicAdd(&a.stats[pair].sum_x, (double)st_x);
            atomicAdd(&a.stats[pair].sum_y, (double)st_y);
            atomicAdd(&a.stats[pair].count, (double)st_c);
        }
    }
}

// ------------------------------------------------------------------------------ kernel
template <int R>
__global__ void __launch_bounds__(128, 3) lk_level_strip_kernel(LevelArgs a, int th)
{
    using C = StripCfg<R>;
    extern __shared__ __align__(16) double sm[];
    double* s_v = sm;
    double* s_ab = s_v + C::V_DOUBLES;
    double* s_lk = s_ab + C::AB_DOUBLES;
    unsigned* s_mk = reinterpret_cast<unsigned*>(s_lk + C::LK_DOUBLES);

    Geo g;
    g.W = a.w;
    g.H = a.h;
    g.xo0 = blockIdx.x * C::OUT_COLS;
    g.c0 = g.xo0 - 4 * C::HL;
    g.y0 = blockIdx.y * th;
    g.y1 = min(g.y0 + th, a.h);
    g.vy_first = max(g.y0 - R, 0);
    g.vy_last = min(g.y1 - 1 + R, a.h - 1);
    g.n_deliv = g.vy_last - g.vy_first + 2;
    g.has_coarse = a.cu != nullptr;
    g.edge = g.xo0 == 0 || g.c0 + 132 >= a.w;
    const int pair = blockIdx.z;

    const int role = threadIdx.x >> 5;
    if (role <= 1) producer_role<R>(a, g, pair, s_ab, role);
    else if (role == 2) consumer_role<R>(a, g, s_ab, s_v, s_lk, s_mk);
    else epilogue_role<R>(a, g, pair, s_lk, s_mk);
}

template <int R>
cudaError_t launch_strip(const LevelArgs& a, int n_pairs, cudaStream_t s, unsigned long long* launches)
{
    using C = StripCfg<R>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(lk_level_strip_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)C::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int strips = (a.w + C::OUT_COLS - 1) / C::OUT_COLS;
    // Chunk height: a multiple of the vertical block length (2R+1), so that the block
    // boundaries of the vertical sums sit at the same absolute rows whatever the chunking --
    // results are then bit-identical for any batch size.  Pick the height that minimises
    // (waves of resident CTAs) x (rows a CTA marches, including its 2R warm-up rows).
    const long resident = 148L * 3;
    const int max_m = (a.h + C::WIN - 1) / C::WIN;
    long best_cost = -1;
    int th = C::WIN * max_m;
    for (int m = 2; m <= max_m; ++m) {
        const int t = C::WIN * m;
        const long chunks = (a.h + t - 1) / t;
        const long ctas = (long)strips * chunks * n_pairs;
        const long waves = (ctas + resident - 1) / resident;
        const long cost = waves * (t + 2 * R + 6);
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; th = t; }
    }
    if (th > a.h) th = a.h;
    const int chunks = (a.h + th - 1) / th;
    dim3 grid(strips, chunks, n_pairs);
    lk_level_strip_kernel<R><<<grid, 128, C::SMEM_BYTES, s>>>(a, th);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace

// Handles R = 2..7 on levels that are at least one window tall and two coarse columns wide;
// everything else (R = 1 bit-exact order, tiny levels) goes to the generic tile kernel.
cudaError_t launch_lk_level_strip(const LevelArgs& a, int n_pairs, cudaStream_t s,
                                  unsigned long long* launches, bool* handled)
{
    *handled = false;
    if (a.w < 4 || a.h < 2 || (a.w & 1)) return cudaSuccess;
    if (a.cu && (a.wc < 2 || a.w != 2 * a.wc || a.h != 2 * a.hc)) return cudaSuccess;
    if ((unsigned long long)a.img_pitch * (unsigned long long)a.h >= (1ull << 31)) return cudaSuccess;
    *handled = true;
    switch (a.radius) {
        case 2: return launch_strip<2>(a, n_pairs, s, launches);
        case 3: return launch_strip<3>(a, n_pairs, s, launches);
        case 4: return launch_strip<4>(a, n_pairs, s, launches);
        case 5: return launch_strip<5>(a, n_pairs, s, launches);
        case 6: return launch_strip<6>(a, n_pairs, s, launches);
        case 7: return launch_strip<7>(a, n_pairs, s, launches);
        default: break;
    }
    *handled = false;
    return cudaSuccess;
}

}  // namespace ofb
