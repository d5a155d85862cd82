// Strip-marching fused pyramid-level kernel (throughput path), window radius R >= 2.
//
// Same contract as lk_level_tile.cu (one launch = one pyramid level of every pair:
// pyrUp(coarse flow) -> float32 maps -> remap of both frames -> fx/fy/ft -> products ->
// (2R+1)^2 window sums -> guarded 2x2 solve -> += upsampled coarse flow; reference
// ParallelVideoPyr.cpp:610-703 and :327-542), different decomposition:
//
//   A CTA owns a vertical strip of OUT_COLS output columns and marches down a chunk of rows.
//   Three warps, three roles, coupled by named barriers over small shared-memory rings:
//
//   producer  (warp 0, lane = column)    per row: upsampled coarse flow (cv::pyrUp rows kept
//             in registers while marching), float32 map, 8 gathers, blend -> warped A', B' row
//             into the AB ring.  All global gathers live here, ahead of the consumer.
//   consumer  (warp 1, lane = 4 columns) per row: 2x2 gradient stencils (previous row kept in
//             registers), five products, HORIZONTAL window sums entirely in registers +
//             shuffles (group prefix/suffix trick, no subtraction), VERTICAL window sums by a
//             van Herk / Gil-Werman block scheme over a (2R)-row shared-memory buffer that only
//             this lane ever touches (no barrier), guarded solve -> LK flow row into the LK ring.
//   epilogue  (warp 2, lane = column)    per output row: re-derives the upsampled coarse flow
//             at the output pixel, adds it (tmpu += upu, :634-637), coalesced stores of u, v,
//             mask, gesture statistics.
//
// Window sums never subtract: an all-zero window sums to exactly 0, so det == 0 pixels keep
// the reference's "zero divisor -> 0" behaviour (cv::divide, :514,518) bit for bit.  The
// association order of the (2R+1)^2 terms differs from the oracle's row-major order by design
// (no reference order exists for R > 1); tests bound the effect at 1e-3 px.
//
// HBM traffic per level-l pixel: read A, B (gathered, L2-friendly), read 1/4 coarse u, v,
// write u, v  = 36 B algorithmic; everything else stays on chip.
#include "of_device.cuh"
#include "of_kernels.h"

namespace ofb {

namespace {

constexpr int MW = 128;   // product columns per strip: 32 lanes x 4
constexpr int NB = 4;     // AB ring depth (rows)
constexpr int NE = 2;     // LK ring depth (rows)
constexpr int AB_PITCH = MW + 4;   // 2 doubles of front padding keep the lane's 4 columns 32-byte aligned

template <int R>
struct MarchCfg {
    static constexpr int Q = R / 4, RHO = R % 4;
    static constexpr int HL = Q + 1;                  // halo lanes on each side (lane 0 col 0 has no left neighbour)
    static constexpr int OUT_LANES = 32 - 2 * HL;
    static constexpr int OUT_COLS = 4 * OUT_LANES;
    static constexpr int WIN = 2 * R + 1;             // vertical block length
    static constexpr int VSLOTS = 2 * R;              // block rows 1 .. 2R are stored
    static constexpr size_t V_DOUBLES = (size_t)VSLOTS * 5 * OUT_COLS;
    static constexpr size_t AB_DOUBLES = (size_t)NB * 2 * AB_PITCH;
    static constexpr size_t LK_DOUBLES = (size_t)NE * 2 * OUT_COLS;
    static constexpr size_t SMEM_BYTES = (V_DOUBLES + AB_DOUBLES + LK_DOUBLES) * 8 + (size_t)NE * OUT_COLS;
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

struct Geo {          // per-CTA geometry, identical in all three roles
    int W, H;
    int c0;           // image column of (lane 0, i 0)
    int xo0;          // first output column of the strip
    int y0, y1;       // output rows [y0, y1)
    int vy_first;     // first real product row = max(y0 - R, 0)
    int vy_last;      // last real product row  = min(y1 - 1 + R, H - 1)
    int n_deliv;      // rows the producer delivers = 1 (the row above vy_first) + real rows
    bool has_coarse;
};

// cv::pyrUp rows for one fine column, kept in registers while marching down.
struct UpCol {
    double r0u, r1u, r2u, r0v, r1v, r2v;
};

__device__ __forceinline__ void upcol_advance(UpCol& c, int& kcur_unused, int k, int kprev, int X,
                                              const double* __restrict__ cu, const double* __restrict__ cv,
                                              size_t cp, int wc, int hc)
{
    (void)kcur_unused;
    if (k == kprev + 1) {
        c.r0u = c.r1u; c.r1u = c.r2u;
        c.r0v = c.r1v; c.r1v = c.r2v;
        const size_t ro = (size_t)pyrup_rho(k + 1, hc) * cp;
        c.r2u = pyrup_row(cu + ro, 0, X, wc);
        c.r2v = pyrup_row(cv + ro, 0, X, wc);
    } else {
        const size_t o0 = (size_t)pyrup_rho(k - 1, hc) * cp, o1 = (size_t)k * cp, o2 = (size_t)pyrup_rho(k + 1, hc) * cp;
        c.r0u = pyrup_row(cu + o0, 0, X, wc); c.r0v = pyrup_row(cv + o0, 0, X, wc);
        c.r1u = pyrup_row(cu + o1, 0, X, wc); c.r1v = pyrup_row(cv + o1, 0, X, wc);
        c.r2u = pyrup_row(cu + o2, 0, X, wc); c.r2v = pyrup_row(cv + o2, 0, X, wc);
    }
}

// Vertical pass of cv::pyrUp.  ((a+b)*4)*(1/64) == (a+b)*(1/16): both scalings are exact.
__device__ __forceinline__ void upcol_eval(const UpCol& c, int y, double& upu, double& upv)
{
    if (y & 1) {
        upu = dmul(dadd(c.r1u, c.r2u), 0.0625);
        upv = dmul(dadd(c.r1v, c.r2v), 0.0625);
    } else {
        upu = dmul(dadd(dadd(c.r0u, dmul(c.r1u, 6.0)), c.r2u), 0.015625);
        upv = dmul(dadd(dadd(c.r0v, dmul(c.r1v, 6.0)), c.r2v), 0.015625);
    }
}

// ------------------------------------------------------------------------------ producer
template <int R>
__device__ __forceinline__ void producer_role(const LevelArgs& a, const Geo& g, int pair, double* s_ab)
{
    const int lane = threadIdx.x & 31;
    const double* __restrict__ A = a.A + (size_t)pair * a.a_pair_stride;
    const double* __restrict__ B = a.B + (size_t)pair * a.b_pair_stride;
    const double* __restrict__ cu = g.has_coarse ? a.cu + (size_t)pair * a.c_pair_stride : nullptr;
    const double* __restrict__ cv = g.has_coarse ? a.cv + (size_t)pair * a.c_pair_stride : nullptr;
    UpCol up[4];
    int kprev = -100;
    int dummy = 0;
    for (int d = 0; d < g.n_deliv; ++d) {
        const int yy = (d == 0) ? r101(g.vy_first - 1, g.H) : g.vy_first + d - 1;
        const int slot = d % NB;
        if (d >= NB) bar_sync(BAR_AB_EMPTY + slot, 64);
        double* ra = s_ab + (size_t)slot * 2 * AB_PITCH + 2;
        double* rb = ra + AB_PITCH;
        const int k = yy >> 1;
        double va[4], vb[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int x = g.c0 + 32 * p + lane;
            va[p] = 0.0;
            vb[p] = 0.0;
            if (x >= 0 && x < g.W) {
                if (g.has_coarse) {
                    if (k != kprev) upcol_advance(up[p], dummy, k, kprev, x, cu, cv, a.c_pitch, a.wc, a.hc);
                    double upu, upv;
                    upcol_eval(up[p], yy, upu, upv);
                    float m1x, m1y, m2x, m2y;
                    map_coords(x, yy, upu, upv, m1x, m1y, m2x, m2y);
                    va[p] = remap_tap4(A, a.img_pitch, g.W, g.H, m1x, m1y);
                    vb[p] = remap_tap4(B, a.img_pitch, g.W, g.H, m2x, m2y);
                } else {
                    va[p] = __ldg(A + (size_t)yy * a.img_pitch + x);
                    vb[p] = __ldg(B + (size_t)yy * a.img_pitch + x);
                }
            }
        }
        kprev = k;
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            ra[32 * p + lane] = va[p];
            rb[32 * p + lane] = vb[p];
        }
        bar_arrive(BAR_AB_FULL + slot, 64);
    }
}

// ------------------------------------------------------------------------------ consumer
template <int OFF>
__device__ __forceinline__ double lane_fetch(double v)
{
    if constexpr (OFF == 0) return v;
    else if constexpr (OFF < 0) return __shfl_up_sync(0xffffffffu, v, -OFF);
    else return __shfl_down_sync(0xffffffffu, v, OFF);
}

// Where the window [c-R, c+R] of output column c = 4j + I starts and ends in terms of the
// 4-column groups owned by the lanes: partial suffix of group j+JL (LENL columns, 0 = none),
// full groups j+FA .. j+FB, partial prefix of group j+JR (LENR columns, 0 = none).
template <int R, int I>
struct HWin {
    static constexpr int Q = R / 4, RHO = R % 4;
    static constexpr int D = I - RHO, E = I + RHO;
    static constexpr int JL = D >= 0 ? -Q : -(Q + 1);
    static constexpr int LENL = D > 0 ? 4 - D : (D == 0 ? 0 : -D);
    static constexpr int FA = D > 0 ? -Q + 1 : -Q;
    static constexpr int JR = E <= 3 ? Q : Q + 1;
    static constexpr int LENR = E < 3 ? E + 1 : (E == 3 ? 0 : E - 3);
    static constexpr int FB = E < 3 ? Q - 1 : Q;
};

template <int R, int I>
__device__ __forceinline__ double hsum_one(const double (&pre)[4], const double (&suf)[4],
                                           const double (&tj)[2 * (R / 4) + 1])
{
    using Wn = HWin<R, I>;
    constexpr int Q = R / 4;
    double acc = 0.0;
    bool have = false;
    if constexpr (Wn::LENL > 0) {
        acc = lane_fetch<Wn::JL>(suf[Wn::LENL]);
        have = true;
    }
#pragma unroll
    for (int j = -Q; j <= Q; ++j) {
        if (j >= Wn::FA && j <= Wn::FB) {
            acc = have ? dadd(acc, tj[j + Q]) : tj[j + Q];
            have = true;
        }
    }
    if constexpr (Wn::LENR > 0) {
        const double pr = lane_fetch<Wn::JR>(pre[Wn::LENR]);
        acc = have ? dadd(acc, pr) : pr;
    }
    return acc;
}

// Horizontal (2R+1)-column window sums for the lane's four columns: group prefixes and
// suffixes in registers, neighbours' pieces through shuffles; additions only.
template <int R>
__device__ __forceinline__ void hsum4(const double (&x)[4], double (&out)[4])
{
    constexpr int Q = R / 4;
    double pre[4], suf[4], tj[2 * Q + 1];
    pre[1] = x[0];
    pre[2] = dadd(x[0], x[1]);
    pre[3] = dadd(pre[2], x[2]);
    const double T = dadd(pre[3], x[3]);
    suf[1] = x[3];
    suf[2] = dadd(x[2], x[3]);
    suf[3] = dadd(x[1], suf[2]);
    pre[0] = suf[0] = 0.0;
    tj[Q] = T;
    if constexpr (Q >= 1) { tj[Q - 1] = lane_fetch<-1>(T); tj[Q + 1] = lane_fetch<1>(T); }
    if constexpr (Q >= 2) { tj[Q - 2] = lane_fetch<-2>(T); tj[Q + 2] = lane_fetch<2>(T); }
    if constexpr (Q >= 3) { tj[Q - 3] = lane_fetch<-3>(T); tj[Q + 3] = lane_fetch<3>(T); }
    out[0] = hsum_one<R, 0>(pre, suf, tj);
    out[1] = hsum_one<R, 1>(pre, suf, tj);
    out[2] = hsum_one<R, 2>(pre, suf, tj);
    out[3] = hsum_one<R, 3>(pre, suf, tj);
}

__device__ __forceinline__ void ld4(const double* p, double (&v)[4])
{
    const double2 a = *reinterpret_cast<const double2*>(p);
    const double2 b = *reinterpret_cast<const double2*>(p + 2);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4(double* p, const double (&v)[4])
{
    *reinterpret_cast<double2*>(p) = make_double2(v[0], v[1]);
    *reinterpret_cast<double2*>(p + 2) = make_double2(v[2], v[3]);
}

template <int R>
__device__ __forceinline__ void consumer_role(const LevelArgs& a, const Geo& g, const double* s_ab,
                                              double* s_v, double* s_lk, unsigned char* s_mk)
{
    using C = MarchCfg<R>;
    const int lane = threadIdx.x & 31;
    const bool active = lane >= C::HL && lane < 32 - C::HL;
    const int xb = g.c0 + 4 * lane;                 // image column of i = 0
    const int vcol = active ? 4 * (lane - C::HL) : 0;
    const int W = g.W, H = g.H;

    double prevA[4], prevB[4], prevLA, prevLB;
    double P[5][4];
#pragma unroll
    for (int c = 0; c < 5; ++c)
#pragma unroll
        for (int i = 0; i < 4; ++i) P[c][i] = 0.0;

    int dd = 0;   // deliveries consumed so far
    {   // the row above the first product row (reflect-101 at the image top)
        bar_sync(BAR_AB_FULL + 0, 64);
        const double* ra = s_ab + 2;
        ld4(ra + 4 * lane, prevA);
        ld4(ra + AB_PITCH + 4 * lane, prevB);
        prevLA = ra[4 * lane - 1];
        prevLB = ra[AB_PITCH + 4 * lane - 1];
        if (dd + NB < g.n_deliv) bar_arrive(BAR_AB_EMPTY + 0, 64);
        dd = 1;
    }

    const int th = g.y1 - g.y0;
    const int n_virtual = th + 2 * R;
    int ib = 0;
    for (int n = 0; n < n_virtual; ++n, ib = (ib + 1 == C::WIN) ? 0 : ib + 1) {
        const int vy = g.y0 - R + n;
        double fx[4], fy[4], ft[4];
        if (vy >= 0 && vy < H) {
            const int slot = dd % NB;
            bar_sync(BAR_AB_FULL + slot, 64);
            const double* ra = s_ab + (size_t)slot * 2 * AB_PITCH + 2;
            double curA[4], curB[4];
            ld4(ra + 4 * lane, curA);
            ld4(ra + AB_PITCH + 4 * lane, curB);
            double curLA = ra[4 * lane - 1];
            double curLB = ra[AB_PITCH + 4 * lane - 1];
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
                const bool in = (unsigned)(xb + i) < (unsigned)W;
                fx[i] = in ? dadd(fxa, fxb) : 0.0;
                fy[i] = in ? dadd(fya, fyb) : 0.0;
                ft[i] = in ? dadd(fta, ftb) : 0.0;
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
        double S[5][4];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            double x[4], h[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                // fx2, fy2, fxfy, fxft, fyft (:383-405)
                const double p = c == 0 ? fx[i] : (c == 1 ? fy[i] : (c == 2 ? fx[i] : (c == 3 ? fx[i] : fy[i])));
                const double q = c == 0 ? fx[i] : (c == 1 ? fy[i] : (c == 2 ? fy[i] : ft[i]));
                x[i] = dmul(p, q);
            }
            hsum4<R>(x, h);
            if (active) {
                // vertical van Herk / Gil-Werman: P = running prefix of the current block,
                // s_v holds the suffix sums of the previous block, overwritten row by row.
                double* vrow = s_v + ((size_t)ib * 5 + c) * C::OUT_COLS + vcol;   // slot ib  <-> block row ib+1
                if (ib == 0) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) P[c][i] = h[i];
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) P[c][i] = dadd(P[c][i], h[i]);
                }
                if (emit) {
                    if (ib == C::WIN - 1) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) S[c][i] = P[c][i];
                    } else {
                        double sp[4];
                        ld4(vrow, sp);                                   // suffix of previous block from row ib+1
#pragma unroll
                        for (int i = 0; i < 4; ++i) S[c][i] = dadd(sp[i], P[c][i]);
                    }
                }
                if (ib >= 1) st4(vrow - (size_t)5 * C::OUT_COLS, h);     // raw row ib -> slot ib-1
            }
        }

        if (ib == C::WIN - 1 && active) {
            // block finished: turn the stored raw rows 1..2R into suffix sums, in place
#pragma unroll
            for (int c = 0; c < 5; ++c) {
                double acc[4];
                ld4(s_v + ((size_t)(C::VSLOTS - 1) * 5 + c) * C::OUT_COLS + vcol, acc);
#pragma unroll 2
                for (int t = C::VSLOTS - 2; t >= 0; --t) {
                    double* vr = s_v + ((size_t)t * 5 + c) * C::OUT_COLS + vcol;
                    double raw[4];
                    ld4(vr, raw);
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[i] = dadd(raw[i], acc[i]);
                    st4(vr, acc);
                }
            }
        }

        if (emit) {
            const int m = n - (C::WIN - 1);
            const int y = g.y0 + m;
            const int eslot = m % NE;
            if (m >= NE) bar_sync(BAR_LK_EMPTY + eslot, 64);
            if (active) {
                const bool yin = y >= R && y <= H - 1 - R;
                double lu[4], lv[4];
                unsigned okbits = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int x = xb + i;
                    const bool interior = yin && x >= R && x <= W - 1 - R;
                    const double sxx = interior ? S[0][i] : 0.0, syy = interior ? S[1][i] : 0.0;
                    const double sxy = interior ? S[2][i] : 0.0, sxt = interior ? S[3][i] : 0.0;
                    const double syt = interior ? S[4][i] : 0.0;
                    const double det = dsub(dmul(sxx, syy), dmul(sxy, sxy));
                    const double nu = dsub(dmul(sxy, syt), dmul(syy, sxt));
                    const double nv = dsub(dmul(sxt, sxy), dmul(sxx, syt));
                    const bool ok = fabs(det) > a.det_eps;
                    lu[i] = ok ? __ddiv_rn(nu, det) : 0.0;
                    lv[i] = ok ? __ddiv_rn(nv, det) : 0.0;
                    okbits |= ok ? (1u << (8 * i)) : 0u;
                }
                double* lk = s_lk + (size_t)eslot * 2 * C::OUT_COLS + vcol;
                st4(lk, lu);
                st4(lk + C::OUT_COLS, lv);
                *reinterpret_cast<unsigned*>(s_mk + (size_t)eslot * C::OUT_COLS + vcol) = okbits;
            }
            bar_arrive(BAR_LK_FULL + eslot, 64);
        }
    }
}

// ------------------------------------------------------------------------------ epilogue
template <int R>
__device__ __forceinline__ void epilogue_role(const LevelArgs& a, const Geo& g, int pair,
                                              const double* s_lk, const unsigned char* s_mk)
{
    using C = MarchCfg<R>;
    constexpr int NP = (C::OUT_COLS + 31) / 32;
    const int lane = threadIdx.x & 31;
    const double* __restrict__ cu = g.has_coarse ? a.cu + (size_t)pair * a.c_pair_stride : nullptr;
    const double* __restrict__ cv = g.has_coarse ? a.cv + (size_t)pair * a.c_pair_stride : nullptr;
    double* __restrict__ U = a.u ? a.u + (size_t)pair * a.uv_pair_stride : nullptr;
    double* __restrict__ V = a.v ? a.v + (size_t)pair * a.uv_pair_stride : nullptr;
    unsigned char* __restrict__ M = a.mask ? a.mask + (size_t)pair * a.mask_pair_stride : nullptr;
    UpCol up[NP];
    int kprev = -100;
    int dummy = 0;
    double st_x = 0.0, st_y = 0.0, st_c = 0.0;
    const int th = g.y1 - g.y0;
    for (int m = 0; m < th; ++m) {
        const int y = g.y0 + m;
        const int eslot = m % NE;
        const int k = y >> 1;
        // upsampled coarse flow first (global loads), then wait for the LK row
        double upu[NP], upv[NP];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            const int oc = 32 * p + lane;
            const int x = g.xo0 + oc;
            upu[p] = upv[p] = 0.0;
            if (g.has_coarse && oc < C::OUT_COLS && x < g.W) {
                if (k != kprev) upcol_advance(up[p], dummy, k, kprev, x, cu, cv, a.c_pitch, a.wc, a.hc);
                upcol_eval(up[p], y, upu[p], upv[p]);
            }
        }
        kprev = k;
        bar_sync(BAR_LK_FULL + eslot, 64);
        const double* lk = s_lk + (size_t)eslot * 2 * C::OUT_COLS;
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            const int oc = 32 * p + lane;
            const int x = g.xo0 + oc;
            if (oc < C::OUT_COLS && x < g.W) {
                double fu = lk[oc], fv = lk[C::OUT_COLS + oc];
                const unsigned char ok = s_mk[(size_t)eslot * C::OUT_COLS + oc];
                if (g.has_coarse) {                    // tmpu += upu; tmpv += upv  (:634-637)
                    fu = dadd(fu, upu[p]);
                    fv = dadd(fv, upv[p]);
                }
                if (U) {
                    U[(size_t)y * a.uv_pitch + x] = fu;
                    V[(size_t)y * a.uv_pitch + x] = fv;
                }
                if (M) M[(size_t)y * a.mask_pitch + x] = ok;
                if (a.stats) {
                    const double val = sqrt(dadd(dmul(fu, fu), dmul(fv, fv)));
                    if (!(val < a.mag_thresh)) { st_x += (double)x; st_y += (double)y; st_c += 1.0; }
                }
            }
        }
        if (m + NE < th) bar_arrive(BAR_LK_EMPTY + eslot, 64);
    }
    if (a.stats) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_x += __shfl_xor_sync(0xffffffffu, st_x, o);
            st_y += __shfl_xor_sync(0xffffffffu, st_y, o);
            st_c += __shfl_xor_sync(0xffffffffu, st_c, o);
        }
        if (lane == 0 && st_c != 0.0) {      // integer-valued sums: exact in any order
            atomicAdd(&a.stats[pair].sum_x, st_x);
            atomicAdd(&a.stats[pair].sum_y, st_y);
            atomicAdd(&a.stats[pair].count, st_c);
        }
    }
}

// ------------------------------------------------------------------------------ kernel
template <int R>
__global__ void __launch_bounds__(96) lk_level_march_kernel(LevelArgs a, int th)
{
    using C = MarchCfg<R>;
    extern __shared__ __align__(16) double sm[];
    double* s_v = sm;
    double* s_ab = s_v + C::V_DOUBLES;
    double* s_lk = s_ab + C::AB_DOUBLES;
    unsigned char* s_mk = reinterpret_cast<unsigned char*>(s_lk + C::LK_DOUBLES);

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
    const int pair = blockIdx.z;

    // padding in front of every AB row is read (never used) by lane 0: keep it defined
    if (threadIdx.x < NB * 2) {
        s_ab[(size_t)threadIdx.x * AB_PITCH] = 0.0;
        s_ab[(size_t)threadIdx.x * AB_PITCH + 1] = 0.0;
    }
    __syncthreads();

    const int role = threadIdx.x >> 5;
    if (role == 0) producer_role<R>(a, g, pair, s_ab);
    else if (role == 1) consumer_role<R>(a, g, s_ab, s_v, s_lk, s_mk);
    else epilogue_role<R>(a, g, pair, s_lk, s_mk);
}

template <int R>
cudaError_t launch_march(const LevelArgs& a, int n_pairs, cudaStream_t s, unsigned long long* launches)
{
    using C = MarchCfg<R>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(lk_level_march_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)C::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int strips = (a.w + C::OUT_COLS - 1) / C::OUT_COLS;
    // Chunk height: a multiple of the vertical block length (2R+1), so that the block
    // boundaries of the vertical sums sit at the same absolute rows whatever the chunking --
    // results are then bit-identical for any batch size.  Enough CTAs for a few waves of
    // 148 SMs x 3 CTAs, but tall enough that the 2R warm-up rows of a chunk stay a small
    // fraction of its work.
    int th = 8 * C::WIN;
    const long target = 148L * 3 * 4;
    while (th > 3 * C::WIN && (long)strips * ((a.h + th - 1) / th) * n_pairs < target) th -= C::WIN;
    if (th > a.h) th = a.h;
    const int chunks = (a.h + th - 1) / th;
    dim3 grid(strips, chunks, n_pairs);
    lk_level_march_kernel<R><<<grid, 96, C::SMEM_BYTES, s>>>(a, th);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_lk_level_march(const LevelArgs& a, int n_pairs, cudaStream_t s,
                                  unsigned long long* launches, bool* handled)
{
    *handled = true;
    switch (a.radius) {
        case 2: return launch_march<2>(a, n_pairs, s, launches);
        case 3: return launch_march<3>(a, n_pairs, s, launches);
        case 4: return launch_march<4>(a, n_pairs, s, launches);
        case 5: return launch_march<5>(a, n_pairs, s, launches);
        case 6: return launch_march<6>(a, n_pairs, s, launches);
        case 7: return launch_march<7>(a, n_pairs, s, launches);
        default: break;
    }
    *handled = false;
    return cudaSuccess;
}

}  // namespace ofb
