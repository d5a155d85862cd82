// Phase-synchronous fused pyramid-level kernel (throughput path), window radius R = 2..7.
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
//   P0  horizontal pass of cv::pyrUp for the coarse-flow rows this step touches  -> HR
//   P1  per pixel: vertical pass of pyrUp, float32 maps, 8 gathers, two blends   -> AB
//       (the upsampled flow of output pixels is parked in u, v for the epilogue)
//   P2  per pixel: 2x2 gradient stencils of both frames, five products           -> PROD
//   P3  per (column, channel): vertical window sums, van Herk / Gil-Werman with the
//       suffix sums of the previous block held in registers                      -> PROD
//   P4a per (row, channel) warp line: horizontal window sums, registers+shuffles -> PROD
//   P4b per output pixel: guarded solve, += parked up-flow, stores, mask, statistics
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

namespace ofb {

namespace {

constexpr int PH_NT = 640;       // threads per CTA: 128 columns x 5 channels in P3
constexpr int PH_PC = 128;       // product columns per strip
constexpr int PH_PITCH = 136;    // doubles per PROD row: lo pairs [0,64), hi pairs [72,136)

template <int R>
struct PhaseCfg {
    static constexpr int Q = R / 4;
    static constexpr int HL = Q + 1;                  // halo lanes (4 columns each) per side
    static constexpr int OUT_LANES = 32 - 2 * HL;
    static constexpr int OUT_COLS = 4 * OUT_LANES;
    static constexpr int WIN = 2 * R + 1;             // rows per step = vertical block length
    static constexpr int AB_ROWS = WIN + 1;           // + the row above the step
    static constexpr int HR_ROWS = (WIN + 1) / 2 + 3; // coarse rows a step can touch
    static constexpr int PROD_D = WIN * 5 * PH_PITCH;
    static constexpr int AB_D = AB_ROWS * 2 * PH_PC;
    static constexpr int HR_D = HR_ROWS * 2 * PH_PC;
    static constexpr size_t SMEM_BYTES = (size_t)(PROD_D + AB_D + HR_D) * 8;
};

// Position of strip column X (0..127) inside a PROD row.  Columns 4j, 4j+1 of lane j sit at
// lo[2j], columns 4j+2, 4j+3 at hi[2j]; hi is skewed by 8 doubles so that both the
// column-per-thread accesses (P2, P3, P4b) and the 4-columns-per-lane 128-bit accesses (P4a)
// are bank-conflict-free.
__device__ __forceinline__ int ppos(int X) { return ((X & 2) ? 72 : 0) + 2 * (X >> 2) + (X & 1); }

struct PGeo {
    int W, H;
    int c0;           // image column of strip column 0 (multiple of 4, may be negative)
    int xo0;          // first output column
    int y0, y1;       // output rows [y0, y1)
    bool has_coarse;
};

// ------------------------------------------------------------------------------ P0
// Horizontal pass of cv::pyrUp (:646-647) of coarse rows rho(ka) .. rho(ka + nk - 1), both
// flow planes, for the strip's 128 fine columns.  Columns outside the image hold 0.
__device__ __forceinline__ void phase_up_rows(const LevelArgs& a, const PGeo& g, const double* __restrict__ cu,
                                              const double* __restrict__ cv, int ka, int nk, double* s_hr)
{
    const int wc = a.wc, hc = a.hc;
    for (int t = threadIdx.x; t < nk * 2 * PH_PC; t += PH_NT) {
        const int X = t & (PH_PC - 1);
        const int p = (t >> 7) & 1;
        const int r = t >> 8;
        const int gx = g.c0 + X;
        double val = 0.0;
        if ((unsigned)gx < (unsigned)g.W) {
            const double* row = (p ? cv : cu) + (size_t)pyrup_rho(ka + r, hc) * a.c_pitch;
            val = pyrup_row(row, 0, gx, wc);
        }
        s_hr[t] = val;
    }
}

// ------------------------------------------------------------------------------ P1
// Warped frames A', B' (:667-700) of fine rows vy0 + r - 1, r = r_lo .. WIN, into the AB block
// (slot of r = (sb + r) mod AB_ROWS).  Row -1 is the reflect-101 row above row 0 (image row 1),
// other rows outside the image are never read and are skipped.
template <int R>
__device__ __forceinline__ void phase_remap(const LevelArgs& a, const PGeo& g, const double* __restrict__ A,
                                            const double* __restrict__ B, double* __restrict__ U,
                                            double* __restrict__ V, int vy0, int r_lo, int sb, int ka,
                                            const double* s_hr, double* s_ab)
{
    using C = PhaseCfg<R>;
    const int W = g.W, H = g.H;
    const int pitch = (int)a.img_pitch;
    const int ntask = (C::WIN + 1 - r_lo) * PH_PC;
    for (int t = threadIdx.x; t < ntask; t += PH_NT) {
        const int X = t & (PH_PC - 1);
        const int r = (t >> 7) + r_lo;
        const int vy = vy0 + r - 1;
        int yy = vy;
        if (vy == -1) yy = H > 1 ? 1 : 0;
        else if ((unsigned)vy >= (unsigned)H) continue;
        int slot = sb + r;
        if (slot >= C::AB_ROWS) slot -= C::AB_ROWS;
        double* dst = s_ab + slot * (2 * PH_PC) + X;
        const int gx = g.c0 + X;
        double va = 0.0, vb = 0.0;
        if ((unsigned)gx < (unsigned)W) {
            if (g.has_coarse) {
                const int k = yy >> 1;
                const double* hu = s_hr + ((k - ka) * 2) * PH_PC + X;      // row k, plane u
                const double* hv = hu + PH_PC;
                double upu, upv;
                if (yy & 1) {
                    upu = dmul(dmul(dadd(hu[0], hu[2 * PH_PC]), 4.0), 1.0 / 64.0);
                    upv = dmul(dmul(dadd(hv[0], hv[2 * PH_PC]), 4.0), 1.0 / 64.0);
                } else {
                    upu = dmul(dadd(dadd(hu[-2 * PH_PC], dmul(hu[0], 6.0)), hu[2 * PH_PC]), 1.0 / 64.0);
                    upv = dmul(dadd(dadd(hv[-2 * PH_PC], dmul(hv[0], 6.0)), hv[2 * PH_PC]), 1.0 / 64.0);
                }
                // park the upsampled flow of output pixels for the epilogue (tmpu += upu, :634-637)
                if (vy == yy && vy >= g.y0 && vy < g.y1 && gx >= g.xo0 && gx < g.xo0 + C::OUT_COLS) {
                    U[(size_t)vy * a.uv_pitch + gx] = upu;
                    V[(size_t)vy * a.uv_pitch + gx] = upv;
                }
                float m1x, m1y, m2x, m2y;
                map_coords(gx, yy, upu, upv, m1x, m1y, m2x, m2y);
                const bool finite = fabsf(m1x) + fabsf(m1y) + fabsf(m2x) + fabsf(m2y) < __int_as_float(0x7f800000);
                const Tap t1 = make_tap(m1x, m1y, W, H, pitch, finite);
                const Tap t2 = make_tap(m2x, m2y, W, H, pitch, finite);
                const double* pa = A + t1.off;
                const double* pb = B + t2.off;
                const double a00 = __ldg(pa), a01 = __ldg(pa + 1), a10 = __ldg(pa + pitch), a11 = __ldg(pa + pitch + 1);
                const double b00 = __ldg(pb), b01 = __ldg(pb + 1), b10 = __ldg(pb + pitch), b11 = __ldg(pb + pitch + 1);
                va = blend4(a00, a01, a10, a11, t1.fr);
                vb = blend4(b00, b01, b10, b11, t2.fr);
                if (!(t1.fr & t2.fr & 0x10000)) {        // rare: a tap outside the image or a non-finite map
                    if (!(t1.fr & 0x10000)) va = remap_tap4(A, a.img_pitch, W, H, m1x, m1y);
                    if (!(t2.fr & 0x10000)) vb = remap_tap4(B, a.img_pitch, W, H, m2x, m2y);
                }
            } else {
                va = __ldg(A + (size_t)yy * a.img_pitch + gx);
                vb = __ldg(B + (size_t)yy * a.img_pitch + gx);
            }
        }
        dst[0] = va;
        dst[PH_PC] = vb;
    }
}

// ------------------------------------------------------------------------------ P2
// Gradients (2x2 filter2D stencils over both frames, accumulation order
// ((k00 p00 + k01 p01) + k10 p10) + k11 p11, :44-84) and the five products (:383-405) of
// product rows vy0 + i, i = 0 .. WIN-1.  Rows and columns outside the image contribute 0.
template <int R>
__device__ __forceinline__ void phase_products(const PGeo& g, int vy0, int sb, const double* s_ab, double* s_prod)
{
    using C = PhaseCfg<R>;
    const int W = g.W, H = g.H;
    for (int t = threadIdx.x; t < C::WIN * PH_PC; t += PH_NT) {
        const int X = t & (PH_PC - 1);
        const int i = t >> 7;
        const int vy = vy0 + i;
        const int gx = g.c0 + X;
        double pxx = 0.0, pyy = 0.0, pxy = 0.0, pxt = 0.0, pyt = 0.0;
        if ((unsigned)vy < (unsigned)H && (unsigned)gx < (unsigned)W) {
            int s1 = sb + i + 1, s0 = sb + i;
            if (s1 >= C::AB_ROWS) s1 -= C::AB_ROWS;
            if (s0 >= C::AB_ROWS) s0 -= C::AB_ROWS;
            // left neighbour: reflect-101 at the image border; strip column 0 has none (its
            // products are never used: it only exists as the neighbour of column 1)
            const int XL = gx == 0 ? (W > 1 ? X + 1 : X) : (X == 0 ? X : X - 1);
            const double* up = s_ab + s0 * (2 * PH_PC);
            const double* cr = s_ab + s1 * (2 * PH_PC);
            const double a00 = up[XL], a01 = up[X], a10 = cr[XL], a11 = cr[X];
            const double b00 = up[PH_PC + XL], b01 = up[PH_PC + X], b10 = cr[PH_PC + XL], b11 = cr[PH_PC + X];
            const double fxa = dadd(dsub(dsub(a01, a00), a10), a11);
            const double fxb = dadd(dsub(dsub(b01, b00), b10), b11);
            const double sa = dadd(a00, a01), sbb = dadd(b00, b01);
            const double fya = dadd(dsub(a10, sa), a11);
            const double fyb = dadd(dsub(b10, sbb), b11);
            const double fta = -dadd(dadd(sa, a10), a11);
            const double ftb = dadd(dadd(sbb, b10), b11);
            const double fx = dadd(fxa, fxb), fy = dadd(fya, fyb), ft = dadd(fta, ftb);
            pxx = dmul(fx, fx);
            pyy = dmul(fy, fy);
            pxy = dmul(fx, fy);
            pxt = dmul(fx, ft);
            pyt = dmul(fy, ft);
        }
        double* d = s_prod + i * (5 * PH_PITCH) + ppos(X);
        d[0] = pxx;
        d[PH_PITCH] = pyy;
        d[2 * PH_PITCH] = pxy;
        d[3 * PH_PITCH] = pxt;
        d[4 * PH_PITCH] = pyt;
    }
}

// ------------------------------------------------------------------------------ P3
// Vertical (2R+1)-row window sums of one (column, channel), in place.  van Herk / Gil-Werman:
// the window of block row i is the suffix of the previous block from row i+1 plus the prefix
// of this block up to row i.  ring[t] holds that suffix for row t+1 on entry and is refilled
// with this block's raw rows, turned into suffix sums at the end -- all in registers.
template <int R>
__device__ __forceinline__ void phase_vertical(double* col, double (&ring)[2 * R])
{
    constexpr int WIN = 2 * R + 1;
    double P = 0.0;
#pragma unroll
    for (int i = 0; i < WIN; ++i) {
        const double p = col[i * (5 * PH_PITCH)];
        P = i == 0 ? p : dadd(P, p);
        double S = P;
        if (i < WIN - 1) S = dadd(ring[i], P);
        col[i * (5 * PH_PITCH)] = S;
        if (i >= 1) ring[i - 1] = p;
    }
#pragma unroll
    for (int t = 2 * R - 2; t >= 0; --t) ring[t] = dadd(ring[t], ring[t + 1]);
}

// ------------------------------------------------------------------------------ P4a
// Horizontal (2R+1)-column window sums of the lines (row i, channel c) whose output row exists,
// in place; one warp per line, lane = 4 consecutive columns.
template <int R>
__device__ __forceinline__ void phase_horizontal(double* s_prod, int i_lo, int i_hi)
{
    using C = PhaseCfg<R>;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const bool active = lane >= C::HL && lane < 32 - C::HL;
    for (int L = i_lo * 5 + warp; L < i_hi * 5; L += PH_NT / 32) {
        double* row = s_prod + L * PH_PITCH + 2 * lane;
        const double2 lo = *reinterpret_cast<const double2*>(row);
        const double2 hi = *reinterpret_cast<const double2*>(row + 72);
        const double x[4] = {lo.x, lo.y, hi.x, hi.y};
        double h[4];
        hsum4<R>(x, h);
        if (active) {
            *reinterpret_cast<double2*>(row) = make_double2(h[0], h[1]);
            *reinterpret_cast<double2*>(row + 72) = make_double2(h[2], h[3]);
        }
    }
}

// ------------------------------------------------------------------------------ P4b
// Guarded 2x2 solve (:407-542), += upsampled coarse flow (:634-637), stores, mask, gesture
// statistics (:845-873) of output rows y0 + m, m = m0 + i, i = i_lo .. i_hi-1.
struct StatAcc { unsigned long long x, y, c; };

template <int R>
__device__ __forceinline__ void phase_solve(const LevelArgs& a, const PGeo& g, double* __restrict__ U,
                                            double* __restrict__ V, unsigned char* __restrict__ M, int m0,
                                            int i_lo, int i_hi, const double* s_prod, StatAcc& st)
{
    using C = PhaseCfg<R>;
    const int W = g.W, H = g.H;
    const double det_eps = a.det_eps, mag2 = a.mag2_thresh;
    for (int t = threadIdx.x + i_lo * C::OUT_COLS; t < i_hi * C::OUT_COLS; t += PH_NT) {
        const int i = t / C::OUT_COLS;
        const int oc = t - i * C::OUT_COLS;
        const int gx = g.xo0 + oc;
        if (gx >= W) continue;
        const int y = g.y0 + m0 + i;
        const double* sp = s_prod + i * (5 * PH_PITCH) + ppos(4 * C::HL + oc);
        const double sxx = sp[0], syy = sp[PH_PITCH], sxy = sp[2 * PH_PITCH];
        const double sxt = sp[3 * PH_PITCH], syt = sp[4 * PH_PITCH];
        const size_t o = (size_t)y * a.uv_pitch + gx;
        double upu = 0.0, upv = 0.0;
        if (g.has_coarse) { upu = U[o]; upv = V[o]; }
        const bool interior = y >= R && y <= H - 1 - R && gx >= R && gx <= W - 1 - R;
        const double det = dsub(dmul(sxx, syy), dmul(sxy, sxy));
        const double nu = dsub(dmul(sxy, syt), dmul(syy, sxt));
        const double nv = dsub(dmul(sxt, sxy), dmul(sxx, syt));
        const bool ok = interior && fabs(det) > det_eps;
        double lu, lv;
        if (!div2_fast(nu, nv, det, lu, lv) && ok) {     // rare: exponent range outside the fast path
            lu = ddiv_slow(nu, det);
            lv = ddiv_slow(nv, det);
        }
        lu = ok ? lu : 0.0;
        lv = ok ? lv : 0.0;
        const double fu = g.has_coarse ? dadd(lu, upu) : lu;
        const double fv = g.has_coarse ? dadd(lv, upv) : lv;
        U[o] = fu;
        V[o] = fv;
        if (M) M[(size_t)y * a.mask_pitch + gx] = ok ? 1 : 0;
        if (a.stats) {
            // val = sqrt(u^2 + v^2); !(val < mag_thresh)  <=>  !(u^2 + v^2 < mag2)
            const double s2 = dadd(dmul(fu, fu), dmul(fv, fv));
            if (!(s2 < mag2)) { st.x += (unsigned)gx; st.y += (unsigned)y; st.c += 1; }
        }
    }
}

// ------------------------------------------------------------------------------ kernel
template <int R>
__global__ void __launch_bounds__(PH_NT, 1) lk_level_phase_kernel(LevelArgs a, int th)
{
    using C = PhaseCfg<R>;
    extern __shared__ __align__(16) double sm[];
    double* s_prod = sm;
    double* s_ab = s_prod + C::PROD_D;
    double* s_hr = s_ab + C::AB_D;

    PGeo g;
    g.W = a.w;
    g.H = a.h;
    g.xo0 = blockIdx.x * C::OUT_COLS;
    g.c0 = g.xo0 - 4 * C::HL;
    g.y0 = blockIdx.y * th;
    g.y1 = min(g.y0 + th, a.h);
    g.has_coarse = a.cu != nullptr;
    const int pair = blockIdx.z;
    const int H = g.H;

    const double* __restrict__ A = a.A + (size_t)pair * a.a_pair_stride;
    const double* __restrict__ B = a.B + (size_t)pair * a.b_pair_stride;
    const double* __restrict__ cu = g.has_coarse ? a.cu + (size_t)pair * a.c_pair_stride : nullptr;
    const double* __restrict__ cv = g.has_coarse ? a.cv + (size_t)pair * a.c_pair_stride : nullptr;
    double* U = a.u + (size_t)pair * a.uv_pair_stride;
    double* V = a.v + (size_t)pair * a.uv_pair_stride;
    unsigned char* M = a.mask ? a.mask + (size_t)pair * a.mask_pair_stride : nullptr;

    // P3 ownership: thread -> (channel, strip column)
    double* vcol = s_prod + (threadIdx.x >> 7) * PH_PITCH + ppos(threadIdx.x & (PH_PC - 1));
    double ring[2 * R];
#pragma unroll
    for (int t = 0; t < 2 * R; ++t) ring[t] = 0.0;
    StatAcc st = {0, 0, 0};

    const int rows = g.y1 - g.y0;
    const int n_steps = (rows + 2 * R + C::WIN - 1) / C::WIN;
    int sb = 0;                                        // (s * WIN) mod AB_ROWS
    for (int s = 0; s < n_steps; ++s) {
        const int vy0 = g.y0 - R + s * C::WIN;         // product row of block row 0
        const int r_lo = s == 0 ? 0 : 1;
        // fine rows this step warps: [vlo, vhi]; source rows after border handling: [lo, hi]
        const int vlo = vy0 + r_lo - 1, vhi = vy0 + C::WIN - 1;
        int lo = max(vlo, 0), hi = min(vhi, H - 1);
        if (vlo <= -1 && vhi >= -1) { lo = 0; hi = max(hi, min(1, H - 1)); }
        const int ka = (lo >> 1) - 1;
        if (lo <= hi) {
            if (g.has_coarse) {
                phase_up_rows(a, g, cu, cv, ka, (hi >> 1) + 1 - ka + 1, s_hr);
                __syncthreads();
            }
            phase_remap<R>(a, g, A, B, U, V, vy0, r_lo, sb, ka, s_hr, s_ab);
        }
        __syncthreads();
        phase_products<R>(g, vy0, sb, s_ab, s_prod);
        __syncthreads();
        phase_vertical<R>(vcol, ring);
        __syncthreads();
        const int m0 = C::WIN * (s - 1) + 1;           // output row (relative to y0) of block row 0
        const int i_lo = max(0, -m0), i_hi = min(C::WIN, rows - m0);
        phase_horizontal<R>(s_prod, i_lo, i_hi);
        __syncthreads();
        phase_solve<R>(a, g, U, V, M, m0, i_lo, i_hi, s_prod, st);
        __syncthreads();
        sb += C::WIN;
        if (sb >= C::AB_ROWS) sb -= C::AB_ROWS;
    }

    if (a.stats) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st.x += __shfl_xor_sync(0xffffffffu, st.x, o);
            st.y += __shfl_xor_sync(0xffffffffu, st.y, o);
            st.c += __shfl_xor_sync(0xffffffffu, st.c, o);
        }
        if ((threadIdx.x & 31) == 0 && st.c != 0) {    // integer-valued sums: exact in any order
            atomicAdd(&a.stats[pair].sum_x, (double)st.x);
            atomicAdd(&a.stats[pair].sum_y, (double)st.y);
            atomicAdd(&a.stats[pair].count, (double)st.c);
        }
    }
}

template <int R>
cudaError_t launch_phase(const LevelArgs& a, int n_pairs, cudaStream_t s, unsigned long long* launches)
{
    using C = PhaseCfg<R>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(lk_level_phase_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)C::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int strips = (a.w + C::OUT_COLS - 1) / C::OUT_COLS;
    // Chunk height: a multiple of the vertical block length, so that block boundaries sit at
    // the same absolute rows whatever the chunking (bit-identical results for any batch size).
    // Pick the height that minimises (waves of resident CTAs) x (steps a CTA runs).
    const long resident = 148;
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
    lk_level_phase_kernel<R><<<grid, PH_NT, C::SMEM_BYTES, s>>>(a, th);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace

// Handles R = 2..7 on even-width levels with a flow output; everything else (R = 1 bit-exact
// order, tiny levels, statistics-only calls) goes to the other kernels.
cudaError_t launch_lk_level_phase(const LevelArgs& a, int n_pairs, cudaStream_t s,
                                  unsigned long long* launches, bool* handled)
{
    *handled = false;
    if (!a.u || !a.v) return cudaSuccess;
    if (a.w < 4 || a.h < 2 || (a.w & 1)) return cudaSuccess;
    if (a.cu && (a.wc < 2 || a.w != 2 * a.wc || a.h != 2 * a.hc)) return cudaSuccess;
    if ((unsigned long long)a.img_pitch * (unsigned long long)a.h >= (1ull << 31)) return cudaSuccess;
    *handled = true;
    switch (a.radius) {
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
