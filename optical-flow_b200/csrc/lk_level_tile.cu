// Generic fused pyramid-level kernel ("tile" variant): any window radius.
//
// One launch does, for one pyramid level l of every pair in the batch, everything the
// reference does between two pyramid levels (ParallelVideoPyr.cpp:610-703) plus the
// single-level LK itself (:327-542):
//
//   pyrUp of the coarse flow (:646-647)  -> map build in float32 (:667-684)
//   -> remap of frame 1 at +f/2 and frame 2 at -f/2 (:691-703)       [skipped at the
//   -> fx/fy/ft 2x2 stencils on the warped frames (:44-84)             coarsest level]
//   -> five products (:383-405) -> (2r+1)^2 windowed sums (:114-191)
//   -> guarded 2x2 solve (:470-472, 510-520) -> += upsampled coarse flow (:634-637)
//
// Upsampled flows, maps, warped frames, gradients, products and sums live only in shared
// memory / registers; HBM sees: read A_l, B_l (gather), read u_{l+1}, v_{l+1}, write u_l, v_l.
//
// Summation order: r == 1 uses the reference's row-major 9-term order from 0.0
// (get_Sum9, :122-157) and is bit-exact with it; r > 1 (no reference order exists,
// SURVEY.md H7) sums rows first, then columns.
//
// This variant favours generality (runtime radius, levels smaller than a tile, 1-pixel
// wide images); the phase-synchronous kernel in lk_level_phase.cu is the throughput path
// for the radii it is instantiated for.
#include "of_device.cuh"
#include "of_kernels.h"

namespace ofb {

namespace {

constexpr int TL_TW = 32;
constexpr int TL_NT = 256;

struct TileGeom {
    int th;             // output tile height
    int rw, rh;         // warped region  (tw + 2r + 1) x (th + 2r + 1)
    int pw, ph;         // product region (tw + 2r) x (th + 2r)
    int cw, ch;         // coarse-flow tile
    __host__ __device__ TileGeom(int r, int th_) : th(th_)
    {
        rw = TL_TW + 2 * r + 1;
        rh = th + 2 * r + 1;
        pw = TL_TW + 2 * r;
        ph = th + 2 * r;
        cw = rw / 2 + 3;
        ch = rh / 2 + 3;
    }
    __host__ __device__ size_t smem_doubles() const
    {
        return (size_t)2 * rw * rh + (size_t)5 * pw * ph + (size_t)5 * TL_TW * ph + (size_t)2 * cw * ch;
    }
};

__global__ void __launch_bounds__(TL_NT) lk_level_tile_kernel(LevelArgs a, int th)
{
    extern __shared__ double sm[];
    const int R = a.radius;
    const TileGeom g(R, th);
    double* sA = sm;
    double* sB = sA + g.rw * g.rh;
    double* sP = sB + g.rw * g.rh;                 // 5 planes pw x ph
    double* sH = sP + 5 * g.pw * g.ph;             // 5 planes TL_TW x ph
    double* sCu = sH + 5 * TL_TW * g.ph;           // cw x ch
    double* sCv = sCu + g.cw * g.ch;
    __shared__ double s_red[3];

    const int pair = blockIdx.z;
    const double* __restrict__ A = a.A + (size_t)pair * a.a_pair_stride;
    const double* __restrict__ B = a.B + (size_t)pair * a.b_pair_stride;
    const bool has_coarse = a.cu != nullptr;
    const int W = a.w, H = a.h;
    const int x0 = blockIdx.x * TL_TW, y0 = blockIdx.y * th;
    const int ox = x0 - R - 1, oy = y0 - R - 1;     // image coords of warped-region (0,0)
    const int tid = threadIdx.x;

    // ---- phase 0: coarse flow tile (1-pixel apron for the 3-tap pyrUp) -----------------
    int cx0 = 0, cy0 = 0;
    if (has_coarse) {
        cx0 = max(0, (max(ox, 0) >> 1) - 1);
        cy0 = max(0, (max(oy, 0) >> 1) - 1);
        const double* __restrict__ cu = a.cu + (size_t)pair * a.c_pair_stride;
        const double* __restrict__ cv = a.cv + (size_t)pair * a.c_pair_stride;
        for (int i = tid; i < g.cw * g.ch; i += TL_NT) {
            const int ry = i / g.cw, rx = i - ry * g.cw;
            const int cy = cy0 + ry, cx = cx0 + rx;
            double fu = 0.0, fv = 0.0;
            if (cy < a.hc && cx < a.wc) {
                fu = __ldg(cu + (size_t)cy * a.c_pitch + cx);
                fv = __ldg(cv + (size_t)cy * a.c_pitch + cx);
            }
            sCu[i] = fu;
            sCv[i] = fv;
        }
    }
    if (tid < 3) s_red[tid] = 0.0;
    __syncthreads();

    // ---- phase 1: warped frames over the region ----------------------------------------
    for (int i = tid; i < g.rw * g.rh; i += TL_NT) {
        const int ry = i / g.rw, rx = i - ry * g.rw;
        const int y = oy + ry, x = ox + rx;
        double va = 0.0, vb = 0.0;
        if (x >= 0 && x < W && y >= 0 && y < H) {
            if (has_coarse) {
                const double upu = pyrup_at(sCu, g.cw, cy0, cx0, y, x, a.wc, a.hc);
                const double upv = pyrup_at(sCv, g.cw, cy0, cx0, y, x, a.wc, a.hc);
                float m1x, m1y, m2x, m2y;
                map_coords(x, y, upu, upv, m1x, m1y, m2x, m2y);
                va = remap_tap4(A, a.img_pitch, W, H, m1x, m1y);
                vb = remap_tap4(B, a.img_pitch, W, H, m2x, m2y);
            } else {
                va = __ldg(A + (size_t)y * a.img_pitch + x);
                vb = __ldg(B + (size_t)y * a.img_pitch + x);
            }
        }
        sA[i] = va;
        sB[i] = vb;
    }
    __syncthreads();

    // ---- phase 2: gradients and the five products over the product region ---------------
    const int pplane = g.pw * g.ph;
    for (int i = tid; i < pplane; i += TL_NT) {
        const int py = i / g.pw, px = i - py * g.pw;
        const int y = y0 - R + py, x = x0 - R + px;
        double fx2 = 0.0, fy2 = 0.0, fxfy = 0.0, fxft = 0.0, fyft = 0.0;
        if (x >= 0 && x < W && y >= 0 && y < H) {
            const int yu = (y == 0) ? (H > 1 ? 1 : 0) : y - 1;     // reflect-101
            const int xl = (x == 0) ? (W > 1 ? 1 : 0) : x - 1;
            const int r_u = (yu - oy) * g.rw, r_c = (y - oy) * g.rw;
            const int c_l = xl - ox, c_c = x - ox;
            const double a00 = sA[r_u + c_l], a01 = sA[r_u + c_c], a10 = sA[r_c + c_l], a11 = sA[r_c + c_c];
            const double b00 = sB[r_u + c_l], b01 = sB[r_u + c_c], b10 = sB[r_c + c_l], b11 = sB[r_c + c_c];
            // filter2D accumulation order ((k00 p00 + k01 p01) + k10 p10) + k11 p11, k = +-1
            const double fxa = dadd(dsub(dsub(a01, a00), a10), a11);
            const double fxb = dadd(dsub(dsub(b01, b00), b10), b11);
            const double sa = dadd(a00, a01), sb = dadd(b00, b01);
            const double fya = dadd(dsub(a10, sa), a11);
            const double fyb = dadd(dsub(b10, sb), b11);
            const double fta = -dadd(dadd(sa, a10), a11);
            const double ftb = dadd(dadd(sb, b10), b11);
            const double fx = dadd(fxa, fxb), fy = dadd(fya, fyb), ft = dadd(fta, ftb);
            fx2 = dmul(fx, fx);
            fy2 = dmul(fy, fy);
            fxfy = dmul(fx, fy);
            fxft = dmul(fx, ft);
            fyft = dmul(fy, ft);
        }
        sP[i] = fx2;
        sP[pplane + i] = fy2;
        sP[2 * pplane + i] = fxfy;
        sP[3 * pplane + i] = fxft;
        sP[4 * pplane + i] = fyft;
    }
    __syncthreads();

    // ---- phase 3 (r > 1): horizontal window sums ------------------------------------------
    const int hplane = TL_TW * g.ph;
    if (R > 1) {
        for (int i = tid; i < hplane; i += TL_NT) {
            const int py = i / TL_TW, tx = i - py * TL_TW;
            const double* p = sP + py * g.pw + tx;
#pragma unroll 1
            for (int c = 0; c < 5; ++c) {
                const double* q = p + c * pplane;
                double s = q[0];
                for (int j = 1; j <= 2 * R; ++j) s = dadd(s, q[j]);
                sH[c * hplane + i] = s;
            }
        }
        __syncthreads();
    }

    // ---- phase 4: vertical sums, solve, accumulate, store ---------------------------------
    double st_x = 0.0, st_y = 0.0, st_c = 0.0;
    for (int i = tid; i < TL_TW * th; i += TL_NT) {
        const int ty = i / TL_TW, tx = i - ty * TL_TW;
        const int y = y0 + ty, x = x0 + tx;
        if (x >= W || y >= H) continue;
        double S[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
        if (x >= R && x <= W - 1 - R && y >= R && y <= H - 1 - R) {
            if (R == 1) {
#pragma unroll
                for (int c = 0; c < 5; ++c) {
                    const double* q = sP + c * pplane + ty * g.pw + tx;
                    double s = 0.0;
#pragma unroll
                    for (int di = 0; di < 3; ++di)
#pragma unroll
                        for (int dj = 0; dj < 3; ++dj) s = dadd(s, q[di * g.pw + dj]);
                    S[c] = s;
                }
            } else {
#pragma unroll
                for (int c = 0; c < 5; ++c) {
                    const double* q = sH + c * hplane + ty * TL_TW + tx;
                    double s = q[0];
                    for (int di = 1; di <= 2 * R; ++di) s = dadd(s, q[di * TL_TW]);
                    S[c] = s;
                }
            }
        }
        const double sxx = S[0], syy = S[1], sxy = S[2], sxt = S[3], syt = S[4];
        const double det = dsub(dmul(sxx, syy), dmul(sxy, sxy));
        const double nu = dsub(dmul(sxy, syt), dmul(syy, sxt));
        const double nv = dsub(dmul(sxt, sxy), dmul(sxx, syt));
        const bool ok = fabs(det) > a.det_eps;
        double fu = ok ? __ddiv_rn(nu, det) : 0.0;
        double fv = ok ? __ddiv_rn(nv, det) : 0.0;
        if (has_coarse) {
            fu = dadd(fu, pyrup_at(sCu, g.cw, cy0, cx0, y, x, a.wc, a.hc));
            fv = dadd(fv, pyrup_at(sCv, g.cw, cy0, cx0, y, x, a.wc, a.hc));
        }
        if (a.u) {
            a.u[(size_t)pair * a.uv_pair_stride + (size_t)y * a.uv_pitch + x] = fu;
            a.v[(size_t)pair * a.uv_pair_stride + (size_t)y * a.uv_pitch + x] = fv;
        }
        if (a.mask) a.mask[(size_t)pair * a.mask_pair_stride + (size_t)y * a.mask_pitch + x] = ok ? 1 : 0;
        if (a.stats) {
            const double val = sqrt(dadd(dmul(fu, fu), dmul(fv, fv)));
            if (!(val < a.mag_thresh)) { st_x += (double)x; st_y += (double)y; st_c += 1.0; }
        }
    }
    if (a.stats) {
        // integer-valued sums: exact in any order below 2^53, so atomics keep parity
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_x += __shfl_xor_sync(0xffffffffu, st_x, o);
            st_y += __shfl_xor_sync(0xffffffffu, st_y, o);
            st_c += __shfl_xor_sync(0xffffffffu, st_c, o);
        }
        if ((tid & 31) == 0 && st_c != 0.0) {
            atomicAdd(&s_red[0], st_x);
            atomicAdd(&s_red[1], st_y);
            atomicAdd(&s_red[2], st_c);
        }
        __syncthreads();
        if (tid == 0 && s_red[2] != 0.0) {
            atomicAdd(&a.stats[pair].sum_x, s_red[0]);
            atomicAdd(&a.stats[pair].sum_y, s_red[1]);
            atomicAdd(&a.stats[pair].count, s_red[2]);
        }
    }
}

}  // namespace

cudaError_t launch_lk_level_tile(const LevelArgs& a, int n_pairs, cudaStream_t s,
                                 unsigned long long* launches)
{
    int max_smem = 0, dev = 0;                      // per device: queried on every launch (this is the rare path)
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    int th = 32;
    while (th > 1 && TileGeom(a.radius, th).smem_doubles() * sizeof(double) + 1024 > (size_t)max_smem) th >>= 1;
    const size_t smem = TileGeom(a.radius, th).smem_doubles() * sizeof(double);
    if (smem + 1024 > (size_t)max_smem) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(lk_level_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    dim3 grid((a.w + TL_TW - 1) / TL_TW, (a.h + th - 1) / th, n_pairs);
    lk_level_tile_kernel<<<grid, TL_NT, smem, s>>>(a, th);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace ofb
