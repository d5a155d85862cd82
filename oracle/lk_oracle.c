/* CPU oracle (plain C + OpenMP) for the dense coarse-to-fine Lucas-Kanade path.
 *
 * TEST INFRASTRUCTURE ONLY -- never linked into, loaded by, or called from the
 * product (optical-flow_b200/).  Used by tests/ as the checker at sizes the numpy
 * oracle is too slow for, and by bench.py as the timed CPU baseline
 * (cpu_baseline.kind = "port": the reference itself needs OpenCV 2.4 C++ headers
 * and libraries, which are not installed, so it cannot be compiled here).
 *
 * Parity status: PINNED -- tests/test_oracle_golden.py checks this file against the
 * reference's own golden vectors (U2Test/V2Test, UPyrTest/VPyrTest, <= 1e-6 as the
 * reference's harness does, ParallelBasicPyr.cpp:787,818) and bit-for-bit against
 * oracle/lk_oracle.py, whose OpenCV restatements were pinned against cv2.
 *
 * Structure follows the reference stage by stage, one materialised temporary per
 * cv::Mat expression (file:line into /root/reference/ParallelVideoPyr.cpp; serial
 * twin ctfLKof.cpp:31-215).  OpenMP: row-parallel loops (the reference's
 * get_Sum9_Mat placement, :164-188) instead of the reference's "thread k runs Mat
 * op k" sections, which silently skip work when fewer threads are granted
 * (SURVEY.md 3.5 Q5).
 *
 * Build: gcc -O3 -fopenmp -ffp-contract=off -fPIC -shared  (see oracle/Makefile).
 * -ffp-contract=off: the reference materialises every product, so nothing here may
 * be fused into an FMA.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define LKO_OK 0
#define LKO_E_BADARG 1
#define LKO_E_NOT_DIVISIBLE 2
#define LKO_E_NOMEM 4

/* BORDER_REFLECT_101: -1 -> 1, n -> n-2 (n == 1 -> 0). */
static inline int r101(int i, int n)
{
    if (n == 1) return 0;
    int p = 2 * (n - 1);
    i %= p;
    if (i < 0) i += p;
    return i >= n ? p - i : i;
}

/* getMaxLayer, :545-564 */
int lko_get_max_layer(int w, int h)
{
    if (w <= 0 || h <= 0) return 0;
    int pw = 1, ph = 1;
    while (w % (1 << pw) == 0) ++pw;
    while (h % (1 << ph) == 0) ++ph;
    return pw < ph ? pw : ph;
}

/* get_fx / get_fy / get_ft, :44-84: six 2x2 filter2D (anchor (1,1), reflect-101,
 * accumulation ((k00 p00 + k01 p01) + k10 p10) + k11 p11) and three adds. */
static void gradients(const double *a, const double *b, int w, int h,
                      double *fx, double *fy, double *ft)
{
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; ++y) {
        int yu = r101(y - 1, h);
        for (int x = 0; x < w; ++x) {
            int xl = r101(x - 1, w);
            double a00 = a[(size_t)yu * w + xl], a01 = a[(size_t)yu * w + x];
            double a10 = a[(size_t)y * w + xl], a11 = a[(size_t)y * w + x];
            double b00 = b[(size_t)yu * w + xl], b01 = b[(size_t)yu * w + x];
            double b10 = b[(size_t)y * w + xl], b11 = b[(size_t)y * w + x];
            double fxa = ((-a00 + a01) + -a10) + a11;
            double fxb = ((-b00 + b01) + -b10) + b11;
            double fya = ((-a00 + -a01) + a10) + a11;
            double fyb = ((-b00 + -b01) + b10) + b11;
            double fta = ((-a00 + -a01) + -a10) + -a11;
            double ftb = ((b00 + b01) + b10) + b11;
            size_t i = (size_t)y * w + x;
            fx[i] = fxa + fxb;
            fy[i] = fya + fyb;
            ft[i] = fta + ftb;
        }
    }
}

/* Mat::mul, :383-405 */
static void mul(const double *p, const double *q, size_t n, double *o)
{
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) o[i] = p[i] * q[i];
}

/* get_Sum9_Mat / get_Sum9, :114-191, generalised to radius r: row-major direct sum
 * from 0.0 over the (2r+1)^2 window on the interior, ring of width r left at 0. */
static void box_sum(const double *m, int w, int h, int r, double *o)
{
    memset(o, 0, (size_t)w * h * sizeof(double));
#pragma omp parallel for schedule(static)
    for (int y = r; y < h - r; ++y)
        for (int x = r; x < w - r; ++x) {
            double s = 0.0;
            for (int i = -r; i <= r; ++i)
                for (int j = -r; j <= r; ++j)
                    s += m[(size_t)(y + i) * w + (x + j)];
            o[(size_t)y * w + x] = s;
        }
}

/* getLucasKanadeOpticalFlow, :327-542 */
int lko_lk_single(const double *a, const double *b, int w, int h, int r,
                  double det_eps, double *u, double *v, unsigned char *mask)
{
    if (!a || !b || !u || !v || w < 1 || h < 1 || r < 1) return LKO_E_BADARG;
    size_t n = (size_t)w * h;
    double *buf = (double *)malloc(13 * n * sizeof(double));
    if (!buf) return LKO_E_NOMEM;
    double *fx = buf, *fy = buf + n, *ft = buf + 2 * n;
    double *p[5], *s[5];
    for (int k = 0; k < 5; ++k) { p[k] = buf + (3 + k) * n; s[k] = buf + (8 + k) * n; }
    gradients(a, b, w, h, fx, fy, ft);
    mul(fx, fx, n, p[0]);           /* fx2  */
    mul(fy, fy, n, p[1]);           /* fy2  */
    mul(fx, fy, n, p[2]);           /* fxfy */
    mul(fx, ft, n, p[3]);           /* fxft */
    mul(fy, ft, n, p[4]);           /* fyft */
    for (int k = 0; k < 5; ++k) box_sum(p[k], w, h, r, s[k]);
    const double *sxx = s[0], *syy = s[1], *sxy = s[2], *sxt = s[3], *syt = s[4];
    /* :470-472 then cv::divide (:514,518; zero divisor -> 0 in OpenCV 2.4) */
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) {
        double t1 = sxx[i] * syy[i], t2 = sxy[i] * sxy[i];
        double det = t1 - t2;
        double n1 = sxy[i] * syt[i], n2 = syy[i] * sxt[i];
        double n3 = sxt[i] * sxy[i], n4 = sxx[i] * syt[i];
        double nu = n1 - n2, nv = n3 - n4;
        int ok = fabs(det) > det_eps;
        u[i] = ok ? nu / det : 0.0;
        v[i] = ok ? nv / det : 0.0;
        if (mask) mask[i] = (unsigned char)ok;
    }
    free(buf);
    return LKO_OK;
}

/* cv::pyrDown (:571): separable [1 4 6 4 1], decimate, reflect-101, * 1/256. */
void lko_pyr_down(const double *src, int w, int h, double *dst)
{
    int ow = (w + 1) / 2, oh = (h + 1) / 2;
    double *hor = (double *)malloc((size_t)ow * h * sizeof(double));
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; ++y) {
        const double *s = src + (size_t)y * w;
        for (int x = 0; x < ow; ++x) {
            int c = 2 * x;
            hor[(size_t)y * ow + x] = s[c] * 6.0 + (s[r101(c - 1, w)] + s[r101(c + 1, w)]) * 4.0
                                      + s[r101(c - 2, w)] + s[r101(c + 2, w)];
        }
    }
#pragma omp parallel for schedule(static)
    for (int y = 0; y < oh; ++y) {
        int c = 2 * y;
        const double *r0 = hor + (size_t)r101(c - 2, h) * ow, *r1 = hor + (size_t)r101(c - 1, h) * ow;
        const double *r2 = hor + (size_t)c * ow, *r3 = hor + (size_t)r101(c + 1, h) * ow;
        const double *r4 = hor + (size_t)r101(c + 2, h) * ow;
        for (int x = 0; x < ow; ++x)
            dst[(size_t)y * ow + x] = (r2[x] * 6.0 + (r1[x] + r3[x]) * 4.0 + r0[x] + r4[x]) * (1.0 / 256.0);
    }
    free(hor);
}

/* cv::pyrUp (:646-647): out 2w x 2h; see oracle/lk_oracle.py:pyr_up for the rule. */
void lko_pyr_up(const double *src, int w, int h, double *dst)
{
    int ow = 2 * w, oh = 2 * h;
    double *rows = (double *)malloc((size_t)ow * h * sizeof(double));
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; ++y) {
        const double *s = src + (size_t)y * w;
        double *r = rows + (size_t)y * ow;
        if (w == 1) { r[0] = r[1] = s[0] * 8.0; continue; }
        r[0] = s[0] * 6.0 + s[1] * 2.0;
        r[1] = (s[0] + s[1]) * 4.0;
        for (int x = 1; x < w - 1; ++x) {
            r[2 * x] = s[x - 1] + s[x] * 6.0 + s[x + 1];
            r[2 * x + 1] = (s[x] + s[x + 1]) * 4.0;
        }
        r[ow - 2] = s[w - 2] + s[w - 1] * 7.0;
        r[ow - 1] = s[w - 1] * 8.0;
    }
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; ++y) {
        const double *r0 = rows + (size_t)(r101(2 * (y - 1), oh) / 2) * ow;
        const double *r1 = rows + (size_t)y * ow;
        const double *r2 = rows + (size_t)(r101(2 * (y + 1), oh) / 2) * ow;
        double *d0 = dst + (size_t)(2 * y) * ow, *d1 = d0 + ow;
        for (int x = 0; x < ow; ++x) {
            d0[x] = (r0[x] + r1[x] * 6.0 + r2[x]) * (1.0 / 64.0);
            d1[x] = ((r1[x] + r2[x]) * 4.0) * (1.0 / 64.0);
        }
    }
    free(rows);
}

/* OpenCV's fixed-point map coordinate: s = rint_half_even(m * 32.f) in float;
 * integer part s >> 5 saturated to int16 (remap stores it as CV_16SC2). */
static inline void q5(float m, int *ip, int *frac)
{
    float t = m * 32.0f;
    if (!(t > -1073741824.0f)) t = (t != t) ? 1073741824.0f : -1073741824.0f;
    if (t > 1073741824.0f) t = 1073741824.0f;
    int s = (int)lrintf(t);               /* default rounding mode: half to even */
    int i = s >> 5;
    *ip = i < -32768 ? -32768 : (i > 32767 ? 32767 : i);
    *frac = s & 31;
}

/* Map build (:667-684, float32) + cv::remap INTER_LINEAR / BORDER_CONSTANT 0
 * (:691-703).  sign = +1 samples at (x + fu/2, y + fv/2) (frame 1), -1 frame 2. */
void lko_warp(const double *src, int w, int h, const double *upu, const double *upv,
              float sign, double *dst)
{
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            size_t i = (size_t)y * w + x;
            float fu = (float)upu[i], fv = (float)upv[i];
            float mx = sign > 0 ? (float)x + fu / 2.0f : (float)x - fu / 2.0f;
            float my = sign > 0 ? (float)y + fv / 2.0f : (float)y - fv / 2.0f;
            int ix, iy, ax, ay;
            q5(mx, &ix, &ax);
            q5(my, &iy, &ay);
            float w00 = (float)((32 - ay) * (32 - ax)) * (1.0f / 1024.0f);
            float w01 = (float)((32 - ay) * ax) * (1.0f / 1024.0f);
            float w10 = (float)(ay * (32 - ax)) * (1.0f / 1024.0f);
            float w11 = (float)(ay * ax) * (1.0f / 1024.0f);
            int x0 = ix >= 0 && ix < w, x1 = ix + 1 >= 0 && ix + 1 < w;
            int y0 = iy >= 0 && iy < h, y1 = iy + 1 >= 0 && iy + 1 < h;
            double s00 = (y0 && x0) ? src[(size_t)iy * w + ix] : 0.0;
            double s01 = (y0 && x1) ? src[(size_t)iy * w + ix + 1] : 0.0;
            double s10 = (y1 && x0) ? src[(size_t)(iy + 1) * w + ix] : 0.0;
            double s11 = (y1 && x1) ? src[(size_t)(iy + 1) * w + ix + 1] : 0.0;
            dst[i] = s00 * (double)w00 + s01 * (double)w01 + s10 * (double)w10 + s11 * (double)w11;
        }
}

/* coarseToFineEstimation, :578-725.  Inputs are not modified (Q1). */
int lko_ctf(const double *a, const double *b, int w, int h, int n_levels, int r,
            double det_eps, double *u, double *v, unsigned char *mask)
{
    if (!a || !b || !u || !v || w < 1 || h < 1 || n_levels < 1 || n_levels > 30 || r < 1)
        return LKO_E_BADARG;
    int f = 1 << (n_levels - 1);
    if (w % f || h % f) return LKO_E_NOT_DIVISIBLE;
    double *pa[32], *pb[32];
    int lw[32], lh[32];
    int rc = LKO_OK;
    memset(pa, 0, sizeof pa);
    memset(pb, 0, sizeof pb);
    for (int l = 0; l < n_levels; ++l) {
        lw[l] = w >> l; lh[l] = h >> l;
        size_t n = (size_t)lw[l] * lh[l];
        pa[l] = (double *)malloc(n * sizeof(double));
        pb[l] = (double *)malloc(n * sizeof(double));
        if (!pa[l] || !pb[l]) { rc = LKO_E_NOMEM; goto done; }
        if (l == 0) { memcpy(pa[0], a, n * sizeof(double)); memcpy(pb[0], b, n * sizeof(double)); }
        else { lko_pyr_down(pa[l - 1], lw[l - 1], lh[l - 1], pa[l]); lko_pyr_down(pb[l - 1], lw[l - 1], lh[l - 1], pb[l]); }
    }
    {
        size_t n0 = (size_t)w * h;
        double *tu = (double *)malloc(n0 * sizeof(double)), *tv = (double *)malloc(n0 * sizeof(double));
        double *upu = (double *)malloc(n0 * sizeof(double)), *upv = (double *)malloc(n0 * sizeof(double));
        double *wr = (double *)malloc(n0 * sizeof(double));
        if (!tu || !tv || !upu || !upv || !wr) rc = LKO_E_NOMEM;
        for (int l = n_levels - 1; l >= 0 && rc == LKO_OK; --l) {
            size_t n = (size_t)lw[l] * lh[l];
            rc = lko_lk_single(pa[l], pb[l], lw[l], lh[l], r, det_eps, tu, tv, l == 0 ? mask : NULL);
            if (rc) break;
            if (l != n_levels - 1) {                      /* :634-637 */
#pragma omp parallel for schedule(static)
                for (size_t i = 0; i < n; ++i) { tu[i] += upu[i]; tv[i] += upv[i]; }
            }
            if (l == 0) {                                 /* :638-645 */
                memcpy(u, tu, n * sizeof(double));
                memcpy(v, tv, n * sizeof(double));
                break;
            }
            lko_pyr_up(tu, lw[l], lh[l], upu);            /* :646-647 */
            lko_pyr_up(tv, lw[l], lh[l], upv);
            size_t nn = (size_t)lw[l - 1] * lh[l - 1];
            lko_warp(pa[l - 1], lw[l - 1], lh[l - 1], upu, upv, +1.0f, wr);   /* :695-696 */
            memcpy(pa[l - 1], wr, nn * sizeof(double));
            lko_warp(pb[l - 1], lw[l - 1], lh[l - 1], upu, upv, -1.0f, wr);   /* :700-701 */
            memcpy(pb[l - 1], wr, nn * sizeof(double));
        }
        free(tu); free(tv); free(upu); free(upv); free(wr);
    }
done:
    for (int l = 0; l < n_levels; ++l) { free(pa[l]); free(pb[l]); }
    return rc;
}

/* Gesture statistics of the live loop, :845-873. */
void lko_gesture_stats(const double *u, const double *v, int w, int h, double mag_thresh,
                       double *sum_x, double *sum_y, long long *count)
{
    double sx = 0.0, sy = 0.0;
    long long c = 0;
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            double xf = u[(size_t)i * w + j], yf = v[(size_t)i * w + j];
            double val = sqrt(xf * xf + yf * yf);
            if (!(val < mag_thresh)) { sx += j; sy += i; ++c; }
        }
    *sum_x = sx; *sum_y = sy; *count = c;
}
