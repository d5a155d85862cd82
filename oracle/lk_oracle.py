"""CPU oracle (numpy) for the dense coarse-to-fine Lucas-Kanade path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker / timed CPU
baseline.  The product path (``optical-flow_b200/``) never imports this file and
has no CPU fallback.

Parity status: PINNED.  This restatement reproduces the reference's four golden
files (``U2Test.txt``/``V2Test.txt`` single level, ``UPyrTest.txt``/``VPyrTest.txt``
7 levels) to their print precision (<= 5e-7) -- see ``tests/test_oracle_golden.py``
and ``tests/golden/make_golden.py``.  Every OpenCV primitive restated here
(``filter2D``, ``pyrDown``, ``pyrUp``, ``remap``) was additionally checked against
``cv2`` 4.13 in the build container by ``tests/golden/make_golden.py``.

The arithmetic the reference delegates to OpenCV 2.4.x (un-vendored, version
identified only through ``/root/reference/Makefile:8-10``) is restated from its
published algorithm; the call sites are cited per function.  All line numbers are
into ``/root/reference/ParallelVideoPyr.cpp`` unless prefixed (serial twin:
``ctfLKof.cpp:31-215``).

Conventions: images are C-contiguous float64 ``(H, W)``; ``rn`` = one IEEE-754
binary64 rounding; numpy never contracts to FMA, which is what the reference's
materialised ``Mat`` expressions give.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "get_max_layer", "gradients", "box_sum", "solve", "lk_single",
    "pyr_down", "pyr_up", "build_maps", "remap_q5", "gaussian_pyramid",
    "coarse_to_fine", "gesture_stats",
]


# --------------------------------------------------------------------------- a2
def get_max_layer(width: int, height: int) -> int:
    """``getMaxLayer`` (:545-564, ``ctfLKof.cpp:196-215``).

    p_n = 1 + (number of times 2 divides n); result = min(p_W, p_H).
    """
    def p(n: int) -> int:
        k = 1
        while n % (2 ** k) == 0:
            k += 1
        return k
    if width <= 0 or height <= 0:
        raise ValueError("empty image")
    return min(p(width), p(height))


# --------------------------------------------------------------------------- a4
def _r101(idx: np.ndarray, n: int) -> np.ndarray:
    """BORDER_REFLECT_101 index map (``-1 -> 1``, ``n -> n-2``); n == 1 -> 0."""
    if n == 1:
        return np.zeros_like(idx)
    period = 2 * (n - 1)
    m = np.mod(idx, period)
    return np.where(m >= n, period - m, m)


def _taps(img: np.ndarray):
    """The four taps of a 2x2 ``filter2D`` with default anchor (1,1) and
    BORDER_REFLECT_101: p00=I[y-1][x-1], p01=I[y-1][x], p10=I[y][x-1], p11=I[y][x]."""
    h, w = img.shape
    ym = _r101(np.arange(-1, h - 1), h)
    xm = _r101(np.arange(-1, w - 1), w)
    up = img[ym, :]
    p00 = up[:, xm]
    p01 = up
    p10 = img[:, xm]
    p11 = img
    return p00, p01, p10, p11


def _corr2x2(img, k00, k01, k10, k11):
    """``filter2D(src, dst, -1, K)`` for a 2x2 kernel of +-1 entries, accumulated
    ((k00*p00 + k01*p01) + k10*p10) + k11*p11 (:51-52, 65-66, 78, 80)."""
    p00, p01, p10, p11 = _taps(img)
    return ((k00 * p00 + k01 * p01) + k10 * p10) + k11 * p11


def gradients(a: np.ndarray, b: np.ndarray):
    """``get_fx`` / ``get_fy`` / ``get_ft`` (:44-84).  No 1/4 scaling."""
    fx = _corr2x2(a, -1.0, 1.0, -1.0, 1.0) + _corr2x2(b, -1.0, 1.0, -1.0, 1.0)
    fy = _corr2x2(a, -1.0, -1.0, 1.0, 1.0) + _corr2x2(b, -1.0, -1.0, 1.0, 1.0)
    ft = _corr2x2(a, -1.0, -1.0, -1.0, -1.0) + _corr2x2(b, 1.0, 1.0, 1.0, 1.0)
    return fx, fy, ft


# --------------------------------------------------------------------------- a6
def box_sum(m: np.ndarray, r: int = 1) -> np.ndarray:
    """``get_Sum9_Mat`` (:162-191) generalised to radius ``r``.

    Interior ``r <= y <= H-1-r``: sum of the (2r+1)^2 window accumulated row-major
    from 0.0 (the order of ``get_Sum9`` :122-157).  Ring of width r is 0
    (``Mat::zeros`` :163, loop bounds :183,185).
    """
    h, w = m.shape
    out = np.zeros_like(m)
    if h <= 2 * r or w <= 2 * r:
        return out
    acc = np.zeros((h - 2 * r, w - 2 * r), dtype=np.float64)
    for i in range(-r, r + 1):
        for j in range(-r, r + 1):
            acc = acc + m[r + i:h - r + i, r + j:w - r + j]
    out[r:h - r, r:w - r] = acc
    return out


# --------------------------------------------------------------------------- a7
def solve(sxx, syy, sxy, sxt, syt, det_eps: float = 0.0):
    """2x2 normal-equation solve (:470-472, 510-520) with OpenCV-2.4 ``divide``
    semantics (zero divisor -> 0).  Valid <=> |det| > det_eps (det != 0 at 0.0)."""
    det = sxx * syy - sxy * sxy
    nu = sxy * syt - syy * sxt
    nv = sxt * sxy - sxx * syt
    mask = np.abs(det) > det_eps
    safe = np.where(mask, det, 1.0)
    u = np.where(mask, nu / safe, 0.0)
    v = np.where(mask, nv / safe, 0.0)
    return u, v, mask


# --------------------------------------------------------------------------- a8
def lk_single(a, b, r: int = 1, det_eps: float = 0.0, want_mask: bool = False):
    """``getLucasKanadeOpticalFlow`` (:327-542, ``ctfLKof.cpp:121-144``)."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    if a.shape != b.shape:
        raise ValueError("frame sizes differ")
    fx, fy, ft = gradients(a, b)
    sxx = box_sum(fx * fx, r)
    syy = box_sum(fy * fy, r)
    sxt = box_sum(fx * ft, r)
    sxy = box_sum(fx * fy, r)
    syt = box_sum(fy * ft, r)
    u, v, mask = solve(sxx, syy, sxy, sxt, syt, det_eps)
    if want_mask:
        return u, v, mask
    return u, v


# --------------------------------------------------------------------------- a3
def pyr_down(img: np.ndarray) -> np.ndarray:
    """``cv::pyrDown`` as called at :571 -- separable [1 4 6 4 1], decimate by 2,
    BORDER_REFLECT_101, out size ((W+1)/2, (H+1)/2).

    Horizontal first: h = s[2x]*6 + (s[2x-1] + s[2x+1])*4 + s[2x-2] + s[2x+2];
    vertical: (h[2y]*6 + (h[2y-1] + h[2y+1])*4 + h[2y-2] + h[2y+2]) * (1/256).
    """
    img = np.ascontiguousarray(img, dtype=np.float64)
    h, w = img.shape
    ow, oh = (w + 1) // 2, (h + 1) // 2
    xs = 2 * np.arange(ow)
    c = [img[:, _r101(xs + d, w)] for d in (-2, -1, 0, 1, 2)]
    hor = c[2] * 6.0 + (c[1] + c[3]) * 4.0 + c[0] + c[4]
    ys = 2 * np.arange(oh)
    rws = [hor[_r101(ys + d, h), :] for d in (-2, -1, 0, 1, 2)]
    return (rws[2] * 6.0 + (rws[1] + rws[3]) * 4.0 + rws[0] + rws[4]) * (1.0 / 256.0)


def gaussian_pyramid(img: np.ndarray, n_levels: int):
    """``getGaussianPyramid`` (:566-575).  Level 0 is a COPY here (the reference
    aliases the caller's buffer; the oracle never mutates inputs)."""
    pyr = [np.array(img, dtype=np.float64, copy=True)]
    for _ in range(n_levels - 1):
        pyr.append(pyr_down(pyr[-1]))
    return pyr


# -------------------------------------------------------------------------- a10
def _pyr_up_rows(s: np.ndarray) -> np.ndarray:
    """Horizontal pass of ``cv::pyrUp`` on every row of ``s`` -> width 2n."""
    h, n = s.shape
    r = np.empty((h, 2 * n), dtype=np.float64)
    if n == 1:
        r[:, 0] = s[:, 0] * 8.0
        r[:, 1] = s[:, 0] * 8.0
        return r
    r[:, 0] = s[:, 0] * 6.0 + s[:, 1] * 2.0
    r[:, 1] = (s[:, 0] + s[:, 1]) * 4.0
    if n > 2:
        r[:, 2:2 * n - 2:2] = s[:, 0:n - 2] + s[:, 1:n - 1] * 6.0 + s[:, 2:n]
        r[:, 3:2 * n - 2:2] = (s[:, 1:n - 1] + s[:, 2:n]) * 4.0
    r[:, 2 * n - 2] = s[:, n - 2] + s[:, n - 1] * 7.0
    r[:, 2 * n - 1] = s[:, n - 1] * 8.0
    return r


def pyr_up(img: np.ndarray) -> np.ndarray:
    """``cv::pyrUp`` as called at :646-647: out 2W x 2H, kernel [1 4 6 4 1] x 4.

    Vertical pass uses source rows rho(k) = R101(2k, 2H) / 2  (rho(-1)=1,
    rho(H)=H-1): dst[2y] = (r[rho(y-1)] + r[rho(y)]*6 + r[rho(y+1)]) / 64,
    dst[2y+1] = (r[rho(y)] + r[rho(y+1)]) * 4 / 64.  Flow VALUES are not doubled.
    """
    img = np.ascontiguousarray(img, dtype=np.float64)
    h, w = img.shape
    rows = _pyr_up_rows(img)
    k = np.arange(h)
    def rho(i):
        return _r101(2 * i, 2 * h) // 2
    r0, r1, r2 = rows[rho(k - 1)], rows[rho(k)], rows[rho(k + 1)]
    out = np.empty((2 * h, 2 * w), dtype=np.float64)
    out[0::2] = (r0 + r1 * 6.0 + r2) * (1.0 / 64.0)
    out[1::2] = ((r1 + r2) * 4.0) * (1.0 / 64.0)
    return out


# -------------------------------------------------------------------------- a11
def build_maps(upu: np.ndarray, upv: np.ndarray):
    """Map construction (:667-684): float32 arithmetic throughout.
    map1 = (x + fu/2, y + fv/2) samples frame 1; map2 = (x - fu/2, y - fv/2)."""
    h, w = upu.shape
    with np.errstate(over="ignore"):
        fu = upu.astype(np.float32)
        fv = upv.astype(np.float32)
    xs = np.arange(w, dtype=np.float32)[None, :]
    ys = np.arange(h, dtype=np.float32)[:, None]
    half = np.float32(2.0)
    m1 = (xs + fu / half, ys + fv / half)
    m2 = (xs - fu / half, ys - fv / half)
    return m1, m2


# -------------------------------------------------------------------------- a12
_COORD_LIM = np.float32(2.0 ** 30)


def _q5(m: np.ndarray):
    """OpenCV's fixed-point coordinate: s = rint_half_even(m * 32.f) (float mul),
    integer part s >> 5 saturated to int16, 5-bit fraction s & 31."""
    with np.errstate(over="ignore", invalid="ignore"):
        t = m.astype(np.float32) * np.float32(32.0)
    t = np.where(np.isfinite(t), t, _COORD_LIM)          # non-finite -> far outside
    t = np.clip(t, -_COORD_LIM, _COORD_LIM)
    s = np.rint(t).astype(np.int64)
    i = np.clip(s >> 5, -32768, 32767)
    a = s & 31
    return i, a


def remap_q5(src: np.ndarray, mapx: np.ndarray, mapy: np.ndarray) -> np.ndarray:
    """``cv::remap(src, dst, map, Mat(), INTER_LINEAR)`` (:691-703) for CV_64FC1
    data and a CV_32FC2 map: 1/32-px coordinate quantisation, float32 weights,
    double blend left to right, BORDER_CONSTANT 0 tested per tap."""
    src = np.ascontiguousarray(src, dtype=np.float64)
    h, w = src.shape
    ix, ax = _q5(mapx)
    iy, ay = _q5(mapy)
    one = np.float32(1.0 / 1024.0)
    w00 = ((32 - ay) * (32 - ax)).astype(np.float32) * one
    w01 = ((32 - ay) * ax).astype(np.float32) * one
    w10 = (ay * (32 - ax)).astype(np.float32) * one
    w11 = (ay * ax).astype(np.float32) * one

    def tap(yy, xx):
        ok = (yy >= 0) & (yy < h) & (xx >= 0) & (xx < w)
        return np.where(ok, src[np.clip(yy, 0, h - 1), np.clip(xx, 0, w - 1)], 0.0)

    return (tap(iy, ix) * w00.astype(np.float64)
            + tap(iy, ix + 1) * w01.astype(np.float64)
            + tap(iy + 1, ix) * w10.astype(np.float64)
            + tap(iy + 1, ix + 1) * w11.astype(np.float64))


# -------------------------------------------------------------------------- a13
def coarse_to_fine(a, b, n_levels: int, r: int = 1, det_eps: float = 0.0,
                   want_mask: bool = False, trace: dict | None = None):
    """``coarseToFineEstimation`` (:578-725, ``ctfLKof.cpp:157-194``).

    Inputs are not modified (Q1).  Raises ValueError where the reference would
    throw a size-mismatch ``cv::Exception`` (W or H not divisible by 2^(L-1), Q2).
    ``trace`` (optional dict) receives per-level intermediates for parity triage.
    """
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    if a.shape != b.shape:
        raise ValueError("frame sizes differ")
    if n_levels < 1:
        raise ValueError("n_levels < 1")
    h, w = a.shape
    f = 1 << (n_levels - 1)
    if w % f or h % f:
        raise ValueError("size not divisible by 2^(n_levels-1)")
    pa = gaussian_pyramid(a, n_levels)
    pb = gaussian_pyramid(b, n_levels)
    upu = upv = None
    for lvl in range(n_levels - 1, -1, -1):
        tu, tv, mask = lk_single(pa[lvl], pb[lvl], r, det_eps, want_mask=True)
        if lvl != n_levels - 1:
            tu = tu + upu
            tv = tv + upv
        if trace is not None:
            trace[("A", lvl)] = pa[lvl].copy()
            trace[("B", lvl)] = pb[lvl].copy()
            trace[("u", lvl)] = tu.copy()
            trace[("v", lvl)] = tv.copy()
            trace[("mask", lvl)] = mask.copy()
        if lvl == 0:
            return (tu, tv, mask) if want_mask else (tu, tv)
        upu = pyr_up(tu)
        upv = pyr_up(tv)
        (m1x, m1y), (m2x, m2y) = build_maps(upu, upv)
        pa[lvl - 1] = remap_q5(pa[lvl - 1], m1x, m1y)
        pb[lvl - 1] = remap_q5(pb[lvl - 1], m2x, m2y)


# ------------------------------------------------------------------ 8(f) rank 1
def gesture_stats(u: np.ndarray, v: np.ndarray, mag_thresh: float = 20.0,
                  count_thresh: int = 500):
    """Gesture statistics of the live loop (:845-888): val = sqrt(u^2+v^2);
    val < 20 -> dropped, else sum_x += j, sum_y += i, count++; gesture present
    iff count > 500; centroid = sums / count (integer Point in the reference)."""
    val = np.sqrt(u * u + v * v)
    keep = ~(val < mag_thresh)
    ys, xs = np.nonzero(keep)
    count = int(keep.sum())
    sx = float(xs.sum())
    sy = float(ys.sum())
    return {"sum_x": sx, "sum_y": sy, "count": count,
            "detected": bool(count > count_thresh)}
