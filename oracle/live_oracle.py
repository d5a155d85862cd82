"""TEST INFRASTRUCTURE -- CPU restatement (numpy) of the live loop's pre-processing between the
gray frame and the LK input, /root/reference/ParallelVideoPyr.cpp:791-813:

    cvtColor(BGR2GRAY) :791, diff = cur - 0.3 * prev :800 (8-bit, saturating), threshold(TOZERO, 10)
    :803, Sobel x + Sobel y -> CV_64F :808-811, dilate x2 :812, erode x2 :813.

plus the two steps before it on the 8-bit colour frame: resize(1080 x 1080, INTER_CUBIC) :782 and
GaussianBlur(9 x 9, sigma 1.5) :790.

The arithmetic lives in OpenCV (2.4 in the reference, un-vendored); the reference ships no golden
vector for this stage.  tests/golden/make_live_golden.py pins every function below against cv2 4.13
in the build container (exhaustively for the two 8-bit steps) and commits fixtures.  Pin status of the
two colour-frame steps: GaussianBlur -- cv2 4.13's fixed-point path, and OpenCV 2.4's published source
gives the same integer kernel for this size and sigma; resize -- cv2 4.13's OWN implementation
(cv2.ipp.setUseIPP(False)): with IPP enabled cv2 hands 8-bit bicubic to Intel IPP, whose result differs
by one grey level on about 5 % of the pixels; OpenCV 2.4's SSE2 vertical pass is float32 as well, its
association order is unverified ("parity unpinned" against the reference's own library for that step).
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import numpy as np


GRAY_COEFFS = {14: (1868, 9617, 4899), 15: (3735, 19235, 9798)}     # (B, G, R) weights


def bgr2gray(bgr, bits=14):
    """cv::cvtColor(CV_BGR2GRAY) on 8-bit (:791), fixed point:
    (B*cb + G*cg + R*cr + 2^(bits-1)) >> bits.  bits = 14 (1868, 9617, 4899) is OpenCV 2.4's
    table, the reference's library (published source, not runnable here -- unpinned);
    bits = 15 (3735, 19235, 9798) is OpenCV 3.x/4.x's, pinned against cv2 4.13.  The two differ
    by one grey level on about 0.3 % of random pixels."""
    cb, cg, cr = GRAY_COEFFS[bits]
    b = bgr[..., 0].astype(np.int64)
    g = bgr[..., 1].astype(np.int64)
    r = bgr[..., 2].astype(np.int64)
    return ((b * cb + g * cg + r * cr + (1 << (bits - 1))) >> bits).astype(np.uint8)


def weighted_diff_u8(cur, prev, learning_rate=0.3):
    """Mat diff = current_frame - LEARNING_RATE * prevFrame (:800) on CV_8U: the MatExpr is
    cv::addWeighted(cur, 1, prev, -rate, 0): float32 products and sums, round half to even,
    saturate to [0, 255]."""
    a = np.float32(1.0)
    b = np.float32(-learning_rate)
    t = cur.astype(np.float32) * a + prev.astype(np.float32) * b      # each op rounded to float32
    return np.clip(np.rint(t), 0, 255).astype(np.uint8)


def threshold_tozero(img, thresh=10):
    """cv::threshold(THRESH_TOZERO) (:803): v > thresh ? v : 0."""
    return np.where(img > thresh, img, 0).astype(img.dtype)


def _r101(i, n):
    if n == 1:
        return np.zeros_like(i)
    i = np.abs(i)
    return np.where(i >= n, 2 * (n - 1) - i, i)


def sobel_xy_sum(img_u8):
    """Sobel(dx) + Sobel(dy), ksize 3, CV_64F output, BORDER_REFLECT_101 (:808-811).  Integer valued."""
    h, w = img_u8.shape
    s = img_u8.astype(np.int64)
    ys = np.arange(h)
    xs = np.arange(w)
    def at(dy, dx):
        return s[_r101(ys + dy, h)][:, _r101(xs + dx, w)]
    sx = (at(-1, 1) + 2 * at(0, 1) + at(1, 1)) - (at(-1, -1) + 2 * at(0, -1) + at(1, -1))
    sy = (at(1, -1) + 2 * at(1, 0) + at(1, 1)) - (at(-1, -1) + 2 * at(-1, 0) + at(-1, 1))
    return (sx + sy).astype(np.float64)


def _morph(img, radius, op):
    """max / min over the (2*radius+1)^2 square; pixels outside the image do not take part
    (cv::dilate / cv::erode with the default border value)."""
    h, w = img.shape
    fill = -np.inf if op is np.maximum else np.inf
    p = np.full((h + 2 * radius, w + 2 * radius), fill)
    p[radius:radius + h, radius:radius + w] = img
    out = p[radius:radius + h, radius:radius + w].copy()
    tmp = p[:, radius:radius + w].copy()
    for d in range(-radius, radius + 1):
        tmp = op(tmp, p[:, radius + d:radius + d + w])
    out = tmp[radius:radius + h]
    for d in range(-radius, radius + 1):
        out = op(out, tmp[radius + d:radius + d + h])
    return out


def dilate_erode(img, iterations=2):
    """dilate(3x3, iterations) then erode(3x3, iterations) (:812-813); n iterations of the 3x3
    rectangle equal one pass of the (2n+1)^2 rectangle."""
    return _morph(_morph(img, iterations, np.maximum), iterations, np.minimum)


def live_diff(cur_gray, prev_gray, learning_rate=0.3, diff_thresh=10, iterations=2):
    """:800-813: the LK input image of one live-loop iteration from two consecutive gray frames."""
    d = threshold_tozero(weighted_diff_u8(cur_gray, prev_gray, learning_rate), diff_thresh)
    return dilate_erode(sobel_xy_sum(d), iterations)


# ------------------------------------------------------------------ colour-frame steps :782, :790
GAUSS9_S15 = (2, 9, 28, 55, 68, 55, 28, 9, 2)      # round(256 * exp(-x^2 / 4.5) / sum), sums to 256


def gaussian_kernel_q8(ksize=9, sigma=1.5):
    """8-bit fixed-point kernel of cv::GaussianBlur on CV_8U: exp(-x^2 / (2 sigma^2)) normalised, times
    256, rounded half to even with the rounding error carried to the next tap (cv2 4.x
    getGaussianKernelFixedPoint_ED; for 9 / 1.5 plain rounding -- OpenCV 2.4's cvRound(k * 256) --
    gives the same integers)."""
    x = np.arange(ksize) - (ksize - 1) / 2.0
    k = np.exp(-x * x / (2.0 * sigma * sigma))
    k /= k.sum()
    out = np.zeros(ksize, np.int64)
    err = 0.0
    for i in range(ksize // 2 + 1):
        adj = k[i] * 256.0 + err
        v = int(np.rint(adj))
        err = adj - v
        out[i] = out[ksize - 1 - i] = v
    return out


def _r101_any(i, n):
    """BORDER_REFLECT_101 for any distance (cv::borderInterpolate reflects repeatedly)."""
    if n == 1:
        return np.zeros_like(i)
    p = 2 * (n - 1)
    i = np.abs(i) % p
    return np.where(i >= n, p - i, i)


def gaussian_blur_u8(img, ksize=9, sigma=1.5):
    """cv::GaussianBlur(img, Size(k, k), sigma, sigma) on 8-bit, any channel count (:790):
    separable, BORDER_REFLECT_101, integer kernel / 256 in both passes, one rounding at the end:
    (sum + 2^15) >> 16."""
    k = gaussian_kernel_q8(ksize, sigma)
    r = ksize // 2
    h, w = img.shape[:2]
    src = img.astype(np.int64)
    xs, ys = np.arange(w), np.arange(h)
    row = np.zeros_like(src)
    for d in range(-r, r + 1):
        row += src[:, _r101_any(xs + d, w)] * k[d + r]
    col = np.zeros_like(src)
    for d in range(-r, r + 1):
        col += row[_r101_any(ys + d, h)] * k[d + r]
    return np.clip((col + (1 << 15)) >> 16, 0, 255).astype(np.uint8)


def _cubic_coeffs(fx):
    """cv::interpolateCubic, A = -0.75, float32 arithmetic in OpenCV's order."""
    f = np.float32
    x, A = f(fx), f(-0.75)
    c0 = ((A * (x + f(1)) - f(5) * A) * (x + f(1)) + f(8) * A) * (x + f(1)) - f(4) * A
    c1 = ((A + f(2)) * x - (A + f(3))) * x * x + f(1)
    c2 = ((A + f(2)) * (f(1) - x) - (A + f(3))) * (f(1) - x) * (f(1) - x) + f(1)
    c3 = f(1) - c0 - c1 - c2
    return c0, c1, c2, c3


def cubic_table(src_n, dst_n):
    """Per destination index: the four source indices (clamped: BORDER_REPLICATE) and the four
    coefficients as 11-bit fixed point (INTER_RESIZE_COEF_BITS), cv::resize's table for INTER_CUBIC
    on 8-bit: fx = (float)((d + 0.5) * scale - 0.5) with scale = 1 / ((double)dst / src)."""
    scale = 1.0 / (float(dst_n) / src_n)
    ofs = np.zeros((dst_n, 4), np.int64)
    co = np.zeros((dst_n, 4), np.int64)
    for d in range(dst_n):
        fx = np.float32((d + 0.5) * scale - 0.5)
        sx = int(np.floor(fx))
        c = _cubic_coeffs(np.float32(fx - np.float32(sx)))
        for k in range(4):
            ofs[d, k] = min(max(sx - 1 + k, 0), src_n - 1)
            co[d, k] = int(np.rint(np.float32(c[k] * np.float32(2048))))
    return ofs, co


def resize_cubic_u8(img, dst_w, dst_h):
    """cv::resize(img, Size(dst_w, dst_h), 0, 0, INTER_CUBIC) on 8-bit, any channel count (:782), as
    cv2 4.13 computes it without IPP: horizontal pass in int32 (source * 11-bit coefficients), vertical
    pass in float32 -- row k times (beta_k / 2^22), accumulated from row 3 down to row 0, every product
    and sum rounded -- then round half to even and saturate."""
    h, w = img.shape[:2]
    xo, xc = cubic_table(w, dst_w)
    yo, yc = cubic_table(h, dst_h)
    src = img.astype(np.int64)
    sh = (1, dst_w) + (1,) * (img.ndim - 2)
    sv = (dst_h, 1) + (1,) * (img.ndim - 2)
    hor = sum(src[:, xo[:, k]] * xc[:, k].reshape(sh) for k in range(4))
    scale = np.float32(1.0) / np.float32(2048 * 2048)
    acc = None
    for k in (3, 2, 1, 0):
        t = hor[yo[:, k]].astype(np.float32) * (yc[:, k].astype(np.float32) * scale).reshape(sv)
        acc = t if acc is None else t + acc
    return np.clip(np.rint(acc), 0, 255).astype(np.uint8)


def live_frame_gray(frame_bgr, width=1080, height=1080, bits=15, ksize=9, sigma=1.5):
    """:782-791: one captured colour frame -> the blurred gray working frame (cv::resize copies when the
    sizes agree)."""
    f = frame_bgr if frame_bgr.shape[:2] == (height, width) else resize_cubic_u8(frame_bgr, width, height)
    return bgr2gray(gaussian_blur_u8(f, ksize, sigma), bits)
