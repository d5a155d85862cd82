"""TEST INFRASTRUCTURE -- CPU restatement (numpy) of the live loop's pre-processing between the
gray frame and the LK input, /root/reference/ParallelVideoPyr.cpp:791-813:

    cvtColor(BGR2GRAY) :791, diff = cur - 0.3 * prev :800 (8-bit, saturating), threshold(TOZERO, 10)
    :803, Sobel x + Sobel y -> CV_64F :808-811, dilate x2 :812, erode x2 :813.

The arithmetic lives in OpenCV (2.4 in the reference, un-vendored); the reference ships no golden
vector for this stage.  tests/golden/make_live_golden.py pins every function below against cv2 4.13
in the build container (exhaustively for the two 8-bit steps) and commits fixtures.  resize (:782)
and GaussianBlur (:790) are left to the capture front end (host OpenCV) and are not restated.
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
