#!/usr/bin/env python
"""Pins oracle/live_oracle.py against cv2 (build container only) and writes the fixture
tests/golden/live_golden.npz: three consecutive pacman.mp4 frames (BGR, 8-bit) with cv2's gray
frames and the :800-813 outputs as int16 (the values are integers, |v| <= 2040).

Usage: python tests/golden/make_live_golden.py"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import live_oracle as L  # noqa: E402


def cv_live_diff(cur, prev):
    d = cv2.addWeighted(cur, 1.0, prev, -0.3, 0.0)
    _, d = cv2.threshold(d, 10, 255, cv2.THRESH_TOZERO)
    s = cv2.Sobel(d, cv2.CV_64F, 1, 0) + cv2.Sobel(d, cv2.CV_64F, 0, 1)
    s = cv2.dilate(s, None, iterations=2)
    return cv2.erode(s, None, iterations=2)


def main():
    # exhaustive: the two 8-bit steps
    a, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    assert np.array_equal(cv2.addWeighted(a, 1.0, b, -0.3, 0.0), L.weighted_diff_u8(a, b)), "addWeighted"
    rng = np.random.default_rng(3)
    bgr = rng.integers(0, 256, (64, 4096, 3), dtype=np.uint8)
    assert np.array_equal(cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY), L.bgr2gray(bgr, 15)), "bgr2gray"
    for (h, w) in [(1, 7), (7, 1), (2, 2), (5, 9), (33, 47), (160, 240)]:
        cur = rng.integers(0, 256, (h, w), dtype=np.uint8)
        prev = rng.integers(0, 256, (h, w), dtype=np.uint8)
        d = L.threshold_tozero(L.weighted_diff_u8(cur, prev))
        assert np.array_equal(cv2.Sobel(d, cv2.CV_64F, 1, 0) + cv2.Sobel(d, cv2.CV_64F, 0, 1), L.sobel_xy_sum(d)), ("sobel", h, w)
        assert np.array_equal(cv_live_diff(cur, prev), L.live_diff(cur, prev)), ("live_diff", h, w)
    cap = cv2.VideoCapture("/root/reference/pacman.mp4")
    frames = []
    for _ in range(40):
        ok, f = cap.read()
        assert ok
        frames.append(f)
    frames = np.stack(frames[30:33])
    gray = np.stack([cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in frames])
    assert np.array_equal(gray, L.bgr2gray(frames, 15))
    out = np.stack([cv_live_diff(gray[k + 1], gray[k]) for k in range(2)])
    assert np.array_equal(out, np.stack([L.live_diff(gray[k + 1], gray[k]) for k in range(2)]))
    assert np.abs(out).max() <= 2040 and np.array_equal(out, np.rint(out))
    np.savez_compressed(os.path.join(HERE, "live_golden.npz"), bgr=frames, gray=gray, out=out.astype(np.int16))
    # ---- colour-frame steps :782, :790: cv2's own code (IPP hands 8-bit bicubic to a different algorithm)
    cv2.ipp.setUseIPP(False)
    for shp, (dw, dh) in [((72, 128, 3), (108, 108)), ((37, 53, 3), (64, 40)), ((120, 200, 1), (50, 70)), ((9, 5, 3), (20, 20)),
                          ((160, 240, 3), (1080, 1080)), ((480, 640, 3), (1080, 1080)), ((720, 1280, 3), (1080, 1080))]:
        img = rng.integers(0, 256, shp, dtype=np.uint8)
        a = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_CUBIC).reshape(dh, dw, shp[2])
        assert np.array_equal(a, L.resize_cubic_u8(img, dw, dh)), ("resize", shp, dw, dh)
        g = cv2.GaussianBlur(img, (9, 9), 1.5).reshape(shp)
        assert np.array_equal(g, L.gaussian_blur_u8(img)), ("blur", shp)
    assert tuple(L.gaussian_kernel_q8(9, 1.5)) == L.GAUSS9_S15
    # OpenCV 2.4's kernel for the same call: cvRound(k * 256) of the float kernel (no error diffusion)
    x = np.arange(9) - 4.0
    k = np.exp(-x * x / 4.5)
    k /= k.sum()
    assert tuple(int(v) for v in np.rint(k.astype(np.float32) * 256)) == L.GAUSS9_S15
    small = frames[:, 16:112, 40:168]                          # 3 frames, 96 x 128 BGR
    res = np.stack([cv2.resize(f, (160, 120), interpolation=cv2.INTER_CUBIC) for f in small])
    blur = np.stack([cv2.GaussianBlur(f, (9, 9), 1.5) for f in res])
    gray2 = np.stack([cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in blur])
    assert np.array_equal(res, np.stack([L.resize_cubic_u8(f, 160, 120) for f in small]))
    assert np.array_equal(gray2, np.stack([L.bgr2gray(L.gaussian_blur_u8(f), 15) for f in res]))
    np.savez_compressed(os.path.join(HERE, "live_frame_golden.npz"), bgr=small, resized=res, blurred=blur, gray=gray2)
    print("oracle == cv2 %s on every check; fixture written (nonzero outputs: %d)" % (cv2.__version__, int((out != 0).sum())))


if __name__ == "__main__":
    main()
