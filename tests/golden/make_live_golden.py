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
    print("oracle == cv2 %s on every check; fixture written (nonzero outputs: %d)" % (cv2.__version__, int((out != 0).sum())))


if __name__ == "__main__":
    main()
