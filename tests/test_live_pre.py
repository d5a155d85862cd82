"""Live-loop pre-processing (ParallelVideoPyr.cpp:791-813): cvtColor :791,
``cur - 0.3 * prev`` :800, threshold :803, Sobel x + Sobel y :808-811, dilate x2 / erode x2 :812-813.

CPU part: oracle/live_oracle.py against the committed cv2 4.13 fixture
(tests/golden/live_golden.npz, written by tests/golden/make_live_golden.py).
GPU part (-m gpu): the CUDA kernels through the C ABI against the oracle and the fixture.
Bar: bit-exact (8-bit / integer-valued work)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import live_oracle as LO


@pytest.fixture(scope="module")
def live_golden():
    d = np.load(os.path.join(GOLDEN, "live_golden.npz"))
    return d["bgr"], d["gray"], d["out"].astype(np.float64)


# ------------------------------------------------------------------ oracle vs cv2 fixture (CPU)
def test_oracle_gray_matches_cv2_fixture(live_golden):
    bgr, gray, _ = live_golden
    assert np.array_equal(LO.bgr2gray(bgr, 15), gray)
    g14 = LO.bgr2gray(bgr, 14)                       # OpenCV 2.4 table: at most one grey level away
    assert np.abs(g14.astype(int) - gray.astype(int)).max() <= 1


def test_oracle_live_diff_matches_cv2_fixture(live_golden):
    _, gray, out = live_golden
    for k in range(2):
        assert np.array_equal(LO.live_diff(gray[k + 1], gray[k]), out[k])
    assert (out != 0).sum() > 1000                   # the fixture exercises the non-trivial branch


def test_oracle_weighted_diff_rounding_and_saturation():
    cur = np.array([[0, 255, 10, 11, 100, 5]], np.uint8)
    prev = np.array([[255, 0, 0, 5, 50, 255]], np.uint8)
    # 0-76.5 -> 0 (saturate); 255; 10; 11-1.5=9.5 -> 10 (half to even in float32: 9.5 -> 10);
    # 100-15=85; 5-76.5 -> 0
    assert LO.weighted_diff_u8(cur, prev).tolist() == [[0, 255, 10, 10, 85, 0]]
    assert LO.threshold_tozero(np.array([[9, 10, 11]], np.uint8)).tolist() == [[0, 0, 11]]


def test_oracle_morph_is_iterated_3x3():
    rng = np.random.default_rng(5)
    x = rng.integers(-50, 50, (13, 17)).astype(np.float64)
    once = LO._morph(LO._morph(x, 1, np.maximum), 1, np.maximum)
    assert np.array_equal(once, LO._morph(x, 2, np.maximum))
    once = LO._morph(LO._morph(x, 1, np.minimum), 1, np.minimum)
    assert np.array_equal(once, LO._morph(x, 2, np.minimum))


# ------------------------------------------------------------------ CUDA path (GPU)
@pytest.mark.gpu
def test_gpu_fixture_bgr_frames(ofb, live_golden):
    bgr, gray, out = live_golden
    with ofb.Context(0) as ctx:
        for k in range(2):
            got = ctx.live_diff(gray[k + 1], gray[k])
            assert np.array_equal(got, out[k]), "gray input, pair %d" % k
            got = ctx.live_diff(bgr[k + 1], bgr[k], coeff_bits=15)
            assert np.array_equal(got, out[k]), "BGR input (OpenCV >= 3 table), pair %d" % k
            want14 = LO.live_diff(LO.bgr2gray(bgr[k + 1], 14), LO.bgr2gray(bgr[k], 14))
            got = ctx.live_diff(bgr[k + 1], bgr[k])          # reference's library: OpenCV 2.4 table
            assert np.array_equal(got, want14), "BGR input (OpenCV 2.4 table), pair %d" % k


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [14, 15])
@pytest.mark.parametrize("shape", [(1, 1), (1, 7), (7, 1), (5, 9), (33, 47), (64, 257), (160, 240), (1080, 1080)])
def test_gpu_bgr2gray_dev(ofb, shape, bits):
    import torch
    h, w = shape
    rng = np.random.default_rng(h * 7 + w + bits)
    bgr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    d_bgr = torch.from_numpy(bgr).cuda()
    d_gray = torch.zeros((h, w), dtype=torch.uint8, device="cuda")
    with ofb.Context(0) as ctx:
        ctx.bgr2gray_dev(d_bgr, d_gray, coeff_bits=bits)
        ctx.synchronize()
    assert np.array_equal(d_gray.cpu().numpy(), LO.bgr2gray(bgr, bits))


@pytest.mark.gpu
@pytest.mark.parametrize("iters", [1, 2, 3, 4])
@pytest.mark.parametrize("shape", [(1, 1), (1, 7), (7, 1), (2, 2), (5, 9), (33, 47), (31, 130), (160, 240), (540, 540)])
def test_gpu_live_diff_random(ofb, shape, iters):
    h, w = shape
    rng = np.random.default_rng(h * 11 + w + iters)
    # smooth-ish frames so that the thresholded difference has both zero and non-zero regions
    cur = rng.integers(0, 256, (h, w), dtype=np.uint8)
    prev = np.where(rng.random((h, w)) < 0.5, cur, rng.integers(0, 256, (h, w), dtype=np.uint8)).astype(np.uint8)
    with ofb.Context(0) as ctx:
        got = ctx.live_diff(cur, prev, morph_iterations=iters)
    want = LO.live_diff(cur, prev, iterations=iters)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_live_diff_params_and_dev_entry(ofb):
    import torch
    rng = np.random.default_rng(9)
    h, w = 97, 203
    cur = rng.integers(0, 256, (h, w), dtype=np.uint8)
    prev = rng.integers(0, 256, (h, w), dtype=np.uint8)
    with ofb.Context(0) as ctx:
        got = ctx.live_diff(cur, prev, learning_rate=0.55, diff_thresh=40)
        assert np.array_equal(got, LO.live_diff(cur, prev, learning_rate=0.55, diff_thresh=40))
        # device entry with padded rows (step > width)
        d_cur = torch.zeros((h, 256), dtype=torch.uint8, device="cuda")
        d_prev = torch.zeros((h, 256), dtype=torch.uint8, device="cuda")
        d_cur[:, :w] = torch.from_numpy(cur).cuda()
        d_prev[:, :w] = torch.from_numpy(prev).cuda()
        d_out = torch.full((h, 208), -7.0, dtype=torch.float64, device="cuda")
        ctx.live_diff_dev(d_cur[:, :w], d_prev[:, :w], d_out[:, :w])
        ctx.synchronize()
        o = d_out.cpu().numpy()
        assert np.array_equal(o[:, :w], LO.live_diff(cur, prev))
        assert (o[:, w:] == -7.0).all()                      # padding untouched
        with pytest.raises(ofb.OfError):
            ctx.live_diff(cur, prev, morph_iterations=5)
        with pytest.raises(ofb.OfError):
            ctx.live_diff(cur, prev[:, :-1])


@pytest.mark.gpu
def test_gpu_live_pair_feeds_lk(ofb, c_oracle, live_golden):
    """One live-loop iteration, headless: frames -> live_diff -> coarseToFineEstimation (:831-832),
    against the oracle chain on the same frames."""
    bgr, _, _ = live_golden
    g = LO.bgr2gray(bgr, 14)
    want_prev, want_cur = LO.live_diff(g[1], g[0]), LO.live_diff(g[2], g[1])
    L = ofb.getMaxLayer(240, 160)
    with ofb.Context(0) as ctx:
        prev_diff = ctx.live_diff(bgr[1], bgr[0])
        diff = ctx.live_diff(bgr[2], bgr[1])
        assert np.array_equal(prev_diff, want_prev) and np.array_equal(diff, want_cur)
        u, v, m = ctx.coarseToFineEstimation(prev_diff, diff, L, want_mask=True)
    uo, vo, mo = c_oracle.coarse_to_fine(want_prev, want_cur, L, want_mask=True)
    assert np.array_equal(m, mo) and np.array_equal(u, uo) and np.array_equal(v, vo)
    assert (~mo).sum() > 2 * (240 + 160)                     # interior det == 0 pixels exist (sparse input)
