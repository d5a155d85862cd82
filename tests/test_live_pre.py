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


# ------------------------------------------------------------------ colour-frame steps :782, :790
@pytest.fixture(scope="module")
def frame_golden():
    d = np.load(os.path.join(GOLDEN, "live_frame_golden.npz"))
    return d["bgr"], d["resized"], d["blurred"], d["gray"]


def test_oracle_resize_blur_match_cv2_fixture(frame_golden):
    """cv2 4.13 (IPP off) wrote the fixture: resize(INTER_CUBIC) 128x96 -> 160x120, GaussianBlur(9x9, 1.5), gray."""
    bgr, res, blur, gray = frame_golden
    for k in range(len(bgr)):
        assert np.array_equal(LO.resize_cubic_u8(bgr[k], 160, 120), res[k])
        assert np.array_equal(LO.gaussian_blur_u8(res[k]), blur[k])
        assert np.array_equal(LO.bgr2gray(blur[k], 15), gray[k])
        assert np.array_equal(LO.live_frame_gray(bgr[k], 160, 120, 15), gray[k])
    assert tuple(LO.gaussian_kernel_q8(9, 1.5)) == LO.GAUSS9_S15 and sum(LO.GAUSS9_S15) == 256


def test_oracle_cubic_table_properties():
    ofs, co = LO.cubic_table(100, 150)
    assert (co.sum(axis=1) >= 2046).all() and (co.sum(axis=1) <= 2050).all()      # about 1 in 11-bit fixed point
    assert ofs.min() == 0 and ofs.max() == 99                                     # taps are clamped to the image
    ofs, co = LO.cubic_table(64, 64)                                              # same size: identity taps
    assert (co[:, 1] == 2048).all() and (co[:, [0, 2, 3]] == 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("case", [((72, 128, 3), (108, 108)), ((37, 53, 3), (64, 40)), ((120, 200, 1), (50, 70)),
                                  ((9, 5, 3), (20, 20)), ((160, 240, 3), (1080, 1080)), ((720, 1280, 3), (1080, 1080)),
                                  ((1080, 1920, 3), (1080, 1080)), ((33, 47, 1), (33, 47))])
def test_gpu_resize_cubic(ofb, case):
    import torch
    shp, (dw, dh) = case
    rng = np.random.default_rng(shp[0] * 31 + shp[1] + dw)
    img = rng.integers(0, 256, shp, dtype=np.uint8)
    src = torch.from_numpy(img if shp[2] == 3 else img[:, :, 0]).cuda()
    dst = torch.zeros((dh, dw, 3) if shp[2] == 3 else (dh, dw), dtype=torch.uint8, device="cuda")
    with ofb.Context(0) as ctx:
        ctx.resize_cubic_dev(src, dst)
        ctx.synchronize()
    want = LO.resize_cubic_u8(img, dw, dh) if (dh, dw) != shp[:2] else img
    assert np.array_equal(dst.cpu().numpy().reshape(dh, dw, shp[2]), want)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 1, 3), (3, 2, 3), (9, 5, 3), (37, 53, 3), (64, 257, 1), (160, 240, 3), (1080, 1080, 3)])
def test_gpu_gaussian_blur(ofb, shape):
    import torch
    rng = np.random.default_rng(shape[0] * 13 + shape[1])
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    h, w, cn = shape
    src = torch.from_numpy(img if cn == 3 else img[:, :, 0]).cuda()
    with ofb.Context(0) as ctx:
        dst = torch.zeros_like(src)
        ctx.gaussian_blur_dev(src, dst)
        ctx.synchronize()
        want = LO.gaussian_blur_u8(img)
        assert np.array_equal(dst.cpu().numpy().reshape(shape), want)
        if cn == 3:
            for bits in (14, 15):
                g = torch.zeros((h, w), dtype=torch.uint8, device="cuda")
                ctx.gaussian_blur_dev(src, g, gray_coeff_bits=bits)
                ctx.synchronize()
                assert np.array_equal(g.cpu().numpy(), LO.bgr2gray(want, bits)), bits
            d5 = torch.zeros_like(src)
            ctx.gaussian_blur_dev(src, d5, ksize=5, sigma=1.1)
            ctx.synchronize()
            assert np.array_equal(d5.cpu().numpy(), LO.gaussian_blur_u8(img, 5, 1.1))
        with pytest.raises(ofb.OfError):
            ctx.gaussian_blur_dev(src, dst, ksize=4)


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [14, 15])
@pytest.mark.parametrize("write_back", [False, True])
def test_gpu_live_step_matches_the_oracle_chain(ofb, c_oracle, frame_golden, bits, write_back):
    """of_live_step_u8 over five frames == resize -> blur -> gray -> difference image -> pyramidal LK ->
    statistics of the oracle, iteration by iteration (main() :769-905); with write_back_warped the next
    iteration's prevDiff is the warped difference image (quirk Q1, :700-701 + :843)."""
    from oracle import lk_oracle as O
    bgr = frame_golden[0]
    frames = [bgr[0], bgr[1], bgr[2], bgr[1], bgr[0]]
    W, H = 160, 120
    L = ofb.getMaxLayer(W, H)
    gray = [LO.live_frame_gray(f, W, H, bits) for f in frames]
    with ofb.Context(0, mag_thresh=0.25, count_thresh=20, write_back_warped=write_back) as ctx:
        prev_diff = None
        for k, f in enumerate(frames):
            stage, stats, u, v = ctx.live_step(f, want_flow=True, work_width=W, work_height=H, coeff_bits=bits)
            assert stage == min(k, 2)
            if k == 0:
                continue
            diff = LO.live_diff(gray[k], gray[k - 1])
            if k >= 2:
                tr = {}
                uo, vo = O.coarse_to_fine(prev_diff, diff, L, trace=tr)
                assert np.array_equal(u, uo) and np.array_equal(v, vo), "iteration %d" % k
                s = c_oracle.gesture_stats(uo, vo, 0.25, 20)
                assert (stats["sum_x"], stats["sum_y"], stats["count"], bool(stats["detected"])) == \
                    (s["sum_x"], s["sum_y"], s["count"], s["detected"])
                if write_back:
                    diff = tr[("B", 0)]                       # the reference's diff is its warped version by now
            prev_diff = diff
        ctx.live_reset()
        assert ctx.live_step(frames[0], work_width=W, work_height=H)[0] == 0
