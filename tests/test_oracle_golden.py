"""CPU suite: pins the oracle (numpy and C restatements) to the reference's own golden
vectors and checks the two restatements against each other bit for bit."""
import numpy as np
import pytest

from oracle import lk_oracle as O


def test_get_max_layer_table(c_oracle):
    # values from the reference's configurations
    for (w, h), want in {(640, 448): 7, (1080, 1080): 4, (1920, 1080): 4, (3840, 2160): 5,
                         (3264, 2448): 5, (240, 160): 5, (320, 180): 3, (641, 448): 1, (8, 7): 1}.items():
        assert O.get_max_layer(w, h) == want
        assert c_oracle.get_max_layer(w, h) == want


def test_kat1_single_level_vs_reference_golden(table_pair, kat_golden, c_oracle):
    # serialOrig.cpp:575-635 / parallelBasic.cpp:585-647: |d| > 1e-6 fails
    a, b = table_pair
    for impl in (O, c_oracle):
        u, v = impl.lk_single(a, b)
        assert np.abs(u - kat_golden["U2Test"]).max() <= 1e-6
        assert np.abs(v - kat_golden["V2Test"]).max() <= 1e-6
    # Q3: ring is exactly zero for a single level
    assert not u[0].any() and not u[-1].any() and not u[:, 0].any() and not u[:, -1].any()


def test_kat7_pyramid_vs_reference_golden(table_pair, kat_golden, c_oracle):
    # ParallelBasicPyr.cpp:774-832, SerialPyr.cpp:261-319
    a, b = table_pair
    L = O.get_max_layer(640, 448)
    assert L == 7
    un, vn = O.coarse_to_fine(a, b, L)
    uc, vc = c_oracle.coarse_to_fine(a, b, L)
    for u, v in ((un, vn), (uc, vc)):
        assert np.abs(u - kat_golden["UPyrTest"]).max() <= 1e-6
        assert np.abs(v - kat_golden["VPyrTest"]).max() <= 1e-6
    assert np.array_equal(un, uc) and np.array_equal(vn, vc)


@pytest.mark.parametrize("r", [1, 2, 7])
@pytest.mark.parametrize("L", [1, 3])
def test_c_oracle_bit_exact_with_numpy_oracle(table_pair, c_oracle, r, L):
    a, b = table_pair
    a, b = a[:128, :160], b[:128, :160]
    un, vn, mn = O.coarse_to_fine(a, b, L, r, want_mask=True)
    uc, vc, mc = c_oracle.coarse_to_fine(a, b, L, r, want_mask=True)
    assert np.array_equal(un, uc) and np.array_equal(vn, vc) and np.array_equal(mn, mc)


@pytest.mark.parametrize("shape", [(7, 10), (1, 9), (9, 1), (2, 2), (31, 17), (1, 1), (3, 2)])
def test_pyramid_primitives_edge_shapes(c_oracle, shape):
    rng = np.random.default_rng(3)
    x = rng.standard_normal(shape)
    assert np.array_equal(c_oracle.pyr_down(x), O.pyr_down(x))
    assert np.array_equal(c_oracle.pyr_up(x), O.pyr_up(x))
    fu, fv = rng.standard_normal(shape) * 4, rng.standard_normal(shape) * 4
    (m1x, m1y), (m2x, m2y) = O.build_maps(fu, fv)
    w1 = O.remap_q5(x, np.broadcast_to(m1x, shape), np.broadcast_to(m1y, shape))
    w2 = O.remap_q5(x, np.broadcast_to(m2x, shape), np.broadcast_to(m2y, shape))
    assert np.array_equal(c_oracle.warp(x, fu, fv, +1), w1)
    assert np.array_equal(c_oracle.warp(x, fu, fv, -1), w2)


def test_pyr_up_preserves_constant_and_is_not_doubled():
    # a10: flow VALUES are not multiplied by 2 by the upsample
    x = np.full((5, 6), 3.25)
    assert np.array_equal(O.pyr_up(x), np.full((10, 12), 3.25))


def test_zero_divisor_gives_zero_and_mask(c_oracle):
    # OpenCV 2.4 divide semantics (:514,518): det == 0 -> flow 0; interior zeros exercise it
    rng = np.random.default_rng(5)
    a = rng.random((48, 64))
    b = rng.random((48, 64))
    a[:, 20:50] = 0.0
    b[:, 20:50] = 0.0
    for impl in (O, c_oracle):
        u, v, m = impl.lk_single(a, b, want_mask=True)
        assert np.isfinite(u).all() and np.isfinite(v).all()
        assert not m[5:40, 23:47].any()
        assert not u[5:40, 23:47].any()
        assert m[5:40, 3:17].all()


def test_not_divisible_raises(c_oracle):
    a = np.zeros((30, 40))
    for impl in (O, c_oracle):
        with pytest.raises(ValueError):
            impl.coarse_to_fine(a, a, 3)          # 30 % 4 != 0  (Q2)
        with pytest.raises(ValueError):
            impl.coarse_to_fine(a, np.zeros((30, 41)), 1)


def test_inputs_not_mutated(table_pair):
    a, b = table_pair
    a0, b0 = a.copy(), b.copy()
    O.coarse_to_fine(a, b, 3)
    assert np.array_equal(a, a0) and np.array_equal(b, b0)


def test_gesture_stats(c_oracle):
    u = np.zeros((10, 12))
    v = np.zeros((10, 12))
    u[2, 3] = 25.0
    v[7, 9] = -20.0            # val == 20 is NOT < 20 -> counted (:859)
    u[5, 5] = 19.999
    for impl in (O, c_oracle):
        s = impl.gesture_stats(u, v)
        assert s["count"] == 2 and s["sum_x"] == 12.0 and s["sum_y"] == 9.0 and not s["detected"]
