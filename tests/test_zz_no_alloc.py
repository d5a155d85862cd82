"""GPU test (-m gpu) of of_ctf_f64_dev_static, the allocation-free variant of of_ctf_f64_dev: scratch in a static
device array inside the library (include/of_b200.h; no counterpart in the reference).
Called through the C ABI (ctypes) and compared with the existing entry point on two calls
in a row with different inputs (the array keeps the previous call's contents).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FLOW_TOL = 1e-3      # px, the r > 1 bar of tests/test_gpu_parity.py; r = 1 is bit-exact there and here


def _pair(h, w, seed):
    rng = np.random.default_rng(seed)
    base = rng.random((h + 8, w + 8))
    a = base[4:4 + h, 4:4 + w].copy()
    b = (0.7 * base[4:4 + h, 5:5 + w] + 0.3 * base[5:5 + h, 4:4 + w]) + rng.normal(0, 0.01, (h, w))
    return a, b


@pytest.mark.parametrize("r", [1, 7])
def test_static_scratch_entry_point_matches_the_arena_one(ofb, r):
    import torch
    assert torch.cuda.is_available()
    L = ofb.lib()
    h, w, levels = 272, 480, 4
    with ofb.Context(0, win_radius=r) as ctx:
        for call, seed in enumerate((11, 12)):                 # two calls in a row, different frames
            a, b = _pair(h, w, seed + r)
            d = torch.from_numpy(np.stack([a, b])).cuda()
            u0 = torch.empty((h, w), dtype=torch.float64, device="cuda")
            v0 = torch.empty_like(u0)
            ctx.coarseToFineEstimation_dev(d[0], d[1], levels, u0, v0)            # existing entry point
            u1 = torch.full((h, w), 7.0, dtype=torch.float64, device="cuda")
            v1 = torch.full((h, w), 7.0, dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            st = L.of_ctf_f64_dev_static(ctx._h, d[0].data_ptr(), d[1].data_ptr(), w, h, w * 8, levels,
                                         u1.data_ptr(), v1.data_ptr(), w * 8, C.c_void_p(1))   # legacy default stream
            assert st == ofb.OF_OK, (L.of_last_error_string(ctx._h) or b"").decode()
            torch.cuda.synchronize()
            if r == 1:
                assert torch.equal(u0, u1) and torch.equal(v0, v1), "call %d" % call
            else:
                err = max(float((u0 - u1).abs().max()), float((v0 - v1).abs().max()))
                assert err <= FLOW_TOL, "call %d: %g px" % (call, err)
            assert float(u1.abs().max()) > 0.0


def test_static_scratch_rejects_what_does_not_fit(ofb):
    import torch
    L = ofb.lib()
    h, w = 2160, 3840                                           # 4K, 4 levels: about 88 MB of scratch > 32 MiB
    z = torch.zeros((2, h, w), dtype=torch.float64, device="cuda")
    u = torch.empty((h, w), dtype=torch.float64, device="cuda")
    with ofb.Context(0) as ctx:
        st = L.of_ctf_f64_dev_static(ctx._h, z[0].data_ptr(), z[1].data_ptr(), w, h, w * 8, 4, u.data_ptr(),
                                     u.data_ptr(), w * 8, None)
        assert st == ofb.OF_E_BADARG
        msg = (L.of_last_error_string(ctx._h) or b"").decode()
        assert "static array holds" in msg and str(32 << 20) in msg
        st = L.of_ctf_f64_dev_static(ctx._h, z[0].data_ptr(), z[1].data_ptr(), w, h, w * 8, 1, u.data_ptr(),
                                     z[1].data_ptr(), w * 8, None)      # one level needs no scratch at all
        assert st == ofb.OF_OK
        torch.cuda.synchronize()
        assert not bool(u.any())                                # identical (zero) frames: zero flow
