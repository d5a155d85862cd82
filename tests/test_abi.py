"""CPU suite: the C-ABI library loads and exports every symbol include/of_b200.h declares;
host-only entry points behave; nothing computes without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "of_b200.h")).read()
    return sorted(set(re.findall(r"OF_API[^;(]*?\b(of_\w+)\s*\(", txt)))


def test_header_symbols_all_exported(ofb):
    names = declared_symbols()
    assert len(names) >= 20
    L = ofb.lib()
    for n in names:
        assert hasattr(L, n), "libof_b200.so does not export %s" % n
    assert sorted(ofb.EXPORTS) == names
    assert L.of_abi_version() == 1


def test_get_max_layer_host_function(ofb):
    # getMaxLayer, ParallelVideoPyr.cpp:545-564
    for (w, h), want in {(640, 448): 7, (1080, 1080): 4, (1920, 1080): 4, (3840, 2160): 5,
                         (3264, 2448): 5, (240, 160): 5, (320, 180): 3, (641, 448): 1}.items():
        assert ofb.getMaxLayer(w, h) == want
    assert ofb.getMaxLayer(0, 5) == 0


def test_params_default_are_the_reference_constants(ofb):
    p = ofb.default_params()
    assert (p.win_radius, p.det_eps, p.write_back_warped, p.mag_thresh, p.count_thresh) == (1, 0.0, 0, 20.0, 500)


def test_no_cpu_fallback_without_gpu(ofb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(ofb.OfError) as ei:
        ofb.Context(0)
    assert ei.value.status == ofb.OF_E_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "optical-flow_b200")
    for dp, _, fs in os.walk(pkg):
        if os.path.basename(dp) == "build":
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                txt = open(os.path.join(dp, f)).read()
                assert "lk_oracle" not in txt and "c_oracle" not in txt and "live_oracle" not in txt, f
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M), f


def test_mgpu_partition_host_function(ofb):
    """of_mgpu_partition (host arithmetic): contiguous blocks that cover every pair once, sizes within one."""
    for n in (0, 1, 7, 63, 194):
        for nd in (1, 2, 3, 4, 8):
            got, sizes = [], []
            for i in range(nd):
                a, b = ofb.MultiGpu.partition(n, nd, i)
                assert 0 <= a <= b <= n
                got += list(range(a, b))
                sizes.append(b - a)
            assert got == list(range(n)) and max(sizes) - min(sizes) <= 1
    with pytest.raises(ofb.BadArgError):
        ofb.MultiGpu.partition(4, 2, 2)


def test_live_params_default_are_the_reference_constants(ofb):
    # ParallelVideoPyr.cpp:728-729 (DIFF_THRESH 10, LEARNING_RATE 0.3), :782 (1080 x 1080), :790 (9, 1.5), :812-813 (2)
    lp = ofb.OfLiveParams()
    ofb.lib().of_live_params_default(C.byref(lp))
    assert (lp.work_width, lp.work_height, lp.blur_ksize, lp.blur_sigma, lp.learning_rate, lp.diff_thresh,
            lp.morph_iterations, lp.n_levels) == (1080, 1080, 9, 1.5, 0.3, 10, 2, 0)
    assert lp.coeff_bits == 14                       # OpenCV 2.4's cvtColor table, the reference's library


def test_static_scratch_limit_is_declared(ofb):
    txt = open(os.path.join(ROOT, "include", "of_b200.h")).read()
    assert re.search(r"#define\s+OF_STATIC_SCRATCH_BYTES\s+\(32u << 20\)", txt)
