"""The header-only C++ shim (include/of_b200_shim.hpp) keeps the reference's entry-point names
(getMaxLayer / getLucasKanadeOpticalFlow / coarseToFineEstimation, ParallelVideoPyr.cpp:545,327,578).
CPU: it compiles and links against libof_b200.so.  GPU: a C++ harness shaped like the reference's
ParallelBasicPyr.cpp main() reproduces the reference's golden files through it."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "shim_check.cpp")


def build_shim_check(ofb, out_dir):
    ofb.lib()                                                 # raises if the library is not built
    exe = os.path.join(str(out_dir), "shim_check")
    libdir = os.path.dirname(ofb.LIB_PATH)
    cmd = ["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe,
           "-L", libdir, "-lof_b200", "-Wl,-rpath," + libdir]
    subprocess.check_call(cmd)
    return exe


def test_shim_compiles_and_links(ofb, tmp_path):
    exe = build_shim_check(ofb, tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


@pytest.mark.gpu
def test_shim_reproduces_reference_goldens(ofb, tmp_path, table_pair, kat_golden):
    exe = build_shim_check(ofb, tmp_path)
    a, b = table_pair
    pa, pb = str(tmp_path / "a.f64"), str(tmp_path / "b.f64")
    a.tofile(pa)
    b.tofile(pb)
    r = subprocess.run([exe, "640", "448", pa, pb, str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "448, 640, Max layer = 7" in r.stdout              # SerialPyr.cpp:234
    rd = lambda n: np.fromfile(str(tmp_path / ("o_%s.f64" % n))).reshape(448, 640)
    assert np.abs(rd("u") - kat_golden["UPyrTest"]).max() <= 1e-6      # ParallelBasicPyr.cpp:787
    assert np.abs(rd("v") - kat_golden["VPyrTest"]).max() <= 1e-6      # ParallelBasicPyr.cpp:818
    assert np.abs(rd("u1") - kat_golden["U2Test"]).max() <= 1e-6       # serialOrig.cpp:588
    assert np.abs(rd("v1") - kat_golden["V2Test"]).max() <= 1e-6       # serialOrig.cpp:618


@pytest.mark.gpu
def test_shim_rejects_non_divisible_size(ofb, tmp_path):
    exe = build_shim_check(ofb, tmp_path)
    a = np.random.default_rng(0).random((30, 50))
    pa = str(tmp_path / "a.f64")
    a.tofile(pa)
    r = subprocess.run([exe, "50", "30", pa, pa, str(tmp_path / "o"), "3"], capture_output=True, text=True)
    assert r.returncode == 1 and "not divisible" in r.stderr
