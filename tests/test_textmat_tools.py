"""Host-side I/O around the path: the reference's ``%lf`` text matrices
(saveMat, ParallelBasicPyr.cpp:193-204) and the headless command-line tools."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from of_b200_loader import ofb  # noqa: F401
import importlib

textmat = importlib.import_module("optical_flow_b200.textmat")


def test_save_mat_is_the_reference_format(tmp_path):
    m = np.array([[0.0, -1.5, 2.123456789], [1e-7, 123456.5, -0.0000004]])
    path = textmat.save_mat(m, str(tmp_path / "M"))
    assert path.endswith("M.txt")
    text = open(path).read()
    assert text == "0.000000 -1.500000 2.123457\n0.000000 123456.500000 -0.000000\n"   # C's %lf
    back = textmat.load_mat(path)
    assert back.shape == (2, 3)
    assert np.abs(back - m).max() <= 5e-7                         # print precision
    assert np.array_equal(textmat.load_mat(path, (2, 3)), back)   # fscanf-style read with known size
    with pytest.raises(ValueError):
        textmat.load_mat(path, (3, 3))


def test_golden_round_trip_within_harness_tolerance(tmp_path, kat_golden):
    g = kat_golden["UPyrTest"][:40, :64]
    back = textmat.load_mat(textmat.save_mat(g, str(tmp_path / "g")))
    assert textmat.compare(back, g) == (0, pytest.approx(0.0, abs=1e-6))
    bad, worst = textmat.compare(back + 2e-6, g)
    assert bad == g.size and worst > 1e-6


def test_load_mat_rejects_ragged(tmp_path):
    p = tmp_path / "r.txt"
    p.write_text("1.0 2.0\n3.0\n")
    with pytest.raises(ValueError):
        textmat.load_mat(str(p))


@pytest.mark.gpu
def test_of_pair_cli_reproduces_kat1(tmp_path, table_pair, kat_golden):
    a, b = table_pair
    np.save(tmp_path / "a.npy", a)
    np.save(tmp_path / "b.npy", b)
    gu = textmat.save_mat(kat_golden["U2Test"], str(tmp_path / "gu"))
    gv = textmat.save_mat(kat_golden["V2Test"], str(tmp_path / "gv"))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "of_pair.py"), str(tmp_path / "a.npy"),
                        str(tmp_path / "b.npy"), "--single", "--out", str(tmp_path / "o"),
                        "--check-u", gu, "--check-v", gv], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "448, 640, Max layer = 1" in r.stdout
    assert "0 entries differ" in r.stdout
    u = textmat.load_mat(str(tmp_path / "o_u.txt"), (448, 640))
    # both sides went through %lf: two roundings to 6 decimals can differ by one last digit
    assert np.abs(u - kat_golden["U2Test"]).max() <= 1.5e-6


@pytest.mark.gpu
def test_of_pair_cli_timing_json_and_level_dump(tmp_path, table_pair):
    """per-stage device timings as JSON and per-level triage planes."""
    a, b = table_pair
    np.save(tmp_path / "a.npy", a[:256, :320])
    np.save(tmp_path / "b.npy", b[:256, :320])
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "of_pair.py"), str(tmp_path / "a.npy"),
                        str(tmp_path / "b.npy"), "--levels", "3", "--format", "npy", "--out", str(tmp_path / "o"),
                        "--timing-json", str(tmp_path / "t.json"), "--dump-levels", str(tmp_path / "lv")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    t = json.load(open(tmp_path / "t.json"))
    assert [(x["stage"], x["level"]) for x in t["stages"]] == [
        ("pyrDown", 1), ("pyrDown", 1), ("pyrDown", 2), ("lk_level", 2), ("lk_level", 1), ("lk_level", 0)]
    assert all(x["ms"] > 0 for x in t["stages"])
    from oracle import lk_oracle as O
    pa = O.gaussian_pyramid(a[:256, :320], 3)
    for l in (1, 2):
        assert np.array_equal(np.load(tmp_path / ("lv_L%d_frame1.npy" % l)), pa[l])
        assert np.load(tmp_path / ("lv_L%d_u.npy" % l)).shape == pa[l].shape


@pytest.mark.gpu
def test_of_video_cli_stats(tmp_path, c_oracle):
    synth = importlib.import_module("optical_flow_b200.synth")
    fr = synth.synth_frames(4, 96, 128, seed=3, gesture=True) * 40.0
    np.save(tmp_path / "f.npy", fr)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "of_video.py"), str(tmp_path / "f.npy"),
                        "--levels", "3", "--radius", "2", "--mag-thresh", "0.5", "--count-thresh", "10",
                        "--flow-out", str(tmp_path / "fl")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = [json.loads(x) for x in r.stdout.strip().splitlines()]
    assert [x["pair"] for x in lines] == [0, 1, 2]
    u = np.load(tmp_path / "fl_u.npy")
    v = np.load(tmp_path / "fl_v.npy")
    for k, x in enumerate(lines):
        uo, vo = c_oracle.coarse_to_fine(fr[k], fr[k + 1], 3, 2)
        assert max(np.abs(u[k] - uo).max(), np.abs(v[k] - vo).max()) <= 1e-3
        sel = ~(np.sqrt(u[k] ** 2 + v[k] ** 2) < 0.5)
        assert x["count"] == int(sel.sum())
        assert x["detected"] == (x["count"] > 10)
        if x["count"]:
            yy, xx = np.nonzero(sel)
            assert x["cx"] == pytest.approx(xx.mean()) and x["cy"] == pytest.approx(yy.mean())
