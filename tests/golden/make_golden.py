#!/usr/bin/env python
"""Generate the committed fixtures under tests/golden/ (run in the BUILD container only).

Needs /root/reference (read-only) and cv2; neither exists at test time on the GPU
box, which is why the outputs are committed.  What it does:

1. pins the numpy oracle's OpenCV restatements (filter2D / pyrDown / pyrUp / remap)
   against cv2 on random inputs of awkward shapes  -> printed max-abs (must be 0);
2. converts the reference's golden text matrices (U2Test/V2Test, UPyrTest/VPyrTest,
   ``%lf`` = 6 decimals) into int32 micro-pixel arrays  -> kat_golden.npz;
3. stores the decoded inputs of every BASELINE config that is a file:
   table crop (``ParallelBasicPyr.cpp:720-723``), hand pair (EXIF ignored, as
   OpenCV 2.4 did), all decodable pacman.mp4 frames as 8-bit gray;
4. checks the oracle against the goldens (<= 5e-7) and writes oracle outputs for the
   hand pair / pacman pairs as float64 checksums  -> oracle_digest.json.

Usage:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import lk_oracle as O  # noqa: E402

REF = "/root/reference"
GRAY = cv2.IMREAD_GRAYSCALE | cv2.IMREAD_IGNORE_ORIENTATION


def pin_primitives():
    rng = np.random.default_rng(7)
    worst = {}
    for (h, w) in [(7, 10), (8, 8), (16, 2), (2, 16), (31, 17), (64, 96), (1, 9), (9, 1), (2, 2)]:
        a = rng.standard_normal((h, w))
        b = rng.standard_normal((h, w))
        # filter2D restatement
        kx = np.array([[-1, 1], [-1, 1]], np.float64)
        ky = np.array([[-1, -1], [1, 1]], np.float64)
        k1 = np.ones((2, 2))
        fx = cv2.filter2D(a, -1, kx) + cv2.filter2D(b, -1, kx)
        fy = cv2.filter2D(a, -1, ky) + cv2.filter2D(b, -1, ky)
        ft = cv2.filter2D(a, -1, -k1) + cv2.filter2D(b, -1, k1)
        ofx, ofy, oft = O.gradients(a, b)
        for n, x, y in (("fx", fx, ofx), ("fy", fy, ofy), ("ft", ft, oft)):
            worst[n] = max(worst.get(n, 0.0), float(np.abs(x - y).max()))
        # pyrDown / pyrUp
        worst["pyrDown"] = max(worst.get("pyrDown", 0.0), float(np.abs(cv2.pyrDown(a) - O.pyr_down(a)).max()))
        worst["pyrUp"] = max(worst.get("pyrUp", 0.0), float(np.abs(cv2.pyrUp(a) - O.pyr_up(a)).max()))
        # remap with large, partly out-of-image displacements
        fu = rng.standard_normal((h, w)) * 3.0
        fv = rng.standard_normal((h, w)) * 3.0
        (m1x, m1y), (m2x, m2y) = O.build_maps(fu, fv)
        for mx, my in ((m1x, m1y), (m2x, m2y)):
            mp = np.dstack([np.broadcast_to(mx, (h, w)), np.broadcast_to(my, (h, w))]).astype(np.float32)
            c = cv2.remap(a, mp, None, cv2.INTER_LINEAR)
            worst["remap"] = max(worst.get("remap", 0.0), float(np.abs(c - O.remap_q5(a, mp[..., 0], mp[..., 1])).max()))
    return worst


def read_txt(name):
    m = np.loadtxt(os.path.join(REF, name))
    q = np.rint(m * 1e6)
    assert np.abs(q / 1e6 - m).max() < 1e-9 and np.abs(q).max() < 2 ** 31
    return q.astype(np.int32)


def digest(x):
    return hashlib.sha256(np.ascontiguousarray(x, dtype=np.float64).tobytes()).hexdigest()


def main():
    worst = pin_primitives()
    print("oracle primitives vs cv2 %s (max-abs):" % cv2.__version__, worst)

    # --- KAT inputs and goldens ------------------------------------------------
    t1 = cv2.imread(os.path.join(REF, "testdata/table1.jpg"), GRAY)[0:448, 0:640]
    t2 = cv2.imread(os.path.join(REF, "testdata/table2.jpg"), GRAY)[0:448, 0:640]
    np.savez_compressed(os.path.join(HERE, "table_pair.npz"), img1=t1, img2=t2)
    gold = {n: read_txt(n + ".txt") for n in ("U2Test", "V2Test", "UPyrTest", "VPyrTest")}
    np.savez_compressed(os.path.join(HERE, "kat_golden.npz"), **gold)

    a = t1.astype(np.float64) / 255.0
    b = t2.astype(np.float64) / 255.0
    u1, v1 = O.lk_single(a, b)
    u7, v7 = O.coarse_to_fine(a, b, O.get_max_layer(640, 448))
    err = {
        "U2Test": float(np.abs(u1 - gold["U2Test"] / 1e6).max()),
        "V2Test": float(np.abs(v1 - gold["V2Test"] / 1e6).max()),
        "UPyrTest": float(np.abs(u7 - gold["UPyrTest"] / 1e6).max()),
        "VPyrTest": float(np.abs(v7 - gold["VPyrTest"] / 1e6).max()),
    }
    print("oracle vs reference goldens (max-abs):", err)
    assert max(err.values()) <= 5.0e-7 + 1e-12, err

    # --- hand pair (config 0) and pacman stream (config 2) -----------------------
    h1 = cv2.imread(os.path.join(REF, "testdata/hand1.jpg"), GRAY)
    h2 = cv2.imread(os.path.join(REF, "testdata/hand2.jpg"), GRAY)
    assert h1.shape == (2448, 3264)
    cv2.imwrite(os.path.join(HERE, "hand1_gray.png"), h1, [cv2.IMWRITE_PNG_COMPRESSION, 9])
    cv2.imwrite(os.path.join(HERE, "hand2_gray.png"), h2, [cv2.IMWRITE_PNG_COMPRESSION, 9])
    cap = cv2.VideoCapture(os.path.join(REF, "pacman.mp4"))
    frames = []
    while True:
        ok, f = cap.read()
        if not ok:
            break
        frames.append(cv2.cvtColor(f, cv2.COLOR_BGR2GRAY))
    frames = np.stack(frames)
    assert frames.shape == (195, 160, 240), frames.shape
    np.savez_compressed(os.path.join(HERE, "pacman_gray.npz"), frames=frames)

    info = {
        "cv2": cv2.__version__,
        "primitives_vs_cv2_maxabs": worst,
        "oracle_vs_reference_goldens_maxabs": err,
        "table_L1": {"u": digest(u1), "v": digest(v1)},
        "table_L7": {"u": digest(u7), "v": digest(v7)},
    }
    with open(os.path.join(HERE, "oracle_digest.json"), "w") as f:
        json.dump(info, f, indent=1, sort_keys=True)
    print("wrote fixtures to", HERE)


if __name__ == "__main__":
    main()
