#!/usr/bin/env python
"""Headless replacement of the reference's image-pair harness mains (ParallelBasicPyr.cpp:717-836,
serialOrig.cpp:538-635): image pair -> gray / 255 -> optional crop -> coarse-to-fine flow on the
GPU -> u, v written in the reference's ``%lf`` text format (saveMat) or as .npy, and optionally
compared with golden text files at the harness tolerance 1e-6.

  python tools/of_pair.py img1 img2 --out flow [--levels N | --single] [--crop X Y W H]
                          [--radius R] [--format txt|npy] [--check-u U.txt --check-v V.txt]

Inputs: .npy (2-D float64, used as is), or any image cv2 / PIL can read (converted to gray and
divided by 255 like ParallelBasicPyr.cpp:722-730).
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from of_b200_loader import ofb  # noqa: E402
import importlib  # noqa: E402

textmat = importlib.import_module("optical_flow_b200.textmat")


def read_gray(path):
    if path.endswith(".npy"):
        return np.ascontiguousarray(np.load(path), dtype=np.float64)
    try:
        import cv2
        g = cv2.imread(path, cv2.IMREAD_GRAYSCALE | cv2.IMREAD_IGNORE_ORIENTATION)   # EXIF-blind like OpenCV 2.4
        if g is None:
            raise IOError("cannot read " + path)
    except ImportError:
        from PIL import Image
        g = np.array(Image.open(path).convert("L"))
    return g.astype(np.float64) / 255.0


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("img1")
    ap.add_argument("img2")
    ap.add_argument("--out", default="flow", help="output stem: <out>_u.txt / <out>_v.txt (or .npy)")
    ap.add_argument("--levels", type=int, default=0, help="pyramid levels; 0 = getMaxLayer(img1)")
    ap.add_argument("--single", action="store_true", help="getLucasKanadeOpticalFlow only (one level)")
    ap.add_argument("--crop", type=int, nargs=4, metavar=("X", "Y", "W", "H"), help="Rect like ParallelBasicPyr.cpp:723")
    ap.add_argument("--radius", type=int, default=1, help="window radius (reference: 1 = 3x3)")
    ap.add_argument("--format", choices=("txt", "npy"), default="txt")
    ap.add_argument("--check-u")
    ap.add_argument("--check-v")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--timing-json", help="write the device time of every stage (pyrDown / fused level kernel per "
                                          "pyramid level) as JSON, like the reference's per-stage timers")
    ap.add_argument("--dump-levels", metavar="STEM", help="parity triage: save the pyramid images and the accumulated "
                                                          "flow of every level >= 1 as <STEM>_L<l>_<plane>.npy")
    a = ap.parse_args(argv)

    i1, i2 = read_gray(a.img1), read_gray(a.img2)
    if a.crop:
        x, y, w, h = a.crop
        i1, i2 = i1[y:y + h, x:x + w], i2[y:y + h, x:x + w]
    i1, i2 = np.ascontiguousarray(i1), np.ascontiguousarray(i2)
    levels = 1 if a.single else (a.levels or ofb.getMaxLayer(i1))
    print("%d, %d, Max layer = %d" % (i1.shape[0], i1.shape[1], levels))       # the harness's first line
    with ofb.Context(a.device, win_radius=a.radius) as ctx:
        if a.timing_json:
            ctx.set_timing(True)
        u, v = ctx.getLucasKanadeOpticalFlow(i1, i2) if a.single else ctx.coarseToFineEstimation(i1, i2, levels)
        if a.timing_json:
            import json
            stages = [{"stage": "pyrDown" if k == 0 else "lk_level", "level": l, "ms": ms} for k, l, ms in ctx.get_timing()]
            json.dump({"width": int(i1.shape[1]), "height": int(i1.shape[0]), "levels": levels, "win_radius": a.radius,
                       "stages": stages, "device_ms_total": sum(x["ms"] for x in stages)}, open(a.timing_json, "w"), indent=1)
            print("wrote", a.timing_json)
        if a.dump_levels and not a.single:
            for l in range(1, levels):
                for plane in ("frame1", "frame2", "u", "v"):
                    np.save("%s_L%d_%s.npy" % (a.dump_levels, l, plane), ctx.get_level(plane, l))
            print("wrote %s_L*_*.npy" % a.dump_levels)
    if a.format == "txt":
        print("wrote", textmat.save_mat(u, a.out + "_u"), textmat.save_mat(v, a.out + "_v"))
    else:
        np.save(a.out + "_u.npy", u)
        np.save(a.out + "_v.npy", v)
        print("wrote %s_u.npy %s_v.npy" % (a.out, a.out))
    rc = 0
    for name, got, path in (("u", u, a.check_u), ("v", v, a.check_v)):
        if path:
            bad, worst = textmat.compare(got, textmat.load_mat(path, got.shape))
            print("%s vs %s: %d entries differ by more than 1e-6 (max %.3g)" % (name, path, bad, worst))
            rc |= bad != 0
    return rc


if __name__ == "__main__":
    sys.exit(main())
