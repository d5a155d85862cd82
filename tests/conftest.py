import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def table_pair():
    """KAT input: table1/table2 -> gray -> Rect(0,0,640,448) -> /255 (ParallelBasicPyr.cpp:720-730)."""
    d = np.load(os.path.join(GOLDEN, "table_pair.npz"))
    return d["img1"].astype(np.float64) / 255.0, d["img2"].astype(np.float64) / 255.0


@pytest.fixture(scope="session")
def kat_golden():
    """The reference's golden matrices, stored as int32 micro-pixels (they are %lf text)."""
    d = np.load(os.path.join(GOLDEN, "kat_golden.npz"))
    return {k: d[k].astype(np.float64) / 1e6 for k in d.files}


@pytest.fixture(scope="session")
def c_oracle():
    from oracle import c_oracle as CO
    CO.build()
    return CO


@pytest.fixture(scope="session")
def ofb():
    from of_b200_loader import ofb as m
    return m


def load_gray_png(name):
    path = os.path.join(GOLDEN, name)
    try:
        import cv2
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        if img is not None:
            return img
    except ImportError:
        pass
    from PIL import Image
    return np.array(Image.open(path))
