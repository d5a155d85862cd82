"""The reference's on-disk matrix format (``saveMat``, ParallelBasicPyr.cpp:193-204): one text
line per row, values printed with C's ``%lf`` (six decimals), separated by single spaces, file
name = given stem + ``.txt``.  The reference harnesses compare such files value by value within
1e-6 (ParallelBasicPyr.cpp:787,818; serialOrig.cpp:588,618); ``compare`` does the same.
"""
import numpy as np


def save_mat(m, stem: str) -> str:
    """Writes ``m`` (2-D) as ``stem + '.txt'`` exactly as the reference's saveMat prints it."""
    a = np.asarray(m, dtype=np.float64)
    if a.ndim != 2:
        raise ValueError("save_mat needs a 2-D array")
    path = stem + ".txt"
    with open(path, "w") as f:
        for row in a:
            f.write(" ".join("%f" % x for x in row))   # Python's %f == C's %lf
            f.write("\n")
    return path


def load_mat(path: str, shape=None) -> np.ndarray:
    """Reads a matrix written by saveMat (``%lf`` tokens, rows on lines).  The reference's own
    readers ``fscanf`` a known rows x cols (ParallelBasicPyr.cpp:771-789): pass ``shape`` to do
    the same, otherwise the shape comes from the line structure."""
    with open(path) as f:
        if shape is None:
            rows = [np.array(line.split(), dtype=np.float64) for line in f if line.strip()]
            if not rows:
                return np.zeros((0, 0))
            if any(len(r) != len(rows[0]) for r in rows):
                raise ValueError("%s: ragged rows" % path)
            return np.stack(rows)
        vals = np.array(f.read().split(), dtype=np.float64)
    h, w = shape
    if vals.size < h * w:
        raise ValueError("%s: %d values, need %d x %d" % (path, vals.size, h, w))
    return vals[: h * w].reshape(h, w)


def compare(m, golden, tol: float = 1e-6):
    """The harness check: number of entries with |m - golden| > tol, and the maximum deviation."""
    d = np.abs(np.asarray(m, np.float64) - np.asarray(golden, np.float64))
    return int((d > tol).sum()), float(d.max()) if d.size else 0.0
