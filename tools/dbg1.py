import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
from of_b200_loader import ofb
rng = np.random.default_rng(0)
a = rng.random((128, 256)); b = rng.random((128, 256))
r = int(os.environ.get("R", "7"))
with ofb.Context(0, win_radius=r) as ctx:
    u, v = ctx.coarseToFineEstimation(a, b, 2)[:2]
    print("ok", float(u.sum()))
