import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from test_engine_gpu import run_case
rows, outs, got, ref = run_case(1, 480, 640, 15)
for nm, g, r in (("cls", got[1], ref[1]), ("mask", got[2], ref[2])):
    d = np.abs(g - r)
    print(nm, "max_abs", d.max(), "p99.99", np.quantile(d, 0.9999), "p99.9", np.quantile(d, 0.999), "mean", d.mean())
    lg = np.log(g / (1 - g + 1e-12) + 1e-30); lr = np.log(r / (1 - r + 1e-12) + 1e-30)
    m = (r > 1e-4) & (r < 1 - 1e-4)
    print(nm, "logit abs err max", np.abs(lg - lr)[m].max(), "p99.9", np.quantile(np.abs(lg - lr)[m], 0.999))
for thr in (0.5, 0.05):
    for band in (1e-2, 2e-2, 5e-2, 1e-1):
        safe = np.abs(ref[1] - thr) > thr * band
        mism = ((got[1] > thr) != (ref[1] > thr)) & safe
        print("thr", thr, "band", band, "mismatches", int(mism.sum()), "of", int(safe.sum()), "total flips", int(((got[1] > thr) != (ref[1] > thr)).sum()))
