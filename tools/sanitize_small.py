"""Smallest run that touches every kernel family once, for compute-sanitizer (no torch: ctypes + numpy only, so the tool has
little to instrument).    compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from flink_jpmml_b200 import capi, pmmlgen as g

eng = capi.Engine(0)
cases = [("kmeans", g.config_c2(k=16, F=32)[0], 32, 0.05), ("gbt", g.gbt_pmml(T=24, depth=8, F=20, seed=1)[0], 20, 0.1),
         ("rf", g.rf_pmml(T=16, depth=6, F=16, n_labels=3, seed=2)[0], 16, 0.0), ("regression", g.regression_pmml(F=96, C=5, seed=3)[0], 96, 0.0),
         ("mlp", g.nn_pmml(n_in=16, hidden=(32, 32), n_out=4, seed=4)[0], 16, 0.0)]
for name, pm, F, pmiss in cases:
    m = eng.compile(pm)
    for n in (7, 1500):
        X, pres, pb = g.events(n, F, 5, pmiss)
        r = eng.score(m, X, pb if pmiss else None)
        print(name, n, "ok rows", int((r["status"] == 0).sum()), flush=True)
# the two-kernel throughput path of the forests wants >= 148 tiles
X, pres, pb = g.events(148 * 640 + 5, 20, 6, 0.1)
m = eng.compile(cases[1][1]); r = eng.score(m, X, pb); print("gbt throughput path ok rows", int((r["status"] == 0).sum()))
eng.close()
