"""One host core's copy rate into pinned memory (what the operator's task thread does per event), as plain copies and as
double -> float narrowing copies, in micro-batch sized blocks.  Prints one JSON line."""
import json, time
import numpy as np, torch

N, F, MB = 1 << 20, 100, 16384
X = np.random.default_rng(1).standard_normal((N, F))
pin64 = torch.empty((2, MB, F), dtype=torch.float64).pin_memory().numpy()
pin32 = torch.empty((2, MB, F), dtype=torch.float32).pin_memory().numpy()
page64 = np.empty((2, MB, F))
out = {}
for name, dst in (("f64_to_pinned", pin64), ("f64_to_pageable", page64), ("f64_to_f32_pinned", pin32)):
    best = 0.0
    for rep in range(3):
        t = time.perf_counter()
        for k, lo in enumerate(range(0, N, MB)):
            np.copyto(dst[k & 1], X[lo:lo + MB], casting="same_kind")
        dt = time.perf_counter() - t
        best = max(best, N / dt)
    out[name] = {"events_per_s": best, "read_GBs": best * F * 8 / 1e9}
print(json.dumps(out))
