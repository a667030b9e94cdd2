"""Operator-level throughput through the host mirror (pinned micro-batches, submit/wait double buffering):
  C4  fixed-model operator (quickEvaluate shape), 2^20 events
  C5  dynamic serving: 32 ModelIds (16 RF 200x8 F=64 + 16 MLP 64-256-256-8), events addressed uniformly,
      every 50 ms a random id gets Del then Add(version + 1) while scoring runs (BASELINE.json config 5)
One JSON line per config.  Host-side packing runs on ONE task thread, as one Flink subtask would.
bench.py imports c5_dynamic() / one_core_copy() for its `operator` record."""
import json, os, sys, tempfile, time, uuid
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from flink_jpmml_b200 import api, capi, pmmlgen as g


def one_core_copy(F=100, rows=1 << 20, block=16384):
    """What one host core copies into pinned memory in micro-batch sized blocks (events/s at F doubles per event): the
    bound of any operator whose task thread has to place every event in a pinned batch."""
    import torch
    X = np.random.default_rng(1).standard_normal((rows, F))
    pin = torch.empty((2, block, F), dtype=torch.float64).pin_memory().numpy()
    best = 0.0
    for _ in range(3):
        t = time.perf_counter()
        for k, lo in enumerate(range(0, rows, block)):
            np.copyto(pin[k & 1], X[lo:lo + block])
        best = max(best, rows / (time.perf_counter() - t))
    return {"events_per_s": best, "GBs": best * F * 8 / 1e9, "F": F}


def c5_models(tmp):
    names, files = [], []
    for k in range(32):
        pmk = g.rf_pmml(seed=1000 + k)[0] if k < 16 else g.nn_pmml(seed=1000 + k)[0]
        p = os.path.join(tmp, f"m{k}.pmml"); open(p, "wb").write(pmk)
        names.append(str(uuid.UUID(int=7000 + k))); files.append(p)
    return names, files


def c5_dynamic(eng, names, files, swap_s, passes=4, N=1 << 21, CH=32768, seed=20261106 + 5):
    """`passes` passes over N events addressed uniformly to the 32 ids; every swap_s seconds one id gets Del + Add(version + 1)."""
    rng = np.random.default_rng(seed)
    version = {k: 1 for k in range(32)}
    tag = uuid.uuid4().int & 0xFFFF                                   # fresh ids per call: the engine's model table is shared
    nm = [str(uuid.UUID(int=(tag << 32) + k)) for k in range(32)]
    op = api.EvaluationCoFunction(eng, max_batch=65536)
    for k in range(32): op.processElement2(api.AddMessage(nm[k], 1, files[k]))
    X = rng.standard_normal((N, 64)); idx = rng.integers(0, 32, N).astype(np.int32)
    ids = lambda: [f"{nm[k]}_{version[k]}" for k in range(32)]
    for _ in range(3):                                                              # warm-up: lazy loads of all 32 models, pinned buckets of two generations
        op.processDense1(ids(), idx[:4 * CH], X[:4 * CH]); op.flush(); op.collect_arrays()
    swaps, emitted, empties = 0, 0, 0
    t0 = time.perf_counter(); last = t0
    for _ in range(passes):
        for lo in range(0, N, CH):
            now = time.perf_counter()
            if now - last >= swap_s:
                k = int(rng.integers(32)); op.processElement2(api.DelMessage(nm[k], version[k])); version[k] += 1
                op.processElement2(api.AddMessage(nm[k], version[k], files[k])); swaps += 1; last = now
            op.processDense1(ids(), idx[lo:lo + CH], X[lo:lo + CH])
            sc, st = op.collect_arrays(); emitted += len(st); empties += int((st != 0).sum())   # a task thread emits as it goes
    op.flush(); dt = time.perf_counter() - t0
    sc, st = op.collect_arrays(); emitted += len(st); empties += int((st != 0).sum())
    parked = op.parkedEvents; flush_s = op.flushSeconds
    op.close()
    for k in range(32): eng.table_del(nm[k], version[k])
    return {"events_per_s": passes * N / dt, "hot_swaps": swaps, "emitted": emitted, "empty_scores": empties, "parked_events": parked, "seconds": dt, "seconds_in_flush": flush_s}


if __name__ == "__main__":
    eng = capi.Engine(0)
    tmp = tempfile.mkdtemp()
    rng = np.random.default_rng(20261106 + 5)
    print(json.dumps({"config": "one host core copying 800-byte events into pinned micro-batches (the operator's bound)", **one_core_copy()}), flush=True)
    # ---------------- C4 through EvaluationFunction
    pm, _ = g.gbt_pmml()
    path = os.path.join(tmp, "c4.pmml"); open(path, "wb").write(pm)
    X = rng.standard_normal((1 << 20, 100))
    for mb in (16384, 262144):
        fn = api.EvaluationFunction(eng, path, max_batch=mb)
        fn.processDense(X[:mb]); fn.flush(); fn.collect()
        t = time.perf_counter(); fn.processDense(X); fn.flush(); dt = time.perf_counter() - t
        n_out = len(fn.collect()); fn.close()
        print(json.dumps({"config": "C4 fixed-model operator (dense events, 1 task thread)", "max_batch": mb, "events_per_s": len(X) / dt, "emitted": n_out}), flush=True)
    # ---------------- C5 dynamic serving
    names, files = c5_models(tmp)
    for swap_s, what in ((float(os.environ.get("FJP_BENCH_SWAP_S", "0.050")), "Del/Add(version+1) every 50 ms"), (1e9, "no control traffic")):
        r = c5_dynamic(eng, names, files, swap_s)
        print(json.dumps({"config": f"C5 dynamic serving: 32 ModelIds (16 RF + 16 MLP), {what}, 1 task thread", **r}), flush=True)
