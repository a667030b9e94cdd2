"""GPU: BASELINE.json's config 5 as written -- 32 ModelIds (16 random forests of 200 trees x depth 8 over 64 fields with
majority vote, 16 MLPs 64-256-256-8), events addressed to a random id, Del / Add(version + 1) interleaved with the events --
driven through the dynamic operator's C entry points (fjh_cofunction_process_dense / fjh_cofunction_process_control) and
compared, EVERY emitted score and status in arrival order, with the oracle.

Reference behaviour being matched: a control message takes effect between the events around it
(S/api/functions/EvaluationCoFunction.scala:69-74), an event addressed to an id that is not (or no longer) in the metadata
map scores against the empty model (:106-110 -> EmptyScore), T/RichConnectedStreamSpec.scala:187-286.
(S/, T/ = flink-jpmml-scala/src/{main,test}/scala/io/radicalbit/flink/pmml/scala/)"""
import os
import uuid

import numpy as np
import pytest

from flink_jpmml_b200 import api

pytestmark = pytest.mark.gpu

N_EVENTS = 1 << 18
N_MODELS = 32
F = 64


def test_c5_32_models_hot_swapped_every_event_against_the_oracle(engine, gen, oracle, tmp_path):
    rng = np.random.default_rng(20261111)
    threads = os.cpu_count() or 4
    files, oracles = [], []
    for k in range(N_MODELS):
        if k % 2 == 0:
            pm = gen.rf_pmml(T=200, depth=8, F=F, n_labels=4, seed=5000 + k)[0]
        else:
            pm = gen.nn_pmml(n_in=F, hidden=(256, 256), n_out=8, seed=5000 + k)[0]
        p = tmp_path / f"c5_{k}.pmml"
        p.write_bytes(pm)
        files.append(str(p)); oracles.append(oracle.OracleModel(pm))
    names = [str(uuid.UUID(int=7000 + k)) for k in range(N_MODELS)]
    version = [1] * N_MODELS
    op = api.EvaluationCoFunction(engine, max_batch=16384)
    for k in range(N_MODELS):
        op.processElement2(api.AddMessage(names[k], 1, files[k]))

    X = rng.standard_normal((N_EVENTS, F))
    model_of = rng.integers(0, N_MODELS, N_EVENTS)
    live = np.ones(N_EVENTS, dtype=bool)          # False: the event names a version that is not being served -> EmptyScore
    # the stream: stretches of events (each 3 % addressed to a stale or future version) separated by one hot swap
    cuts = np.sort(rng.choice(np.arange(1, N_EVENTS), size=63, replace=False)).tolist() + [N_EVENTS]
    lo = 0
    for hi in cuts:
        ids, index_of, idx = [], {}, np.empty(hi - lo, dtype=np.int32)
        odd = rng.random(hi - lo) < 0.03
        for j in range(hi - lo):
            k = int(model_of[lo + j]); v = version[k]
            if odd[j]:
                v = v + 1 if (j & 1) else v - 1      # not yet added / already deleted (version 0 was never added)
                live[lo + j] = False
            key = (k, v)
            if key not in index_of:
                index_of[key] = len(ids); ids.append(f"{names[k]}_{v}")
            idx[j] = index_of[key]
        op.processDense1(ids, idx, X[lo:hi])
        if hi < N_EVENTS:                             # hot swap of a random id: Del, then Add(version + 1) of the same document
            k = int(rng.integers(N_MODELS))
            op.processElement2(api.DelMessage(names[k], version[k]))
            version[k] += 1
            op.processElement2(api.AddMessage(names[k], version[k], files[k]))
        lo = hi
    op.flush()
    score, status = op.collect_arrays()
    op.close()
    assert len(score) == N_EVENTS

    exp_score = np.full(N_EVENTS, np.nan); exp_status = np.full(N_EVENTS, 5, dtype=np.uint8)
    for k in range(N_MODELS):
        sel = np.nonzero((model_of == k) & live)[0]
        r = oracles[k].predict(X[sel], nthreads=threads)
        exp_score[sel] = r["score"]; exp_status[sel] = r["status"]
    assert np.array_equal(status, exp_status), f"{(status != exp_status).sum()} statuses differ"
    ok = exp_status == 0
    assert ok.sum() > 0.9 * N_EVENTS and (status[~live] == 5).all()
    assert np.array_equal(score[ok], exp_score[ok]), f"{(score[ok] != exp_score[ok]).sum()} class labels differ from the oracle"
    assert np.isnan(score[~ok]).all()
    for k in range(N_MODELS):
        engine.table_del(names[k], version[k])
