"""GPU: the reference's operator specs replayed against the batching operators through the C ABI.

  T/RichConnectedStreamSpec.scala:71-335   every lifecycle sequence, (1,1,1,1) <-> Score(3.0) / EmptyScore
  T/QuickDataStreamSpec.scala:62-105       fixed-model operator, dense + sparse mix keeps order
  T/RichDataStreamSpec.scala:73-105        bad path / empty PMML fail the job (ModelLoadingException)
  T/api/functions/EvaluationCoFunctionSpec.scala:122-144
(T/ = flink-jpmml-scala/src/test/scala/io/radicalbit/flink/pmml/scala/)
Each sequence runs twice: flushing after every element (the reference's per-event timing) and with one flush at the end
(everything in one micro-batch) -- the outputs must be identical."""
import os
import uuid

import numpy as np
import pytest

from conftest import GOLDEN
from flink_jpmml_b200 import api

pytestmark = pytest.mark.gpu

KMEANS = os.path.join(GOLDEN, "iris_kmeans43.pmml")
NO_TARGET = os.path.join(GOLDEN, "iris_kmeans_nooutput_notarget.pmml")
EMPTY_DOC = os.path.join(GOLDEN, "iris_kmeans_empty.pmml")
MISSING = "/not/existing/path.pmml"
N1, N2 = str(uuid.UUID(int=1)), str(uuid.UUID(int=2))
ID1, ID2 = f"{N1}_1", f"{N2}_1"
DEFAULT, EMPTY = api.Prediction(api.Score(3.0)), api.Prediction(api.EmptyScore)
ONES = api.DenseVector([1.0, 1.0, 1.0, 1.0])


def ev(mid, vec=ONES):
    return ("event", mid, vec)


def add(name, path=KMEANS):
    return ("ctrl", api.AddMessage(name, 1, path))


def dele(name):
    return ("ctrl", api.DelMessage(name, 1))


SEQUENCES = {
    "model->event": ([add(N1), ev(ID1)], [DEFAULT]),                                                           # :71-81
    "event->model->event": ([ev(ID1), add(N1), ev(ID1)], [EMPTY, DEFAULT]),                                      # :83-95
    "two models, second without target": ([add(N1), add(N2, NO_TARGET), ev(ID1), ev(ID2)], [DEFAULT, EMPTY]),  # :97-116
    "interleaved": ([ev(ID1), add(N1), ev(ID2), add(N2), ev(ID1), ev(ID2)], [EMPTY, EMPTY, DEFAULT, DEFAULT]),  # :118-139
    "only events": ([ev(ID1), ev(ID2)], [EMPTY, EMPTY]),                                                        # :141-153
    "only models": ([add(N1), add(N2, NO_TARGET)], []),                                                         # :155-172
    "del with no model": ([dele(N1), ev(ID1)], [EMPTY]),                                                        # :174-185
    "del current model": ([add(N1), ev(ID1), dele(N1), ev(ID1)], [DEFAULT, EMPTY]),                             # :187-204
    "del other model": ([add(N1), ev(ID1), dele(N2), ev(ID1)], [DEFAULT, DEFAULT]),                             # :206-223
    "add-del-add": ([add(N1), ev(ID1), dele(N1), ev(ID1), add(N1), ev(ID1)], [DEFAULT, EMPTY, DEFAULT]),        # :225-244
    "del-add-del": ([dele(N1), ev(ID1), add(N1), ev(ID1), dele(N1), ev(ID1)], [EMPTY, DEFAULT, EMPTY]),         # :246-265
    "add on same id is discarded": ([add(N1), ev(ID1), add(N1, NO_TARGET), ev(ID1)], [DEFAULT, DEFAULT]),       # :267-286
    "wrong-size input": ([add(N1), ev(ID1, api.DenseVector([1.0, 1.0]))], [EMPTY]),                             # :302-320
}


@pytest.mark.parametrize("name", list(SEQUENCES))
@pytest.mark.parametrize("mode", ["flush-each", "one-batch"])
def test_connected_stream_sequences(engine, name, mode):
    ops, expected = SEQUENCES[name]
    op = api.EvaluationCoFunction(engine, max_batch=64)
    try:
        for o in ops:
            if o[0] == "event":
                op.processElement1(o[1], o[2])
            else:
                op.processElement2(o[1])
            if mode == "flush-each":
                op.flush()
        op.flush()
        assert op.collect() == expected
    finally:
        op.close()
    engine.table_del(N1, 1), engine.table_del(N2, 1)   # operators of later tests start from an empty cache


def test_unloadable_model_fails_the_job_on_the_first_event(engine):
    # :288-300 (missing path) and :322-335 (empty PMML): Add succeeds, the first event for that id is fatal
    for path in (MISSING, EMPTY_DOC):
        op = api.EvaluationCoFunction(engine)
        op.processElement2(api.AddMessage(N1, 1, path))
        with pytest.raises(api.ModelLoadingException):
            op.processElement1(ID1, ONES)
        op.close()


def test_from_metadata_edge_cases(engine):
    # EvaluationCoFunctionSpec.scala:122-144: empty metadata -> empty model; malformed id -> WrongModelIdFormat
    op = api.EvaluationCoFunction(engine)
    op.processElement1(ID1, ONES)
    op.flush()
    assert op.collect() == [EMPTY]
    with pytest.raises(api.WrongModelIdFormat):
        op.processElement1("23441-_1", ONES)
    op.close()


def test_buffered_events_keep_the_model_they_arrived_under(engine):
    """Control messages act between events even though nothing is flushed: 200 events, Del in the middle."""
    op = api.EvaluationCoFunction(engine, max_batch=1024)
    op.processElement2(api.AddMessage(N1, 1, KMEANS))
    for _ in range(100):
        op.processElement1(ID1, ONES)
    op.processElement2(api.DelMessage(N1, 1))
    for _ in range(100):
        op.processElement1(ID1, ONES)
    assert op.servingModels == 0            # evicted at once ...
    op.flush()
    out = op.collect()
    assert out == [DEFAULT] * 100 + [EMPTY] * 100   # ... yet the buffered events still scored against it
    op.close()


def test_checkpoint_restore_reloads_models_lazily(engine):
    op = api.EvaluationCoFunction(engine)
    op.processElement2(api.AddMessage(N1, 1, KMEANS))
    op.processElement1(ID1, ONES)
    snap = op.snapshotState()               # flushes the pending micro-batch first
    assert op.collect() == [DEFAULT]
    op.close()
    engine.table_del(N1, 1)
    op2 = api.EvaluationCoFunction(engine)
    op2.initializeState([snap])
    assert op2.servingMetadata == 1 and op2.servingModels == 0
    op2.processElement1(ID1, ONES)
    op2.flush()
    assert op2.collect() == [DEFAULT] and op2.servingModels == 1
    op2.close()
    engine.table_del(N1, 1)


def test_quick_evaluate_keeps_order_on_heterogeneous_input(engine):
    # QuickDataStreamSpec.scala:62-66: Seq(dense (1,1,1,1), sparse idx 0-3 (1,2,3,4)) -> Seq(3.0, 4.0)
    fn = api.EvaluationFunction(engine, KMEANS, max_batch=3)
    assert fn.modelName == "k-means"
    seq = [ONES, api.SparseVector(4, [0, 1, 2, 3], [1.0, 2.0, 3.0, 4.0]), api.SparseVector(2, [0, 3], [1.0, 1.0]), ONES,
           api.SparseVector(4, [0, 2], [1.0, 2.0])]
    for v in seq:
        fn.processElement(v)
    fn.flush()
    assert fn.collect() == [DEFAULT, api.Prediction(api.Score(4.0)), EMPTY, DEFAULT, DEFAULT]
    fn.close()
    for bad in (MISSING, EMPTY_DOC):        # RichDataStreamSpec.scala:82-105
        with pytest.raises(api.ModelLoadingException):
            api.EvaluationFunction(engine, bad)


def test_pmml_model_predict_single_event(engine, golden_vectors):
    # api/PmmlModelSpec.scala:50-83 through PmmlModel.predict (one event at a time, the literal drop-in)
    model = api.PmmlModel.fromReader(engine, KMEANS)
    assert model.predict(ONES) == DEFAULT
    assert model.predict(api.SparseVector(4, [0, 1, 2, 3], [1.0, 2.0, 3.0, 4.0])) == api.Prediction(api.Score(4.0))
    assert model.predict(api.SparseVector(4, [0, 2], [1.0, 2.0]), 0.0) == DEFAULT
    assert model.predict(api.SparseVector(4, [0, 2], [1.0, 2.0])) == DEFAULT
    assert model.predict(api.DenseVector([1.0, 2.0, 3.0])) == EMPTY
    assert api.PmmlModel.empty(engine).predict(ONES) == EMPTY
    with pytest.raises(api.ModelLoadingException):
        api.PmmlModel.fromReader(engine, MISSING)
    assert DEFAULT.value.get() == 3.0 and EMPTY.value.getOrElse(-1.0) == -1.0
    with pytest.raises(LookupError):
        EMPTY.value.get()                    # TargetSpec.scala:42-44


def test_hot_swap_under_load(engine, gen, oracle, tmp_path):
    """Config C5 in miniature: several ModelIds (RF + MLP), events interleaved with Del / Add(version+1)."""
    rng = np.random.default_rng(5)
    files, oracles = {}, {}
    for k in range(4):
        pm = gen.rf_pmml(T=20, depth=4, F=16, n_labels=4, seed=100 + k)[0] if k % 2 == 0 else gen.nn_pmml(n_in=16, hidden=(32, 32), n_out=4, seed=100 + k)[0]
        p = tmp_path / f"m{k}.pmml"
        p.write_bytes(pm)
        files[k], oracles[k] = str(p), oracle.OracleModel(pm)
    names = [str(uuid.UUID(int=100 + k)) for k in range(4)]
    version = [1] * 4
    op = api.EvaluationCoFunction(engine, max_batch=256)
    for k in range(4):
        op.processElement2(api.AddMessage(names[k], 1, files[k]))
    expected = []
    X = rng.standard_normal((3000, 16))
    for i in range(3000):
        k = int(rng.integers(4))
        if i % 97 == 96:                     # hot swap: Del then Add(version + 1)
            op.processElement2(api.DelMessage(names[k], version[k]))
            version[k] += 1
            op.processElement2(api.AddMessage(names[k], version[k], files[k]))
        stale = rng.random() < 0.05          # an event still addressed to the previous version: EmptyScore
        v = version[k] - 1 if (stale and version[k] > 1) else version[k]
        op.processElement1(f"{names[k]}_{v}", api.DenseVector(X[i]))
        if v == version[k]:
            r = oracles[k].predict(X[i:i + 1])
            expected.append(api.Prediction.of(r["score"][0], r["status"][0]))
        else:
            expected.append(EMPTY)
    op.flush()
    assert op.collect() == expected
    op.close()


def test_double_buffered_operator_keeps_order_over_many_batches(engine, gen, oracle):
    """Fixed-model operator with small pinned micro-batches: 10 000 events cross ~40 submit/wait hand-overs; the output
    must be the oracle's answers in arrival order, wrong-size vectors included."""
    fn = api.EvaluationFunction(engine, KMEANS, max_batch=256)
    X = gen.iris_events(10000)
    o = oracle.OracleModel(open(KMEANS, "rb").read()).predict(X)
    fn.processDense(X[:5000])
    fn.processElement(api.DenseVector([1.0, 2.0]))          # size mismatch in the middle of the stream
    fn.processDense(X[5000:])
    fn.flush()
    out = fn.collect()
    exp = [api.Prediction.of(o["score"][i], o["status"][i]) for i in range(10000)]
    assert out == exp[:5000] + [EMPTY] + exp[5000:]
    fn.close()


def test_dynamic_operator_bulk_ingest(engine, gen, oracle, tmp_path):
    pms = [gen.rf_pmml(T=10, depth=4, F=16, n_labels=3, seed=50 + k)[0] for k in range(3)]
    names = [str(uuid.UUID(int=500 + k)) for k in range(3)]
    op = api.EvaluationCoFunction(engine, max_batch=512)
    for k, pm in enumerate(pms):
        p = tmp_path / f"d{k}.pmml"; p.write_bytes(pm)
        op.processElement2(api.AddMessage(names[k], 1, str(p)))
    rng = np.random.default_rng(1)
    X = rng.standard_normal((5000, 16)); idx = rng.integers(0, 4, 5000)
    ids = [f"{n}_1" for n in names] + [f"{names[0]}_9"]       # the 4th id is unknown -> EmptyScore
    op.processDense1(ids, idx, X)
    op.flush()
    out = op.collect()
    exp = []
    oms = [oracle.OracleModel(pm) for pm in pms]
    for i in range(5000):
        if idx[i] == 3: exp.append(EMPTY)
        else:
            r = oms[idx[i]].predict(X[i:i + 1]); exp.append(api.Prediction.of(r["score"][0], r["status"][0]))
    assert out == exp
    op.close()
    for n in names: engine.table_del(n, 1)


def test_producer_writes_rows_into_the_pinned_batch_itself(engine, gen, oracle, tmp_path):
    """reserve / commit: the producer fills the micro-batch in place (no operator-side copy); mixed with ordinary events and a
    wrong-size event the output order and values are those of processElement."""
    pm = gen.gbt_pmml(T=20, depth=4, F=12, seed=70)[0]
    p = tmp_path / "g.pmml"; p.write_bytes(pm)
    om = oracle.OracleModel(pm)
    rng = np.random.default_rng(3); X = rng.standard_normal((3000, 12))
    fn = api.EvaluationFunction(engine, str(p), max_batch=256)
    assert fn.reserve(5, 10) is None                                  # not the model's field count
    def produce(lo, hi):                                              # grants stop at the batch boundary
        while lo < hi:
            rows = fn.reserve(12, hi - lo)
            assert rows is not None and 0 < rows.shape[0] <= 256
            rows[:] = X[lo:lo + rows.shape[0]]; fn.commit(rows.shape[0]); lo += rows.shape[0]
    produce(0, 1000)
    fn.processElement(api.DenseVector(X[1000]))
    fn.processElement(api.DenseVector([1.0, 2.0]))                    # wrong size: EmptyScore, in place
    produce(1001, 1101)
    fn.processDense(X[1101:])
    fn.flush()
    out = fn.collect(); fn.close()
    r = om.predict(X)
    exp = [api.Prediction.of(r["score"][i], r["status"][i]) for i in range(3000)]
    assert out == exp[:1001] + [EMPTY] + exp[1001:]


def test_bulk_ingest_parks_events_behind_a_running_model_load(engine, gen, oracle, tmp_path):
    """Hot swap without stopping the stream: Add starts the load in the background; bulk events for that id that arrive
    before it has finished are parked (they keep their output position), the other models' events are scored meanwhile.
    Order and values must be what a blocking load would have produced; a Del while parked does not resurrect the model."""
    small = gen.rf_pmml(T=10, depth=4, F=64, n_labels=3, seed=60)[0]
    big = gen.rf_pmml(seed=61)[0]                                  # 200 trees x depth 8: tens of ms to parse and compile
    ps, pb_ = tmp_path / "small.pmml", tmp_path / "big.pmml"; ps.write_bytes(small); pb_.write_bytes(big)
    ns, nb = str(uuid.UUID(int=600)), str(uuid.UUID(int=601))
    oms, omb = oracle.OracleModel(small), oracle.OracleModel(big)
    rng = np.random.default_rng(2)
    n = 40000; X = rng.standard_normal((n, 64)); idx = rng.integers(0, 2, n).astype(np.int32)
    exp_s, exp_b = oms.predict(X, nthreads=8), omb.predict(X, nthreads=8)
    expect = lambda second: [api.Prediction.of((exp_s if k == 0 else second)["score"][i], (exp_s if k == 0 else second)["status"][i]) for i, k in enumerate(idx)]
    op = api.EvaluationCoFunction(engine, max_batch=4096)
    base = op.servingModels                                        # the engine's table is shared with the other tests' leftovers
    op.processElement2(api.AddMessage(ns, 1, str(ps)))
    op.processDense1([f"{ns}_1"], np.zeros(16, np.int32), X[:16]); op.flush(); op.collect()     # the small model is loaded and cached
    op.processElement2(api.AddMessage(nb, 1, str(pb_)))            # background load starts ...
    op.processDense1([f"{ns}_1", f"{nb}_1"], idx, X)               # ... and the events do not wait for it
    parked = op.parkedEvents
    early = op.collect()                                           # whatever is ready: a prefix of the expected output
    op.flush()
    out = early + op.collect()
    assert parked > 0, "the load finished before the first event: nothing was parked (make the model bigger)"
    assert len(early) < n and out == expect(exp_b)
    assert op.servingModels == base + 2
    # Del while events are parked: they are still scored with the model they were bound to; the model is not cached afterwards
    op.processElement2(api.DelMessage(nb, 1))
    op.processElement2(api.AddMessage(nb, 2, str(pb_)))
    op.processDense1([f"{ns}_1", f"{nb}_2"], idx, X)
    parked2 = op.parkedEvents - parked
    op.processElement2(api.DelMessage(nb, 2))
    op.processDense1([f"{ns}_1", f"{nb}_2"], idx[:100], X[:100])   # after the Del: EmptyScore for that id
    op.flush()
    out2 = op.collect()
    assert parked2 > 0 and out2[:n] == expect(exp_b)
    assert out2[n:] == [EMPTY if k == 1 else api.Prediction.of(exp_s["score"][i], exp_s["status"][i]) for i, k in enumerate(idx[:100])]
    assert op.servingModels == base + 1
    # a load that fails is fatal for the job from the call that settles the parked events
    op.processElement2(api.AddMessage(nb, 3, MISSING))
    with pytest.raises(api.ModelLoadingException):
        op.processDense1([f"{nb}_3"], np.zeros(8, np.int32), X[:8])
        op.flush()
    op.close()
    engine.table_del(ns, 1)


def test_engine_is_usable_from_several_host_threads(engine, gen, oracle):
    """SURVEY.md §8b Threading: several subtasks (threads) of one TaskManager share an engine.  Four threads score
    different models concurrently while a fifth keeps adding / deleting table entries; every result must be right."""
    import threading
    from flink_jpmml_b200 import capi
    pms = [gen.gbt_pmml(T=30, depth=6, F=20, seed=300 + k)[0] for k in range(4)]
    models = [engine.compile(pm) for pm in pms]
    refs, Xs = [], []
    for k, pm in enumerate(pms):
        X, pres, pb = gen.events(30000, 20, 400 + k, 0.1)
        Xs.append((X, pb)); refs.append(oracle.OracleModel(pm).predict(X[:3000], pb[:3000], nthreads=4)["score"])
    errors, stop = [], threading.Event()

    def scorer(k):
        try:
            for _ in range(15):
                r = engine.score(models[k], *Xs[k])
                if not np.array_equal(r["score"][:3000], refs[k]) or (r["status"] != 0).any():
                    errors.append(f"thread {k}: wrong result")
        except Exception as e:          # noqa: BLE001
            errors.append(repr(e))

    def churn():
        name = str(uuid.UUID(int=900))
        v = 0
        while not stop.is_set():
            v += 1
            engine.table_add(name, v, models[v % 4]); got = engine.table_get(name, v)
            if got is None: errors.append("table_get lost an entry")
            else: got.release()
            engine.table_del(name, v)

    threads = [threading.Thread(target=scorer, args=(k,)) for k in range(4)]
    ch = threading.Thread(target=churn); ch.start()
    [t.start() for t in threads]; [t.join() for t in threads]
    stop.set(); ch.join()
    assert not errors, errors[:3]


def test_operators_of_several_threads_share_one_engine_without_deadlock(engine, gen, oracle, tmp_path):
    """Four dynamic operators (four task threads of one TaskManager) on ONE engine, each with five model buckets per flush:
    together they want more stream slots than the engine has.  fjp_submit never blocks (it answers FJP_BUSY and the operator
    gives back its own oldest ticket), so every thread finishes and every prediction is the oracle's."""
    import threading
    pms = [gen.gbt_pmml(T=12, depth=4, F=10, seed=800 + k)[0] for k in range(5)]
    paths = []
    for k, pm in enumerate(pms):
        p = tmp_path / f"s{k}.pmml"; p.write_bytes(pm); paths.append(str(p))
    rng = np.random.default_rng(3)
    X = rng.standard_normal((6000, 10)); idx = rng.integers(0, 5, 6000).astype(np.int32)
    oms = [oracle.OracleModel(pm) for pm in pms]
    exp = np.empty(6000)
    for k in range(5):
        exp[idx == k] = oms[k].predict(X[idx == k])["score"]
    errors = []

    def task(t):
        try:
            names = [str(uuid.UUID(int=8000 + 10 * t + k)) for k in range(5)]
            op = api.EvaluationCoFunction(engine, max_batch=300)
            for k in range(5):
                op.processElement2(api.AddMessage(names[k], 1, paths[k]))
            for _ in range(4):
                op.processDense1([f"{n}_1" for n in names], idx, X)
            op.flush()
            score, status = op.collect_arrays()
            op.close()
            if len(score) != 4 * 6000 or (status != 0).any() or not np.array_equal(score, np.tile(exp, 4)):
                errors.append(f"thread {t}: wrong predictions")
            for n in names:
                engine.table_del(n, 1)
        except Exception as e:          # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=task, args=(t,), daemon=True) for t in range(4)]
    [t.start() for t in threads]
    [t.join(timeout=120) for t in threads]
    assert not any(t.is_alive() for t in threads), "an operator thread is stuck in fjp_submit / fjp_wait"
    assert not errors, errors[:3]
