"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): cluster ids, class labels, node ids and statuses bit-exact; continuous scores and
probabilities within 1e-6 relative (they are in fact expected to be bit-identical for sums, the tolerance is the
contract).  Reference-pinned vectors: api/PmmlModelSpec.scala:50-83 (k-means only)."""
import numpy as np
import pytest

from conftest import golden_bytes, vector_to_batch

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def compare(res, ref, rtol=RTOL, exact_scores=False):
    assert np.array_equal(res["status"], ref["status"]), (res["status"][:20], ref["status"][:20])
    ok = ref["status"] == 0
    if exact_scores:
        assert np.array_equal(res["score"][ok], ref["score"][ok])
    else:
        np.testing.assert_allclose(res["score"][ok], ref["score"][ok], rtol=rtol, atol=0)
    assert np.isnan(res["score"][~ok]).all()


def test_golden_vectors_through_c_abi(engine, oracle, gen, golden_vectors):
    for v in golden_vectors["predict"]:
        m = engine.compile(golden_bytes(v["model"]))
        X, pb, rep = vector_to_batch(gen, v)
        r = engine.score(m, X, pb, rep)
        if v["score"] is None:
            assert r["status"][0] != 0 and np.isnan(r["score"][0]), v
        else:
            assert r["status"][0] == 0 and r["score"][0] == v["score"], (v, r)


@pytest.mark.parametrize("name", ["iris_kmeans43.pmml", "iris_kmeans42.pmml", "iris_kmeans41.pmml", "iris_kmeans40.pmml",
                                  "iris_kmeans32.pmml", "iris_kmeans_nooutput.pmml", "iris_kmeans_nooutput_notarget.pmml",
                                  "iris_kmeans_stringfields.pmml"])
def test_iris_models_vs_oracle(engine, oracle, gen, name):
    pm = golden_bytes(name)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert m.active_fields == om.active_fields and m.model_name == om.model_name
    X = gen.iris_events(5000)
    rng = np.random.default_rng(7)
    pres = rng.random(X.shape) > 0.2
    pb = gen.pack_present(pres)
    for args in [(X, None, None), (X, pb, None), (X, pb, 0.0)]:
        r, o = engine.score(m, *args), om.predict(*args)
        compare(r, o, exact_scores=True)
        assert np.array_equal(r["aux"][o["status"] == 0], o["aux"][o["status"] == 0])


def test_c2_kmeans_k64_f128(engine, oracle, gen):
    pm, arr = gen.config_c2()
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    for p_missing in (0.0, 0.05):
        X, pres, pb = gen.events(8192, 128, gen.SEED_BASE + 2, p_missing)
        r, o = engine.score(m, X, pb), om.predict(X, pb, nthreads=8)
        compare(r, o, exact_scores=True)
        assert np.array_equal(r["aux"], o["aux"])
        assert np.array_equal(r["score"], gen.kmeans_numpy(arr["centroids"], X, pres))


@pytest.mark.parametrize("dtype,method", [("double", "sum"), ("float", "sum"), ("double", "average")])
def test_gbt_small_vs_oracle(engine, oracle, gen, dtype, method):
    pm, arr = gen.gbt_pmml(T=37, depth=5, F=13, seed=11, dtype=dtype, method=method)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "format=complete" in m.describe and "enc=float" in m.describe
    X, pres, pb = gen.events(20000, 13, 12, 0.15)
    X[::97, 3] = np.nan            # present NaN: greatest in compareTo order
    X[::89, 5] = np.inf
    X[::83, 7] = -0.0
    r = engine.score(m, X, pb, node_ids=True)
    o = om.predict(X, pb, node_ids=True, nthreads=8)
    compare(r, o, exact_scores=True)
    assert np.array_equal(r["node_ids"], o["node_ids"])   # leaf ids bit-exact
    np.testing.assert_allclose(r["score"], gen.gbt_numpy(arr, np.nan_to_num(X, nan=np.inf), pres), rtol=1e-12)


def test_c4_gbt_500_trees_vs_oracle_and_numpy(engine, oracle, gen):
    pm, arr = gen.gbt_pmml()
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(4096, 100, gen.SEED_BASE + 4, 0.10)
    r = engine.score(m, X, pb, node_ids=True)
    o = om.predict(X, pb, node_ids=True, nthreads=8)
    compare(r, o, exact_scores=True)
    assert np.array_equal(r["node_ids"], o["node_ids"])
    # size-independent check at a larger size: numpy restatement of the same walk
    X, pres, pb = gen.events(1 << 17, 100, 99, 0.10)
    r = engine.score(m, X, pb)
    assert (r["status"] == 0).all()
    assert np.array_equal(r["score"], gen.gbt_numpy(arr, X, pres))


def test_rf_majority_vote(engine, oracle, gen):
    pm, arr = gen.rf_pmml(T=50, depth=6, F=20, n_labels=4, seed=21)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(20000, 20, 22, 0.1)
    r, o = engine.score(m, X, pb, node_ids=True), om.predict(X, pb, node_ids=True, nthreads=8)
    compare(r, o, exact_scores=True)
    assert np.array_equal(r["node_ids"], o["node_ids"])
    assert np.array_equal(r["score"], gen.rf_numpy(arr, X, pres))


def test_rf_average_of_leaf_distributions(engine, oracle, gen):
    # scikit-learn style random forest: ScoreDistribution per leaf, multipleModelMethod="average"
    pm, arr = gen.rf_pmml(T=30, depth=5, F=12, n_labels=3, seed=41, method="average")
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(8000, 12, 42, 0.1)
    r, o = engine.score(m, X, pb, node_ids=True), om.predict(X, pb, node_ids=True, nthreads=8)
    compare(r, o, exact_scores=True)
    assert np.array_equal(r["node_ids"], o["node_ids"])
    lab, _ = gen.rf_average_numpy(arr, X, pres)
    assert (r["score"] == lab).mean() > 0.999      # numpy sums the same terms; argmax ties aside they agree


@pytest.mark.parametrize("n_classes", [2, 3])
def test_gbt_classifier_model_chain(engine, oracle, gen, n_classes):
    # boosted-tree classifier as exporters write it: modelChain(sum of trees per class -> logit / softmax RegressionModel)
    pm, arr = gen.gbt_classifier_pmml(T=40, depth=6, F=30, n_classes=n_classes, seed=51)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "format=complete" in m.describe and "method=9" in m.describe
    X, pres, pb = gen.events(20000, 30, 52, 0.1)
    r, o = engine.score(m, X, pb, n_probs=n_classes), om.predict(X, pb, n_probs=n_classes, nthreads=8)
    compare(r, o, exact_scores=True)                        # labels bit-exact
    np.testing.assert_allclose(r["probs"], o["probs"], rtol=RTOL)
    lab, probs = gen.gbt_classifier_numpy(arr, X, pres)
    assert np.array_equal(r["score"], lab)
    np.testing.assert_allclose(r["probs"], probs, rtol=1e-9)


def test_c3_multinomial_regression(engine, oracle, gen):
    pm, arr = gen.regression_pmml()
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(2048, 1000, gen.SEED_BASE + 3)
    r, o = engine.score(m, X, n_probs=20), om.predict(X, n_probs=20, nthreads=8)
    compare(r, o, exact_scores=True)                       # labels bit-exact
    np.testing.assert_allclose(r["probs"], o["probs"], rtol=RTOL)


def test_regression_generic_and_missing(engine, oracle, gen):
    pm, arr = gen.linear_regression_pmml(F=8, seed=7, normalization="none", exponents=True)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(4000, 8, 5, 0.05)
    r, o = engine.score(m, X, pb), om.predict(X, pb)
    compare(r, o)
    assert (o["status"] == 3).any() and (o["status"] == 0).any()   # a missing predictor leaves no result


def test_neural_network(engine, oracle, gen):
    pm, arr = gen.nn_pmml()   # 64 -> 256 -> 256 -> 8
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(4096, 64, gen.SEED_BASE + 5)
    r, o = engine.score(m, X, n_probs=8), om.predict(X, n_probs=8, nthreads=8)
    compare(r, o, exact_scores=True)
    np.testing.assert_allclose(r["probs"], o["probs"], rtol=RTOL)
    for act in ("tanh", "logistic"):
        pm, arr = gen.nn_pmml(n_in=10, hidden=(12,), n_out=3, seed=3, activation=act)
        m, om = engine.compile(pm), oracle.OracleModel(pm)
        X, pres, pb = gen.events(1000, 10, 4)
        r, o = engine.score(m, X, n_probs=3), om.predict(X, n_probs=3)
        compare(r, o, exact_scores=True)
        np.testing.assert_allclose(r["probs"], o["probs"], rtol=RTOL)


def test_empty_model_and_size_mismatch(engine, gen):
    X = np.ones((5, 4))
    r = engine.score(None, X)
    assert (r["status"] == 5).all() and np.isnan(r["score"]).all()
    m = engine.compile(golden_bytes("iris_kmeans43.pmml"))
    r = engine.score(m, np.ones((3, 3)))
    assert (r["status"] == 1).all()
    assert engine.launch_count > 0
    before = engine.status_counts
    engine.score(None, np.ones((7, 4))); engine.score(m, np.ones((3, 3))); engine.score(m, np.ones((11, 4)))
    after = engine.status_counts
    assert [a - b for a, b in zip(after, before)] == [11, 3, 0, 0, 0, 7]      # per-batch outcome counters


def test_sparse_csr_ingest_matches_dense_path(engine, oracle, gen):
    """SparseVector batches as CSR: expanded on the device, same answers as the dense + present-bits path."""
    pm, arr = gen.gbt_pmml(T=40, depth=6, F=70, seed=31)
    m = engine.compile(pm)
    rng = np.random.default_rng(3)
    n, F = 30000, 70
    X, pres, pb = gen.events(n, F, 32, 0.6)            # 60 % of the slots absent
    X[pres & (rng.random(X.shape) < 0.1)] = 0.0         # explicit zeros stay present
    counts = pres.sum(1)
    row_ptr = np.concatenate([[0], np.cumsum(counts)])
    rows, cols = np.nonzero(pres)
    indices, values = cols.astype(np.int32), X[rows, cols]
    # out-of-range indices are ignored (T/QuickDataStreamSpec.scala:89-93 builds such a vector)
    indices = np.concatenate([indices, [F + 5, -1]]); values = np.concatenate([values, [9.0, 9.0]]); row_ptr[-1] += 2
    dense = engine.score(m, X, pb)
    for rep in (None, 0.25):
        dense = engine.score(m, X, pb, rep)
        sparse = engine.score_sparse(m, row_ptr, indices, values, F, rep)
        assert np.array_equal(dense["status"], sparse["status"]) and np.array_equal(dense["score"], sparse["score"])
    o = oracle.OracleModel(pm).predict(X[:4000], pb[:4000], 0.25, nthreads=8)
    assert np.array_equal(sparse["score"][:4000], o["score"])
    empty = engine.score_sparse(m, np.zeros(4, dtype=np.int64), np.zeros(0, dtype=np.int32), np.zeros(0), F)   # three all-missing rows
    assert np.array_equal(empty["score"], engine.score(m, np.zeros((3, F)), np.zeros((3, 3), dtype=np.uint32))["score"])
    assert (engine.score_sparse(m, row_ptr, indices, values, F + 1)["status"] == 1).all()                        # size mismatch


def test_model_replicate_peer_copy(gen, oracle):
    """fjp_model_replicate: a model compiled on GPU 0 is peer-copied to GPU 1 (no re-compile) and scores identically.
    Needs two GPUs (gpurun --gpus 2); skipped otherwise."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from flink_jpmml_b200 import capi
    e0, e1 = capi.Engine(0), capi.Engine(1)
    pm, arr = gen.gbt_pmml(T=64, depth=8, F=100, seed=9)
    m0 = e0.compile(pm)
    m1 = e1.replicate(m0)
    X, pres, pb = gen.events(50000, 100, 10, 0.1)
    r0, r1 = e0.score(m0, X, pb), e1.score(m1, X, pb)
    assert np.array_equal(r0["score"], r1["score"]) and np.array_equal(r0["status"], r1["status"])
    assert np.array_equal(r1["score"], gen.gbt_numpy(arr, X, pres))
    with pytest.raises(capi.EngineError, match="another device"):
        e1.score(m0, X, pb)
    e0.close(); e1.close()


def test_forest_on_derived_fields_all_entry_points(engine, oracle, gen):
    """sklearn-pipeline shape: standardised inputs s_i = (x_i - mu_i) / sd_i as LocalTransformations feeding a GBT.
    Host, sparse and device-pointer entry points all run the derive pre-pass."""
    import ctypes as C
    import torch
    from flink_jpmml_b200 import capi
    F = 12
    pm, arr = gen.gbt_pmml(T=24, depth=5, F=F, seed=61)
    rng = np.random.default_rng(62)
    mu, sd = rng.normal(size=F), rng.uniform(0.5, 2.0, size=F)
    lt = "<LocalTransformations>" + "".join(
        f'<DerivedField name="s{i}" optype="continuous" dataType="double"><Apply function="/"><Apply function="-"><FieldRef field="x{i}"/>'
        f'<Constant>{float(mu[i])!r}</Constant></Apply><Constant>{float(sd[i])!r}</Constant></Apply></DerivedField>' for i in range(F)) + "</LocalTransformations>"
    text = pm.decode().replace('<SimplePredicate field="x', '<SimplePredicate field="s')
    head, tail = text.split("</MiningSchema>", 1)            # the top-level MiningSchema is the first one
    pm2 = (head + "</MiningSchema>" + lt + tail).encode()
    m, om = engine.compile(pm2), oracle.OracleModel(pm2)
    assert "derived=12" in m.describe and m.n_fields == F
    X, pres, pb = gen.events(20000, F, 63, 0.1)
    r, o = engine.score(m, X, pb), om.predict(X, pb, nthreads=8)
    compare(r, o, exact_scores=True)
    assert np.array_equal(r["score"], gen.gbt_numpy(arr, (X - mu) / sd, pres))
    rows, cols = np.nonzero(pres)
    row_ptr = np.concatenate([[0], np.cumsum(pres.sum(1))])
    sp = engine.score_sparse(m, row_ptr, cols.astype(np.int32), X[rows, cols], F)
    assert np.array_equal(sp["score"], r["score"])
    dX, dP = torch.from_numpy(X).cuda(), torch.from_numpy(pb.view(np.int32)).cuda()
    dS, dT = torch.empty(len(X), dtype=torch.float64, device="cuda"), torch.empty(len(X), dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream()
    rc = capi.lib().fjp_score_device(engine._h, m._h, dX.data_ptr(), dP.data_ptr(), None, len(X), F, dS.data_ptr(), dT.data_ptr(), None, C.c_void_p(st.cuda_stream))
    st.synchronize()
    assert rc == 0 and np.array_equal(dS.cpu().numpy(), r["score"])


def test_pruned_trees_are_padded_into_the_on_chip_format(engine, oracle, gen):
    """XGBoost-style export (float fields, `<`, pruned trees of unequal depth): the compiler pads every tree to the
    ensemble's depth so the ensemble still runs in the shared-memory kernel; node ids and scores match the oracle."""
    pm = gen.gbt_pruned_pmml(T=60, max_depth=6, F=20, seed=5)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "format=complete" in m.describe and "depth=6" in m.describe
    X, pres, pb = gen.events(20000, 20, 6, 0.1)
    X[::50, 2] = np.nan
    r, o = engine.score(m, X, pb, node_ids=True), om.predict(X, pb, node_ids=True, nthreads=8)
    compare(r, o, exact_scores=True)
    assert np.array_equal(r["node_ids"], o["node_ids"])
    r2 = engine.score(m, X, pb)                      # the fast instantiation (no node ids)
    assert np.array_equal(r2["score"], o["score"])


def test_double_thresholds_use_rank_keys(engine, oracle, gen, monkeypatch):
    """Thresholds that are not float32 values (scikit-learn style midpoints): keys are ranks among the feature's sorted
    thresholds (binary search in the kernel), still in the on-chip complete-tree kernel, still bit-exact."""
    monkeypatch.setattr(gen, "FLOAT_THRESHOLDS", False)
    pm, arr = gen.gbt_pmml(T=64, depth=6, F=24, seed=71)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "enc=rank" in m.describe and "format=complete" in m.describe
    X, pres, pb = gen.events(30000, 24, 72, 0.1)
    X[::7, 0] = arr["thr"][0, 0]                      # values exactly on a threshold
    X[::11, 1] = np.nextafter(arr["thr"][0, 1], np.inf)
    X[::13, 2] = np.nan
    r, o = engine.score(m, X, pb, node_ids=True), om.predict(X, pb, node_ids=True, nthreads=8)
    compare(r, o, exact_scores=True)
    assert np.array_equal(r["node_ids"], o["node_ids"])
    assert np.array_equal(engine.score(m, X, pb)["score"], gen.gbt_numpy(arr, np.nan_to_num(X, nan=np.inf), pres))


@pytest.mark.parametrize("dtype,operator", [("double", "lessOrEqual"), ("float", "lessOrEqual"), ("double", "lessThan")])
def test_float_rows_score_identically(engine, gen, oracle, dtype, operator):
    """fjp_score_batch_f32: rows narrowed to float the way fjp_model_float_row_modes prescribes (round up for `<=` tests, down
    for `<`, plain cast for float fields) give the double rows' scores bit for bit -- including values that sit exactly on, one
    ulp above and one ulp below a threshold, huge values, infinities and NaN."""
    if operator == "lessThan":
        pm = gen.gbt_pruned_pmml(T=40, max_depth=6, F=20, seed=9, operator="lessThan", dtype=dtype)
        F = 20
    else:
        pm, _ = gen.gbt_pmml(T=64, depth=8, F=100, seed=11, dtype=dtype)
        F = 100
    m = engine.compile(pm)
    modes = engine.float_row_modes(m)
    assert modes is not None and len(modes) == F and set(modes.tolist()) <= ({0} if dtype == "float" else {1, 2})
    X, pres, pb = gen.events(148 * 640 + 777, F, 12, 0.1)
    thr = np.float32(X[:2000, :].astype(np.float32)).astype(np.float64)        # plant values on / next to float grid points
    X[:2000] = thr; X[2000:4000] = np.nextafter(X[2000:4000].astype(np.float32).astype(np.float64), np.inf)
    X[4000:6000] = np.nextafter(X[4000:6000].astype(np.float32).astype(np.float64), -np.inf)
    X[6000, :5] = [1e300, -1e300, np.inf, -np.inf, np.nan]
    ref = engine.score(m, X, pb)
    r = engine.score_f32(m, engine.narrow_rows(X, modes), pb)
    assert np.array_equal(r["status"], ref["status"]) and np.array_equal(r["score"], ref["score"], equal_nan=True)
    o = oracle.OracleModel(pm).predict(X[:8000], pb[:8000], nthreads=4)
    assert np.array_equal(r["score"][:8000], o["score"], equal_nan=True) and np.array_equal(r["status"][:8000], o["status"])
    # a model whose tests need the doubles says so
    assert engine.float_row_modes(engine.compile(gen.config_c2(k=4, F=8)[0])) is None


def test_submit_many_scores_every_batch_with_its_own_model(engine, oracle, gen):
    """fjp_submit_many: several micro-batches, each for its own model (tree ensemble, k-means, MLP, an empty batch in between),
    behind one ticket; after fjp_wait every batch holds what its model gives through fjp_score_batch and the oracle."""
    import ctypes as C
    from flink_jpmml_b200 import capi
    L = capi.lib()

    class Desc(C.Structure):
        _fields_ = [("model", C.c_void_p), ("X", C.c_void_p), ("present", C.c_void_p), ("replace", C.c_void_p), ("n", C.c_long),
                    ("vec_size", C.c_int), ("score", C.c_void_p), ("status", C.c_void_p)]
    pms = [gen.gbt_pmml(T=30, depth=5, F=12, seed=801)[0], gen.kmeans_pmml(np.random.default_rng(802).standard_normal((5, 12))),
           gen.nn_pmml(n_in=12, hidden=(16,), n_out=3, seed=803)[0]]
    models = [engine.compile(pm) for pm in pms]
    sizes = [3000, 0, 1777, 512]; which = [0, 1, 1, 2]
    keep, descs = [], (Desc * len(sizes))()
    for i, (n, k) in enumerate(zip(sizes, which)):
        X, pres, pb = gen.events(max(n, 1), 12, 810 + i, 0.1 if k == 0 else 0.0)
        X = np.ascontiguousarray(X[:n]); pb = np.ascontiguousarray(pb[:n]); sc = np.full(n, -1.0); st = np.full(n, 9, np.uint8)
        keep.append((X, pb, sc, st))
        descs[i] = Desc(models[k]._h, X.ctypes.data, pb.ctypes.data, None, n, 12, sc.ctypes.data, st.ctypes.data)
    ticket = C.c_int(-1)
    assert L.fjp_submit_many(engine._h, descs, len(sizes), C.byref(ticket)) == 0 and ticket.value >= 0
    assert L.fjp_wait(engine._h, ticket.value) == 0
    for (X, pb, sc, st), n, k in zip(keep, sizes, which):
        if n == 0: continue
        o = oracle.OracleModel(pms[k]).predict(X, pb, nthreads=4); r = engine.score(models[k], X, pb)
        assert np.array_equal(st, o["status"]) and np.array_equal(st, r["status"])
        ok = st == 0
        assert np.array_equal(sc[ok].view(np.int64), r["score"][ok].view(np.int64)) and np.array_equal(sc[ok], o["score"][ok])
    bad = (Desc * 1)(Desc(models[0]._h, keep[0][0].ctypes.data, None, None, 10, 7, keep[0][2].ctypes.data, keep[0][3].ctypes.data))
    assert L.fjp_submit_many(engine._h, bad, 1, C.byref(ticket)) != 0 and ticket.value == -1      # wrong vector size: refused as a whole


@pytest.mark.parametrize("rows,n", [(4096, 3 * 4096 + 1), (100000, 290000), (100000, 265537), (70000, 70001)])
def test_host_entry_points_chunk_pipeline(engine, oracle, gen, monkeypatch, rows, n):
    """The host-buffer entry points cut a batch into chunks over two stream slots, chain the copy-ins with an event and taper the
    last chunk (half, then the rest): whatever the cut, every row gets its own score -- double rows and float rows against
    the oracle (S/api/PmmlModel.scala:109-119 scores one event; here the batch boundary must not show)."""
    monkeypatch.setenv("FJP_CHUNK_ROWS", str(rows))
    pm, _ = gen.gbt_pmml(T=12, depth=4, F=8, seed=901)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(n, 8, 902, 0.1)
    o = om.predict(X, pb, nthreads=8)
    r = engine.score(m, X, pb)
    assert np.array_equal(r["status"], o["status"]) and np.array_equal(r["score"].view(np.int64), o["score"].view(np.int64))
    modes = engine.float_row_modes(m)
    assert modes is not None
    f = engine.score_f32(m, engine.narrow_rows(X, modes), pb)
    assert np.array_equal(f["status"], o["status"]) and np.array_equal(f["score"].view(np.int64), o["score"].view(np.int64))
