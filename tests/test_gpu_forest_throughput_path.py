"""GPU parity for the forest throughput path (csrc/forest_compact.cuh): rank-key pre-pass + compact walk.

Large batches of an eligible ensemble take this path on their own; here FJP_FOREST_COMPACT=2 forces it for batches small
enough for the oracle, FJP_FOREST_COMPACT=0 gives the one-kernel walk to compare with.  Bar: statuses, labels and scores
bit-identical to the oracle and to the other path (sums are added in segment order on both)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def score_both(engine, m, X, pb, monkeypatch, replace=None):
    monkeypatch.setenv("FJP_FOREST_COMPACT", "0")
    classic = engine.score(m, X, pb, replace)
    monkeypatch.setenv("FJP_FOREST_COMPACT", "2")
    launches = engine.launch_count
    compact = engine.score(m, X, pb, replace)
    assert engine.launch_count - launches >= 2          # pre-pass + walk really ran
    return classic, compact


def same(a, b):
    assert np.array_equal(a["status"], b["status"])
    assert np.array_equal(a["score"].view(np.int64), b["score"].view(np.int64))


@pytest.mark.parametrize("dtype,method,n", [("double", "sum", 20000), ("float", "sum", 7777), ("double", "average", 641)])
def test_gbt_vs_oracle_and_one_kernel_walk(engine, oracle, gen, monkeypatch, dtype, method, n):
    pm, arr = gen.gbt_pmml(T=37, depth=5, F=13, seed=11, dtype=dtype, method=method)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "compact=1" in m.describe
    X, pres, pb = gen.events(n, 13, 12, 0.15)
    X[::97, 3] = np.nan            # present NaN: greatest in compareTo order
    X[::89, 5] = np.inf
    X[::83, 7] = -0.0
    X[::7, 0] = arr["thr"][0, 0]   # exactly on a threshold, and one float step either side of it
    X[::11, int(arr["feat"][1, 0])] = np.nextafter(np.float32(arr["thr"][1, 0]), np.float32(np.inf))
    X[::13, int(arr["feat"][2, 0])] = np.nextafter(arr["thr"][2, 0], -np.inf)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    o = om.predict(X, pb, nthreads=8)
    same(compact, o)
    same(compact, classic)


def test_c4_depth8_vs_oracle(engine, oracle, gen, monkeypatch):
    pm, arr = gen.gbt_pmml()
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(4096, 100, gen.SEED_BASE + 4, 0.10)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)
    # at a size where the path is taken unforced: numpy restatement of the same walk
    monkeypatch.delenv("FJP_FOREST_COMPACT")
    X, pres, pb = gen.events(148 * 640 + 333, 100, 99, 0.10)
    launches = engine.launch_count
    r = engine.score(m, X, pb)
    assert engine.launch_count - launches >= 2
    assert (r["status"] == 0).all() and np.array_equal(r["score"], gen.gbt_numpy(arr, X, pres))


def test_replace_nan_and_dense_rows(engine, oracle, gen, monkeypatch):
    pm, arr = gen.gbt_pmml(T=24, depth=6, F=20, seed=3)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(5000, 20, 4, 0.2)
    rep = np.full(5000, np.nan); rep[::3] = 0.25; rep[1::3] = -1.5
    classic, compact = score_both(engine, m, X, pb, monkeypatch, rep)
    same(compact, om.predict(X, pb, rep, nthreads=8))
    same(compact, classic)
    classic, compact = score_both(engine, m, np.nan_to_num(X), None, monkeypatch)      # no present bits at all
    same(compact, om.predict(np.nan_to_num(X), None, nthreads=8))


@pytest.mark.parametrize("method", ["weightedAverage", "weightedSum"])
def test_segment_weights(engine, oracle, gen, monkeypatch, method):
    pm, arr = gen.gbt_pmml(T=29, depth=7, F=17, seed=14, method=method, segment_weights=True)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "compact=1" in m.describe
    X, pres, pb = gen.events(6000, 17, 15, 0.1)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)


def test_rf_majority_vote(engine, oracle, gen, monkeypatch):
    pm, arr = gen.rf_pmml(T=50, depth=6, F=20, n_labels=4, seed=21)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "compact=1" in m.describe
    X, pres, pb = gen.events(20000, 20, 22, 0.1)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)
    assert np.array_equal(compact["score"], gen.rf_numpy(arr, X, pres))


def test_pruned_trees_strict_tests_float_fields(engine, oracle, gen, monkeypatch):
    pm = gen.gbt_pruned_pmml(T=60, max_depth=6, F=20, seed=5)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "compact=1" in m.describe and "depth=6" in m.describe
    X, pres, pb = gen.events(20000, 20, 6, 0.1)
    X[::50, 2] = np.nan
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)


def test_invalid_values_fail_the_row(engine, oracle, gen, monkeypatch):
    """An Interval on a DataField makes values outside it invalid; the default invalidValueTreatment (returnInvalid) turns
    the row into EmptyScore (status 2).  The pre-pass flags those rows for the walk kernel."""
    pm, arr = gen.gbt_pmml(T=16, depth=4, F=9, seed=8)
    pm = pm.replace(b'<DataField name="x2" optype="continuous" dataType="double"/>',
                    b'<DataField name="x2" optype="continuous" dataType="double"><Interval closure="closedClosed" leftMargin="-1" rightMargin="1"/></DataField>')
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "prep_simple=0" in m.describe and "compact=1" in m.describe
    X, pres, pb = gen.events(9000, 9, 9, 0.1)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    o = om.predict(X, pb, nthreads=8)
    assert (o["status"] == 2).sum() > 1000 and (o["status"] == 0).sum() > 1000
    same(compact, o)
    same(compact, classic)


def test_device_entry_point_takes_the_path_for_large_batches(engine, gen):
    import torch
    from flink_jpmml_b200 import capi
    pm, arr = gen.gbt_pmml(T=40, depth=8, F=32, seed=31)
    m = engine.compile(pm)
    n = 148 * 1024 + 17
    X, pres, pb = gen.events(n, 32, 32, 0.1)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        dX = torch.from_numpy(X).cuda(); dP = torch.from_numpy(pb.view(np.int32)).cuda()
        dS = torch.empty(n, dtype=torch.float64, device="cuda"); dT = torch.empty(n, dtype=torch.uint8, device="cuda")
        st.synchronize()
        launches = engine.launch_count
        for _ in range(2):   # twice: the work area comes back from the stream-ordered pool
            rc = capi.lib().fjp_score_device(engine._h, m._h, dX.data_ptr(), dP.data_ptr(), None, n, 32, dS.data_ptr(), dT.data_ptr(), None, C.c_void_p(st.cuda_stream))
            assert rc == 0
        st.synchronize()
    assert engine.launch_count - launches == 4
    assert np.array_equal(dS.cpu().numpy(), gen.gbt_numpy(arr, X, pres)) and int(dT.sum().item()) == 0


def test_hand_derived_cases_with_the_path_forced(engine, gen, monkeypatch):
    """Every hand-derived semantic case (tests/semantic_cases.py) again with FJP_FOREST_COMPACT=2: ensembles that qualify
    (weighted sums / averages, votes, compound and set predicates, padded trees ...) go through the throughput path,
    everything else is unaffected; statuses and scores must still be the expected ones."""
    import semantic_cases
    monkeypatch.setenv("FJP_FOREST_COMPACT", "2")
    taken = 0
    for name, pm, X, P, rep, expected in semantic_cases.cases():
        m = engine.compile(pm)
        before = engine.launch_count
        r = engine.score(m, X, gen.pack_present(P), rep)
        taken += engine.launch_count - before >= 2 and "compact=1" in m.describe
        for i, (st, score, aux) in enumerate(expected):
            assert r["status"][i] == st, (name, i, r["status"])
            if st == 0 and name in semantic_cases.TOLERANT:
                assert abs(r["score"][i] - score) <= 1e-9 * abs(score), (name, i)
            elif st == 0:
                assert r["score"][i] == score, (name, i, r["score"][i], score)
    assert taken >= 2, taken


@pytest.mark.parametrize("dtype", ["double", "float"])
def test_thresholds_that_are_not_floats(engine, oracle, gen, monkeypatch, dtype):
    """scikit-learn / LightGBM style midpoints: thresholds are compared as doubles; the pre-pass tables then hold 64-bit
    images (compact=2).  float-typed fields parse their thresholds as floats and stay on the 32-bit tables."""
    monkeypatch.setattr(gen, "FLOAT_THRESHOLDS", False)
    pm, arr = gen.gbt_pmml(T=64, depth=6, F=24, seed=71, dtype=dtype)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert ("compact=2" if dtype == "double" else "compact=1") in m.describe, m.describe
    X, pres, pb = gen.events(30000, 24, 72, 0.1)
    X[::7, int(arr["feat"][0, 0])] = arr["thr"][0, 0]                      # exactly on a threshold
    X[::11, int(arr["feat"][1, 0])] = np.nextafter(arr["thr"][1, 0], np.inf)
    X[::17, int(arr["feat"][2, 0])] = np.nextafter(arr["thr"][2, 0], -np.inf)
    X[::13, 2] = np.nan
    X[::19, 3] = -np.inf
    X[::23, 4] = 1e300
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)


@pytest.mark.parametrize("n_classes", [2, 3])
def test_boosted_classifier_model_chain(engine, oracle, gen, monkeypatch, n_classes):
    """modelChain (trees per class -> logit / softmax RegressionModel): channel sums close in the walk kernel, the final
    tables run in its epilogue; labels bit-exact, probabilities identical to the one-kernel walk."""
    pm, arr = gen.gbt_classifier_pmml(T=40, depth=6, F=30, n_classes=n_classes, seed=51)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "method=9" in m.describe and "compact=1" in m.describe
    X, pres, pb = gen.events(20000, 30, 52, 0.1)
    monkeypatch.setenv("FJP_FOREST_COMPACT", "0")
    classic = engine.score(m, X, pb, n_probs=n_classes)
    monkeypatch.setenv("FJP_FOREST_COMPACT", "2")
    before = engine.launch_count
    compact = engine.score(m, X, pb, n_probs=n_classes)
    assert engine.launch_count - before >= 2
    o = om.predict(X, pb, n_probs=n_classes, nthreads=8)
    same(compact, o)
    same(compact, classic)
    assert np.array_equal(compact["probs"].view(np.int64), classic["probs"].view(np.int64))
    np.testing.assert_allclose(compact["probs"], o["probs"], rtol=1e-6)


@pytest.mark.parametrize("max_depth,p_leaf,floats,method", [(14, 0.3, True, "sum"), (12, 0.2, False, "average"), (9, 0.4, True, "sum")])
def test_unbalanced_trees_keep_their_shape(engine, oracle, gen, monkeypatch, max_depth, p_leaf, floats, method):
    """Leaf-wise / heavily pruned trees do not fit one common depth: the compiler keeps explicit children (sparse=1) and the
    throughput path walks them as they are; the general kernel (FJP_FOREST_COMPACT=0) is the one to compare with."""
    pm = gen.gbt_pruned_pmml(T=48, max_depth=max_depth, F=23, seed=max_depth, p_leaf=p_leaf, dtype="double",
                             operator="lessOrEqual", method=method, float_thresholds=floats)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "format=general" in m.describe and "sparse=1" in m.describe, m.describe
    X, pres, pb = gen.events(148 * 640 + 1000, 23, 8, 0.1)       # large enough to take the path unforced
    X[::50, 2] = np.nan
    classic, compact = score_both(engine, m, X[:20000], pb[:20000], monkeypatch)
    o = om.predict(X[:20000], pb[:20000], nthreads=8)
    same(compact, o)
    same(compact, classic)
    monkeypatch.delenv("FJP_FOREST_COMPACT")
    import torch
    from flink_jpmml_b200 import capi
    n = len(X); st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        dX = torch.from_numpy(X).cuda(); dP = torch.from_numpy(pb.view(np.int32)).cuda()
        dS = torch.empty(n, dtype=torch.float64, device="cuda"); dT = torch.empty(n, dtype=torch.uint8, device="cuda")
        st.synchronize()
        before = engine.launch_count
        assert capi.lib().fjp_score_device(engine._h, m._h, dX.data_ptr(), dP.data_ptr(), None, n, 23, dS.data_ptr(), dT.data_ptr(), None, C.c_void_p(st.cuda_stream)) == 0
        st.synchronize()
    assert engine.launch_count - before >= 2
    full = om.predict(X, pb, nthreads=8)
    assert np.array_equal(dS.cpu().numpy().view(np.int64), full["score"].view(np.int64)) and np.array_equal(dT.cpu().numpy(), full["status"])


def test_majority_vote_over_trees_deeper_than_the_common_depth_limit(engine, oracle, gen, monkeypatch):
    pm, arr = gen.rf_pmml(T=7, depth=11, F=16, n_labels=3, seed=77)       # depth 11 > 10: no common-depth layout
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "format=general" in m.describe and "sparse=1" in m.describe, m.describe
    X, pres, pb = gen.events(9000, 16, 78, 0.1)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)


def test_model_chain_over_trees_that_keep_their_shape(engine, oracle, gen, monkeypatch):
    pm, arr = gen.gbt_classifier_pmml(T=5, depth=11, F=18, n_classes=3, seed=91)     # depth 11: explicit-children form
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "method=9" in m.describe and "sparse=1" in m.describe, m.describe
    X, pres, pb = gen.events(7000, 18, 92, 0.1)
    monkeypatch.setenv("FJP_FOREST_COMPACT", "0")
    classic = engine.score(m, X, pb, n_probs=3)
    monkeypatch.setenv("FJP_FOREST_COMPACT", "2")
    before = engine.launch_count
    compact = engine.score(m, X, pb, n_probs=3)
    assert engine.launch_count - before >= 2
    o = om.predict(X, pb, n_probs=3, nthreads=8)
    same(compact, o)
    same(compact, classic)
    assert np.array_equal(compact["probs"].view(np.int64), classic["probs"].view(np.int64))
    np.testing.assert_allclose(compact["probs"], o["probs"], rtol=1e-6)


@pytest.mark.parametrize("n_labels,depth", [(3, 5), (4, 8), (7, 6), (13, 5)])
def test_random_forest_averaging_leaf_distributions(engine, oracle, gen, monkeypatch, n_labels, depth):
    """scikit-learn style forests: every leaf carries a ScoreDistribution, multipleModelMethod="average".  The compact
    trees carry the distributions inline; the winner is the largest summed probability, ties to the label met first."""
    pm, arr = gen.rf_pmml(T=30, depth=depth, F=12, n_labels=n_labels, seed=41 + n_labels, method="average")
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "method=7" in m.describe and "compact=1" in m.describe
    X, pres, pb = gen.events(12000, 12, 42, 0.1)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)


@pytest.mark.parametrize("n", [1100, 3000, 20000, 47000])
def test_mid_size_batches_of_ensembles_without_a_one_kernel_walk(engine, oracle, gen, n):
    """Averaged-distribution forests and explicit-children trees have no one-kernel walk; from 16 tiles on their batches take
    the throughput path unforced, with a tile small enough to spread over the SMs (host entry point, chunked)."""
    pm, arr = gen.rf_pmml(T=20, depth=6, F=10, n_labels=3, seed=5, method="average")
    pm2 = gen.gbt_pruned_pmml(T=20, max_depth=12, F=10, seed=6, p_leaf=0.3, dtype="double", operator="lessOrEqual")
    X, pres, pb = gen.events(n, 10, 43, 0.1)
    for doc in (pm, pm2):
        m, om = engine.compile(doc), oracle.OracleModel(doc)
        before = engine.launch_count
        r = engine.score(m, X, pb)
        assert engine.launch_count - before >= 2, m.describe
        same(r, om.predict(X, pb, nthreads=8))


@pytest.mark.parametrize("method", ["weightedMajorityVote", "weightedAverage"])
def test_weighted_classification_ensembles(engine, oracle, gen, monkeypatch, method):
    pm, arr = gen.rf_pmml(T=25, depth=6, F=14, n_labels=4, seed=61, method=method, segment_weights=True)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "compact=1" in m.describe
    X, pres, pb = gen.events(9000, 14, 62, 0.1)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)


@pytest.mark.parametrize("method,weights", [("average", False), ("weightedMajorityVote", True)])
def test_distribution_sums_over_trees_that_keep_their_shape(engine, oracle, gen, monkeypatch, method, weights):
    pm, arr = gen.rf_pmml(T=6, depth=11, F=12, n_labels=3, seed=83, method=method, segment_weights=weights)   # depth 11: explicit children
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "format=general" in m.describe and "sparse=1" in m.describe, m.describe
    X, pres, pb = gen.events(8000, 12, 84, 0.1)
    classic, compact = score_both(engine, m, X, pb, monkeypatch)
    same(compact, om.predict(X, pb, nthreads=8))
    same(compact, classic)
