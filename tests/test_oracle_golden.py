"""CPU: pins the oracle (oracle/oracle.cc) against every known-answer vector the reference's own tests hold for the
hot path, and against independent numpy restatements for the model families the reference never tests.

Reference-pinned (k-means only): api/PmmlModelSpec.scala:50-83, :131-136, :329-347; api/converter/VectorConverterSpec.scala:76-78;
RichConnectedStreamSpec.scala:97-116; RichDataStreamSpec.scala:98-105.  (paths under flink-jpmml-scala/src/test/scala/io/radicalbit/flink/pmml/scala/)"""
import numpy as np
import pytest

from conftest import golden_bytes, vector_to_batch


def test_golden_predict_vectors(oracle, gen, golden_vectors):
    for v in golden_vectors["predict"]:
        m = oracle.OracleModel(golden_bytes(v["model"]))
        X, pb, rep = vector_to_batch(gen, v)
        r = m.predict(X, pb, rep)
        if v["score"] is None:
            assert r["status"][0] != 0 and np.isnan(r["score"][0]), v
        else:
            assert r["status"][0] == 0 and r["score"][0] == v["score"], (v, r)


def test_active_field_order(oracle, golden_vectors):
    m = oracle.OracleModel(golden_bytes("iris_kmeans43.pmml"))
    assert m.active_fields == golden_vectors["active_fields"]["value"]
    assert m.model_name == "k-means"


def test_loadable_versions_and_empty_document(oracle, golden_vectors):
    for name in golden_vectors["loadable"]["value"] + ["iris_kmeans32.pmml"]:
        m = oracle.OracleModel(golden_bytes(name))
        assert m.predict(np.ones((1, 4)))["score"][0] == 3.0
    for name in golden_vectors["not_loadable"]["value"]:
        with pytest.raises(oracle.OracleLoadError):
            oracle.OracleModel(golden_bytes(name))


def test_status_taxonomy(oracle):
    m = oracle.OracleModel(golden_bytes("iris_kmeans43.pmml"))
    assert m.predict(np.ones((1, 3)))["status"][0] == 1                      # size mismatch
    s = oracle.OracleModel(golden_bytes("iris_kmeans_stringfields.pmml"))
    assert s.predict(np.ones((1, 4)))["status"][0] == 2                      # string field fed a Double
    t = oracle.OracleModel(golden_bytes("iris_kmeans_nooutput_notarget.pmml"))
    assert t.predict(np.ones((1, 4)))["status"][0] == 4                      # no named target


def test_kmeans_matches_numpy(oracle, gen):
    pm, arr = gen.config_c2(k=16, F=24)
    m = oracle.OracleModel(pm)
    X, pres, pb = gen.events(3000, 24, 1, 0.1)
    r = m.predict(X, pb)
    assert (r["status"] == 0).all()
    assert np.array_equal(r["score"], gen.kmeans_numpy(arr["centroids"], X, pres))
    assert np.array_equal(r["aux"], r["score"].astype(np.int32))


@pytest.mark.parametrize("dtype", ["double", "float"])
def test_gbt_matches_numpy(oracle, gen, dtype):
    pm, arr = gen.gbt_pmml(T=25, depth=5, F=12, seed=3, dtype=dtype)
    m = oracle.OracleModel(pm)
    X, pres, pb = gen.events(4000, 12, 4, 0.1)
    r = m.predict(X, pb, node_ids=True)
    assert (r["status"] == 0).all()
    assert np.array_equal(r["score"], gen.gbt_numpy(arr, X, pres))
    # implicit node numbering: leaf position -> pre-order id, checked on the left-most / right-most leaves
    li = gen.forest_leaf_index_numpy(arr, X, pres)
    assert (r["node_ids"][li == 0] == 6).all()            # root(1) -> 2 -> 3 -> 4 -> 5 -> leaf 6 at depth 5
    assert (r["node_ids"][li == 31] == 63).all()          # last node of a 63-node tree


def test_rf_regression_nn_match_numpy(oracle, gen):
    pm, arr = gen.rf_pmml(T=31, depth=4, F=9, n_labels=3, seed=5)
    X, pres, pb = gen.events(3000, 9, 6, 0.1)
    r = oracle.OracleModel(pm).predict(X, pb)
    assert np.array_equal(r["score"], gen.rf_numpy(arr, X, pres))
    pm, arr = gen.rf_pmml(T=15, depth=4, F=9, n_labels=3, seed=6, method="average")
    r = oracle.OracleModel(pm).predict(X, pb, n_probs=3)
    lab, avg = gen.rf_average_numpy(arr, X, pres)
    assert (r["score"] == lab).all()
    pm, arr = gen.regression_pmml(F=40, C=6, seed=7)
    X, pres, pb = gen.events(1000, 40, 8)
    r = oracle.OracleModel(pm).predict(X, n_probs=6)
    y = gen.regression_numpy(arr, X)
    assert np.array_equal(r["score"], y.argmax(1).astype(float))
    np.testing.assert_allclose(r["probs"], np.exp(y) / np.exp(y).sum(1, keepdims=True), rtol=1e-12)
    pm, arr = gen.nn_pmml(n_in=7, hidden=(9, 5), n_out=4, seed=9)
    X, pres, pb = gen.events(1000, 7, 10)
    r = oracle.OracleModel(pm).predict(X, n_probs=4)
    z = gen.nn_numpy(arr, X)
    assert np.array_equal(r["score"], z.argmax(1).astype(float))
    np.testing.assert_allclose(r["probs"], np.exp(z) / np.exp(z).sum(1, keepdims=True), rtol=1e-12)


def test_sklearn_cross_check(oracle, gen):
    """Independent arithmetic: a scikit-learn GradientBoostingRegressor exported through our emitter's tree layout."""
    sk = pytest.importorskip("sklearn.ensemble")
    rng = np.random.default_rng(0)
    X = rng.standard_normal((600, 6)).astype(np.float32).astype(np.float64)
    y = X[:, 0] * 2 - X[:, 1] ** 2 + rng.normal(0, 0.1, 600)
    gb = sk.GradientBoostingRegressor(n_estimators=12, max_depth=3, random_state=0).fit(X, y)
    from hand_pmml import sklearn_gbr_to_pmml
    pm = sklearn_gbr_to_pmml(gb, 6)
    r = oracle.OracleModel(pm).predict(X)
    assert (r["status"] == 0).all()
    np.testing.assert_allclose(r["score"], gb.predict(X), rtol=1e-9)


def test_sklearn_classifiers_cross_check(oracle):
    """Oracle vs scikit-learn's own predict_proba: random forest (averaged leaf distributions), multinomial logistic
    regression, MLPs with relu / tanh / logistic hidden layers (tests/hand_pmml.py exports them)."""
    pytest.importorskip("sklearn")
    import hand_pmml
    for name, pm, X, proba in hand_pmml.sklearn_suite():
        r = oracle.OracleModel(pm).predict(X, n_probs=proba.shape[1], nthreads=4)
        if name.startswith("rf"):   # the oracle lists a forest's probabilities in first-met label order: compare the winner's
            hand_pmml.check_against_sklearn(name, r, proba, False)
            np.testing.assert_allclose(np.nanmax(r["probs"], axis=1), proba.max(axis=1), rtol=1e-9)
        else:
            hand_pmml.check_against_sklearn(name, r, proba, True)


@pytest.mark.parametrize("n_classes", [2, 4])
def test_model_chain_classifier_matches_numpy(oracle, gen, n_classes):
    pm, arr = gen.gbt_classifier_pmml(T=10, depth=4, F=8, n_classes=n_classes, seed=3)
    X, pres, pb = gen.events(3000, 8, 4, 0.1)
    r = oracle.OracleModel(pm).predict(X, pb, n_probs=n_classes)
    lab, probs = gen.gbt_classifier_numpy(arr, X, pres)
    assert (r["status"] == 0).all() and np.array_equal(r["score"], lab)
    np.testing.assert_allclose(r["probs"], probs, rtol=1e-12)
