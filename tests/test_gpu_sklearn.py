"""GPU: cross-checks against arithmetic this repository did not write.  scikit-learn fits a random forest (predict_proba =
average of the leaves' class distributions), a multinomial logistic regression and three MLPs (relu / tanh / logistic);
test-side emitters (tests/hand_pmml.py) write them as PMML; the CUDA path scores them through the C ABI and must give
scikit-learn's own labels (except where scikit-learn's top-2 probabilities differ by less than 1e-12) and, for the families
whose kernels report probabilities, scikit-learn's probabilities within 1e-6 relative.  The reference holds no fixture for
these families; this is the independent check for configs 3 and both halves of config 5."""
import numpy as np
import pytest

import hand_pmml

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def suite():
    pytest.importorskip("sklearn")
    return hand_pmml.sklearn_suite()


@pytest.mark.parametrize("k", range(5))
def test_gpu_matches_scikit_learn(engine, oracle, suite, k):
    name, pm, X, proba = suite[k]
    m = engine.compile(pm)
    has_probs = not name.startswith("rf")         # forest kernels report the label only
    r = engine.score(m, X, n_probs=proba.shape[1] if has_probs else 0)
    hand_pmml.check_against_sklearn(name, r, proba, has_probs)
    # and the same answers as the oracle, bit for bit on labels
    o = oracle.OracleModel(pm).predict(X, n_probs=proba.shape[1], nthreads=4)
    assert np.array_equal(r["status"], o["status"]) and np.array_equal(r["score"], o["score"]), name
    # a batch large enough for the throughput kernels (tiles over all SMs) must agree with the small one
    reps = 40
    Xb = np.tile(X, (reps, 1))
    rb = engine.score(m, Xb)
    assert np.array_equal(rb["score"], np.tile(r["score"], reps)) and (rb["status"] == 0).all(), name


def test_gpu_matches_scikit_learn_kmeans(engine, oracle, gen):
    """KMeans fitted by scikit-learn (config 2's family): the CUDA path -- the FMA path for the dense batch, the masked FMA path for
    rows with missing values -- gives scikit-learn's own cluster for every event whose two nearest centres are not a rounding tie."""
    pytest.importorskip("sklearn")
    from sklearn.cluster import KMeans
    rng = np.random.default_rng(21)
    F, K = 40, 16
    X = rng.standard_normal((3000, F)) + rng.integers(0, 3, (3000, 1))
    km = KMeans(n_clusters=K, n_init=2, random_state=0).fit(X)
    pm = gen.kmeans_pmml(km.cluster_centers_)
    Xq = rng.standard_normal((64 * 148 * 2 + 77, F)) + rng.integers(0, 3, (64 * 148 * 2 + 77, 1))
    d = ((Xq[:, None, :] - km.cluster_centers_[None, :, :]) ** 2).sum(axis=2); srt = np.sort(d, axis=1)
    clear = srt[:, 1] - srt[:, 0] > 1e-9
    m = engine.compile(pm)
    r = engine.score(m, Xq)
    assert (r["status"] == 0).all() and clear.mean() > 0.999
    assert np.array_equal(r["aux"][clear], km.predict(Xq)[clear] + 1)            # ids are 1-based cluster positions
    o = oracle.OracleModel(pm).predict(Xq, nthreads=8)
    assert np.array_equal(r["aux"], o["aux"]) and np.array_equal(r["score"], o["score"])
    # rows with missing values: JPMML compares over the present fields only (the adjustment factor scales every distance alike)
    pres = rng.random(Xq.shape) > 0.1; pb = gen.pack_present(pres)
    dm = (np.where(pres[:, None, :], (Xq[:, None, :] - km.cluster_centers_[None, :, :]) ** 2, 0.0)).sum(axis=2); srt = np.sort(dm, axis=1)
    clear = srt[:, 1] - srt[:, 0] > 1e-9
    rm = engine.score(m, Xq, pb)
    assert np.array_equal(rm["aux"][clear], np.argmin(dm, axis=1)[clear] + 1)
    om = oracle.OracleModel(pm).predict(Xq, pb, nthreads=8)
    assert np.array_equal(rm["aux"], om["aux"]) and np.array_equal(rm["status"], om["status"])


def test_gpu_matches_scikit_learn_gradient_boosting(engine, oracle, monkeypatch):
    """GradientBoostingRegressor fitted by scikit-learn (config 4's family: a sum of regression trees plus a constant): the CUDA path
    -- one-kernel walk and the rank-key pre-pass + compact walk -- within 1e-9 of scikit-learn's own predict (sklearn adds its stages
    in another order) and bit for bit with the oracle."""
    sk = pytest.importorskip("sklearn.ensemble")
    rng = np.random.default_rng(0)
    F = 10
    X = rng.standard_normal((3000, F)).astype(np.float32).astype(np.float64)
    y = X[:, 0] * 2 - X[:, 1] ** 2 + np.sin(X[:, 2] * 3) + rng.normal(0, 0.1, 3000)
    gb = sk.GradientBoostingRegressor(n_estimators=60, max_depth=5, random_state=0).fit(X, y)
    pm = hand_pmml.sklearn_gbr_to_pmml(gb, F)
    Xq = rng.standard_normal((640 * 148 + 333, F)).astype(np.float32).astype(np.float64)
    want = gb.predict(Xq)
    m = engine.compile(pm)
    o = oracle.OracleModel(pm).predict(Xq, nthreads=8)
    for compact in ("0", "2"):
        monkeypatch.setenv("FJP_FOREST_COMPACT", compact)
        r = engine.score(m, Xq)
        assert (r["status"] == 0).all()
        np.testing.assert_allclose(r["score"], want, rtol=1e-9, atol=1e-12)
        assert np.array_equal(r["score"].view(np.int64), o["score"].view(np.int64))
