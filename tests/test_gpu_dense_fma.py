"""GPU: the fused-multiply-add path of the dense tile kernel (softmax RegressionModel, k-means ClusteringModel) against the
oracle, through the C ABI.  Same contract as the tensor-core path: a label is decided on the fast arithmetic only when the
winner's margin exceeds the rounding bound between the two evaluation orders; everything else is re-scored by the kernel that
follows JPMML's order (unfused multiply and add, predictor order).  Labels / cluster ids bit-exact for every event,
probabilities within 1e-6 relative (S/api/PmmlModel.scala:159-160 hands these models to JPMML's RegressionModelEvaluator and
ClusteringModelEvaluator)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(engine, oracle, pm, X, n_probs=0):
    m = engine.compile(pm)
    c0 = engine.fast_path_counts
    r = engine.score(m, X, n_probs=n_probs)
    c1 = engine.fast_path_counts
    o = oracle.OracleModel(pm).predict(X, n_probs=n_probs, nthreads=os.cpu_count() or 4)
    assert np.array_equal(r["status"], o["status"]) and (r["status"] == 0).all()
    assert np.array_equal(r["score"], o["score"]), f"{(r['score'] != o['score']).sum()} labels differ from the oracle"
    assert np.array_equal(r["aux"], o["aux"])
    return r, o, c1[2] - c0[2], c1[3] - c0[3]


def test_c3_multinomial_on_the_fma_path(engine, gen, oracle, monkeypatch):
    """BASELINE.json config 3 (1000 features x 20 classes, softmax); probabilities too.  (The host entry point feeds the GPU in
    chunks of 4096 such rows, below the size at which the path switches on by itself: forced here, as bench.py's device-resident
    batches reach it unforced.)"""
    monkeypatch.setenv("FJP_DENSE_FMA", "2")
    pm, _ = gen.regression_pmml()
    X, _, _ = gen.events(64 * 148 + 333, 1000, 31, 0.0)
    r, o, took, back = _run(engine, oracle, pm, X, n_probs=20)
    assert took == len(X) and back < 0.01 * took, (took, back)
    np.testing.assert_allclose(r["probs"], o["probs"], rtol=1e-6, atol=1e-300)


@pytest.mark.parametrize("F,C,n", [(7, 3, 1), (64, 8, 63), (100, 11, 4097), (33, 24, 2000)])
def test_regression_shapes(engine, gen, oracle, monkeypatch, F, C, n):
    monkeypatch.setenv("FJP_DENSE_FMA", "2")
    pm, _ = gen.regression_pmml(F=F, C=C, seed=70 + F)
    X, _, _ = gen.events(n, F, 32, 0.0)
    r, o, took, back = _run(engine, oracle, pm, X, n_probs=C)
    assert took == n
    np.testing.assert_allclose(r["probs"], o["probs"], rtol=1e-6, atol=1e-300)


def test_identical_tables_are_rescored(engine, gen, oracle, monkeypatch):
    """Two RegressionTables with the same coefficients: an exact tie for every event where they lead -- only JPMML's order (the
    first maximal probability) may break it."""
    monkeypatch.setenv("FJP_DENSE_FMA", "2")
    import re
    pm, _ = gen.regression_pmml(F=20, C=4, seed=5)
    txt = pm.decode()
    tabs = re.findall(r'<RegressionTable intercept="[^"]*" targetCategory="[^"]*">.*?</RegressionTable>', txt, re.S)
    body0 = re.sub(r'targetCategory="[^"]*"', 'targetCategory="1"', tabs[0])
    txt = txt.replace(tabs[1], body0)
    X, _, _ = gen.events(3000, 20, 33, 0.0)
    r, o, took, back = _run(engine, oracle, txt.encode(), X)
    lead = np.isin(o["aux"], (0, 1)).sum()
    assert back >= lead > 0 and (o["aux"] != 1).all()


def test_c2_kmeans_on_the_fma_path(engine, gen, oracle):
    """BASELINE.json config 2 (k = 64, 128 fields): cluster ids bit-exact with the oracle on a batch that fills the chip."""
    pm, arr = gen.config_c2()
    X, _, _ = gen.events(64 * 148 * 2 + 91, 128, 34, 0.0)
    r, o, took, back = _run(engine, oracle, pm, X)
    assert took == len(X) and back < 0.001 * took, (took, back)
    assert np.array_equal(r["aux"].astype(np.float64), gen.kmeans_numpy(arr["centroids"], X))


@pytest.mark.parametrize("k,F,n", [(2, 3, 5), (4, 4, 1000), (13, 50, 3000), (64, 128, 129)])
def test_kmeans_shapes_and_ties(engine, gen, oracle, monkeypatch, k, F, n):
    monkeypatch.setenv("FJP_DENSE_FMA", "2")
    rng = np.random.default_rng(k * 100 + F)
    cent = rng.standard_normal((k, F))
    if k >= 4:
        cent[3] = cent[1]                    # a duplicated centre: the first of the two must win every tie
    pm = gen.kmeans_pmml(cent)
    X = rng.standard_normal((n, F))
    X[: min(n, 50)] = cent[rng.integers(0, k, min(n, 50))]      # events sitting exactly on a centre
    r, o, took, back = _run(engine, oracle, pm, X)
    assert took == n
    if k >= 4:
        assert (o["aux"] != 4).all()


@pytest.mark.parametrize("k,F,n,p_missing", [(64, 128, 64 * 148 * 2 + 91, 0.05), (13, 50, 3000, 0.3), (8, 32, 2000, 0.9), (20, 100, 4000, 0.02)])
def test_kmeans_with_missing_values_on_the_fma_path(engine, gen, oracle, monkeypatch, k, F, n, p_missing):
    """BASELINE.json config 2, 5 %-missing variant (and sparser / denser masks, a field count that is no multiple of 32, events
    with no field at all): a missing field is x = 0 in the dot products and its c^2 comes back out of |c|^2; cluster ids
    bit-exact with the oracle, which skips the field and applies JPMML's adjustment factor."""
    monkeypatch.setenv("FJP_DENSE_FMA", "2")
    rng = np.random.default_rng(k * 1000 + F)
    cent = rng.standard_normal((k, F))
    if k >= 4:
        cent[3] = cent[1]
    pm = gen.kmeans_pmml(cent)
    X, pres, pb = gen.events(n, F, 35, p_missing)
    X[~pres] = np.nan                                             # whatever sits in a missing slot must not matter
    X[:20] = cent[rng.integers(0, k, 20)]; pres[:20] = True       # on a centre
    pres[20:30] = False                                           # no field at all
    pb = gen.pack_present(pres)
    m = engine.compile(pm)
    c0 = engine.fast_path_counts
    r = engine.score(m, X, pb)
    c1 = engine.fast_path_counts
    o = oracle.OracleModel(pm).predict(X, pb, nthreads=os.cpu_count() or 4)
    assert np.array_equal(r["status"], o["status"])
    ok = o["status"] == 0
    assert np.array_equal(r["score"][ok], o["score"][ok]) and np.array_equal(r["aux"][ok], o["aux"][ok])
    took, back = c1[2] - c0[2], c1[3] - c0[3]
    assert took == n and 10 <= back < (0.05 if p_missing < 0.5 else 0.5) * n + 40, (took, back)   # back: the events with no field, exact ties, thin margins
