"""GPU parity for the register-tiled FP64 kernels at shapes the named configs do not reach: several field chunks, ragged
tiles, present bits / replaceNan / invalid values, padded tables and layer widths.  Bar: labels and statuses bit-exact,
probabilities / distances within 1e-6 relative (in fact expected identical: sums run in the reference's order, unfused)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def check(r, o, probs=False):
    assert np.array_equal(r["status"], o["status"])
    ok = o["status"] == 0
    assert np.array_equal(r["score"][ok], o["score"][ok])
    assert np.isnan(r["score"][~ok]).all()
    if probs:
        np.testing.assert_allclose(r["probs"][ok], o["probs"][ok], rtol=1e-6)


@pytest.mark.parametrize("k,F", [(32, 150), (16, 40), (64, 128), (8, 70), (24, 33), (44, 64), (3, 9), (100, 20)])
def test_kmeans_shapes_dense_missing_and_replace(engine, oracle, gen, k, F):
    rng = np.random.default_rng(k * 1000 + F)
    pm = gen.kmeans_pmml(rng.standard_normal((k, F)))
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    n = 3 * 1024 + 77                                     # ragged last tile for every tile size in use
    X, pres, pb = gen.events(n, F, 5, 0.07)
    for args in ((np.nan_to_num(X), None, None), (X, pb, None), (X, pb, 0.125)):
        r, o = engine.score(m, *args, n_probs=1), om.predict(*args, n_probs=k, nthreads=8)
        check(r, o)
        assert np.array_equal(r["aux"], o["aux"])
        ok = o["status"] == 0                                                          # the engine reports the winning distance,
        win = o["probs"][np.arange(n), np.maximum(o["aux"], 1) - 1]                    # the oracle every cluster's
        np.testing.assert_allclose(r["probs"][ok, 0], win[ok], rtol=1e-12)
    rows_all_missing = np.zeros_like(pb); rows_all_missing[5:] = pb[5:]              # rows that supply nothing at all
    r, o = engine.score(m, X, rows_all_missing), om.predict(X, rows_all_missing, nthreads=8)
    check(r, o)


@pytest.mark.parametrize("F,C", [(64, 20), (200, 3), (97, 9), (1000, 20)])
def test_dense_regression_shapes(engine, oracle, gen, F, C):
    pm, arr = gen.regression_pmml(F=F, C=C, seed=F + C)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "dense=1" in m.describe
    n = 1500 + 13
    X, pres, pb = gen.events(n, F, 6, 0.002)              # a few rows lose a predictor: no result (status 3)
    r, o = engine.score(m, np.nan_to_num(X), n_probs=C), om.predict(np.nan_to_num(X), n_probs=C, nthreads=8)
    check(r, o, probs=True)
    r, o = engine.score(m, X, pb, n_probs=C), om.predict(X, pb, n_probs=C, nthreads=8)
    assert (o["status"] == 3).any() and (o["status"] == 0).any()
    check(r, o, probs=True)
    r, o = engine.score(m, X, pb, -0.5, n_probs=C), om.predict(X, pb, -0.5, n_probs=C, nthreads=8)   # replaceNan fills the gaps
    assert (o["status"] == 0).all()
    check(r, o, probs=True)


@pytest.mark.parametrize("n_in,hidden,n_out", [(64, (256, 256), 8), (10, (12,), 3), (33, (100, 7), 5), (5, (256,), 2)])
def test_mlp_shapes(engine, oracle, gen, n_in, hidden, n_out):
    pm, arr = gen.nn_pmml(n_in=n_in, hidden=hidden, n_out=n_out, seed=n_in)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    n = 64 * 37 + 5
    X, pres, pb = gen.events(n, n_in, 7, 0.01)
    r, o = engine.score(m, np.nan_to_num(X), n_probs=n_out), om.predict(np.nan_to_num(X), n_probs=n_out, nthreads=8)
    check(r, o, probs=True)
    r, o = engine.score(m, X, pb, n_probs=n_out), om.predict(X, pb, n_probs=n_out, nthreads=8)        # missing input: no result
    assert (o["status"] != 0).any()
    check(r, o, probs=True)
