"""GPU differential test: seeded random tree ensembles of random shape (tree count, depth, pruning, field count and type,
comparison operator, float / double thresholds, sum / average) against the oracle, through BOTH forest paths (the
one-kernel walk and the rank-key pre-pass + compact walk).  Scores, statuses and per-tree node ids bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

OPERATORS = ["lessThan", "lessOrEqual", "greaterThan", "greaterOrEqual"]


def params(seed):
    rng = np.random.default_rng(1000 + seed)
    return dict(T=int(rng.integers(1, 48)), max_depth=int(rng.integers(1, 14)), F=int(rng.integers(1, 48)), seed=seed,
                p_leaf=float(rng.choice([0.0, 0.15, 0.4])), operator=str(rng.choice(OPERATORS)), dtype=str(rng.choice(["float", "double"])),
                method=str(rng.choice(["sum", "average"])), float_thresholds=bool(rng.integers(0, 2)))


@pytest.mark.parametrize("seed", range(32))
def test_random_ensemble(engine, oracle, gen, monkeypatch, seed):
    kw = params(seed)
    pm = gen.gbt_pruned_pmml(**kw)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    F = kw["F"]; n = 2500 + 37 * seed
    X, pres, pb = gen.events(n, F, 77 + seed, 0.12)
    rng = np.random.default_rng(seed)
    from scipy.stats import norm
    grid = norm.ppf((np.arange(255) + 1.0) / 256.0)
    if kw["float_thresholds"]:
        grid = grid.astype(np.float32).astype(np.float64)
    hit = rng.random(X.shape) < 0.05                       # values exactly on thresholds, and one ulp either side
    X[hit] = rng.choice(grid, size=int(hit.sum()))
    up = rng.random(X.shape) < 0.01; X[up] = np.nextafter(rng.choice(grid, size=int(up.sum())), np.inf)
    dn = rng.random(X.shape) < 0.01; X[dn] = np.nextafter(rng.choice(grid, size=int(dn.sum())), -np.inf)
    X[rng.random(X.shape) < 0.003] = np.nan                # present NaN
    X[rng.random(X.shape) < 0.003] = np.inf
    X[rng.random(X.shape) < 0.003] = -0.0
    o = om.predict(X, pb, node_ids=True, nthreads=8)
    monkeypatch.setenv("FJP_FOREST_COMPACT", "0")
    r = engine.score(m, X, pb, node_ids=True)
    assert np.array_equal(r["status"], o["status"]), kw
    assert np.array_equal(r["score"].view(np.int64), o["score"].view(np.int64)), kw
    assert np.array_equal(r["node_ids"], o["node_ids"]), kw
    monkeypatch.setenv("FJP_FOREST_COMPACT", "2")
    c = engine.score(m, X, pb)
    assert np.array_equal(c["status"], o["status"]), kw
    assert np.array_equal(c["score"].view(np.int64), o["score"].view(np.int64)), kw
