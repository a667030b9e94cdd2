"""GPU: the tensor-core path of dense layers (csrc/sliced_gemm.cuh: tcgen05.mma kind::i8 on digit planes, margin test, FP64
re-scoring) against the oracle, through the C ABI.

The contract under test: class labels (score, status, aux) are those of the FP64 evaluation in JPMML's order -- bit-exact
with the oracle -- for EVERY event: the path decides a label itself only when the winner's margin exceeds its proven error
bound, and hands everything else (thin margins, missing / invalid / non-finite inputs) to the FP64 kernel.  Reference
hand-off being replaced: S/api/PmmlModel.scala:159-160 (JPMML NeuralNetworkEvaluator)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _check(engine, oracle, pm, X, pb=None, replace=None, min_direct=0.0):
    m = engine.compile(pm)
    t0, f0 = engine.tensor_path_counts
    r = engine.score(m, X, pb, replace)
    t1, f1 = engine.tensor_path_counts
    o = oracle.OracleModel(pm).predict(X, pb, replace, nthreads=os.cpu_count() or 4)
    assert np.array_equal(r["status"], o["status"]), f"{(r['status'] != o['status']).sum()} statuses differ"
    ok = o["status"] == 0
    assert np.array_equal(r["score"][ok], o["score"][ok]), f"{(r['score'][ok] != o['score'][ok]).sum()} labels differ from the oracle"
    assert np.array_equal(r["aux"][ok], o["aux"][ok])
    took, handed_back = t1 - t0, f1 - f0
    assert took == len(X), "the tensor-core path did not take this batch"
    assert handed_back <= (1.0 - min_direct) * took + 1, f"{handed_back} of {took} events were handed back to the FP64 kernel"
    return took, handed_back


def test_c5_mlp_labels_bit_exact_on_the_tensor_path(engine, gen, oracle):
    """BASELINE.json config 5's MLP (64-256-256-8, rectifier, softmax output) on a batch that fills the chip."""
    pm, _ = gen.nn_pmml()
    X, _, _ = gen.events(148 * 128 * 2 + 77, 64, 21, 0.0)
    took, back = _check(engine, oracle, pm, X, min_direct=0.99)
    print(f"tensor path: {took} events, {back} re-scored in FP64 ({back / took:.2e})")


@pytest.mark.parametrize("n", [1, 127, 128, 129, 5000])
def test_ragged_batches(engine, gen, oracle, monkeypatch, n):
    monkeypatch.setenv("FJP_SLICED", "2")          # take the path whatever the batch size
    pm, _ = gen.nn_pmml(n_in=64, hidden=(256, 256), n_out=8, seed=31)
    X, _, _ = gen.events(n, 64, 22 + n, 0.0)
    _check(engine, oracle, pm, X)


@pytest.mark.parametrize("act,hidden,n_in,n_out", [("rectifier", (32,), 20, 3), ("tanh", (24, 16), 12, 4), ("logistic", (100, 60, 30), 70, 5),
                                                    ("identity", (48, 48), 33, 16)])
def test_layer_shapes_and_activations(engine, gen, oracle, monkeypatch, act, hidden, n_in, n_out):
    """Widths that are not multiples of the 32-neuron / 64-input chunks, one to three hidden layers, every supported activation."""
    monkeypatch.setenv("FJP_SLICED", "2")
    pm, _ = gen.nn_pmml(n_in=n_in, hidden=hidden, n_out=n_out, seed=41, activation=act)
    X, _, _ = gen.events(3000, n_in, 23, 0.0)
    _check(engine, oracle, pm, X, min_direct=0.9)


def test_missing_and_wild_inputs_are_rescored(engine, gen, oracle, monkeypatch):
    """Missing inputs (no result: status 3), replaceNan, huge and tiny magnitudes, NaN / Inf: the path must hand over exactly
    the events it cannot bound, and the FP64 kernel's answer is what comes back."""
    monkeypatch.setenv("FJP_SLICED", "2")
    pm, _ = gen.nn_pmml(n_in=64, hidden=(256, 256), n_out=8, seed=51)
    X, pres, pb = gen.events(6000, 64, 24, 0.02)
    X[10, 3] = np.nan; X[11, 5] = np.inf; X[12, :] *= 1e150; X[13, :] *= 1e-200; X[14, :] = 0.0; X[15, 7] = -np.inf; X[16, :] *= 1e6
    _check(engine, oracle, pm, X, pb)
    _check(engine, oracle, pm, X, pb, replace=0.25)


def test_near_ties_go_to_the_fp64_kernel(engine, gen, oracle, monkeypatch):
    """Two output neurons with IDENTICAL weights: every event is an exact tie that only the FP64 evaluation order may break
    (first maximal output).  The path must not decide any of them."""
    monkeypatch.setenv("FJP_SLICED", "2")
    pm, arr = gen.nn_pmml(n_in=16, hidden=(32,), n_out=4, seed=61)
    txt = pm.decode()
    # copy neuron "2,0" onto neuron "2,1" (same bias, same Con weights)
    import re
    n0 = re.search(r'<Neuron id="2,0" bias="([^"]+)">(.*?)</Neuron>', txt, re.S)
    txt = re.sub(r'<Neuron id="2,1" bias="[^"]+">.*?</Neuron>', f'<Neuron id="2,1" bias="{n0.group(1)}">{n0.group(2)}</Neuron>', txt, flags=re.S)
    X, _, _ = gen.events(4000, 16, 25, 0.0)
    m = engine.compile(txt.encode())
    t0, f0 = engine.tensor_path_counts
    r = engine.score(m, X)
    t1, f1 = engine.tensor_path_counts
    o = oracle.OracleModel(txt.encode()).predict(X, nthreads=4)
    assert np.array_equal(r["score"], o["score"]) and np.array_equal(r["status"], o["status"])
    winners_01 = np.isin(o["aux"], (0, 1)).sum()
    assert f1 - f0 >= winners_01 > 0, "tied events were decided by the tensor path"
    assert (o["aux"] != 1).all()          # the first of two equal outputs wins


# ---- the fused single-layer variant: softmax RegressionModels over wide plain dense rows (config 3) ---------------------------
def _check_reg(engine, oracle, pm, X, expect_taken=True):
    m = engine.compile(pm)
    t0, f0 = engine.tensor_path_counts
    r = engine.score(m, X)
    t1, f1 = engine.tensor_path_counts
    o = oracle.OracleModel(pm).predict(X, nthreads=os.cpu_count() or 4)
    assert np.array_equal(r["status"], o["status"]) and (o["status"] == 0).all()
    assert np.array_equal(r["score"], o["score"]), f"{(r['score'] != o['score']).sum()} labels differ from the oracle"
    assert np.array_equal(r["aux"], o["aux"])
    if expect_taken:
        assert t1 - t0 == len(X), "the fused tensor-core layer did not take this batch"
    return t1 - t0, f1 - f0


def test_c3_multinomial_labels_bit_exact_on_the_fused_tensor_layer(engine, gen, oracle, monkeypatch):
    """BASELINE.json config 3 (1000 fields x 20 classes, softmax): rows are read once, sliced per (event, 64-field chunk) with
    their own scales, multiplied on the tensor cores; labels must be the oracle's for every event."""
    monkeypatch.setenv("FJP_SLICED", "2")
    pm, _ = gen.regression_pmml()
    X, _, _ = gen.events(148 * 128 + 55, 1000, 41, 0.0)
    took, back = _check_reg(engine, oracle, pm, X)
    assert back < 0.01 * took, (took, back)
    print(f"fused tensor layer: {took} events, {back} re-scored in FP64")


@pytest.mark.parametrize("F,C,n", [(128, 2, 1), (130, 24, 777), (256, 3, 4096), (1000, 20, 129), (2050, 7, 300)])
def test_fused_layer_shapes(engine, gen, oracle, monkeypatch, F, C, n):
    monkeypatch.setenv("FJP_SLICED", "2")
    pm, _ = gen.regression_pmml(F=F, C=C, seed=80 + C)
    X, _, _ = gen.events(n, F, 42, 0.0)
    _check_reg(engine, oracle, pm, X)


def test_fused_layer_wild_values_and_scales(engine, gen, oracle, monkeypatch):
    """Chunks with very different magnitudes inside one row, zeros, denormals, NaN / Inf (those events are re-scored)."""
    monkeypatch.setenv("FJP_SLICED", "2")
    pm, _ = gen.regression_pmml(F=512, C=6, seed=90)
    X, _, _ = gen.events(5000, 512, 43, 0.0)
    X[:, 64:128] *= 1e-9; X[:, 128:192] *= 1e6; X[:, 192:256] = 0.0; X[100:200, 256:320] *= 1e-300
    X[7, 3] = np.nan; X[8, 500] = np.inf; X[9, :] = 0.0; X[10, 70] = 1e250
    _check_reg(engine, oracle, pm, X)
    # an odd field count keeps the model on the FP64 tiles (rows are not 16-byte aligned): still the oracle's answers
    pm2, _ = gen.regression_pmml(F=257, C=4, seed=91)
    X2, _, _ = gen.events(2000, 257, 44, 0.0)
    _check_reg(engine, oracle, pm2, X2, expect_taken=False)
