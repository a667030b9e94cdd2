"""Deterministic PMML 4.3 emitters for the synthetic configs of BASELINE.json / SURVEY.md §8(d).

The reference ships only Iris k-means documents (flink-jpmml-assets/src/main/resources/kmeans*.xml);
every other model family it can serve exists there only through JPMML.  These emitters produce the
documents for configs C2..C5 plus small adversarial cases, and return the raw parameter arrays next to
the XML so tests can cross-check the oracle and the CUDA path against plain numpy arithmetic
(a third, independent restatement).

Seeds follow SURVEY.md §8(d): 20261106 + config number.
"""
from __future__ import annotations

import numpy as np

SEED_BASE = 20261106
NS = "http://www.dmg.org/PMML-4_3"


def _r(x) -> str:
    """Round-trip decimal text of a double."""
    return repr(float(x))


def _header(version="4.3", ns=NS) -> list[str]:
    return ['<?xml version="1.0" encoding="UTF-8"?>',
            f'<PMML version="{version}" xmlns="{ns}">',
            '<Header><Application name="flink-jpmml-b200-gen" version="1"/></Header>']


def _data_dictionary(fields, target=None) -> list[str]:
    """fields: list of (name, optype, dataType); target: (name, optype, dataType, [values])."""
    n = len(fields) + (1 if target else 0)
    out = [f'<DataDictionary numberOfFields="{n}">']
    for name, optype, dtype in fields:
        out.append(f'<DataField name="{name}" optype="{optype}" dataType="{dtype}"/>')
    if target:
        name, optype, dtype, values = target
        if values:
            out.append(f'<DataField name="{name}" optype="{optype}" dataType="{dtype}">')
            out.extend(f'<Value value="{v}"/>' for v in values)
            out.append('</DataField>')
        else:
            out.append(f'<DataField name="{name}" optype="{optype}" dataType="{dtype}"/>')
    out.append('</DataDictionary>')
    return out


def _mining_schema(names, target=None, extra_attr="") -> list[str]:
    out = ['<MiningSchema>']
    if target:
        out.append(f'<MiningField name="{target}" usageType="target"/>')
    out.extend(f'<MiningField name="{n}"{extra_attr}/>' for n in names)
    out.append('</MiningSchema>')
    return out


# --------------------------------------------------------------------------------------------
# ClusteringModel (C1-shaped small cases, C2)
# --------------------------------------------------------------------------------------------

def kmeans_pmml(centroids: np.ndarray, field_names=None, target="cluster", measure="squaredEuclidean",
                cluster_ids=None, field_weights=None, missing_weights=None, compare="absDiff",
                kind="distance", minkowski_p=None):
    """Centre-based ClusteringModel.  Returns xml bytes."""
    k, F = centroids.shape
    names = field_names or [f"x{i}" for i in range(F)]
    out = _header()
    out += _data_dictionary([(n, "continuous", "double") for n in names],
                            (target, "categorical", "string", None) if target else None)
    out.append(f'<ClusteringModel modelName="kmeans-k{k}-f{F}" functionName="clustering" '
               f'modelClass="centerBased" numberOfClusters="{k}">')
    out.append('<MiningSchema>')
    out.extend(f'<MiningField name="{n}"/>' for n in names)
    if target:
        out.append(f'<MiningField name="{target}" usageType="predicted"/>')
    out.append('</MiningSchema>')
    m = f'<{measure} p-parameter="{_r(minkowski_p)}"/>' if measure == "minkowski" else f'<{measure}/>'
    out.append(f'<ComparisonMeasure kind="{kind}" compareFunction="{compare}">{m}</ComparisonMeasure>')
    for i, n in enumerate(names):
        w = f' fieldWeight="{_r(field_weights[i])}"' if field_weights is not None else ""
        out.append(f'<ClusteringField field="{n}"{w}/>')
    if missing_weights is not None:
        arr = " ".join(_r(v) for v in missing_weights)
        out.append(f'<MissingValueWeights><Array n="{F}" type="real">{arr}</Array></MissingValueWeights>')
    for c in range(k):
        cid = f' id="{cluster_ids[c]}"' if cluster_ids is not None else ""
        arr = " ".join(_r(v) for v in centroids[c])
        out.append(f'<Cluster name="cluster_{c}"{cid}><Array n="{F}" type="real">{arr}</Array></Cluster>')
    out.append('</ClusteringModel></PMML>')
    return "\n".join(out).encode()


def config_c2(k=64, F=128, seed=SEED_BASE + 2):
    rng = np.random.default_rng(seed)
    cent = rng.standard_normal((k, F))
    return kmeans_pmml(cent), {"centroids": cent}


def kmeans_numpy(cent, X, present=None):
    """1-based argmin of sum_i (x_i - z_i)^2 over present fields (first wins); plain numpy restatement."""
    n = X.shape[0]
    if present is None:
        present = np.ones_like(X, dtype=bool)
    best = np.zeros(n, dtype=np.int64)
    bestd = np.full(n, np.inf)
    for c in range(cent.shape[0]):
        d = np.zeros(n)
        for i in range(cent.shape[1]):  # sequential field order, like the interpreter
            t = X[:, i] - cent[c, i]
            d = np.where(present[:, i], d + t * t, d)
        better = d < bestd
        best = np.where(better, c, best)
        bestd = np.where(better, d, bestd)
    return (best + 1).astype(np.float64)


# --------------------------------------------------------------------------------------------
# Tree ensembles (C4 GBT, C5 RF)
# --------------------------------------------------------------------------------------------

def _complete_tree_xml(out, feat, thr, dleft, leaf_txt, names, depth, operator="lessOrEqual", with_default=True):
    """Emit one complete binary tree (level-order arrays) as nested <Node>s.

    Left child carries the SimplePredicate, right child <True/> (sklearn2pmml / LightGBM export style).
    Node ids are 1-based pre-order numbers, which is also JPMML's implicit numbering.
    """
    n_internal = (1 << depth) - 1
    counter = [0]

    def emit(idx, pred):
        counter[0] += 1
        my_id = counter[0]
        if idx >= n_internal:
            out.append(f'<Node id="{my_id}" score="{leaf_txt[idx - n_internal]}">{pred}</Node>')
            return my_id
        # ids of children are known from the pre-order size of the left subtree
        level = (idx + 1).bit_length() - 1
        sub = (1 << (depth - level)) - 1  # nodes in each child subtree incl. leaves: 2^(depth-level)-1
        left_id, right_id = my_id + 1, my_id + 1 + sub
        dc = f' defaultChild="{left_id if dleft[idx] else right_id}"' if with_default else ""
        out.append(f'<Node id="{my_id}"{dc}>{pred}')
        p = f'<SimplePredicate field="{names[feat[idx]]}" operator="{operator}" value="{_r(thr[idx])}"/>'
        emit(2 * idx + 1, p)
        emit(2 * idx + 2, '<True/>')
        out.append('</Node>')
        return my_id

    emit(0, '<True/>')


FLOAT_THRESHOLDS = True   # tests flip this to get thresholds that are NOT float32-representable (rank-encoded keys)


def forest_arrays(T, depth, F, seed, n_quantiles=255):
    """Random complete trees: feature ~ U{0..F-1}; thresholds = standard-normal quantiles rounded to float32
    (exactly representable in both float and double); default child uniform; returns level-order arrays."""
    from scipy.stats import norm
    rng = np.random.default_rng(seed)
    n_int = (1 << depth) - 1
    feat = rng.integers(0, F, size=(T, n_int))
    grid = norm.ppf((np.arange(n_quantiles) + 1.0) / (n_quantiles + 1.0))
    if FLOAT_THRESHOLDS:
        grid = grid.astype(np.float32).astype(np.float64)
    thr = grid[rng.integers(0, n_quantiles, size=(T, n_int))]
    dleft = rng.integers(0, 2, size=(T, n_int)).astype(bool)
    return feat, thr, dleft, rng


def gbt_pmml(T=500, depth=8, F=100, seed=SEED_BASE + 4, dtype="double", method="sum",
             rescale_constant=0.5, rescale_factor=None, missing_strategy="defaultChild", segment_weights=False):
    """MiningModel regression, Segmentation sum|average over T complete depth-`depth` TreeModels (config C4)."""
    feat, thr, dleft, rng = forest_arrays(T, depth, F, seed)
    leaf = rng.normal(0.0, 0.1, size=(T, 1 << depth))
    names = [f"x{i}" for i in range(F)]
    out = _header()
    out += _data_dictionary([(n, "continuous", dtype) for n in names], ("y", "continuous", "double", None))
    out.append(f'<MiningModel modelName="gbt-{T}x{depth}" functionName="regression">')
    out += _mining_schema(names, "y")
    if rescale_constant is not None or rescale_factor is not None:
        a = f' rescaleConstant="{_r(rescale_constant)}"' if rescale_constant is not None else ""
        b = f' rescaleFactor="{_r(rescale_factor)}"' if rescale_factor is not None else ""
        out.append(f'<Targets><Target field="y"{a}{b}/></Targets>')
    out.append(f'<Segmentation multipleModelMethod="{method}">')
    weights = np.round(rng.uniform(0.5, 1.5, size=T), 3) if segment_weights else None   # weightedSum / weightedAverage
    for t in range(T):
        out.append(f'<Segment id="{t + 1}"' + (f' weight="{_r(weights[t])}"' if segment_weights else "") + '><True/>')
        out.append(f'<TreeModel functionName="regression" missingValueStrategy="{missing_strategy}" '
                   'noTrueChildStrategy="returnNullPrediction" splitCharacteristic="binarySplit">')
        out += _mining_schema(names)
        _complete_tree_xml(out, feat[t], thr[t], dleft[t], [_r(v) for v in leaf[t]], names, depth,
                           with_default=(missing_strategy == "defaultChild"))
        out.append('</TreeModel></Segment>')
    out.append('</Segmentation></MiningModel></PMML>')
    arrays = {"feat": feat, "thr": thr, "dleft": dleft, "leaf": leaf, "depth": depth, "method": method,
              "rescale_constant": rescale_constant or 0.0, "rescale_factor": 1.0 if rescale_factor is None else rescale_factor,
              "dtype": dtype, "weights": weights}
    return "\n".join(out).encode(), arrays


def rf_pmml(T=200, depth=8, F=64, n_labels=4, seed=SEED_BASE + 5, target_dtype="string", method="majorityVote",
            segment_weights=False):
    """MiningModel classification over T complete TreeModels whose leaves carry labels "0".."L-1".
    method="majorityVote" (labels only) or "average" (scikit-learn style: every leaf has a ScoreDistribution of record
    counts and the ensemble averages the leaf probabilities)."""
    feat, thr, dleft, rng = forest_arrays(T, depth, F, seed)
    leaf = rng.integers(0, n_labels, size=(T, 1 << depth))
    counts = rng.integers(0, 20, size=(T, 1 << depth, n_labels)) if method in ("average", "weightedAverage") else None
    if counts is not None:
        counts[np.arange(T)[:, None], np.arange(1 << depth)[None, :], leaf] += 21      # the leaf's label has the most records
    weights = np.round(rng.uniform(0.5, 1.5, size=T), 3) if segment_weights else None   # drawn last: earlier draws keep their values
    names = [f"x{i}" for i in range(F)]
    labels = [str(i) for i in range(n_labels)]
    out = _header()
    out += _data_dictionary([(n, "continuous", "double") for n in names], ("y", "categorical", target_dtype, labels))
    out.append(f'<MiningModel modelName="rf-{T}x{depth}" functionName="classification">')
    out += _mining_schema(names, "y")
    out.append(f'<Segmentation multipleModelMethod="{method}">')
    for t in range(T):
        out.append(f'<Segment id="{t + 1}"' + (f' weight="{_r(weights[t])}"' if segment_weights else "") + '><True/>')
        out.append('<TreeModel functionName="classification" missingValueStrategy="defaultChild" '
                   'noTrueChildStrategy="returnNullPrediction" splitCharacteristic="binarySplit">')
        out += _mining_schema(names)
        n0 = len(out)
        _complete_tree_xml(out, feat[t], thr[t], dleft[t], [labels[v] for v in leaf[t]], names, depth)
        if counts is not None:   # splice the ScoreDistributions into the leaf elements just emitted (leaves appear left to right)
            k = 0
            for i in range(n0, len(out)):
                if out[i].startswith('<Node') and ' score="' in out[i] and out[i].endswith('</Node>'):
                    sd = "".join(f'<ScoreDistribution value="{labels[c]}" recordCount="{int(counts[t, k, c])}"/>' for c in range(n_labels))
                    out[i] = out[i][:-len('</Node>')] + sd + '</Node>'
                    k += 1
        out.append('</TreeModel></Segment>')
    out.append('</Segmentation></MiningModel></PMML>')
    arrays = {"feat": feat, "thr": thr, "dleft": dleft, "leaf": leaf, "depth": depth, "n_labels": n_labels, "counts": counts}
    return "\n".join(out).encode(), arrays


def rf_average_numpy(arrays, X, present=None):
    """Averaged leaf probabilities (recordCount / total per leaf), summed in segment order; first maximal label."""
    li = forest_leaf_index_numpy(arrays, X, present)
    n, T = li.shape
    c = arrays["counts"].astype(np.float64)
    prob = c / c.sum(axis=2, keepdims=True)
    acc = np.zeros((n, arrays["n_labels"]))
    for t in range(T):
        acc = acc + prob[t, li[:, t], :]
    return acc.argmax(axis=1).astype(np.float64), acc / T


def forest_leaf_index_numpy(arrays, X, present=None):
    """Leaf position (0..2^depth-1) of every (event, tree); plain numpy restatement of the walk."""
    feat, thr, dleft, depth = arrays["feat"], arrays["thr"], arrays["dleft"], arrays["depth"]
    T = feat.shape[0]
    n = X.shape[0]
    if present is None:
        present = np.ones_like(X, dtype=bool)
    if arrays.get("dtype") == "float":
        X = X.astype(np.float32).astype(np.float64)
    rows = np.arange(n)
    n_int = (1 << depth) - 1
    out = np.empty((n, T), dtype=np.int64)
    for t in range(T):
        idx = np.zeros(n, dtype=np.int64)
        for _ in range(depth):
            f = feat[t, idx]
            x = X[rows, f]
            p = present[rows, f]
            with np.errstate(invalid="ignore"):
                go_left = np.where(p, x <= thr[t, idx], dleft[t, idx])
            idx = 2 * idx + 1 + (~go_left)
        out[:, t] = idx - n_int
    return out


def gbt_numpy(arrays, X, present=None):
    li = forest_leaf_index_numpy(arrays, X, present)
    T = li.shape[1]
    s = np.zeros(X.shape[0])
    for t in range(T):  # sequential segment order
        s = s + arrays["leaf"][t, li[:, t]]
    if arrays["method"] == "average":
        s = s / T
    return s * arrays["rescale_factor"] + arrays["rescale_constant"]


def rf_numpy(arrays, X, present=None):
    """Majority vote; ties -> label first seen in segment order (JPMML insertion order, SURVEY App. A [R])."""
    li = forest_leaf_index_numpy(arrays, X, present)
    n, T = li.shape
    L = arrays["n_labels"]
    votes = np.zeros((n, L), dtype=np.int64)
    first = np.full((n, L), T + 1, dtype=np.int64)
    rows = np.arange(n)
    for t in range(T):
        lab = arrays["leaf"][t, li[:, t]]
        votes[rows, lab] += 1
        first[rows, lab] = np.minimum(first[rows, lab], t)
    best = votes.max(axis=1, keepdims=True)
    cand = np.where(votes == best, first, T + 2)
    return cand.argmin(axis=1).astype(np.float64)


# --------------------------------------------------------------------------------------------
# RegressionModel (C3)
# --------------------------------------------------------------------------------------------

def regression_pmml(F=1000, C=20, seed=SEED_BASE + 3, normalization="softmax", target_dtype="string"):
    rng = np.random.default_rng(seed)
    coef = rng.normal(0.0, 1.0 / np.sqrt(F), size=(C, F))
    icpt = rng.normal(0.0, 0.1, size=C)
    names = [f"x{i}" for i in range(F)]
    labels = [str(i) for i in range(C)]
    out = _header()
    out += _data_dictionary([(n, "continuous", "double") for n in names], ("y", "categorical", target_dtype, labels))
    out.append(f'<RegressionModel modelName="mlogit-{F}x{C}" functionName="classification" '
               f'normalizationMethod="{normalization}">')
    out += _mining_schema(names, "y")
    for c in range(C):
        out.append(f'<RegressionTable intercept="{_r(icpt[c])}" targetCategory="{labels[c]}">')
        out.extend(f'<NumericPredictor name="{names[i]}" coefficient="{_r(coef[c, i])}"/>' for i in range(F))
        out.append('</RegressionTable>')
    out.append('</RegressionModel></PMML>')
    return "\n".join(out).encode(), {"coef": coef, "intercept": icpt}


def regression_numpy(arrays, X):
    coef, icpt = arrays["coef"], arrays["intercept"]
    y = np.tile(icpt, (X.shape[0], 1))
    for i in range(coef.shape[1]):  # sequential predictor order
        y = y + X[:, i:i + 1] * coef[:, i][None, :]
    return y


def linear_regression_pmml(F=8, seed=7, normalization="none", exponents=False):
    """functionName=regression single table (optionally with exponent=2 terms)."""
    rng = np.random.default_rng(seed)
    coef = rng.normal(size=F)
    icpt = float(rng.normal())
    names = [f"x{i}" for i in range(F)]
    out = _header()
    out += _data_dictionary([(n, "continuous", "double") for n in names], ("y", "continuous", "double", None))
    out.append(f'<RegressionModel modelName="linreg" functionName="regression" normalizationMethod="{normalization}">')
    out += _mining_schema(names, "y")
    out.append(f'<RegressionTable intercept="{_r(icpt)}">')
    expo = [(2 if (exponents and i % 3 == 0) else 1) for i in range(F)]
    out.extend(f'<NumericPredictor name="{names[i]}" exponent="{expo[i]}" coefficient="{_r(coef[i])}"/>' for i in range(F))
    out.append('</RegressionTable></RegressionModel></PMML>')
    return "\n".join(out).encode(), {"coef": coef, "intercept": icpt, "exponent": np.array(expo)}


# --------------------------------------------------------------------------------------------
# NeuralNetwork (C5)
# --------------------------------------------------------------------------------------------

def nn_pmml(n_in=64, hidden=(256, 256), n_out=8, seed=SEED_BASE + 6, activation="rectifier",
            out_normalization="softmax", target_dtype="string"):
    """Classification MLP: hidden layers `activation`, output layer identity + softmax normalisation,
    NeuralOutputs via NormDiscrete on the target."""
    rng = np.random.default_rng(seed)
    sizes = [n_in, *hidden, n_out]
    names = [f"x{i}" for i in range(n_in)]
    labels = [str(i) for i in range(n_out)]
    out = _header()
    out += _data_dictionary([(n, "continuous", "double") for n in names], ("y", "categorical", target_dtype, labels))
    out.append(f'<NeuralNetwork modelName="mlp-{"-".join(map(str, sizes))}" functionName="classification" '
               f'activationFunction="{activation}">')
    out += _mining_schema(names, "y")
    out.append(f'<NeuralInputs numberOfInputs="{n_in}">')
    for i, n in enumerate(names):
        out.append(f'<NeuralInput id="0,{i}"><DerivedField optype="continuous" dataType="double">'
                   f'<FieldRef field="{n}"/></DerivedField></NeuralInput>')
    out.append('</NeuralInputs>')
    weights, biases = [], []
    for l in range(1, len(sizes)):
        fan_in, fan_out = sizes[l - 1], sizes[l]
        W = rng.normal(0.0, 1.0 / np.sqrt(fan_in), size=(fan_out, fan_in))
        b = rng.normal(0.0, 0.1, size=fan_out)
        weights.append(W)
        biases.append(b)
        last = l == len(sizes) - 1
        attr = f' activationFunction="identity" normalizationMethod="{out_normalization}"' if last else ""
        out.append(f'<NeuralLayer numberOfNeurons="{fan_out}"{attr}>')
        for j in range(fan_out):
            out.append(f'<Neuron id="{l},{j}" bias="{_r(b[j])}">')
            out.extend(f'<Con from="{l - 1},{i}" weight="{_r(W[j, i])}"/>' for i in range(fan_in))
            out.append('</Neuron>')
        out.append('</NeuralLayer>')
    out.append(f'<NeuralOutputs numberOfOutputs="{n_out}">')
    L = len(sizes) - 1
    for j in range(n_out):
        out.append(f'<NeuralOutput outputNeuron="{L},{j}"><DerivedField optype="categorical" dataType="string">'
                   f'<NormDiscrete field="y" value="{labels[j]}"/></DerivedField></NeuralOutput>')
    out.append('</NeuralOutputs></NeuralNetwork></PMML>')
    return "\n".join(out).encode(), {"weights": weights, "biases": biases, "activation": activation}


def nn_numpy(arrays, X):
    """Pre-softmax outputs of the last layer (sequential connection order)."""
    a = X
    nl = len(arrays["weights"])
    for l, (W, b) in enumerate(zip(arrays["weights"], arrays["biases"])):
        z = np.tile(b, (a.shape[0], 1))
        for i in range(W.shape[1]):
            z = z + a[:, i:i + 1] * W[:, i][None, :]
        if l < nl - 1:
            act = arrays["activation"]
            if act == "rectifier":
                a = np.maximum(z, 0.0)
            elif act == "tanh":
                a = np.tanh(z)
            elif act == "logistic":
                a = 1.0 / (1.0 + np.exp(-z))
            else:
                a = z
        else:
            a = z
    return a


# --------------------------------------------------------------------------------------------
# Synthetic events
# --------------------------------------------------------------------------------------------

def events(n, F, seed, p_missing=0.0):
    """x ~ N(0,1)^F doubles; present-bit matrix n x ceil(F/32) uint32 (bit=1 present)."""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, F))
    present = np.ones((n, F), dtype=bool) if p_missing <= 0 else rng.random((n, F)) >= p_missing
    return X, present, pack_present(present)


def pack_present(present: np.ndarray) -> np.ndarray:
    n, F = present.shape
    W = (F + 31) // 32
    padded = np.zeros((n, W * 32), dtype=np.uint8)
    padded[:, :F] = present
    bits = np.packbits(padded.reshape(n, W, 32), axis=2, bitorder="little")  # n x W x 4 bytes
    return np.ascontiguousarray(bits).view(np.uint32).reshape(n, W)


def iris_events(n, seed=SEED_BASE + 1):
    """4 doubles U[0.2,6.0] floored to one decimal (shape of examples' IrisSource.scala:32-43)."""
    rng = np.random.default_rng(seed)
    return np.floor(rng.uniform(0.2, 6.0, size=(n, 4)) * 10.0) / 10.0


# --------------------------------------------------------------------------------------------
# Boosted-tree CLASSIFIER as exporters write it: modelChain(sum-of-trees per class -> RegressionModel link)
# --------------------------------------------------------------------------------------------

def gbt_classifier_pmml(T=50, depth=6, F=30, n_classes=2, seed=77, learning_rate=0.1, dtype="double"):
    """MiningModel classification, multipleModelMethod="modelChain":
         K inner MiningModels (regression, sum of T complete trees, Targets rescaleFactor=lr rescaleConstant=init_k,
         OutputField decisionFunction(k) feature=predictedValue), K = 1 for two classes, else K = n_classes;
         last segment RegressionModel classification: logit (binary) or softmax over the K decision functions."""
    K = 1 if n_classes == 2 else n_classes
    names = [f"x{i}" for i in range(F)]
    labels = [str(i) for i in range(n_classes)]
    out = _header()
    out += _data_dictionary([(n, "continuous", dtype) for n in names], ("y", "categorical", "string", labels))
    out.append('<MiningModel modelName="gbt-classifier" functionName="classification">')
    out += _mining_schema(names, "y")
    out.append('<Segmentation multipleModelMethod="modelChain">')
    chans = []
    for k in range(K):
        feat, thr, dleft, rng = forest_arrays(T, depth, F, seed + 13 * k)
        leaf = rng.normal(0.0, 1.0, size=(T, 1 << depth))
        init = float(rng.normal(0.0, 0.3))
        chans.append({"feat": feat, "thr": thr, "dleft": dleft, "leaf": leaf, "depth": depth, "method": "sum",
                      "rescale_constant": init, "rescale_factor": learning_rate, "dtype": dtype})
        out.append(f'<Segment id="{k + 1}"><True/><MiningModel functionName="regression">')
        out += _mining_schema(names)
        out.append(f'<Output><OutputField name="decisionFunction({k})" optype="continuous" dataType="double" feature="predictedValue" isFinalResult="false"/></Output>')
        out.append(f'<Targets><Target rescaleFactor="{_r(learning_rate)}" rescaleConstant="{_r(init)}"/></Targets>')
        out.append('<Segmentation multipleModelMethod="sum">')
        for t in range(T):
            out.append(f'<Segment id="{t + 1}"><True/><TreeModel functionName="regression" missingValueStrategy="defaultChild" '
                       'noTrueChildStrategy="returnNullPrediction" splitCharacteristic="binarySplit">')
            out += _mining_schema(names)
            _complete_tree_xml(out, feat[t], thr[t], dleft[t], [_r(v) for v in leaf[t]], names, depth)
            out.append('</TreeModel></Segment>')
        out.append('</Segmentation></MiningModel></Segment>')
    norm = "logit" if n_classes == 2 else "softmax"
    out.append(f'<Segment id="{K + 1}"><True/><RegressionModel functionName="classification" normalizationMethod="{norm}">')
    out.append('<MiningSchema><MiningField name="y" usageType="target"/>' + "".join(f'<MiningField name="decisionFunction({k})"/>' for k in range(K)) + '</MiningSchema>')
    if n_classes == 2:
        out.append('<RegressionTable intercept="0.0" targetCategory="1"><NumericPredictor name="decisionFunction(0)" coefficient="1.0"/></RegressionTable>')
        out.append('<RegressionTable intercept="0.0" targetCategory="0"/>')
    else:
        for k in range(K):
            out.append(f'<RegressionTable intercept="0.0" targetCategory="{k}"><NumericPredictor name="decisionFunction({k})" coefficient="1.0"/></RegressionTable>')
    out.append('</RegressionModel></Segment></Segmentation></MiningModel></PMML>')
    return "\n".join(out).encode(), {"channels": chans, "n_classes": n_classes}


def gbt_classifier_numpy(arrays, X, present=None):
    """(label, probabilities in RegressionTable order) -- plain numpy restatement of the chain."""
    dec = np.stack([gbt_numpy(ch, X, present) for ch in arrays["channels"]], axis=1)
    if arrays["n_classes"] == 2:
        p1 = 1.0 / (1.0 + np.exp(-(0.0 + 1.0 * dec[:, 0])))
        probs = np.stack([p1, 1.0 - p1], axis=1)            # table order: "1", then "0"
        label = np.where(probs[:, 0] >= probs[:, 1], 1.0, 0.0)  # first maximal entry wins ties
        return label, probs
    e = np.exp(dec)
    probs = e / e.sum(axis=1, keepdims=True)
    return probs.argmax(axis=1).astype(np.float64), probs


def gbt_pruned_pmml(T=60, max_depth=6, F=20, seed=5, p_leaf=0.25, operator="lessThan", dtype="float", method="sum",
                    float_thresholds=True):
    """XGBoost-export shape: float fields, `x < t` on the left child, <True/> on the right, defaultChild, trees pruned at
    random (a node below the root becomes a leaf with probability p_leaf), so depths differ within and across trees."""
    from scipy.stats import norm
    rng = np.random.default_rng(seed)
    grid = norm.ppf((np.arange(255) + 1.0) / 256.0)
    if float_thresholds:
        grid = grid.astype(np.float32).astype(np.float64)
    names = [f"x{i}" for i in range(F)]
    out = _header()
    out += _data_dictionary([(n, "continuous", dtype) for n in names], ("y", "continuous", "double", None))
    out.append(f'<MiningModel modelName="gbt-pruned" functionName="regression">')
    out += _mining_schema(names, "y")
    out.append(f'<Segmentation multipleModelMethod="{method}">')
    for t in range(T):
        out.append(f'<Segment id="{t + 1}"><True/><TreeModel functionName="regression" missingValueStrategy="defaultChild" '
                   'noTrueChildStrategy="returnNullPrediction" splitCharacteristic="binarySplit">')
        out += _mining_schema(names)
        counter = [0]

        def emit(depth, pred):
            counter[0] += 1
            my_id = counter[0]
            if depth == max_depth or (depth >= 1 and rng.random() < p_leaf):
                out.append(f'<Node id="{my_id}" score="{_r(rng.normal(0, 0.1))}">{pred}</Node>')
                return
            at = len(out)
            out.append(None)   # patched below once the children's ids are known
            f, thr = names[int(rng.integers(F))], float(grid[int(rng.integers(255))])
            left_id = counter[0] + 1
            emit(depth + 1, f'<SimplePredicate field="{f}" operator="{operator}" value="{_r(thr)}"/>')
            right_id = counter[0] + 1
            emit(depth + 1, '<True/>')
            out[at] = f'<Node id="{my_id}" defaultChild="{left_id if rng.random() < 0.5 else right_id}">{pred}'
            out.append('</Node>')

        emit(0, '<True/>')
        out.append('</TreeModel></Segment>')
    out.append('</Segmentation></MiningModel></PMML>')
    return "\n".join(out).encode()
