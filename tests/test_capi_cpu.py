"""CPU: the C-ABI library loads, exports every symbol include/fjpmml.h declares, and the PMML compiler (host code)
accepts / rejects documents like the reference's model load does.  No compute calls are made here (no GPU)."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden_bytes


@pytest.fixture(scope="module")
def capi():
    from flink_jpmml_b200 import capi as c
    c.build()
    return c


def test_library_exports_every_declared_symbol(capi):
    decl = set()
    for h in os.listdir(os.path.join(ROOT, "include")):
        with open(os.path.join(ROOT, "include", h)) as f:
            text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
        decl |= set(re.findall(r"\b(fj[ph]_[a-z0-9_]+)\s*\(", text))
    assert decl, "no declarations found"
    lib = ctypes.CDLL(capi.LIB_PATH)
    missing = [s for s in sorted(decl) if not hasattr(lib, s)]
    assert not missing, missing
    assert set(capi.EXPORTS) <= decl


def test_version_and_no_cpu_scoring_path(capi):
    assert b"sm_100a" in capi.lib().fjp_version()
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(capi.EngineError, match="no CPU scoring path"):
            capi.Engine(0)


def test_compile_without_engine_reports_model_facts(capi, golden_vectors):
    m = capi.compile_model(golden_bytes("iris_kmeans43.pmml"))
    assert m.active_fields == golden_vectors["active_fields"]["value"]      # VectorConverterSpec.scala:76-78
    assert m.model_name == "k-means" and m.family == "ClusteringModel"
    assert m.labels == ["1", "2", "3", "4"]                                  # 1-based cluster positions
    assert m.device_bytes == 0
    for name in golden_vectors["loadable"]["value"] + ["iris_kmeans32.pmml", "iris_kmeans_nooutput.pmml",
                                                       "iris_kmeans_nooutput_notarget.pmml", "iris_kmeans_stringfields.pmml"]:
        assert capi.compile_model(golden_bytes(name)).n_fields == 4


def test_empty_or_broken_documents_fail_to_load(capi):
    # RichDataStreamSpec.scala:98-105: an empty PMML makes the job fail with ModelLoadingException
    for bad in (golden_bytes("iris_kmeans_empty.pmml"), b"", b"<PMML", b"<NotPMML/>", b"<PMML><DataDictionary/></PMML>"):
        with pytest.raises(capi.ModelLoadingException):
            capi.compile_model(bad)


def test_compiled_layouts(capi, gen):
    pm, _ = gen.gbt_pmml(T=12, depth=4, F=7, seed=1)
    d = capi.compile_model(pm).describe
    assert "format=complete" in d and "depth=4" in d and "enc=float" in d and "trees=12" in d
    from hand_pmml import doc, tree_model, AB, SMALL_TREE
    d = capi.compile_model(doc(AB, tree_model(["a", "b"], SMALL_TREE))).describe
    assert "format=general" in d and "enc=float" in d           # per feature only one kind of test: float keys are exact
    mixed = SMALL_TREE.replace('operator="lessThan" value="0"', 'operator="lessThan" value="0.1"').replace('field="b"', 'field="a"')
    d = capi.compile_model(doc(AB, tree_model(["a", "b"], mixed))).describe
    assert "enc=rank" in d                                       # < and <= on one double feature, and 0.1 is no float
    pm, _ = gen.regression_pmml(F=16, C=3, seed=2)
    assert "dense=1" in capi.compile_model(pm).describe
    pm, _ = gen.nn_pmml(n_in=5, hidden=(6,), n_out=2, seed=2)
    assert "layers=2" in capi.compile_model(pm).describe


def test_deep_forests_compile_on_the_host(capi, gen):
    """Complete-tree blobs of every depth / width the GPU suite and the bench use are built on the host without a device."""
    for T, depth, F in ((20, 8, 100), (6, 6, 128), (6, 6, 130), (6, 5, 10), (3, 10, 40)):
        pm, _ = gen.gbt_pmml(T=T, depth=depth, F=F, seed=depth)
        d = capi.compile_model(pm).describe
        assert "format=complete" in d and f"depth={depth}" in d and f"trees={T}" in d, d
        assert f"nodes={T * ((1 << depth) - 1)}" in d, d
    pm, _ = gen.rf_pmml(T=9, depth=7, F=30, n_labels=3, seed=5)
    assert "format=complete" in capi.compile_model(pm).describe


def test_throughput_path_forms_compile_on_the_host(capi, gen, monkeypatch):
    """Every variant of the compact forest forms is built by host code only: complete / explicit-children trees, float and
    64-bit threshold tables, inline leaf distributions, modelChain channels."""
    d = capi.compile_model(gen.gbt_pmml(T=9, depth=8, F=100, seed=2)[0]).describe
    assert "format=complete" in d and "compact=1" in d and "sparse" not in d
    d = capi.compile_model(gen.gbt_pruned_pmml(T=12, max_depth=13, F=20, seed=3, p_leaf=0.3)).describe
    assert "format=general" in d and "compact=1 sparse=1" in d, d
    d = capi.compile_model(gen.gbt_pruned_pmml(T=12, max_depth=13, F=20, seed=3, p_leaf=0.3, dtype="double", float_thresholds=False)).describe
    assert "compact=2 sparse=1" in d, d
    for method in ("average", "weightedMajorityVote"):
        d = capi.compile_model(gen.rf_pmml(T=5, depth=6, F=9, n_labels=5, seed=4, method=method, segment_weights=method != "average")[0]).describe
        assert "compact=1" in d, d
    d = capi.compile_model(gen.rf_pmml(T=3, depth=11, F=9, n_labels=3, seed=4, method="average")[0]).describe
    assert "sparse=1" in d, d
    d = capi.compile_model(gen.gbt_classifier_pmml(T=4, depth=11, F=9, n_classes=3, seed=5)[0]).describe
    assert "method=9" in d and "sparse=1" in d, d
    monkeypatch.setattr(gen, "FLOAT_THRESHOLDS", False)
    d = capi.compile_model(gen.gbt_pmml(T=9, depth=6, F=12, seed=2)[0]).describe
    assert "enc=rank" in d and "compact=2" in d, d


def test_unsupported_features_are_load_errors_not_wrong_answers(capi, gen):
    pm, _ = gen.gbt_pmml(T=3, depth=2, F=4, seed=1)
    for bad in (pm.replace(b'multipleModelMethod="sum"', b'multipleModelMethod="modelChain"'),
                pm.replace(b'multipleModelMethod="sum"', b'multipleModelMethod="selectAll"'),
                pm.replace(b'missingValueStrategy="defaultChild"', b'missingValueStrategy="weightedConfidence"')):
        with pytest.raises(capi.ModelLoadingException, match="unsupported"):
            capi.compile_model(bad)
