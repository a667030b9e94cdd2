import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_bytes(name: str) -> bytes:
    with open(os.path.join(GOLDEN, name), "rb") as f:
        return f.read()


@pytest.fixture(scope="session")
def golden_vectors():
    with open(os.path.join(GOLDEN, "golden_vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def gen():
    from flink_jpmml_b200 import pmmlgen
    return pmmlgen


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def engine():
    """One engine on cuda:0 for the GPU tests; fails loudly when the CUDA library or a device is missing."""
    from flink_jpmml_b200 import capi
    e = capi.Engine(0)
    yield e
    e.close()


def vector_to_batch(gen, v):
    """golden_vectors.json entry -> (X, present_bits, replace) with the reference's dense / sparse conventions."""
    if v["kind"] == "dense":
        return np.array([v["values"]], dtype=np.float64), None, v["replace"]
    X = np.zeros((1, v["size"]))
    pres = np.zeros((1, v["size"]), dtype=bool)
    for i, val in zip(v["indices"], v["values"]):
        X[0, i] = val
        pres[0, i] = True
    return X, gen.pack_present(pres), v["replace"]
