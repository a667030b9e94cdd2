"""ctypes binding of the CPU oracle (oracle/oracle.cc).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this;
the product package never does (the product has no CPU path).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

STATUS = {0: "score", 1: "size", 2: "prepare", 3: "evaluate", 4: "extract", 5: "nomodel"}


def build(force=False):
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "oracle.cc")):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_load.restype = C.c_void_p
        L.orc_load.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_num_active.argtypes = [C.c_void_p]
        L.orc_active_name.argtypes = [C.c_void_p, C.c_int]
        L.orc_active_name.restype = C.c_char_p
        L.orc_model_name.argtypes = [C.c_void_p]
        L.orc_model_name.restype = C.c_char_p
        L.orc_num_segments.argtypes = [C.c_void_p]
        L.orc_predict.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
        _lib = L
    return _lib


class OracleLoadError(Exception):
    """Counterpart of ModelLoadingException (api/exceptions/package.scala:42-43)."""


class OracleModel:
    def __init__(self, pmml: bytes):
        if isinstance(pmml, str):
            pmml = pmml.encode()
        err = C.create_string_buffer(512)
        self._h = lib().orc_load(pmml, len(pmml), err, 512)
        if not self._h:
            raise OracleLoadError(err.value.decode(errors="replace"))
        self.n_active = lib().orc_num_active(self._h)
        self.active_fields = [lib().orc_active_name(self._h, i).decode() for i in range(self.n_active)]
        self.model_name = lib().orc_model_name(self._h).decode()
        self.n_segments = lib().orc_num_segments(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_free(self._h)
            self._h = None

    def predict(self, X, present_bits=None, replace=None, node_ids=False, n_probs=0, nthreads=1):
        """X: n x vec_size doubles.  Returns dict(score, status, aux[, node_ids][, probs])."""
        X = np.ascontiguousarray(X, dtype=np.float64)
        if X.ndim == 1:
            X = X[None, :]
        n, F = X.shape
        score = np.empty(n, dtype=np.float64)
        status = np.empty(n, dtype=np.uint8)
        aux = np.empty(n, dtype=np.int32)
        pb = None
        if present_bits is not None:
            pb = np.ascontiguousarray(present_bits, dtype=np.uint32)
            assert pb.shape == (n, (F + 31) // 32), pb.shape
        rp = None
        if replace is not None:
            rp = np.ascontiguousarray(np.broadcast_to(np.asarray(replace, dtype=np.float64), (n,)))
        ns = max(1, self.n_segments)
        ids = np.empty((n, ns), dtype=np.int32) if node_ids else None
        pr = np.empty((n, n_probs), dtype=np.float64) if n_probs else None
        p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        lib().orc_predict(self._h, p(X), p(pb), p(rp), n, F, p(score), p(status), p(aux), p(ids), ns, p(pr), n_probs, nthreads)
        out = {"score": score, "status": status, "aux": aux}
        if node_ids:
            out["node_ids"] = ids
        if n_probs:
            out["probs"] = pr
        return out
