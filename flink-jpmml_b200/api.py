"""Python view of the host-side mirror (csrc/host/flinkjpmml.hpp) -- same names as the reference's Scala API, so the
tests read like the reference's specs.  Everything here is a thin call into libfjpmml.so (fjpmml_host.h)."""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import capi

FJH_OK, FJH_ENGINE_ERROR, FJH_WRONG_MODEL_ID_FORMAT, FJH_MODEL_LOADING = 0, -1, -2, -3


class WrongModelIdFormat(Exception):
    """api/exceptions/package.scala:53 -- job-fatal in the reference."""


ModelLoadingException = capi.ModelLoadingException


# ---- prediction ADT (models/prediction/Target.scala:44-84, Prediction.scala:72) ----------------------------------------
@dataclass(frozen=True)
class Score:
    value: float

    def get(self):
        return self.value

    def getOrElse(self, default):
        return self.value


class _EmptyScore:
    def get(self):
        raise LookupError("EmptyScore.nan")

    def getOrElse(self, default):
        return default

    def __repr__(self):
        return "EmptyScore"


EmptyScore = _EmptyScore()


@dataclass(frozen=True)
class Prediction:
    value: object

    @staticmethod
    def of(score: float, status: int) -> "Prediction":
        return Prediction(Score(float(score))) if status == 0 else Prediction(EmptyScore)


# ---- vectors (org.apache.flink.ml.math) ------------------------------------------------------------------------
@dataclass
class DenseVector:
    data: Sequence[float]

    @property
    def size(self):
        return len(self.data)


@dataclass
class SparseVector:
    size: int
    indices: Sequence[int]
    data: Sequence[float]


def _vec_args(v):
    if isinstance(v, DenseVector):
        vals = np.ascontiguousarray(v.data, dtype=np.float64)
        return vals, v.size, None, 0
    vals = np.ascontiguousarray(v.data, dtype=np.float64)
    idx = np.ascontiguousarray(v.indices, dtype=np.int32)
    return vals, v.size, idx, len(idx)


def _lib():
    L = capi.lib()
    if getattr(L, "_fjh_ready", False):
        return L
    vp, cp, sz, i32, i64, lng, dbl = C.c_void_p, C.c_char_p, C.c_size_t, C.c_int, C.c_int64, C.c_long, C.c_double
    sig = {
        "fjh_model_id_parse": (i32, [cp, cp, sz, C.POINTER(i64)]),
        "fjh_model_id_hash": (C.c_int32, [cp, i64]),
        "fjh_predict": (i32, [vp, vp, vp, i32, vp, i32, dbl, C.POINTER(dbl), C.POINTER(C.c_uint8)]),
        "fjh_eval_function_open": (vp, [vp, cp, i32, cp, sz]),
        "fjh_eval_function_model_name": (cp, [vp]),
        "fjh_eval_function_process": (i32, [vp, vp, i32, vp, i32, dbl]),
        "fjh_eval_function_process_dense": (i32, [vp, vp, lng, i32]),
        "fjh_eval_function_reserve": (vp, [vp, i32, i32, C.POINTER(i32)]),
        "fjh_eval_function_commit": (i32, [vp, i32]),
        "fjh_eval_function_flush": (i32, [vp]),
        "fjh_eval_function_poll": (lng, [vp, vp, vp, lng]),
        "fjh_eval_function_close": (None, [vp]),
        "fjh_cofunction_create": (vp, [vp, i32]),
        "fjh_cofunction_process_event": (i32, [vp, cp, vp, i32, vp, i32, dbl, cp, sz]),
        "fjh_cofunction_process_dense": (i32, [vp, C.POINTER(cp), vp, vp, lng, i32, cp, sz]),
        "fjh_cofunction_process_control": (i32, [vp, i32, cp, i64, cp, i64]),
        "fjh_cofunction_flush": (i32, [vp, cp, sz]),
        "fjh_cofunction_poll": (lng, [vp, vp, vp, lng]),
        "fjh_cofunction_snapshot": (lng, [vp, cp, lng]),
        "fjh_cofunction_restore": (i32, [vp, C.POINTER(cp), i32]),
        "fjh_cofunction_metadata_size": (i32, [vp]),
        "fjh_cofunction_cached_models": (i32, [vp]),
        "fjh_cofunction_parked_events": (lng, [vp]),
        "fjh_cofunction_flush_seconds": (dbl, [vp]),
        "fjh_cofunction_destroy": (None, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    L._fjh_ready = True
    return L


HOST_EXPORTS = ["fjh_model_id_parse", "fjh_model_id_hash", "fjh_predict", "fjh_eval_function_open", "fjh_eval_function_model_name",
                "fjh_eval_function_process", "fjh_eval_function_process_dense", "fjh_eval_function_reserve", "fjh_eval_function_commit", "fjh_eval_function_flush", "fjh_eval_function_poll", "fjh_eval_function_close",
                "fjh_cofunction_create", "fjh_cofunction_process_event", "fjh_cofunction_process_dense", "fjh_cofunction_process_control", "fjh_cofunction_flush",
                "fjh_cofunction_poll", "fjh_cofunction_snapshot", "fjh_cofunction_restore", "fjh_cofunction_metadata_size",
                "fjh_cofunction_cached_models", "fjh_cofunction_parked_events", "fjh_cofunction_flush_seconds", "fjh_cofunction_destroy"]


def _raise(rc, err=b""):
    msg = err.decode(errors="replace") if err else ""
    if rc == FJH_WRONG_MODEL_ID_FORMAT:
        raise WrongModelIdFormat(msg)
    if rc == FJH_MODEL_LOADING:
        raise ModelLoadingException(msg)
    if rc != FJH_OK:
        raise capi.EngineError(msg or capi.lib().fjp_last_error().decode(errors="replace"))


# ---- ModelId / control messages (models/core/ModelId.scala, models/control/ServingMessage.scala) -----------------------
@dataclass(frozen=True)
class ModelId:
    name: str
    version: int
    separatorSymbol = "_"

    @staticmethod
    def fromIdentifier(id: str) -> "ModelId":
        name = C.create_string_buffer(64)
        ver = C.c_int64()
        rc = _lib().fjh_model_id_parse(id.encode(), name, 64, C.byref(ver))
        if rc != FJH_OK:
            raise WrongModelIdFormat(id)
        return ModelId(name.value.decode(), ver.value)

    def hashCode(self) -> int:
        return _lib().fjh_model_id_hash(self.name.encode(), self.version)


@dataclass(frozen=True)
class AddMessage:
    name: str
    version: int
    path: str
    occurredOn: int = 0


@dataclass(frozen=True)
class DelMessage:
    name: str
    version: int
    occurredOn: int = 0


def java_string_hash(s: str) -> int:
    h = 0
    for ch in s:
        h = (31 * h + ord(ch)) & 0xFFFFFFFF
    return h - (1 << 32) if h >= (1 << 31) else h


# ---- PmmlModel (api/PmmlModel.scala) ----------------------------------------------------------------------------------
class PmmlModel:
    def __init__(self, engine: capi.Engine, model: Optional[capi.Model]):
        self._engine, self._model = engine, model

    @staticmethod
    def fromReader(engine: capi.Engine, path: str) -> "PmmlModel":
        try:
            with open(path, "rb") as f:
                data = f.read()
        except OSError as e:
            raise ModelLoadingException(str(e))
        return PmmlModel(engine, engine.compile(data))

    @staticmethod
    def empty(engine: capi.Engine) -> "PmmlModel":
        return PmmlModel(engine, None)

    def predict(self, vector, replaceNan: Optional[float] = None) -> Prediction:
        vals, size, idx, nnz = _vec_args(vector)
        score, status = C.c_double(), C.c_uint8()
        rc = _lib().fjh_predict(self._engine._h, self._model._h if self._model else None, capi._ptr(vals), size, capi._ptr(idx), nnz,
                                math.nan if replaceNan is None else float(replaceNan), C.byref(score), C.byref(status))
        _raise(rc)
        return Prediction.of(score.value, status.value)


def _poll(fn, handle):
    out = []
    sc = np.empty(4096, dtype=np.float64)
    st = np.empty(4096, dtype=np.uint8)
    while True:
        n = fn(handle, capi._ptr(sc), capi._ptr(st), 4096)
        out += [Prediction.of(sc[i], st[i]) for i in range(n)]
        if n < 4096:
            return out


def _poll_arrays(fn, handle):
    """Same drain as _poll, as (scores, status) arrays: what a throughput test or benchmark wants."""
    scs, sts = [], []
    sc = np.empty(1 << 16, dtype=np.float64)
    st = np.empty(1 << 16, dtype=np.uint8)
    while True:
        n = fn(handle, capi._ptr(sc), capi._ptr(st), 1 << 16)
        scs.append(sc[:n].copy()); sts.append(st[:n].copy())
        if n < (1 << 16):
            return np.concatenate(scs), np.concatenate(sts)


# ---- operators ---------------------------------------------------------------------------------------------------------
class EvaluationFunction:
    """stream.evaluate(reader) / quickEvaluate(reader) with micro-batching (package.scala:76-82, 138-142)."""

    def __init__(self, engine: capi.Engine, path: str, max_batch: int = 1024):
        err = C.create_string_buffer(512)
        self._h = _lib().fjh_eval_function_open(engine._h, path.encode(), max_batch, err, 512)
        if not self._h:
            raise ModelLoadingException(err.value.decode(errors="replace"))

    @property
    def modelName(self):
        return _lib().fjh_eval_function_model_name(self._h).decode()

    def processElement(self, vector, replaceNan=None):
        vals, size, idx, nnz = _vec_args(vector)
        _raise(_lib().fjh_eval_function_process(self._h, capi._ptr(vals), size, capi._ptr(idx), nnz, math.nan if replaceNan is None else float(replaceNan)))

    def processDense(self, X):
        X = np.ascontiguousarray(X, dtype=np.float64)
        _raise(_lib().fjh_eval_function_process_dense(self._h, capi._ptr(X), X.shape[0], X.shape[1]))

    def reserve(self, size: int, want_rows: int):
        """Rows of the pinned micro-batch the producer may fill itself: a writable (granted, size) array, or None when `size`
        is not the model's field count.  commit(rows) declares how many were written."""
        granted = C.c_int(0)
        p = _lib().fjh_eval_function_reserve(self._h, size, want_rows, C.byref(granted))
        if not p or granted.value <= 0:
            return None
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(granted.value, size))

    def commit(self, rows: int):
        _raise(_lib().fjh_eval_function_commit(self._h, rows))

    def flush(self):
        _raise(_lib().fjh_eval_function_flush(self._h))

    def collect(self):
        return _poll(_lib().fjh_eval_function_poll, self._h)

    def collect_arrays(self):
        return _poll_arrays(_lib().fjh_eval_function_poll, self._h)

    def close(self):
        if self._h:
            _lib().fjh_eval_function_close(self._h)
            self._h = None


class EvaluationCoFunction:
    """events.withSupportStream(ctrl).evaluate(...) with micro-batching (package.scala:63-65, 107-119, EvaluationCoFunction.scala:55-133)."""

    def __init__(self, engine: Optional[capi.Engine], max_batch: int = 1024):
        self._h = _lib().fjh_cofunction_create(engine._h if engine else None, max_batch)

    def processElement1(self, modelId: str, vector, replaceNan=None):
        vals, size, idx, nnz = _vec_args(vector)
        err = C.create_string_buffer(512)
        rc = _lib().fjh_cofunction_process_event(self._h, modelId.encode(), capi._ptr(vals), size, capi._ptr(idx), nnz,
                                                  math.nan if replaceNan is None else float(replaceNan), err, 512)
        _raise(rc, err.value)

    def processDense1(self, ids: Sequence[str], id_index, X):
        X = np.ascontiguousarray(X, dtype=np.float64)
        idx = np.ascontiguousarray(id_index, dtype=np.int32)
        arr = (C.c_char_p * len(ids))(*[i.encode() for i in ids])
        err = C.create_string_buffer(512)
        _raise(_lib().fjh_cofunction_process_dense(self._h, arr, capi._ptr(idx), capi._ptr(X), X.shape[0], X.shape[1], err, 512), err.value)

    def processElement2(self, msg):
        if isinstance(msg, AddMessage):
            rc = _lib().fjh_cofunction_process_control(self._h, 1, msg.name.encode(), msg.version, msg.path.encode(), msg.occurredOn)
        else:
            rc = _lib().fjh_cofunction_process_control(self._h, 0, msg.name.encode(), msg.version, None, msg.occurredOn)
        _raise(rc)

    def flush(self):
        err = C.create_string_buffer(512)
        _raise(_lib().fjh_cofunction_flush(self._h, err, 512), err.value)

    def collect(self):
        return _poll(_lib().fjh_cofunction_poll, self._h)

    def collect_arrays(self):
        return _poll_arrays(_lib().fjh_cofunction_poll, self._h)

    def snapshotState(self) -> bytes:
        buf = C.create_string_buffer(1 << 16)
        n = _lib().fjh_cofunction_snapshot(self._h, buf, 1 << 16)
        if n < 0:
            raise capi.EngineError("snapshot failed")
        return buf.raw[:n]

    def initializeState(self, snapshots: Sequence[bytes]):
        arr = (C.c_char_p * len(snapshots))(*snapshots)
        _raise(_lib().fjh_cofunction_restore(self._h, arr, len(snapshots)))

    @property
    def servingMetadata(self) -> int:
        return _lib().fjh_cofunction_metadata_size(self._h)

    @property
    def servingModels(self) -> int:
        return _lib().fjh_cofunction_cached_models(self._h)

    @property
    def parkedEvents(self) -> int:
        """Events the bulk path parked behind a model load that was still running (they keep their output position)."""
        return _lib().fjh_cofunction_parked_events(self._h)

    @property
    def flushSeconds(self) -> float:
        """Seconds the task thread spent handing micro-batches to the GPU and waiting for them (the rest is packing)."""
        return _lib().fjh_cofunction_flush_seconds(self._h)

    def close(self):
        if self._h:
            _lib().fjh_cofunction_destroy(self._h)
            self._h = None
