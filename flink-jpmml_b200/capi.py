"""ctypes binding of include/fjpmml.h.  Thin on purpose: every call goes straight through the C ABI.

There is no CPU scoring path here or below: without the CUDA library (or without a GPU) scoring raises."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfjpmml.so")
_lib = None

ST_SCORE, ST_SIZE, ST_PREPARE, ST_EVALUATE, ST_EXTRACT, ST_NOMODEL = range(6)
FJP_BUSY = 1   # fjp_submit: no free stream slot, wait for an own ticket and retry


class EngineError(RuntimeError):
    """Engine-fatal condition (CUDA error, no device, bad argument)."""


class ModelLoadingException(Exception):
    """Same role as api/exceptions/package.scala:42-43 in the reference."""


def build(force: bool = False) -> str:
    """Compiles csrc/ for sm_100a (nvcc cross-compiles without a GPU)."""
    args = ["make", "-C", os.path.join(_HERE, "csrc"), "-s"] + (["-B"] if force else [])
    subprocess.check_call(args)
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EngineError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() (make -C flink-jpmml_b200/csrc)")
    L = C.CDLL(LIB_PATH)
    vp, cp, sz, i32, i64, lng = C.c_void_p, C.c_char_p, C.c_size_t, C.c_int, C.c_int64, C.c_long
    sig = {
        "fjp_version": (cp, []),
        "fjp_last_error": (cp, []),
        "fjp_max_inflight": (i32, []),
        "fjp_engine_create": (i32, [i32, C.POINTER(vp), cp, sz]),
        "fjp_engine_destroy": (None, [vp]),
        "fjp_engine_device": (i32, [vp]),
        "fjp_engine_sync": (i32, [vp]),
        "fjp_engine_launch_count": (lng, [vp]),
        "fjp_engine_status_counts": (None, [vp, C.POINTER(C.c_uint64)]),
        "fjp_engine_fast_path_counts": (i32, [vp, C.POINTER(C.c_uint64)]),
        "fjp_model_compile": (i32, [vp, cp, sz, C.POINTER(vp), cp, sz]),
        "fjp_model_retain": (None, [vp]),
        "fjp_model_release": (None, [vp]),
        "fjp_model_replicate": (i32, [vp, vp, C.POINTER(vp), cp, sz]),
        "fjp_model_num_fields": (i32, [vp]),
        "fjp_model_field_name": (cp, [vp, i32]),
        "fjp_model_name": (cp, [vp]),
        "fjp_model_family": (cp, [vp]),
        "fjp_model_num_labels": (i32, [vp]),
        "fjp_model_label": (cp, [vp, i32]),
        "fjp_model_num_segments": (i32, [vp]),
        "fjp_model_describe": (cp, [vp]),
        "fjp_model_device_bytes": (sz, [vp]),
        "fjp_table_add": (i32, [vp, cp, i64, vp]),
        "fjp_table_del": (i32, [vp, cp, i64]),
        "fjp_table_get": (vp, [vp, cp, i64]),
        "fjp_table_size": (i32, [vp]),
        "fjp_table_epoch": (C.c_uint64, [vp]),
        "fjp_host_alloc": (vp, [sz]),
        "fjp_host_free": (None, [vp]),
        "fjp_score_batch": (i32, [vp, vp, vp, vp, vp, lng, i32, vp, vp, vp]),
        "fjp_score_batch_ex": (i32, [vp, vp, vp, vp, vp, lng, i32, vp, vp, vp, vp, i32, vp, i32]),
        "fjp_score_sparse": (i32, [vp, vp, vp, vp, vp, vp, lng, i32, vp, vp, vp]),
        "fjp_model_float_row_modes": (i32, [vp, vp, i32]),
        "fjp_score_batch_f32": (i32, [vp, vp, vp, vp, lng, i32, vp, vp, vp]),
        "fjp_score_device": (i32, [vp, vp, vp, vp, vp, lng, i32, vp, vp, vp, vp]),
        "fjp_score_sparse_noalloc": (i32, [vp, vp, vp, vp, vp, lng, i32, vp, vp, vp]),
        "fjp_submit": (i32, [vp, vp, vp, vp, vp, lng, i32, vp, vp, vp, C.POINTER(i32)]),
        "fjp_submit_many": (i32, [vp, vp, i32, C.POINTER(i32)]),
        "fjp_wait": (i32, [vp, i32]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


EXPORTS = ["fjp_version", "fjp_last_error", "fjp_max_inflight", "fjp_engine_create", "fjp_engine_destroy", "fjp_engine_device",
           "fjp_engine_sync", "fjp_engine_launch_count", "fjp_engine_status_counts", "fjp_engine_fast_path_counts", "fjp_model_compile", "fjp_model_retain", "fjp_model_release",
           "fjp_model_replicate", "fjp_model_num_fields", "fjp_model_field_name", "fjp_model_name", "fjp_model_family",
           "fjp_model_num_labels", "fjp_model_label", "fjp_model_num_segments", "fjp_model_describe", "fjp_model_device_bytes",
           "fjp_table_add", "fjp_table_del", "fjp_table_get", "fjp_table_size", "fjp_table_epoch", "fjp_host_alloc",
           "fjp_host_free", "fjp_score_batch", "fjp_score_batch_ex", "fjp_score_sparse", "fjp_model_float_row_modes", "fjp_score_batch_f32", "fjp_score_sparse_noalloc", "fjp_score_device", "fjp_submit", "fjp_submit_many", "fjp_wait"]


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _check(rc):
    if rc < 0:
        raise EngineError(lib().fjp_last_error().decode(errors="replace"))
    return rc


class Model:
    """Opaque compiled model (fjp_model*)."""

    def __init__(self, handle, owned=True):
        self._h = handle
        self._owned = owned
        L = lib()
        self.n_fields = L.fjp_model_num_fields(handle)
        self.active_fields = [L.fjp_model_field_name(handle, i).decode() for i in range(self.n_fields)]
        self.model_name = L.fjp_model_name(handle).decode()
        self.family = L.fjp_model_family(handle).decode()
        self.labels = [L.fjp_model_label(handle, i).decode() for i in range(L.fjp_model_num_labels(handle))]
        self.n_segments = L.fjp_model_num_segments(handle)
        self.describe = L.fjp_model_describe(handle).decode()
        self.device_bytes = L.fjp_model_device_bytes(handle)

    def release(self):
        if self._h and self._owned:
            lib().fjp_model_release(self._h)
        self._h = None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


def compile_model(pmml: bytes, engine: "Engine | None" = None) -> Model:
    """fjp_model_compile.  engine=None compiles without uploading (inspection; cannot score)."""
    if isinstance(pmml, str):
        pmml = pmml.encode()
    out = C.c_void_p()
    err = C.create_string_buffer(1024)
    rc = lib().fjp_model_compile(engine._h if engine else None, pmml, len(pmml), C.byref(out), err, 1024)
    if rc != 0:
        raise ModelLoadingException(err.value.decode(errors="replace"))
    return Model(out.value)


class Engine:
    def __init__(self, device: int = 0):
        out = C.c_void_p()
        err = C.create_string_buffer(512)
        if lib().fjp_engine_create(device, C.byref(out), err, 512) != 0:
            raise EngineError(err.value.decode(errors="replace"))
        self._h = out.value
        self.device = device

    def close(self):
        if self._h:
            lib().fjp_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def compile(self, pmml: bytes) -> Model:
        return compile_model(pmml, self)

    def replicate(self, model: Model) -> Model:
        out = C.c_void_p()
        err = C.create_string_buffer(512)
        if lib().fjp_model_replicate(self._h, model._h, C.byref(out), err, 512) != 0:
            raise EngineError(err.value.decode(errors="replace"))
        return Model(out.value)

    # ---- table
    def table_add(self, name: str, version: int, model: Model) -> int:
        return _check(lib().fjp_table_add(self._h, name.encode(), version, model._h))

    def table_del(self, name: str, version: int) -> int:
        return _check(lib().fjp_table_del(self._h, name.encode(), version))

    def table_get(self, name: str, version: int):
        h = lib().fjp_table_get(self._h, name.encode(), version)
        return Model(h) if h else None

    @property
    def table_size(self):
        return lib().fjp_table_size(self._h)

    @property
    def table_epoch(self):
        return lib().fjp_table_epoch(self._h)

    @property
    def launch_count(self):
        return lib().fjp_engine_launch_count(self._h)

    @property
    def status_counts(self):
        out = (C.c_uint64 * 6)()
        lib().fjp_engine_status_counts(self._h, out)
        return list(out)

    @property
    def tensor_path_counts(self):
        """(events scored by the tensor-core path so far, events of those re-scored by the FP64 kernels)"""
        return self.fast_path_counts[:2]

    @property
    def fast_path_counts(self):
        """(tensor path: events, handed back; FMA dense path: events, handed back)"""
        out = (C.c_uint64 * 4)()
        _check(lib().fjp_engine_fast_path_counts(self._h, out))
        return tuple(int(v) for v in out)

    def sync(self):
        _check(lib().fjp_engine_sync(self._h))

    def score_sparse(self, model, row_ptr, indices, values, vec_size, replace=None):
        """CSR rows (SparseVector batch) through fjp_score_sparse."""
        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int64); indices = np.ascontiguousarray(indices, dtype=np.int32)
        values = np.ascontiguousarray(values, dtype=np.float64); n = len(row_ptr) - 1
        rp = None if replace is None else np.ascontiguousarray(np.broadcast_to(np.asarray(replace, dtype=np.float64), (n,)))
        score = np.empty(n, dtype=np.float64); status = np.empty(n, dtype=np.uint8); aux = np.empty(n, dtype=np.int32)
        _check(lib().fjp_score_sparse(self._h, model._h if model else None, _ptr(row_ptr), _ptr(indices), _ptr(values), _ptr(rp), n, vec_size,
                                      _ptr(score), _ptr(status), _ptr(aux)))
        return {"score": score, "status": status, "aux": aux}

    def float_row_modes(self, model):
        """Per vector slot: 0 cast, 1 round up, 2 round down -- or None when the model needs double rows."""
        F = lib().fjp_model_float_row_modes(model._h, None, 0)
        if F == 0:
            return None
        modes = np.zeros(F, dtype=np.uint8)
        lib().fjp_model_float_row_modes(model._h, _ptr(modes), F)
        return modes

    @staticmethod
    def narrow_rows(X, modes):
        """What a binding does when it packs a micro-batch of float rows: every double becomes the float the model's tests see."""
        X = np.asarray(X, dtype=np.float64)
        with np.errstate(over="ignore"):
            f = X.astype(np.float32)
        back = f.astype(np.float64)
        up = (back < X) & (modes[None, :] == 1)          # rounded down by the cast but must round up
        dn = (back > X) & (modes[None, :] == 2)
        f = np.where(up, np.nextafter(f, np.float32(np.inf)), f)
        f = np.where(dn, np.nextafter(f, np.float32(-np.inf)), f)
        return np.ascontiguousarray(f, dtype=np.float32)

    def score_f32(self, model, Xf, present_bits=None):
        Xf = np.ascontiguousarray(Xf, dtype=np.float32)
        n, F = Xf.shape
        pb = None if present_bits is None else np.ascontiguousarray(present_bits, dtype=np.uint32)
        score = np.empty(n, dtype=np.float64); status = np.empty(n, dtype=np.uint8); aux = np.empty(n, dtype=np.int32)
        _check(lib().fjp_score_batch_f32(self._h, model._h if model else None, _ptr(Xf), _ptr(pb), n, F, _ptr(score), _ptr(status), _ptr(aux)))
        return {"score": score, "status": status, "aux": aux}

    # ---- scoring (host buffers)
    def score(self, model: "Model | None", X, present_bits=None, replace=None, node_ids=False, n_probs=0, aux=True):
        X = np.ascontiguousarray(X, dtype=np.float64)
        if X.ndim == 1:
            X = X[None, :]
        n, F = X.shape
        pb = None
        if present_bits is not None:
            pb = np.ascontiguousarray(present_bits, dtype=np.uint32)
            assert pb.shape == (n, (F + 31) // 32), pb.shape
        rp = None
        if replace is not None:
            rp = np.ascontiguousarray(np.broadcast_to(np.asarray(replace, dtype=np.float64), (n,)))
        score = np.empty(n, dtype=np.float64)
        status = np.empty(n, dtype=np.uint8)
        ax = np.empty(n, dtype=np.int32) if aux else None
        ns = max(1, model.n_segments) if (model and node_ids) else 0
        ids = np.empty((n, ns), dtype=np.int32) if ns else None
        pr = np.empty((n, n_probs), dtype=np.float64) if n_probs else None
        _check(lib().fjp_score_batch_ex(self._h, model._h if model else None, _ptr(X), _ptr(pb), _ptr(rp), n, F,
                                        _ptr(score), _ptr(status), _ptr(ax), _ptr(ids), ns, _ptr(pr), n_probs))
        out = {"score": score, "status": status}
        if aux:
            out["aux"] = ax
        if ns:
            out["node_ids"] = ids
        if n_probs:
            out["probs"] = pr
        return out
