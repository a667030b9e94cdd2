"""GPU: fjp_score_sparse_noalloc -- the sparse scoring call for callers that cannot use the device allocator.

Some users of this library run where cudaMalloc is off the table while events are flowing: a subtask that shares its GPU
with other jobs under a fixed memory budget (an allocation can fail or stall the whole device), or a latency-bound
operator.  They pre-allocate their CSR input and their outputs once and call the no-alloc entry point; its only working
memory is an 8 MiB area built into the library.  This test drives it through the C ABI, twice in a row with different
batches (the area is shared and keeps the previous call's contents), and requires the same answers -- bit for bit, the
tolerance the other tree tests use -- as the existing fjp_score_sparse, which stages through allocated arenas.

The file name sorts last on purpose: if this addition fails, a run that stops at the first failure has already run
everything else."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _csr(gen, n, F, seed, p_missing):
    X, pres, pb = gen.events(n, F, seed, p_missing)
    counts = pres.sum(1)
    row_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    rows, cols = np.nonzero(pres)
    return row_ptr, cols.astype(np.int32), np.ascontiguousarray(X[rows, cols])


def _noalloc(lib, engine, model, row_ptr, indices, values, F):
    import torch
    n = len(row_ptr) - 1
    # the CALLER owns inputs and outputs (here: torch tensors); the entry point itself allocates nothing
    d_rp, d_ix, d_v = (torch.from_numpy(a).cuda() for a in (row_ptr, indices, values))
    d_score = torch.full((n,), -1.0, dtype=torch.float64, device="cuda")
    d_status = torch.full((n,), 77, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    rc = lib.fjp_score_sparse_noalloc(engine._h, model._h, d_rp.data_ptr(), d_ix.data_ptr(), d_v.data_ptr(), n, F,
                                      d_score.data_ptr(), d_status.data_ptr(), None)
    torch.cuda.synchronize()
    return rc, d_score.cpu().numpy(), d_status.cpu().numpy()


def test_noalloc_sparse_matches_allocating_entry_point(engine, gen):
    from flink_jpmml_b200 import capi
    lib = capi.lib()
    F = 100
    pm, _ = gen.gbt_pmml(T=48, depth=8, F=F, seed=77)
    model = engine.compile(pm)
    # two calls in a row: a large batch, then a smaller and much sparser one that reuses the same built-in area
    for n, seed, p_missing in ((9000, 1, 0.3), (2500, 2, 0.85)):
        row_ptr, indices, values = _csr(gen, n, F, seed, p_missing)
        ref = engine.score_sparse(model, row_ptr, indices, values, F)
        rc, score, status = _noalloc(lib, engine, model, row_ptr, indices, values, F)
        assert rc == 0, lib.fjp_last_error()
        assert np.array_equal(status, ref["status"])
        assert np.array_equal(score, ref["score"]), "no-alloc path differs from fjp_score_sparse"
        assert (status == 0).all()
    # the point of the entry point: once a model's kernels are loaded (the warm-up calls above), a call takes nothing from
    # the device -- the free-memory figure and the default pool's reservation are the same before and after
    import torch
    row_ptr, indices, values = _csr(gen, 4000, F, 9, 0.5)
    d_rp, d_ix, d_v = (torch.from_numpy(a).cuda() for a in (row_ptr, indices, values))
    d_score = torch.empty(4000, dtype=torch.float64, device="cuda"); d_status = torch.empty(4000, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(2):
        assert lib.fjp_score_sparse_noalloc(engine._h, model._h, d_rp.data_ptr(), d_ix.data_ptr(), d_v.data_ptr(), 4000, F,
                                            d_score.data_ptr(), d_status.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the no-alloc entry point changed the device's free memory"
    # a batch that does not fit the built-in area is refused with a message, not truncated; the documented bound is exact
    limit_rows = (8 << 20) // (F * 8 + 4 * 4)
    row_ptr, indices, values = _csr(gen, limit_rows, F, 4, 0.9)
    rc, score, status = _noalloc(lib, engine, model, row_ptr, indices, values, F)
    assert rc == 0 and (status == 0).all(), "a batch of exactly the documented size must fit"
    row_ptr, indices, values = _csr(gen, limit_rows + 500, F, 3, 0.9)
    rc, score, status = _noalloc(lib, engine, model, row_ptr, indices, values, F)
    assert rc < 0 and b"scratch" in lib.fjp_last_error()
    assert (status == 77).all() and (score == -1.0).all()      # outputs untouched
    # wrong vector size is refused as well
    rc, _, _ = _noalloc(lib, engine, model, row_ptr[:11], indices, values, F + 1)
    assert rc < 0 and b"vec_size" in lib.fjp_last_error()
