"""GPU: BASELINE.json's full sizes (2^22 events for C2 / C4, 2^20 for C3), checked through size-independent properties:
  - a permuted batch gives the permuted scores (no dependence on the tile / chunk an event lands in),
  - the host entry point (chunked copy pipeline) and the device entry point (one launch) agree bit for bit,
  - scoring twice gives the same bits,
  - a prefix of the batch equals the plain-numpy restatement of the model."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_score(engine, model, X, P, n, F):
    import torch
    from flink_jpmml_b200 import capi
    S = torch.empty(n, dtype=torch.float64, device="cuda"); T = torch.empty(n, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()          # the inputs were produced on torch's default stream; the call below runs on its own stream
    st = torch.cuda.Stream()
    rc = capi.lib().fjp_score_device(engine._h, model._h, X.data_ptr(), P.data_ptr() if P is not None else None, None, n, F,
                                     S.data_ptr(), T.data_ptr(), None, C.c_void_p(st.cuda_stream))
    assert rc == 0
    st.synchronize()
    return S, T


def _present_bits(torch, n, F, p_missing, gen_):
    W = (F + 31) // 32
    P = torch.empty(n, W, dtype=torch.int32, device="cuda")
    shifts = torch.arange(32, device="cuda", dtype=torch.int64)
    for lo in range(0, n, 1 << 18):
        hi = min(n, lo + (1 << 18))
        bits = torch.zeros(hi - lo, W * 32, dtype=torch.int64, device="cuda")
        bits[:, :F] = (torch.rand(hi - lo, F, generator=gen_, device="cuda") >= p_missing).to(torch.int64)
        w = (bits.view(hi - lo, W, 32) << shifts).sum(-1)
        P[lo:hi] = torch.where(w >= 2 ** 31, w - 2 ** 32, w).to(torch.int32)
    return P


def _unpack(P, F):
    bits = P.cpu().numpy().view(np.uint32)
    return ((bits[:, :, None] >> np.arange(32, dtype=np.uint32)[None, None, :]) & 1).reshape(len(bits), -1)[:, :F].astype(bool)


@pytest.mark.parametrize("config", ["C4", "C2", "C3"])
def test_full_size_properties(engine, gen, config):
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(20261106)
    if config == "C4":
        pm, arr = gen.gbt_pmml(); F, n, p_missing = 100, 1 << 22, 0.10
        ref = lambda X, pres: gen.gbt_numpy(arr, X, pres)
    elif config == "C2":
        pm, arr = gen.config_c2(); F, n, p_missing = 128, 1 << 22, 0.05
        ref = lambda X, pres: gen.kmeans_numpy(arr["centroids"], X, pres)
    else:
        pm, arr = gen.regression_pmml(); F, n, p_missing = 1000, 1 << 20, 0.0
        ref = lambda X, pres: gen.regression_numpy(arr, X).argmax(1).astype(np.float64)
    model = engine.compile(pm)
    X = torch.randn(n, F, generator=g, device="cuda", dtype=torch.float64)
    P = _present_bits(torch, n, F, p_missing, g) if p_missing else None
    S, T = _device_score(engine, model, X, P, n, F)
    assert int((T == 0).sum()) == n
    S2, _ = _device_score(engine, model, X, P, n, F)
    assert torch.equal(S, S2)                                              # idempotent
    # permutation: gather a permuted half-batch on the device (keeps memory in bounds) and compare with the permuted scores
    m = n // 2
    perm = torch.randperm(n, generator=g, device="cuda")[:m]
    Sp, _ = _device_score(engine, model, X[perm].contiguous(), P[perm].contiguous() if P is not None else None, m, F)
    assert torch.equal(Sp, S[perm])
    # host entry point on a 2^18 slice that straddles chunk boundaries, and numpy on a prefix
    k = 1 << 18
    Xh = X[1000:1000 + k].cpu().numpy(); Ph = P[1000:1000 + k].cpu().numpy().view(np.uint32) if P is not None else None
    r = engine.score(model, Xh, Ph)
    assert np.array_equal(r["score"], S[1000:1000 + k].cpu().numpy())
    q = 4096
    pres = _unpack(P[:q], F) if P is not None else None
    assert np.array_equal(S[:q].cpu().numpy(), ref(X[:q].cpu().numpy(), pres))
