"""not-gpu: the arithmetic of the int8 pre-filter's lower bounds, restated in numpy with the same fp32
steps the sidecar build (csrc/lmd_sidecar.cu) and the traversal kernel (csrc/lmd_kernels.cuh, ROUTE = 3) take, checked
against float64 distances on data with awkward scales: a bound that ever exceeds the true distance (plus the kernel's
margin) would let the pre-filter drop a row the reference keeps. The GPU parity tests check the kernels; this one checks
the mathematics they rely on."""
import numpy as np
import pytest

F = np.float32


def _data(n, d, seed, row_scale):
    rng = np.random.default_rng(seed)
    z = rng.standard_normal((n, 16)).astype(F)
    proj = np.linalg.qr(rng.standard_normal((d, 16)))[0].T.astype(F)
    x = (z @ proj + F(0.05) * rng.standard_normal((n, d)).astype(F)).astype(F)
    if row_scale:
        x *= rng.uniform(0.05, 20.0, (n, 1)).astype(F)
    q = (x[:96] + F(0.05) * rng.standard_normal((96, d)).astype(F)) * rng.uniform(0.01, 100.0, (96, 1)).astype(F)
    return x, q.astype(F), rng


def _encode(v, scale):
    """clamp(rint(v / scale)) and the reconstruction error, inflated as the kernels do"""
    c = np.clip(np.rint(v / scale), -127, 127).astype(np.int32)
    r = v - scale * c.astype(F)
    return c, (np.sqrt((r * r).sum(1)).astype(F) * F(1.0001)).astype(F)


def _scale(rows, centred):
    m = rows.astype(np.float64).mean(0) if centred else np.zeros(rows.shape[1])
    var = (rows.astype(np.float64) ** 2).mean(0) - m * m
    return m.astype(F), F(4.0 * np.sqrt(np.maximum(var, 0)).max() / 127.0)


@pytest.mark.parametrize("d,row_scale", [(128, False), (128, True), (768, True), (100, True)])
def test_l2_bound(d, row_scale):
    x, q, _ = _data(2500, d, 3, row_scale)
    q = x[:96] + F(0.3) * (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(F)  # queries on the data's scale
    off, s = _scale(x, True)
    cx, ex = _encode(x - off, s)
    cq, eq = _encode(q - off, s)
    i2 = (cq * cq).sum(1)[:, None] - 2 * (cq @ cx.T) + (cx * cx).sum(1)[None, :]
    grid = s * np.sqrt(np.maximum(i2, 0).astype(F))
    true = np.sqrt(((q[:, None, :].astype(np.float64) - x[None, :, :].astype(np.float64)) ** 2).sum(2))
    # the kernel skips when grid > (worst + eq + ex) * 1.0005, i.e. it relies on  grid / 1.0005 - eq - ex <= d
    lb = grid / F(1.0005) - eq[:, None] - ex[None, :]
    assert (lb <= true).all(), float((lb - true).max())


@pytest.mark.parametrize("d,row_scale", [(128, False), (128, True), (768, True), (100, True)])
def test_cosine_bound(d, row_scale):
    x, q, _ = _data(2500, d, 4, row_scale)
    xu = (x / np.linalg.norm(x, axis=1, keepdims=True).astype(F)).astype(F)
    qu = (q / np.linalg.norm(q, axis=1, keepdims=True).astype(F)).astype(F)
    _, s = _scale(xu, False)
    cx, ex = _encode(xu, s)
    cq, eq = _encode(qu, s)
    ub = s * s * (cq @ cx.T).astype(F) + eq[:, None] * (F(1.0) + ex[None, :]) + ex[None, :]
    lb = F(1.0) - ub
    x64, q64 = x.astype(np.float64), q.astype(np.float64)
    cos = (q64 @ x64.T) / (np.linalg.norm(q64, axis=1)[:, None] * np.linalg.norm(x64, axis=1)[None, :])
    true = 1.0 - np.clip(cos, -1.0, 1.0)
    assert (lb <= true + 2e-4).all(), float((lb - true).max())
    assert np.median(true - lb) < 0.05  # ... and it is tight enough to be worth having


@pytest.mark.parametrize("d,row_scale", [(128, False), (128, True), (768, True), (100, True)])
def test_inner_product_bound(d, row_scale):
    x, q, _ = _data(2500, d, 5, row_scale)
    ss = (x.astype(np.float64) ** 2).mean(0)
    _, s = _scale(x, False)
    rho = F(np.sqrt(ss.sum()))
    qn = np.linalg.norm(q, axis=1).astype(F)
    cx, ex = _encode(x, s)
    cq, eq = _encode((q * (rho / qn)[:, None]).astype(F), s)
    xh = s * np.sqrt((cx * cx).sum(1).astype(F))
    ub = (qn / rho)[:, None] * (s * s * (cq @ cx.T).astype(F) + eq[:, None] * xh[None, :] + rho * ex[None, :])
    margin = F(1e-4) * qn[:, None] * (xh + ex)[None, :]
    true = -(q.astype(np.float64) @ x.astype(np.float64).T)
    assert (-ub <= true + margin).all()
    assert (-ub <= true).all()  # in fact without the margin, which is only there for the kernel's own fp32 dot product
    rel = (true + ub) / (qn[:, None] * np.linalg.norm(x, axis=1)[None, :])
    assert np.median(rel) < 0.08


def test_zero_rows_and_queries_are_never_excluded():
    """err = +inf for a zero (or non-finite-reciprocal) row / query makes every bound -inf or NaN: `bound > worst` is false."""
    inf = F(np.inf)
    with np.errstate(invalid="ignore"):
        for q_err, c_err in ((inf, F(0.01)), (F(0.01), inf), (inf, inf)):
            ub = F(0.5) + q_err * (F(1.0) + c_err) + c_err
            assert not (F(1.0) - ub > F(0.0))
        ub_ip = (F(0.0) / F(1.0)) * (F(0.0) + inf * F(1.0) + F(1.0) * F(0.0))  # zero query: 0 * inf
        assert np.isnan(ub_ip) and not (-ub_ip > F(-1.0))
