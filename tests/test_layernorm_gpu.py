"""GPU parity: LayerNorm forward/backward through the C-ABI vs the oracle and the golden vectors."""

import pytest
import torch

from oracle import regions as O
from parity import assert_parity

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _run_cuda(x, w, b, dy, eps):
    from b200ex import ops

    c = lambda t: None if t is None else t.to(DEV)
    y, mean, rstd = ops.layernorm_fwd(c(x), c(w), c(b), eps)
    dx, dw, db = ops.layernorm_bwd(c(dy), c(x), c(w), mean, rstd, b is not None)
    torch.cuda.synchronize()
    return y, mean, rstd, dx, dw, db


def _scales(x, w, dy, mean, rstd):
    """Magnitudes of the intermediates whose last-bit differences (reduction order of the row statistics, one-ulp
    flips of the storage-rounded terms) bound the error of each output; see parity.py."""
    f32 = x.dtype == torch.float32
    xf, g = x.float(), dy.float()
    wf = w.float() if w is not None else torch.ones(x.shape[-1])
    xc = xf - mean
    gw = wf * g
    H = x.shape[-1]
    # y: an ulp of the mean moves (x - mean) by ulp(|mean|); same for rstd
    y_scale = 4.0 * ((xf.abs() + mean.abs()) * rstd * wf.abs())
    direct = (rstd * gw).abs()
    c1_abs = (xc * gw).abs().sum(-1, keepdim=True)
    c2_abs = direct.sum(-1, keepdim=True)
    stats = c2_abs / H + (c1_abs * rstd**3) * xc.abs() / H
    dx_scale = (direct + stats) * (8.0 if f32 else 1.0)
    terms_w = ((xc * rstd) * g).to(x.dtype).float().abs()
    terms_b = g.abs()
    col = lambda t: t.sum(0) * 4.0 if f32 else t.max(0).values + t.sum(0) * 2.0**-12
    return y_scale, dx_scale, col(terms_w), col(terms_b)


def _check(x, w, b, dy, eps, what):
    y, mean, rstd, dx, dw, db = _run_cuda(x, w, b, dy, eps)
    ey, emean, erstd = O.layernorm_fwd(x, w, b, eps)
    edx, edw, edb = O.layernorm_bwd(dy, x, w, b is not None, emean, erstd)
    ys, dxs, dws, dbs = _scales(x, w, dy, emean, erstd)
    assert_parity(mean, emean, what=what + " mean", ulp_scale=x.float().abs().mean(-1, keepdim=True) * 4.0 if x.dtype == torch.float32 else None, atol_extra=1e-6)
    assert_parity(rstd, erstd, what=what + " rstd")
    assert_parity(y, ey, what=what + " y", ulp_scale=ys)
    assert_parity(dx, edx, what=what + " dx", ulp_scale=dxs)
    if w is not None:
        assert_parity(dw, edw, what=what + " dw", ulp_scale=dws)
    else:
        assert dw is None
    if b is not None:
        assert_parity(db, edb, what=what + " db", ulp_scale=dbs)
    else:
        assert db is None


def test_golden_vectors(golden):
    n = 0
    for name, c in golden.items():
        if not name.startswith("layernorm_"):
            continue
        n += 1
        i, o = c["inputs"], c["outputs"]
        y, mean, rstd, dx, dw, db = _run_cuda(i["x"], i["w"], i.get("b"), i["dy"], c["meta"]["eps"])
        _, emean, erstd = O.layernorm_fwd(i["x"], i["w"], i.get("b"), c["meta"]["eps"])
        ys, dxs, dws, dbs = _scales(i["x"], i["w"], i["dy"], emean, erstd)
        assert_parity(y, o["y"], what=name + " y", ulp_scale=ys)
        assert_parity(dx, o["dx"], what=name + " dx", ulp_scale=dxs)
        assert_parity(dw, o["dw"], what=name + " dw", ulp_scale=dws)
        if c["meta"]["bias"]:
            assert_parity(db, o["db"], what=name + " db", ulp_scale=dbs)
    assert n >= 5


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize(
    "rows,hidden",
    [(1, 8), (3, 32), (33, 256), (17, 768), (7, 1000), (5, 1016), (40, 1024), (12, 1600), (64, 4096), (9, 8192), (4, 16384), (2, 32768), (6, 100), (2, 4097)],
)
def test_shapes(dtype, rows, hidden):
    g = torch.Generator().manual_seed(rows * 131 + hidden)
    mk = lambda *s: torch.randn(*s, generator=g).to(dtype)
    x = (torch.randn(rows, hidden, generator=g) * 1.5 + 0.5).to(dtype)  # a non-zero mean exercises the centring
    _check(x, mk(hidden), mk(hidden), mk(rows, hidden), 1e-5, f"{dtype} {rows}x{hidden}")


@pytest.mark.parametrize("has_w,has_b", [(True, False), (False, False), (False, True)])
def test_optional_operands(has_w, has_b):
    g = torch.Generator().manual_seed(7)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    rows, hidden = 17, 2048
    _check(mk(rows, hidden), mk(hidden) if has_w else None, mk(hidden) if has_b else None, mk(rows, hidden), 1e-6, f"w={has_w} b={has_b}")


def test_many_rows_small_hidden():
    """More row groups than resident CTAs: the ring wraps several times per CTA and every CTA flushes partials."""
    g = torch.Generator().manual_seed(11)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    rows, hidden = 20001, 256
    _check(mk(rows, hidden), mk(hidden), mk(hidden), mk(rows, hidden), 1e-5, "20001x256")


def test_empty_batch():
    from b200ex import ops

    x = torch.empty(0, 512, dtype=torch.bfloat16, device=DEV)
    w = torch.ones(512, dtype=torch.bfloat16, device=DEV)
    y, mean, rstd = ops.layernorm_fwd(x, w, w, 1e-5)
    assert y.shape == (0, 512) and mean.shape == (0, 1) and rstd.shape == (0, 1)
    dx, dw, db = ops.layernorm_bwd(x, x, w, mean, rstd, True)
    torch.cuda.synchronize()
    assert dx.shape == (0, 512) and torch.count_nonzero(dw).item() == 0 and torch.count_nonzero(db).item() == 0


def test_3d_input():
    g = torch.Generator().manual_seed(3)
    x = torch.randn(5, 2, 512, generator=g).to(torch.bfloat16).transpose(0, 1)  # non-contiguous view
    w, b = torch.randn(512, generator=g).to(torch.bfloat16), torch.randn(512, generator=g).to(torch.bfloat16)
    from b200ex import ops

    y, mean, rstd = ops.layernorm_fwd(x.to(DEV), w.to(DEV), b.to(DEV), 1e-5)
    ey, emean, erstd = O.layernorm_fwd(x, w, b, 1e-5)
    assert y.shape == x.shape and rstd.shape == (2, 5, 1) and mean.shape == (2, 5, 1)
    assert_parity(y, ey, what="3d y", ulp_scale=ey.float().abs() + 1.0), assert_parity(rstd, erstd, what="3d rstd")


def test_full_size():
    """bf16 [8192 tokens x 4096 hidden] (the RMSNorm headline shape), fwd + bwd, against the oracle."""
    g = torch.Generator().manual_seed(0)
    rows, hidden = 8192, 4096
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    _check(mk(rows, hidden) + 1.0, mk(hidden), mk(hidden), mk(rows, hidden), 1e-5, "full-size")


def test_full_size_properties():
    """Size-independent properties at the full size: shift invariance and scale equivariance of the forward,
    linearity of the backward in dy, dx rows orthogonal to both 1 and (x - mean), db / dw equal to independent
    column sums."""
    from b200ex import ops

    torch.manual_seed(1)
    rows, hidden = 8192, 4096
    x = torch.randn(rows, hidden, device=DEV, dtype=torch.float32)
    w = torch.randn(hidden, device=DEV, dtype=torch.float32)
    b = torch.randn(hidden, device=DEV, dtype=torch.float32)
    dy = torch.randn(rows, hidden, device=DEV, dtype=torch.float32)
    y1, mean1, rstd1 = ops.layernorm_fwd(x, w, b, 0.0)
    y2, mean2, rstd2 = ops.layernorm_fwd(x * 4.0 + 3.0, w, b, 0.0)
    torch.testing.assert_close(y1, y2, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(mean2, mean1 * 4.0 + 3.0, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(rstd1, rstd2 * 4.0, rtol=1e-5, atol=1e-6)
    dx1, dw1, db1 = ops.layernorm_bwd(dy, x, w, mean1, rstd1, True)
    dx2, dw2, db2 = ops.layernorm_bwd(dy * 2.0, x, w, mean1, rstd1, True)
    torch.testing.assert_close(dx2, dx1 * 2.0, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dw2, dw1 * 2.0, rtol=1e-5, atol=1e-4)
    torch.testing.assert_close(db2, db1 * 2.0, rtol=1e-5, atol=1e-4)
    assert dx1.double().sum(-1).abs().max().item() < 1e-2
    assert (dx1.double() * (x - mean1).double()).sum(-1).abs().max().item() < 1e-2
    torch.testing.assert_close(db1, dy.sum(0), rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(dw1, (((x - mean1) * rstd1) * dy).sum(0), rtol=1e-4, atol=1e-3)
