"""GPU parity: RMSNorm(+residual) forward/backward through the C-ABI vs the oracle and the golden vectors."""

import pytest
import torch

from oracle import regions as O
from parity import assert_parity

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _run_cuda(x, r, w, dy, dh, eps):
    from b200ex import ops

    c = lambda t: None if t is None else t.to(DEV)
    h, y, rstd = ops.rmsnorm_fwd(c(x), c(r), c(w), eps)
    dx, dw = ops.rmsnorm_bwd(c(dy), h, c(w), rstd, c(dh))
    torch.cuda.synchronize()
    return h, y, rstd, dx, dw


def _check(x, r, w, dy, dh, eps, what):
    h, y, rstd, dx, dw = _run_cuda(x, r, w, dy, dh, eps)
    eh, ey, erstd = O.rmsnorm_residual_fwd(x, r, w, eps)
    edx, edw = O.rmsnorm_residual_bwd(dy, dh, eh, w, erstd)
    if r is not None:
        assert_parity(h, eh, what=what + " h")
    assert_parity(y, ey, what=what + " y")
    assert_parity(rstd, erstd, what=what + " rstd")
    # dx = dh_in + round(dh): a one-ulp flip of the rounded dh is bounded by ulp(|dh|), not by |dx|.
    scale = (edx.float() - dh.float()).abs() if dh is not None else None
    assert_parity(dx, edx, what=what + " dx", ulp_scale=scale)
    if w is not None:
        # dw sums `rows` storage-rounded terms in fp32 in a different order than ATen: the error is bounded
        # by ulps of the sum of |terms| (a cancelling column can have |dw| far below that), not of |dw|.
        # In 16-bit storage a last-bit difference in h*rstd can also flip the rounding of one term by an ulp.
        terms = ((eh.float() * erstd) * dy.float()).to(x.dtype).float().abs()
        scale = terms.sum(0) * 4.0 if x.dtype == torch.float32 else terms.max(0).values + terms.sum(0) * 2.0**-12
        assert_parity(dw, edw, what=what + " dw", ulp_scale=scale)
    else:
        assert dw is None


def test_golden_vectors(golden):
    for name, c in golden.items():
        if not name.startswith("rmsnorm_"):
            continue
        i, o = c["inputs"], c["outputs"]
        h, y, rstd, dx, dw = _run_cuda(i["x"], i["r"], i["w"], i["dy"], i["dh"], c["meta"]["eps"])
        assert_parity(h, o["h"], what=name + " h")
        assert_parity(y, o["y"], what=name + " y")
        assert_parity(dx, o["dx"], what=name + " dx", ulp_scale=(o["dx"].float() - i["dh"].float()).abs())
        assert_parity(dw, o["dw"], what=name + " dw")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize(
    "rows,hidden",
    [(1, 8), (3, 32), (33, 256), (7, 1000), (5, 1016), (64, 4096), (9, 8192), (4, 16384), (2, 32768), (3, 65536), (6, 100), (2, 4097)],
)
def test_shapes(dtype, rows, hidden):
    g = torch.Generator().manual_seed(rows * 131 + hidden)
    mk = lambda *s: torch.randn(*s, generator=g).to(dtype)
    _check(mk(rows, hidden), mk(rows, hidden), mk(hidden), mk(rows, hidden), mk(rows, hidden), 1e-5, f"{dtype} {rows}x{hidden}")


@pytest.mark.parametrize("has_res,has_w,has_dh", [(False, True, False), (True, False, True), (False, False, False), (True, True, False)])
def test_optional_operands(has_res, has_w, has_dh):
    g = torch.Generator().manual_seed(7)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    rows, hidden = 17, 2048
    _check(mk(rows, hidden), mk(rows, hidden) if has_res else None, mk(hidden) if has_w else None, mk(rows, hidden),
           mk(rows, hidden) if has_dh else None, 1e-6, f"res={has_res} w={has_w} dh={has_dh}")


def test_empty_batch():
    from b200ex import ops

    x = torch.empty(0, 512, dtype=torch.bfloat16, device=DEV)
    w = torch.ones(512, dtype=torch.bfloat16, device=DEV)
    h, y, rstd = ops.rmsnorm_fwd(x, x, w, 1e-5)
    assert y.shape == (0, 512) and rstd.shape == (0, 1)
    dx, dw = ops.rmsnorm_bwd(x, h, w, rstd, None)
    torch.cuda.synchronize()
    assert dx.shape == (0, 512) and torch.count_nonzero(dw).item() == 0


def test_3d_input_and_noncontiguous():
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 5, 512, generator=g).to(torch.bfloat16)
    r = torch.randn(5, 2, 512, generator=g).to(torch.bfloat16).transpose(0, 1)  # non-contiguous view
    w = torch.randn(512, generator=g).to(torch.bfloat16)
    from b200ex import ops

    h, y, rstd = ops.rmsnorm_fwd(x.to(DEV), r.to(DEV), w.to(DEV), 1e-5)
    eh, ey, erstd = O.rmsnorm_residual_fwd(x, r, w, 1e-5)
    assert h.shape == x.shape and rstd.shape == (2, 5, 1)
    assert_parity(h, eh, what="3d h"), assert_parity(y, ey, what="3d y"), assert_parity(rstd, erstd, what="3d rstd")


def test_full_size_baseline_config():
    """BASELINE.json configs[1]: bf16 [8192 tokens x 4096 hidden], fwd + bwd, against the oracle."""
    g = torch.Generator().manual_seed(0)
    rows, hidden = 8192, 4096
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    x, r, w, dy, dh = mk(rows, hidden), mk(rows, hidden), mk(hidden), mk(rows, hidden), mk(rows, hidden)
    _check(x, r, w, dy, dh, 1e-5, "full-size")


def test_full_size_properties():
    """Size-independent properties at the full size: scale equivariance of y, linearity of dx in dy,
    and dw equal to the column sum of the per-row terms recomputed independently on the GPU."""
    from b200ex import ops

    torch.manual_seed(1)
    rows, hidden = 8192, 4096
    x = torch.randn(rows, hidden, device=DEV, dtype=torch.float32)
    w = torch.randn(hidden, device=DEV, dtype=torch.float32)
    dy = torch.randn(rows, hidden, device=DEV, dtype=torch.float32)
    _, y1, rstd1 = ops.rmsnorm_fwd(x, None, w, 0.0)
    _, y2, rstd2 = ops.rmsnorm_fwd(x * 4.0, None, w, 0.0)
    torch.testing.assert_close(y1, y2, rtol=1e-5, atol=1e-6)  # RMSNorm is scale invariant at eps = 0
    torch.testing.assert_close(rstd1, rstd2 * 4.0, rtol=1e-5, atol=1e-6)
    dx1, dw1 = ops.rmsnorm_bwd(dy, x, w, rstd1)
    dx2, dw2 = ops.rmsnorm_bwd(dy * 2.0, x, w, rstd1)
    torch.testing.assert_close(dx2, dx1 * 2.0, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dw2, dw1 * 2.0, rtol=1e-5, atol=1e-4)
    # each row of dx is orthogonal to h (the normalised direction carries no gradient)
    dots = (dx1.double() * x.double()).sum(-1)
    assert dots.abs().max().item() < 1e-2
    torch.testing.assert_close(dw1, ((x * rstd1) * dy).sum(0), rtol=1e-4, atol=1e-3)
