"""GPU parity: softmax over the last dimension, forward/backward through the C-ABI vs the oracle and golden vectors."""

import pytest
import torch

from oracle import regions as O
from parity import assert_parity

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _check(x, g, what):
    from b200ex import ops

    y = ops.softmax_fwd(x.to(DEV))
    ey = O.softmax_fwd(x)
    # probabilities: an ulp of exp() near the row maximum moves every element by an ulp of its own magnitude
    assert_parity(y, ey, what=what + " y", ulp_scale=ey.float().abs() * 4 if x.dtype == torch.float32 else None)
    # the backward is fed the ORACLE's y so that both sides start from the same bits
    dx = ops.softmax_bwd(g.to(DEV), ey.to(DEV))
    edx = O.softmax_bwd(g, ey)
    # dx = y * (g - s): the row sum s differs by reduction order (fp32) or by an ulp of its rounded value (16-bit)
    s_abs = (ey.float() * g.float()).abs().sum(-1, keepdim=True)
    scale = ey.float() * (g.float().abs() + s_abs) * (8.0 if x.dtype == torch.float32 else 1.0)
    assert_parity(dx, edx, what=what + " dx", ulp_scale=scale + 1e-30)
    torch.cuda.synchronize()


def test_golden_vectors(golden):
    from b200ex import ops

    n = 0
    for name, c in golden.items():
        if not name.startswith("softmax_"):
            continue
        n += 1
        i, o = c["inputs"], c["outputs"]
        y = ops.softmax_fwd(i["x"].to(DEV))
        assert_parity(y, o["y"], what=name + " y", ulp_scale=o["y"].float().abs() * 4 if i["x"].dtype == torch.float32 else None)
        dx = ops.softmax_bwd(i["g"].to(DEV), o["y"].to(DEV))
        s_abs = (o["y"].float() * i["g"].float()).abs().sum(-1, keepdim=True)
        assert_parity(dx, o["dx"], what=name + " dx", ulp_scale=o["y"].float() * (i["g"].float().abs() + s_abs) * 8.0 + 1e-30)
    assert n >= 5


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("rows,cols", [(1, 8), (3, 32), (33, 256), (17, 1000), (5, 1016), (64, 1024), (9, 4096), (4, 16384), (2, 32768), (3, 40000), (6, 100), (2, 4097)])
def test_shapes(dtype, rows, cols):
    g0 = torch.Generator().manual_seed(rows * 131 + cols)
    x = (torch.randn(rows, cols, generator=g0) * 3).to(dtype)
    g = torch.randn(rows, cols, generator=g0).to(dtype)
    _check(x, g, f"{dtype} {rows}x{cols}")


def test_attention_shape_and_masked_rows():
    """[B, H, T, T] scores with a causal -inf mask: leading dims are flattened, masked entries get exactly 0."""
    from b200ex import ops

    g0 = torch.Generator().manual_seed(5)
    B, H, T = 2, 4, 128
    x = torch.randn(B, H, T, T, generator=g0).to(torch.bfloat16)
    mask = torch.ones(T, T, dtype=torch.bool).tril().logical_not()
    x = x.masked_fill(mask, float("-inf"))
    y = ops.softmax_fwd(x.to(DEV))
    assert y.shape == x.shape
    assert_parity(y, O.softmax_fwd(x), what="causal y")
    assert torch.count_nonzero(y.cpu()[..., mask]).item() == 0
    torch.testing.assert_close(y.float().sum(-1).cpu(), torch.ones(B, H, T), rtol=0, atol=2e-2)


def test_properties_at_full_size():
    """Size-independent properties at an attention-sized problem (fp32 [32768 x 2048]): rows sum to one, shift
    invariance, backward rows sum to zero and are linear in dy."""
    from b200ex import ops

    torch.manual_seed(2)
    x = torch.randn(32768, 2048, device=DEV) * 2
    g = torch.randn(32768, 2048, device=DEV)
    y = ops.softmax_fwd(x)
    torch.testing.assert_close(y.sum(-1), torch.ones(32768, device=DEV), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(ops.softmax_fwd(x + 3.0), y, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(y, torch.softmax(x, -1), rtol=1e-5, atol=1e-7)
    dx = ops.softmax_bwd(g, y)
    assert dx.sum(-1).abs().max().item() < 1e-5
    torch.testing.assert_close(ops.softmax_bwd(g * 2, y), dx * 2, rtol=1e-5, atol=1e-7)


def test_empty():
    from b200ex import ops

    x = torch.empty(0, 64, dtype=torch.bfloat16, device=DEV)
    assert ops.softmax_fwd(x).shape == (0, 64) and ops.softmax_bwd(x, x).shape == (0, 64)
