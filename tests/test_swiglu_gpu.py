"""GPU parity: SwiGLU forward/backward through the C-ABI vs the oracle and the golden vectors."""

import pytest
import torch

from oracle import regions as O
from parity import assert_parity

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _check(a, b, g, what):
    from b200ex import ops

    out = ops.swiglu_fwd(a.to(DEV), b.to(DEV))
    da, db = ops.swiglu_bwd(a.to(DEV), b.to(DEV), g.to(DEV))
    torch.cuda.synchronize()
    assert_parity(out, O.swiglu_fwd(a, b), what=what + " out")
    eda, edb = O.swiglu_bwd(a, b, g)
    assert_parity(db, edb, what=what + " db")
    # da = round(p) * round(q) with p, q themselves chains of storage-dtype roundings of exp/div results:
    # a last-bit difference in exp() can flip one of ~7 intermediate roundings by an ulp.
    assert_parity(da, eda, what=what + " da", ulp_scale=eda.float().abs() * 4 + (b.float() * g.float()).abs() * 1.0)


def test_golden_vectors(golden):
    from b200ex import ops

    for name, c in golden.items():
        if not name.startswith("swiglu_"):
            continue
        i, o = c["inputs"], c["outputs"]
        out = ops.swiglu_fwd(i["a"].to(DEV), i["b"].to(DEV))
        da, db = ops.swiglu_bwd(i["a"].to(DEV), i["b"].to(DEV), i["g"].to(DEV))
        assert_parity(out, o["out"], what=name + " out")
        assert_parity(db, o["db"], what=name + " db")
        assert_parity(da, o["da"], what=name + " da", ulp_scale=o["da"].float().abs() * 4 + (i["b"].float() * i["g"].float()).abs())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("shape", [(1,), (7,), (3, 77), (64, 1024), (5, 14336), (2, 3, 1000)])
def test_shapes(dtype, shape):
    g = torch.Generator().manual_seed(sum(shape))
    mk = lambda: (torch.randn(*shape, generator=g) * 2).to(dtype)
    _check(mk(), mk(), mk(), f"{dtype} {shape}")


def test_extreme_values():
    a = torch.tensor([-100.0, -20.0, -1e-3, 0.0, 1e-3, 20.0, 100.0, -88.0], dtype=torch.bfloat16)
    b = torch.ones_like(a) * 3
    _check(a, b, torch.ones_like(a), "extremes")


def test_empty():
    from b200ex import ops

    z = torch.empty(0, 16, dtype=torch.bfloat16, device=DEV)
    assert ops.swiglu_fwd(z, z).shape == (0, 16)
    assert ops.swiglu_bwd(z, z, z)[0].shape == (0, 16)


def test_full_size_block_shape():
    """Llama-3-8B MLP activation: bf16 [8192 tokens, 14336]."""
    g = torch.Generator().manual_seed(0)
    mk = lambda: (torch.randn(8192, 14336, generator=g)).to(torch.bfloat16)
    _check(mk(), mk(), mk(), "full")
