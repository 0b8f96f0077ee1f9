"""GPU parity: bias + activation forward/backward through the C-ABI vs the oracle and golden vectors."""

import pytest
import torch

from oracle import regions as O
from parity import assert_parity

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ORACLE = {"gelu": (O.bias_gelu_fwd, O.bias_gelu_bwd), "relu": (O.bias_relu_fwd, O.bias_relu_bwd)}


def _check(x, b, g, act, what):
    from b200ex import ops

    out = ops.bias_act_fwd(x.to(DEV), b.to(DEV), act)
    dx, db = ops.bias_act_bwd(x.to(DEV), b.to(DEV), g.to(DEV), act)
    torch.cuda.synchronize()
    fwd, bwd = ORACLE[act]
    eout, (edx, edb) = fwd(x, b), bwd(x, b, g)
    lowp = x.dtype != torch.float32
    # gelu is a chain of storage-dtype roundings of erf()/exp(): one flipped intermediate ulp moves the result
    # by an ulp of the pre-activation / cotangent magnitude.
    a_mag = (x.float() + b.float()).abs()
    assert_parity(out, eout, what=what + " out", ulp_scale=a_mag if (lowp and act == "gelu") else None)
    assert_parity(dx, edx, what=what + " dx", ulp_scale=(g.float().abs() * (1 + a_mag)) if (lowp and act == "gelu") else None)
    terms = edx.float().abs()
    if not lowp:
        scale = terms.sum(0) * 4.0
    elif act == "gelu":
        # one flipped intermediate (cdf, pdf term) moves a single dx by an ulp of |g| * (1 + |a|), and dbias with it
        scale = (g.float().abs() * (1 + a_mag)).max(0).values * 2 + terms.sum(0) * 2.0**-12
    else:
        scale = terms.max(0).values + terms.sum(0) * 2.0**-12
    assert_parity(db, edb, what=what + " dbias", ulp_scale=scale)


@pytest.mark.parametrize("act", ["gelu", "relu"])
def test_golden_vectors(golden, act):
    for name, c in golden.items():
        if not name.startswith(f"bias_{act}_"):
            continue
        i = c["inputs"]
        _check(i["x"], i["b"], i["g"], act, name)


@pytest.mark.parametrize("act", ["gelu", "relu"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("rows,hidden", [(1, 8), (9, 60), (33, 768), (64, 4096), (7, 16384), (4, 1001), (1200, 512)])
def test_shapes(act, dtype, rows, hidden):
    g = torch.Generator().manual_seed(rows + hidden)
    mk = lambda *s: (torch.randn(*s, generator=g) * 1.5).to(dtype)
    _check(mk(rows, hidden), mk(hidden), mk(rows, hidden), act, f"{act} {dtype} {rows}x{hidden}")


def test_no_bias_and_identity():
    from b200ex import ops

    x = torch.randn(16, 256).to(torch.bfloat16)
    g = torch.randn(16, 256).to(torch.bfloat16)
    out = ops.bias_act_fwd(x.to(DEV), None, "relu")
    assert torch.equal(out.cpu(), torch.relu(x))
    dx, db = ops.bias_act_bwd(x.to(DEV), None, g.to(DEV), "relu")
    assert db is None and torch.equal(dx.cpu(), torch.where(x > 0, g, torch.zeros_like(g)))
    b = torch.randn(256).to(torch.bfloat16)
    out = ops.bias_act_fwd(x.to(DEV), b.to(DEV), "none")
    assert torch.equal(out.cpu(), x + b)


def test_saved_preactivation_form_and_colsum():
    """What the executor's regions use: the forward also returns the rounded pre-activation, the backward takes it
    instead of (x, bias) and still returns the bias gradient; colsum is the stand-alone column reduction."""
    from b200ex import ops

    g0 = torch.Generator().manual_seed(9)
    mk = lambda *s: (torch.randn(*s, generator=g0) * 1.5).to(torch.bfloat16)
    x, b, g = mk(300, 1024), mk(1024), mk(300, 1024)
    out, pre = ops.bias_act_fwd(x.to(DEV), b.to(DEV), "gelu", return_pre=True)
    assert torch.equal(pre.cpu(), x + b) and torch.equal(out.cpu(), ops.bias_act_fwd(x.to(DEV), b.to(DEV), "gelu").cpu())
    dx1, db1 = ops.bias_act_bwd(x.to(DEV), b.to(DEV), g.to(DEV), "gelu")
    dx2, db2 = ops.bias_act_bwd(pre, None, g.to(DEV), "gelu", want_dbias=True)
    assert torch.equal(dx1, dx2) and torch.equal(db1, db2)
    assert ops.bias_act_bwd(pre, None, g.to(DEV), "gelu")[1] is None
    cs = ops.colsum(dx1)
    assert torch.equal(cs, db1)  # same partial-sum tree
    ref = torch.sum(dx1.cpu(), 0)
    terms = dx1.cpu().float().abs()
    assert_parity(cs, ref, what="colsum", ulp_scale=terms.sum(0) * 2.0**-12)
    for shape, dt in (((7, 5, 96), torch.float32), ((33, 1001), torch.float16), ((2, 8), torch.bfloat16)):
        t = torch.randn(*shape, generator=g0).to(dt)
        assert_parity(ops.colsum(t.to(DEV)), t.reshape(-1, shape[-1]).sum(0), what=f"colsum {shape} {dt}",
                      ulp_scale=t.float().abs().reshape(-1, shape[-1]).sum(0) * (4.0 if dt == torch.float32 else 2.0**-8))


def test_nanogpt_mlp_shape():
    """GPT-2 style MLP activation: bf16 [8192 tokens, 3072], gelu(x + bias)."""
    g = torch.Generator().manual_seed(0)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    _check(mk(8192, 3072), mk(3072), mk(8192, 3072), "gelu", "nanogpt mlp")
