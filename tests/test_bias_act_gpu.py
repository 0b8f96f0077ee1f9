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


_FP32_FORM = r"""
import os, sys, torch
os.environ["B200K_BIAS_ACT_PACKED"] = "0"
sys.path[:0] = [sys.argv[1]]
from b200ex import ops
d = torch.load(sys.argv[2])
x, b, g = (d[k].cuda() for k in ("x", "b", "g"))
out, pre = ops.bias_act_fwd(x, b, "gelu", return_pre=True)
dx, db = ops.bias_act_bwd(x, b, g, "gelu")
torch.save({"out": out.cpu(), "pre": pre.cpu(), "dx": dx.cpu(), "db": db.cpu()}, sys.argv[3])
"""


def test_packed_bf16_gelu_equals_the_fp32_form_bit_for_bit(tmp_path):
    """The bf16 GELU chain runs on packed bf16x2 multiplies / adds where both operands are bf16 values; that is
    claimed to round exactly like fp32-op-then-round.  Run the fp32 form in a second process
    (B200K_BIAS_ACT_PACKED=0 is read once per process) on inputs that cover every bf16 exponent the activation
    sees, including subnormal products, and compare all outputs bitwise."""
    import os
    import subprocess
    import sys

    from b200ex import ops

    g0 = torch.Generator().manual_seed(5)
    rows, hidden = 512, 2048
    mag = torch.exp2(torch.randint(-30, 8, (rows, hidden), generator=g0).float())
    x = (torch.randn(rows, hidden, generator=g0) * mag).to(torch.bfloat16)
    x[:64] = (torch.randn(64, hidden, generator=g0) * 3).to(torch.bfloat16)  # the interesting range of erf
    b = (torch.randn(hidden, generator=g0) * 0.5).to(torch.bfloat16)
    g = (torch.randn(rows, hidden, generator=g0) * torch.exp2(torch.randint(-20, 4, (rows, hidden), generator=g0).float())).to(torch.bfloat16)
    src, dst = tmp_path / "in.pt", tmp_path / "out.pt"
    torch.save({"x": x, "b": b, "g": g}, src)
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lightning-thunder_b200")
    subprocess.run([sys.executable, "-c", _FP32_FORM, pkg, str(src), str(dst)], check=True, timeout=300)
    ref = torch.load(dst)
    out, pre = ops.bias_act_fwd(x.to(DEV), b.to(DEV), "gelu", return_pre=True)
    dx, db = ops.bias_act_bwd(x.to(DEV), b.to(DEV), g.to(DEV), "gelu")
    torch.cuda.synchronize()
    for name, t in (("out", out), ("pre", pre), ("dx", dx), ("db", db)):
        a, e = t.cpu().view(torch.int16), ref[name].view(torch.int16)
        assert torch.equal(a, e), f"{name}: {(a != e).sum().item()} of {a.numel()} words differ between the packed and the fp32 form"
