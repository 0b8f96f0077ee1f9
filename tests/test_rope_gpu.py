"""GPU parity: QKV split + RoPE (+GQA expand) forward/backward through the C-ABI vs the oracle and golden vectors."""

import pytest
import torch

from oracle import regions as O
from parity import assert_parity

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _check(dtype, B, T, n_head, groups, hs, rope_n, seed=0):
    from b200ex import ops

    g = torch.Generator().manual_seed(seed)
    mk = lambda *s: torch.randn(*s, generator=g).to(dtype)
    qkv = mk(B, T, (n_head + 2 * groups) * hs)
    cos, sin = mk(T, rope_n), mk(T, rope_n)
    hkv = n_head if (groups != n_head and groups != 1) else groups
    dq, dk, dv = mk(B, n_head, T, hs), mk(B, hkv, T, hs), mk(B, hkv, T, hs)
    args = (n_head, groups, hs, rope_n)
    q, k, v = ops.qkv_rope_fwd(qkv.to(DEV), cos.to(DEV), sin.to(DEV), *args)
    dqkv = ops.qkv_rope_bwd(dq.to(DEV), dk.to(DEV), dv.to(DEV), cos.to(DEV), sin.to(DEV), *args)
    torch.cuda.synchronize()
    eq, ek, ev = O.qkv_split_rope_fwd(qkv, cos, sin, *args)
    what = f"{dtype} B{B} T{T} nh{n_head} g{groups} hs{hs} rope{rope_n}"
    assert_parity(q, eq, what=what + " q"), assert_parity(k, ek, what=what + " k"), assert_parity(v, ev.contiguous(), what=what + " v")
    edqkv = O.qkv_split_rope_bwd(dq, dk, dv, cos, sin, *args)
    # k/v gradients are fp32 sums of q_per_kv storage-rounded terms; cancellation is bounded by the terms.
    assert_parity(dqkv, edqkv, what=what + " dqkv", atol_extra=0.0 if dtype == torch.float32 else 2.0**-6)


def test_golden_vectors(golden):
    from b200ex import ops

    for name, c in golden.items():
        if not name.startswith("rope_"):
            continue
        i, o, m = c["inputs"], c["outputs"], c["meta"]
        args = (m["n_head"], m["n_query_groups"], m["head_size"], m["rope_n_elem"])
        d = lambda t: t.to(DEV)
        q, k, v = ops.qkv_rope_fwd(d(i["qkv"]), d(i["cos"]), d(i["sin"]), *args)
        dqkv = ops.qkv_rope_bwd(d(i["dq"]), d(i["dk"]), d(i["dv"]), d(i["cos"]), d(i["sin"]), *args)
        # pure products and sums of two rounded products: the CUDA result is bit-identical to the reference
        for got, key in ((q, "q"), (k, "k"), (v, "v"), (dqkv, "dqkv")):
            assert torch.equal(got.cpu(), o[key]), f"{name} {key}: not bit-exact"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize(
    "B,T,n_head,groups,hs,rope_n",
    [(2, 16, 8, 2, 16, 16), (1, 12, 4, 2, 32, 16), (2, 8, 4, 4, 16, 16), (1, 8, 4, 1, 16, 16), (1, 5, 6, 3, 24, 12), (2, 7, 32, 8, 128, 128), (1, 3, 8, 2, 64, 32), (1, 2, 2, 2, 6, 6)],
)
def test_shapes(dtype, B, T, n_head, groups, hs, rope_n):
    _check(dtype, B, T, n_head, groups, hs, rope_n, seed=T + hs)


def test_llama3_8b_shape():
    """Llama-3-8B attention shape: B=2, T=4096, 32 heads / 8 KV groups x 128, full rotary."""
    _check(torch.bfloat16, 2, 4096, 32, 8, 128, 128)


def test_norm_preservation_property():
    """With cos/sin from real angles, RoPE is a rotation: per-position |q| is preserved (fp32)."""
    from b200ex import ops

    B, T, nh, G, hs = 1, 64, 8, 2, 64
    torch.manual_seed(0)
    qkv = torch.randn(B, T, (nh + 2 * G) * hs, device=DEV)
    ang = torch.rand(T, hs // 2, device=DEV) * 6.28
    cos, sin = torch.cat([ang.cos(), ang.cos()], -1), torch.cat([ang.sin(), ang.sin()], -1)
    q, k, v = ops.qkv_rope_fwd(qkv, cos, sin, nh, G, hs, hs)
    q_in = qkv.view(B, T, G, nh // G + 2, hs)[:, :, :, : nh // G].permute(0, 2, 3, 1, 4).reshape(B, nh, T, hs)
    torch.testing.assert_close(q.norm(dim=-1), q_in.norm(dim=-1), rtol=1e-5, atol=1e-5)


_FP32_FORM = r"""
import os, sys, torch
os.environ["B200K_ROPE_PACKED"] = "0"
sys.path[:0] = [sys.argv[1]]
from b200ex import ops
d = torch.load(sys.argv[2])
c = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
q, k, v = ops.qkv_rope_fwd(c["qkv"], c["cos"], c["sin"], *c["args"])
dqkv = ops.qkv_rope_bwd(c["dq"], c["dk"], c["dv"], c["cos"], c["sin"], *c["args"])
torch.save({"q": q.cpu(), "k": k.cpu(), "v": v.cpu(), "dqkv": dqkv.cpu()}, sys.argv[3])
"""


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_packed_form_equals_the_fp32_form_bit_for_bit(tmp_path, dtype):
    """16-bit storage runs the rotation on packed 2 x 16-bit multiplies / adds / subtracts (one rounding of the exact
    result) instead of fp32-op-then-round; the two are claimed to give the same bits.  The fp32 form runs in a
    second process (B200K_ROPE_PACKED=0 is read once per process); inputs span many exponents, including products
    that are subnormal or overflow in fp16, and cos/sin beyond [-1, 1]."""
    import os
    import subprocess
    import sys

    from b200ex import ops

    g = torch.Generator().manual_seed(17)
    B, T, n_head, groups, hs, rope_n = 2, 96, 8, 2, 64, 48
    wide = lambda *s: (torch.randn(*s, generator=g) * torch.exp2(torch.randint(-14, 12, s, generator=g).float())).to(dtype)
    d = {"qkv": wide(B, T, (n_head + 2 * groups) * hs), "cos": wide(T, rope_n), "sin": wide(T, rope_n),
         "dq": wide(B, n_head, T, hs), "dk": wide(B, n_head, T, hs), "dv": wide(B, n_head, T, hs), "args": (n_head, groups, hs, rope_n)}
    src, dst = tmp_path / "in.pt", tmp_path / "out.pt"
    torch.save(d, src)
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lightning-thunder_b200")
    subprocess.run([sys.executable, "-c", _FP32_FORM, pkg, str(src), str(dst)], check=True, timeout=300)
    ref = torch.load(dst)
    c = {k: (v.to(DEV) if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    q, k, v = ops.qkv_rope_fwd(c["qkv"], c["cos"], c["sin"], *c["args"])
    dqkv = ops.qkv_rope_bwd(c["dq"], c["dk"], c["dv"], c["cos"], c["sin"], *c["args"])
    torch.cuda.synchronize()
    for name, t in (("q", q), ("k", k), ("v", v), ("dqkv", dqkv)):
        a, e = t.cpu().contiguous().view(torch.int16), ref[name].contiguous().view(torch.int16)
        # NaN payloads may differ between the two forms (inf - inf); compare them as "both NaN"
        both_nan = torch.isnan(t.cpu().float()) & torch.isnan(ref[name].float())
        assert torch.equal(a[~both_nan], e[~both_nan]), f"{name}: {(a != e).sum().item()} of {a.numel()} words differ"
