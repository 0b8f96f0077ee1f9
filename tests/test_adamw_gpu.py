"""GPU: b200ex.optim.AdamW (b200k_adamw) against torch.optim.AdamW(fused=True) — the optimizer of the reference's
training benchmark — and against the oracle's restatement of ATen's arithmetic."""

import pytest
import torch

from oracle import regions as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SHAPES = [(1,), (7,), (8,), (1000,), (8192,), (8193,), (3, 4096), (257, 129), (1 << 20,), (8192 * 3 + 5,)]


def _close(x, y, dtype, what):
    """Same fp32 expressions as ATen's kernel, but which multiply of `beta * m + (1 - beta) * g` the compiler fuses into
    the add may differ: the last fp32 bit of a moment (measured: v differs by <= 3.5e-7 relative; m, a cancelling sum,
    by that much of its terms).  A 16-bit store shows it on under 1 % of the elements per step as one storage ulp of
    the terms' magnitude."""
    ulp = {torch.bfloat16: 2.0**-7, torch.float16: 2.0**-10, torch.float32: 2.0**-20}[dtype]
    diff = (x.float() - y.float()).abs()
    scale = y.float().abs().max().item() if y.numel() else 0.0
    assert (diff <= 2 * ulp * y.float().abs() + ulp * scale + 1e-30).all(), f"{what}: max abs diff {diff.max().item():.3e} at scale {scale:.3e}"
    if dtype is not torch.float32 and x.numel() >= 1000:
        assert (diff != 0).float().mean().item() <= 0.05, f"{what}: {(diff != 0).sum().item()} of {x.numel()} elements differ"


def _params(dtype, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter((torch.randn(s, generator=g) * 0.5).to(dtype).to(DEV)) for s in SHAPES]


def _set_grads(params, seed):
    g = torch.Generator().manual_seed(1000 + seed)
    for p in params:
        p.grad = (torch.randn(p.shape, generator=g) * 0.1).to(p.dtype).to(DEV)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("wd", [0.0, 0.1])
def test_equals_torch_fused_adamw(dtype, wd):
    from b200ex.optim import AdamW

    ours, theirs = _params(dtype), _params(dtype)
    a = AdamW(ours, lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=wd)
    b = torch.optim.AdamW(theirs, lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=wd, fused=True)
    for step in range(5):
        _set_grads(ours, step), _set_grads(theirs, step)
        a.step(), b.step()
    torch.cuda.synchronize()
    for i, (p, q) in enumerate(zip(ours, theirs)):
        sa, sb = a.state[p], b.state[q]
        for name, x, y in (("param", p.data, q.data), ("exp_avg", sa["exp_avg"], sb["exp_avg"]), ("exp_avg_sq", sa["exp_avg_sq"], sb["exp_avg_sq"])):
            _close(x, y, dtype, f"{name} of tensor {i} {tuple(p.shape)}")
        assert float(sa["step"]) == float(sb["step"]) == 5.0


def test_equals_oracle_and_state_dict_layout():
    from b200ex.optim import AdamW

    params = _params(torch.bfloat16, seed=3)[:6]
    ref = [(p.detach().clone(), torch.zeros_like(p), torch.zeros_like(p)) for p in params]
    opt = AdamW(params, lr=1e-2, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.05)
    for step in range(1, 4):
        _set_grads(params, step)
        opt.step()
        ref = [O.adamw_step(p, q.grad, m, v, step, 1e-2, 0.9, 0.999, 1e-6, 0.05) for (p, m, v), q in zip(ref, params)]
    torch.cuda.synchronize()
    for q, (p, m, v) in zip(params, ref):
        for a, b in ((q.detach().cpu(), p), (opt.state[q]["exp_avg"].cpu(), m), (opt.state[q]["exp_avg_sq"].cpu(), v)):
            _close(a, b, torch.bfloat16, "oracle")
    sd = opt.state_dict()
    assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"}  # torch's keys: checkpoints are interchangeable
    other = torch.optim.AdamW(params, lr=1e-2, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.05, fused=True)
    other.load_state_dict(sd)
    assert float(other.state[params[0]]["step"]) == 3.0


def test_rejects_what_it_cannot_do():
    from b200ex.optim import AdamW

    p = torch.nn.Parameter(torch.randn(8))
    p.grad = torch.randn(8)
    with pytest.raises(RuntimeError, match="CUDA"):
        AdamW([p]).step()
