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


def test_load_state_dict_after_a_step_switches_to_the_loaded_moments():
    """step, load_state_dict, step: the second step must read and write the LOADED moment tensors (load_state_dict
    replaces them), exactly as torch.optim.AdamW(fused=True) does."""
    from b200ex.optim import AdamW

    dtype = torch.bfloat16
    ours, theirs, donor = _params(dtype), _params(dtype), _params(dtype, seed=5)
    kw = dict(lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    a, b, d = AdamW(ours, **kw), torch.optim.AdamW(theirs, fused=True, **kw), torch.optim.AdamW(donor, fused=True, **kw)
    for step in range(3):  # a checkpoint from another run
        _set_grads(donor, 50 + step)
        d.step()
    _set_grads(ours, 0), _set_grads(theirs, 0)
    a.step(), b.step()  # warm-up step before the resume
    import copy

    sd = d.state_dict()
    # load_state_dict keeps tensors that already have the right device / dtype: each optimizer gets its own copy
    a.load_state_dict(copy.deepcopy(sd)), b.load_state_dict(copy.deepcopy(sd))
    for step in range(1, 3):
        _set_grads(ours, step), _set_grads(theirs, step)
        a.step(), b.step()
    torch.cuda.synchronize()
    for i, (p, q) in enumerate(zip(ours, theirs)):
        sa, sb = a.state[p], b.state[q]
        assert float(sa["step"]) == float(sb["step"]) == 5.0
        for name, x, y in (("param", p.data, q.data), ("exp_avg", sa["exp_avg"], sb["exp_avg"]), ("exp_avg_sq", sa["exp_avg_sq"], sb["exp_avg_sq"])):
            _close(x, y, dtype, f"{name} of tensor {i} after resume")
    # and what is saved afterwards is what was updated
    saved = a.state_dict()["state"]
    assert all(torch.equal(saved[i]["exp_avg"], a.state[p]["exp_avg"]) for i, p in enumerate(ours))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_offset_views_are_updated_element_by_element(dtype):
    """Contiguous parameters that do not start on a 16-byte address (a slice of a flat buffer, p[1:]) take the scalar
    path of the kernel instead of faulting on a misaligned vector access."""
    from b200ex.optim import AdamW

    g = torch.Generator().manual_seed(11)
    flat_a = (torch.randn(3 * 8192 + 64, generator=g) * 0.5).to(dtype).to(DEV)
    flat_b = flat_a.clone()
    cuts = [(1, 1000), (1001, 9194), (9195, 9195 + 8192 * 2 + 3)]
    ours = [torch.nn.Parameter(flat_a[s:e]) for s, e in cuts]
    theirs = [torch.nn.Parameter(flat_b[s:e].clone()) for s, e in cuts]
    assert any(p.data_ptr() % 16 for p in ours)
    kw = dict(lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    a, b = AdamW(ours, **kw), torch.optim.AdamW(theirs, fused=True, **kw)
    for step in range(3):
        gg = torch.Generator().manual_seed(70 + step)
        gbuf = (torch.randn(flat_a.shape, generator=gg) * 0.1).to(dtype).to(DEV)
        for (s, e), p, q in zip(cuts, ours, theirs):
            p.grad, q.grad = gbuf[s:e], gbuf[s:e].clone()
        a.step(), b.step()
    torch.cuda.synchronize()
    for i, (p, q) in enumerate(zip(ours, theirs)):
        _close(p.data, q.data, dtype, f"offset param {i}")
        _close(a.state[p]["exp_avg_sq"], b.state[q]["exp_avg_sq"], dtype, f"offset exp_avg_sq {i}")
    # bytes outside the parameter views are untouched
    assert torch.equal(flat_a[:1], flat_b[:1]) and torch.equal(flat_a[cuts[-1][1]:], flat_b[cuts[-1][1]:])
