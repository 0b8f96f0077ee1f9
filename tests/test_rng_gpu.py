"""GPU: counter-based uniform random numbers (b200k_uniform_philox) and dropout through the executor.

The oracle's Philox-4x32-10 is pinned on the CPU against the published known-answer vectors
(tests/test_oracle_golden.py) and here against ATen's own CUDA generator; the kernel must equal the oracle bit for
bit.  The executor tests follow thunder/tests/test_randomness.py: reproducibility, agreement between `uniform` and
`uniform_philox` + state prims, generator offset advancing by 4 per call, and a dropout backward that regenerates
the forward's mask."""

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import regions as O
from thunder_env import fusions, import_thunder

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_oracle_philox_matches_atens_cuda_generator():
    """torch.rand draws element i of a small fp32 tensor from Philox(seed, subsequence = i, counter = offset / 4),
    component 0 (one element per thread while numel <= the launch's thread count): an independent implementation of
    the same generator and of curand's integer -> (0, 1] conversion."""
    gen = torch.Generator(device=DEV)
    n = 1024
    for seed, offset in ((0, 0), (20261106, 8), (2**40 + 12345, 4 * 1000003)):
        gen.manual_seed(seed)
        gen.set_offset(offset)
        got = torch.rand(n, device=DEV, generator=gen).cpu().numpy()
        words = O.philox4x32_10(seed, np.arange(n, dtype=np.uint64), offset // 4)[:, 0]
        u = words.astype(np.float32) * np.float32(2.0**-32) + np.float32(2.0**-33)
        u = np.where(u == np.float32(1.0), np.float32(0.0), u)
        assert np.array_equal(got, u), f"seed {seed} offset {offset}"


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(1,), (7,), (8,), (1025,), (33, 17), (4, 128, 96), (1 << 20,)])
def test_kernel_equals_oracle(dtype, shape):
    from b200ex import ops

    seed, offset = 123456789012345, 4242
    got = ops.uniform_philox(shape, 0.0, 1.0, device=DEV, dtype=dtype, seed=seed, offset=offset)
    want = O.uniform_philox(shape, 0.0, 1.0, seed, offset, dtype)
    assert got.dtype == dtype and tuple(got.shape) == tuple(shape)
    assert torch.equal(got.cpu(), want)
    assert got.min().item() >= 0.0 and got.max().item() < 1.0


def test_kernel_range_and_statistics():
    from b200ex import ops

    u = ops.uniform_philox((4096, 4096), -2.0, 6.0, device=DEV, dtype=torch.float32, seed=7, offset=3)
    assert torch.equal(u.cpu()[:4], O.uniform_philox((4096, 4096), -2.0, 6.0, 7, 3, torch.float32)[:4])
    assert u.min().item() >= -2.0 and u.max().item() < 6.0
    assert abs(u.mean().item() - 2.0) < 5e-3 and abs(u.var().item() - 64.0 / 12.0) < 2e-2
    # different counters and different seeds give unrelated streams
    v = ops.uniform_philox((4096, 4096), -2.0, 6.0, device=DEV, dtype=torch.float32, seed=7, offset=4)
    w = ops.uniform_philox((4096, 4096), -2.0, 6.0, device=DEV, dtype=torch.float32, seed=8, offset=3)
    for other in (v, w):
        c = torch.corrcoef(torch.stack([u.flatten(), other.flatten()]))[0, 1].item()
        assert abs(c) < 1e-3


def test_empty_and_unaligned_output():
    from b200ex import ops

    assert ops.uniform_philox((0, 5), 0.0, 1.0, device=DEV, dtype=torch.float32, seed=1, offset=0).shape == (0, 5)
    L = __import__("b200ex._lib", fromlist=["lib"]).lib()
    buf = torch.zeros(64, dtype=torch.bfloat16, device=DEV)
    view = buf[3:3 + 21]  # 2-byte aligned only: the kernel must fall back to element stores
    rc = L.b200k_uniform_philox(view.data_ptr(), 21, 5, 9, 0.0, 1.0, 1, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(view.cpu(), O.uniform_philox((21,), 0.0, 1.0, 5, 9, torch.bfloat16))
    assert buf[:3].abs().sum().item() == 0 and buf[24:].abs().sum().item() == 0


# ---- through the executor ---------------------------------------------------------------------------------------
thunder, b200 = import_thunder()  # raises when baseline/_ref is missing: the plugin path must not skip silently


def test_dropout_backward_regenerates_the_forward_mask():
    torch.manual_seed(0)
    x = torch.randn(256, 1024, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    p = 0.25
    jf = thunder.jit(lambda x: F.dropout(x, p), executors=[b200.ex])
    torch.cuda.manual_seed(11)
    y = jf(x)
    (g,) = torch.autograd.grad(y, x, torch.ones_like(y))
    fw, bw = thunder.last_traces(jf)[-1], thunder.last_backward_traces(jf)[-1]
    names = [bs.sym.name for bs in fw.bound_symbols]
    assert "get_and_update_rng_state_impl" in names and "uniform" not in names  # stateless form, state prim outside
    # nothing but two integers is saved for the mask: the backward region contains the generator again
    subsyms = lambda tr: [s.sym.name for bs in tr.bound_symbols if bs.sym.is_fusion for s in bs.subsymbols]
    assert "uniform_philox" in subsyms(fw) and "uniform_philox" in subsyms(bw)
    ret = next(bs for bs in fw.bound_symbols if bs.sym.name == "python_return")
    saved_tensors = ret.args[1][0]
    assert not any(tuple(getattr(p, "shape", ())) == (256, 1024) for p in saved_tensors), saved_tensors
    kept = y != 0
    torch.testing.assert_close(y[kept].float(), (x.detach()[kept].float() / (1 - p)).to(torch.bfloat16).float(), rtol=1.6e-2, atol=1e-5)
    # the gradient is 1/(1-p) exactly where the forward kept the element (zeros of x itself aside)
    assert torch.equal((g != 0) | (x.detach() == 0), kept | (x.detach() == 0))
    frac = kept.float().mean().item()
    assert abs(frac - (1 - p)) < 5e-3


def test_reproducible_and_generator_advances_by_four_per_call():
    def f(a):
        b = thunder.torch.uniform_like(a, device=a.device, dtype=a.dtype)
        c = thunder.torch.uniform_like(a, device=a.device, dtype=a.dtype)
        return a * b + c

    a = torch.randn(64, 64, device=DEV)
    gen = torch.cuda.default_generators[0]
    jf = thunder.jit(f, executors=[b200.ex])
    outs = []
    for _ in range(2):
        gen.manual_seed(20)
        first = jf(a)
        second = jf(a)
        assert gen.get_offset() == 4 * 4  # four uniform calls, 4 per call (thunder/tests/test_randomness.py:174)
        outs.append((first, second))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert not torch.equal(outs[0][0], outs[0][1])
    # the same program with the transform disabled draws the same numbers through prims.uniform
    jg = thunder.jit(f, executors=[b200.ex], disable_replace_uniform=True)
    gen.manual_seed(20)
    assert torch.equal(jg(a), outs[0][0]) and torch.equal(jg(a), outs[0][1])
    assert gen.get_offset() == 4 * 4
    assert not any(bs.sym.name == "get_and_update_rng_state_impl" for bs in thunder.last_traces(jg)[-1].bound_symbols)


def test_uniform_philox_symbol_with_explicit_seed():
    import thunder.torch as ltorch

    def f(shape, dtype, device, seed, offset):
        return ltorch.uniform_philox(shape, device=device, dtype=dtype, seed=seed, offset=offset)

    jf = thunder.jit(f, executors=[b200.ex])
    out = jf((10, 30), thunder.float32, "cuda:0", 1234, 77)
    assert torch.equal(out.cpu(), O.uniform_philox((10, 30), 0.0, 1.0, 1234, 77, torch.float32))
    assert torch.equal(out, jf((10, 30), thunder.float32, "cuda:0", 1234, 77))


def test_dropout_is_one_kernel_and_equals_the_materialised_form():
    """The mask never exists in memory: the region's specialised kernel generates the random numbers in registers.
    The result must equal, bit for bit, the same region run with the random tensor materialised by
    b200k_uniform_philox first (the path launches that cannot be specialised take)."""
    from b200ex.pointwise import PwLaunch

    x = torch.randn(512, 768, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    cot = torch.randn(512, 768, device=DEV, dtype=torch.bfloat16)
    f = lambda x: F.dropout(F.gelu(x), 0.1) + x
    res = {}
    for fused in (True, False):
        PwLaunch.use_jit = True
        jf = thunder.jit(f, executors=[b200.ex])
        torch.cuda.manual_seed(5)
        if not fused:
            orig = PwLaunch.run
            PwLaunch.run = lambda self, env, _in_kernel_rng=True: orig(self, env, False)
        try:
            y = jf(x)
            (g,) = torch.autograd.grad(y, x, cot)
        finally:
            if not fused:
                PwLaunch.run = orig
        res[fused] = (y.detach().clone(), g.clone())
        if fused:
            fw, bw = thunder.last_traces(jf)[-1], thunder.last_backward_traces(jf)[-1]
            # forward: one launch (gelu, mask, scale, residual add together); the generator is not a launch of its own
            assert [ks for _, ks in fusions(fw)] == [["b200k_pointwise"]], fusions(fw)
            assert all(k != "uniform_philox" for _, ks in fusions(bw) for k in ks)
    assert torch.equal(res[True][0], res[False][0]) and torch.equal(res[True][1], res[False][1])
    dropped = (res[True][0] == x.detach()).float().mean().item()  # y == x exactly where the gelu branch was dropped
    assert abs(dropped - 0.1) < 0.01


def test_dropout_with_unaligned_width_falls_back_to_the_materialised_mask():
    x = torch.randn(33, 77, device=DEV, dtype=torch.bfloat16, requires_grad=True)  # 77 is not a multiple of 8
    jf = thunder.jit(lambda x: F.dropout(x, 0.5), executors=[b200.ex])
    torch.cuda.manual_seed(9)
    y1 = jf(x)
    torch.cuda.manual_seed(9)
    y2 = jf(x)
    assert torch.equal(y1, y2)
    u = O.uniform_philox((33, 77), 0.0, 1.0, 9, 0, torch.bfloat16)
    # thunder's dropout keeps an element where u < 1 - p (thunder/torch/__init__.py dropout)
    assert torch.equal((y1 != 0).cpu() | (x.detach().cpu() == 0), (u < 0.5) | (x.detach().cpu() == 0))
