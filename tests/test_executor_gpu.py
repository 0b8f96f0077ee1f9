"""GPU: ``thunder.jit(fn, executors=[b200])`` against ``thunder.jit(fn, executors=["torch"])`` — the reference's
torch-executor output on the same inputs is the parity target BASELINE.json names — forward and backward,
for every fused-region family and for a small Llama-style model."""

import pytest
import torch
import torch.nn.functional as F

from parity import TOL, assert_parity, assert_parity_or_closer, assert_tensor_scale_parity
from thunder_env import fusions, import_thunder

pytestmark = pytest.mark.gpu

thunder, b200 = import_thunder()  # raises when baseline/_ref is missing: the plugin path must not skip silently

DEV = "cuda:0"
BF, F32 = torch.bfloat16, torch.float32


def _both(fn, args, cot_seed=0, opts=None):
    """Run fn under b200 and under the torch executor on clones of args; return (outs, grads) for both."""
    res = []
    for execs in ([b200.ex], ["torch"]):
        a = [t.detach().clone().requires_grad_(t.is_floating_point() and t.requires_grad) if isinstance(t, torch.Tensor) else t for t in args]
        jf = thunder.jit(fn, executors=execs, **(opts or {}))
        out = jf(*a)
        outs = list(out) if isinstance(out, (tuple, list)) else [out]
        g = torch.Generator(device=DEV).manual_seed(cot_seed)
        cots = [torch.randn(o.shape, generator=g, device=DEV, dtype=torch.float32).to(o.dtype) for o in outs]
        diff = [t for t in a if isinstance(t, torch.Tensor) and t.requires_grad]
        grads = torch.autograd.grad(outs, diff, grad_outputs=cots, allow_unused=True) if diff else []
        res.append((outs, list(grads), jf))
    # fp64 eager run of the same function: the arbiter of the reference's assert_closer protocol
    a = [t.detach().clone().double().requires_grad_(t.requires_grad) if isinstance(t, torch.Tensor) and t.is_floating_point() else t for t in args]
    out = fn(*a)
    outs = list(out) if isinstance(out, (tuple, list)) else [out]
    g = torch.Generator(device=DEV).manual_seed(cot_seed)
    cots = [torch.randn(o.shape, generator=g, device=DEV, dtype=torch.float32).to(r.dtype).double() for o, r in zip(outs, res[0][0])]
    diff = [t for t in a if isinstance(t, torch.Tensor) and t.requires_grad]
    grads = torch.autograd.grad(outs, diff, grad_outputs=cots, allow_unused=True) if diff else []
    res.append((outs, list(grads), None))
    torch.cuda.synchronize()
    return res


def _mk(*shape, dtype=BF, grad=True, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed + sum(shape))
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(DEV).requires_grad_(grad)


def _compare(res, what):
    """b200 against the torch executor at north_star's tolerances (bf16 rtol 1.6e-2 / atol 1e-5, fp32 rtol 1e-5 / atol
    1e-6); an element outside them must be at least as close to the fp64 run as the torch executor's."""
    (o1, g1, _), (o2, g2, _), (o3, g3, _) = res
    for i, (a, b, r) in enumerate(zip(o1, o2, o3)):
        assert_parity_or_closer(a, b, r, what=f"{what} out[{i}]")
    for i, (a, b, r) in enumerate(zip(g1, g2, g3)):
        if a is None and b is None:
            continue
        assert_parity_or_closer(a, b, r, what=f"{what} grad[{i}]")


@pytest.mark.parametrize("dtype", [BF, F32])
def test_rmsnorm_residual(dtype):
    def f(x, r, w):
        h = x + r
        return h, F.rms_norm(h, (512,), w, 1e-5)

    res = _both(f, [_mk(64, 512, dtype=dtype, seed=1), _mk(64, 512, dtype=dtype, seed=2), _mk(512, dtype=dtype, seed=3)])
    kinds = [k for _, ks in fusions(thunder.last_traces(res[0][2])[-1]) for k in ks]
    assert "b200_rms_norm_fwd+residual" in kinds
    _compare(res, f"rmsnorm {dtype}")


@pytest.mark.parametrize("dtype", [BF, F32])
@pytest.mark.parametrize("bias", [True, False])
def test_layer_norm(dtype, bias):
    def f(x, r, w, b):
        h = x + r
        return F.layer_norm(h, (768,), w, b if bias else None, 1e-5) * 2.0

    args = [_mk(48, 768, dtype=dtype, seed=1), _mk(48, 768, dtype=dtype, seed=2), _mk(768, dtype=dtype, seed=3), _mk(768, dtype=dtype, seed=4, grad=bias)]
    res = _both(f, args)
    jf = res[0][2]
    assert "b200_layer_norm_fwd" in [k for _, ks in fusions(thunder.last_traces(jf)[-1]) for k in ks]
    assert "b200_layer_norm_bwd" in [k for _, ks in fusions(thunder.last_backward_traces(jf)[-1]) for k in ks]
    _compare(res, f"layer_norm {dtype}")


@pytest.mark.parametrize("dtype", [BF, F32])
def test_softmax_with_scale_and_mask(dtype):
    def f(q, k):
        att = (q @ k.transpose(-2, -1)) * 0.125
        att = att.masked_fill(torch.ones(96, 96, device=q.device, dtype=torch.bool).tril().logical_not(), float("-inf"))
        return F.softmax(att, dim=-1)

    res = _both(f, [_mk(2, 4, 96, 64, dtype=dtype, seed=31), _mk(2, 4, 96, 64, dtype=dtype, seed=32)])
    jf = res[0][2]
    assert "b200_softmax_fwd" in [k for _, ks in fusions(thunder.last_traces(jf)[-1]) for k in ks]
    assert "b200_softmax_bwd" in [k for _, ks in fusions(thunder.last_backward_traces(jf)[-1]) for k in ks]
    _compare(res, f"softmax {dtype}")


def test_swiglu_inline_expression():
    res = _both(lambda a, b: F.silu(a) * b, [_mk(128, 1024, seed=4, scale=2.0), _mk(128, 1024, seed=5)])
    kinds = [k for _, ks in fusions(thunder.last_traces(res[0][2])[-1]) for k in ks]
    assert kinds, "swiglu was not claimed"
    _compare(res, "swiglu")


def test_swiglu_function_is_native():
    res = _both(lambda a, b: b200.swiglu(a, b), [_mk(64, 2048, seed=6, scale=2.0), _mk(64, 2048, seed=7)])
    fw = fusions(thunder.last_traces(res[0][2])[-1])
    assert [k for _, ks in fw for k in ks] == ["b200_swiglu"]
    _compare(res, "b200.swiglu")


@pytest.mark.parametrize("dtype", [BF, F32])
def test_cross_entropy(dtype):
    logits = _mk(256, 32064, dtype=dtype, seed=8, scale=2.0)
    tgt = torch.randint(0, 32064, (256,), device=DEV)
    tgt[5] = -100

    def f(lg, t):
        return F.cross_entropy(lg, t, ignore_index=-100)

    res = _both(f, [logits, tgt])
    assert "b200_cross_entropy_fwd" in [k for _, ks in fusions(thunder.last_traces(res[0][2])[-1]) for k in ks]
    _compare(res, f"cross_entropy {dtype}")


def test_qkv_split_rope_function():
    qkv = _mk(2, 64, (8 + 4) * 32, seed=9)
    cos, sin = _mk(64, 32, grad=False, seed=10), _mk(64, 32, grad=False, seed=11)
    res = _both(lambda q, c, s: b200.qkv_split_rope(q, c, s, 8, 2, 32, 32), [qkv, cos, sin])
    assert [k for _, ks in fusions(thunder.last_traces(res[0][2])[-1]) for k in ks] == ["b200_qkv_split_rope"]
    _compare(res, "qkv_split_rope")


@pytest.mark.parametrize("act", ["gelu", "relu"])
def test_bias_activation(act):
    fn = (lambda x, b: F.gelu(x + b)) if act == "gelu" else (lambda x, b: F.relu(x + b))
    res = _both(fn, [_mk(256, 768, seed=12, scale=1.5), _mk(768, seed=13)])
    jf = res[0][2]
    # gelu saves the pre-activation and runs on the family kernels; relu's autograd saves the bool mask, so its regions
    # keep the trace's dtypes and run as generic programs (see _native_for)
    fwd_kind, bwd_kind = (f"b200k_bias_act_fwd[{act}]", f"b200k_bias_act_bwd[{act}]") if act == "gelu" else ("b200k_pointwise", "b200k_pointwise")
    assert [k for _, ks in fusions(thunder.last_traces(jf)[-1]) for k in ks] == [fwd_kind]
    assert [k for _, ks in fusions(thunder.last_backward_traces(jf)[-1]) for k in ks] == [bwd_kind, "b200k_colsum"]
    _compare(res, f"bias_{act}")


def test_tanh_chain_goes_through_the_generic_interpreter_and_colsum():
    res = _both(lambda x, b: torch.tanh(x * b) + 0.5 * x, [_mk(256, 768, seed=21), _mk(768, seed=22)])
    kinds = [k for _, ks in fusions(thunder.last_backward_traces(res[0][2])[-1]) for k in ks]
    assert "b200k_pointwise" in kinds and "b200k_colsum" in kinds
    _compare(res, "tanh chain")


def test_generic_pointwise_fp32_with_broadcasts():
    def f(x, y, s):
        z = torch.tanh(x) * y + torch.exp(-x * x) / (1.0 + y * y)
        return torch.where(z > 0.1, z, 0.01 * z) * s, torch.sqrt(torch.abs(x)) + s

    res = _both(f, [_mk(8, 33, 65, dtype=F32, seed=14), _mk(8, 33, 65, dtype=F32, seed=15), _mk(65, dtype=F32, seed=16)])
    _compare(res, "generic pointwise")


def test_region_callable_is_reused_and_handles_new_strides():
    jf = thunder.jit(lambda a, b: F.silu(a) * b, executors=[b200.ex])
    a, b = _mk(64, 256, grad=False, seed=17), _mk(64, 256, grad=False, seed=18)
    ref = F.silu(a) * b
    out1 = jf(a, b)
    at = a.t().contiguous().t()  # same shape, column-major strides
    out2 = jf(at, b)
    assert_parity(out1, ref, what="contiguous", ulp_scale=ref.float().abs())
    assert_parity(out2, ref, what="transposed strides", ulp_scale=ref.float().abs())
    assert thunder.cache_misses(jf) == 1


@pytest.mark.parametrize("fusion_type", ["consecutive", "dataflow"])
def test_tiny_llama_forward_backward(fusion_type):
    """A 2-layer Llama-style GPT (RMSNorm, fused-QKV RoPE with GQA, SwiGLU, cross-entropy inside the jitted callable):
    loss and every parameter gradient against the torch executor."""
    from baseline.llama_model import GPT, build_rope_cache, tiny_config
    from thunder.executors.sdpaex import sdpa_ex

    cfg = tiny_config(spelling="ops", n_embd=128, intermediate_size=256, n_head=4, n_query_groups=2, padded_vocab_size=512)
    torch.manual_seed(0)
    model = GPT(cfg).to(DEV).to(BF)
    B, T = 2, 64
    cos, sin = build_rope_cache(T, cfg.rope_n_elem, device=DEV, dtype=BF)
    idx = torch.randint(0, cfg.padded_vocab_size, (B, T), device=DEV)
    tgt = torch.randint(0, cfg.padded_vocab_size, (B, T), device=DEV)
    tgt[0, :5] = -1
    results = {}
    for name, execs in (("b200", [sdpa_ex, b200.ex]), ("torch", [sdpa_ex, "torch"])):
        model.zero_grad(set_to_none=True)
        jm = thunder.jit(model, executors=execs, fusion_type=fusion_type)
        loss = jm(idx, cos, sin, tgt)
        loss.backward()
        results[name] = (loss.detach().clone(), {n: p.grad.detach().clone() for n, p in model.named_parameters()}, jm)
    torch.cuda.synchronize()
    jm = results["b200"][2]
    fk = [k for _, ks in fusions(thunder.last_traces(jm)[-1]) for k in ks]
    bk = [k for _, ks in fusions(thunder.last_backward_traces(jm)[-1]) for k in ks]
    assert fk.count("b200_qkv_split_rope") == cfg.n_layer and "b200_cross_entropy_fwd" in fk and "b200_cross_entropy_bwd" in bk
    assert fk.count("b200_rms_norm_fwd+residual") == 2 * cfg.n_layer
    torch.testing.assert_close(results["b200"][0].float(), results["torch"][0].float(), rtol=1.6e-2, atol=1e-5)
    _grads_close(results)


def test_llama3_8b_block_full_size():
    """BASELINE.json configs[2]: one LitGPT-style Llama-3-8B transformer block (n_embd 4096, 32 heads / 8 KV groups x 128,
    intermediate 14336), bf16, batch 2 x seq 4096, forward + backward, b200 regions against the torch executor."""
    from baseline.llama_model import LLAMA3_8B, Block, build_rope_cache
    from thunder.executors.sdpaex import sdpa_ex

    import dataclasses

    cfg = dataclasses.replace(LLAMA3_8B, spelling="ops")  # the op-level route; the litgpt spelling is tests/test_regionc_gpu.py
    torch.manual_seed(0)
    block = Block(cfg).to(DEV).to(BF)
    B, T = 2, 4096
    cos, sin = build_rope_cache(T, cfg.rope_n_elem, base=cfg.rope_base, device=DEV, dtype=BF)
    x0 = (torch.randn(B, T, cfg.n_embd, device=DEV) * 0.5).to(BF)
    cot = torch.randn(B, T, cfg.n_embd, device=DEV).to(BF)
    res = {}
    for name, execs in (("b200", [sdpa_ex, b200.ex]), ("torch", [sdpa_ex, "torch"])):
        block.zero_grad(set_to_none=True)
        x = x0.clone().requires_grad_(True)
        jb = thunder.jit(block, executors=execs, fusion_type="dataflow")
        y = jb(x, cos, sin)
        y.backward(cot)
        res[name] = (y.detach(), x.grad.detach(), {n: p.grad.detach().clone() for n, p in block.named_parameters()}, jb)
    torch.cuda.synchronize()
    jb = res["b200"][3]
    fk = [k for _, ks in fusions(thunder.last_traces(jb)[-1]) for k in ks]
    bk = [k for _, ks in fusions(thunder.last_backward_traces(jb)[-1]) for k in ks]
    # region census of SURVEY Appendix C, all on hand-written kernels
    assert sorted(fk) == sorted(["b200_rms_norm_fwd", "b200_qkv_split_rope", "b200_rms_norm_fwd+residual", "b200k_swiglu_fwd", "b200k_pointwise"]), fk
    assert "b200k_swiglu_bwd" in bk and "b200_qkv_rope_bwd" in bk and bk.count("b200_rms_norm_bwd") + bk.count("b200_rms_norm_bwd+residual") == 2

    def rel(a, b):
        return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()

    # outputs of GEMM/attention chains over bf16 activations that differ by an ulp here and there: aggregate error
    assert rel(res["b200"][0], res["torch"][0]) < 1e-2
    assert rel(res["b200"][1], res["torch"][1]) < 1.6e-2
    assert_tensor_scale_parity(res["b200"][0], res["torch"][0], what="block output")
    assert_tensor_scale_parity(res["b200"][1], res["torch"][1], what="block input gradient")
    for n, g in res["b200"][2].items():
        assert rel(g, res["torch"][2][n]) < 1.6e-2, f"grad {n}"
        assert_tensor_scale_parity(g, res["torch"][2][n], what=f"grad {n}")


class _Gpt2Block(torch.nn.Module):
    def __init__(self, d, heads):
        super().__init__()
        self.ln1, self.ln2 = torch.nn.LayerNorm(d), torch.nn.LayerNorm(d)
        self.attn, self.proj = torch.nn.Linear(d, 3 * d), torch.nn.Linear(d, d)
        self.fc, self.out = torch.nn.Linear(d, 4 * d), torch.nn.Linear(4 * d, d)
        self.heads = heads

    def forward(self, x):
        B, T, C = x.shape
        q, k, v = self.attn(self.ln1(x)).split(C, dim=2)
        q, k, v = (t.view(B, T, self.heads, C // self.heads).transpose(1, 2) for t in (q, k, v))
        att = (q @ k.transpose(-2, -1)) * (1.0 / (C // self.heads) ** 0.5)
        att = att.masked_fill(torch.ones(T, T, device=x.device, dtype=torch.bool).tril().logical_not(), float("-inf"))
        y = (F.softmax(att, dim=-1) @ v).transpose(1, 2).contiguous().view(B, T, C)
        x = x + self.proj(y)
        return x + self.out(F.gelu(self.fc(self.ln2(x))))


class _Gpt2(torch.nn.Module):
    def __init__(self, vocab=256, block=64, d=128, heads=4, layers=2):
        super().__init__()
        self.wte, self.wpe = torch.nn.Embedding(vocab, d), torch.nn.Embedding(block, d)
        self.h = torch.nn.ModuleList(_Gpt2Block(d, heads) for _ in range(layers))
        self.ln_f = torch.nn.LayerNorm(d)
        self.head = torch.nn.Linear(d, vocab, bias=False)

    def forward(self, idx, targets):
        pos = torch.arange(0, idx.shape[1], device=idx.device)
        x = self.wte(idx) + self.wpe(pos)
        for b in self.h:
            x = b(x)
        logits = self.head(self.ln_f(x))
        return F.cross_entropy(logits.view(-1, logits.size(-1)), targets.view(-1), ignore_index=-1)


def test_tiny_gpt2_forward_backward():
    """A second model family (nanoGPT-style: LayerNorm, biased linears, GELU, learned positions, explicit masked softmax
    attention): LayerNorm, bias + GELU and cross-entropy run on our kernels, everything b200 does not claim stays with the
    torch executor, and loss and parameter gradients agree with the all-torch run."""
    torch.manual_seed(0)
    model = _Gpt2().to(DEV).to(BF)
    idx = torch.randint(0, 256, (2, 64), device=DEV)
    tgt = torch.randint(0, 256, (2, 64), device=DEV)
    tgt[1, :7] = -1
    results = {}
    for name, execs in (("b200", [b200.ex]), ("torch", ["torch"])):
        model.zero_grad(set_to_none=True)
        jm = thunder.jit(model, executors=execs)
        loss = jm(idx, tgt)
        loss.backward()
        results[name] = (loss.detach().clone(), {n: p.grad.detach().clone() for n, p in model.named_parameters()}, jm)
    torch.cuda.synchronize()
    jm = results["b200"][2]
    fk = [k for _, ks in fusions(thunder.last_traces(jm)[-1]) for k in ks]
    bk = [k for _, ks in fusions(thunder.last_backward_traces(jm)[-1]) for k in ks]
    assert "b200_cross_entropy_fwd" in fk and "b200_cross_entropy_bwd" in bk
    assert fk.count("b200_layer_norm_fwd") == 5 and bk.count("b200_layer_norm_bwd") == 5
    assert fk.count("b200_softmax_fwd") == 2 and bk.count("b200_softmax_bwd") == 2
    torch.testing.assert_close(results["b200"][0].float(), results["torch"][0].float(), rtol=1.6e-2, atol=1e-5)
    _grads_close(results)


def test_readme_mlp_fp32_on_the_gpu():
    """BASELINE.json configs[0] moved to the GPU: the bias + ReLU between the two linears and its backward are b200
    regions; loss and gradients against the torch executor at the fp32 tolerance."""
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(2048, 4096), torch.nn.ReLU(), torch.nn.Linear(4096, 64)).to(DEV)
    x = torch.randn(64, 2048, device=DEV)
    outs = []
    for execs in ([b200.ex], ["torch"]):
        model.zero_grad(set_to_none=True)
        jm = thunder.jit(model, executors=execs)
        y = jm(x)
        y.square().mean().backward()
        outs.append((y.detach().clone(), [p.grad.clone() for p in model.parameters()], jm))
    assert fusions(thunder.last_traces(outs[0][2])[-1]) and fusions(thunder.last_backward_traces(outs[0][2])[-1])
    torch.testing.assert_close(outs[0][0], outs[1][0], rtol=1e-5, atol=1e-6)
    for a, b in zip(outs[0][1], outs[1][1]):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)


def _grads_close(results, tol=1.6e-2, zero_in_exact_arithmetic=()):
    """Parameter gradients: elementwise at rtol 1.6e-2 plus one bf16 ulp of the gradient tensor's largest magnitude (one
    differently rounded term of the sums they are), and norm-wise within rtol."""
    for n, g in results["b200"][1].items():
        ref = results["torch"][1][n]
        if n in zero_in_exact_arithmetic:
            # both runs hold nothing but rounding noise here (e.g. a key-projection bias under softmax): same size of
            # noise, no element-wise claim
            assert g.float().abs().max().item() <= 4.0 * ref.float().abs().max().item() + 1e-12, f"grad {n}"
            continue
        # norm-wise, with one bf16 ulp of the tensor's scale on every element as the floor (a gradient that is zero
        # in exact arithmetic — a key-projection bias under softmax — is rounding noise in both runs)
        floor = 2.0**-8 * ref.float().abs().max() * ref.numel() ** 0.5
        err = (g.float() - ref.float()).norm() / (ref.float().norm() + floor + 1e-12)
        assert err.item() < tol, f"grad {n}: relative error {err.item():.3e}"
        assert_tensor_scale_parity(g, ref, what=f"grad {n}")


def test_reference_llama2_model():
    """thunder/tests/llama2_model.py (the llama2.c-style Transformer of the reference's own test-suite: RMSNorm spelled
    with pow/mean/rsqrt, rotary embedding on real pairs, SwiGLU, masked softmax attention): loss and parameter
    gradients with b200 against the torch executor."""
    from thunder.tests import llama2_model

    args = llama2_model.ModelArgs(dim=128, n_layers=2, n_heads=4, n_kv_heads=2, vocab_size=256, multiple_of=32, max_seq_len=64, dropout=0.0)
    torch.manual_seed(0)
    model = llama2_model.Transformer(args).to(DEV).to(BF)
    ids = torch.randint(0, 256, (2, 64), device=DEV)
    tgt = torch.randint(0, 256, (2, 64), device=DEV)
    results = {}
    for name, execs in (("b200", [b200.ex]), ("torch", ["torch"])):
        model.zero_grad(set_to_none=True)
        jm = thunder.jit(model, executors=execs)
        logits = jm(ids, tgt)
        loss = F.cross_entropy(logits.float().view(-1, logits.size(-1)), tgt.view(-1))  # outside the jitted module in this model
        loss.backward()
        results[name] = (loss.detach().clone(), {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}, jm)
    torch.cuda.synchronize()
    fk = [k for _, ks in fusions(thunder.last_traces(results["b200"][2])[-1]) for k in ks]
    assert fk.count("b200k_swiglu_fwd") == args.n_layers and len(fk) > 10
    torch.testing.assert_close(results["b200"][0].float(), results["torch"][0].float(), rtol=1.6e-2, atol=1e-5)
    _grads_close(results)


def test_reference_bart_self_attention():
    """thunder/tests/hf_bart_self_attn.py (the reference's copy of HF BART's attention block: scaled q, bmm, softmax,
    dropout 0, bmm, biased projections)."""
    from thunder.tests import hf_bart_self_attn

    torch.manual_seed(0)
    model = hf_bart_self_attn.BartAttention(256, 8, dropout=0.0).to(DEV).to(BF)
    x = torch.randn(4, 96, 256, device=DEV, dtype=BF)
    cot = torch.randn(4, 96, 256, device=DEV, dtype=BF)
    mask = torch.zeros(4, 1, 96, 96, device=DEV, dtype=BF)
    mask[:, :, :, 80:] = -10000.0  # padded keys
    results = {}
    for name, execs in (("b200", [b200.ex]), ("torch", ["torch"])):
        model.zero_grad(set_to_none=True)
        xx = x.clone().requires_grad_(True)
        jm = thunder.jit(model, executors=execs)
        out = jm(xx, mask)
        y = out[0] if isinstance(out, (tuple, list)) else out
        y.backward(cot)
        grads = {n: p.grad.detach().clone() for n, p in model.named_parameters()}
        grads["x"] = xx.grad.detach().clone()
        results[name] = (y.detach().clone(), grads, jm)
    torch.cuda.synchronize()
    fk = [k for _, ks in fusions(thunder.last_traces(results["b200"][2])[-1]) for k in ks]
    assert "b200_softmax_fwd" in fk
    err = (results["b200"][0].float() - results["torch"][0].float()).norm() / results["torch"][0].float().norm()
    assert err.item() < 1e-2
    _grads_close(results, zero_in_exact_arithmetic=("k_proj.bias",))


def test_strided_operands_of_a_family_region_are_read_in_place():
    """SwiGLU on the two halves of a fused gate/up projection (`x.chunk(2, -1)`: row stride 2*I): the region still is the
    SwiGLU family, but the launch goes to the program's generated kernel, which reads the strided halves in place —
    no dense copy (ops.COPY_STATS stays put), same bits as the dense route."""
    from b200ex import ops

    g = torch.Generator(device=DEV).manual_seed(5)
    gu = (torch.randn(256, 2 * 1024, generator=g, device=DEV) * 2).to(BF)
    fn = lambda a, b: F.silu(a) * b
    jf = thunder.jit(fn, executors=[b200.ex])
    a, b = gu.chunk(2, -1)
    assert not a.is_contiguous()
    before = dict(ops.COPY_STATS)
    got = jf(a, b)
    assert ops.COPY_STATS == before, "a strided operand was copied"
    dense = jf(a.contiguous(), b.contiguous())
    assert [k for _, ks in fusions(thunder.last_traces(jf)[-1]) for k in ks] == ["b200k_swiglu_fwd"]
    assert torch.equal(got, dense)
    assert torch.equal(got, thunder.jit(fn, executors=["torch"])(a, b))
