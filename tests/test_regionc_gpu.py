"""GPU tests of the generated region kernels (b200ex/regionc.py) through the plugin boundary: the reference's own
plain-torch model pieces jitted with ``executors=[b200]`` against the reference's torch executor on the same CUDA
inputs.  Layout-only regions must be bit-equal (they move bytes); regions with reductions reassociate fp32 sums, so
they are held to north_star's tolerances: fp32 rtol 1e-5 / atol 1e-6 (scaled by the row length for sums), bf16 rtol
1.6e-2 plus one bf16 ulp of the tensor's largest magnitude where a sum can cancel."""

import pytest
import torch
import torch.nn.functional as F

from ref_models import litgpt_rmsnorm, llama2_rmsnorm, qkv_split_rope
from thunder_env import fusions, import_thunder

pytestmark = pytest.mark.gpu

thunder, b200 = import_thunder()  # raises when baseline/_ref is missing: the plugin path must not skip silently

DEV = "cuda:0"
BF, F32 = torch.bfloat16, torch.float32


def run_both(fn, args, diff=None, seed=7):
    diff = range(len(args)) if diff is None else diff
    res = {}
    for tag, execs in (("b200", [b200.ex]), ("torch", ["torch"])):
        a = [x.detach().clone().requires_grad_(i in diff and x.is_floating_point()) for i, x in enumerate(args)]
        if tag == "b200":  # keep the caller's memory layout (clone() makes strided inputs dense)
            a = [torch.empty_strided(x.shape, x.stride(), dtype=x.dtype, device=x.device).copy_(x).requires_grad_(y.requires_grad) for x, y in zip(args, a)]
        jf = thunder.jit(fn, executors=execs)
        o = jf(*a)
        o = list(o) if isinstance(o, (tuple, list)) else [o]
        g = torch.Generator(device=DEV).manual_seed(seed)
        cot = [torch.randn(t.shape, generator=g, device=DEV).to(t.dtype) for t in o]
        ins = [x for x in a if x.requires_grad]
        grads = list(torch.autograd.grad(o, ins, cot, allow_unused=True)) if ins else []
        res[tag] = (o, grads, jf)
    torch.cuda.synchronize()
    return res


def kinds(jf):
    fw = fusions(thunder.last_traces(jf)[-1])
    bwt = thunder.last_backward_traces(jf)
    return fw, (fusions(bwt[-1]) if bwt else [])


def pairs(res):
    return list(zip(res["b200"][0] + res["b200"][1], res["torch"][0] + res["torch"][1]))


def assert_bit_equal(res):
    for got, want in pairs(res):
        assert got.shape == want.shape and got.dtype == want.dtype
        assert torch.equal(got, want), f"max diff {(got.float() - want.float()).abs().max().item()}"


def assert_reduction_close(res):
    for got, want in pairs(res):
        assert got.shape == want.shape and got.dtype == want.dtype
        w = want.float()
        if want.dtype is F32:
            torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6 * max(1.0, float(w.abs().max())))
        else:
            torch.testing.assert_close(got.float(), w, rtol=1.6e-2, atol=float(w.abs().max()) * 2.0**-8)


ROPE = [(32, 8, 128, 128, 2, 64), (8, 8, 64, 64, 2, 32), (8, 1, 64, 64, 3, 16), (8, 2, 64, 16, 2, 24), (32, 8, 128, 128, 2, 4096)]


@pytest.mark.parametrize("n_head,groups,hs,rope_n,B,T", ROPE)
def test_reference_qkv_split_rope_one_region_each_way(n_head, groups, hs, rope_n, B, T):
    """The reference pins this slice to 1 forward + 1 backward fusion (thunder/tests/test_torch_compile_executor.py:
    66-89).  The last case is the Llama-3-8B shape of BASELINE.json configs[2]: qkv bf16 [2, 4096, 6144]."""
    g = torch.Generator(device=DEV).manual_seed(0)
    qkv = torch.randn(B, T, (n_head + 2 * groups) * hs, generator=g, device=DEV).to(BF)
    cos, sin = torch.randn(T, rope_n, generator=g, device=DEV).to(BF), torch.randn(T, rope_n, generator=g, device=DEV).to(BF)
    fn = lambda qkv, cos, sin: qkv_split_rope(qkv, cos, sin, n_head, groups, hs, rope_n)
    res = run_both(fn, [qkv, cos, sin], diff=[0])
    fw, bw = kinds(res["b200"][2])
    assert len(fw) == 1 and len(bw) == 1, (fw, bw)
    assert all(k.startswith("b200k_region") for k in fw[0][1] + bw[0][1])
    assert_bit_equal(res)


@pytest.fixture(params=["auto", "1", "v1", "0"])
def staging(request, monkeypatch):
    """How row kernels stage their rows: through the cp.async ring in shared memory ("1": with the geometry that
    assumes it, "v1": at the register geometry), prefetched in registers ("0"), or whichever the first call measures
    fastest ("auto", the default).  The arithmetic is the same; every row
    test runs all three so that a fault in the build that loses the timing cannot hide."""
    from b200ex import regionc

    monkeypatch.setattr(regionc, "_RING_MODE", request.param)
    return request.param


@pytest.mark.parametrize("dtype", [BF, F32])
@pytest.mark.parametrize("rows,hidden", [(64, 256), (33, 96), (8192, 4096), (5, 8200)])
def test_hand_spelled_rmsnorm_with_residual(staging, dtype, rows, hidden):
    """litgpt.model.RMSNorm spelled with torch ops: one row kernel forward, one backward (dx plus the weight
    gradient's column sum folded from per-CTA partials)."""
    g = torch.Generator(device=DEV).manual_seed(0)
    x, r, w = (torch.randn(*s, generator=g, device=DEV).to(dtype) for s in ((rows, hidden), (rows, hidden), (hidden,)))

    def fn(x, r, w):
        h = x + r
        return h, litgpt_rmsnorm(h, w)

    res = run_both(fn, [x, r, w])
    fw, bw = kinds(res["b200"][2])
    if hidden <= 8192:  # longer rows with this many resident operands do not fit the registers: declined, still correct
        assert len(fw) == 1 and fw[0][1] == ["b200k_region[row]"], fw
        assert len(bw) == 1 and bw[0][1][0] == "b200k_region[row]", bw
    assert_reduction_close(res)


def test_llama2_rmsnorm(staging):
    g = torch.Generator(device=DEV).manual_seed(0)
    x, w = torch.randn(4, 128, 512, generator=g, device=DEV).to(BF), torch.randn(512, generator=g, device=DEV)
    res = run_both(llama2_rmsnorm, [x, w])
    fw, bw = kinds(res["b200"][2])
    assert len(fw) == 1 and len(bw) == 1
    assert_reduction_close(res)


def test_hand_spelled_row_reductions(staging):
    g = torch.Generator(device=DEV).manual_seed(0)
    x = torch.randn(6, 40, 1000, generator=g, device=DEV)
    w, b = torch.randn(1000, generator=g, device=DEV), torch.randn(1000, generator=g, device=DEV)

    def ln(x, w, b):
        mu = x.mean(-1, keepdim=True)
        var = ((x - mu) ** 2).mean(-1, keepdim=True)
        return (x - mu) * torch.rsqrt(var + 1e-5) * w + b

    def sm(x):
        e = torch.exp(x - x.amax(-1, keepdim=True))
        return e / e.sum(-1, keepdim=True)

    def lsm(x):
        return x - torch.logsumexp(x, -1, keepdim=True)

    def varmean(x):
        v, m = torch.var_mean(x, dim=-1, keepdim=True, correction=1)
        return (x - m) / torch.sqrt(v + 1e-3)

    for fn, args in ((ln, [x, w, b]), (sm, [x]), (lsm, [x]), (varmean, [x])):
        res = run_both(fn, args)
        fw, _ = kinds(res["b200"][2])
        assert len(fw) == 1 and "b200k_region[row]" in fw[0][1], (fn.__name__, fw)
        for got, want in pairs(res):
            torch.testing.assert_close(got, want, rtol=1e-5, atol=2e-5)


def test_layout_prims_as_index_maps():
    g = torch.Generator(device=DEV).manual_seed(0)
    x, y = torch.randn(4, 6, 16, generator=g, device=DEV), torch.randn(4, 6, 16, generator=g, device=DEV)
    cases = {
        "transpose": lambda x, y: x.permute(2, 0, 1) * 2.0 + y.permute(2, 0, 1),
        "slice_stride": lambda x, y: x[:, 1:5:2, 4:12] * y[:, 0:2, 8:16],
        "pad": lambda x, y: F.pad(x * y, (2, 6, 1, 0), value=1.5) + 1.0,
        "negative_pad": lambda x, y: F.pad(x, (-2, -6)) * 3.0,
        "cat_outer": lambda x, y: torch.cat((x * 2.0, y, x + y), dim=1) - 1.0,
        "cat_inner_unaligned": lambda x, y: torch.cat((x[..., :3], y[..., 5:]), dim=-1) * 2.0,
        "reshape_split_merge": lambda x, y: (x.reshape(4, 2, 3, 4, 4) * 2.0).permute(0, 2, 1, 3, 4).reshape(12, 32) + 1.0,
        "expand_sum_small": lambda x, y: (x.unsqueeze(1).expand(4, 3, 6, 16) * y.unsqueeze(1)).sum(1),
        "squeeze_full": lambda x, y: x[:, :1].squeeze(1) + torch.full((4, 16), 0.25, device=DEV),
        "chunk_glu": lambda x, y: F.silu(x.chunk(2, -1)[0]) * x.chunk(2, -1)[1],
        "bias_grad_colsum": lambda x, y: F.gelu(x + y[0, 0]),
    }
    for name, fn in cases.items():
        res = run_both(fn, [x, y])
        fw, _ = kinds(res["b200"][2])
        assert len(fw) == 1, (name, fw)
        for got, want in pairs(res):
            torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6, msg=lambda m: f"{name}: {m}")


def test_strided_and_misaligned_inputs_are_read_in_place():
    """Runtime strides pick the load mode per leaf (vector / scalar); nothing is copied to make an input dense."""
    g = torch.Generator(device=DEV).manual_seed(0)
    base = torch.randn(64, 520, generator=g, device=DEV).to(BF)
    x = base[:, 3:515]                     # misaligned start, row stride 520
    y = torch.randn(512, 64, generator=g, device=DEV).to(BF).t()  # transposed: inner stride 64
    fn = lambda x, y: torch.cat((x[:, :256] * y[:, 256:], x[:, 256:] + y[:, :256]), dim=-1)
    res = run_both(fn, [x, y])
    fw, _ = kinds(res["b200"][2])
    assert len(fw) == 1 and fw[0][1] == ["b200k_region[gather]"]
    assert_bit_equal(res)


def test_output_views_cost_no_kernel():
    """Outputs that are reshapes / transposes of another output come back as torch views of it (SURVEY §7 "view ops
    glued onto regions"): one launch, and the views share its storage."""
    g = torch.Generator(device=DEV).manual_seed(0)
    x, y = torch.randn(8, 32, 64, generator=g, device=DEV).to(BF), torch.randn(8, 32, 64, generator=g, device=DEV).to(BF)

    def fn(x, y):
        z = x * y + 1.0
        return z, z.reshape(256, 64), z.reshape(256, 64).t()

    jf = thunder.jit(fn, executors=[b200.ex], fusion_type="dataflow")
    z, z2, z2t = jf(x, y)
    fw = fusions(thunder.last_traces(jf)[-1])
    assert len(fw) == 1 and len(fw[0][1]) == 1, fw
    assert z2.data_ptr() == z.data_ptr() and z2t.data_ptr() == z.data_ptr() and z2t.stride() == (1, 64)
    want = fn(x, y)
    for a, b in zip((z, z2, z2t), want):
        assert torch.equal(a, b)


def test_llama3_8b_block_in_litgpt_spelling_full_size():
    """BASELINE.json configs[2] with the model written as litgpt writes it (inline RMSNorm, plain-ops QKV split + RoPE;
    nothing in it names this repository): b200 finds the regions itself — row kernels for the norms, one gather kernel
    each way for the attention slice, the native SwiGLU kernels — and agrees with the torch executor."""
    import dataclasses

    from baseline.llama_model import LLAMA3_8B, Block, build_rope_cache
    from parity import assert_tensor_scale_parity
    from thunder.executors.sdpaex import sdpa_ex

    cfg = dataclasses.replace(LLAMA3_8B, spelling="litgpt")
    torch.manual_seed(0)
    block = Block(cfg).to(DEV).to(BF)
    B, T = 2, 4096
    cos, sin = build_rope_cache(T, cfg.rope_n_elem, base=cfg.rope_base, device=DEV, dtype=BF)
    x0 = (torch.randn(B, T, cfg.n_embd, device=DEV) * 0.5).to(BF)
    cot = torch.randn(B, T, cfg.n_embd, device=DEV).to(BF)
    res = {}
    from baseline.attention import cudnn_sdpa_ex  # the attention executor list of the training benchmark (both arms)

    for name, execs in (("b200", [cudnn_sdpa_ex, sdpa_ex, b200.ex]), ("torch", [cudnn_sdpa_ex, sdpa_ex, "torch"])):
        block.zero_grad(set_to_none=True)
        x = x0.clone().requires_grad_(True)
        jb = thunder.jit(block, executors=execs)
        y = jb(x, cos, sin)
        y.backward(cot)
        res[name] = (y.detach(), x.grad.detach(), {n: p.grad.detach().clone() for n, p in block.named_parameters()}, jb)
    torch.cuda.synchronize()
    jb = res["b200"][3]
    fk = [k for _, ks in fusions(thunder.last_traces(jb)[-1]) for k in ks]
    bk = [k for _, ks in fusions(thunder.last_backward_traces(jb)[-1]) for k in ks]
    assert fk.count("b200k_region[row]") == 2 and fk.count("b200k_region[gather]") == 1 and "b200k_swiglu_fwd" in fk, fk
    assert bk.count("b200k_region[row]") == 2 and bk.count("b200k_region[gather]") == 1 and "b200k_swiglu_bwd" in bk, bk

    def rel(a, b):
        return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()

    assert rel(res["b200"][0], res["torch"][0]) < 1e-2 and rel(res["b200"][1], res["torch"][1]) < 1.6e-2
    assert_tensor_scale_parity(res["b200"][0], res["torch"][0], what="block output")
    assert_tensor_scale_parity(res["b200"][1], res["torch"][1], what="block input gradient")
    for n, g in res["b200"][2].items():
        assert rel(g, res["torch"][2][n]) < 1.6e-2, f"grad {n}"
        assert_tensor_scale_parity(g, res["torch"][2][n], what=f"grad {n}")


def test_unvectorisable_pointwise_launches_use_the_gather_kernel_not_the_interpreter():
    """Inner extents that are not a multiple of 8, transposed operands, misaligned views: the elementwise path hands
    such launches to a generated gather kernel (narrower or element-wise loads) instead of the ahead-of-time
    interpreter, which decodes its program per element."""
    from b200ex.pointwise import PwLaunch

    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(37, 33, generator=g, device=DEV).to(BF)
    yt = torch.randn(33, 37, generator=g, device=DEV).to(BF).t()   # inner stride 37
    base = torch.randn(37 * 33 + 3, generator=g, device=DEV).to(BF)
    z = base[3:].view(37, 33)                                        # 6-byte aligned start
    fn = lambda x, y, z: torch.tanh(x) * y + z * 0.5
    jf = thunder.jit(fn, executors=[b200.ex])
    got = jf(x, yt, z)
    want = thunder.jit(fn, executors=["torch"])(x, yt, z)
    assert torch.equal(got, want)
    launches = []
    for bs in thunder.last_traces(jf)[-1].bound_symbols:
        if bs.sym.is_fusion and bs.sym.name.startswith("b200Fusion"):
            for st in bs._call_ctx[bs.sym.name].steps:
                d = getattr(st.run, "__defaults__", None) or ()
                launches += [o for o in d if isinstance(o, PwLaunch)]
    assert launches and all(l.gather_launches >= 1 and l.jit_launches == 0 for l in launches)
