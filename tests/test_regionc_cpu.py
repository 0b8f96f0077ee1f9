"""CPU tests of the region compiler (b200ex/regionc.py): layout prims as index maps, row reductions, column sums.

The compiler front end is exercised end to end through the plugin boundary: ``thunder.jit(fn, executors=[b200])``
with claiming switched on for CPU proxies, and every generated program evaluated by the CPU evaluator in
tests/regionc_ref.py (test infrastructure) in place of the CUDA launch.  Results are compared with the reference's
torch executor on the same inputs.  The CUDA source of every program is also compiled with NVRTC for sm_100a, which
needs no GPU.  The GPU counterpart (the generated kernels themselves) is tests/test_regionc_gpu.py."""

import math

import pytest
import torch
import torch.nn.functional as F

import regionc_ref
from thunder_env import fusions, import_thunder

thunder, b200 = import_thunder()
from b200ex import jit, regionc  # noqa: E402

BF, F32 = torch.bfloat16, torch.float32


@pytest.fixture(autouse=True)
def _cpu_claiming(monkeypatch):
    programs = []

    def ref_run(self, env):
        programs.append(self.P)
        res = regionc_ref.run_program(self.P, {n: env[n] for n in self.in_names})
        for name, (src, shape) in self.views.items():
            res[name] = res[src].view(shape)
        return res

    monkeypatch.setattr(regionc.RegionLaunch, "run", ref_run)
    monkeypatch.setattr(b200.B200Region, "force_region_compiler", True)
    b200._testing_claim_cpu(True)
    yield programs
    b200._testing_claim_cpu(False)


def run_both(fn, args, diff=None):
    """fn through b200 (CPU evaluator) and through the torch executor: outputs + gradients of `diff` args."""
    diff = range(len(args)) if diff is None else diff
    res = {}
    for tag, execs in (("b200", [b200.ex]), ("torch", ["torch"])):
        a = [x.clone().requires_grad_(i in diff and x.is_floating_point()) for i, x in enumerate(args)]
        jf = thunder.jit(fn, executors=execs)
        o = jf(*a)
        o = list(o) if isinstance(o, (tuple, list)) else [o]
        g = torch.Generator().manual_seed(7)
        cot = [torch.randn(t.shape, generator=g).to(t.dtype) for t in o]
        grads = torch.autograd.grad(o, [x for x in a if x.requires_grad], cot, allow_unused=True) if any(x.requires_grad for x in a) else []
        res[tag] = (o, list(grads), jf)
    return res


def assert_same(res, exact=True, rtol=0.0, atol=0.0):
    for got, want in zip(res["b200"][0] + res["b200"][1], res["torch"][0] + res["torch"][1]):
        assert (got is None) == (want is None)
        if got is None:
            continue
        assert got.shape == want.shape and got.dtype == want.dtype
        if exact:
            assert torch.equal(got, want), f"max diff {(got.float() - want.float()).abs().max().item()}"
        else:
            torch.testing.assert_close(got, want, rtol=rtol, atol=atol)


def region_kinds(jf):
    fw = fusions(thunder.last_traces(jf)[-1])
    bw = fusions(thunder.last_backward_traces(jf)[-1]) if thunder.last_backward_traces(jf) else []
    return fw, bw


def compiles(programs):
    """Every program's CUDA source must build for sm_100a (NVRTC needs no device)."""
    if not jit.enabled():
        pytest.skip("cuda-python / NVRTC not importable here")
    seen = set()
    for P in programs:
        if P.signature in seen:
            continue
        seen.add(P.signature)

        class _T:  # contiguous, aligned stand-in for leaf_modes
            def __init__(self, shape):
                self.shape = shape
                self._st = [0] * len(shape)
                acc = 1
                for i in range(len(shape) - 1, -1, -1):
                    self._st[i] = acc
                    acc *= shape[i]

            def stride(self, i):
                return self._st[i]

            def data_ptr(self):
                return 0

            dtype = None

        ts = []
        for lf in P.leaves:
            t = _T(P.in_shapes[lf.name])
            t.dtype = lf.dtype
            ts.append(t)
        modes = regionc.leaf_modes(P, ts)
        # every build of a row program (register prefetch, cp.async ring at either geometry) must compile
        for variant in regionc.row_variants(P, modes) if P.kind == "row" else [0]:
            cubin = jit.compile_cubin(regionc.source_for(P, modes, variant=variant))
            assert len(cubin) > 1000


from ref_models import litgpt_rmsnorm, llama2_rmsnorm, qkv_split_rope  # noqa: E402

ROPE_CONFIGS = [
    (32, 8, 128, 128),  # Llama-3-8B: GQA, full rotary
    (8, 8, 64, 64),     # MHA
    (8, 1, 64, 64),     # MQA: k, v keep one head
    (8, 2, 64, 16),     # partial rotary (rope_n_elem < head_size): the second `cat` is live
]


@pytest.mark.parametrize("n_head,groups,hs,rope_n", ROPE_CONFIGS)
def test_reference_qkv_split_rope_is_one_region_each_way(_cpu_claiming, n_head, groups, hs, rope_n):
    """thunder/tests/test_torch_compile_executor.py:66-89 pins the LitGPT QKV-split + RoPE slice to exactly one fusion
    forward and one backward.  Same here, with bit-equal results versus the torch executor."""
    torch.manual_seed(0)
    B, T = 2, 16
    qkv = torch.randn(B, T, (n_head + 2 * groups) * hs).to(BF)
    cos, sin = torch.randn(T, rope_n).to(BF), torch.randn(T, rope_n).to(BF)
    fn = lambda qkv, cos, sin: qkv_split_rope(qkv, cos, sin, n_head, groups, hs, rope_n)
    res = run_both(fn, [qkv, cos, sin], diff=[0])
    fw, bw = region_kinds(res["b200"][2])
    assert len(fw) == 1 and len(bw) == 1, (fw, bw)
    assert_same(res)
    compiles(_cpu_claiming)


# ---- norms spelled with torch ops ------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [BF, F32])
@pytest.mark.parametrize("hidden", [256, 96])
def test_hand_spelled_rmsnorm_with_residual(_cpu_claiming, dtype, hidden):
    torch.manual_seed(0)
    x, r, w = torch.randn(4, 16, hidden).to(dtype), torch.randn(4, 16, hidden).to(dtype), torch.randn(hidden).to(dtype)

    def fn(x, r, w):
        h = x + r
        return h, litgpt_rmsnorm(h, w)

    res = run_both(fn, [x, r, w])
    fw, bw = region_kinds(res["b200"][2])
    assert len(fw) == 1 and fw[0][1] == ["b200k_region[row]"], fw
    assert len(bw) == 1, bw
    # one row kernel (dx, with the weight gradient's column sum folded in) and at most the fold's follow-up convert
    assert bw[0][1][0] == "b200k_region[row]" and len(bw[0][1]) <= 2, bw
    if dtype is F32:
        assert_same(res, exact=False, rtol=1e-5, atol=1e-5)
    else:
        # sums are reassociated (fp32), so a bf16 result can land on the neighbouring value: rtol 1.6e-2 (north_star)
        # plus one bf16 ulp of the tensor's largest magnitude for the cancelling sums in dx
        o, g, _ = res["torch"]
        for got, want in zip(res["b200"][0] + res["b200"][1], o + g):
            torch.testing.assert_close(got.float(), want.float(), rtol=1.6e-2, atol=float(want.float().abs().max()) * 2.0**-8)
    compiles(_cpu_claiming)


def test_llama2_rmsnorm(_cpu_claiming):
    torch.manual_seed(0)
    x, w = torch.randn(3, 8, 128).to(BF), torch.randn(128)
    res = run_both(llama2_rmsnorm, [x, w])
    fw, bw = region_kinds(res["b200"][2])
    assert len(fw) == 1 and len(bw) == 1, (fw, bw)
    assert "b200k_region[row]" in fw[0][1] and "b200k_region[row]" in bw[0][1]
    o, g, _ = res["torch"]
    for got, want in zip(res["b200"][0] + res["b200"][1], o + g):
        torch.testing.assert_close(got.float(), want.float(), rtol=1.6e-2, atol=float(want.float().abs().max()) * 2.0**-8)
    compiles(_cpu_claiming)


def test_hand_spelled_layernorm_softmax_logsoftmax(_cpu_claiming):
    torch.manual_seed(0)
    x = torch.randn(6, 10, 64)
    w, b = torch.randn(64), torch.randn(64)

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
        fw, bw = region_kinds(res["b200"][2])
        assert len(fw) == 1 and any(k == "b200k_region[row]" for k in fw[0][1]), (fn.__name__, fw)
        # amax's backward counts ties with an int64 sum, which stays with the torch executor and cuts the region in two
        assert len(bw) == (2 if fn in (sm, lsm) else 1), (fn.__name__, bw)
        assert_same(res, exact=False, rtol=1e-5, atol=2e-6)
    compiles(_cpu_claiming)


# ---- layout prims one by one ------------------------------------------------------------------------------------------
def test_layout_prims_as_index_maps(_cpu_claiming):
    torch.manual_seed(0)
    x = torch.randn(4, 6, 16)
    y = torch.randn(4, 6, 16)
    cases = {
        "transpose": lambda x, y: x.permute(2, 0, 1) * 2.0 + y.permute(2, 0, 1),
        "slice_stride": lambda x, y: x[:, 1:5:2, 4:12] * y[:, 0:2, 8:16],
        "pad": lambda x, y: F.pad(x * y, (2, 6, 1, 0), value=1.5) + 1.0,
        "negative_pad": lambda x, y: F.pad(x, (-2, -6)) * 3.0,
        "cat_outer": lambda x, y: torch.cat((x * 2.0, y, x + y), dim=1) - 1.0,
        "cat_inner_unaligned": lambda x, y: torch.cat((x[..., :3], y[..., 5:]), dim=-1) * 2.0,
        "reshape_split_merge": lambda x, y: (x.reshape(4, 2, 3, 4, 4) * 2.0).permute(0, 2, 1, 3, 4).reshape(12, 32) + 1.0,
        "expand_sum_small": lambda x, y: (x.unsqueeze(1).expand(4, 3, 6, 16) * y.unsqueeze(1)).sum(1),
        "squeeze_full": lambda x, y: x[:, :1].squeeze(1) + torch.full((4, 16), 0.25),
        "chunk_glu": lambda x, y: F.silu(x.chunk(2, -1)[0]) * x.chunk(2, -1)[1],
        "bias_grad_colsum": lambda x, y: F.gelu(x + y[0, 0]),
    }
    for name, fn in cases.items():
        res = run_both(fn, [x, y])
        fw, bw = region_kinds(res["b200"][2])
        assert len(fw) == 1, (name, fw)
        assert_same(res, exact=False, rtol=1e-5, atol=1e-6)
    compiles(_cpu_claiming)


def test_unsupported_layouts_decline_but_keep_the_elementwise_part(_cpu_claiming):
    """A reshape that cuts a dimension unevenly cannot become an index map: the group is declined as a whole, its
    elementwise runs still fuse and the layout prims go to the torch executor.  Results stay correct."""
    torch.manual_seed(0)
    x = torch.randn(6, 10)
    fn = lambda x: (x * 2.0)[:, 1:].reshape(9, 6) + 1.0  # [6, 9] -> [9, 6]: digits do not line up
    res = run_both(fn, [x])
    assert_same(res, exact=False, rtol=1e-6, atol=1e-6)


def test_row_constant_factors_leave_the_row_sums(_cpu_claiming):
    """sum(s * y) = s * sum(y) for s constant along the row: the hand-spelled LayerNorm backward then needs ONE pass
    over the row for its sums (every reduction in phase 0) instead of two with a barrier in between, and its builds
    include the shared-memory ring (16-bit operands, three phases' worth of work)."""
    torch.manual_seed(0)
    x = torch.randn(64, 512).to(BF)
    w, b = torch.randn(512).to(BF), torch.randn(512).to(BF)

    def ln(x, w, b):
        xf = x.float()
        mu = xf.mean(-1, keepdim=True)
        var = ((xf - mu) ** 2).mean(-1, keepdim=True)
        return ((xf - mu) * torch.rsqrt(var + 1e-5) * w.float() + b.float()).to(x.dtype)

    res = run_both(ln, [x, w, b])
    o, g, _ = res["torch"]
    for got, want in zip(res["b200"][0] + res["b200"][1], o + g):
        torch.testing.assert_close(got.float(), want.float(), rtol=1.6e-2, atol=float(want.float().abs().max()) * 2.0**-8)
    bwd = [P for P in _cpu_claiming if P.kind == "row" and any(o.kind == "col" for o in P.outputs)]
    assert len(bwd) == 1
    P = bwd[0]
    reduces = [i for i, s_ in enumerate(P.ssa) if s_[0].startswith("reduce_")]
    assert len(reduces) == 3 and all(P.phase[P.ssa[i][1]] == 0 for i in reduces), [(P.ssa[i], P.phase[i]) for i in reduces]
    compiles(_cpu_claiming)


def test_row_ring_plan_fits_the_sm(_cpu_claiming):
    """The staging ring of a row kernel (regionc.row_ring) is planned on the host: its shared memory must leave room for
    the CTAs per SM the launch bounds promise (200 KB split between them), it needs at least two stages, and the
    builds offered to the first-call timing start with the register build (the one used under stream capture)."""
    torch.manual_seed(0)
    x = torch.randn(64, 4096).to(BF)
    w, b = torch.randn(4096).to(BF), torch.randn(4096).to(BF)

    def ln(x, w, b):
        xf = x.float()
        mu = xf.mean(-1, keepdim=True)
        var = ((xf - mu) ** 2).mean(-1, keepdim=True)
        return ((xf - mu) * torch.rsqrt(var + 1e-5) * w.float() + b.float()).to(x.dtype)

    run_both(ln, [x, w, b])
    rows = [P for P in _cpu_claiming if P.kind == "row"]
    assert rows
    seen_ring = False
    for P in rows:
        class _T:
            def __init__(self, shape, dt):
                self.shape, self.dtype, self._st, acc = shape, dt, [0] * len(shape), 1
                for i in range(len(shape) - 1, -1, -1):
                    self._st[i] = acc
                    acc *= shape[i]

            def stride(self, i):
                return self._st[i]

            def data_ptr(self):
                return 0

        modes = regionc.leaf_modes(P, [_T(P.in_shapes[lf.name], lf.dtype) for lf in P.leaves])
        variants = regionc.row_variants(P, modes)
        assert variants[0] == 0
        for v in variants[1:]:
            seen_ring = True
            stages, units, smem = regionc.row_ring(P, modes, v == 2)
            T, C, RPC, block = regionc.row_geometry(P, staged=v == 2)
            assert stages >= 2 and smem == stages * units * block * 16
            assert smem * max(1, 512 // block) <= 200 * 1024
    assert seen_ring, "a 4096-wide bf16 LayerNorm should offer the ring"
