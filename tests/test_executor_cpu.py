"""CPU: the b200 executor behind the reference's plugin API — registration, claiming, region lowering, fuel,
shape-only rule, signature table, and the no-fallback guarantee.  Structure only: claiming of CPU proxies is
switched on for these tests (b200._testing_claim_cpu), executing a region on CPU tensors must raise."""

import pytest
import torch
import torch.nn.functional as F

from thunder_env import fusions, import_thunder

thunder, b200 = import_thunder()  # raises when baseline/_ref is missing: the plugin path must not skip silently

from thunder.extend import FUEL_LEVEL, FusionExecutor, get_executor


@pytest.fixture(autouse=True)
def _claim_cpu():
    b200._testing_claim_cpu(True)
    b200.ex.set_fuel(FUEL_LEVEL.UNLIMITED)
    yield
    b200._testing_claim_cpu(False)
    b200.ex.set_fuel(FUEL_LEVEL.UNLIMITED)


def _trace(fn, *args, **opts):
    jf = thunder.jit(fn, executors=[b200.ex], **opts)
    with pytest.raises(RuntimeError, match="CUDA"):
        jf(*args)  # lowering succeeded; launching on CPU tensors is refused (no CPU path)
    bw = thunder.last_backward_traces(jf)
    return thunder.last_traces(jf)[-1], (bw[-1] if bw else None)


def test_registered_under_the_reference_plugin_api():
    assert isinstance(b200.ex, FusionExecutor) and b200.ex.name == "b200"
    assert get_executor("b200") is b200.ex
    b200._testing_claim_cpu(False)
    jf = thunder.jit(lambda x: x + 1, executors=["b200"])  # name lookup works once the module is imported
    assert jf(torch.ones(2)).tolist() == [2.0, 2.0]  # CPU tensors are left to the torch executor


def test_cpu_tensors_are_not_claimed_by_default():
    b200._testing_claim_cpu(False)
    jf = thunder.jit(lambda x, y: F.silu(x) * y, executors=[b200.ex])
    out = jf(torch.randn(4, 8), torch.randn(4, 8))
    assert out.shape == (4, 8) and fusions(thunder.last_traces(jf)[-1]) == []


def test_readme_mlp_on_cpu_is_the_reference_bit_for_bit():
    """BASELINE.json configs[0]: the reference's README MLP (Linear 2048->4096, ReLU, Linear 4096->64, batch 64, fp32)
    on the CPU.  b200 claims nothing there, so with b200 in the executor list the model must give exactly what the
    reference's torch executor gives, forward and backward."""
    b200._testing_claim_cpu(False)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(2048, 4096), torch.nn.ReLU(), torch.nn.Linear(4096, 64))
    x = torch.randn(64, 2048)
    outs = []
    for execs in ([b200.ex], ["torch"]):
        model.zero_grad(set_to_none=True)
        jm = thunder.jit(model, executors=execs)
        y = jm(x)
        y.square().mean().backward()
        outs.append((y.detach().clone(), [p.grad.clone() for p in model.parameters()]))
        if execs != ["torch"]:
            assert fusions(thunder.last_traces(jm)[-1]) == []
    (y0, g0), (y1, g1) = outs
    assert torch.equal(y0, y1) and all(torch.equal(a, b) for a, b in zip(g0, g1))
    torch.testing.assert_close(y0, model(x), rtol=1e-5, atol=1e-6)


def test_rmsnorm_residual_and_swiglu_regions():
    bf = torch.bfloat16

    def f(x, r, w, a, b):
        h = x + r
        return h, F.rms_norm(h, (64,), w, 1e-5), F.silu(a) * b

    mk = lambda *s: torch.randn(*s).to(bf).requires_grad_(True)
    fw, bw = _trace(f, mk(4, 64), mk(4, 64), mk(64), mk(4, 64), mk(4, 64))
    (fname, fk), = fusions(fw)
    (bname, bk), = fusions(bw)
    assert sorted(fk) == ["b200_rms_norm_fwd+residual", "b200k_swiglu_fwd"]
    assert sorted(bk) == ["b200_rms_norm_bwd+residual", "b200k_swiglu_bwd"]
    # fusion names end in the counter that update_fusion_call_ctx parses (passes.py:167-172)
    assert fname[-1].isdigit() and bname[-1].isdigit()


def test_rmsnorm_fp32_and_no_residual():
    def f(x, w):
        return F.rms_norm(x, (32,), w)

    fw, bw = _trace(f, torch.randn(3, 5, 32, requires_grad=True), torch.randn(32, requires_grad=True))
    assert fusions(fw)[0][1] == ["b200_rms_norm_fwd"] and fusions(bw)[0][1] == ["b200_rms_norm_bwd"]


def test_cross_entropy_is_claimed_at_op_level():
    def f(logits, t):
        return F.cross_entropy(logits, t, ignore_index=-1)

    logits = torch.randn(8, 50).to(torch.bfloat16).requires_grad_(True)
    fw, bw = _trace(f, logits, torch.randint(0, 50, (8,)))
    assert fusions(fw)[0][1] == ["b200_cross_entropy_fwd"] and fusions(bw)[0][1] == ["b200_cross_entropy_bwd"]


def test_cross_entropy_declines_what_the_kernel_does_not_cover():
    b200._testing_claim_cpu(True)

    def f(logits, t):
        return F.cross_entropy(logits, t, label_smoothing=0.1)

    jf = thunder.jit(f, executors=[b200.ex])
    try:
        jf(torch.randn(8, 50), torch.randint(0, 50, (8,)))
    except RuntimeError:
        pass
    kinds = [k for _, ks in fusions(thunder.last_traces(jf)[-1]) for k in ks]
    assert "b200_cross_entropy_fwd" not in kinds


def test_qkv_split_rope_lookaside():
    def f(qkv, cos, sin):
        return b200.qkv_split_rope(qkv, cos, sin, 4, 2, 16, 16)

    qkv = torch.randn(2, 8, 128).to(torch.bfloat16).requires_grad_(True)
    cos, sin = torch.randn(8, 16).to(torch.bfloat16), torch.randn(8, 16).to(torch.bfloat16)
    jf = thunder.jit(f, executors=[b200.ex])
    with pytest.raises(RuntimeError, match="no eager fallback"):
        jf(qkv, cos, sin)  # the claimed operator has no CPU / eager path: non-CUDA inputs are an error
    fw, bw = thunder.last_traces(jf)[-1], thunder.last_backward_traces(jf)[-1]
    assert fusions(fw)[0][1] == ["b200_qkv_split_rope"] and fusions(bw)[0][1] == ["b200_qkv_rope_bwd"]
    # without the executor the same function is ordinary PyTorch
    q, k, v = b200.qkv_split_rope(qkv, cos, sin, 4, 2, 16, 16)
    assert q.shape == (2, 4, 8, 16) and k.shape == (2, 4, 8, 16)


def test_bias_activation_regions_route_to_the_native_kernels():
    mk = lambda *s: torch.randn(*s).to(torch.bfloat16).requires_grad_(True)
    fw, bw = _trace(lambda x, b: F.gelu(x + b), mk(6, 32), mk(32))
    assert [k for _, ks in fusions(fw) for k in ks] == ["b200k_bias_act_fwd[gelu]"]
    # dx from the fused backward kernel, the bias gradient from its column sums (no second pass over dx)
    assert [k for _, ks in fusions(bw) for k in ks] == ["b200k_bias_act_bwd[gelu]", "b200k_colsum"]
    # relu under autograd saves the BOOL mask `a > 0`, not the pre-activation the family kernel works on: the region
    # keeps the trace's dtypes and runs as a generic program (one launch forward, where(mask, g, 0) backward)
    fw, bw = _trace(lambda x, b: F.relu(x + b), mk(6, 32), mk(32))
    assert [k for _, ks in fusions(fw) for k in ks] == ["b200k_pointwise"]
    assert [k for _, ks in fusions(bw) for k in ks] == ["b200k_pointwise", "b200k_colsum"]
    # outside autograd there is nothing to save and relu(x + b) is the family kernel
    jf = thunder.jit(lambda x, b: F.relu(x + b), executors=[b200.ex])
    with pytest.raises(RuntimeError):
        jf(torch.randn(6, 32).to(torch.bfloat16), torch.randn(32).to(torch.bfloat16))
    assert [k for _, ks in fusions(thunder.last_traces(jf)[-1]) for k in ks] == ["b200k_bias_act_fwd[relu]"]


def test_generic_pointwise_region_with_broadcast():
    def f(x, b):
        return torch.tanh(x * b) + 0.5 * x

    fw, bw = _trace(f, torch.randn(6, 32).to(torch.bfloat16).requires_grad_(True), torch.randn(32).to(torch.bfloat16).requires_grad_(True))
    assert all(k == "b200k_pointwise" for _, ks in fusions(fw) for k in ks) and fusions(fw)
    kinds = [k for _, ks in fusions(bw) for k in ks]
    assert "b200k_pointwise" in kinds and "b200k_colsum" in kinds  # db = sum over rows is claimed too


def test_optimization_fuel():
    def f(x, y):
        a = F.silu(x) * y
        b = torch.matmul(a, a.t())
        return F.gelu(b)

    x, y = torch.randn(8, 8), torch.randn(8, 8)
    fw, _ = _trace(f, x, y)
    assert len(fusions(fw)) == 2
    b200.ex.set_fuel(1)
    fw, _ = _trace(f, x, y)
    assert len(fusions(fw)) == 1
    b200.ex.set_fuel(0)
    jf = thunder.jit(f, executors=[b200.ex])
    jf(x, y)  # nothing claimed -> runs on the torch executor
    assert fusions(thunder.last_traces(jf)[-1]) == []
    with pytest.raises(ValueError):
        b200.ex.set_fuel(-1)


def test_shape_only_groups_are_not_fused():
    def f(x):
        return x.unsqueeze(0).expand(2, 4, 8)

    jf = thunder.jit(f, executors=[b200.ex])
    jf(torch.randn(4, 8))
    assert fusions(thunder.last_traces(jf)[-1]) == []


def test_llama_block_region_census():
    """Region census of a Llama-style model: with b200's default partitioner (dataflow) every
    named family is lowered to its hand-written kernel."""
    from baseline.llama_model import GPT, build_rope_cache, tiny_config

    cfg = tiny_config(spelling="ops")
    torch.manual_seed(0)
    m = GPT(cfg).to(torch.bfloat16)
    cos, sin = build_rope_cache(16, cfg.rope_n_elem, dtype=torch.bfloat16)
    idx, tgt = torch.randint(0, 256, (2, 16)), torch.randint(0, 256, (2, 16))
    jm = thunder.jit(m, executors=[b200.ex])
    with pytest.raises(RuntimeError, match="CUDA"):
        jm(idx, cos, sin, tgt)
    fk = [k for _, ks in fusions(thunder.last_traces(jm)[-1]) for k in ks]
    bk = [k for _, ks in fusions(thunder.last_backward_traces(jm)[-1]) for k in ks]
    L = cfg.n_layer
    assert fk.count("b200_qkv_split_rope") == L and fk.count("b200k_swiglu_fwd") == L
    assert fk.count("b200_rms_norm_fwd") + fk.count("b200_rms_norm_fwd+residual") == 2 * L + 1
    assert fk.count("b200_rms_norm_fwd+residual") == 2 * L  # every norm but the first sees a residual add
    assert fk.count("b200_cross_entropy_fwd") == 1
    assert bk.count("b200_qkv_rope_bwd") == L and bk.count("b200k_swiglu_bwd") == L
    assert bk.count("b200_rms_norm_bwd") + bk.count("b200_rms_norm_bwd+residual") == 2 * L + 1
    assert bk.count("b200_cross_entropy_bwd") == 1


def test_partitioner_options():
    """`fusion_type` (the reference's option) wins; otherwise `b200_fusion_type` picks, default dataflow.  With the
    consecutive partitioner a matmul between the two halves of the SwiGLU backward splits it in two regions."""
    lin = torch.nn.Linear(16, 16, bias=False).to(torch.bfloat16)

    def f(x, w1, w2):
        return lin(F.silu(x @ w1) * (x @ w2))

    mk = lambda *s: torch.randn(*s).to(torch.bfloat16).requires_grad_(True)
    args = (mk(8, 16), mk(16, 16), mk(16, 16))

    def kinds(**opts):
        _, bw = _trace(f, *args, **opts)
        return [k for _, ks in fusions(bw) for k in ks]

    assert "b200k_swiglu_bwd" in kinds()
    assert "b200k_swiglu_bwd" in kinds(fusion_type="dataflow")
    for opts in ({"fusion_type": "consecutive"}, {"b200_fusion_type": "consecutive"}):
        ks = kinds(**opts)
        assert "b200k_swiglu_bwd" not in ks and ks.count("b200k_pointwise") >= 2


def test_signature_table_is_current():
    """tools/gen_signatures.py regenerates the table of pointwise programs that have a native kernel; the
    committed table must match (drift means thunder's decomposition or our canonicalisation changed)."""
    import importlib.util
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("gen_signatures", os.path.join(root, "tools", "gen_signatures.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    committed = dict(b200._NATIVE_SIGNATURES)
    path = os.path.join(root, "lightning-thunder_b200", "b200ex", "_signatures.py")
    before = open(path).read()
    try:
        table = gen.main()
    finally:
        open(path, "w").write(before)
        b200._NATIVE_SIGNATURES.clear()
        b200._NATIVE_SIGNATURES.update(committed)
    assert table == committed and len(table) >= 4


def test_trace_level_cse_removes_duplicate_subsymbols():
    """thunder/tests/test_nvfuser.py:190-215: (t0 + 5) is computed twice in the source, once in the fused trace."""
    b200._testing_claim_cpu(True)
    try:
        def func(x):
            t0 = x.relu()
            t1 = t0 + 5
            t2 = t0 + 5
            t3 = t0 @ t0
            return torch.where(t3 > t1, t1, t2)

        jf = thunder.jit(func, executors=[b200.ex])
        with pytest.raises(RuntimeError, match="CUDA"):
            jf(torch.randn(5, 5))
        tr = thunder.last_traces(jf)[-1]
        fs = [b for b in tr.bound_symbols if b.sym.is_fusion]
        assert len(fs) == 2  # split by the matmul
        adds = [s for f in fs for s in f.subsymbols if s.sym.name == "add"]
        assert len(adds) == 1, [s.sym.name for f in fs for s in f.subsymbols]
    finally:
        b200._testing_claim_cpu(False)


def test_trace_level_cse_dedupes_fusion_arguments():
    """thunder/tests/test_nvfuser.py:218-245: x @ y appears twice; after CSE the single fusion takes one matmul result
    plus z and w, and returns one tensor."""
    b200._testing_claim_cpu(True)
    try:
        def func(w, x, y, z):
            t1 = (x @ y) + z
            t3 = (x @ y) + w
            return t1 + t3

        jf = thunder.jit(func, executors=[b200.ex], fusion_type="dataflow")
        with pytest.raises(RuntimeError, match="CUDA"):
            jf(*[torch.randn(5, 5) for _ in range(4)])
        tr = thunder.last_traces(jf)[-1]
        fs = [b for b in tr.bound_symbols if b.sym.is_fusion]
        assert len(fs) == 1
        assert len([b for b in tr.bound_symbols if b.sym.name == "matmul"]) == 1
        assert len(fs[0].flat_proxy_args) == 3 and {a.name for a in fs[0].flat_proxy_args} >= {"z", "w"}
        assert len(fs[0].flat_proxy_outs) == 1
    finally:
        b200._testing_claim_cpu(False)
