"""GPU: the reference's own OpInfo samples through ``executors=[b200]`` against each op's torch reference — the b200
counterpart of ``thunder/tests/test_ops.py::test_core_vs_torch_consistency`` (same sample generators from
``thunder/tests/opinfos.py``, same comparison: the sample's comparator or ``assert_close``, and when that fails the
reference's ``assert_closer`` against an fp64 evaluation).  It sweeps the pointwise vocabulary of SURVEY §8 row a4
(elementwise unary / binary ops, conditionals, casts, broadcasts) plus the ops with op-level rules, in fp32, bf16
and fp16, and checks that b200 claimed a region wherever every operand is a CUDA float tensor."""

import pytest
import torch
from torch.testing import assert_close

from thunder_env import fusions, import_thunder

pytestmark = pytest.mark.gpu

thunder, b200 = import_thunder()  # raises when baseline/_ref is missing: the plugin path must not skip silently
import thunder.core.dtypes as dtypes
from thunder.core.pytree import tree_flatten, tree_map
from thunder.tests import opinfos as _opinfos
from thunder.tests.framework import assert_closer
from thunder.tests.opinfos import custom_comparator

DEV = "cuda"
MAX_SAMPLES = 8
_OTHER = (
    "where", "clamp", "lerp", "addcmul", "addcdiv", "masked_fill", "type_as", "to", "gelu", "layer_norm", "rms_norm",
    "cross_entropy", "softmax", "log_softmax", "dropout", "zeros_like", "ones_like", "full_like", "expand", "unsqueeze", "squeeze", "sum", "mean",
)
_OPS = {o.name: o for o in _opinfos.opinfos if o.name in _OTHER and o.torch_reference is not None}
_OPS.update({o.name: o for o in list(_opinfos.elementwise_unary_ops) + list(_opinfos.elementwise_binary_ops) if o.torch_reference is not None})
_DTYPES = {"f32": dtypes.float32, "bf16": dtypes.bfloat16, "f16": dtypes.float16}
# regions claimed per (op, dtype); filled by the sweep and read by the coverage test that runs after it
_CLAIMED: dict = {}


def _cases():
    for name in sorted(_OPS):
        for tag, dt in _DTYPES.items():
            if dt in _OPS[name].dtypes():
                yield pytest.param(name, tag, id=f"{name}-{tag}")


class _TorchEx:
    """The reference's per-op test directives are keyed by executor name; the ones written for its torch executor
    describe what the ATen kernels themselves cannot do (e.g. no bf16 ndtri), which holds for any executor that
    leaves those ops to torch."""

    name = "torch"


def _directives(op, dt):
    """(comparator, skip reason | None, expected-failure flag) from the reference's own directives for this test."""
    comp, skip, xfail = assert_close, None, False
    for d in op.test_decorators("test_core_vs_torch_consistency", _TorchEx, DEV, dt):
        if isinstance(d, custom_comparator):
            comp = d.comparator
        elif getattr(d, "name", None) == "skip" or getattr(getattr(d, "mark", None), "name", None) == "skip":
            skip = "skipped by the reference's own directive for this op / dtype"
        elif getattr(d, "name", None) == "xfail" or getattr(getattr(d, "mark", None), "name", None) == "xfail":
            xfail = True
    return comp, skip, xfail


def _run(op, execs, sample):
    jf = thunder.jit(op.op, executors=execs, cache="no caching", disable_torch_autograd=True)
    return jf, jf(*sample.args, **sample.kwargs)


def _check(op, sample, comp, got):
    want = op.torch_reference(*sample.args, **sample.kwargs)
    try:
        comp(got, want)
    except AssertionError:
        up = lambda x: x.to(torch.double) if isinstance(x, torch.Tensor) and x.is_floating_point() else (torch.double if isinstance(x, torch.dtype) and x.is_floating_point else x)
        ref = op.torch_reference(*tree_map(up, sample.args), **tree_map(up, sample.kwargs))
        assert_closer(reference=(ref,), candidate=(got,), competitor=(want,), comparator=comp)


@pytest.mark.parametrize("name,tag", list(_cases()))
def test_b200_vs_torch_reference(name, tag):
    op, dt = _OPS[name], _DTYPES[tag]
    default_comp, skip, xfail = _directives(op, dt)
    if skip:
        pytest.skip(skip)
    claimed = 0
    shared = []  # samples that the reference's own torch executor fails in the same way: not ours, keep going
    for i, sample in enumerate(op.sample_inputs(DEV, dt)):
        if i >= MAX_SAMPLES:
            break
        comp = sample.comp if sample.comp is not None else default_comp
        try:
            jf, got = _run(op, [b200.ex], sample)
            claimed += len(fusions(thunder.last_traces(jf)[-1]))
            _check(op, sample, comp, got)
        except Exception as e:
            if xfail:
                pytest.xfail(f"expected failure per the reference's directive: {type(e).__name__}")
            try:
                _, got_torch = _run(op, ["torch"], sample)
                _check(op, sample, comp, got_torch)
            except Exception as e2:
                shared.append(f"sample {i}: torch executor {type(e2).__name__}, b200 {type(e).__name__}")
                continue
            raise
    torch.cuda.synchronize()
    _CLAIMED[(name, tag)] = claimed
    if shared:
        pytest.xfail("the reference's torch executor fails the same samples: " + "; ".join(shared[:3]))


def _grad_cases():
    # every op of the sweep, not only the ones the reference flags `supports_grad` (that flag selects the ops of its
    # finite-difference tests); samples the torch executor cannot differentiate are skipped at run time
    for name in sorted(_OPS):
        for tag in ("f32", "bf16"):
            if _DTYPES[tag] in _OPS[name].dtypes():
                yield pytest.param(name, tag, id=f"{name}-{tag}")


def _fwd_bwd(op, execs, sample, seed):
    """Forward + backward of one sample under an executor list: (outputs, grads w.r.t. every tensor that requires grad)."""
    jf = thunder.jit(op.op, executors=execs, cache="no caching")
    out = jf(*sample.args, **sample.kwargs)
    flat = torch.utils._pytree.tree_leaves
    outs = [o for o in flat(out) if isinstance(o, torch.Tensor) and o.requires_grad]
    leaves = [a for a in flat((sample.args, sample.kwargs)) if isinstance(a, torch.Tensor) and a.requires_grad]
    if not outs or not leaves:
        return jf, out, None
    g = torch.Generator(device=DEV).manual_seed(seed)
    cots = [torch.randn(o.shape, generator=g, device=DEV, dtype=torch.float32).to(o.dtype) for o in outs]
    return jf, out, torch.autograd.grad(outs, leaves, grad_outputs=cots, allow_unused=True)


@pytest.mark.parametrize("name,tag", list(_grad_cases()))
def test_b200_vs_torch_executor_forward_backward(name, tag):
    """Under autograd both executors run the same decomposed trace, so outputs and input gradients of the b200 regions
    must agree with the torch executor's on the reference's samples (the pattern of thunder/tests/test_grad.py's
    executor-consistency checks)."""
    op, dt = _OPS[name], _DTYPES[tag]
    default_comp, skip, _ = _directives(op, dt)
    if skip:
        pytest.skip(skip)
    ran = 0
    for i, sample in enumerate(op.sample_inputs(DEV, dt, requires_grad=True)):
        if i >= 3:
            break
        try:
            _, want_out, want_grads = _fwd_bwd(op, ["torch"], sample, i)
        except Exception:
            continue  # the reference's torch executor cannot differentiate this sample
        jf, out, grads = _fwd_bwd(op, [b200.ex], sample, i)
        comp = sample.comp if sample.comp is not None else default_comp
        comp(out, want_out, equal_nan=True)
        if want_grads is not None:
            for g, w in zip(grads, want_grads):
                if w is None:
                    assert g is None
                else:
                    comp(g, w, equal_nan=True)
        ran += 1
    torch.cuda.synchronize()
    if ran == 0:
        pytest.skip("no differentiable sample")


def test_zz_vocabulary_was_claimed():
    """Runs after the sweep: the ops below are the pointwise vocabulary b200 must fuse (not decline) for float CUDA
    tensors; a silent decline would make the sweep above a test of the torch executor."""
    if not _CLAIMED:
        pytest.skip("sweep did not run in this session")
    must = ["abs", "exp", "exp2", "expm1", "log", "log1p", "log2", "tanh", "sigmoid", "silu", "erf", "rsqrt", "sqrt", "neg", "reciprocal",
            "sin", "cos", "relu", "gelu", "add", "sub", "mul", "true_divide", "maximum", "minimum", "pow", "where", "clamp",
            "lt", "gt", "ge", "le", "eq", "ne", "square", "layer_norm", "rms_norm"]
    missing = [n for n in must if (n, "f32") in _CLAIMED and _CLAIMED[(n, "f32")] == 0]
    assert not missing, f"b200 claimed nothing for: {missing}"
    total = sum(1 for v in _CLAIMED.values() if v > 0)
    assert total >= 150, f"only {total} of {len(_CLAIMED)} (op, dtype) cases had a b200 region"
