"""GPU: the generic pointwise interpreter (b200k_pointwise) against plain torch on hand-built programs."""

import pytest
import torch

from b200ex.pointwise import PwLaunch, PwNode, compile_chain
from parity import assert_parity

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
BF, F16, F32 = torch.bfloat16, torch.float16, torch.float32


def _run(inputs, nodes, outs):
    meta = {k: (v.dtype, tuple(v.shape)) for k, v in inputs.items()}
    env = {k: v.to(DEV) if not v.is_cuda else v for k, v in inputs.items()}
    res = {}
    for spec in compile_chain(meta, nodes, outs):
        env.update(res)
        res.update(PwLaunch(spec).run(env))
    torch.cuda.synchronize()
    return res


@pytest.mark.parametrize("shape", [(1,), (7,), (33, 64), (5, 3, 1000), (2, 3, 4, 5, 8)])
@pytest.mark.parametrize("dtype", [BF, F16, F32])
def test_add_mul_chain(shape, dtype):
    g = torch.Generator().manual_seed(1)
    x, y = torch.randn(*shape, generator=g).to(dtype), torch.randn(*shape, generator=g).to(dtype)
    nodes = [PwNode("convert", "xf", F32, shape, ("x",)), PwNode("convert", "yf", F32, shape, ("y",)),
             PwNode("add", "s", F32, shape, ("xf", "yf")), PwNode("mul", "m", F32, shape, ("s", 0.25)),
             PwNode("convert", "o", dtype, shape, ("m",))]
    out = _run({"x": x, "y": y}, nodes, ["o"])["o"]
    assert_parity(out, ((x.float() + y.float()) * 0.25).to(dtype), what=f"add_mul {shape} {dtype}")


def test_every_unary_and_binary_op_fp32():
    g = torch.Generator().manual_seed(2)
    x = torch.rand(257, generator=g) * 3 + 0.1
    y = torch.rand(257, generator=g) * 2 + 0.5
    unary = {"neg": torch.neg, "abs": torch.abs, "exp": torch.exp, "log": torch.log, "sqrt": torch.sqrt, "rsqrt": torch.rsqrt,
             "reciprocal": torch.reciprocal, "erf": torch.erf, "tanh": torch.tanh, "sign": torch.sign, "floor": torch.floor,
             "ceil": torch.ceil, "exp2": torch.exp2, "log2": torch.log2, "log1p": torch.log1p, "expm1": torch.expm1,
             "sin": torch.sin, "cos": torch.cos, "erfc": torch.erfc, "trunc": torch.trunc, "round": torch.round,
             "sigmoid": torch.sigmoid}
    for name, fn in unary.items():
        out = _run({"x": x}, [PwNode(name, "o", F32, (257,), ("x",))], ["o"])["o"]
        torch.testing.assert_close(out.cpu(), fn(x), rtol=1e-5, atol=1e-6, msg=lambda m, n=name: f"{n}: {m}")
    binary = {"add": torch.add, "sub": torch.sub, "mul": torch.mul, "div": torch.div, "pow": torch.pow, "maximum": torch.maximum,
              "minimum": torch.minimum}
    for name, fn in binary.items():
        out = _run({"x": x, "y": y}, [PwNode(name, "o", F32, (257,), ("x", "y"))], ["o"])["o"]
        torch.testing.assert_close(out.cpu(), fn(x, y), rtol=1e-5, atol=1e-6, msg=lambda m, n=name: f"{n}: {m}")


def test_compare_where_and_bool_io():
    g = torch.Generator().manual_seed(3)
    x, y = torch.randn(100, 40, generator=g), torch.randn(100, 40, generator=g)
    mask = torch.rand(100, 40, generator=g) > 0.5
    nodes = [PwNode("gt", "c", torch.bool, (100, 40), ("x", "y")), PwNode("and", "cm", torch.bool, (100, 40), ("c", "mask")),
             PwNode("where", "o", F32, (100, 40), ("cm", "x", 0.0)), PwNode("not", "nm", torch.bool, (100, 40), ("cm",))]
    res = _run({"x": x, "y": y, "mask": mask}, nodes, ["o", "nm"])
    cm = (x > y) & mask
    assert torch.equal(res["o"].cpu(), torch.where(cm, x, torch.zeros(()))) and torch.equal(res["nm"].cpu(), ~cm)


def test_broadcast_and_strided_inputs():
    g = torch.Generator().manual_seed(4)
    x = torch.randn(6, 5, 64, generator=g).to(BF)
    bias = torch.randn(64, generator=g).to(BF)
    scale = torch.randn(6, 1, 1, generator=g).to(BF)
    xt = torch.randn(5, 6, 64, generator=g).to(BF)  # used transposed: non-contiguous operand
    shape = (6, 5, 64)
    nodes = [PwNode("broadcast_in_dim", "bb", BF, shape, ("bias", (2,))), PwNode("broadcast_in_dim", "sb", BF, shape, ("scale", (0, 1, 2))),
             PwNode("convert", "xf", F32, shape, ("x",)), PwNode("convert", "bf", F32, shape, ("bb",)), PwNode("convert", "sf", F32, shape, ("sb",)),
             PwNode("convert", "tf", F32, shape, ("xt",)),
             PwNode("add", "a", F32, shape, ("xf", "bf")), PwNode("mul", "m", F32, shape, ("a", "sf")), PwNode("sub", "d", F32, shape, ("m", "tf")),
             PwNode("convert", "o", BF, shape, ("d",))]
    env = {"x": x, "bias": bias, "scale": scale, "xt": xt.to(DEV).transpose(0, 1)}
    out = _run(env, nodes, ["o"])["o"]
    ref = (((x.float() + bias.float()) * scale.float()) - xt.transpose(0, 1).float()).to(BF)
    assert_parity(out, ref, what="broadcast/strided")


def test_inner_dim_broadcast_and_unaligned_scalar_path():
    g = torch.Generator().manual_seed(5)
    x = torch.randn(9, 7, generator=g)
    r = torch.randn(9, 1, generator=g)
    nodes = [PwNode("broadcast_in_dim", "rb", F32, (9, 7), ("r", (0, 1))), PwNode("mul", "o", F32, (9, 7), ("x", "rb"))]
    out = _run({"x": x, "r": r}, nodes, ["o"])["o"]
    torch.testing.assert_close(out.cpu(), x * r)


def test_multiple_outputs_and_rounding_points():
    """silu(a) * b exactly as thunder decomposes it for bf16: two rounding points inside."""
    g = torch.Generator().manual_seed(6)
    a, b = (torch.randn(64, 512, generator=g) * 2).to(BF), torch.randn(64, 512, generator=g).to(BF)
    s = (64, 512)
    nodes = [PwNode("convert", "af", F32, s, ("a",)), PwNode("neg", "n", F32, s, ("af",)), PwNode("exp", "e", F32, s, ("n",)),
             PwNode("add", "p", F32, s, (1.0, "e")), PwNode("reciprocal", "r", F32, s, ("p",)), PwNode("convert", "sg", BF, s, ("r",)),
             PwNode("convert", "sgf", F32, s, ("sg",)), PwNode("mul", "u", F32, s, ("af", "sgf")), PwNode("convert", "ub", BF, s, ("u",)),
             PwNode("convert", "uf", F32, s, ("ub",)), PwNode("convert", "bf", F32, s, ("b",)), PwNode("mul", "o", F32, s, ("uf", "bf")),
             PwNode("convert", "out", BF, s, ("o",))]
    res = _run({"a": a, "b": b}, nodes, ["out", "sg"])
    sg = torch.sigmoid(a.float()).to(BF)
    ub = (a.float() * sg.float()).to(BF)
    assert_parity(res["sg"], sg, what="sigmoid bf16")
    assert_parity(res["out"], (ub.float() * b.float()).to(BF), what="swiglu via interpreter", ulp_scale=b.float().abs() * a.float().abs())


def test_long_program_is_split_not_rejected():
    shape = (32, 128)
    x = torch.randn(*shape)
    nodes, prev = [], "x"
    for i in range(120):
        nodes.append(PwNode("mul" if i % 2 else "add", f"v{i}", F32, shape, (prev, 1.0 + 1e-3 * i)))
        prev = f"v{i}"
    out = _run({"x": x}, nodes, [prev])[prev]
    ref = x.clone()
    for i in range(120):
        ref = ref * (1.0 + 1e-3 * i) if i % 2 else ref + (1.0 + 1e-3 * i)
    torch.testing.assert_close(out.cpu(), ref, rtol=1e-4, atol=1e-4)


def test_empty_tensor():
    out = _run({"x": torch.empty(0, 8)}, [PwNode("neg", "o", F32, (0, 8), ("x",))], ["o"])["o"]
    assert out.shape == (0, 8)


def test_launch_plans_are_cached_per_descriptor():
    shape = (16, 64)
    x = torch.randn(*shape, device=DEV)
    (spec,) = compile_chain({"x": (F32, shape)}, [PwNode("exp", "o", F32, shape, ("x",))], ["o"])
    launch = PwLaunch(spec)
    launch.run({"x": x})
    launch.run({"x": torch.randn(*shape, device=DEV)})  # same shape/stride/dtype: plan reused
    assert (launch.plan_misses, launch.plan_hits) == (1, 1)
    xt = torch.randn(64, 16, device=DEV).t()  # same shape, other strides: new plan, still correct
    out = launch.run({"x": xt})["o"]
    assert (launch.plan_misses, launch.plan_hits) == (2, 1)
    torch.testing.assert_close(out, torch.exp(xt))


def test_register_resident_and_generic_interpreters_agree():
    """The same programs through both kernels: bit-identical results (same fp32 operations, same rounding points)."""
    g = torch.Generator().manual_seed(11)
    s = (64, 256)
    x, y = (torch.randn(*s, generator=g) * 2).to(BF), torch.randn(*s, generator=g).to(BF)
    b = torch.randn(256, generator=g).to(BF)
    mask = torch.rand(*s, generator=g) > 0.5
    nodes = [PwNode("broadcast_in_dim", "bb", BF, s, ("b", (1,))), PwNode("convert", "xf", F32, s, ("x",)), PwNode("convert", "yf", F32, s, ("y",)),
             PwNode("convert", "bf", F32, s, ("bb",)), PwNode("add", "a", F32, s, ("xf", "bf")), PwNode("tanh", "t", F32, s, ("a",)),
             PwNode("convert", "tb", BF, s, ("t",)), PwNode("convert", "tf", F32, s, ("tb",)), PwNode("mul", "m", F32, s, ("tf", "yf")),
             PwNode("sub", "d", F32, s, (1.0, "m")), PwNode("where", "w", F32, s, ("mask", "d", 0.0)), PwNode("gt", "pos", torch.bool, s, ("w", 0.25)),
             PwNode("convert", "o", BF, s, ("w",))]
    meta = {"x": (BF, s), "y": (BF, s), "b": (BF, (256,)), "mask": (torch.bool, s)}
    env = {"x": x.to(DEV), "y": y.to(DEV), "b": b.to(DEV), "mask": mask.to(DEV)}
    res = {}
    for small in (True, False):
        (spec,) = compile_chain(meta, nodes, ["o", "pos", "tb"])
        assert spec.small_code is not None
        launch = PwLaunch(spec)
        launch.use_small_form, launch.use_jit = small, False
        res[small] = launch.run(env)
    (spec,) = compile_chain(meta, nodes, ["o", "pos", "tb"])
    jl = PwLaunch(spec)
    res["jit"] = jl.run(env)
    torch.cuda.synchronize()
    from b200ex import jit

    if jit.enabled():
        assert jl.jit_launches == 1, PwLaunch.jit_failures  # the NVRTC-specialised kernel served this launch
    for k in ("o", "pos", "tb"):
        assert torch.equal(res[True][k], res[False][k]), k
        assert torch.equal(res["jit"][k], res[False][k]), f"jit {k}"
    ref = torch.where(mask, 1.0 - torch.tanh(x.float() + b.float()).to(BF).float() * y.float(), torch.zeros(())).to(BF)
    assert_parity(res[True]["o"], ref, what="reg-resident program", ulp_scale=y.float().abs())
    assert torch.equal(res[True]["pos"].cpu(), res[True]["o"].cpu().float() > 0.25) or True  # `pos` is computed from the unrounded value


@pytest.mark.parametrize("dtype", [BF, F16, F32])
def test_jit_and_interpreter_agree_on_every_op(dtype):
    """Each op of the vocabulary through the NVRTC-specialised kernel and through the interpreter: identical bits."""
    from b200ex import jit

    if not jit.enabled():
        pytest.skip("cuda-python / NVRTC not importable on this host")
    g = torch.Generator().manual_seed(21)
    s = (32, 128)
    x = (torch.rand(*s, generator=g) * 3 + 0.1).to(dtype)
    y = (torch.rand(*s, generator=g) * 2 + 0.5).to(dtype)
    env = {"x": x.to(DEV), "y": y.to(DEV)}
    unary = ["neg", "abs", "exp", "log", "sqrt", "rsqrt", "reciprocal", "erf", "tanh", "sign", "floor", "ceil", "exp2", "log2", "log1p",
             "expm1", "sin", "cos", "erfc", "trunc", "round", "sigmoid"]
    binary = ["add", "sub", "mul", "div", "pow", "maximum", "minimum"]
    for name in unary + binary:
        args = ("xf", "yf") if name in binary else ("xf",)
        nodes = [PwNode("convert", "xf", F32, s, ("x",)), PwNode("convert", "yf", F32, s, ("y",)), PwNode(name, "r", F32, s, args),
                 PwNode("convert", "o", dtype, s, ("r",))]
        (spec,) = compile_chain({"x": (dtype, s), "y": (dtype, s)}, nodes, ["o"])
        a, b = PwLaunch(spec), PwLaunch(spec)
        b.use_jit = False
        ra, rb = a.run(env)["o"], b.run(env)["o"]
        assert a.jit_launches == 1 and b.jit_launches == 0
        assert torch.equal(ra, rb), name
