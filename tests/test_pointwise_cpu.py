"""CPU: the pointwise-program compiler (no thunder, no GPU): code generation, register allocation limits,
rounding insertion and the canonical signature."""

import pytest
import torch

from b200ex.pointwise import MAX_REG, OPCODE, PwNode, PwUnsupported, _collapse, compile_region

BF, F32 = torch.bfloat16, torch.float32


def _ops(spec):
    out, i = [], 0
    names = {v: k for k, v in OPCODE.items()}
    while i < len(spec.code):
        op = names[spec.code[i] & 0xFF]
        out.append(op)
        i += 2 if op in ("const", "where") else 1
    return out


def test_add_bf16_inserts_one_rounding():
    nodes = [PwNode("convert", "xf", F32, (4, 8), ("x",)), PwNode("convert", "yf", F32, (4, 8), ("y",)),
             PwNode("add", "s", F32, (4, 8), ("xf", "yf")), PwNode("convert", "o", BF, (4, 8), ("s",))]
    (spec,) = compile_region({"x": (BF, (4, 8)), "y": (BF, (4, 8))}, nodes, ["o"])
    assert _ops(spec) == ["add", "rnd_bf16"] and len(spec.leaves) == 2 and spec.iter_shape == (4, 8)


def test_native_16bit_op_is_rounded_but_exact_ops_are_not():
    nodes = [PwNode("exp", "e", BF, (8,), ("x",)), PwNode("neg", "n", BF, (8,), ("e",))]
    (spec,) = compile_region({"x": (BF, (8,))}, nodes, ["n"])
    assert _ops(spec) == ["exp", "rnd_bf16", "neg"]


def test_broadcast_becomes_a_leaf_mapping_and_scalars_become_constants():
    nodes = [PwNode("broadcast_in_dim", "bb", F32, (4, 8), ("b", (1,))), PwNode("add", "s", F32, (4, 8), ("x", "bb")),
             PwNode("mul", "o", F32, (4, 8), ("s", 0.5))]
    (spec,) = compile_region({"x": (F32, (4, 8)), "b": (F32, (8,))}, nodes, ["o"])
    assert dict(spec.leaves) == {"x": (0, 1), "b": (1,)}
    assert _ops(spec) == ["add", "const", "mul"]


def test_pow_with_small_constant_exponent_is_strength_reduced():
    (spec,) = compile_region({"x": (F32, (8,))}, [PwNode("pow", "o", F32, (8,), ("x", 3.0))], ["o"])
    assert _ops(spec) == ["mul", "mul"]
    (spec,) = compile_region({"x": (F32, (8,))}, [PwNode("pow", "o", F32, (8,), ("x", 2.5))], ["o"])
    assert _ops(spec) == ["const", "pow"]


def test_outputs_of_different_shape_get_separate_programs():
    nodes = [PwNode("mul", "a", F32, (4, 1), ("r", 2.0)), PwNode("broadcast_in_dim", "ab", F32, (4, 8), ("a", (0, 1))),
             PwNode("mul", "o", F32, (4, 8), ("x", "ab"))]
    specs = compile_region({"r": (F32, (4, 1)), "x": (F32, (4, 8))}, nodes, ["a", "o"])
    assert sorted(s.iter_shape for s in specs) == [(4, 1), (4, 8)]


def test_signature_is_independent_of_names_rank_and_output_order():
    def prog(names, shape, outs):
        x, y, s, d = names
        nodes = [PwNode("add", s, F32, shape, (x, y)), PwNode("sub", d, F32, shape, (x, y))]
        return compile_region({x: (F32, shape), y: (F32, shape)}, nodes, outs)[0]

    a = prog(("x", "y", "s", "d"), (4, 8), ["s", "d"])
    b = prog(("p", "q", "t1", "t2"), (2, 3, 8), ["t2", "t1"])
    assert a.signature == b.signature
    c = compile_region({"x": (BF, (4, 8)), "y": (BF, (4, 8))}, [PwNode("add", "s", BF, (4, 8), ("x", "y"))], ["s"])[0]
    assert c.signature != a.signature


def test_unsupported_dtype_and_register_pressure():
    with pytest.raises(PwUnsupported):
        compile_region({"x": (torch.int64, (4,))}, [PwNode("add", "o", torch.int64, (4,), ("x", 1))], ["o"])
    # a chain that keeps more than MAX_REG values alive
    n = MAX_REG + 8
    nodes = [PwNode("mul", f"m{i}", F32, (4,), ("x", float(i + 2))) for i in range(n)]
    acc = "m0"
    for i in range(1, n):
        nodes.append(PwNode("add", f"a{i}", F32, (4,), (acc, f"m{i}")))
        acc = f"a{i}"
    outs = [f"m{i}" for i in range(0, n, 9)][:6]
    with pytest.raises(PwUnsupported):
        compile_region({"x": (F32, (4,))}, nodes, [acc] + [f"m{i}" for i in range(5)] + [f"m{i}" for i in range(5, 11)])


def test_collapse_merges_contiguous_dims():
    shape, strides = _collapse([2, 3, 4], [[12, 4, 1], [0, 0, 1], [12, 4, 1]])
    assert shape == [6, 4] and strides == [[4, 1], [0, 1], [4, 1]]
    shape, strides = _collapse([2, 3, 4], [[12, 4, 1], [12, 4, 1]])
    assert shape == [24] and strides == [[1], [1]]
    shape, strides = _collapse([1, 1], [[0, 0]])
    assert shape == [1]


def test_register_resident_encoding():
    nodes = [PwNode("convert", "xf", F32, (4, 8), ("x",)), PwNode("div", "d", F32, (4, 8), ("xf", 8.0)), PwNode("convert", "o", BF, (4, 8), ("d",))]
    (spec,) = compile_region({"x": (BF, (4, 8))}, nodes, ["o"])
    assert spec.small_code is not None and len(spec.small_code) == 6  # div with an immediate, rounding; no const instruction
    assert spec.small_code[0] >> 24 == 0xFE and spec.small_out_reg == [spec.small_code[3] >> 8 & 0xFF]
    # more than 8 values alive at once: only the generic form exists
    nodes = [PwNode("mul", f"m{i}", F32, (4,), ("x", float(i + 2))) for i in range(10)]
    acc = "m0"
    for i in range(1, 10):
        nodes.append(PwNode("add", f"a{i}", F32, (4,), (acc, f"m{i}")))
        acc = f"a{i}"
    nodes2 = nodes + [PwNode("mul", "z", F32, (4,), (acc, "m0"))]
    outs = ["z"] + [f"m{i}" for i in range(1, 6)]
    (spec,) = compile_region({"x": (F32, (4,))}, nodes2, outs)
    assert spec.small_code is None and spec.code


def test_jit_source_generation_and_nvrtc_compile():
    """The run-time specialised path: straight-line CUDA C for a program, compiled by NVRTC for sm_100a (no GPU needed)."""
    from b200ex import jit

    nodes = [PwNode("broadcast_in_dim", "bb", BF, (4, 64), ("b", (1,))), PwNode("convert", "xf", F32, (4, 64), ("x",)),
             PwNode("convert", "bf", F32, (4, 64), ("bb",)), PwNode("add", "a", F32, (4, 64), ("xf", "bf")),
             PwNode("gt", "m", torch.bool, (4, 64), ("a", 0.0)), PwNode("where", "w", F32, (4, 64), ("m", "a", 0.0)),
             PwNode("convert", "o", BF, (4, 64), ("w",))]
    (spec,) = compile_region({"x": (BF, (4, 64)), "b": (BF, (64,))}, nodes, ["o", "m"])
    assert spec.ssa and spec.out_ids and spec.leaf_dtypes == [BF, BF]
    src = jit.generate_source(spec, (True, False), False, jit.unroll_for(2))
    assert "pw_kernel" in src and "__fadd_rn" in src and "st_bool" in src and src.count("ld_bf16(P.in[0]") == jit.unroll_for(2)
    if not jit.enabled():
        pytest.skip("cuda-python / NVRTC not importable here")
    cubin = jit.compile_cubin(src)
    assert cubin[:4] == b"\x7fELF"
    with pytest.raises(jit.JitUnavailable):
        jit.compile_cubin("this is not CUDA")
