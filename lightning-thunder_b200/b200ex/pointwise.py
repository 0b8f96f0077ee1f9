"""Compiler and launcher for generic pointwise regions (the b200k_pointwise entry point).

A fused region that contains only elementwise prims, ``convert_element_type`` and
``broadcast_in_dim`` is described by a small neutral IR (:class:`PwNode`); :func:`compile_region`
turns it into one register program per distinct output shape, :class:`PwLaunch` binds runtime
tensors (any strides; broadcasts become zero strides) and calls the C-ABI.  This is the generic
counterpart of nvFuser's pointwise scheduler behind ``FusionDefinitionWrapper.__call__``
(thunder/executors/nvfuserex_impl.py:526-576); it knows nothing about thunder itself, the executor
(b200ex/executor.py) translates bound symbols into the IR.

Numerics: every value is fp32 inside the program.  A node whose result dtype is bf16/f16 gets an
explicit rounding instruction, which reproduces the reference's torch-executor behaviour of computing
each op in fp32 and storing the declared dtype (thunder/executors/torchex.py:80-87).
"""

from __future__ import annotations

import ctypes
import hashlib
import itertools
import struct
import threading
from dataclasses import dataclass

import torch

from . import _lib, jit
from .ops import _on

_SM = {}


def _sm_count() -> int:
    d = torch.cuda.current_device()
    if d not in _SM:
        _SM[d] = _lib.lib().b200k_sm_count()
    return _SM[d]

MAX_IN, MAX_OUT, MAX_REG, MAX_CODE = 10, 6, 48, 192
SMALL_REGS, SMALL_MAX_INSTR, IMM = 8, 40, 0xFE  # register-resident form (B200K_PW_SMALL_* in include/b200k.h)

OPS = [
    "const", "mov", "add", "sub", "mul", "div", "neg", "abs", "exp", "log", "sqrt", "rsqrt", "reciprocal", "erf", "tanh", "pow",
    "maximum", "minimum", "gt", "ge", "lt", "le", "eq", "ne", "where", "rnd_bf16", "rnd_f16", "and", "or", "not", "sign",
    "floor", "ceil", "exp2", "log2", "log1p", "expm1", "sin", "cos", "erfc", "sigmoid", "trunc", "round",
]
OPCODE = {n: i for i, n in enumerate(OPS)}
BINARY = {"add", "sub", "mul", "div", "pow", "maximum", "minimum", "gt", "ge", "lt", "le", "eq", "ne", "and", "or"}
COMPARE = {"gt", "ge", "lt", "le", "eq", "ne"}
UNARY = {"neg", "abs", "exp", "log", "sqrt", "rsqrt", "reciprocal", "erf", "tanh", "not", "sign", "floor", "ceil", "exp2", "log2",
         "log1p", "expm1", "sin", "cos", "erfc", "sigmoid", "trunc", "round"}
# ops whose fp32 result is exactly representable whenever the operands are: no rounding instruction needed
EXACT = {"neg", "abs", "where", "mov", "maximum", "minimum", "sign", "gt", "ge", "lt", "le", "eq", "ne", "and", "or", "not",
         "floor", "ceil", "trunc", "round"}

DTYPE_CODE = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2, torch.bool: 3}


class PwUnsupported(Exception):
    """The region cannot be expressed as a b200k_pointwise program (the executor then declines it)."""


@dataclass(frozen=True)
class PwNode:
    op: str  # an entry of OPS, or "convert" / "broadcast_in_dim"
    out: str
    dtype: torch.dtype
    shape: tuple
    args: tuple  # value names (str) or Python numbers; broadcast_in_dim: (a, broadcast_dims)


@dataclass
class PwProgramSpec:
    iter_shape: tuple
    leaves: list  # [(input name, mapping)]  mapping[i] = iteration dim of the input's dim i
    code: list  # uint32 words
    n_reg: int
    outputs: list  # [(name, dtype, register)]
    signature: str = ""  # canonical text of the program (used to route known families to native kernels)
    small_code: list | None = None  # register-resident encoding (3 words per instruction) when the program fits it
    small_out_reg: list | None = None
    ssa: list | None = None  # (op, a, b, c, imm) in evaluation order; what b200ex.jit turns into straight-line CUDA C
    out_ids: list | None = None
    leaf_dtypes: list | None = None
    # leaf index -> (minval, maxval) for leaves that are counter-based random tensors spanning the whole iteration
    # space: the specialised kernels generate them in registers instead of loading them (see PwLaunch.gen_info)
    generated: dict | None = None


class _Operand(ctypes.Structure):
    _fields_ = [("ptr", ctypes.c_void_p), ("s0", ctypes.c_int64), ("s1", ctypes.c_int64), ("inner", ctypes.c_int32), ("dtype", ctypes.c_int32)]


class _Program(ctypes.Structure):
    _fields_ = [
        ("d0", ctypes.c_int64), ("d1", ctypes.c_int64), ("d2", ctypes.c_int64),
        ("n_in", ctypes.c_int32), ("n_out", ctypes.c_int32), ("n_code", ctypes.c_int32), ("n_reg", ctypes.c_int32),
        ("inp", _Operand * MAX_IN), ("out", _Operand * MAX_OUT), ("out_reg", ctypes.c_int32 * MAX_OUT), ("code", ctypes.c_uint32 * MAX_CODE),
        ("n_small", ctypes.c_int32), ("small_out_reg", ctypes.c_int32 * MAX_OUT), ("small_code", ctypes.c_uint32 * (3 * SMALL_MAX_INSTR)),
    ]


def _f32_bits(x: float) -> int:
    x = float(x)
    try:
        return struct.unpack("<I", struct.pack("<f", x))[0]
    except OverflowError:  # |x| beyond fp32 range rounds to infinity, as a float32 cast would
        return 0x7F800000 if x > 0 else 0xFF800000


# --------------------------------------------------------------------------------------------------
# Compilation
# --------------------------------------------------------------------------------------------------
def compile_region(inputs: dict, nodes: list, outputs: list, generated: dict | None = None) -> list:
    """inputs: name -> (dtype, shape); nodes in execution order; outputs: value names; generated: input name ->
    (minval, maxval) for inputs that are uniform_philox tensors (b200k_uniform_philox convention).
    Returns one :class:`PwProgramSpec` per distinct output shape.  Raises :class:`PwUnsupported`."""
    producer = {n.out: n for n in nodes}
    meta = {k: v for k, v in inputs.items()}
    for n in nodes:
        meta[n.out] = (n.dtype, tuple(n.shape))
    for name, (dt, _) in meta.items():
        if dt not in DTYPE_CODE:
            raise PwUnsupported(f"dtype {dt} of {name}")
    groups: dict[tuple, list] = {}
    for o in outputs:
        if o not in meta:
            raise PwUnsupported(f"unknown output {o}")
        groups.setdefault(tuple(meta[o][1]), []).append(o)
    specs = []
    for shape, outs in groups.items():
        # canonical output order: the permutation with the smallest signature (programs with up to 3 outputs),
        # so that the same dataflow always yields the same signature regardless of proxy naming
        best = None
        perms = itertools.permutations(outs) if len(outs) <= 3 else [tuple(outs)]
        for perm in perms:
            spec = _compile_group(shape, list(perm), inputs, producer, meta, generated or {})
            if best is None or spec.signature < best.signature:
                best = spec
        specs.append(best)
    return specs


def compile_chain(inputs: dict, nodes: list, outputs: list, generated: dict | None = None) -> list:
    """Like :func:`compile_region`, but a region too large for one program (registers, code words, operand
    slots) is cut into consecutive programs that hand intermediates over through memory.  The returned
    specs must be launched in order."""
    try:
        return compile_region(inputs, nodes, outputs, generated)
    except PwUnsupported:
        if len(nodes) <= 1:
            raise
    mid = len(nodes) // 2
    first, second = nodes[:mid], nodes[mid:]
    made_first = {n.out for n in first}
    made_second = {n.out for n in second}
    used_by_second = set()
    for n in second:
        args = n.args[:1] if n.op == "broadcast_in_dim" else n.args
        used_by_second.update(a for a in args if isinstance(a, str))
    outs_first = [n.out for n in first if n.out in used_by_second or n.out in outputs]
    meta_first = {n.out: (n.dtype, tuple(n.shape)) for n in first}
    inputs_second = dict(inputs)
    inputs_second.update({o: meta_first[o] for o in outs_first})
    outs_second = [o for o in outputs if o in made_second]
    specs = compile_chain(inputs, first, outs_first, generated) if outs_first else []
    if outs_second:
        specs += compile_chain(inputs_second, second, outs_second, generated)
    return specs


def _compile_group(iter_shape, outs, inputs, producer, meta, generated) -> PwProgramSpec:
    ssa = []  # (op, a, b, c, imm)  operands are ssa ids
    memo: dict = {}
    leaves: list = []
    leaf_ids: dict = {}
    consts: dict = {}

    cse: dict = {}

    def emit(op, a=-1, b=-1, c=-1, imm=None):
        # common-subexpression elimination inside the region (the reference does it with nvFuserExecutor.cse,
        # nvfuserex_impl.py:743-833): the same op on the same values is computed once
        key = (op, a, b, c, imm)
        if op != "leaf" and key in cse:
            return cse[key]
        ssa.append(key)
        if op != "leaf":
            cse[key] = len(ssa) - 1
        return len(ssa) - 1

    def const(x):
        key = _f32_bits(x)
        if key not in consts:
            consts[key] = emit("const", imm=key)
        return consts[key]

    def value(name, mapping):
        key = (name, mapping)
        if key in memo:
            return memo[key]
        if name in inputs:
            if key not in leaf_ids:
                leaf_ids[key] = emit("leaf", imm=len(leaves))
                leaves.append((name, mapping))
            memo[key] = leaf_ids[key]
            return memo[key]
        n = producer.get(name)
        if n is None:
            raise PwUnsupported(f"value {name} has no producer in the region")
        memo[key] = r = node(n, mapping)
        return r

    def arg(a, mapping):
        if isinstance(a, str):
            return value(a, mapping)
        if isinstance(a, (bool, int, float)):
            return const(float(a))
        raise PwUnsupported(f"operand {a!r}")

    def as_dtype(x, dt):
        """A Python number as ATen sees it next to a tensor of dtype dt (rounded to 16-bit storage types)."""
        if isinstance(x, (int, float)) and not isinstance(x, bool) and dt in (torch.bfloat16, torch.float16):
            return float(torch.tensor(float(x), dtype=dt))
        return x

    def rounded(r, n):
        if n.op in EXACT:
            return r
        if n.dtype == torch.bfloat16:
            return emit("rnd_bf16", r)
        if n.dtype == torch.float16:
            return emit("rnd_f16", r)
        return r

    def node(n: PwNode, mapping):
        if n.op == "broadcast_in_dim":
            a, bdims = n.args
            return value(a, tuple(mapping[d] for d in bdims))
        if n.op == "convert":
            r = arg(n.args[0], mapping)
            src_dt = meta[n.args[0]][0] if isinstance(n.args[0], str) else torch.float32
            if n.dtype == torch.bfloat16 and src_dt != torch.bfloat16:
                return emit("rnd_bf16", r)
            if n.dtype == torch.float16 and src_dt != torch.float16:
                return emit("rnd_f16", r)
            if n.dtype == torch.bool and src_dt != torch.bool:
                return emit("ne", r, const(0.0))
            return r
        if n.op == "mov":
            return arg(n.args[0], mapping)
        if n.op == "where":
            c = arg(n.args[0], mapping)
            a, b = (arg(as_dtype(x, n.dtype), mapping) for x in n.args[1:])
            return emit("where", c, a, b)
        if n.op == "pow" and isinstance(n.args[1], (int, float)) and not isinstance(n.args[1], bool):
            # ATen evaluates small constant exponents by multiplication / sqrt (pow_tensor_scalar_optimized_kernel)
            a, e = arg(n.args[0], mapping), float(n.args[1])
            if e == 2.0:
                return rounded(emit("mul", a, a), n)
            if e == 3.0:
                return rounded(emit("mul", emit("mul", a, a), a), n)
            if e == 0.5:
                return rounded(emit("sqrt", a), n)
            if e == -0.5:
                return rounded(emit("rsqrt", a), n)
            if e == -1.0:
                return rounded(emit("reciprocal", a), n)
            if e == 1.0:
                return a
            return rounded(emit("pow", a, const(e)), n)
        if n.op in BINARY:
            x, y = n.args[0], n.args[1]
            if n.op in COMPARE:
                # ATen compares in the tensor operand's dtype: a Python number is cast to it first (a wrapped scalar does
                # not take part in type promotion), so `u_bf16 < 0.9` tests against bf16(0.9).  Arithmetic ops get their
                # scalar in fp32 (opmath), which is what the fp32 constant already is.
                dts = [meta[v][0] for v in (x, y) if isinstance(v, str)]
                if len(dts) == 1:
                    x, y = as_dtype(x, dts[0]), as_dtype(y, dts[0])
            a, b = arg(x, mapping), arg(y, mapping)
            return rounded(emit(n.op, a, b), n)
        if n.op in UNARY:
            return rounded(emit(n.op, arg(n.args[0], mapping)), n)
        raise PwUnsupported(f"op {n.op}")

    ident = tuple(range(len(iter_shape)))
    out_ids = [value(o, ident) for o in outs]

    # canonical signature: the SSA text with leaves numbered in first-use order (input names erased)
    sig = ";".join(f"{op}({a},{b},{c},{imm if op != 'leaf' else ''})" for op, a, b, c, imm in ssa)
    sig += "|" + ",".join(f"{oid}:{meta[o][0]}" for oid, o in zip(out_ids, outs))
    # leaf layout relative to the iteration space, rank independent: "*" = spans every dim, else the dims it
    # covers counted from the innermost one
    nd = len(iter_shape)
    sig += "|" + ",".join(f"{meta[nm][0]}:" + ("*" if tuple(mp) == ident else "/".join(str(d - nd) for d in mp)) for nm, mp in leaves)
    # random leaves that cover the iteration space one to one can be produced inside the kernel
    gen = {i: tuple(float(v) for v in generated[nm]) for i, (nm, mp) in enumerate(leaves)
           if nm in generated and tuple(mp) == ident and tuple(meta[nm][1]) == tuple(iter_shape)}
    if gen:
        sig += "|gen:" + ",".join(f"{i}:{lo!r}:{hi!r}" for i, (lo, hi) in sorted(gen.items()))

    # register allocation: leaves are preloaded into registers 0..n_in-1, everything else linear scan
    if len(leaves) > MAX_IN or len(outs) > MAX_OUT:
        raise PwUnsupported("too many operands for one pointwise program")
    last_use = {}
    for i, (op, a, b, c, _) in enumerate(ssa):
        for x in (a, b, c):
            if x >= 0:
                last_use[x] = i
    for oid in out_ids:
        last_use[oid] = len(ssa)
    reg_of, free, n_reg = {}, [], len(leaves)
    code: list = []
    for i, (op, a, b, c, imm) in enumerate(ssa):
        if op == "leaf":
            reg_of[i] = imm
        srcs = [reg_of[x] for x in (a, b, c) if x >= 0]
        # operands dying here can be reused for the destination (reads happen before the write per element)
        for x in {a, b, c}:
            if x >= 0 and last_use.get(x) == i and x not in out_ids:
                free.append(reg_of[x])
        if op == "leaf":
            continue
        if i not in last_use:
            # dead value (can happen through memoised broadcasts); still needs a scratch destination
            pass
        if free:
            dst = free.pop()
        else:
            dst = n_reg
            n_reg += 1
        reg_of[i] = dst
        if n_reg > MAX_REG:
            raise PwUnsupported("pointwise program needs too many registers")
        if op == "const":
            code += [OPCODE["const"] | dst << 8, imm]
        elif op == "where":
            code += [OPCODE["where"] | dst << 8 | srcs[0] << 16 | srcs[1] << 24, srcs[2]]
        elif op in BINARY:
            code.append(OPCODE[op] | dst << 8 | srcs[0] << 16 | srcs[1] << 24)
        else:
            code.append(OPCODE[op] | dst << 8 | srcs[0] << 16)
    if len(code) > MAX_CODE:
        raise PwUnsupported("pointwise program too long")
    # two outputs may not share a register with a later-written value: outputs were excluded from reuse above
    outputs = [(o, meta[o][0], reg_of[oid]) for o, oid in zip(outs, out_ids)]
    spec = PwProgramSpec(tuple(iter_shape), leaves, code, max(n_reg, 1), outputs, hashlib.sha1(sig.encode()).hexdigest())
    spec.ssa, spec.out_ids, spec.leaf_dtypes = ssa, out_ids, [meta[nm][0] for nm, _ in leaves]
    spec.generated = gen or None
    small = _encode_small(ssa, len(leaves), out_ids)
    if small is not None:
        spec.small_code, spec.small_out_reg = small
    return spec


def _encode_small(ssa, n_leaves, out_ids):
    """Second encoding for the register-resident interpreter: constants become immediates, at most SMALL_REGS values
    live at once.  Returns (words, out registers) or None when the program does not fit."""
    if n_leaves > SMALL_REGS:
        return None
    last_use = {}
    for i, (op, a, b, c, _) in enumerate(ssa):
        for x in (a, b, c):
            if x >= 0:
                last_use[x] = i
    for oid in out_ids:
        if ssa[oid][0] == "const":
            return None
        last_use[oid] = len(ssa)
    reg_of, free, n_reg, words = {}, [], n_leaves, []
    n_instr = 0
    for i, (op, a, b, c, imm) in enumerate(ssa):
        if op == "const":
            continue
        if op == "leaf":
            reg_of[i] = imm
            continue
        ops_ = [x for x in (a, b, c) if x >= 0]
        consts = {ssa[x][4] for x in ops_ if ssa[x][0] == "const"}
        if len(consts) > 1:
            return None  # one immediate slot per instruction
        immw = consts.pop() if consts else 0
        enc = lambda x: IMM if ssa[x][0] == "const" else reg_of[x]
        srcs = [enc(x) for x in ops_]
        for x in set(ops_):
            if ssa[x][0] != "const" and last_use.get(x) == i and x not in out_ids:
                free.append(reg_of[x])
        if free:
            dst = free.pop()
        else:
            dst = n_reg
            n_reg += 1
        if n_reg > SMALL_REGS:
            return None
        reg_of[i] = dst
        ra = srcs[0]
        rb = srcs[1] if len(srcs) > 1 else 0
        rc = srcs[2] if len(srcs) > 2 else 0
        words += [OPCODE[op] | dst << 8 | ra << 16 | rb << 24, rc, immw]
        n_instr += 1
        if n_instr > SMALL_MAX_INSTR:
            return None
    return words, [reg_of[oid] for oid in out_ids]


# --------------------------------------------------------------------------------------------------
# Launch
# --------------------------------------------------------------------------------------------------
def _collapse(shape, strides_list):
    """Merge adjacent iteration dims that every operand traverses contiguously (or broadcasts in both).
    Returns (shape, strides_list) with size-1 dims dropped."""
    dims = [d for d in range(len(shape)) if shape[d] != 1]
    if not dims:
        return [1], [[0] for _ in strides_list]
    out_shape = [shape[dims[0]]]
    out_strides = [[s[dims[0]]] for s in strides_list]
    for d in dims[1:]:
        n = shape[d]
        if all(os[-1] == s[d] * n for os, s in zip(out_strides, strides_list)):
            out_shape[-1] *= n
            for os, s in zip(out_strides, strides_list):
                os[-1] = s[d]
        else:
            out_shape.append(n)
            for os, s in zip(out_strides, strides_list):
                os.append(s[d])
    return out_shape, out_strides


class _Plan:
    """Everything about one launch that depends only on the operands' (shape, stride, dtype): the collapsed
    iteration box, per-operand strides, which operands need a contiguous copy first, and a prefilled program
    struct whose pointers are patched per call."""

    __slots__ = ("P", "contig_fix", "lock", "jit", "jit_params", "items")

    def __init__(self, P, contig_fix):
        self.P, self.contig_fix, self.lock = P, contig_fix, threading.Lock()
        self.jit = None         # (inner flags, one_d) when the NVRTC-specialised kernel can serve this descriptor
        self.jit_params = None  # jit.Params mirror of P (dims + operands)
        self.items = 0


class PwLaunch:
    """A compiled program bound to its region: ``run(env)`` reads input tensors from ``env`` by name and returns
    {output name: new contiguous tensor}.  Launch plans are cached per operand descriptor — the analogue of the
    reference's ``lru_cache`` keyed on ``to_runtime_descriptors`` (nvfuserex_impl.py:505-521, 622-625), needed
    because trace proxies carry no strides."""

    MAX_PLANS = 2048
    use_small_form = True  # tests flip this to compare the register-resident and the generic interpreter
    use_jit = True         # False: always interpret (b200k_pointwise), never specialise with NVRTC
    force_interpreter = False  # tests: keep non-vectorisable launches on the interpreter instead of the gather kernel
    jit_failures: list = []

    def __init__(self, spec: PwProgramSpec, gen_info: dict | None = None):
        self.spec = spec
        # name -> (shape, dtype, minval, maxval, seed, offset) for leaves that are uniform_philox tensors; seed / offset
        # are ints or names of host integers in `env`.  Such a leaf is generated inside the specialised kernel when the
        # spec allows it (spec.generated) and that kernel serves the launch; otherwise it is materialised first.
        self.gen_info = gen_info or {}
        self._plans: dict = {}
        self.plan_hits = 0
        self.plan_misses = 0
        self.jit_launches = 0
        self.gather_launches = 0
        self._gather = None

    @staticmethod
    def _leaf_strides(t: torch.Tensor, mapping, ndim):
        st = [0] * ndim
        for i, d in enumerate(mapping):
            if t.shape[i] != 1:
                st[d] = t.stride(i)
        return st

    def _build_plan(self, tensors) -> _Plan:
        spec = self.spec
        shape = list(spec.iter_shape)
        nd = len(shape)
        contig = [0] * nd
        acc = 1
        for d in range(nd - 1, -1, -1):
            contig[d] = acc
            acc *= shape[d]
        in_strides = [self._leaf_strides(t, m, nd) if t is not None else list(contig) for t, (_, m) in zip(tensors, spec.leaves)]
        cshape, cstrides = _collapse(shape, in_strides + [contig])
        contig_fix = False
        if len(cshape) > 3:
            # more than three independent strides: every operand gets the contiguous broadcast layout first
            # (a strided copy by torch; rare — needs e.g. several differently permuted inputs)
            contig_fix = True
            cshape, cstrides = _collapse(shape, [list(contig) for _ in tensors] + [contig])
        while len(cshape) < 3:
            cshape.insert(0, 1)
            for st in cstrides:
                st.insert(0, 0)
        P = _Program()
        P.d0, P.d1, P.d2 = cshape
        P.n_in, P.n_out, P.n_code, P.n_reg = len(tensors), len(spec.outputs), len(spec.code), spec.n_reg
        for k, (t, st) in enumerate(zip(tensors, cstrides[:-1])):
            P.inp[k] = _Operand(0, st[0], st[1], st[2], DTYPE_CODE[t.dtype if t is not None else spec.leaf_dtypes[k]])
        so = [cshape[1] * cshape[2], cshape[2], 1]  # outputs are freshly allocated, contiguous over the collapsed box
        for k, (name, dt, reg) in enumerate(spec.outputs):
            P.out[k] = _Operand(0, so[0], so[1], so[2], DTYPE_CODE[dt])
            P.out_reg[k] = reg
        for k, w in enumerate(spec.code):
            P.code[k] = w
        P.n_small = 0
        if spec.small_code is not None and self.use_small_form:
            P.n_small = len(spec.small_code) // 3
            for k, w in enumerate(spec.small_code):
                P.small_code[k] = w
            for k, r in enumerate(spec.small_out_reg):
                P.small_out_reg[k] = r
        plan = _Plan(P, contig_fix)
        self._plan_jit(plan, tensors if not contig_fix else None, cshape, cstrides)
        return plan

    def _plan_jit(self, plan, tensors, cshape, cstrides):
        """Decide whether this descriptor can use the run-time specialised kernel (b200ex.jit): 8 elements per
        work item, so the inner extent must be a multiple of 8 and unit-stride operands 16-byte addressable."""
        spec = self.spec
        if not (self.use_jit and jit.enabled()) or cshape[2] % 8 != 0 or len(spec.leaves) > jit.MAX_IN:
            return
        ins = cstrides[:-1]
        for st in ins:
            if st[2] not in (0, 1) or (st[2] == 1 and (st[0] % 8 or st[1] % 8)):
                return
        if tensors is not None and any(t.data_ptr() % 16 for t, st in zip(tensors, ins) if st[2] == 1 and t is not None):
            return
        if spec.generated and len(spec.generated) > jit.MAX_GEN:
            return
        JP = jit.Params()
        JP.d0, JP.d1, JP.d2 = cshape
        for k in range(len(spec.leaves)):
            o = plan.P.inp[k]
            JP.inp[k] = jit._Operand(0, o.s0, o.s1, o.inner, o.dtype)
        for k in range(len(spec.outputs)):
            o = plan.P.out[k]
            JP.out[k] = jit._Operand(0, o.s0, o.s1, o.inner, o.dtype)
        plan.jit = (tuple(st[2] == 1 for st in ins), cshape[0] * cshape[1] == 1)
        plan.jit_params = JP
        plan.items = cshape[0] * cshape[1] * (cshape[2] // 8)

    def _materialise(self, name: str, env: dict, dev) -> torch.Tensor:
        from . import ops

        shape, dt, lo, hi, seed, offset = self.gen_info[name]
        val = lambda v: env[v] if isinstance(v, str) else v
        return ops.uniform_philox(shape, lo, hi, device=dev, dtype=dt, seed=val(seed), offset=val(offset))

    def run(self, env: dict, _in_kernel_rng: bool = True) -> dict:
        spec = self.spec
        shape = spec.iter_shape
        in_kernel = spec.generated if (spec.generated and _in_kernel_rng and self.use_jit and jit.enabled()) else {}
        tensors = []
        dev = None
        for idx, (name, mapping) in enumerate(spec.leaves):
            if idx in in_kernel:
                tensors.append(None)
                continue
            t = env.get(name)
            if t is None and name in self.gen_info:
                if dev is None:
                    dev = next((v.device for v in env.values() if isinstance(v, torch.Tensor) and v.is_cuda), torch.device("cuda", torch.cuda.current_device()))
                t = env[name] = self._materialise(name, env, dev)  # kept in env: other programs of the region reuse it
            if not isinstance(t, torch.Tensor) or not t.is_cuda:
                raise RuntimeError(f"b200ex pointwise: operand {name} is not a CUDA tensor")
            if t.dim() != len(mapping):
                raise RuntimeError(f"b200ex pointwise: operand {name} has rank {t.dim()}, expected {len(mapping)}")
            if t.dtype not in DTYPE_CODE:
                # The reference's traces are not always dtype-consistent: the backward of `amax` hands an int64 tie
                # count to a proxy declared bool.  Integers enter the program as the fp32 values they are.
                if t.dtype in (torch.int64, torch.int32, torch.int16, torch.int8, torch.uint8):
                    t = t.to(torch.float32)
                else:
                    raise RuntimeError(f"b200ex pointwise: operand {name} has unsupported dtype {t.dtype}")
            tensors.append(t)
            dev = dev or t.device
        if dev is None:
            dev = next((v.device for v in env.values() if isinstance(v, torch.Tensor) and v.is_cuda), torch.device("cuda", torch.cuda.current_device()))
        outs = {name: torch.empty(shape, dtype=dt, device=dev) for name, dt, _ in spec.outputs}
        if 0 in shape:
            return outs
        # pointers enter the key only through their alignment class (vector path needs 16 B / 8 B alignment)
        key = tuple((t.shape, t.stride(), t.dtype, t.data_ptr() % 16 == 0) if t is not None else None for t in tensors)
        plan = self._plans.get(key)
        if plan is None:
            self.plan_misses += 1
            plan = self._build_plan(tensors)
            if len(self._plans) < self.MAX_PLANS:
                self._plans[key] = plan
        else:
            self.plan_hits += 1
        if in_kernel and (plan.jit is None or plan.contig_fix):
            return self.run(env, _in_kernel_rng=False)  # this descriptor is not served by a specialised kernel
        if plan.jit is None and not in_kernel and jit.enabled() and self.use_jit and not self.force_interpreter:
            # not vectorisable by the skeleton above (inner extent, strides, alignment): the region compiler's gather
            # kernel reads such operands with narrower or element-wise loads; the interpreter below is only for hosts
            # without NVRTC
            if self._gather is None:
                from . import regionc

                self._gather = regionc.RegionLaunch(regionc.Launch(regionc.program_from_pointwise(spec), {}))
            self.gather_launches += 1
            return self._gather.run({name: t for (name, _), t in zip(spec.leaves, tensors)})
        if plan.contig_fix:
            nd = len(shape)
            fixed = []
            for t, (_, m) in zip(tensors, spec.leaves):
                full = [1] * nd
                for i, d in enumerate(m):
                    full[d] = t.shape[i]
                fixed.append(t.reshape(full).expand(shape).contiguous())
            tensors = fixed
        if plan.jit is not None:
            try:
                kernel = jit.get_kernel(spec, plan.jit[0], plan.jit[1], dev.index, bool(in_kernel))
            except jit.JitUnavailable as e:
                plan.jit = None  # this host cannot build it: the interpreter below serves the descriptor instead
                PwLaunch.jit_failures.append(str(e)[:400])
                if in_kernel:
                    return self.run(env, _in_kernel_rng=False)
            else:
                with plan.lock:
                    JP = plan.jit_params
                    for k, t in enumerate(tensors):
                        JP.inp[k].ptr = t.data_ptr() if t is not None else 0
                    for slot, idx in enumerate(sorted(in_kernel)):
                        _, _, _, _, seed, offset = self.gen_info[spec.leaves[idx][0]]
                        JP.sc[2 * slot] = ctypes.c_int64((env[seed] if isinstance(seed, str) else seed) & 0xFFFFFFFFFFFFFFFF).value
                        JP.sc[2 * slot + 1] = ctypes.c_int64((env[offset] if isinstance(offset, str) else offset) & 0xFFFFFFFFFFFFFFFF).value
                    for k, (name, _, _) in enumerate(spec.outputs):
                        JP.out[k].ptr = outs[name].data_ptr()
                    with _on(dev):
                        kernel.launch(JP, plan.items, _sm_count(), torch.cuda.current_stream(dev).cuda_stream)
                self.jit_launches += 1
                return outs
        with plan.lock:  # the struct is shared between calls; the launch copies it before returning
            P = plan.P
            for k, t in enumerate(tensors):
                P.inp[k].ptr = t.data_ptr()
            for k, (name, _, _) in enumerate(spec.outputs):
                P.out[k].ptr = outs[name].data_ptr()
            with _on(dev):
                rc = _lib.lib().b200k_pointwise(ctypes.byref(P), torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(rc, "b200k_pointwise")
        return outs
