"""Region compiler: index maps, row reductions and column sums in one generated kernel.

``b200ex.pointwise`` handles regions that are purely elementwise.  The regions the reference's own models
produce are wider: the LitGPT attention slice (thunder/benchmarks/__init__.py:2282-2328) is ``reshape`` /
``transpose`` / ``split`` / ``expand`` / ``slice`` / ``cat`` around a few multiplies, its backward adds ``pad`` and a
``sum`` over the query-group dimension, and a norm spelled with torch ops (thunder/tests/llama2_model.py:22-35) is a
row reduction between two elementwise halves plus, in the backward, a column reduction for the weight gradient.
nvFuser claims all of these (thunder/executors/nvfuserex_impl.py:1229-1401 layout prims, :2132-2244 reductions).
This module is their b200 lowering:

* **Index maps instead of copies.**  Every value is demanded at *coordinates* that are affine forms over the kernel's
  iteration variables.  ``transpose`` permutes them, ``slice`` scales and shifts, ``broadcast_in_dim`` drops, ``pad``
  and ``cat`` shift and add a range predicate, ``reshape`` re-digitises the linear index (splitting an iteration
  variable when a digit boundary falls inside it), a ``sum`` over a small extent unrolls.  What reaches memory is a
  set of *gather leaves* — (input, coordinates, guard) — so layout prims cost no bytes, exactly as in the oracle
  where they are views.
* **Row kernels.**  Reductions over the trailing dimensions make the kernel row-persistent: each row lives in the
  registers of one warp or one CTA, reductions are warp shuffles plus one shared-memory exchange, and values on the
  far side of a reduction are recomputed from the resident leaves rather than stored.  Sums over all leading
  dimensions (weight / bias gradients) accumulate per CTA in registers across the rows it owns and are folded by a
  second tiny kernel from fp32 partials (deterministic order).
* **Code generation.**  Straight-line CUDA C with shapes, index arithmetic and the SSA program baked in, compiled with
  NVRTC for sm_100a (``b200ex.jit.compile_cubin``), W consecutive elements of the innermost variable per work item
  with 16-byte loads and stores wherever the runtime strides and alignment allow.

The IR (:class:`Node`) knows nothing about thunder; ``b200ex/executor.py`` translates bound symbols into it.
"""

from __future__ import annotations

import ctypes
import hashlib
import math
import os
import struct
import threading
from dataclasses import dataclass, field

import torch

from . import jit
from .pointwise import BINARY, COMPARE, EXACT, UNARY, PwUnsupported, _f32_bits

MAX_RANK = 8
MAX_LEAVES = 96
MAX_OUTS = 8
SMALL_REDUCE = 16  # a reduction over at most this many elements is unrolled into the consuming expression
MAX_ROW_FLOATS = 192  # per-thread fp32 registers a row kernel may spend on resident leaves and accumulators

_DT = {torch.float32: "f32", torch.bfloat16: "bf16", torch.float16: "f16", torch.bool: "bool"}
# What a leaf may be at run time.  The reference's traces are not always dtype-consistent (the backward of `amax`
# hands an int64 tie count to a proxy declared bool), so loads follow the tensor that actually arrives; integers are
# read as fp32 values.
_LOAD_DT = {**_DT, torch.int64: "i64", torch.int32: "i32"}
_DT_BYTES = {"f32": 4, "bf16": 2, "f16": 2, "bool": 1, "i64": 8, "i32": 4}


# --------------------------------------------------------------------------------------------------
# IR
# --------------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class Node:
    """op: an elementwise op of b200ex.pointwise.OPS, or one of convert / broadcast_in_dim / reshape / transpose /
    slice / pad / cat / full / reduce.  args are value names or Python numbers; attrs carries the op's parameters:
    broadcast_in_dim {bdims}, transpose {perm}, slice {starts, strides}, pad {value, config=[(lo, hi, dilation)]},
    cat {dim}, full {value}, reduce {kind: sum|amax|amin|prod, dims}."""

    op: str
    out: str
    dtype: torch.dtype
    shape: tuple
    args: tuple
    attrs: dict = field(default_factory=dict, compare=False, hash=False)


class _SplitVar(Exception):
    def __init__(self, var: int, inner: int):
        self.var, self.inner = var, inner


# ---- affine forms over iteration variables: (terms, const) with terms = ((var, coef), ...) sorted by var ------------
def _aff(terms=(), c=0):
    return (tuple(sorted((v, k) for v, k in terms if k != 0)), int(c))


def _aff_add(a, b):
    d = dict(a[0])
    for v, k in b[0]:
        d[v] = d.get(v, 0) + k
    return _aff(d.items(), a[1] + b[1])


def _aff_scale(a, k):
    return _aff(((v, c * k) for v, c in a[0]), a[1] * k)


def _aff_shift(a, c):
    return (a[0], a[1] + int(c))


_ZERO = _aff()


@dataclass
class Leaf:
    name: str
    dtype: torch.dtype
    coords: tuple  # one affine form per dimension of the input tensor
    guard: tuple  # sorted condition indices that must all hold for the element to exist


@dataclass
class Out:
    name: str
    dtype: torch.dtype
    shape: tuple
    sid: int
    kind: str  # "full" (one element per iteration point), "row" (one per row), "col" (sum over rows, one per column)


@dataclass
class Program:
    kind: str  # "flat" | "row"
    ext: list  # extent of every iteration variable
    n_outer: int  # row kernels: variables [0, n_outer) index the row, the rest run inside it; flat: == len(ext)
    leaves: list
    conds: list  # (linear terms, lo, hi): lo <= sum(coef * var) < hi
    ssa: list  # (op, a, b, c, imm)
    outputs: list
    W: int = 1
    signature: str = ""
    phase: list | None = None  # row kernels: phase of every ssa value
    full: list | None = None  # row kernels: does the value vary inside a row?
    in_shapes: dict | None = None  # input name -> shape, for the inputs the leaves read
    order: list | None = None  # flat kernels: variables from fastest to slowest in the thread mapping

    @property
    def row_len(self):
        return math.prod(self.ext[self.n_outer:]) if self.kind == "row" else 1

    @property
    def n_rows(self):
        return math.prod(self.ext[: self.n_outer])


# --------------------------------------------------------------------------------------------------
# Front end: IR -> Program
# --------------------------------------------------------------------------------------------------
class _Builder:
    def __init__(self, kind, dims, n_outer_dims, inputs, producer, meta):
        # dims: list of factor lists, one per iteration dimension; variables are numbered over the flattened factors
        self.kind, self.dims, self.n_outer_dims = kind, dims, n_outer_dims
        self.inputs, self.producer, self.meta = inputs, producer, meta
        self.ext = [f for d in dims for f in d]
        self.n_outer = sum(len(d) for d in dims[:n_outer_dims]) if kind == "row" else len(self.ext)
        self.ssa, self.cse, self.consts = [], {}, {}
        self.leaves, self.leaf_ids = [], {}
        self.conds, self.cond_ids = [], {}
        self.memo = {}
        self.full, self.phase = [], []

    # ---- iteration space -------------------------------------------------------------------------------------
    def dim_coord(self, d):
        """Affine coordinate of iteration dimension d in terms of its factor variables."""
        base = sum(len(x) for x in self.dims[:d])
        terms, stride = [], 1
        for i in range(len(self.dims[d]) - 1, -1, -1):
            terms.append((base + i, stride))
            stride *= self.dims[d][i]
        return _aff(terms)

    def var_dim(self, var):
        acc = 0
        for d, fs in enumerate(self.dims):
            if var < acc + len(fs):
                return d, var - acc
            acc += len(fs)
        raise AssertionError

    # ---- ssa ----------------------------------------------------------------------------------------------------
    def emit(self, op, a=-1, b=-1, c=-1, imm=None):
        key = (op, a, b, c, imm)
        if key in self.cse:
            return self.cse[key]
        self.ssa.append(key)
        sid = len(self.ssa) - 1
        self.cse[key] = sid
        args = [x for x in (a, b, c) if x >= 0]
        if op == "leaf":
            lf = self.leaves[imm]
            inner = set(range(self.n_outer, len(self.ext)))
            uses = {v for co in lf.coords for v, _ in co[0]} | {v for ci in lf.guard for v, _ in self.conds[ci][0]}
            self.full.append(bool(uses & inner))
            self.phase.append(0)
        elif op == "pred":
            inner = set(range(self.n_outer, len(self.ext)))
            self.full.append(bool({v for v, _ in self.conds[imm][0]} & inner))
            self.phase.append(0)
        elif op.startswith("reduce_"):
            self.full.append(False)
            self.phase.append(self.phase[a] + 1)
        else:
            self.full.append(any(self.full[x] for x in args))
            self.phase.append(max((self.phase[x] for x in args), default=0))
        return sid

    def const(self, x):
        return self.emit("const", imm=_f32_bits(x))

    # ---- ranges and conditions --------------------------------------------------------------------------------
    def rng(self, a, guard=()):
        """Inclusive range of an affine form.  Guard conditions whose linear part is a sub-sum of the form tighten the
        range of that sub-sum (e.g. a `cat` piece's `64 <= d < 128` inside `768 g + 128 j + d - 64`)."""
        lo = hi = a[1]
        rem = dict(a[0])

        vlo, vhi = {}, {}

        def plain(terms):
            l = h = 0
            for v, k in terms:
                a0, a1 = vlo.get(v, 0), vhi.get(v, self.ext[v] - 1)
                l, h = l + min(k * a0, k * a1), h + max(k * a0, k * a1)
            return l, h

        if guard:
            by_terms: dict = {}
            for ci in guard:
                terms, clo, chi = self.conds[ci]
                if len(terms) == 1 and terms[0][1] == 1:  # a condition on one variable bounds it under any coefficient
                    v = terms[0][0]
                    vlo[v], vhi[v] = max(vlo.get(v, 0), clo), min(vhi.get(v, self.ext[v] - 1), chi - 1)
                    continue
                by_terms.setdefault(terms, []).append((clo, chi))
            for terms in sorted(by_terms, key=len, reverse=True):
                if terms and all(rem.get(v) == k for v, k in terms):
                    l, h = plain(terms)
                    for clo, chi in by_terms[terms]:
                        l, h = max(l, clo), min(h, chi - 1)
                    lo, hi = lo + l, hi + h
                    for v, _ in terms:
                        del rem[v]
        l, h = plain(rem.items())
        return lo + l, hi + h

    def cond(self, a, lo, hi, guard):
        """Condition lo <= a < hi.  Returns True / False when the ranges decide it, else a condition index."""
        rlo, rhi = self.rng(a, guard)
        if rlo >= lo and rhi < hi:
            return True
        if rhi < lo or rlo >= hi:
            return False
        terms, clo, chi = a[0], max(lo, rlo) - a[1], min(hi, rhi + 1) - a[1]
        # lo <= g*rest + k*v < hi with lo, hi multiples of g and 0 <= k*v < g does not depend on v: drop the term and
        # divide by g (a `cat` boundary at a multiple of the low digit's span only tests the high digits)
        while len(terms) > 1:
            v, k = min(terms, key=lambda t: abs(t[1]))
            g = math.gcd(*[c for u, c in terms if u != v], clo, chi)
            if k <= 0 or g <= k * (self.ext[v] - 1):
                break
            terms = tuple((u, c // g) for u, c in terms if u != v)
            clo, chi = clo // g, chi // g
        key = (terms, clo, chi)
        if key not in self.cond_ids:
            self.cond_ids[key] = len(self.conds)
            self.conds.append(key)
        return self.cond_ids[key]

    # ---- values ---------------------------------------------------------------------------------------------------
    def value(self, name, coords, guard=()):
        key = (name, coords, guard)
        r = self.memo.get(key)
        if r is None:
            if name in self.inputs:
                r = self.leaf(name, coords, guard)
            else:
                n = self.producer.get(name)
                if n is None:
                    raise PwUnsupported(f"value {name} has no producer in the region")
                r = self.node(n, coords, guard)
            self.memo[key] = r
        return r

    def leaf(self, name, coords, guard):
        dt, shape = self.inputs[name]
        coords = tuple(_ZERO if shape[i] == 1 else c for i, c in enumerate(coords))
        for i, c in enumerate(coords):
            lo, hi = self.rng(c, guard)
            if lo < 0 or hi >= shape[i]:
                raise PwUnsupported(f"index map of {name} leaves the tensor along dim {i}: [{lo}, {hi}] vs {shape[i]}")
        # only the conditions that protect this access are kept: those that share a variable with the coordinates
        used = {v for c in coords for v, _ in c[0]}
        g = tuple(sorted(ci for ci in guard if {v for v, _ in self.conds[ci][0]} & used))
        key = (name, coords, g)
        if key not in self.leaf_ids:
            if len(self.leaves) >= MAX_LEAVES:
                raise PwUnsupported("too many gather leaves for one kernel")
            self.leaves.append(Leaf(name, dt, coords, g))
            self.leaf_ids[key] = len(self.leaves) - 1
        return self.emit("leaf", imm=self.leaf_ids[key])

    def arg(self, a, coords, guard):
        if isinstance(a, str):
            return self.value(a, coords, guard)
        if isinstance(a, (bool, int, float)):
            return self.const(float(a))
        raise PwUnsupported(f"operand {a!r}")

    def rounded(self, r, n):
        if n.op in EXACT:
            return r
        if n.dtype == torch.bfloat16:
            return self.emit("rnd_bf16", r)
        if n.dtype == torch.float16:
            return self.emit("rnd_f16", r)
        return r

    @staticmethod
    def as_dtype(x, dt):
        if isinstance(x, (int, float)) and not isinstance(x, bool) and dt in (torch.bfloat16, torch.float16):
            return float(torch.tensor(float(x), dtype=dt))
        return x

    def node(self, n: Node, coords, guard):
        op = n.op
        meta = self.meta
        if op == "broadcast_in_dim":
            a = n.args[0]
            ashape = meta[a][1]
            return self.value(a, tuple(_ZERO if ashape[i] == 1 else coords[d] for i, d in enumerate(n.attrs["bdims"])), guard)
        if op == "transpose":
            perm = n.attrs["perm"]
            inv = [0] * len(perm)
            for i, p in enumerate(perm):
                inv[p] = i
            return self.value(n.args[0], tuple(coords[inv[d]] for d in range(len(perm))), guard)
        if op == "slice":
            st, sr = n.attrs["starts"], n.attrs["strides"]
            return self.value(n.args[0], tuple(_aff_shift(_aff_scale(c, sr[i]), st[i]) for i, c in enumerate(coords)), guard)
        if op == "reshape":
            a = n.args[0]
            return self.value(a, self.redigit(coords, n.shape, meta[a][1], guard), guard)
        if op == "full":
            return self.const(float(self.as_dtype(n.attrs["value"], n.dtype)))
        if op == "pad":
            a = n.args[0]
            padv = self.const(float(self.as_dtype(n.attrs["value"], n.dtype)))
            ashape = meta[a][1]
            if 0 in ashape:
                return padv
            g, preds, inner = list(guard), [], []
            for i, (lo, hi, dil) in enumerate(n.attrs["config"]):
                if dil != 0:
                    raise PwUnsupported("pad with dilation")
                c = _aff_shift(coords[i], -lo)
                inner.append(c)
                r = self.cond(c, 0, ashape[i], tuple(g))
                if r is False:
                    return padv
                if r is not True:
                    g.append(r)
                    preds.append(r)
            v = self.value(a, tuple(inner), tuple(sorted(set(g))))
            if not preds:
                return v
            p = self.all_of(preds)
            return self.emit("where", p, v, padv)
        if op == "cat":
            dim = n.attrs["dim"]
            off, pieces = 0, []
            for a in n.args:
                ext = meta[a][1][dim]
                if ext:
                    pieces.append((a, off, ext))
                off += ext
            live = []
            for a, o, ext in pieces:
                c = _aff_shift(coords[dim], -o)
                r = self.cond(c, 0, ext, guard)
                if r is False:
                    continue
                cs = coords[:dim] + (c,) + coords[dim + 1:]
                if r is True:
                    return self.value(a, cs, guard)
                live.append((r, self.value(a, cs, tuple(sorted(set(guard) | {r})))))
            if not live:
                raise PwUnsupported("cat: no reachable piece")
            res = live[-1][1]
            for r, v in reversed(live[:-1]):
                res = self.emit("where", self.emit("pred", imm=r), v, res)
            return res
        if op == "reduce":
            r = self.reduce(n, coords, guard)
            if n.dtype in (torch.bfloat16, torch.float16):
                r = self.emit("rnd_bf16" if n.dtype == torch.bfloat16 else "rnd_f16", r)
            return r
        if op == "convert":
            r = self.arg(n.args[0], coords, guard)
            src_dt = meta[n.args[0]][0] if isinstance(n.args[0], str) else torch.float32
            if n.dtype == torch.bfloat16 and src_dt != torch.bfloat16:
                return self.emit("rnd_bf16", r)
            if n.dtype == torch.float16 and src_dt != torch.float16:
                return self.emit("rnd_f16", r)
            if n.dtype == torch.bool and src_dt != torch.bool:
                return self.emit("ne", r, self.const(0.0))
            return r
        if op == "mov":
            return self.arg(n.args[0], coords, guard)
        # elementwise: operands of a smaller rank / broadcast extent follow numpy rules on the trailing dims
        def sub(a):
            if not isinstance(a, str):
                return coords
            shp = meta[a][1]
            if tuple(shp) == tuple(n.shape):
                return coords
            if len(shp) > len(coords):
                raise PwUnsupported(f"{op}: operand rank {len(shp)} above result rank {len(coords)}")
            tail = coords[len(coords) - len(shp):]
            return tuple(_ZERO if shp[i] == 1 else c for i, c in enumerate(tail))

        A = lambda a: self.arg(a, sub(a), guard)
        if op == "where":
            c = A(n.args[0])
            a, b = (A(self.as_dtype(x, n.dtype)) for x in n.args[1:])
            return self.emit("where", c, a, b)
        if op == "pow" and isinstance(n.args[1], (int, float)) and not isinstance(n.args[1], bool):
            a, e = A(n.args[0]), float(n.args[1])
            if e == 2.0:
                return self.rounded(self.emit("mul", a, a), n)
            if e == 3.0:
                return self.rounded(self.emit("mul", self.emit("mul", a, a), a), n)
            if e == 0.5:
                return self.rounded(self.emit("sqrt", a), n)
            if e == -0.5:
                return self.rounded(self.emit("rsqrt", a), n)
            if e == -1.0:
                return self.rounded(self.emit("reciprocal", a), n)
            if e == 1.0:
                return a
            return self.rounded(self.emit("pow", a, self.const(e)), n)
        if op in BINARY:
            x, y = n.args[0], n.args[1]
            if op in COMPARE:
                dts = [meta[v][0] for v in (x, y) if isinstance(v, str)]
                if len(dts) == 1:
                    x, y = self.as_dtype(x, dts[0]), self.as_dtype(y, dts[0])
            return self.rounded(self.emit(op, A(x), A(y)), n)
        if op in UNARY:
            return self.rounded(self.emit(op, A(n.args[0])), n)
        raise PwUnsupported(f"op {op}")

    def all_of(self, conds):
        p = self.emit("pred", imm=conds[0])
        for ci in conds[1:]:
            p = self.emit("and", p, self.emit("pred", imm=ci))
        return p

    # ---- reshape: re-digitise the linear index ---------------------------------------------------------------
    def redigit(self, coords, shape, ashape, guard):
        if tuple(shape) == tuple(ashape):
            return coords
        lin, stride = _ZERO, 1
        for i in range(len(shape) - 1, -1, -1):
            lin = _aff_add(lin, _aff_scale(coords[i], stride))
            stride *= shape[i]
        astr = [1] * len(ashape)
        for j in range(len(ashape) - 2, -1, -1):
            astr[j] = astr[j + 1] * ashape[j + 1]
        digit_terms = [[] for _ in ashape]
        for v, k in lin[0]:
            if k < 0:
                raise PwUnsupported("reshape of a reversed index")
            # the digit whose weight range contains k
            j = next((j for j in range(len(ashape)) if ashape[j] > 1 and astr[j] <= k < astr[j] * ashape[j]), None)
            if j is None:
                if self.ext[v] == 1:
                    continue
                raise PwUnsupported("reshape: index term outside the source tensor")
            if k % astr[j]:
                raise PwUnsupported("reshape: a variable straddles two source dimensions unevenly")
            top = astr[j] * ashape[j]
            vlo, vhi = self.rng(_aff([(v, 1)]), guard)  # the guard may confine the variable to one `cat` piece
            if k * (vhi - vlo + 1) > top:
                if top % k or self.ext[v] % (top // k):
                    raise PwUnsupported("reshape: extents do not factor along the source dimensions")
                raise _SplitVar(v, top // k)
            digit_terms[j].append((v, k // astr[j]))
        # constant part: distribute over the digits from the innermost one so that every digit stays in range
        out = [None] * len(ashape)
        carry = lin[1]
        for j in range(len(ashape) - 1, -1, -1):
            t = _aff(digit_terms[j])
            if ashape[j] == 1:
                out[j] = _ZERO
                continue
            lo, _ = self.rng(t, guard)
            q = (lo + carry) // ashape[j]
            c = carry - q * ashape[j]
            out[j] = _aff_shift(t, c)
            carry = q
        if carry != 0:
            raise PwUnsupported("reshape: index leaves the source tensor")
        for j, c in enumerate(out):
            lo, hi = self.rng(c, guard)
            if lo < 0 or hi >= ashape[j]:
                raise PwUnsupported("reshape: a digit can carry into the next source dimension")
        return tuple(out)

    # ---- reductions ---------------------------------------------------------------------------------------------
    def row_sum(self, x):
        """Row sum of value x with row-constant factors pulled out: sum(s * y) = s * sum(y), sum(-y) = -sum(y),
        sum(y / s) = sum(y) / s for s constant along the row.  The sum is reassociated across lanes anyway, so this
        stays inside the same fp32 rounding noise; what it buys is one multiply per element less and — the real gain
        — sums whose factor comes out of an EARLIER reduction no longer wait for it: a hand-spelled LayerNorm backward
        has all three of its row sums in the first pass instead of two passes with a barrier between them."""
        if os.environ.get("B200EX_HOIST", "1") != "0":
            op, a, b = self.ssa[x][0], self.ssa[x][1], self.ssa[x][2]
            if op == "mul" and self.full[a] != self.full[b]:
                s_, y = (a, b) if not self.full[a] else (b, a)
                return self.emit("mul", s_, self.row_sum(y))
            if op == "neg":
                return self.emit("neg", self.row_sum(a))
            if op == "div" and self.full[a] and not self.full[b]:
                return self.emit("div", self.row_sum(a), b)
        return self.emit("reduce_sum", x)

    def reduce(self, n: Node, coords, guard):
        a = n.args[0]
        ashape = self.meta[a][1]
        dims = sorted(d % len(ashape) for d in n.attrs["dims"])
        kind = n.attrs["kind"]
        keep = [d for d in range(len(ashape)) if d not in dims]
        count = math.prod(ashape[d] for d in dims)

        def operand_coords(fixed):
            cs, it = [None] * len(ashape), iter(coords)
            for d in range(len(ashape)):
                cs[d] = fixed[d] if d in fixed else next(it)
            return tuple(cs)

        trailing = dims == list(range(len(ashape) - len(dims), len(ashape)))
        if self.kind == "row" and trailing and count > 1:
            red_dims = list(range(self.n_outer_dims, len(self.dims)))
            if [math.prod(self.dims[d]) for d in red_dims] == [ashape[d] for d in dims]:
                inner = set(range(self.n_outer, len(self.ext)))
                if any(v in inner for c in coords for v, _ in c[0]):
                    raise PwUnsupported("row reduction demanded at a position inside the row")
                cs = operand_coords({d: self.dim_coord(rd) for d, rd in zip(dims, red_dims)})
                x = self.value(a, cs, guard)
                if not self.full[x]:
                    raise PwUnsupported("row reduction of a value that is constant along the row")
                return self.row_sum(x) if kind == "sum" else self.emit("reduce_" + kind, x)
        if count > SMALL_REDUCE:
            raise PwUnsupported(f"reduction over {count} elements along dims {dims} is neither a row nor a small reduction")
        comb = {"sum": "add", "prod": "mul", "amax": "maximum", "amin": "minimum"}[kind]
        acc = None
        idx = [0] * len(dims)
        for flat in range(count):
            r = flat
            for i in range(len(dims) - 1, -1, -1):
                idx[i] = r % ashape[dims[i]]
                r //= ashape[dims[i]]
            v = self.value(a, operand_coords({d: _aff(c=i) for d, i in zip(dims, idx)}), guard)
            acc = v if acc is None else self.emit(comb, acc, v)
        return acc


def _sig(P: Program, in_shapes) -> str:
    leafname = {}
    parts = ["K:" + P.kind, "E:" + ",".join(map(str, P.ext)), f"O:{P.n_outer}"]
    for lf in P.leaves:
        k = leafname.setdefault(lf.name, len(leafname))
        parts.append(f"L{k}:{_DT[lf.dtype]}:{tuple(in_shapes[lf.name])}:{lf.coords}:{lf.guard}")
    parts.append("C:" + repr(P.conds))
    parts.append("S:" + ";".join(f"{op}({a},{b},{c},{imm})" for op, a, b, c, imm in P.ssa))
    parts.append("R:" + ",".join(f"{o.sid}:{_DT[o.dtype]}:{o.kind}" for o in P.outputs))
    return hashlib.sha1("|".join(parts).encode()).hexdigest()


def _build(kind, shape, n_outer_dims, inputs, producer, meta, outs_spec):
    """outs_spec: [(name, kind)].  Iterates variable splitting to a fixed point."""
    dims = [[int(s)] for s in shape]
    for _ in range(32):
        B = _Builder(kind, dims, n_outer_dims, inputs, producer, meta)
        try:
            outputs = []
            for name, okind in outs_spec:
                dt, oshape = meta[name]
                if okind == "full":
                    coords = tuple(B.dim_coord(d) for d in range(len(dims)))
                    sid = B.value(name, coords)
                elif okind == "row":
                    # the output's dims are the outer dims, possibly with extra 1s
                    it = iter(range(n_outer_dims))
                    coords = []
                    for s in oshape:
                        if s == 1:
                            coords.append(_ZERO)
                            continue
                        d = next(it)
                        while math.prod(dims[d]) == 1:
                            d = next(it)
                        coords.append(B.dim_coord(d))
                    sid = B.value(name, tuple(coords))
                    if B.full[sid]:
                        raise PwUnsupported(f"output {name} varies inside a row but has a row shape")
                else:  # "col": name is a reduce node over the outer dims; its operand is demanded at the identity
                    n = producer[name]
                    coords = tuple(B.dim_coord(d) for d in range(len(dims)))
                    sid = B.value(n.args[0], coords)
                outputs.append(Out(name, dt, tuple(oshape), sid, okind))
            P = Program(kind, B.ext, B.n_outer, B.leaves, B.conds, B.ssa, outputs, phase=B.phase, full=B.full)
            if kind == "flat" and B.ext:
                # conditions on the innermost variable whose boundaries share a granule G: split it into (hi, lo) so
                # that they test `hi` only and the thread mapping can keep a warp on one side of every boundary
                vd = len(B.ext) - 1
                bounds = [b for terms, lo, hi in B.conds if terms == ((vd, 1),) for b in (lo, hi)]
                if bounds and not any(vd in dict(terms) and terms != ((vd, 1),) for terms, _, _ in B.conds):
                    G = math.gcd(B.ext[vd], *bounds)
                    if 8 <= G < B.ext[vd] and G % 8 == 0:
                        raise _SplitVar(vd, G)
            P.W = _choose_width(P, inputs)
            P.order = _thread_order(P)
            P.in_shapes = {lf.name: tuple(inputs[lf.name][1]) for lf in P.leaves}
            P.signature = _sig(P, P.in_shapes)
            return P
        except _SplitVar as s:
            d, i = B.var_dim(s.var)
            n = dims[d][i]
            if s.inner <= 0 or n % s.inner or s.inner == n or s.inner == 1:
                raise PwUnsupported("reshape: cannot split the iteration variable") from None
            dims = [list(x) for x in dims]
            dims[d][i:i + 1] = [n // s.inner, s.inner]
    raise PwUnsupported("reshape: variable splitting did not converge")


def _thread_order(P: Program) -> list:
    """Flat kernels: how the linear work-item index maps to the iteration variables, fastest first, as
    (variable, part, size) with part "all" | "lo" | "hi".  Without conditions: the natural order.  With conditions a
    CTA's worth of items (innermost chunks x a block of the next unconditioned variables) comes first, then the
    conditioned variables, then the rest: a warp never straddles a `cat` / `pad` boundary (no divergence), and the
    two sides of a boundary are still visited close together in time (the partner halves of a rotary pair hit L2)."""
    nv = len(P.ext)
    if not nv:
        return []
    cond_vars = {v for terms, _, _ in P.conds for v, _ in terms}
    if not cond_vars:
        return [(v, "all", P.ext[v]) for v in range(nv - 1, -1, -1)]
    order = [(nv - 1, "all", P.ext[nv - 1])]
    have = max(1, P.ext[nv - 1] // max(P.W, 1))
    free = [v for v in range(nv - 2, -1, -1) if v not in cond_vars]
    tail = []
    while free and have < 256:
        v = free.pop(0)
        n = P.ext[v]
        if have * n <= 256:
            order.append((v, "all", n))
            have *= n
            continue
        b = next(d for d in range(1, n + 1) if n % d == 0 and have * d >= 256)
        if b == n:
            order.append((v, "all", n))
        else:
            order.append((v, "lo", b))
            tail.append((v, "hi", n // b))
        have *= b
    order += [(v, "all", P.ext[v]) for v in range(nv - 2, -1, -1) if v in cond_vars]
    order += tail + [(v, "all", P.ext[v]) for v in free]
    return order


def _choose_width(P: Program, inputs) -> int:
    """Elements per work item along the innermost variable: the largest power of two <= 8 such that every condition is
    uniform over an item and every leaf reads the item as one aligned run, one broadcast value, or element by element."""
    if not P.ext:
        return 1
    vd = len(P.ext) - 1
    if P.kind == "row" and vd < P.n_outer:
        return 1
    W = 8
    while W > 1 and P.ext[vd] % W:
        W //= 2
    for terms, lo, hi in P.conds:
        k = dict(terms).get(vd, 0)
        if k == 0:
            continue
        if abs(k) != 1:
            return 1
        # the other terms must move the boundary in multiples of W as well
        while W > 1 and (lo % W or hi % W or any(c % W for v, c in terms if v != vd)):
            W //= 2
    return W


# --------------------------------------------------------------------------------------------------
# Component -> ordered launches
# --------------------------------------------------------------------------------------------------
@dataclass
class Launch:
    program: Program
    views: dict  # output name -> (source name computed by the program, shape): zero-copy reshapes of kernel outputs


def _classify_reduce(n: Node, meta):
    a = n.args[0]
    ashape = meta[a][1]
    dims = sorted(d % len(ashape) for d in n.attrs["dims"])
    count = math.prod(ashape[d] for d in dims)
    if count <= 1:
        return "small"
    if dims == list(range(len(ashape) - len(dims), len(ashape))) and count > SMALL_REDUCE:
        return "row"
    if dims == list(range(len(dims))) and n.attrs["kind"] == "sum" and count > SMALL_REDUCE:
        return "col"
    if count <= SMALL_REDUCE:
        return "small"
    if dims == list(range(len(ashape) - len(dims), len(ashape))):
        return "row"
    raise PwUnsupported(f"reduction over dims {dims} of shape {tuple(ashape)}")


def compile_component(inputs: dict, nodes: list, outputs: list) -> list:
    """inputs: name -> (dtype, shape); nodes in execution order; outputs: names that must be materialised.
    Returns :class:`Launch` objects to run in order.  Raises PwUnsupported."""
    meta = dict(inputs)
    producer = {}
    for n in nodes:
        meta[n.out] = (n.dtype, tuple(n.shape))
        producer[n.out] = n
    for name, (dt, _) in meta.items():
        if dt not in _DT:
            raise PwUnsupported(f"dtype {dt} of {name}")
    # outputs that are contiguous reshapes of another computed value are views of that value's buffer
    views, want = {}, []
    for o in outputs:
        n = producer.get(o)
        src = o
        while n is not None and n.op == "reshape" and n.args[0] in producer:
            src = n.args[0]
            n = producer[src]
        if src != o and src in outputs:  # free only when the source is written anyway
            views[o] = (src, tuple(meta[o][1]))
        else:
            src = o
        if src not in want:
            want.append(src)

    kinds = {n.out: _classify_reduce(n, meta) for n in nodes if n.op == "reduce"}
    # a column sum whose only use is a dtype conversion is folded straight into that dtype (ltorch.sum of a 16-bit
    # tensor is convert -> sum -> convert): the fold kernel rounds once, no extra launch
    for o in [o for o, k in kinds.items() if k == "col" and o not in outputs]:
        users = [n for n in nodes if any(isinstance(a, str) and a == o for a in n.args)]
        if len(users) == 1 and users[0].op == "convert":
            cv, red = users[0], producer[o]
            merged = Node("reduce", cv.out, cv.dtype, cv.shape, red.args, red.attrs)
            nodes = [merged if n is red else n for n in nodes if n is not cv]
            producer = {n.out: n for n in nodes}
            meta[cv.out] = (cv.dtype, tuple(cv.shape))
            kinds[cv.out] = kinds.pop(o)
    cols = [o for o, k in kinds.items() if k == "col"]
    # everything downstream of a column sum waits for the fold: second stage
    after = set()
    for n in nodes:
        if any(isinstance(a, str) and (a in after or a in cols) for a in n.args):
            after.add(n.out)
    stage1 = [n for n in nodes if n.out not in after]
    stage2 = [n for n in nodes if n.out in after]
    s1_names = {n.out for n in stage1}
    need1 = [o for o in want if o in s1_names or o in inputs]
    for n in stage2:
        for a in n.args:
            if isinstance(a, str) and a in s1_names and a not in need1:
                need1.append(a)
    if any(o in inputs for o in need1):
        raise PwUnsupported("a region input is returned unchanged")

    launches = []
    rows = [producer[o] for o, k in kinds.items() if k == "row" and o in s1_names]
    if rows or cols:
        if rows:
            a = rows[0].args[0]
            shape = tuple(meta[a][1])
            n_outer = len(shape) - len(rows[0].attrs["dims"])
        else:
            a = producer[cols[0]].args[0]
            shape = tuple(meta[a][1])
            n_outer = len(producer[cols[0]].attrs["dims"])
        outer, red = shape[:n_outer], shape[n_outer:]
        spec, flat_outs = [], []
        for o in need1:
            oshape = tuple(meta[o][1])
            if o in cols:
                if tuple(meta[producer[o].args[0]][1]) != shape:
                    raise PwUnsupported("column sum over a different iteration space")
                spec.append((o, "col"))
            elif oshape == shape:
                spec.append((o, "full"))
            elif tuple(s for s in oshape if s != 1) == tuple(s for s in outer if s != 1) and math.prod(red) > 1:
                spec.append((o, "row"))
            else:
                flat_outs.append(o)  # e.g. an upcast weight saved for the backward: its own small elementwise kernel
        if spec:
            launches.append(Launch(_build("row", shape, n_outer, inputs, producer, meta, spec), {}))
    else:
        flat_outs = need1
    if flat_outs:
        groups: dict = {}
        for o in flat_outs:
            groups.setdefault(tuple(meta[o][1]), []).append(o)
        for shape, outs in groups.items():
            for i in range(0, len(outs), MAX_OUTS):
                launches.append(Launch(_build("flat", shape, len(shape), inputs, producer, meta, [(o, "full") for o in outs[i:i + MAX_OUTS]]), {}))
    if stage2:
        inputs2 = dict(inputs)
        inputs2.update({o: meta[o] for o in need1})
        want2 = [o for o in want if o in after]
        if want2:
            launches += compile_component(inputs2, stage2, want2)
    for o, (src, shape) in views.items():
        owner = next((ln for ln in launches if any(out.name == src for out in ln.program.outputs)), None)
        if owner is None:
            if any(o in ln.views for ln in launches):
                continue  # already attached by the recursive call for the second stage
            raise PwUnsupported("view of a value that no kernel of the region materialises")
        owner.views[o] = (src, shape)
    return launches


# --------------------------------------------------------------------------------------------------
# Code generation
# --------------------------------------------------------------------------------------------------
_HEAD = r"""
typedef unsigned int u32; typedef unsigned short u16; typedef unsigned char u8; typedef long long i64;
struct Op { void* ptr; i64 st[%(MAX_RANK)d]; };
struct Params { Op in[%(NIN)d]; Op out[%(NOUT)d]; float* scratch; int nparts; };
struct V4 { u32 x, y, z, w; };
struct V2 { u32 x, y; };
__device__ __forceinline__ V4 ldg16(const void* p) { V4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%%0,%%1,%%2,%%3}, [%%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r; }
__device__ __forceinline__ V2 ldg8(const void* p) { V2 r; asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%%0,%%1}, [%%2];" : "=r"(r.x), "=r"(r.y) : "l"(p)); return r; }
__device__ __forceinline__ u32 ldg4(const void* p) { u32 r; asm volatile("ld.global.nc.L1::no_allocate.u32 %%0, [%%1];" : "=r"(r) : "l"(p)); return r; }
// 16-byte asynchronous global -> shared copy (LDGSTS, L2 only); ok == false copies nothing and zero-fills
__device__ __forceinline__ void cpa16(void* dst, const void* src, bool ok) {
  const u32 d = (u32)__cvta_generic_to_shared(dst); const int n = ok ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%%0], [%%1], 16, %%2;" :: "r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cpa_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cpa_wait() { asm volatile("cp.async.wait_group %%0;" :: "n"(N) : "memory"); }
__device__ __forceinline__ void stg16(void* p, V4 v) { asm volatile("st.global.L1::no_allocate.v4.u32 [%%0], {%%1,%%2,%%3,%%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
__device__ __forceinline__ void stg8(void* p, V2 v) { asm volatile("st.global.L1::no_allocate.v2.u32 [%%0], {%%1,%%2};" :: "l"(p), "r"(v.x), "r"(v.y) : "memory"); }
__device__ __forceinline__ void stg4(void* p, u32 v) { asm volatile("st.global.L1::no_allocate.u32 [%%0], %%1;" :: "l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ float bf2f(u32 h) { return __uint_as_float(h << 16); }
__device__ __forceinline__ u32 f2bf2(float lo, float hi) { u32 r; asm("cvt.rn.bf16x2.f32 %%0, %%1, %%2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ float h2f(u32 h) { float f; asm("{.reg .b16 t; cvt.u16.u32 t, %%1; cvt.f32.f16 %%0, t;}" : "=f"(f) : "r"(h)); return f; }
__device__ __forceinline__ u32 f2h2(float lo, float hi) { u32 r; asm("cvt.rn.f16x2.f32 %%0, %%1, %%2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ float rnd_bf16(float x) { return __uint_as_float(f2bf2(x, 0.f) << 16); }
__device__ __forceinline__ float rnd_f16(float x) { return h2f(f2h2(x, 0.f) & 0xffffu); }
__device__ __forceinline__ void rnd_bf16_2(float a, float b, float& x, float& y) { const u32 p = f2bf2(a, b); x = __uint_as_float(p << 16); y = __uint_as_float(p & 0xffff0000u); }
__device__ __forceinline__ void rnd_f16_2(float a, float b, float& x, float& y) { const u32 p = f2h2(a, b); x = h2f(p & 0xffffu); y = h2f(p >> 16); }
// packed 16-bit arithmetic on raw words (two bf16 / f16 values per register).  mul/add/sub with an explicit .rn are
// never contracted into an fma, so each is exactly one rounding of the exact result.
__device__ __forceinline__ u32 hmul2_bf16(u32 a, u32 b) { u32 d; asm("mul.rn.bf16x2 %%0, %%1, %%2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ u32 hadd2_bf16(u32 a, u32 b) { u32 d; asm("add.rn.bf16x2 %%0, %%1, %%2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ u32 hsub2_bf16(u32 a, u32 b) { u32 d; asm("sub.rn.bf16x2 %%0, %%1, %%2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ u32 hmul2_f16(u32 a, u32 b) { u32 d; asm("mul.rn.f16x2 %%0, %%1, %%2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ u32 hadd2_f16(u32 a, u32 b) { u32 d; asm("add.rn.f16x2 %%0, %%1, %%2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ u32 hsub2_f16(u32 a, u32 b) { u32 d; asm("sub.rn.f16x2 %%0, %%1, %%2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ u32 hneg2(u32 a) { return a ^ 0x80008000u; }
__device__ __forceinline__ float2 unpk_bf16(u32 w) { return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)); }
__device__ __forceinline__ float2 unpk_f16(u32 w) { return make_float2(h2f(w & 0xffffu), h2f(w >> 16)); }
__device__ __forceinline__ u32 wsel(const V4& v, int p) { return p == 0 ? v.x : (p == 1 ? v.y : (p == 2 ? v.z : v.w)); }
__device__ __forceinline__ float nanmax(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fmaxf(a, b); }
__device__ __forceinline__ float nanmin(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fminf(a, b); }
__device__ __forceinline__ float sgn(float a) { return a > 0.f ? 1.f : (a < 0.f ? -1.f : a); }
__device__ __forceinline__ float ld1_f32(const void* p, i64 o) { return ((const float*)p)[o]; }
__device__ __forceinline__ float ld1_bf16(const void* p, i64 o) { return bf2f(((const u16*)p)[o]); }
__device__ __forceinline__ float ld1_f16(const void* p, i64 o) { return h2f(((const u16*)p)[o]); }
__device__ __forceinline__ float ld1_bool(const void* p, i64 o) { return ((const u8*)p)[o] != 0 ? 1.f : 0.f; }
__device__ __forceinline__ float ld1_i64(const void* p, i64 o) { return (float)((const i64*)p)[o]; }
__device__ __forceinline__ float ld1_i32(const void* p, i64 o) { return (float)((const int*)p)[o]; }
__device__ __forceinline__ void st1_f32(void* p, i64 o, float v) { ((float*)p)[o] = v; }
__device__ __forceinline__ void st1_bf16(void* p, i64 o, float v) { ((u16*)p)[o] = (u16)(f2bf2(v, 0.f) & 0xffffu); }
__device__ __forceinline__ void st1_f16(void* p, i64 o, float v) { ((u16*)p)[o] = (u16)(f2h2(v, 0.f) & 0xffffu); }
__device__ __forceinline__ void st1_bool(void* p, i64 o, float v) { ((u8*)p)[o] = v != 0.f ? 1 : 0; }
// W consecutive elements -> fp32
template <int W> __device__ __forceinline__ void ldv_f32(const void* p, i64 o, float* r) { const float* q = (const float*)p + o;
  if (W == 8) { V4 a = ldg16(q), b = ldg16(q + 4); r[0]=__uint_as_float(a.x); r[1]=__uint_as_float(a.y); r[2]=__uint_as_float(a.z); r[3]=__uint_as_float(a.w); r[4]=__uint_as_float(b.x); r[5]=__uint_as_float(b.y); r[6]=__uint_as_float(b.z); r[7]=__uint_as_float(b.w); }
  else if (W == 4) { V4 a = ldg16(q); r[0]=__uint_as_float(a.x); r[1]=__uint_as_float(a.y); r[2]=__uint_as_float(a.z); r[3]=__uint_as_float(a.w); }
  else if (W == 2) { V2 a = ldg8(q); r[0]=__uint_as_float(a.x); r[1]=__uint_as_float(a.y); }
  else r[0] = q[0]; }
template <int W> __device__ __forceinline__ void ldv_bf16(const void* p, i64 o, float* r) { const u16* q = (const u16*)p + o;
  if (W == 8) { V4 a = ldg16(q); u32 w[4] = {a.x, a.y, a.z, a.w};
    #pragma unroll
    for (int i = 0; i < 4; ++i) { r[2*i] = __uint_as_float(w[i] << 16); r[2*i+1] = __uint_as_float(w[i] & 0xffff0000u); } }
  else if (W == 4) { V2 a = ldg8(q); r[0] = __uint_as_float(a.x << 16); r[1] = __uint_as_float(a.x & 0xffff0000u); r[2] = __uint_as_float(a.y << 16); r[3] = __uint_as_float(a.y & 0xffff0000u); }
  else if (W == 2) { u32 a = ldg4(q); r[0] = __uint_as_float(a << 16); r[1] = __uint_as_float(a & 0xffff0000u); }
  else r[0] = bf2f(q[0]); }
template <int W> __device__ __forceinline__ void ldv_f16(const void* p, i64 o, float* r) { const u16* q = (const u16*)p + o;
  if (W == 8) { V4 a = ldg16(q); u32 w[4] = {a.x, a.y, a.z, a.w};
    #pragma unroll
    for (int i = 0; i < 4; ++i) { r[2*i] = h2f(w[i] & 0xffffu); r[2*i+1] = h2f(w[i] >> 16); } }
  else if (W == 4) { V2 a = ldg8(q); r[0] = h2f(a.x & 0xffffu); r[1] = h2f(a.x >> 16); r[2] = h2f(a.y & 0xffffu); r[3] = h2f(a.y >> 16); }
  else if (W == 2) { u32 a = ldg4(q); r[0] = h2f(a & 0xffffu); r[1] = h2f(a >> 16); }
  else r[0] = h2f(q[0]); }
template <int W> __device__ __forceinline__ void ldv_bool(const void* p, i64 o, float* r) { const u8* q = (const u8*)p + o;
  #pragma unroll
  for (int i = 0; i < W; ++i) r[i] = q[i] != 0 ? 1.f : 0.f; }
template <int W> __device__ __forceinline__ void ldv_i64(const void* p, i64 o, float* r) { const i64* q = (const i64*)p + o;
  #pragma unroll
  for (int i = 0; i < W; ++i) r[i] = (float)q[i]; }
template <int W> __device__ __forceinline__ void ldv_i32(const void* p, i64 o, float* r) { const int* q = (const int*)p + o;
  #pragma unroll
  for (int i = 0; i < W; ++i) r[i] = (float)q[i]; }
template <int W> __device__ __forceinline__ void stv_f32(void* p, i64 o, const float* r) { float* q = (float*)p + o;
  if (W == 8) { V4 a = {__float_as_uint(r[0]), __float_as_uint(r[1]), __float_as_uint(r[2]), __float_as_uint(r[3])}, b = {__float_as_uint(r[4]), __float_as_uint(r[5]), __float_as_uint(r[6]), __float_as_uint(r[7])}; stg16(q, a); stg16(q + 4, b); }
  else if (W == 4) { V4 a = {__float_as_uint(r[0]), __float_as_uint(r[1]), __float_as_uint(r[2]), __float_as_uint(r[3])}; stg16(q, a); }
  else if (W == 2) { V2 a = {__float_as_uint(r[0]), __float_as_uint(r[1])}; stg8(q, a); }
  else q[0] = r[0]; }
template <int W> __device__ __forceinline__ void stv_bf16(void* p, i64 o, const float* r) { u16* q = (u16*)p + o;
  if (W == 8) { V4 a = {f2bf2(r[0], r[1]), f2bf2(r[2], r[3]), f2bf2(r[4], r[5]), f2bf2(r[6], r[7])}; stg16(q, a); }
  else if (W == 4) { V2 a = {f2bf2(r[0], r[1]), f2bf2(r[2], r[3])}; stg8(q, a); }
  else if (W == 2) stg4(q, f2bf2(r[0], r[1]));
  else q[0] = (u16)(f2bf2(r[0], 0.f) & 0xffffu); }
template <int W> __device__ __forceinline__ void stv_f16(void* p, i64 o, const float* r) { u16* q = (u16*)p + o;
  if (W == 8) { V4 a = {f2h2(r[0], r[1]), f2h2(r[2], r[3]), f2h2(r[4], r[5]), f2h2(r[6], r[7])}; stg16(q, a); }
  else if (W == 4) { V2 a = {f2h2(r[0], r[1]), f2h2(r[2], r[3])}; stg8(q, a); }
  else if (W == 2) stg4(q, f2h2(r[0], r[1]));
  else q[0] = (u16)(f2h2(r[0], 0.f) & 0xffffu); }
template <int W> __device__ __forceinline__ void stv_bool(void* p, i64 o, const float* r) { u8* q = (u8*)p + o;
  #pragma unroll
  for (int i = 0; i < W; ++i) q[i] = r[i] != 0.f ? 1 : 0; }
"""

_EXPR = dict(jit._EXPR)
_RED_ID = {"sum": "0.f", "prod": "1.f", "amax": "__int_as_float(0xff800000)", "amin": "__int_as_float(0x7f800000)"}
_RED_COMB = {"sum": "__fadd_rn({a}, {b})", "prod": "__fmul_rn({a}, {b})", "amax": "nanmax({a}, {b})", "amin": "nanmin({a}, {b})"}


def leaf_modes(P: Program, leaf_tensors) -> tuple:
    """Per leaf (mode, load dtype): mode 'v' (one aligned run of W elements), 'b' (one value for the whole item) or 's'
    (element by element).  leaf_tensors[i] is the runtime tensor of leaf i: its strides and alignment decide between
    'v' and 's', its dtype is the one the kernel loads."""
    W, vd = P.W, len(P.ext) - 1
    modes = []
    for lf, t in zip(P.leaves, leaf_tensors):
        dt = _LOAD_DT.get(getattr(t, "dtype", lf.dtype))
        if dt is None:
            raise RuntimeError(f"b200ex region: operand {lf.name} has unsupported dtype {t.dtype}")
        ks = [dict(c[0]).get(vd, 0) for c in lf.coords]
        if not any(ks) or W == 1:
            modes.append(("b" if not any(ks) else "s", dt))
            continue
        last = len(ks) - 1
        c = lf.coords[last]
        ok = ks[last] == 1 and not any(ks[:last]) and t.stride(last) == 1 and c[1] % W == 0 and not any(k % W for v, k in c[0] if v != vd)
        if ok:
            esz = _DT_BYTES[dt]
            align = min(16, W * esz)
            ok = t.data_ptr() % align == 0 and all((t.stride(i) * esz) % align == 0 for i in range(last) if t.shape[i] != 1)
        modes.append(("v" if ok else "s", dt))
    # strides are compiled into the kernel (shapes already are): index arithmetic folds to a few 32-bit operations
    out = []
    for (m, dt), lf, t in zip(modes, P.leaves, leaf_tensors):
        st = tuple(0 if t.shape[i] == 1 else int(t.stride(i)) for i in range(len(lf.coords)))
        out.append((m, dt, st))
    return tuple(out)


def _aff_c(a, names, cast="") -> str:
    parts = []
    for v, k in a[0]:
        parts.append(f"{names[v]}" if k == 1 else f"{k} * {names[v]}")
    if a[1] or not parts:
        parts.append(str(a[1]))
    return "(" + " + ".join(parts).replace("+ -", "- ") + ")"


class _Gen:
    """Shared pieces of the two kernel shapes."""

    def __init__(self, P: Program, modes: tuple):
        self.P, self.modes, self.W = P, modes, P.W
        self.vd = len(P.ext) - 1
        self.L: list = []
        # packed fp32x2 arithmetic (FADD2 / FMUL2 / FFMA2) halves the issue slots of add / sub / mul but wants aligned
        # register pairs.  Measured: it pays in kernels with registers to spare (RoPE forward 53.2 -> 51.2 us, LayerNorm
        # forward 35.8 -> 32.3 us) and costs occupancy in the big ones (RoPE backward 67.5 -> 69.8 us, double-buffered
        # row backwards +2..6 %), so those keep scalar arithmetic.
        self.packed_math = len(P.leaves) <= 12

    def head(self, n_in):
        return _HEAD % {"MAX_RANK": MAX_RANK, "NIN": max(n_in, 1), "NOUT": max(len(self.P.outputs), 1)}

    def cond_expr(self, ci, names, e=None):
        terms, lo, hi = self.P.conds[ci]
        lin = _aff_c((terms, 0), names)
        if e is not None and dict(terms).get(self.vd, 0):
            lin = f"({lin} + {dict(terms)[self.vd]} * {e})"
        return f"({lin} >= {lo} && {lin} < {hi})"

    def leaf_aff(self, k):
        """Element offset of leaf k as one affine form over the iteration variables (strides are compile-time)."""
        lf, st = self.P.leaves[k], self.modes[k][2]
        a = _ZERO
        for i, c in enumerate(lf.coords):
            a = _aff_add(a, _aff_scale(c, st[i]))
        return a

    def leaf_offset(self, k, names):
        a = self.leaf_aff(k)
        span = abs(a[1]) + sum(abs(c) * (self.P.ext[v] - 1) for v, c in a[0])
        if span < 2**31:
            return f"(i64){_aff_c(a, names)}"
        return _aff_c(a, [f"(i64){n}" for n in names])

    def leaf_estride(self, k):
        return str(dict(self.leaf_aff(k)[0]).get(self.vd, 0))

    def emit_elements(self, L, ind, W, ids, ref, leafref, predref, tail):
        """The per-element loop: values `ids` for every element of the chunk, two elements per trip.  A pair lives in
        one float2: add / sub / mul are sm_100's packed fp32x2 instructions (`FADD2` / `FFMA2` / `FMUL2`, bit-identical
        to the scalar round-to-nearest ops, half the issue slots), roundings to a 16-bit format convert the pair at
        once (`cvt.rn.bf16x2`), everything else runs per lane.  ref(a, lane) names an operand (lane ".x" / ".y" / ""),
        leafref(k, e) reads element e of leaf k, tail(lane, e) -> extra lines per element (stores, partial sums)."""
        pair = W >= 2
        packed = pair and self.packed_math  # float2 pairs cost registers: only where the kernel has them to spare
        lanes = [(".x", "e"), (".y", "e + 1")] if pair else [("", "e")]
        L.append(f"{ind}#pragma unroll")
        L.append(f"{ind}for (int e = 0; e < {W}; e += {len(lanes)}) {{")
        both = lambda a: f"make_float2({ref(a, '.x')}, {ref(a, '.y')})"
        for i in ids:
            op, a, b, c, imm = self.P.ssa[i]
            if pair and op in ("rnd_bf16", "rnd_f16"):
                L.append(f"{ind}  float2 t{i}; {op}_2({ref(a, '.x')}, {ref(a, '.y')}, t{i}.x, t{i}.y);")
                continue
            if packed and op in ("add", "mul"):
                L.append(f"{ind}  const float2 t{i} = __f{op}2_rn({both(a)}, {both(b)});")
                continue
            if packed and op == "sub":
                L.append(f"{ind}  const float2 t{i} = __ffma2_rn({both(b)}, make_float2(-1.f, -1.f), {both(a)});")
                continue
            exprs = []
            for suf, e in lanes:
                if op == "leaf":
                    exprs.append(leafref(imm, e))
                elif op == "const":
                    exprs.append(f"__uint_as_float(0x{imm:08x}u)")
                elif op == "pred":
                    exprs.append(predref(imm))
                else:
                    exprs.append(_EXPR[op].format(a=ref(a, suf), b=ref(b, suf), c=ref(c, suf)))
            if pair:
                L.append(f"{ind}  const float2 t{i} = make_float2({exprs[0]}, {exprs[1]});")
            else:
                L.append(f"{ind}  const float t{i} = {exprs[0]};")
        for suf, e in lanes:
            L.extend(f"{ind}  {line}" for line in tail(suf, e))
        L.append(f"{ind}}}")

    def emit_pairs16(self, L, ind, live, res, fits, dt, rawleaf, floatleaf, predref, outs):
        """Element loop of a flat kernel whose 16-bit values stay PACKED (two per 32-bit register) wherever the program
        allows.  thunder's 16-bit traces round after every op, so `rnd(a * b)`, `rnd(a + b)`, `rnd(a - b)` with a and b
        already representable in the format are ONE packed instruction on the raw words (`mul/add/sub.rn.bf16x2`): the
        fp32 result of such an op is exact (or differs from the exact one far below the format's rounding point), so one
        rounding of the exact result is what upcast-operate-round gives, bit for bit — without unpacking the operands
        or converting the result.  Negation flips two sign bits, a `where` on a uniform condition selects a word,
        outputs are stored from the words.  Everything else is computed per lane in fp32 as in emit_elements.

        dt: "bf16" | "f16"; rawleaf(k) -> V4 expression of leaf k's 8 packed elements or None; floatleaf(k, e) -> fp32
        element expression for leaves that are not packed; outs: [(kind "h"|"f", target array name, ssa id)]."""
        P = self.P
        ids = sorted(live)
        users: dict = {}
        R = lambda a: res[a]
        for i in ids:
            for x in P.ssa[i][1:4]:
                if x >= 0 and R(x)[0] == "v":
                    users.setdefault(R(x)[1], []).append(i)
        out_ids = {R(sid)[1] for _, _, sid in outs if R(sid)[0] == "v"}
        rnd = "rnd_" + dt
        hop = {"mul": f"hmul2_{dt}", "add": f"hadd2_{dt}", "sub": f"hsub2_{dt}"}

        def fits16(a):
            r = R(a)
            if r[0] == "k":
                f = struct.unpack("<f", struct.pack("<I", r[1]))[0]
                return f == f and float(torch.tensor(f, dtype=torch.bfloat16 if dt == "bf16" else torch.float16)) == f
            return dt in fits[r[1]]

        # mul / add / sub that only feed a rounding to dt and whose operands are exact in dt: absorbed by that rounding
        absorbed = set()
        for i in ids:
            op, a, b = P.ssa[i][0], P.ssa[i][1], P.ssa[i][2]
            if op in hop and i not in out_ids and users.get(i) and all(P.ssa[u][0] == rnd for u in users[i]) and fits16(a) and fits16(b):
                absorbed.add(i)
        L.append(f"{ind}#pragma unroll")
        L.append(f"{ind}for (int p = 0; p < 4; ++p) {{")
        hname, fname = {}, {}
        body: list = []

        def lit16(bits):
            f = struct.unpack("<f", struct.pack("<I", bits))[0]
            h = int(torch.tensor(f, dtype=torch.bfloat16 if dt == "bf16" else torch.float16).view(torch.int16)) & 0xFFFF
            return f"0x{h | (h << 16):08x}u"

        def H(a):  # packed word holding value a for this pair
            r = R(a)
            if r[0] == "k":
                return lit16(r[1])
            i = r[1]
            if i not in hname:
                pack = "f2bf2" if dt == "bf16" else "f2h2"
                body.append(f"const u32 h{i} = {pack}({F(a)}.x, {F(a)}.y);")  # exact: the value fits the format
                hname[i] = f"h{i}"
            return hname[i]

        def F(a):  # float2 holding value a for this pair
            r = R(a)
            if r[0] == "k":
                return f"make_float2(__uint_as_float(0x{r[1]:08x}u), __uint_as_float(0x{r[1]:08x}u))"
            i = r[1]
            if i not in fname:
                body.append(f"const float2 t{i} = unpk_{dt}({hname[i]});")
                fname[i] = f"t{i}"
            return fname[i]

        for i in ids:
            op, a, b, c, imm = P.ssa[i]
            if i in absorbed:
                continue
            if op == "leaf":
                raw = rawleaf(imm)
                if raw is not None:
                    body.append(f"const u32 h{i} = wsel({raw}, p);")
                    hname[i] = f"h{i}"
                else:
                    body.append(f"const float2 t{i} = make_float2({floatleaf(imm, '2 * p')}, {floatleaf(imm, '2 * p + 1')});")
                    fname[i] = f"t{i}"
            elif op == "const":
                continue  # literals are materialised where they are used
            elif op == "pred":
                body.append(f"const float2 t{i} = make_float2({predref(imm)}, {predref(imm)});")
                fname[i] = f"t{i}"
            elif op == rnd:
                src = R(a)
                if src[0] == "v" and src[1] in absorbed:
                    sop, sa, sb = P.ssa[src[1]][0], P.ssa[src[1]][1], P.ssa[src[1]][2]
                    body.append(f"const u32 h{i} = {hop[sop]}({H(sa)}, {H(sb)});")
                else:
                    pack = "f2bf2" if dt == "bf16" else "f2h2"
                    fa = F(a)
                    body.append(f"const u32 h{i} = {pack}({fa}.x, {fa}.y);")
                hname[i] = f"h{i}"
            elif op == "neg" and fits16(a) and (R(a)[0] == "k" or R(a)[1] in hname):
                body.append(f"const u32 h{i} = hneg2({H(a)});")
                hname[i] = f"h{i}"
            elif op == "where" and R(a)[0] == "v" and P.ssa[R(a)[1]][0] == "pred" and fits16(b) and fits16(c):
                body.append(f"const u32 h{i} = {predref(P.ssa[R(a)[1]][4])} != 0.f ? {H(b)} : {H(c)};")
                hname[i] = f"h{i}"
            else:
                if self.packed_math and op in ("add", "mul"):
                    body.append(f"const float2 t{i} = __f{op}2_rn({F(a)}, {F(b)});")
                elif self.packed_math and op == "sub":
                    body.append(f"const float2 t{i} = __ffma2_rn({F(b)}, make_float2(-1.f, -1.f), {F(a)});")
                else:
                    fa, fb, fc = (F(x) if x >= 0 else None for x in (a, b, c))
                    lane = lambda s_: _EXPR[op].format(a=f"{fa}{s_}" if fa else "", b=f"{fb}{s_}" if fb else "", c=f"{fc}{s_}" if fc else "")
                    body.append(f"const float2 t{i} = make_float2({lane('.x')}, {lane('.y')});")
                fname[i] = f"t{i}"
        for kind, target, sid in outs:
            if kind == "h":
                body.append(f"{target}[p] = {H(sid)};")
            else:
                fs = F(sid)
                body.append(f"{target}[2 * p] = {fs}.x; {target}[2 * p + 1] = {fs}.y;")
        L.extend(f"{ind}  {line}" for line in body)
        L.append(f"{ind}}}")

    def ssa_line(self, i, ref, leafref, predref):
        op, a, b, c, imm = self.P.ssa[i]
        if op == "leaf":
            return leafref(imm)
        if op == "const":
            return f"__uint_as_float(0x{imm:08x}u)"
        if op == "pred":
            return predref(imm)
        return _EXPR[op].format(a=ref(a), b=ref(b), c=ref(c))


def input_slots(P: Program):
    """Distinct input tensors of a program in first-use order, and the slot of every leaf."""
    names, slot = [], []
    for lf in P.leaves:
        if lf.name not in names:
            names.append(lf.name)
        slot.append(names.index(lf.name))
    return names, slot


def _fold(P: Program, assume: dict):
    """Simplify the program under decided conditions.  Returns (res, live, undecided): res[i] is ("k", bits) when
    value i is a known constant, ("a", j) when it is value j, else ("v", i); live = values still needed for the
    outputs; undecided = conditions the live part still depends on."""
    ONE, ZERO = _f32_bits(1.0), _f32_bits(0.0)
    res: list = []
    # conditions on the same linear form are ranges of one quantity: what is decided narrows that quantity's range,
    # which decides others (after "not [0, 64)" on a 128-wide dim, "[64, 128)" is known to hold)
    assume = dict(assume)
    by_terms: dict = {}
    for ci, (terms, lo, hi) in enumerate(P.conds):
        by_terms.setdefault(terms, []).append(ci)
    for terms, cis in by_terms.items():
        lo = sum(k * (P.ext[v] - 1) for v, k in terms if k < 0)
        hi = sum(k * (P.ext[v] - 1) for v, k in terms if k > 0)
        changed = True
        while changed:
            changed = False
            for ci in cis:
                if ci not in assume:
                    continue
                clo, chi = P.conds[ci][1], P.conds[ci][2]
                if assume[ci]:
                    nlo, nhi = max(lo, clo), min(hi, chi - 1)
                else:
                    nlo, nhi = (chi if clo <= lo < chi else lo), (clo - 1 if clo <= hi < chi else hi)
                if (nlo, nhi) != (lo, hi):
                    lo, hi, changed = nlo, nhi, True
        for ci in cis:
            if ci not in assume:
                clo, chi = P.conds[ci][1], P.conds[ci][2]
                if lo >= clo and hi < chi:
                    assume[ci] = True
                elif hi < clo or lo >= chi:
                    assume[ci] = False

    def R(x):
        return res[x] if x >= 0 else None

    NZERO = _f32_bits(-0.0)
    # which 16-bit formats every value is exactly representable in: rounding such a value again is the identity
    # (`-x` of a bf16 x, a `where` between bf16 values, a sum whose other operand is a zero pad, ...)
    fits: list = []
    small = {ZERO: {"bf16", "f16"}, NZERO: {"bf16", "f16"}, ONE: {"bf16", "f16"}, _f32_bits(-1.0): {"bf16", "f16"}}

    def F(x):
        r_ = res[x]
        return small.get(r_[1], set()) if r_[0] == "k" else fits[r_[1]]

    for i, (op, a, b, c, imm) in enumerate(P.ssa):
        r = ("v", i)
        fit: set = set()
        if op == "const":
            r = ("k", imm)
        elif op == "pred" and imm in assume:
            r = ("k", ONE if assume[imm] else ZERO)
        elif op == "pred":
            fit = {"bf16", "f16"}
        elif op == "leaf" and any(assume.get(ci) is False for ci in P.leaves[imm].guard):
            r = ("k", ZERO)
        elif op == "leaf":
            fit = {"bf16": {"bf16"}, "f16": {"f16"}, "bool": {"bf16", "f16"}}.get(_DT[P.leaves[imm].dtype], set())
        elif op in ("rnd_bf16", "rnd_f16"):
            kind = op[4:]
            if kind in F(a):
                r = R(a)
            fit = {kind}
        elif op in ("neg", "abs", "mov"):
            fit = set(F(a))
        elif op in ("maximum", "minimum"):
            fit = F(a) & F(b)
        elif op in COMPARE or op in ("not", "or", "sign"):
            fit = {"bf16", "f16"}
        elif op == "add" and (R(a) in (("k", ZERO), ("k", NZERO)) or R(b) in (("k", ZERO), ("k", NZERO))):
            r = R(b) if R(a) in (("k", ZERO), ("k", NZERO)) else R(a)  # x + 0 (a zero pad on the other side of a cat)
        elif op == "and":
            ra, rb = R(a), R(b)
            if ra == ("k", ZERO) or rb == ("k", ZERO):
                r = ("k", ZERO)
            elif ra == ("k", ONE):
                r = rb
            elif rb == ("k", ONE):
                r = ra
        elif op == "where":
            rc = R(a)
            if rc[0] == "k":
                r = R(b) if rc[1] != ZERO and rc[1] != NZERO else R(c)
            fit = F(b) & F(c)
        if op == "and":
            fit = {"bf16", "f16"}
        res.append(r)
        fits.append(F(i) if r != ("v", i) else fit)
    live, undecided = set(), set()
    stack = [o.sid for o in P.outputs]
    while stack:
        i = stack.pop()
        r = res[i]
        if r[0] == "k":
            continue
        i = r[1]
        if i in live:
            continue
        live.add(i)
        op, a, b, c, imm = P.ssa[i]
        if op == "pred":
            undecided.add(imm)
        elif op == "leaf":
            undecided.update(ci for ci in P.leaves[imm].guard if ci not in assume)
        stack.extend(x for x in (a, b, c) if x >= 0)
    _fold.last_fits = fits  # per-value set of 16-bit formats the value is exactly representable in (for this call)
    return res, live, undecided


MAX_BLOCKS = 48


def generate_flat(P: Program, modes: tuple, unroll: int, min_ctas: int = 0) -> str:
    """Flat (gather) kernel.  Conditions are uniform over a work item, so `cat` / `pad` selections become control
    flow: the program is specialised per reachable combination of conditions and every specialisation loads only
    the leaves it needs (register pressure is the widest branch, not the sum of all pieces)."""
    G = _Gen(P, modes)
    in_names, G.slot = input_slots(P)
    W, ext, nv = P.W, P.ext, len(P.ext)
    L = [G.head(len(in_names))]
    total = math.prod(ext) // W
    it = "u32" if math.prod(ext) < 2**31 else "i64"
    _, _, und0 = _fold(P, {})
    U = unroll
    bounds = f"256, {min_ctas}" if min_ctas else "256"
    L.append(f'extern "C" __global__ void __launch_bounds__({bounds}) rk_main(const __grid_constant__ Params P) {{')
    L.append(f"  const {it} TOTAL = {total};")
    L.append(f"  for ({it} base = ({it})blockIdx.x * {256 * U} + threadIdx.x; base < TOTAL; base += ({it})gridDim.x * {256 * U}) {{")
    L.append(f"    {it} x[{U}][{max(nv, 1)}]; bool live[{U}];")
    if P.conds:
        L.append(f"    bool cn[{U}][{len(P.conds)}];")
    L.append("    #pragma unroll")
    L.append(f"    for (int u = 0; u < {U}; ++u) {{")
    L.append(f"      {it} r = base + ({it})u * 256; live[u] = r < TOTAL; if (!live[u]) r = 0;")
    order = P.order if P.order is not None else [(k, "all", ext[k]) for k in range(nv - 1, -1, -1)]
    lo_size = {k: n for k, part, n in order if part == "lo"}
    for pos, (k, part, n) in enumerate(order):  # fastest first
        e = n // W if k == nv - 1 else n
        mul = f" * {W}" if (k == nv - 1 and W > 1) else ""
        val = "r" if pos == len(order) - 1 else f"(r % {e})"
        step = "" if pos == len(order) - 1 else f" r /= {e};"
        if part == "hi":
            L.append(f"      x[u][{k}] += {val} * {lo_size[k]};{step}")
        else:
            L.append(f"      x[u][{k}] = {val}{mul};{step}")
    inames = [f"(int)x[u][{k}]" for k in range(nv)] if it == "u32" else [f"x[u][{k}]" for k in range(nv)]
    for ci in range(len(P.conds)):
        L.append(f"      cn[u][{ci}] = {G.cond_expr(ci, inames)};")
    L.append("    }")
    # linear offset of the item in every (fresh, contiguous) output
    lin, stride = [], 1
    cast = "" if math.prod(ext) < 2**31 else "(i64)"
    for k in range(nv - 1, -1, -1):
        lin.append(f"{cast}x[u][{k}] * {stride}u" if stride != 1 else f"{cast}x[u][{k}]")
        stride *= ext[k]
    out_off = "(i64)(" + (" + ".join(lin) if lin else "0") + ")"

    def load(k, ind, dst):
        lf = P.leaves[k]
        dt, ptr, off = modes[k][1], f"P.in[{G.slot[k]}].ptr", G.leaf_offset(k, inames)
        if modes[k][0] == "v":
            return f"{ind}ldv_{dt}<{W}>({ptr}, {off}, {dst});"
        if modes[k][0] == "b":
            return f"{ind}{{ const float b = ld1_{dt}({ptr}, {off}); for (int e = 0; e < {W}; ++e) {dst}[e] = b; }}"
        return f"{ind}{{ const i64 o = {off}, es = {G.leaf_estride(k)}; for (int e = 0; e < {W}; ++e) {dst}[e] = ld1_{dt}({ptr}, o + e * es); }}"

    # 16-bit values stay packed when the kernel works on one 16-bit format with full 16-byte items (emit_pairs16)
    sixteen = {modes[k][1] for k in range(len(P.leaves)) if modes[k][1] in ("bf16", "f16")} | {_DT[o.dtype] for o in P.outputs if _DT[o.dtype] in ("bf16", "f16")}
    dt16 = next(iter(sixteen)) if (len(sixteen) == 1 and W == 8 and os.environ.get("B200EX_PACKED16", "1") != "0") else None
    is_raw = lambda k: dt16 is not None and modes[k][0] == "v" and modes[k][1] == dt16

    def math_and_store(res, live, ind, leafref, rawref):
        if dt16 is not None:
            outs = []
            for k, o in enumerate(P.outputs):
                if _DT[o.dtype] == dt16:
                    L.append(f"{ind}u32 ow{k}[4];")
                    outs.append(("h", f"ow{k}", o.sid))
                else:
                    L.append(f"{ind}float o{k}[{W}];")
                    outs.append(("f", f"o{k}", o.sid))
            G.emit_pairs16(L, ind, live, res, _fold.last_fits, dt16, lambda k: rawref(k) if is_raw(k) else None, leafref,
                           lambda ci: f"(cn[u][{ci}] ? 1.f : 0.f)", outs)
            for k, o in enumerate(P.outputs):
                if _DT[o.dtype] == dt16:
                    L.append(f"{ind}stg16((u16*)P.out[{k}].ptr + {out_off}, V4{{ow{k}[0], ow{k}[1], ow{k}[2], ow{k}[3]}});")
                else:
                    L.append(f"{ind}stv_{_DT[o.dtype]}<{W}>(P.out[{k}].ptr, {out_off}, o{k});")
            return
        for k in range(len(P.outputs)):
            L.append(f"{ind}float o{k}[{W}];")

        def ref(a, suf):
            r = res[a]
            return f"__uint_as_float(0x{r[1]:08x}u)" if r[0] == "k" else f"t{r[1]}{suf}"

        tail = lambda suf, e: [f"o{k}[{e}] = {ref(o.sid, suf)};" for k, o in enumerate(P.outputs)]
        G.emit_elements(L, ind, W, sorted(live), ref, leafref, lambda ci: f"(cn[u][{ci}] ? 1.f : 0.f)", tail)
        for k, o in enumerate(P.outputs):
            L.append(f"{ind}stv_{_DT[o.dtype]}<{W}>(P.out[{k}].ptr, {out_off}, o{k});")

    def load_raw(k, ind, dst):
        return f"{ind}{dst} = ldg16((const u16*)P.in[{G.slot[k]}].ptr + {G.leaf_offset(k, inames)});"

    if not und0:
        # one straight-line block: all loads of all U items first (every request in flight), then math and stores
        res, live, _ = _fold(P, {})
        fits0 = _fold.last_fits
        used = sorted(P.ssa[i][4] for i in live if P.ssa[i][0] == "leaf")
        if any(not is_raw(k) for k in used):
            L.append(f"    float v[{U}][{len(P.leaves)}][{W}];")
        if any(is_raw(k) for k in used):
            L.append(f"    V4 rw[{U}][{len(P.leaves)}];")
        L.append("    #pragma unroll")
        L.append(f"    for (int u = 0; u < {U}; ++u) if (live[u]) {{")
        for k in used:
            L.append(load_raw(k, "      ", f"rw[u][{k}]") if is_raw(k) else load(k, "      ", f"v[u][{k}]"))
        L.append("    }")
        L.append("    #pragma unroll")
        L.append(f"    for (int u = 0; u < {U}; ++u) if (live[u]) {{")
        _fold.last_fits = fits0
        math_and_store(res, live, "      ", lambda k, e: f"v[u][{k}][{e}]", lambda k: f"rw[u][{k}]")
        L.append("    }")
    else:
        blocks = [0]

        def tree(assume, ind):
            res, live, und = _fold(P, assume)
            fits_here = _fold.last_fits
            if not und:
                blocks[0] += 1
                if blocks[0] > MAX_BLOCKS:
                    raise PwUnsupported("too many condition combinations for one kernel")
                L.append(f"{ind}{{")
                for k in sorted(P.ssa[i][4] for i in live if P.ssa[i][0] == "leaf"):
                    if is_raw(k):
                        L.append(f"{ind}  V4 r{k};")
                        L.append(load_raw(k, ind + "  ", f"r{k}"))
                    else:
                        L.append(f"{ind}  float v{k}[{W}];")
                        L.append(load(k, ind + "  ", f"v{k}"))
                _fold.last_fits = fits_here
                math_and_store(res, live, ind + "  ", lambda k, e: f"v{k}[{e}]", lambda k: f"r{k}")
                L.append(f"{ind}}}")
                return
            ci = min(und)
            L.append(f"{ind}if (cn[u][{ci}]) {{")
            tree({**assume, ci: True}, ind + "  ")
            L.append(f"{ind}}} else {{")
            tree({**assume, ci: False}, ind + "  ")
            L.append(f"{ind}}}")

        L.append("    { const int u = 0; if (live[0]) {")
        tree({}, "      ")
        L.append("    } }")
    L.append("  }")
    L.append("}")
    return "\n".join(L)


def _leaf_sids(P: Program) -> dict:
    return {s[4]: i for i, s in enumerate(P.ssa) if s[0] == "leaf"}


def row_geometry(P: Program, staged: bool = False):
    """(threads per row, chunks per thread, rows per CTA, block size).  A row lives in the registers of T threads,
    each owning C interleaved chunks of W elements; the more operands stay resident, the fewer chunks per thread.
    staged: the per-row operands wait in the shared-memory ring (row_ring), so only the row-independent operands and
    the column accumulators (which also bring the most arithmetic: counted twice) hold registers — measured: the
    LayerNorm forward at 4 chunks x 128 threads 25.7 us against 28.0 us at 2 x 256, while kernels with column sums
    spill there and lose."""
    R, W = P.row_len, P.W
    items = (R + W - 1) // W
    sid = _leaf_sids(P)
    ncols = sum(1 for o in P.outputs if o.kind == "col")
    resident = sum(1 for k in range(len(P.leaves)) if P.full[sid[k]]) + ncols
    if staged:
        resident = len(_row_operands(P)[0]) + 2 * ncols
    cmax = 4 if resident <= 2 else (2 if resident <= 5 else 1)
    if resident > 5:
        # many operands, but 16-bit ones stay packed (4 registers per 8 elements): two chunks per thread still fit
        # next to the second (prefetch) buffer, and a 256-thread CTA leaves room for two CTAs per SM
        no = P.n_outer
        per_row = [k for k in range(len(P.leaves)) if P.full[sid[k]] and any(v < no for c in P.leaves[k].coords for v, _ in c[0])]
        cost = sum(4 if (P.leaves[k].dtype in (torch.bfloat16, torch.float16) and W == 8) else W for k in per_row)
        fixed = (resident - len(per_row)) * W
        if 2 * (2 * cost + fixed) <= 112:
            cmax = 2
    T = 32
    while T < 1024 and (items + T - 1) // T > cmax:
        T *= 2
    C = (items + T - 1) // T
    if resident * C * W > MAX_ROW_FLOATS:
        raise PwUnsupported(f"row of {R} elements with {resident} resident operands does not fit in registers")
    block = T if T > 32 else 128  # block-wide barriers need every thread of the CTA on the same row
    return T, C, block // T, block


def _row_operands(P: Program):
    """(hoisted, per_row, scalars): leaf indices that are row-independent full operands, per-row full operands, and
    per-row scalars."""
    no = P.n_outer
    leaf_sid = _leaf_sids(P)

    def uses_outer(lf):
        vs = {v for c in lf.coords for v, _ in c[0]} | {v for ci in lf.guard for v, _ in P.conds[ci][0]}
        return any(v < no for v in vs)

    full_leaves = [k for k in range(len(P.leaves)) if P.full[leaf_sid[k]]]
    hoisted = [k for k in full_leaves if not uses_outer(P.leaves[k])]
    per_row = [k for k in full_leaves if k not in hoisted]
    scalars = [k for k in range(len(P.leaves)) if k not in full_leaves]
    return hoisted, per_row, scalars


_ES = {"bf16": 2, "f16": 2, "f32": 4}
# B200EX_ROW_RING: "auto" (default: time the builds on the first call), "1" (always the ring, widest geometry), "v1" (the
# ring at the register geometry), "0" (never)
_RING_MODE = os.environ.get("B200EX_ROW_RING", "auto")


def row_ring(P: Program, modes: tuple, staged: bool = True):
    """(stages, 16-byte units per thread and stage, dynamic shared-memory bytes) of the row kernel's staging ring, or
    (0, 0, 0) when the kernel keeps its rows in registers.

    A row kernel at two 256-thread CTAs per SM that prefetches one row ahead in registers has 2 x 8 KB per operand in
    flight per SM — a quarter of what HBM3e needs to stay busy (ncu: the LayerNorm forward ran at 2 TB/s with its warps
    waiting on the loads).  With the ring every thread copies its own chunks of the next stages - 1 rows into its own
    shared-memory slots (cp.async, 16 bytes each) and reads them back itself, so no barrier is involved and the depth
    is bounded by shared memory, not registers."""
    if _RING_MODE == "0" or P.kind != "row":
        return 0, 0, 0
    hoisted, per_row, _ = _row_operands(P)
    T, C, RPC, block = row_geometry(P, staged=staged)
    cols = sum(1 for o in P.outputs if o.kind == "col")
    nph = max(P.phase) + 1
    if not per_row or not (len(per_row) >= 3 or cols or nph >= 3):
        return 0, 0, 0
    units = 0
    for k in per_row:
        mode, dt = modes[k][0], modes[k][1]
        if mode != "v" or dt not in _ES or (P.W * _ES[dt]) % 16:
            return 0, 0, 0
        units += C * (P.W * _ES[dt] // 16)
    stage = units * block * 16
    budget = 200 * 1024 // max(1, 512 // block)
    nb = min(6, max(3, -(-32 * 1024 // stage) + 1), budget // stage)
    if nb < 2:
        return 0, 0, 0
    return nb, units, nb * stage


def generate_row(P: Program, modes: tuple, variant: int = 2) -> str:
    """Row-persistent kernel.  One row lives in the registers of T threads (a warp, or a whole CTA), C interleaved
    chunks of W elements per thread.  16-bit operands stay packed (one 16-byte register quad per chunk) and are
    unpacked where they are used, which halves their register cost and pays for a second buffer: the next row's loads
    are issued before the current row's reductions, so the barriers and shuffles of one row overlap the memory
    latency of the next.  Values on the far side of a reduction are recomputed from the resident operands."""
    G = _Gen(P, modes)
    in_names, G.slot = input_slots(P)
    W, ext, nv, no = P.W, P.ext, len(P.ext), P.n_outer
    R, NR = P.row_len, P.n_rows
    # variant 0: rows prefetched in registers; 1: staged through the ring at the register geometry; 2: ring with the
    # geometry that assumes the per-row operands hold no registers (row_geometry(staged=True))
    NB, ring_units, _ = row_ring(P, modes, variant == 2) if variant else (0, 0, 0)
    ring = NB > 0
    T, C, RPC, block = row_geometry(P, staged=ring and variant == 2)
    cols = [k for k, o in enumerate(P.outputs) if o.kind == "col"]
    nph = max(P.phase) + 1
    reduces = [i for i, s in enumerate(P.ssa) if s[0].startswith("reduce_")]
    leaf_sid = _leaf_sids(P)
    n_inner = nv - no
    cond_cls = []
    for terms, _, _ in P.conds:
        o, i = any(v < no for v, _ in terms), any(v >= no for v, _ in terms)
        cond_cls.append("mixed" if (o and i) else ("inner" if i else "outer"))

    def uses_outer(lf):
        vs = {v for c in lf.coords for v, _ in c[0]} | {v for ci in lf.guard for v, _ in P.conds[ci][0]}
        return any(v < no for v in vs)

    full_leaves = [k for k in range(len(P.leaves)) if P.full[leaf_sid[k]]]
    hoisted = [k for k in full_leaves if not uses_outer(P.leaves[k])]
    per_row = [k for k in full_leaves if k not in hoisted]
    scalars = [k for k in range(len(P.leaves)) if k not in full_leaves]
    packed = {k for k in per_row if modes[k][0] == "v" and modes[k][1] in ("bf16", "f16") and W == 8}
    cost = sum(C * (4 if k in packed else W) for k in per_row)
    fixed = (len(hoisted) + len(cols)) * C * W
    # the second buffer pays when a row's critical path is long (several operands, a column sum, several reductions);
    # a two-operand single-reduction row (RMSNorm forward) is already at the HBM limit without it
    worth = len(per_row) >= 3 or bool(cols) or nph >= 3
    nbuf = 2 if (worth and 2 * cost + fixed <= 150 and os.environ.get("B200EX_ROW_PREFETCH", "1") != "0") else 1
    if ring:
        nbuf = 1  # rows wait in shared memory, not in a second set of registers
        # 16-byte unit of (leaf k, chunk c, part j) inside a stage: unit index * block + threadIdx.x
        ring_base, ring_n16, u_ = {}, {}, 0
        for k in per_row:
            ring_n16[k] = W * _ES[modes[k][1]] // 16
            ring_base[k] = u_
            u_ += C * ring_n16[k]
    G.packed_math = nbuf == 1  # float2 arithmetic where no second register buffer competes for the registers

    names_i = lambda b: [f"xo{b}[{k}]" for k in range(no)] + [f"xi[c][{k - no}]" for k in range(no, nv)]
    names_r = lambda b: [f"xo{b}[{k}]" for k in range(no)] + ["0"] * n_inner

    def cref(ci, b):
        return {"outer": f"cr{b}_{ci}", "inner": f"ci_[c][{ci}]", "mixed": f"cm{b}_[c][{ci}]"}[cond_cls[ci]]

    def load_full(k, b, ind, dst):
        lf = P.leaves[k]
        dt, ptr = modes[k][1], f"P.in[{G.slot[k]}].ptr"
        g = " && ".join(["lc[c]"] + [cref(ci, b) for ci in lf.guard])
        off = G.leaf_offset(k, names_i(b))
        out = [f"{ind}#pragma unroll", f"{ind}for (int c = 0; c < C_; ++c) {{"]
        if k in packed:
            out.append(f"{ind}  {dst}[c] = ({g}) ? ldg16((const u16*){ptr} + {off}) : V4{{0u, 0u, 0u, 0u}};")
        elif modes[k][0] == "v":
            out.append(f"{ind}  if ({g}) ldv_{dt}<W_>({ptr}, {off}, {dst}[c]); else {{ for (int e = 0; e < W_; ++e) {dst}[c][e] = 0.f; }}")
        elif modes[k][0] == "b":
            out.append(f"{ind}  {{ const float b_ = ({g}) ? ld1_{dt}({ptr}, {off}) : 0.f; for (int e = 0; e < W_; ++e) {dst}[c][e] = b_; }}")
        else:
            out.append(f"{ind}  if ({g}) {{ const i64 o = {off}, es = {G.leaf_estride(k)}; for (int e = 0; e < W_; ++e) {dst}[c][e] = ld1_{dt}({ptr}, o + e * es); }} else {{ for (int e = 0; e < W_; ++e) {dst}[c][e] = 0.f; }}")
        out.append(f"{ind}}}")
        return out

    def leaf_ref(k, b, e="e"):
        if k in hoisted:
            return f"vh{k}[c][{e}]"
        if k in packed:
            return f"unp_{modes[k][1]}(z{k}, {e})"
        if ring and k in per_row:
            return f"zf{k}[{e}]"
        return f"v{b}_{k}[c][{e}]"

    # identities known without any condition decided (re-rounding an exact value, x + 0): res[i] names what value i is
    res, _, _ = _fold(P, {})
    rid = lambda a: res[a][1] if res[a][0] == "v" else None  # resolved ssa id, None for a constant

    L = [G.head(len(in_names))]
    L.append("__device__ __forceinline__ float unp_bf16(const V4& v, int e) { const u32 w = (e >> 1) == 0 ? v.x : ((e >> 1) == 1 ? v.y : ((e >> 1) == 2 ? v.z : v.w)); return (e & 1) ? __uint_as_float(w & 0xffff0000u) : __uint_as_float(w << 16); }")
    L.append("__device__ __forceinline__ float unp_f16(const V4& v, int e) { const u32 w = (e >> 1) == 0 ? v.x : ((e >> 1) == 1 ? v.y : ((e >> 1) == 2 ? v.z : v.w)); return (e & 1) ? h2f(w >> 16) : h2f(w & 0xffffu); }")
    L.append(f"#define T_ {T}\n#define C_ {C}\n#define W_ {W}\n#define R_ {R}")
    # at most 128 registers per thread: two 256-thread CTAs (or four 128-thread ones) stay resident per SM
    L.append(f'extern "C" __global__ void __launch_bounds__({block}, {max(1, 512 // block)}) rk_main(const __grid_constant__ Params P) {{')
    if T > 32:
        L.append(f"  __shared__ float red_s[{max(len(reduces), 1)}][{T // 32}];")
    L.append("  const int tr = threadIdx.x % T_, rslot = threadIdx.x / T_;")
    for k in cols:
        L.append(f"  float acc{k}[C_][W_];")
        L.append(f"  for (int c = 0; c < C_; ++c) for (int e = 0; e < W_; ++e) acc{k}[c][e] = 0.f;")
    # position of every chunk this thread owns inside the row (does not depend on the row)
    L.append(f"  int xi[C_][{max(n_inner, 1)}]; bool lc[C_];")
    if "inner" in cond_cls:
        L.append(f"  bool ci_[C_][{len(P.conds)}];")
    L.append("  #pragma unroll")
    L.append("  for (int c = 0; c < C_; ++c) {")
    L.append("    int r = (tr + c * T_) * W_; lc[c] = r < R_; if (!lc[c]) r = 0;")
    for k in range(nv - 1, no - 1, -1):
        if k == no:
            L.append(f"    xi[c][{k - no}] = r;")
        else:
            L.append(f"    xi[c][{k - no}] = r % {ext[k]}; r /= {ext[k]};")
    for ci, cls in enumerate(cond_cls):
        if cls == "inner":
            L.append(f"    ci_[c][{ci}] = {G.cond_expr(ci, names_i(0))};")
    L.append("  }")
    # operands that do not depend on the row (weights, biases) are loaded once per thread
    for k in hoisted:
        L.append(f"  float vh{k}[C_][W_];")
        was, packed_k = modes[k], k in packed
        L += load_full(k, 0, "  ", f"vh{k}")
    # per-buffer state of a row: coordinates, conditions, operands
    for b in range(nbuf):
        L.append(f"  int xo{b}[{max(no, 1)}];")
        for ci, cls in enumerate(cond_cls):
            if cls == "outer":
                L.append(f"  bool cr{b}_{ci} = false;")
        if "mixed" in cond_cls:
            L.append(f"  bool cm{b}_[C_][{len(P.conds)}];")
        if not ring:
            for k in per_row:
                L.append(f"  V4 w{b}_{k}[C_];" if k in packed else f"  float v{b}_{k}[C_][W_];")
        for k in scalars:
            L.append(f"  float s{b}_{k} = 0.f;")
    if ring:
        L.append("  extern __shared__ __align__(16) V4 ring_[];")
        L.append(f"  int xo1[{max(no, 1)}];  // coordinates of a row being prefetched")
        for ci, cls in enumerate(cond_cls):
            if cls == "outer":
                L.append(f"  bool cr1_{ci} = false;")
        if "mixed" in cond_cls:
            L.append(f"  bool cm1_[C_][{len(P.conds)}];")
        for k in scalars:
            L.append(f"  float s1_{k} = 0.f;  // the next row's")

    def emit_load(b, rowvar, ind, coords=True, scal=True, full=True, slot=None):
        """coords: the row's outer coordinates and conditions into set b; scal: its scalars; full: its operands — into
        registers, or (slot given) into ring stage `slot` with cp.async."""
        if coords:
            L.append(f"{ind}{{ i64 r = {rowvar};")
            for k in range(no - 1, -1, -1):
                if k == 0:
                    L.append(f"{ind}  xo{b}[0] = (int)r;")
                else:
                    L.append(f"{ind}  xo{b}[{k}] = (int)(r % {ext[k]}); r /= {ext[k]};")
            L.append(f"{ind}}}")
            for ci, cls in enumerate(cond_cls):
                if cls == "outer":
                    L.append(f"{ind}cr{b}_{ci} = {G.cond_expr(ci, names_r(b))};")
            if "mixed" in cond_cls:
                L.append(f"{ind}#pragma unroll")
                L.append(f"{ind}for (int c = 0; c < C_; ++c) {{")
                for ci, cls in enumerate(cond_cls):
                    if cls == "mixed":
                        L.append(f"{ind}  cm{b}_[c][{ci}] = {G.cond_expr(ci, names_i(b))};")
                L.append(f"{ind}}}")
        for k in scalars if scal else []:
            lf = P.leaves[k]
            g = " && ".join([f"cr{b}_{ci}" for ci in lf.guard]) or "true"
            L.append(f"{ind}s{b}_{k} = ({g}) ? ld1_{modes[k][1]}(P.in[{G.slot[k]}].ptr, {G.leaf_offset(k, names_r(b))}) : 0.f;")
        for k in per_row if full else []:
            if slot is None:
                L.extend(load_full(k, b, ind, f"w{b}_{k}" if k in packed else f"v{b}_{k}"))
                continue
            lf = P.leaves[k]
            es, ptr = _ES[modes[k][1]], f"P.in[{G.slot[k]}].ptr"
            g = " && ".join(["lc[c]"] + [cref(ci, b) for ci in lf.guard])
            L.append(f"{ind}#pragma unroll")
            L.append(f"{ind}for (int c = 0; c < C_; ++c) {{")
            L.append(f"{ind}  const bool ok_ = {g};")
            L.append(f"{ind}  const char* src_ = (const char*){ptr} + (ok_ ? (i64)({G.leaf_offset(k, names_i(b))}) * {es} : (i64)0);")
            for j in range(ring_n16[k]):
                L.append(f"{ind}  cpa16(ring_ + ((i64)({slot}) * {ring_units} + {ring_base[k]} + c * {ring_n16[k]} + {j}) * {block} + threadIdx.x, src_ + {16 * j}, ok_);")
            L.append(f"{ind}}}")

    def needed_for(targets):
        """Values that vary inside the row and must be evaluated in the element loop to produce `targets`."""
        need, stack = set(), list(targets)
        while stack:
            i = rid(stack.pop())
            if i is None or i in need or not P.full[i]:
                continue
            need.add(i)
            stack.extend(x for x in P.ssa[i][1:4] if x >= 0)
        return sorted(need)

    def konst(a):
        return f"__uint_as_float(0x{res[a][1]:08x}u)"

    def emit_body(b, rowvar, ind):
        L.append(f"{ind}{{")
        ind2 = ind + "  "
        emitted_row = set()

        def emit_row_scalars(upto_phase):
            for i, s_ in enumerate(P.ssa):
                if P.full[i] or i in emitted_row or P.phase[i] > upto_phase or s_[0].startswith("reduce_") or rid(i) != i:
                    continue
                qref = lambda a: konst(a) if rid(a) is None else f"q{rid(a)}"
                expr = G.ssa_line(i, qref, lambda k: f"s{b}_{k}", lambda ci: f"(cr{b}_{ci} ? 1.f : 0.f)")
                L.append(f"{ind2}const float q{i} = {expr};")
                emitted_row.add(i)

        def ref(a, suf=""):
            i = rid(a)
            if i is None:
                return konst(a)
            return f"t{i}{suf}" if P.full[i] else f"q{i}"

        for ph in range(nph):
            emit_row_scalars(ph)
            last = ph == nph - 1
            reds = [i for i in reduces if P.phase[P.ssa[i][1]] == ph]
            targets = [P.ssa[i][1] for i in reds]
            out_full = [(k, o) for k, o in enumerate(P.outputs) if o.kind in ("full", "col")] if last else []
            targets += [o.sid for _, o in out_full]
            need = needed_for(targets)
            for i in reds:
                L.append(f"{ind2}float p{i} = {_RED_ID[P.ssa[i][0][7:]]};")
            if need or out_full:
                L.append(f"{ind2}#pragma unroll")
                L.append(f"{ind2}for (int c = 0; c < C_; ++c) if (lc[c]) {{")
                for k, o in out_full:
                    if o.kind == "full":
                        L.append(f"{ind2}  float o{k}[W_];")
                used = {P.ssa[i][4] for i in need if P.ssa[i][0] == "leaf"}
                for k in sorted(used & set(per_row)) if ring else []:
                    unit = lambda j, k=k: f"st_[({ring_base[k]} + c * {ring_n16[k]} + {j}) * {block}]"
                    if k in packed:
                        L.append(f"{ind2}  V4 z{k} = {unit(0)}; asm volatile(\"\" : \"+r\"(z{k}.x), \"+r\"(z{k}.y), \"+r\"(z{k}.z), \"+r\"(z{k}.w));")
                    else:
                        L.append(f"{ind2}  float zf{k}[W_];")
                        for j in range(ring_n16[k]):
                            per = 16 // _ES[modes[k][1]]
                            assert modes[k][1] == "f32"
                            L.append(f"{ind2}  {{ const V4 a_ = {unit(j)}; zf{k}[{per * j}] = __uint_as_float(a_.x); zf{k}[{per * j + 1}] = __uint_as_float(a_.y); zf{k}[{per * j + 2}] = __uint_as_float(a_.z); zf{k}[{per * j + 3}] = __uint_as_float(a_.w); }}")
                for k in sorted(used & packed) if not ring else []:
                    # a fresh copy per phase that the compiler cannot relate to the other phases' copies: the operand
                    # stays packed across the reduction instead of being kept unpacked (twice the registers)
                    L.append(f"{ind2}  V4 z{k} = w{b}_{k}[c]; asm volatile(\"\" : \"+r\"(z{k}.x), \"+r\"(z{k}.y), \"+r\"(z{k}.z), \"+r\"(z{k}.w));")

                def tail(suf, e, reds=reds, out_full=out_full):
                    lines = [f"p{i} = {_RED_COMB[P.ssa[i][0][7:]].format(a=f'p{i}', b=ref(P.ssa[i][1], suf))};" for i in reds]
                    for k, o in out_full:
                        if o.kind == "full":
                            lines.append(f"o{k}[{e}] = {ref(o.sid, suf)};")
                        else:
                            lines.append(f"acc{k}[c][{e}] = __fadd_rn(acc{k}[c][{e}], {ref(o.sid, suf)});")
                    return lines

                G.emit_elements(L, ind2 + "  ", W, need, ref, lambda k, e: leaf_ref(k, b, e), lambda ci: f"({cref(ci, b)} ? 1.f : 0.f)", tail)
                for k, o in out_full:
                    if o.kind == "full":
                        L.append(f"{ind2}  stv_{_DT[o.dtype]}<W_>(P.out[{k}].ptr, {rowvar} * R_ + (tr + c * T_) * W_, o{k});")
                L.append(f"{ind2}}}")
            for i in reds:
                comb = _RED_COMB[P.ssa[i][0][7:]]
                L.append(f"{ind2}#pragma unroll")
                L.append(f"{ind2}for (int m = {min(T, 32) // 2}; m > 0; m >>= 1) p{i} = {comb.format(a=f'p{i}', b=f'__shfl_xor_sync(0xffffffffu, p{i}, m)')};")
                if T > 32:
                    L.append(f"{ind2}if ((threadIdx.x & 31) == 0) red_s[{reduces.index(i)}][threadIdx.x >> 5] = p{i};")
            if reds and T > 32:
                L.append(f"{ind2}__syncthreads();")
            for i in reds:
                if T > 32:
                    rs, comb = reduces.index(i), _RED_COMB[P.ssa[i][0][7:]]
                    L.append(f"{ind2}float q{i} = red_s[{rs}][0];")
                    L.append(f"{ind2}#pragma unroll")
                    L.append(f"{ind2}for (int w = 1; w < {T // 32}; ++w) q{i} = {comb.format(a=f'q{i}', b=f'red_s[{rs}][w]')};")
                else:
                    L.append(f"{ind2}const float q{i} = p{i};")
                emitted_row.add(i)
        emit_row_scalars(nph)
        for k, o in enumerate(P.outputs):
            if o.kind == "row":
                L.append(f"{ind2}if (tr == 0) st1_{_DT[o.dtype]}(P.out[{k}].ptr, {rowvar}, {ref(o.sid)});")
        if reduces and T > 32:
            L.append(f"{ind2}__syncthreads();  // red_s is reused by the next row")
        L.append(f"{ind}}}")

    L.append(f"  const i64 STEP = (i64)gridDim.x * {RPC};")
    L.append(f"  i64 row = (i64)blockIdx.x * {RPC} + rslot;")
    if ring:
        D = NB - 1
        L.append("  #pragma unroll")
        L.append(f"  for (int j = 0; j < {D}; ++j) {{  // the first {D} rows of this thread are on their way before any is used")
        L.append("    const i64 rp = row + j * STEP;")
        L.append(f"    if (rp < {NR}) {{")
        emit_load(1, "rp", "      ", scal=False, slot="j")
        L.append("    }")
        L.append("    cpa_commit();")
        L.append("  }")
        if scalars:
            L.append(f"  if (row < {NR}) {{")
            emit_load(1, "row", "    ", full=False)
            L.append("  }")
        L.append("  int slot = 0;")
        L.append(f"  for (; row < {NR}; row += STEP) {{")
        L.append(f"    {{ const i64 rp = row + {D} * STEP;  // refill the stage consumed one row ago")
        L.append(f"      if (rp < {NR}) {{")
        emit_load(1, "rp", "        ", scal=False, slot=f"(slot == 0 ? {D} : slot - 1)")
        L.append("      }")
        L.append("      cpa_commit();")
        L.append("    }")
        if scalars:
            for k in scalars:
                L.append(f"    s0_{k} = s1_{k};")
            L.append(f"    {{ const i64 rn = row + STEP; if (rn < {NR}) {{")
            emit_load(1, "rn", "      ", full=False)
            L.append("    } }")
        L.append(f"    cpa_wait<{D}>();  // this row's copies (the oldest group) have landed")
        emit_load(0, "row", "    ", scal=False, full=False)
        L.append(f"    const V4* st_ = ring_ + (i64)slot * {ring_units * block} + threadIdx.x;")
        emit_body(0, "row", "    ")
        L.append(f"    slot = slot == {D} ? 0 : slot + 1;")
        L.append("  }")
    elif nbuf == 1:
        L.append(f"  for (; row < {NR}; row += STEP) {{")
        emit_load(0, "row", "    ")
        emit_body(0, "row", "    ")
        L.append("  }")
    else:
        L.append(f"  if (row < {NR}) {{")
        emit_load(0, "row", "    ")
        L.append("  }")
        L.append(f"  while (row < {NR}) {{")
        L.append("    const i64 nxt = row + STEP;")
        L.append(f"    if (nxt < {NR}) {{")
        emit_load(1, "nxt", "      ")
        L.append("    }")
        emit_body(0, "row", "    ")
        L.append(f"    if (nxt >= {NR}) break;")
        L.append("    row = nxt + STEP;")
        L.append(f"    if (row < {NR}) {{")
        emit_load(0, "row", "      ")
        L.append("    }")
        emit_body(1, "nxt", "    ")
        L.append("  }")
    if cols:
        # partial column sums of every row slot of this CTA: scratch[col output][part][column]
        L.append("  #pragma unroll")
        L.append("  for (int c = 0; c < C_; ++c) if (lc[c]) {")
        for j_, k in enumerate(cols):
            L.append(f"    stv_f32<W_>(P.scratch, ((i64){j_} * P.nparts + (i64)blockIdx.x * {RPC} + rslot) * R_ + (tr + c * T_) * W_, acc{k}[c]);")
        L.append("  }")
    L.append("}")
    if cols:
        # fold of the per-CTA partials: 32 columns x 8 slices of the partials per CTA, fixed order (deterministic)
        L.append('extern "C" __global__ void __launch_bounds__(256) rk_fold(const __grid_constant__ Params P) {')
        L.append("  __shared__ float part[8][33];")
        L.append("  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5, col = blockIdx.x * 32 + lane;")
        for j_, k in enumerate(cols):
            dt = _DT[P.outputs[k].dtype]
            L.append(f"  {{ float s = 0.f; const float* p = P.scratch + (i64){j_} * P.nparts * R_ + col;")
            L.append("    if (col < R_) for (int i = g; i < P.nparts; i += 8) s = __fadd_rn(s, p[(i64)i * R_]);")
            L.append("    part[g][lane] = s; __syncthreads();")
            L.append("    if (g == 0 && col < R_) { float t = part[0][lane];")
            L.append("      #pragma unroll")
            L.append("      for (int w = 1; w < 8; ++w) t = __fadd_rn(t, part[w][lane]);")
            L.append(f"      st1_{dt}(P.out[{k}].ptr, col, t); }}")
            L.append("    __syncthreads(); }")
        L.append("}")
    return "\n".join(L)


# --------------------------------------------------------------------------------------------------
# Launch
# --------------------------------------------------------------------------------------------------
class _Op(ctypes.Structure):
    _fields_ = [("ptr", ctypes.c_void_p), ("st", ctypes.c_int64 * MAX_RANK)]


def _params_type(n_in, n_out):
    class Params(ctypes.Structure):
        _fields_ = [("inp", _Op * max(n_in, 1)), ("out", _Op * max(n_out, 1)), ("scratch", ctypes.c_void_p), ("nparts", ctypes.c_int32)]

    return Params


_kernels: dict = {}
_choice: dict = {}  # (program signature, leaf modes, device) -> build chosen by the first call's timing
_klock = threading.Lock()


class _Kernel:
    def __init__(self, cubin: bytes, has_fold: bool, smem: int = 0):
        from cuda.bindings import driver

        self.driver = driver
        self.smem = int(smem)  # dynamic shared memory of rk_main (the row kernels' staging ring)
        self.variant = 0
        (self.module,) = jit._check(driver.cuModuleLoadData(cubin), "cuModuleLoadData")
        (self.main,) = jit._check(driver.cuModuleGetFunction(self.module, b"rk_main"), "cuModuleGetFunction")
        if self.smem:
            jit._check(driver.cuFuncSetAttribute(self.main, driver.CUfunction_attribute.CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, self.smem), "cuFuncSetAttribute")
        self.fold = jit._check(driver.cuModuleGetFunction(self.module, b"rk_fold"), "cuModuleGetFunction")[0] if has_fold else None
        self.launches = 0
        self._occ: dict = {}

    def local_bytes(self) -> int:
        """Per-thread local memory of rk_main: non-zero means the register cap made it spill."""
        try:
            (n,) = jit._check(self.driver.cuFuncGetAttribute(self.driver.CUfunction_attribute.CU_FUNC_ATTRIBUTE_LOCAL_SIZE_BYTES, self.main), "cuFuncGetAttribute")
            return int(n)
        except Exception:
            return 1

    def ctas_per_sm(self, block: int) -> int:
        """Resident CTAs per SM for `rk_main` at this block size.  Persistent grids are sized with it: a grid that is
        not a whole number of resident waves leaves part of the GPU idle for the last one."""
        if block not in self._occ:
            try:
                (n,) = jit._check(self.driver.cuOccupancyMaxActiveBlocksPerMultiprocessor(self.main, block, self.smem), "cuOccupancyMaxActiveBlocksPerMultiprocessor")
                self._occ[block] = max(1, int(n))
            except Exception:
                self._occ[block] = 4
        return self._occ[block]

    def launch(self, func, params, grid, block, stream):
        jit.fast_launch(self.driver, func, grid, block, stream, ctypes.addressof(params), self.smem if func is self.main else 0)
        self.launches += 1


# Resident CTAs (256 threads) per SM a gather kernel's grid is sized for.  Measured on the RoPE forward (7 loads + 3
# stores of 16 bytes per thread, 32 registers): 8 CTAs per SM 59.3 us, 6 -> 51.3 us, 5 -> 53.2 us, 4 -> 57.1 us — past
# 1536 threads per SM the extra requests in flight only lengthen the DRAM queues (profiles/r2_bench_regionc.txt).
_GRID_CTAS = int(os.environ.get("B200EX_FLAT_GRIDCTAS", "6"))


def flat_unroll(P: Program) -> int:
    """Work items per thread and trip of a flat kernel.  Several only for straight-line programs (all loads of all
    items are issued first); a program with `cat` / `pad` branches takes one item at a time."""
    if _fold(P, {})[2]:
        return 1
    n = len(P.leaves)
    return 4 if n <= 2 else (2 if n <= 5 else 1)


def source_for(P: Program, modes: tuple, min_ctas: int = 0, variant: int = 2) -> str:
    return generate_flat(P, modes, flat_unroll(P), min_ctas) if P.kind == "flat" else generate_row(P, modes, variant)


def row_variants(P: Program, modes: tuple) -> list:
    """Builds of a row program worth timing against each other (see generate_row): [0] alone when the ring does not
    apply, else 0, 1 and — where it gives another geometry — 2."""
    if P.kind != "row" or not row_ring(P, modes, False)[0]:
        return [0]
    out = [0, 1]
    if row_ring(P, modes, True)[0] and row_geometry(P, True) != row_geometry(P, False):
        out.append(2)
    return out


def _get_kernel(P: Program, modes: tuple, dev_index: int, variant: int = -1) -> _Kernel:
    if P.kind != "row":
        variant = 0
    elif variant < 0 or variant not in row_variants(P, modes):
        vs = row_variants(P, modes)
        variant = 0 if _RING_MODE == "0" else (1 if (_RING_MODE == "v1" and 1 in vs) else vs[-1])
    key = (P.signature, modes, dev_index, variant)
    k = _kernels.get(key)
    if k is None:
        with _klock:
            k = _kernels.get(key)
            if k is None:
                with torch.cuda.device(dev_index):
                    k = None
                    if P.kind == "flat" and P.conds:
                        # Branchy gather kernels are issue-bound and short on registers by a few: ask for full
                        # occupancy (8 then 6 CTAs of 256 threads per SM) and keep the first build that does not spill.
                        for want in (8, 6, 5):
                            cand = _Kernel(jit.compile_cubin(source_for(P, modes, want)), False)
                            if cand.local_bytes() == 0:
                                k = cand
                                break
                    if k is None:
                        k = _Kernel(jit.compile_cubin(source_for(P, modes, variant=variant)), any(o.kind == "col" for o in P.outputs),
                                    row_ring(P, modes, variant == 2)[2] if variant else 0)
                        k.variant = variant
                _kernels[key] = k
    return k


def program_from_pointwise(spec) -> Program:
    """A b200ex.pointwise program (purely elementwise, leaves mapped onto the iteration dims) as a flat region program.
    PwLaunch uses it for launches its own vector skeleton cannot serve — inner extents that are not a multiple of 8,
    odd strides, misaligned views — which the gather kernel reads with narrower or element-wise loads instead of
    handing them to the ahead-of-time interpreter."""
    ext = [int(n) for n in spec.iter_shape]
    var = lambda d: _aff([(d, 1)])
    leaves, in_shapes = [], {}
    for (name, mapping), dt in zip(spec.leaves, spec.leaf_dtypes):
        leaves.append(Leaf(name, dt, tuple(var(d) for d in mapping), ()))
    ssa = [tuple(e) for e in spec.ssa]
    outputs = [Out(name, dt, tuple(ext), sid, "full") for (name, dt, _), sid in zip(spec.outputs, spec.out_ids)]
    P = Program("flat", ext, len(ext), leaves, [], ssa, outputs)
    P.W = 1
    if ext:
        P.W = 8
        while P.W > 1 and ext[-1] % P.W:
            P.W //= 2
    P.order = _thread_order(P)
    P.signature = "pw:" + spec.signature + ":" + ",".join(map(str, ext))
    return P


class RegionLaunch:
    """One generated kernel bound to a region: ``run(env)`` reads the input tensors by name and returns
    {output name: new tensor}.  Plans (leaf access modes, parameter block) are cached per runtime descriptor of the
    inputs — shape, strides, dtype, alignment class — like the reference's FusionDefinitionWrapper cache
    (nvfuserex_impl.py:505-521)."""

    MAX_PLANS = 512

    def __init__(self, launch: Launch):
        self.P = launch.program
        self.views = launch.views
        self.in_names, self.slot = input_slots(self.P)
        self.Params = _params_type(len(self.in_names), len(self.P.outputs))
        self._plans: dict = {}
        self.tuned: dict = {}  # plan key -> (ms in registers, ms through the ring) where both were timed
        self.kind = "b200k_region[row]" if self.P.kind == "row" else "b200k_region[gather]"

    def run(self, env: dict) -> dict:
        from .ops import _on
        from .pointwise import _sm_count

        P = self.P
        tensors = []
        for name in self.in_names:
            t = env.get(name)
            if not isinstance(t, torch.Tensor) or not t.is_cuda:
                raise RuntimeError(f"b200ex region: operand {name} is not a CUDA tensor")
            tensors.append(t)
        dev = tensors[0].device if tensors else torch.device("cuda", torch.cuda.current_device())
        outs = {}
        for o in P.outputs:
            outs[o.name] = torch.empty(o.shape, dtype=o.dtype, device=dev)
        result = dict(outs)
        for name, (src, shape) in self.views.items():
            result[name] = outs[src].view(shape)
        if any(0 in o.shape for o in P.outputs) or 0 in P.ext:
            return result
        key = tuple((t.stride(), t.dtype, t.data_ptr() & 15) for t in tensors)
        plan = self._plans.get(key)
        if plan is None:
            for t, name in zip(tensors, self.in_names):
                if t.dim() > MAX_RANK:
                    raise RuntimeError(f"b200ex region: operand {name} has rank {t.dim()} > {MAX_RANK}")
                if P.in_shapes is not None and tuple(t.shape) != tuple(P.in_shapes[name]):
                    raise RuntimeError(f"b200ex region: operand {name} has shape {tuple(t.shape)}, the region was traced for {P.in_shapes[name]}")
            modes = leaf_modes(P, [tensors[s] for s in self.slot])
            sms = _sm_count()

            def make_plan(kernel):
                if P.kind == "flat":
                    U = flat_unroll(P)
                    items = math.prod(P.ext) // P.W
                    geo = (max(1, min((items + 256 * U - 1) // (256 * U), sms * min(_GRID_CTAS, kernel.ctas_per_sm(256)))), 256, 0, 0)
                else:
                    T, C, RPC, block = row_geometry(P, staged=kernel.variant == 2)
                    ncol = sum(1 for o in P.outputs if o.kind == "col")
                    groups = (P.n_rows + RPC - 1) // RPC
                    grid = max(1, min(groups, sms * kernel.ctas_per_sm(block)))
                    geo = (grid, block, grid * RPC if ncol else 0, ncol * grid * RPC * P.row_len)
                return (kernel, self.Params(), threading.Lock(), geo)

            plan = None
            variants = row_variants(P, modes) if _RING_MODE != "0" else [0]
            if len(variants) > 1 and _RING_MODE not in ("1", "v1"):
                plan = make_plan(_get_kernel(P, modes, dev.index, 0))
                if _RING_MODE == "auto" and not torch.cuda.is_current_stream_capturing():
                    # Builds of the same arithmetic — rows prefetched one ahead in registers, or staged through the
                    # shared-memory ring (at the register geometry, or with more chunks per thread).  Which wins
                    # depends on whether the row's critical path, its register budget or the bytes in flight bound
                    # the kernel (all three happen, measured on B200), so the first call times them on its own operands
                    # and keeps the fastest.  Every build sums in the same order per geometry; results differ at most
                    # by the reassociation of the row sums.
                    ckey = (P.signature, modes, dev.index)
                    best = _choice.get(ckey)  # the same program in another region (every layer of a model) is timed once
                    if best is None:
                        plans = [plan] + [make_plan(_get_kernel(P, modes, dev.index, v)) for v in variants[1:]]
                        times = [self._time_plan(pl, tensors, outs, dev) for pl in plans]
                        self.tuned[key] = dict(zip(variants, times))
                        best = _choice[ckey] = variants[times.index(min(times))]
                        plan = plans[variants.index(best)]
                    elif best != 0:
                        plan = make_plan(_get_kernel(P, modes, dev.index, best))
            if plan is None:
                plan = make_plan(_get_kernel(P, modes, dev.index))
            if len(self._plans) < self.MAX_PLANS:
                self._plans[key] = plan
        self._launch_plan(plan, tensors, outs, dev)
        return result

    def _launch_plan(self, plan, tensors, outs, dev):
        from .ops import _on

        P = self.P
        kernel, prm, lock, (grid, block, nparts, scratch_elems) = plan
        stream = torch.cuda.current_stream(dev).cuda_stream
        scratch = torch.empty(scratch_elems, dtype=torch.float32, device=dev) if scratch_elems else None
        with lock, _on(dev):
            for k, t in enumerate(tensors):
                prm.inp[k].ptr = t.data_ptr()
            for k, o in enumerate(P.outputs):
                prm.out[k].ptr = outs[o.name].data_ptr()
            if scratch is not None:
                prm.nparts = nparts
                prm.scratch = scratch.data_ptr()
            kernel.launch(kernel.main, prm, grid, block, stream)
            if scratch is not None:
                kernel.launch(kernel.fold, prm, (P.row_len + 31) // 32, 256, stream)

    def _time_plan(self, plan, tensors, outs, dev, reps: int = 3) -> float:
        """Milliseconds per launch of `plan` on the caller's operands (one warm-up, then `reps` launches between two
        events on the current stream).  The launches write the region's real outputs; they are idempotent."""
        with torch.cuda.device(dev):
            self._launch_plan(plan, tensors, outs, dev)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                self._launch_plan(plan, tensors, outs, dev)
            b.record()
            b.synchronize()
            return a.elapsed_time(b) / reps
