"""Test infrastructure: a CPU evaluator for b200ex.regionc programs.

It walks the same Program the CUDA code generator consumes — iteration variables, gather leaves with their affine
coordinates and guards, conditions, the SSA list, row reductions, column sums — with torch CPU tensors over the whole
iteration space at once.  CPU tests use it to check the compiler front end (index maps, variable splitting, guard
logic, phases) against the reference's torch executor without a GPU; the generated kernels are then checked against
the same reference on the GPU.  Nothing in the product imports this file."""

import math

import torch


def _aff_eval(a, grids):
    r = torch.full_like(grids[0], a[1]) if grids else torch.tensor(a[1])
    for v, k in a[0]:
        r = r + k * grids[v]
    return r


def _rnd(x, dt):
    return x.to(dt).to(torch.float32)


_UN = {
    "neg": torch.neg, "abs": torch.abs, "exp": torch.exp, "log": torch.log, "sqrt": torch.sqrt, "rsqrt": torch.rsqrt,
    "reciprocal": torch.reciprocal, "erf": torch.erf, "tanh": torch.tanh, "sign": torch.sign, "floor": torch.floor, "ceil": torch.ceil,
    "exp2": torch.exp2, "log2": torch.log2, "log1p": torch.log1p, "expm1": torch.expm1, "sin": torch.sin, "cos": torch.cos,
    "erfc": torch.erfc, "sigmoid": torch.sigmoid, "trunc": torch.trunc, "round": torch.round, "mov": lambda x: x,
    "not": lambda x: (x == 0).float(), "rnd_bf16": lambda x: _rnd(x, torch.bfloat16), "rnd_f16": lambda x: _rnd(x, torch.float16),
}
_BIN = {
    "add": torch.add, "sub": torch.sub, "mul": torch.mul, "div": torch.div, "pow": torch.pow, "maximum": torch.maximum, "minimum": torch.minimum,
    "gt": lambda a, b: (a > b).float(), "ge": lambda a, b: (a >= b).float(), "lt": lambda a, b: (a < b).float(), "le": lambda a, b: (a <= b).float(),
    "eq": lambda a, b: (a == b).float(), "ne": lambda a, b: (a != b).float(),
    "and": lambda a, b: ((a != 0) & (b != 0)).float(), "or": lambda a, b: ((a != 0) | (b != 0)).float(),
}


def run_program(P, env: dict) -> dict:
    """env: input name -> torch CPU tensor.  Returns output name -> tensor of the declared dtype and shape."""
    ext = list(P.ext)
    shape = tuple(ext) if ext else ()
    grids = list(torch.meshgrid(*[torch.arange(n) for n in ext], indexing="ij")) if ext else []
    one = torch.zeros(shape, dtype=torch.float32)
    conds = []
    for terms, lo, hi in P.conds:
        lin = _aff_eval((terms, 0), grids)
        conds.append((lin >= lo) & (lin < hi))
    leaves = []
    for lf in P.leaves:
        x = env[lf.name]
        # the runtime dtype wins over the declared one (see regionc._LOAD_DT)
        mask = torch.ones(shape, dtype=torch.bool)
        for ci in lf.guard:
            mask = mask & conds[ci]
        idx = []
        for d, c in enumerate(lf.coords):
            i = _aff_eval(c, grids) if grids else torch.tensor(c[1])
            ok = (i >= 0) & (i < x.shape[d])
            assert bool((ok | ~mask).all()), f"leaf {lf.name}: guarded access leaves dim {d}"
            idx.append(i.clamp(0, max(x.shape[d] - 1, 0)))
        val = x[tuple(idx)].to(torch.float32) if x.dim() else x.to(torch.float32).expand(shape)
        leaves.append(torch.where(mask, val, torch.zeros_like(val)))
    inner_axes = tuple(range(P.n_outer, len(ext)))
    vals = []
    for op, a, b, c, imm in P.ssa:
        if op == "leaf":
            r = leaves[imm]
        elif op == "const":
            r = one + torch.tensor([imm], dtype=torch.int64).to(torch.int32).view(torch.float32)
        elif op == "pred":
            r = conds[imm].float()
        elif op == "where":
            r = torch.where(vals[a] != 0, vals[b], vals[c])
        elif op.startswith("reduce_"):
            x = vals[a]
            kind = op[7:]
            if kind == "sum":
                r = x.sum(dim=inner_axes, keepdim=True)
            elif kind == "prod":
                r = x
                for ax in inner_axes:
                    r = r.prod(dim=ax, keepdim=True)
            elif kind == "amax":
                r = x.amax(dim=inner_axes, keepdim=True)
            else:
                r = x.amin(dim=inner_axes, keepdim=True)
            r = r.expand(shape)
        elif op in _BIN:
            r = _BIN[op](vals[a], vals[b])
        else:
            r = _UN[op](vals[a])
        vals.append(r.to(torch.float32))
    out = {}
    for o in P.outputs:
        v = vals[o.sid].expand(shape) if shape else vals[o.sid]
        if o.kind == "full":
            r = v.reshape(o.shape)
        elif o.kind == "row":
            r = v[(Ellipsis,) + (0,) * len(inner_axes)].reshape(o.shape)
        else:
            r = v.sum(dim=tuple(range(P.n_outer))).reshape(o.shape)
        out[o.name] = (r != 0) if o.dtype == torch.bool else r.to(o.dtype)
    return out


def run_launches(launches, env: dict) -> dict:
    env = dict(env)
    for ln in launches:
        res = run_program(ln.program, env)
        for name, (src, shape) in ln.views.items():
            res[name] = res[src].view(shape)
        env.update(res)
    return env
