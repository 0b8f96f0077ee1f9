"""CPU restatement of the reference's fused-region math — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this module, and only as the checker or as the timed CPU baseline.  The product
path (``lightning-thunder_b200/``) never imports it and has no CPU fallback.

What it restates: the per-region programs that the reference's torch executor runs for the
hot-path regions of SURVEY.md §8(a) — one ``torch`` call per ltorch operation, in the order and
with the dtype round trips of the execution trace that ``thunder.jit(fn, executors=["torch"])``
produces (``thunder/executors/torchex.py``; ``convert_element_type`` -> ``Tensor.to(copy=True)``
at torchex.py:80-87).  Every function cites the reference lines that define the math.

Parity pin: ``oracle/gen_golden.py`` runs the UNMODIFIED reference (imported from
``baseline/_ref`` or ``/root/reference`` with the shims in ``baseline/shims``) on seeded inputs
and stores inputs + reference outputs under ``tests/golden/``; ``tests/test_oracle_golden.py``
checks this module against those vectors BIT-EXACTLY on CPU (same ATen kernels, same order).
"""

from __future__ import annotations

import math

import torch

__all__ = [
    "rmsnorm_residual_fwd",
    "rmsnorm_residual_bwd",
    "swiglu_fwd",
    "swiglu_bwd",
    "qkv_split_rope_fwd",
    "qkv_split_rope_bwd",
    "bias_gelu_fwd",
    "bias_gelu_bwd",
    "bias_relu_fwd",
    "bias_relu_bwd",
    "cross_entropy_fwd",
    "cross_entropy_bwd",
    "layernorm_fwd",
    "layernorm_bwd",
    "softmax_fwd",
    "softmax_bwd",
    "adamw_step",
    "philox4x32_10",
    "uniform_philox",
]


def _f32(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.float32, copy=True)


# ----------------------------------------------------------------------------------------------
# RMSNorm (+ residual add) — SURVEY §8 rows a5 / a6
# ----------------------------------------------------------------------------------------------
def rmsnorm_residual_fwd(x, residual, weight, eps: float):
    """``h = x + residual; y = rms_norm(h, (H,), weight, eps)``.

    Follows ltorch.add (thunder/torch/__init__.py, elementwise binary with fp32 upcast for
    bf16/fp16) + ltorch.rms_norm (thunder/torch/__init__.py:4450-4479): upcast, ``mean(h*h)`` as
    ``sum / H``, ``rsqrt(. + eps)``, scale by the upcast weight, downcast.
    Returns (h, y, rstd[rows,1] fp32).  ``residual``/``weight`` may be None.
    """
    dt = x.dtype
    if residual is not None:
        h = torch.add(_f32(x), _f32(residual)).to(dt) if dt != torch.float32 else torch.add(x, residual)
    else:
        h = x
    hf = _f32(h)
    sq = torch.mul(hf, hf)
    ssum = torch.sum(sq, (-1,))
    ms = torch.div(ssum.unsqueeze(-1), float(h.shape[-1]))
    rstd = torch.rsqrt(torch.add(ms, eps))
    n = torch.mul(hf, rstd.expand(hf.shape))
    if weight is not None:
        n = torch.mul(n, _f32(weight.unsqueeze(0).expand(h.shape) if h.dim() == 2 else weight.expand(h.shape)))
    return h, n.to(dt), rstd


def rmsnorm_residual_bwd(dy, dh_in, h, weight, rstd):
    """Backward of :func:`rmsnorm_residual_fwd` as autodiff derives it from the prim vjps
    (thunder/core/transforms.py mul/sum/rsqrt/div/broadcast rules; trace census SURVEY App. C).

    ``dh_in`` is the cotangent of ``h`` arriving from the residual stream (may be None).
    Returns (dx, dw); ``dr`` equals ``dx``.
    """
    dt = h.dtype
    H = float(h.shape[-1])
    hf = _f32(h)
    dyf = _f32(dy)
    rstd_b = rstd.expand(hf.shape)
    if weight is not None:
        wf = _f32(weight.expand(h.shape))
        g = torch.mul(wf, dyf)
        n = torch.mul(hf, rstd_b)
        dw_terms = torch.mul(n, dyf).to(dt)
        red_dims = tuple(range(h.dim() - 1))
        dw = torch.sum(dw_terms, red_dims, False, dtype=None)  # bf16 in -> fp32 accumulate -> bf16 out
    else:
        g = dyf
        dw = None
    t205 = torch.mul(rstd_b, g)
    t206 = torch.mul(hf, g)
    c = torch.sum(t206, (-1,), False).unsqueeze(-1)
    t212 = torch.mul(-0.5, c)
    t213 = torch.pow(rstd, 3.0)
    t214 = torch.mul(t212, t213)
    s = torch.true_divide(t214, H)
    s_b = s.expand(hf.shape)
    t222 = torch.mul(hf, s_b)
    t223 = torch.mul(hf, s_b)
    if dt == torch.float32 and dh_in is not None:
        # No rounding boundary in fp32: autodiff accumulates the incoming cotangent first.
        return torch.add(torch.add(torch.add(dh_in, t205), t222), t223), dw
    dh = torch.add(torch.add(t205, t222), t223).to(dt)
    dx = torch.add(dh_in, dh) if dh_in is not None else dh
    return dx, dw


# ----------------------------------------------------------------------------------------------
# SwiGLU — SURVEY §8 rows a7 / a8
# ----------------------------------------------------------------------------------------------
def swiglu_fwd(a, b):
    """``silu(a) * b``: ltorch.silu (thunder/torch/__init__.py:2266; torchex runs
    ``torch.nn.functional.silu``) then ltorch.mul with fp32 upcast."""
    u = torch.nn.functional.silu(a, False)
    if a.dtype == torch.float32:
        return torch.mul(u, b)
    return torch.mul(_f32(u), _f32(b)).to(a.dtype)


def swiglu_bwd(a, b, g):
    """Backward of :func:`swiglu_fwd`: mul vjp, then the silu vjp of clang.silu
    (thunder/clang/__init__.py:1417, ``a * sigmoid(a)``) evaluated op by op in the storage dtype.
    Returns (da, db)."""
    dt = a.dtype
    u = torch.nn.functional.silu(a, False)
    if dt == torch.float32:
        db = torch.mul(u, g)
        gu = torch.mul(b, g)
    else:
        gf = _f32(g)
        db = torch.mul(_f32(u), gf).to(dt)
        gu = torch.mul(_f32(b), gf).to(dt)
    t152 = torch.neg(a)
    t157 = torch.exp(_f32(t152)).to(dt) if dt != torch.float32 else torch.exp(t152)
    t160 = torch.add(1, t157)
    t163 = torch.true_divide(1, t160)
    t167 = torch.mul(gu, t163)
    t170 = torch.sub(1, t163)
    t174 = torch.mul(a, t170)
    t177 = torch.add(1, t174)
    da = torch.mul(t167, t177)
    return da, db


# ----------------------------------------------------------------------------------------------
# QKV split + RoPE + GQA expand — SURVEY §8 rows a9 / a10
# ----------------------------------------------------------------------------------------------
def _apply_rope(x, cos, sin):
    """litgpt ``apply_rope`` as exercised by thunder/benchmarks/__init__.py:2322-2323:
    ``roped = x * cos + cat(-x2, x1) * sin`` with a storage-dtype round after every ltorch op."""
    dt = x.dtype
    hs = x.shape[-1]
    x1 = x[..., : hs // 2]
    x2 = x[..., hs // 2 :]
    if dt == torch.float32:
        rotated = torch.cat([torch.neg(x2), x1], -1)
        return torch.add(torch.mul(x, cos.expand(x.shape)), torch.mul(rotated, sin.expand(x.shape)))
    rotated = torch.cat([torch.neg(_f32(x2)).to(dt), x1], -1)
    xc = torch.mul(_f32(x), _f32(cos.expand(x.shape))).to(dt)
    rs = torch.mul(_f32(rotated), _f32(sin.expand(x.shape))).to(dt)
    return torch.add(_f32(xc), _f32(rs)).to(dt)


def qkv_split_rope_fwd(qkv, cos, sin, n_head: int, n_query_groups: int, head_size: int, rope_n_elem: int):
    """thunder/benchmarks/__init__.py:2282-2328 (vendored LitGPT attention slice):
    view -> permute -> split(q_per_kv,1,1) -> GQA expand -> reshape -> RoPE on the first
    ``rope_n_elem`` features of q and k -> cat with the un-roped tail.
    qkv [B,T,(n_head+2*groups)*hs]; cos, sin [T, rope_n_elem].  Returns q, k, v [B, n_head, T, hs]."""
    B, T, _ = qkv.shape
    q_per_kv = n_head // n_query_groups
    total = q_per_kv + 2
    t = qkv.view(B, T, n_query_groups, total, head_size).permute(0, 2, 3, 1, 4)
    q, k, v = t.split((q_per_kv, 1, 1), dim=2)
    if n_query_groups != n_head and n_query_groups != 1:
        k = k.expand(B, n_query_groups, q_per_kv, T, head_size)
        v = v.expand(B, n_query_groups, q_per_kv, T, head_size)
    q = q.reshape(B, -1, T, head_size)
    k = k.reshape(B, -1, T, head_size)
    v = v.reshape(B, -1, T, head_size)
    q_roped = _apply_rope(q[..., :rope_n_elem], cos, sin)
    k_roped = _apply_rope(k[..., :rope_n_elem], cos, sin)
    q = torch.cat((q_roped, q[..., rope_n_elem:]), dim=-1)
    k = torch.cat((k_roped, k[..., rope_n_elem:]), dim=-1)
    return q, k, v


def _apply_rope_bwd(g, cos, sin):
    """vjp of :func:`_apply_rope` w.r.t. x with the reference's rounding points: mul vjps rounded to
    the storage dtype, cat vjp = slice, slice vjp = pad, neg vjp in fp32, then two storage-dtype adds."""
    dt = g.dtype
    hs = g.shape[-1]
    half = hs // 2
    if dt == torch.float32:
        g_rot = torch.mul(sin.expand(g.shape), g)
        g_x = torch.mul(cos.expand(g.shape), g)
        lo, hi = g_rot[..., :half], g_rot[..., half:]
        pad_neg = torch.nn.functional.pad(torch.neg(lo), (half, 0))
        pad_hi = torch.nn.functional.pad(hi, (0, half))
        return torch.add(torch.add(g_x, pad_neg), pad_hi)
    gf = _f32(g.to(dt))
    g_rot = torch.mul(_f32(sin.expand(g.shape)), gf).to(dt)
    g_x = torch.mul(_f32(cos.expand(g.shape)), gf).to(dt)
    lo, hi = g_rot[..., :half], g_rot[..., half:]
    pad_neg = torch.nn.functional.pad(torch.neg(_f32(lo)).to(dt), (half, 0))
    pad_hi = torch.nn.functional.pad(hi, (0, half))
    return torch.add(torch.add(g_x, pad_neg), pad_hi)


def qkv_split_rope_bwd(dq, dk, dv, cos, sin, n_head: int, n_query_groups: int, head_size: int, rope_n_elem: int):
    """Backward of :func:`qkv_split_rope_fwd` (autodiff of the slice above): RoPE vjp on dq and dk,
    GQA expand vjp = sum over q_per_kv (fp32 accumulate, storage-dtype result), cat into
    [B, groups, q_per_kv+2, T, hs], permute back and flatten.  Returns dqkv [B,T,(n_head+2*groups)*hs]."""
    B, _, T, _ = dq.shape
    q_per_kv = n_head // n_query_groups

    def rope_part(g):
        head = _apply_rope_bwd(g[..., :rope_n_elem], cos, sin)
        if rope_n_elem == head_size:
            return torch.add(torch.zeros_like(head), head)
        tail = torch.nn.functional.pad(g[..., rope_n_elem:], (rope_n_elem, 0))
        return torch.add(tail, torch.nn.functional.pad(head, (0, head_size - rope_n_elem)))

    dq_x = rope_part(dq).reshape(B, n_query_groups, q_per_kv, T, head_size)
    dk_x = rope_part(dk)
    if n_query_groups != n_head and n_query_groups != 1:
        dk_g = torch.sum(dk_x.reshape(B, n_query_groups, q_per_kv, T, head_size), (2,), False).unsqueeze(2)
        dv_g = torch.sum(dv.reshape(B, n_query_groups, q_per_kv, T, head_size), (2,), False).unsqueeze(2)
    else:
        dk_g = dk_x.reshape(B, n_query_groups, 1, T, head_size)
        dv_g = dv.reshape(B, n_query_groups, 1, T, head_size)
    cat = torch.cat((dq_x, dk_g, dv_g), 2)
    return cat.permute(0, 3, 1, 2, 4).reshape(B, T, -1)


# ----------------------------------------------------------------------------------------------
# bias + activation — SURVEY §8 row a11
# ----------------------------------------------------------------------------------------------
_SQRT2 = 1.4142135623730951
_TWO_OVER_SQRT_PI = 1.1283791670955126


def bias_gelu_fwd(x, bias):
    """``gelu(x + bias)`` in the erf form of ltorch.gelu (thunder/torch/__init__.py:2071-2082):
    ``a * (0.5 + 0.5 * erf(a / sqrt(2)))`` with one storage-dtype round per ltorch op."""
    dt = x.dtype
    if dt == torch.float32:
        a = torch.add(x, bias.expand(x.shape))
        return torch.mul(a, torch.add(0.5, torch.mul(0.5, torch.erf(torch.div(a, _SQRT2)))))
    a = torch.add(_f32(x), _f32(bias.expand(x.shape))).to(dt)
    t191 = torch.div(_f32(a), _SQRT2).to(dt)
    t196 = torch.erf(_f32(t191)).to(dt)
    t201 = torch.mul(0.5, _f32(t196)).to(dt)
    t206 = torch.add(0.5, _f32(t201)).to(dt)
    return torch.mul(_f32(a), _f32(t206)).to(dt)


def bias_gelu_bwd(x, bias, g):
    """Backward of :func:`bias_gelu_fwd`; returns (dx, dbias) with ``dbias = sum_rows dx``."""
    dt = x.dtype
    red_dims = tuple(range(x.dim() - 1))
    if dt == torch.float32:
        a = torch.add(x, bias.expand(x.shape))
        z = torch.div(a, _SQRT2)
        cdf = torch.add(0.5, torch.mul(0.5, torch.erf(z)))
        t216 = torch.mul(cdf, g)
        t217 = torch.mul(a, g)
        t228 = torch.mul(0.5, t217)
        t236 = torch.mul(_TWO_OVER_SQRT_PI, torch.exp(torch.neg(torch.pow(z, 2))))
        t242 = torch.true_divide(torch.mul(t236, t228), _SQRT2)
        dx = torch.add(t216, t242)
        return dx, torch.sum(dx, red_dims, False)
    a = torch.add(_f32(x), _f32(bias.expand(x.shape))).to(dt)
    t191 = torch.div(_f32(a), _SQRT2).to(dt)
    t193 = _f32(t191)
    t196 = torch.erf(t193).to(dt)
    t201 = torch.mul(0.5, _f32(t196)).to(dt)
    t206 = torch.add(0.5, _f32(t201)).to(dt)
    gf = _f32(g)
    t221 = torch.mul(_f32(t206), gf).to(dt)
    t219 = torch.mul(_f32(a), gf).to(dt)
    t230 = torch.mul(0.5, _f32(t219)).to(dt)
    t236 = torch.mul(_TWO_OVER_SQRT_PI, torch.exp(torch.neg(torch.pow(t193, 2))))
    t239 = torch.mul(t236, _f32(t230)).to(dt)
    t244 = torch.true_divide(_f32(t239), _SQRT2).to(dt)
    dx = torch.add(t221, t244)
    return dx, torch.sum(dx, red_dims, False, dtype=None)


def bias_relu_fwd(x, bias):
    """``relu(x + bias)``; ltorch.relu decomposes to ``where(a > 0, a, 0)`` (README MLP trace)."""
    dt = x.dtype
    a = torch.add(x, bias.expand(x.shape)) if dt == torch.float32 else torch.add(_f32(x), _f32(bias.expand(x.shape))).to(dt)
    return torch.where(torch.gt(a, 0), a, torch.zeros((), dtype=dt))


def bias_relu_bwd(x, bias, g):
    dt = x.dtype
    a = torch.add(x, bias.expand(x.shape)) if dt == torch.float32 else torch.add(_f32(x), _f32(bias.expand(x.shape))).to(dt)
    dx = torch.where(torch.gt(a, 0), g, torch.zeros((), dtype=dt))
    return dx, torch.sum(dx, tuple(range(x.dim() - 1)), False, dtype=None)


# ----------------------------------------------------------------------------------------------
# softmax cross-entropy — SURVEY §8 rows a12 / a13
# ----------------------------------------------------------------------------------------------
def cross_entropy_fwd(logits, target, ignore_index: int = -100):
    """``F.cross_entropy(logits, target, ignore_index=..., reduction="mean")`` as ltorch.cross_entropy
    (thunder/torch/__init__.py:5226-5270) decomposes it: log_softmax (:5839-5851, result in the
    logits dtype) then nll_loss (:6018-6031; neg, ne(ignore), where, pad one column,
    take_along_axis, sum, count, true_divide).  Returns (loss, logp, n_valid)."""
    V = logits.shape[1]
    logp = torch.nn.functional.log_softmax(logits, 1)
    neg = torch.neg(logp)
    t = target.unsqueeze(1)
    valid = torch.ne(t, ignore_index)
    idx = torch.where(valid, t, V)
    padded = torch.nn.functional.pad(neg, (0, 1), value=0.0)
    picked = torch.take_along_dim(padded, idx, 1)
    total = torch.sum(picked, None, False, dtype=None)
    n_valid = torch.sum(valid, None, False, dtype=None)
    loss = torch.true_divide(total, n_valid)
    return loss, logp, n_valid


def cross_entropy_bwd(g, logp, target, n_valid, ignore_index: int = -100):
    """``nll_loss_backward`` prim (thunder/torch/__init__.py:6035-6047; ATen impl
    thunder/executors/torchex.py:1731-1783) followed by ``log_softmax_backward``
    (:5859-5864, ATen ``_log_softmax_backward_data``).  Returns dlogits in the logits dtype."""
    dlogp = torch.ops.aten.nll_loss_backward(
        g, logp, target, None, 1, ignore_index, n_valid.to(logp.dtype) if n_valid.dtype != logp.dtype else n_valid
    )
    return torch.ops.aten._log_softmax_backward_data(dlogp, logp, 1, logp.dtype)


# ----------------------------------------------------------------------------------------------
# LayerNorm — SURVEY §8 row f.1 (the normalisation of the GPT-2 / BERT style blocks)
# ----------------------------------------------------------------------------------------------
def layernorm_fwd(x, weight, bias, eps: float):
    """``F.layer_norm(x, (H,), weight, bias, eps)`` over the last dimension.

    Follows ltorch.layer_norm -> _native_layer_norm -> _normalize (thunder/torch/__init__.py:4406-4447):
    upcast, ``var_mean(correction=0)``, ``rsqrt(var + eps)``, ``(x - mean) * rstd``, times the upcast weight,
    plus the upcast bias, downcast.  Returns (y, mean[rows,1] fp32, rstd[rows,1] fp32); weight / bias may be None.
    """
    dt = x.dtype
    xf = _f32(x)
    var, mean = torch.var_mean(xf, (-1,), correction=0)
    var, mean = var.unsqueeze(-1), mean.unsqueeze(-1)
    rstd = torch.rsqrt(torch.add(var, eps))
    xc = torch.sub(xf, mean.expand(xf.shape))
    out = torch.mul(xc, rstd.expand(xf.shape))
    if weight is not None:
        out = torch.mul(out, _f32(weight.expand(x.shape)))
    if bias is not None:
        out = torch.add(out, _f32(bias.expand(x.shape)))
    return out.to(dt), mean, rstd


def layernorm_bwd(dy, x, weight, has_bias: bool, mean, rstd):
    """Backward of :func:`layernorm_fwd` as autodiff derives it: the mul/add/sub/rsqrt vjps plus the
    var_mean rule (thunder/core/transforms.py:1292-1312: ``mean_grad = (1/H) * gm`` and
    ``var_grad = (2 * gv * (x - mean)) / H``, summed first, then added to the direct term).

    16-bit storage: the direct term ``rstd * (w * dy)`` and the statistics term are each rounded to the
    storage dtype where the trace holds them as tensors, then added in that dtype.  dw sums the ROUNDED
    products ``round(xhat * dy)``, db sums dy (both with fp32 accumulation inside ``torch.sum``).
    Returns (dx, dw, db); dw / db are None when the parameter is absent.
    """
    dt = x.dtype
    H = x.shape[-1]
    lead = tuple(range(x.dim() - 1))
    xf = _f32(x)
    g = _f32(dy)
    mean_b, rstd_b = mean.expand(xf.shape), rstd.expand(xf.shape)
    xc = torch.sub(xf, mean_b)
    xhat = torch.mul(xc, rstd_b)
    db = torch.sum(g.to(dt), lead, False, dtype=None) if has_bias else None
    if weight is not None:
        gw = torch.mul(_f32(weight.expand(x.shape)), g)
        dw = torch.sum(torch.mul(xhat, g).to(dt), lead, False, dtype=None)
    else:
        gw, dw = g, None
    direct = torch.mul(rstd_b, gw)
    c1 = torch.sum(torch.mul(xc, gw), (-1,), False).unsqueeze(-1)
    c2 = torch.sum(torch.neg(direct), (-1,), False).unsqueeze(-1)
    gvar = torch.mul(torch.mul(-0.5, c1), torch.pow(rstd, 3.0))
    mean_grad = torch.mul(1.0 / H, c2.expand(xf.shape))
    var_grad = torch.true_divide(torch.mul(torch.mul(2, gvar.expand(xf.shape)), torch.sub(xf, mean_b)), H)
    stats = torch.add(mean_grad, var_grad)
    if dt == torch.float32:
        return torch.add(direct, stats), dw, db
    return torch.add(direct.to(dt), stats.to(dt)), dw, db


# ----------------------------------------------------------------------------------------------
# Counter-based uniform random numbers — SURVEY §8 row f.1 (dropout)
# ----------------------------------------------------------------------------------------------
_PHILOX_M0, _PHILOX_M1 = 0xD2511F53, 0xCD9E8D57
_PHILOX_W0, _PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(seed: int, subsequence, counter: int):
    """Philox-4x32-10 (Salmon et al., SC'11; the generator behind curand's Philox4_32_10 and ATen's CUDA RNG):
    key = the 64-bit seed, 128-bit counter = (counter, subsequence).  ``subsequence`` is a numpy uint64 array; returns
    a uint32 array [..., 4].  Ten rounds of ``(hi(M1*c2) ^ c1 ^ k0, lo(M1*c2), hi(M0*c0) ^ c3 ^ k1, lo(M0*c0))`` with the
    key bumped by the Weyl constants between rounds."""
    import numpy as np

    sub = np.asarray(subsequence, dtype=np.uint64)
    mask = np.uint64(0xFFFFFFFF)
    c0 = np.full(sub.shape, counter & 0xFFFFFFFF, dtype=np.uint64)
    c1 = np.full(sub.shape, (counter >> 32) & 0xFFFFFFFF, dtype=np.uint64)
    c2, c3 = sub & mask, (sub >> np.uint64(32)) & mask
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = np.uint64(_PHILOX_M0) * c0, np.uint64(_PHILOX_M1) * c2
        c0, c1, c2, c3 = ((p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0)) & mask, p1 & mask, ((p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1)) & mask, p0 & mask
        k0, k1 = (k0 + _PHILOX_W0) & 0xFFFFFFFF, (k1 + _PHILOX_W1) & 0xFFFFFFFF
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def uniform_philox(shape, minval: float, maxval: float, seed: int, offset: int, dtype):
    """``prims.uniform_philox`` (thunder/core/prims.py:3007-3045) in the fusion executor's convention
    (thunder/executors/nvfuserex_impl.py:1147-1181 and torchex.py:478-493, "the nvFuser way"): ``offset`` is the
    generator's Philox offset divided by 4 and is used as the counter as it is; element ``i`` (row-major) takes
    component ``i % 4`` of the block with subsequence ``i // 4``, so one call consumes a single counter value and the
    state prim can advance the generator by 4 per call.  u = x * 2^-32 + 2^-33 in fp32 (curand_uniform), mapped from
    (0, 1] to [0, 1), scaled to [minval, maxval) and rounded to ``dtype`` (a value that rounds up to maxval becomes minval)."""
    import numpy as np

    n = 1
    for s in shape:
        n *= int(s)
    idx = np.arange(n, dtype=np.uint64)
    blocks = philox4x32_10(int(seed), idx >> np.uint64(2), int(offset))
    x = blocks[np.arange(n), (idx & np.uint64(3)).astype(np.int64)] if n else np.zeros(0, dtype=np.uint32)
    u = x.astype(np.float32) * np.float32(2.0**-32) + np.float32(2.0**-33)
    u = np.where(u == np.float32(1.0), np.float32(0.0), u).astype(np.float32)
    lo, hi = np.float32(minval), np.float32(maxval)
    v = (lo + (hi - lo) * u.astype(np.float64)).astype(np.float32) if (minval, maxval) != (0.0, 1.0) else u
    t = torch.from_numpy(v).to(dtype)
    t = torch.where(t == torch.tensor(maxval, dtype=dtype), torch.tensor(minval, dtype=dtype), t)
    return t.reshape(tuple(shape))


# ----------------------------------------------------------------------------------------------
# softmax over the last dimension — SURVEY §8 row f.1
# ----------------------------------------------------------------------------------------------
def softmax_fwd(x):
    """ltorch.softmax (thunder/torch/__init__.py `_softmax`): the torch executor runs it as one ``torch.softmax`` call."""
    return torch.softmax(x, -1)


def softmax_bwd(g, y):
    """Backward as autodiff derives it from the decomposition and the torch executor runs it: four torch calls on
    storage-dtype tensors (``mul``, ``sum(keepdim)``, ``sub``, ``mul``), each rounding its result."""
    t = torch.mul(y, g)
    s = torch.sum(t, -1, True, dtype=None)
    return torch.mul(y, torch.sub(g, s))


# ----------------------------------------------------------------------------------------------
# AdamW step — SURVEY §8 row f.4
# ----------------------------------------------------------------------------------------------
def adamw_step(p, g, m, v, step: int, lr: float, beta1: float, beta2: float, eps: float, weight_decay: float):
    """One ``torch.optim.AdamW(fused=True)`` update (the optimizer of thunder/benchmarks/benchmark_litgpt.py:278-282),
    i.e. ATen's ``_fused_adamw_``: every hyper-parameter is narrowed to fp32, the whole update is fp32, parameters and
    moments are rounded to their storage dtype on store.  Returns (p, m, v) as new tensors."""
    import numpy as np

    dt = p.dtype
    f = np.float32
    arr = lambda t: t.detach().float().cpu().numpy().astype(np.float32)
    P, G, M, V = arr(p), arr(g), arr(m), arr(v)
    if weight_decay != 0:
        P = P - (f(lr) * f(weight_decay)) * P
    M = f(beta1) * M + (f(1) - f(beta1)) * G
    V = f(beta2) * V + ((f(1) - f(beta2)) * G) * G
    step_size = f(lr) / f(1 - beta1**step)
    denom = np.sqrt(V) / f(math.sqrt(1 - beta2**step)) + f(eps)
    P = P - (step_size * M) / denom
    back = lambda a: torch.from_numpy(a.astype(np.float32)).to(dt)
    return back(P), back(M), back(V)
