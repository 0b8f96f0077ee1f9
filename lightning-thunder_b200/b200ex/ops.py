"""Tensor-level launchers: torch CUDA tensors in, torch CUDA tensors out, one C-ABI call each.

These are the callables the b200 executor puts into a trace in place of a fused region
(the role ``FusionDefinitionWrapper.__call__`` plays for nvFuser,
thunder/executors/nvfuserex_impl.py:526-576).  Outputs and scratch come from torch's caching
allocator (graph- and stream-safe), launches go to the current stream of the inputs' device, and
nothing here synchronises.  Inputs must be CUDA tensors; anything else raises.
"""

from __future__ import annotations

import torch

from . import _lib

__all__ = [
    "rmsnorm_fwd", "rmsnorm_bwd", "swiglu_fwd", "swiglu_bwd", "cross_entropy_fwd", "cross_entropy_bwd",
    "qkv_rope_fwd", "qkv_rope_bwd", "bias_act_fwd", "bias_act_bwd", "colsum", "layernorm_fwd", "layernorm_bwd", "uniform_philox", "softmax_fwd", "softmax_bwd",
    "ACT_CODE", "DTYPE_CODE",
]

DTYPE_CODE = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


def _code(t: torch.Tensor) -> int:
    try:
        return DTYPE_CODE[t.dtype]
    except KeyError:
        raise TypeError(f"b200ex: unsupported dtype {t.dtype}") from None


def _require_cuda(*ts) -> torch.device:
    dev = None
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("b200ex launchers take CUDA tensors only (there is no CPU path)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"b200ex: tensors on different devices ({dev} vs {t.device})")
    return dev


class _NoCtx:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NOCTX = _NoCtx()


def _on(dev: torch.device):
    """Context that makes `dev` current for the launch; free when it already is (the common case: switching
    devices through torch costs more host time than the launch itself)."""
    return _NOCTX if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


# The family kernels take dense row-major operands.  A strided operand is copied first — a full extra read + write
# that no roofline line of the kernel counts — so every such copy is counted here and reported by the benchmarks
# (train_step's `operand_copies`); a correct model-level layout keeps both numbers at zero.  The generated region
# kernels (b200ex/regionc.py) read strided operands in place instead.
COPY_STATS = {"copies": 0, "bytes": 0}


def _contig(t):
    if t.is_contiguous():
        return t
    COPY_STATS["copies"] += 1
    COPY_STATS["bytes"] += 2 * t.numel() * t.element_size()
    return t.contiguous()


def _rows2d(t: torch.Tensor) -> torch.Tensor:
    """View as [rows, hidden]; makes a contiguous copy only if the layout is not already row-major."""
    t = _contig(t)
    return t.view(-1, t.shape[-1]) if t.dim() > 0 else t.view(1, 1)


def rmsnorm_fwd(x, residual, weight, eps: float):
    """(h, y, rstd) = RMSNorm(x + residual) — see b200k_rmsnorm_fwd.  ``h is x`` when residual is None.
    rstd has shape x.shape[:-1] + (1,) in fp32."""
    dev = _require_cuda(x, residual, weight)
    shape = x.shape
    x2 = _rows2d(x)
    r2 = _rows2d(residual) if residual is not None else None
    w = _contig(weight) if weight is not None else None
    rows, hidden = x2.shape
    y = torch.empty_like(x2)
    h = torch.empty_like(x2) if residual is not None else None
    rstd = torch.empty(rows, dtype=torch.float32, device=dev)
    with _on(dev):
        rc = _lib.lib().b200k_rmsnorm_fwd(
            _ptr(x2), _ptr(r2), _ptr(w), _ptr(h), _ptr(y), _ptr(rstd), rows, hidden, float(eps), _code(x), _stream(dev)
        )
    _lib.check(rc, "b200k_rmsnorm_fwd")
    h_out = h.view(shape) if h is not None else x
    return h_out, y.view(shape), rstd.view(*shape[:-1], 1)


def rmsnorm_bwd(dy, h, weight, rstd, dh_in=None):
    """(dx, dw) — see b200k_rmsnorm_bwd.  dw is None when weight is None."""
    dev = _require_cuda(dy, h, weight, rstd, dh_in)
    shape = h.shape
    dy2, h2 = _rows2d(dy), _rows2d(h)
    d2 = _rows2d(dh_in) if dh_in is not None else None
    w = _contig(weight) if weight is not None else None
    rs = _contig(rstd).view(-1)
    rows, hidden = h2.shape
    dx = torch.empty_like(h2)
    dw = torch.empty_like(w) if w is not None else None
    L = _lib.lib()
    with _on(dev):
        nbytes = L.b200k_rmsnorm_bwd_scratch_bytes(rows, hidden) if w is not None else 0
        scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev) if w is not None else None
        rc = L.b200k_rmsnorm_bwd(
            _ptr(dy2), _ptr(h2), _ptr(w), _ptr(rs), _ptr(d2), _ptr(dx), _ptr(dw), _ptr(scratch), nbytes,
            rows, hidden, _code(h), _stream(dev),
        )
    _lib.check(rc, "b200k_rmsnorm_bwd")
    return dx.view(shape), dw


def layernorm_fwd(x, weight, bias, eps: float):
    """(y, mean, rstd) = LayerNorm over the last dimension — see b200k_layernorm_fwd.
    mean and rstd have shape x.shape[:-1] + (1,) in fp32."""
    dev = _require_cuda(x, weight, bias)
    xc = _contig(x)
    w = _contig(weight) if weight is not None else None
    b = _contig(bias) if bias is not None else None
    hidden = xc.shape[-1]
    rows = xc.numel() // hidden if hidden else 0
    y = torch.empty_like(xc)
    stat_shape = tuple(xc.shape[:-1]) + (1,)
    mean = torch.empty(stat_shape, dtype=torch.float32, device=dev)
    rstd = torch.empty(stat_shape, dtype=torch.float32, device=dev)
    with _on(dev):
        rc = _lib.lib().b200k_layernorm_fwd(
            xc.data_ptr(), _ptr(w), _ptr(b), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), rows, hidden, float(eps), _code(x), _stream(dev)
        )
    _lib.check(rc, "b200k_layernorm_fwd")
    return y, mean, rstd


def layernorm_bwd(dy, x, weight, mean, rstd, want_dbias: bool):
    """(dx, dw, db) — see b200k_layernorm_bwd.  dw is None when weight is None, db is None unless want_dbias."""
    dev = _require_cuda(dy, x, weight, mean, rstd)
    dyc, xc, mu, rs = _contig(dy), _contig(x), _contig(mean), _contig(rstd)
    w = _contig(weight) if weight is not None else None
    hidden = xc.shape[-1]
    rows = xc.numel() // hidden if hidden else 0
    dx = torch.empty_like(xc)
    dw = torch.empty_like(w) if w is not None else None
    db = torch.empty(hidden, dtype=x.dtype, device=dev) if want_dbias else None
    L = _lib.lib()
    with _on(dev):
        need = w is not None or want_dbias
        nbytes = L.b200k_layernorm_bwd_scratch_bytes(rows, hidden) if need else 0
        scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev) if need else None
        rc = L.b200k_layernorm_bwd(
            dyc.data_ptr(), xc.data_ptr(), _ptr(w), mu.data_ptr(), rs.data_ptr(), dx.data_ptr(), _ptr(dw), _ptr(db), _ptr(scratch), nbytes,
            rows, hidden, _code(x), _stream(dev),
        )
    _lib.check(rc, "b200k_layernorm_bwd")
    return dx, dw, db


def swiglu_fwd(a, b):
    """silu(a) * b — see b200k_swiglu_fwd."""
    dev = _require_cuda(a, b)
    a, b = _contig(a), _contig(b)
    out = torch.empty_like(a)
    with _on(dev):
        rc = _lib.lib().b200k_swiglu_fwd(_ptr(a), _ptr(b), _ptr(out), a.numel(), _code(a), _stream(dev))
    _lib.check(rc, "b200k_swiglu_fwd")
    return out


def swiglu_bwd(a, b, g):
    """(da, db) for out = silu(a) * b — see b200k_swiglu_bwd."""
    dev = _require_cuda(a, b, g)
    a, b, g = _contig(a), _contig(b), _contig(g)
    da, db = torch.empty_like(a), torch.empty_like(a)
    with _on(dev):
        rc = _lib.lib().b200k_swiglu_bwd(_ptr(a), _ptr(b), _ptr(g), _ptr(da), _ptr(db), a.numel(), _code(a), _stream(dev))
    _lib.check(rc, "b200k_swiglu_bwd")
    return da, db


REDUCTION_CODE = {"mean": 1, "sum": 2}


def cross_entropy_fwd(logits, target, ignore_index: int = -100, reduction: str = "mean"):
    """(loss, row_max, row_logsum, n_valid) — see b200k_cross_entropy_fwd.  logits [rows, vocab], target int64 [rows]."""
    dev = _require_cuda(logits, target)
    if logits.dim() != 2 or target.dim() != 1 or target.dtype != torch.int64 or target.shape[0] != logits.shape[0]:
        raise ValueError("b200ex.cross_entropy_fwd: expected logits [rows, vocab] and int64 target [rows]")
    logits, target = _contig(logits), _contig(target)
    rows, vocab = logits.shape
    f32 = dict(dtype=torch.float32, device=dev)
    row_max, row_logsum, row_nll = torch.empty(rows, **f32), torch.empty(rows, **f32), torch.empty(rows, **f32)
    loss = torch.empty((), dtype=logits.dtype, device=dev)
    n_valid = torch.empty((), dtype=torch.int64, device=dev)
    with _on(dev):
        rc = _lib.lib().b200k_cross_entropy_fwd(
            _ptr(logits), _ptr(target), _ptr(row_max), _ptr(row_logsum), _ptr(row_nll), _ptr(loss), _ptr(n_valid),
            rows, vocab, int(ignore_index), REDUCTION_CODE[reduction], _code(logits), _stream(dev),
        )
    _lib.check(rc, "b200k_cross_entropy_fwd")
    return loss, row_max, row_logsum, n_valid


def cross_entropy_bwd(g, logits, target, row_max, row_logsum, n_valid, ignore_index: int = -100, reduction: str = "mean"):
    """dlogits — see b200k_cross_entropy_bwd.  g is the 0-dim cotangent of the loss."""
    dev = _require_cuda(g, logits, target, row_max, row_logsum, n_valid)
    logits, target = _contig(logits), _contig(target)
    rows, vocab = logits.shape
    g = g.to(logits.dtype) if g.dtype != logits.dtype else g
    dlogits = torch.empty_like(logits)
    with _on(dev):
        rc = _lib.lib().b200k_cross_entropy_bwd(
            _ptr(g), _ptr(logits), _ptr(target), _ptr(row_max), _ptr(row_logsum), _ptr(n_valid), _ptr(dlogits),
            rows, vocab, int(ignore_index), REDUCTION_CODE[reduction], _code(logits), _stream(dev),
        )
    _lib.check(rc, "b200k_cross_entropy_bwd")
    return dlogits


def _expand_kv(n_head: int, n_query_groups: int) -> bool:
    return n_query_groups != n_head and n_query_groups != 1


def qkv_rope_fwd(qkv, cos, sin, n_head: int, n_query_groups: int, head_size: int, rope_n_elem: int):
    """(q, k, v) from the fused qkv projection — see b200k_qkv_rope_fwd."""
    dev = _require_cuda(qkv, cos, sin)
    qkv, cos, sin = _contig(qkv), _contig(cos), _contig(sin)
    B, T, _ = qkv.shape
    qpk = n_head // n_query_groups
    expand = _expand_kv(n_head, n_query_groups)
    hkv = n_head if expand else n_query_groups
    q = torch.empty(B, n_head, T, head_size, dtype=qkv.dtype, device=dev)
    k = torch.empty(B, hkv, T, head_size, dtype=qkv.dtype, device=dev)
    v = torch.empty(B, hkv, T, head_size, dtype=qkv.dtype, device=dev)
    with _on(dev):
        rc = _lib.lib().b200k_qkv_rope_fwd(
            _ptr(qkv), _ptr(cos), _ptr(sin), _ptr(q), _ptr(k), _ptr(v), B, T, n_query_groups, qpk, head_size, rope_n_elem,
            int(expand), _code(qkv), _stream(dev),
        )
    _lib.check(rc, "b200k_qkv_rope_fwd")
    return q, k, v


def qkv_rope_bwd(dq, dk, dv, cos, sin, n_head: int, n_query_groups: int, head_size: int, rope_n_elem: int):
    """dqkv [B, T, (n_head + 2 * n_query_groups) * head_size] — see b200k_qkv_rope_bwd."""
    dev = _require_cuda(dq, dk, dv, cos, sin)
    dq, dk, dv, cos, sin = _contig(dq), _contig(dk), _contig(dv), _contig(cos), _contig(sin)
    B, _, T, _ = dq.shape
    qpk = n_head // n_query_groups
    expand = _expand_kv(n_head, n_query_groups)
    dqkv = torch.empty(B, T, (n_head + 2 * n_query_groups) * head_size, dtype=dq.dtype, device=dev)
    with _on(dev):
        rc = _lib.lib().b200k_qkv_rope_bwd(
            _ptr(dq), _ptr(dk), _ptr(dv), _ptr(cos), _ptr(sin), _ptr(dqkv), B, T, n_query_groups, qpk, head_size, rope_n_elem,
            int(expand), _code(dq), _stream(dev),
        )
    _lib.check(rc, "b200k_qkv_rope_bwd")
    return dqkv


ACT_CODE = {"none": 0, "relu": 1, "gelu": 2}


def bias_act_fwd(x, bias, act: str, return_pre: bool = False):
    """act(x + bias) — see b200k_bias_act_fwd.  With return_pre also returns the rounded pre-activation x + bias."""
    dev = _require_cuda(x, bias)
    shape = x.shape
    x2 = _rows2d(x)
    b = _contig(bias) if bias is not None else None
    out = torch.empty_like(x2)
    pre = torch.empty_like(x2) if return_pre else None
    rows, hidden = x2.shape
    with _on(dev):
        rc = _lib.lib().b200k_bias_act_fwd(_ptr(x2), _ptr(b), _ptr(out), _ptr(pre), rows, hidden, ACT_CODE[act], _code(x), _stream(dev))
    _lib.check(rc, "b200k_bias_act_fwd")
    return (out.view(shape), pre.view(shape)) if return_pre else out.view(shape)


def bias_act_bwd(x, bias, g, act: str, want_dbias: bool | None = None):
    """(dx, dbias) — see b200k_bias_act_bwd.  ``x`` is the input when ``bias`` is given, the saved pre-activation when
    ``bias`` is None.  dbias (column sum of dx) is returned when want_dbias, which defaults to ``bias is not None``."""
    dev = _require_cuda(x, bias, g)
    shape = x.shape
    x2, g2 = _rows2d(x), _rows2d(g)
    b = _contig(bias) if bias is not None else None
    rows, hidden = x2.shape
    want = (bias is not None) if want_dbias is None else want_dbias
    dx = torch.empty_like(x2)
    db = torch.empty(hidden, dtype=x.dtype, device=dev) if want else None
    L = _lib.lib()
    with _on(dev):
        nbytes = L.b200k_bias_act_bwd_scratch_bytes(rows, hidden) if want else 0
        scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev) if want else None
        rc = L.b200k_bias_act_bwd(
            _ptr(x2), _ptr(b), _ptr(g2), _ptr(dx), _ptr(db), _ptr(scratch), nbytes, rows, hidden, ACT_CODE[act], _code(x), _stream(dev)
        )
    _lib.check(rc, "b200k_bias_act_bwd")
    return dx.view(shape), db


def colsum(x):
    """Sum over all leading dims: x [..., hidden] -> [hidden] in x's dtype (fp32 accumulation) — see b200k_colsum."""
    dev = _require_cuda(x)
    x2 = _rows2d(x)
    rows, hidden = x2.shape
    out = torch.empty(hidden, dtype=x.dtype, device=dev)
    L = _lib.lib()
    with _on(dev):
        nbytes = L.b200k_bias_act_bwd_scratch_bytes(rows, hidden)
        scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
        rc = L.b200k_colsum(_ptr(x2), _ptr(out), _ptr(scratch), nbytes, rows, hidden, _code(x), _stream(dev))
    _lib.check(rc, "b200k_colsum")
    return out


def uniform_philox(shape, minval: float, maxval: float, *, device, dtype: torch.dtype, seed: int, offset: int):
    """Uniform random tensor in [minval, maxval) from (seed, offset) — see b200k_uniform_philox."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("b200ex launchers take CUDA tensors only (there is no CPU path)")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    out = torch.empty(tuple(shape), dtype=dtype, device=dev)
    with _on(dev):
        rc = _lib.lib().b200k_uniform_philox(
            out.data_ptr(), out.numel(), int(seed) & 0xFFFFFFFFFFFFFFFF, int(offset) & 0xFFFFFFFFFFFFFFFF, float(minval), float(maxval),
            _code(out), _stream(dev),
        )
    _lib.check(rc, "b200k_uniform_philox")
    return out


def softmax_fwd(x):
    """softmax over the last dimension — see b200k_softmax_fwd."""
    dev = _require_cuda(x)
    xc = _contig(x)
    cols = xc.shape[-1]
    y = torch.empty_like(xc)
    with _on(dev):
        rc = _lib.lib().b200k_softmax_fwd(xc.data_ptr(), y.data_ptr(), xc.numel() // cols if cols else 0, cols, _code(x), _stream(dev))
    _lib.check(rc, "b200k_softmax_fwd")
    return y


def softmax_bwd(dy, y):
    """dx of softmax over the last dimension from the saved output — see b200k_softmax_bwd."""
    dev = _require_cuda(dy, y)
    gc, yc = _contig(dy), _contig(y)
    cols = yc.shape[-1]
    dx = torch.empty_like(yc)
    with _on(dev):
        rc = _lib.lib().b200k_softmax_bwd(gc.data_ptr(), yc.data_ptr(), dx.data_ptr(), yc.numel() // cols if cols else 0, cols, _code(y), _stream(dev))
    _lib.check(rc, "b200k_softmax_bwd")
    return dx
