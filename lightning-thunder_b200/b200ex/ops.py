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

__all__ = ["rmsnorm_fwd", "rmsnorm_bwd", "DTYPE_CODE"]

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


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def _rows2d(t: torch.Tensor) -> torch.Tensor:
    """View as [rows, hidden]; makes a contiguous copy only if the layout is not already row-major."""
    t = t if t.is_contiguous() else t.contiguous()
    return t.view(-1, t.shape[-1])


def rmsnorm_fwd(x, residual, weight, eps: float):
    """(h, y, rstd) = RMSNorm(x + residual) — see b200k_rmsnorm_fwd.  ``h is x`` when residual is None.
    rstd has shape x.shape[:-1] + (1,) in fp32."""
    dev = _require_cuda(x, residual, weight)
    shape = x.shape
    x2 = _rows2d(x)
    r2 = _rows2d(residual) if residual is not None else None
    w = weight.contiguous() if weight is not None else None
    rows, hidden = x2.shape
    y = torch.empty_like(x2)
    h = torch.empty_like(x2) if residual is not None else None
    rstd = torch.empty(rows, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
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
    w = weight.contiguous() if weight is not None else None
    rs = rstd.contiguous().view(-1)
    rows, hidden = h2.shape
    dx = torch.empty_like(h2)
    dw = torch.empty_like(w) if w is not None else None
    L = _lib.lib()
    with torch.cuda.device(dev):
        nbytes = L.b200k_rmsnorm_bwd_scratch_bytes(rows, hidden) if w is not None else 0
        scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev) if w is not None else None
        rc = L.b200k_rmsnorm_bwd(
            _ptr(dy2), _ptr(h2), _ptr(w), _ptr(rs), _ptr(d2), _ptr(dx), _ptr(dw), _ptr(scratch), nbytes,
            rows, hidden, _code(h), _stream(dev),
        )
    _lib.check(rc, "b200k_rmsnorm_bwd")
    return dx.view(shape), dw
