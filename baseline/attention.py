"""Benchmark-side attention route: ``scaled_dot_product_attention`` -> cuDNN's fused attention through ATen.

Attention is outside this repository's path (BASELINE.json: GEMMs and attention stay on library kernels).  The
reference's ``sdpa`` executor only knows ATen's flash / memory-efficient kernels (thunder/executors/sdpaex.py:431-468),
which are sm_80-era code: on B200 they take 3.5 ms per Llama-3-8B layer (fwd + bwd) where cuDNN's kernel takes 1.0 ms
(tools/bench_sdpa.py).  This operator executor asks ATen for the cuDNN kernel instead; it is used by BOTH arms of the
training-step benchmark (b200 and the reference's torch executor), so the comparison between them is unchanged.
Library call only — no kernel of this repository is involved.
"""

from __future__ import annotations

import torch

import thunder.core.dtypes as dtypes
import thunder.torch as ltorch
from thunder.core.proxies import TensorProxy
from thunder.core.transforms import get_grad, put_grads
from thunder.extend import OperatorExecutor, register_executor

cudnn_sdpa_ex = OperatorExecutor("cudnn_sdpa", version="0.1")
register_executor(cudnn_sdpa_ex)


def _fwd_meta(q, k, v, is_causal, scale):
    B, H, T, _ = q.shape
    out = TensorProxy(like=q, shape=(B, H, T, v.shape[-1]))
    lse = TensorProxy(like=q, shape=(B, H, T, 1), dtype=dtypes.float32)
    seed = TensorProxy(like=q, shape=(), dtype=dtypes.int64)
    offset = TensorProxy(like=q, shape=(), dtype=dtypes.int64)
    return out, lse, seed, offset


def _fwd_impl(q, k, v, is_causal, scale):
    out, lse, _, _, _, _, seed, offset, _ = torch.ops.aten._scaled_dot_product_cudnn_attention(
        q, k, v, None, True, 0.0, is_causal, False, scale=scale
    )
    return out, lse, seed, offset


def _bwd_meta(g, q, k, v, out, lse, seed, offset, is_causal, scale):
    return TensorProxy(like=q), TensorProxy(like=k), TensorProxy(like=v)


def _bwd_impl(g, q, k, v, out, lse, seed, offset, is_causal, scale):
    T, S = q.shape[2], k.shape[2]
    return torch.ops.aten._scaled_dot_product_cudnn_attention_backward(
        g.contiguous(), q, k, v, out, lse, seed, offset, None, None, None, T, S, 0.0, is_causal, scale=scale
    )


cudnn_sdpa_fwd = cudnn_sdpa_ex.register_operator("cudnn_sdpa_fwd", meta=_fwd_meta, fn=_fwd_impl)
cudnn_sdpa_bwd = cudnn_sdpa_ex.register_operator("cudnn_sdpa_bwd", meta=_bwd_meta, fn=_bwd_impl)


def _checker(query, key, value, attn_mask=None, dropout_p=0.0, is_causal=False, *, scale=None) -> bool:
    if attn_mask is not None or dropout_p != 0.0:
        return False
    ts = (query, key, value)
    if any(t.ndim != 4 or str(t.device).startswith("cpu") for t in ts):
        return False
    return all(dtypes.to_strong_dtype(t.dtype) in (dtypes.bfloat16, dtypes.float16) for t in ts) and query.shape[-1] <= 256 and query.shape[-1] % 8 == 0


def _transform(query, key, value, attn_mask=None, dropout_p=0.0, is_causal=False, *, scale=None):
    out, _, _, _ = cudnn_sdpa_fwd(query, key, value, is_causal, scale)
    return out


def _grad(query, key, value, attn_mask=None, dropout_p=0.0, is_causal=False, *, scale=None):
    out, lse, seed, offset = cudnn_sdpa_fwd(query, key, value, is_causal, scale)
    g = get_grad(out)
    dq, dk, dv = cudnn_sdpa_bwd(g, query, key, value, out, lse, seed, offset, is_causal, scale)
    put_grads((query, key, value), (dq, dk, dv))
    return out


cudnn_sdpa_ex.register_implementation(
    ltorch.scaled_dot_product_attention, checker=_checker, execution_transform=_transform, grad_transform=_grad
)
