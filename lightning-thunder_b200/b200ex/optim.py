"""AdamW on the b200k multi-tensor kernel: a drop-in for ``torch.optim.AdamW(..., fused=True)``, the optimizer of the
reference's training benchmark (thunder/benchmarks/benchmark_litgpt.py:278-282).

One launch per (device, dtype) group updates every parameter (b200k_adamw, csrc/adamw.cu) with ATen's arithmetic;
state layout and ``state_dict`` keys are torch's (``step``, ``exp_avg``, ``exp_avg_sq``), so checkpoints move between
the two.  CUDA parameters of fp32 / bf16 / fp16 only; anything else raises — there is no eager fallback.
"""

from __future__ import annotations

import math

import torch

from . import _lib
from .ops import DTYPE_CODE, _on

__all__ = ["AdamW"]


class _Group:
    """Launch data of one (device, dtype) group: a pinned host table whose grad column is refreshed every step (grads
    are new tensors after every backward), its device copy, and the static chunk list."""

    def __init__(self, params, states, device, dtype):
        self.params, self.states, self.device, self.dtype = params, states, device, dtype
        chunk = _lib.lib().b200k_adamw_chunk_elems(DTYPE_CODE[dtype])
        n = len(params)
        # two pinned staging buffers used alternately; an event per buffer says when its upload has been consumed
        self.hosts = [torch.zeros(n, 5, dtype=torch.int64).pin_memory() for _ in range(2)]
        self.events = [None, None]
        self.turn = 0
        tensor_of, index_of = [], []
        for i, p in enumerate(params):
            for h in self.hosts:
                h[i, 0], h[i, 2], h[i, 3], h[i, 4] = p.data_ptr(), states[i]["exp_avg"].data_ptr(), states[i]["exp_avg_sq"].data_ptr(), p.numel()
            k = (p.numel() + chunk - 1) // chunk
            tensor_of += [i] * k
            index_of += range(k)
        self.n_chunks = len(tensor_of)
        self.table = torch.empty(n, 5, dtype=torch.int64, device=device)
        self.chunk_tensor = torch.tensor(tensor_of, dtype=torch.int32, device=device)
        self.chunk_index = torch.tensor(index_of, dtype=torch.int32, device=device)
        self.ptrs = _table_key(params, states)


def _table_key(params, states) -> list:
    """What the cached device table depends on: the addresses of the parameters AND of their moments.
    ``load_state_dict`` / ``state.clear()`` replace the moment tensors; a table keyed on the parameters alone would
    keep updating the old ones."""
    return [(p.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()) for p, st in zip(params, states)]


class AdamW(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2, decoupled_weight_decay: bool = True):
        if lr < 0 or eps < 0 or not 0 <= betas[0] < 1 or not 0 <= betas[1] < 1 or weight_decay < 0:
            raise ValueError("invalid AdamW hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, decoupled_weight_decay=decoupled_weight_decay))
        self._groups: dict = {}

    def _state_for(self, p):
        st = self.state[p]
        if not st:
            st["step"] = torch.zeros((), dtype=torch.float32)  # host scalar, same key and dtype as torch's
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _lib.lib()
        for gi, group in enumerate(self.param_groups):
            buckets: dict = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype not in DTYPE_CODE or p.grad.dtype != p.dtype or p.grad.is_sparse:
                    raise RuntimeError("b200ex.optim.AdamW takes dense CUDA parameters of fp32 / bf16 / fp16 with same-dtype gradients")
                if not p.is_contiguous() or not p.grad.is_contiguous():
                    raise RuntimeError("b200ex.optim.AdamW needs contiguous parameters and gradients")
                buckets.setdefault((p.device, p.dtype), []).append(p)
            beta1, beta2 = group["betas"]
            for (device, dtype), params in buckets.items():
                states = [self._state_for(p) for p in params]
                for p, st in zip(params, states):
                    for k in ("exp_avg", "exp_avg_sq"):
                        m = st[k]
                        if m.device != p.device or m.dtype != p.dtype or m.shape != p.shape or not m.is_contiguous():
                            raise RuntimeError(f"b200ex.optim.AdamW: state '{k}' must be a contiguous tensor of the parameter's shape, dtype and device")
                steps = {float(st["step"]) for st in states}
                if len(steps) != 1:
                    raise RuntimeError("b200ex.optim.AdamW: parameters of one group must share their step count")
                step = steps.pop() + 1.0
                for st in states:
                    st["step"] += 1
                key = (gi, device, dtype)
                g = self._groups.get(key)
                if g is None or g.ptrs != _table_key(params, states):
                    g = self._groups[key] = _Group(params, states, device, dtype)
                host, ev = g.hosts[g.turn], g.events[g.turn]
                if ev is not None:
                    ev.synchronize()  # the upload that last used this staging buffer has finished
                host[:, 1] = torch.tensor([p.grad.data_ptr() for p in params], dtype=torch.int64)
                with _on(device):
                    g.table.copy_(host, non_blocking=True)
                    g.events[g.turn] = torch.cuda.Event()
                    g.events[g.turn].record(torch.cuda.current_stream(device))
                    g.turn ^= 1
                    rc = L.b200k_adamw(
                        g.table.data_ptr(), g.chunk_tensor.data_ptr(), g.chunk_index.data_ptr(), g.n_chunks,
                        float(group["lr"]), float(beta1), float(beta2), float(group["eps"]), float(group["weight_decay"]),
                        1.0 - beta1**step, math.sqrt(1.0 - beta2**step), int(bool(group["decoupled_weight_decay"])), DTYPE_CODE[dtype],
                        torch.cuda.current_stream(device).cuda_stream,
                    )
                _lib.check(rc, "b200k_adamw")
        return loss
