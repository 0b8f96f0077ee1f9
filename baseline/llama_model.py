"""LitGPT-style Llama model restated for the benchmarks and tests (litgpt is an un-vendored dependency of the
reference and is not installable here).

Two spellings of the same model (``LlamaConfig.spelling``):

* ``"litgpt"`` (default) — written the way litgpt writes it: ``RMSNorm.forward`` as ``x.float(); mean(x*x);
  x * rsqrt(.. + eps); * weight.float(); .to(dtype)`` and the attention slice as the plain reshape / permute / split /
  expand / slice / cat code the reference vendors at thunder/benchmarks/__init__.py:2282-2328 with litgpt's
  ``apply_rope``.  Nothing in it names this repository: b200 has to find and fuse the regions itself.
* ``"ops"`` — ``F.rms_norm`` (thunder/torch/__init__.py:4450) and ``b200ex.executor.qkv_split_rope`` (a plain
  PyTorch function that b200 swaps for its fused operator): the op-level route to the hand-written kernels.

``LLaMAMLP`` = ``proj(silu(fc_1(x)) * fc_2(x))``, pre-norm residual blocks, and a loss computed with
``F.cross_entropy`` on the flattened logits.  Baseline / test infrastructure, not product code.
"""

from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class LlamaConfig:
    n_layer: int = 32
    n_head: int = 32
    n_query_groups: int = 8
    n_embd: int = 4096
    intermediate_size: int = 14336
    padded_vocab_size: int = 128256
    block_size: int = 8192
    norm_eps: float = 1e-5
    rope_base: int = 500000
    rotary_percentage: float = 1.0
    spelling: str = "litgpt"

    @property
    def head_size(self) -> int:
        return self.n_embd // self.n_head

    @property
    def rope_n_elem(self) -> int:
        return int(self.rotary_percentage * self.head_size)


LLAMA3_8B = LlamaConfig()


def tiny_config(**kw) -> LlamaConfig:
    base = dict(n_layer=2, n_head=4, n_query_groups=2, n_embd=64, intermediate_size=128, padded_vocab_size=256, block_size=64)
    base.update(kw)
    return LlamaConfig(**base)


def build_rope_cache(seq_len: int, n_elem: int, base: int = 10000, device=None, dtype=torch.float32):
    theta = 1.0 / (base ** (torch.arange(0, n_elem, 2, device=device).float() / n_elem))
    idx = torch.outer(torch.arange(seq_len, device=device).float(), theta).repeat(1, 2)
    return torch.cos(idx).to(dtype), torch.sin(idx).to(dtype)


def _default_qkv_split_rope(spelling: str = "litgpt"):
    if spelling == "litgpt":
        return litgpt_qkv_split_rope
    from b200ex.executor import qkv_split_rope  # plain PyTorch definition; b200 swaps in its fused operator

    return qkv_split_rope


def litgpt_apply_rope(x, cos, sin):
    head_size = x.size(-1)
    x1 = x[..., : head_size // 2]
    x2 = x[..., head_size // 2:]
    rotated = torch.cat((-x2, x1), dim=-1)
    roped = (x * cos) + (rotated * sin)
    return roped.to(dtype=x.dtype)


def litgpt_qkv_split_rope(qkv, cos, sin, n_head, n_query_groups, head_size, rope_n_elem):
    """The slice between the qkv projection and scaled_dot_product_attention, as litgpt writes it
    (thunder/benchmarks/__init__.py:2282-2328)."""
    B, T, _ = qkv.shape
    q_per_kv = n_head // n_query_groups
    qkv = qkv.view(B, T, n_query_groups, q_per_kv + 2, head_size)
    qkv = qkv.permute(0, 2, 3, 1, 4)
    q, k, v = qkv.split((q_per_kv, 1, 1), dim=2)
    if n_query_groups != n_head and n_query_groups != 1:
        k = k.expand(B, n_query_groups, q_per_kv, T, head_size)
        v = v.expand(B, n_query_groups, q_per_kv, T, head_size)
    q = q.reshape(B, -1, T, head_size)
    k = k.reshape(B, -1, T, head_size)
    v = v.reshape(B, -1, T, head_size)
    q_roped = litgpt_apply_rope(q[..., :rope_n_elem], cos, sin)
    k_roped = litgpt_apply_rope(k[..., :rope_n_elem], cos, sin)
    q = torch.cat((q_roped, q[..., rope_n_elem:]), dim=-1)
    k = torch.cat((k_roped, k[..., rope_n_elem:]), dim=-1)
    return q, k, v


class RMSNorm(nn.Module):
    def __init__(self, size: int, eps: float, spelling: str = "litgpt"):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(size))
        self.eps, self.size, self.spelling = eps, size, spelling

    def forward(self, x):
        if self.spelling != "litgpt":
            return F.rms_norm(x, (self.size,), self.weight, self.eps)
        dtype = x.dtype  # litgpt.model.RMSNorm.forward
        x = x.float()
        norm_x = torch.mean(x * x, dim=-1, keepdim=True)
        x_normed = x * torch.rsqrt(norm_x + self.eps)
        return (x_normed * self.weight.float()).to(dtype=dtype)


class CausalSelfAttention(nn.Module):
    def __init__(self, cfg: LlamaConfig, qkv_split_rope=None):
        super().__init__()
        self.cfg = cfg
        self.qkv = nn.Linear(cfg.n_embd, (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size, bias=False)
        self.proj = nn.Linear(cfg.n_head * cfg.head_size, cfg.n_embd, bias=False)
        self.qkv_split_rope = qkv_split_rope or _default_qkv_split_rope(cfg.spelling)

    def forward(self, x, cos, sin):
        B, T, _ = x.shape
        c = self.cfg
        q, k, v = self.qkv_split_rope(self.qkv(x), cos, sin, c.n_head, c.n_query_groups, c.head_size, c.rope_n_elem)
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.proj(y.transpose(1, 2).reshape(B, T, c.n_head * c.head_size))


class LLaMAMLP(nn.Module):
    def __init__(self, cfg: LlamaConfig):
        super().__init__()
        self.fc_1 = nn.Linear(cfg.n_embd, cfg.intermediate_size, bias=False)
        self.fc_2 = nn.Linear(cfg.n_embd, cfg.intermediate_size, bias=False)
        self.proj = nn.Linear(cfg.intermediate_size, cfg.n_embd, bias=False)

    def forward(self, x):
        return self.proj(F.silu(self.fc_1(x)) * self.fc_2(x))


class Block(nn.Module):
    def __init__(self, cfg: LlamaConfig, qkv_split_rope=None):
        super().__init__()
        self.norm_1 = RMSNorm(cfg.n_embd, cfg.norm_eps, cfg.spelling)
        self.attn = CausalSelfAttention(cfg, qkv_split_rope)
        self.norm_2 = RMSNorm(cfg.n_embd, cfg.norm_eps, cfg.spelling)
        self.mlp = LLaMAMLP(cfg)

    def forward(self, x, cos, sin):
        x = self.attn(self.norm_1(x), cos, sin) + x
        return self.mlp(self.norm_2(x)) + x


class GPT(nn.Module):
    """Returns the mean cross-entropy loss when ``targets`` is given (loss inside the jitted callable)."""

    def __init__(self, cfg: LlamaConfig, qkv_split_rope=None):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.padded_vocab_size, cfg.n_embd)
        self.h = nn.ModuleList(Block(cfg, qkv_split_rope) for _ in range(cfg.n_layer))
        self.ln_f = RMSNorm(cfg.n_embd, cfg.norm_eps, cfg.spelling)
        self.lm_head = nn.Linear(cfg.n_embd, cfg.padded_vocab_size, bias=False)

    def forward(self, idx, cos, sin, targets=None):
        x = self.wte(idx)
        for block in self.h:
            x = block(x, cos, sin)
        logits = self.lm_head(self.ln_f(x))
        if targets is None:
            return logits
        return F.cross_entropy(logits.view(-1, logits.size(-1)), targets.view(-1), ignore_index=-1)
