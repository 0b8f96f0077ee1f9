"""LitGPT-style Llama model restated for the benchmarks and tests (litgpt is an un-vendored dependency of the
reference and is not installable here).

Semantics follow what the reference exercises: RMSNorm via ``F.rms_norm`` (thunder/torch/__init__.py:4450),
the fused-QKV attention slice vendored at thunder/benchmarks/__init__.py:2282-2328 (here the plain-PyTorch
function ``b200ex.executor.qkv_split_rope``, which traces to the same ops when the b200 executor is absent),
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


def _default_qkv_split_rope():
    from b200ex.executor import qkv_split_rope  # plain PyTorch definition; b200 swaps in its fused operator

    return qkv_split_rope


class RMSNorm(nn.Module):
    def __init__(self, size: int, eps: float):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(size))
        self.eps, self.size = eps, size

    def forward(self, x):
        return F.rms_norm(x, (self.size,), self.weight, self.eps)


class CausalSelfAttention(nn.Module):
    def __init__(self, cfg: LlamaConfig, qkv_split_rope=None):
        super().__init__()
        self.cfg = cfg
        self.qkv = nn.Linear(cfg.n_embd, (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size, bias=False)
        self.proj = nn.Linear(cfg.n_head * cfg.head_size, cfg.n_embd, bias=False)
        self.qkv_split_rope = qkv_split_rope or _default_qkv_split_rope()

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
        self.norm_1 = RMSNorm(cfg.n_embd, cfg.norm_eps)
        self.attn = CausalSelfAttention(cfg, qkv_split_rope)
        self.norm_2 = RMSNorm(cfg.n_embd, cfg.norm_eps)
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
        self.ln_f = RMSNorm(cfg.n_embd, cfg.norm_eps)
        self.lm_head = nn.Linear(cfg.n_embd, cfg.padded_vocab_size, bias=False)

    def forward(self, idx, cos, sin, targets=None):
        x = self.wte(idx)
        for block in self.h:
            x = block(x, cos, sin)
        logits = self.lm_head(self.ln_f(x))
        if targets is None:
            return logits
        return F.cross_entropy(logits.view(-1, logits.size(-1)), targets.view(-1), ignore_index=-1)
