"""Plain-PyTorch restatements of the reference's own model pieces that the tests jit through b200 and through the
torch executor: the LitGPT attention slice the reference vendors (thunder/benchmarks/__init__.py:2282-2328 with
litgpt.model.apply_rope), litgpt.model.RMSNorm, and the RMSNorm of thunder/tests/llama2_model.py:22-35.  litgpt itself
is not installed here."""

import torch


def litgpt_apply_rope(x, cos, sin):
    head_size = x.size(-1)
    x1 = x[..., : head_size // 2]
    x2 = x[..., head_size // 2:]
    rotated = torch.cat((-x2, x1), dim=-1)
    roped = (x * cos) + (rotated * sin)
    return roped.to(dtype=x.dtype)


def qkv_split_rope(qkv, cos, sin, n_head, n_query_groups, head_size, rope_n_elem):
    B, T, _ = qkv.shape
    q_per_kv = n_head // n_query_groups
    total_qkv = q_per_kv + 2
    qkv = qkv.view(B, T, n_query_groups, total_qkv, head_size)
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


def litgpt_rmsnorm(x, weight, eps=1e-5, add_unit_offset=False):
    """litgpt.model.RMSNorm.forward as the reference's models use it."""
    dtype = x.dtype
    x = x.float()
    norm_x = torch.mean(x * x, dim=-1, keepdim=True)
    x_normed = x * torch.rsqrt(norm_x + eps)
    weight = (1 + weight) if add_unit_offset else weight
    return (x_normed * weight.float()).to(dtype=dtype)


def llama2_rmsnorm(x, weight, eps=1e-5):
    """thunder/tests/llama2_model.py:22-35 (RMSNorm._norm + forward)."""
    xf = x.float()
    output = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).type_as(x)
    return output * weight


