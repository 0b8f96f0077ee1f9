"""Attention backends of torch.nn.functional.scaled_dot_product_attention at the Llama-3-8B shape (library kernels;
attention is outside this repo's path — the numbers say which backend the training step should ask torch for)."""
import torch, torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel
q, k, v = (torch.randn(2, 32, 4096, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True) for _ in range(3))
g = torch.randn_like(q)
for name, be in (("flash", SDPBackend.FLASH_ATTENTION), ("cudnn", SDPBackend.CUDNN_ATTENTION), ("efficient", SDPBackend.EFFICIENT_ATTENTION)):
    try:
        with sdpa_kernel(be):
            for _ in range(3):
                o = F.scaled_dot_product_attention(q, k, v, is_causal=True); o.backward(g)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10):
                o = F.scaled_dot_product_attention(q, k, v, is_causal=True); o.backward(g)
            b.record(); torch.cuda.synchronize()
        print(f"{name:10s} fwd+bwd {a.elapsed_time(b) / 10:7.3f} ms")
    except Exception as e:
        print(f"{name:10s} unavailable: {type(e).__name__}: {str(e)[:120]}")
