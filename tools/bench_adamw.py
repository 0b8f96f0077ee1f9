"""AdamW step over a Llama-3-8B-like parameter set (bf16, 1/4 of the layers): b200k_adamw vs torch fused AdamW."""
import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(root, "lightning-thunder_b200"), root]
import torch
from b200ex.optim import AdamW
shapes = [(128256, 4096)] + [(4096,), (6144, 4096), (4096, 4096), (4096,), (14336, 4096), (14336, 4096), (4096, 14336)] * 8 + [(4096,)]
def make():
    ps = [torch.nn.Parameter(torch.randn(s, device="cuda", dtype=torch.bfloat16) * 0.02) for s in shapes]
    for p in ps: p.grad = torch.randn_like(p) * 0.01
    return ps
n = sum(torch.Size(s).numel() for s in shapes)
for name, ctor in (("b200k_adamw", lambda ps: AdamW(ps, lr=1e-5)), ("torch fused", lambda ps: torch.optim.AdamW(ps, lr=1e-5, fused=True))):
    ps = make(); opt = ctor(ps)
    for _ in range(3): opt.step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): opt.step()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print(f"{name:12s} {n / 1e9:.2f} B params: {ms:7.3f} ms/step = {n * 14 / ms / 1e6:7.0f} GB/s (14 B/param)")
    del ps, opt; torch.cuda.empty_cache()
