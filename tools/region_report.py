"""Per-region report of a b200-compiled module: every b200 fusion of the forward and backward trace with its launch
kinds, the shapes / strides / dtypes its inputs actually had at run time, and its device time (CUDA events around 5
calls on those same inputs).  Used to find the region that is slower than its bytes explain.

    python tools/region_report.py [litgpt|ops] [layers]
"""

import dataclasses
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from thunder_env import import_thunder  # noqa: E402

thunder, b200 = import_thunder()
from baseline.llama_model import LLAMA3_8B, Block, build_rope_cache  # noqa: E402
from thunder.executors.sdpaex import sdpa_ex  # noqa: E402

DEV = "cuda:0"


def main():
    spelling = sys.argv[1] if len(sys.argv) > 1 else "litgpt"
    cfg = dataclasses.replace(LLAMA3_8B, spelling=spelling)
    torch.manual_seed(0)
    block = Block(cfg).to(DEV).to(torch.bfloat16)
    B, T = 2, 4096
    cos, sin = build_rope_cache(T, cfg.rope_n_elem, base=cfg.rope_base, device=DEV, dtype=torch.bfloat16)
    x = (torch.randn(B, T, cfg.n_embd, device=DEV) * 0.5).to(torch.bfloat16).requires_grad_(True)
    seen = {}
    orig = b200.B200Region.__call__

    def spy(self, *args):
        if self.name not in seen:
            seen[self.name] = (self, [a.detach() if isinstance(a, torch.Tensor) else a for a in args])
        return orig(self, *args)

    b200.B200Region.__call__ = spy
    attn = [sdpa_ex]
    try:
        from baseline.attention import cudnn_sdpa_ex

        attn = [cudnn_sdpa_ex, sdpa_ex]
    except Exception:
        pass
    jb = thunder.jit(block, executors=attn + [b200.ex])
    y = jb(x, cos, sin)
    y.backward(torch.randn_like(y))
    torch.cuda.synchronize()
    b200.B200Region.__call__ = orig
    for name, (region, args) in seen.items():
        for _ in range(2):
            region(*args)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(5):
            outs = region(*args)
        b.record()
        torch.cuda.synchronize()
        us = a.elapsed_time(b) * 1e3 / 5
        nbytes = sum(t.numel() * t.element_size() for t in args if isinstance(t, torch.Tensor))
        seen_ptr = set()
        for o in outs:
            if isinstance(o, torch.Tensor) and o.data_ptr() not in seen_ptr:
                seen_ptr.add(o.data_ptr())
                nbytes += o.numel() * o.element_size()
        print(f"{name:14s} {us:9.1f} us {nbytes / 1e6:9.1f} MB {nbytes / us / 1e3:8.0f} GB/s  {region.launch_kinds}")
        for t in args:
            if isinstance(t, torch.Tensor):
                print(f"      in  {tuple(t.shape)} stride {t.stride()} {t.dtype} ptr%16={t.data_ptr() % 16}")
        for o in outs:
            if isinstance(o, torch.Tensor):
                print(f"      out {tuple(o.shape)} stride {o.stride()} {o.dtype}")


if __name__ == "__main__":
    main()
