"""Times the generated region kernels (b200ex/regionc.py) that serve the reference's own plain-torch model pieces:
the LitGPT QKV-split + RoPE slice at the Llama-3-8B shape and litgpt's RMSNorm with residual at [8192 x 4096], each
forward and backward, as thunder.jit(executors=[b200]) callables.  CUDA events over a CUDA graph of back-to-back
calls on rotating buffers (> L2); bytes are the region's inputs + outputs, each tensor once."""

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from ref_models import litgpt_rmsnorm, qkv_split_rope  # noqa: E402
from thunder_env import fusions, import_thunder  # noqa: E402

thunder, b200 = import_thunder()
DEV = "cuda:0"
NSETS = 4


def region_callables(jf):
    """[(name, B200Region, input proxies, output proxies)] of the forward and backward trace."""
    out = []
    for tr in (thunder.last_traces(jf)[-1], thunder.last_backward_traces(jf)[-1]):
        for bs in tr.bound_symbols:
            if bs.sym.is_fusion and bs.sym.name.startswith("b200Fusion"):
                out.append((bs.sym.name, bs._call_ctx[bs.sym.name], list(bs.flat_proxy_args), list(bs.flat_proxy_outs)))
    return out


def time_region(region, ins, outs, iters=20):
    sets = []
    for s in range(NSETS):
        g = torch.Generator(device=DEV).manual_seed(s)
        args = []
        for p in ins:
            dt = thunder.core.dtypes.to_torch_dtype(p.dtype)
            t = torch.randn(tuple(p.shape), generator=g, device=DEV, dtype=torch.float32)
            args.append((t.abs() + 0.5).to(dt) if "rstd" in p.name else t.to(dt))
        sets.append(args)
    nbytes = sum(a.numel() * a.element_size() for a in sets[0])
    res = region(*sets[0])
    seen = set()
    for r in res:  # outputs that are views of another output are counted once
        if r.data_ptr() not in seen:
            seen.add(r.data_ptr())
            nbytes += r.numel() * r.element_size()
    if os.environ.get("B200_PROFILE"):  # under ncu: one launch per region is all that is wanted
        torch.cuda.synchronize()
        return float("nan"), nbytes
    for i in range(3):
        region(*sets[i % NSETS])
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        keep = [region(*sets[i % NSETS]) for i in range(NSETS * 2)]
    graph.replay()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        graph.replay()
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / (iters * NSETS * 2)
    return us, nbytes


def main():
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    g = torch.Generator(device=DEV).manual_seed(0)
    rows = []
    # QKV split + RoPE, Llama-3-8B
    B, T, nh, G, hs = 2, 4096, 32, 8, 128
    qkv = torch.randn(B, T, (nh + 2 * G) * hs, generator=g, device=DEV).to(torch.bfloat16).requires_grad_(True)
    cos, sin = (torch.randn(T, hs, generator=g, device=DEV).to(torch.bfloat16) for _ in range(2))
    jf = thunder.jit(lambda qkv, cos, sin: qkv_split_rope(qkv, cos, sin, nh, G, hs, hs), executors=[b200.ex])
    q, k, v = jf(qkv, cos, sin)
    torch.autograd.grad([q, k, v], [qkv], [torch.ones_like(q), torch.ones_like(k), torch.ones_like(v)])
    for name, region, ins, outs in region_callables(jf):
        us, nb = time_region(region, ins, outs)
        rows.append(("qkv_split_rope (plain torch ops) " + name, region.launch_kinds, us, nb))
    if os.environ.get("B200_BENCH_ONLY") == "rope":
        return report(rows, peak)
    # hand-spelled RMSNorm + residual
    R, H = 8192, 4096
    x, r = (torch.randn(R, H, generator=g, device=DEV).to(torch.bfloat16).requires_grad_(True) for _ in range(2))
    w = torch.randn(H, generator=g, device=DEV).to(torch.bfloat16).requires_grad_(True)

    def fn(x, r, w):
        h = x + r
        return h, litgpt_rmsnorm(h, w)

    jf = thunder.jit(fn, executors=[b200.ex])
    h, y = jf(x, r, w)
    torch.autograd.grad([h, y], [x, r, w], [torch.ones_like(h), torch.ones_like(y)])
    for name, region, ins, outs in region_callables(jf):
        us, nb = time_region(region, ins, outs)
        rows.append(("litgpt RMSNorm + residual (plain torch ops) " + name, region.launch_kinds, us, nb))
    # LayerNorm spelled with torch ops (mean / var / rsqrt), [8192 x 4096] bf16
    xl = torch.randn(R, H, generator=g, device=DEV).to(torch.bfloat16).requires_grad_(True)
    wl, bl = (torch.randn(H, generator=g, device=DEV).to(torch.bfloat16).requires_grad_(True) for _ in range(2))

    def ln(x, w, b):
        xf = x.float()
        mu = xf.mean(-1, keepdim=True)
        var = ((xf - mu) ** 2).mean(-1, keepdim=True)
        return ((xf - mu) * torch.rsqrt(var + 1e-5) * w.float() + b.float()).to(x.dtype)

    jf = thunder.jit(ln, executors=[b200.ex])
    yl = jf(xl, wl, bl)
    torch.autograd.grad([yl], [xl, wl, bl], [torch.ones_like(yl)])
    for name, region, ins, outs in region_callables(jf):
        us, nb = time_region(region, ins, outs)
        rows.append(("LayerNorm (plain torch ops) " + name, region.launch_kinds, us, nb))
    # softmax over the last dim spelled with exp / sum, [16384 x 2048] bf16 (amax's backward stays partly with torch)
    xs = torch.randn(16384, 2048, generator=g, device=DEV).to(torch.bfloat16).requires_grad_(True)

    def sm(x):
        xf = x.float()
        e = torch.exp(xf - xf.amax(-1, keepdim=True))
        return (e / e.sum(-1, keepdim=True)).to(x.dtype)

    jf = thunder.jit(sm, executors=[b200.ex])
    ys = jf(xs)
    torch.autograd.grad([ys], [xs], [torch.ones_like(ys)])
    for name, region, ins, outs in region_callables(jf):
        us, nb = time_region(region, ins, outs)
        rows.append(("softmax (plain torch ops) " + name, region.launch_kinds, us, nb))
    report(rows, peak)


def report(rows, peak):
    print(f"{'region':64s} {'launches':40s} {'us':>8s} {'MB':>8s} {'GB/s':>8s} {'frac':>6s}")
    for name, k, us, nb in rows:
        print(f"{name:64s} {','.join(k):40s} {us:8.1f} {nb / 1e6:8.1f} {nb / us / 1e3:8.0f} {nb / us / 1e3 / peak:6.2f}")


if __name__ == "__main__":
    main()
