"""Host-side cost of one launch through each path (tiny tensors, so the GPU time is negligible)."""
import os, sys, time
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [root, os.path.join(root, "lightning-thunder_b200")]
import torch
from b200ex import ops
from b200ex.pointwise import PwLaunch, PwNode, compile_chain

dev = "cuda:0"
x = torch.randn(8, 256, device=dev, dtype=torch.bfloat16)
w = torch.randn(256, device=dev, dtype=torch.bfloat16)
(spec,) = compile_chain({"x": (torch.bfloat16, (8, 256)), "y": (torch.bfloat16, (8, 256))}, [PwNode("add", "o", torch.bfloat16, (8, 256), ("x", "y"))], ["o"])
pw = PwLaunch(spec)
env = {"x": x, "y": x}
cases = {
    "torch.add (eager reference point)": lambda: torch.add(x, x),
    "b200k_pointwise via PwLaunch.run": lambda: pw.run(env),
    "ops.rmsnorm_fwd": lambda: ops.rmsnorm_fwd(x, x, w, 1e-5),
    "ops.swiglu_fwd": lambda: ops.swiglu_fwd(x, x),
}
for name, fn in cases.items():
    for _ in range(200):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 3000
    for _ in range(n):
        fn()
    dt = time.perf_counter() - t0
    torch.cuda.synchronize()
    print(f"{name:40s} {dt / n * 1e6:7.1f} us/call (host)")
