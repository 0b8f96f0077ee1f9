"""Dropout region timing: b200 (mask generated in registers) vs the reference's torch executor, bf16 [8192 x 4096]."""
import os, sys, warnings
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(root, "tests"), root]
warnings.simplefilter("ignore")
import torch, torch.nn.functional as F
from thunder_env import import_thunder, fusions
thunder, b200 = import_thunder()
x = torch.randn(8192, 4096, device="cuda", dtype=torch.bfloat16, requires_grad=True)
r = torch.randn(8192, 4096, device="cuda", dtype=torch.bfloat16)
cot = torch.randn(8192, 4096, device="cuda", dtype=torch.bfloat16)
f = lambda x, r: F.dropout(x, 0.1) + r
for name, execs in (("b200", [b200.ex]), ("torch executor", ["torch"])):
    jf = thunder.jit(f, executors=execs)
    def step():
        y = jf(x, r)
        torch.autograd.grad(y, x, cot)
    for _ in range(5): step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50): step()
    b.record(); torch.cuda.synchronize()
    print(f"{name:16s} fwd+bwd {a.elapsed_time(b) / 50 * 1e3:8.1f} us   peak mem {torch.cuda.max_memory_allocated() / 2**20:.0f} MiB", 
          fusions(thunder.last_traces(jf)[-1]) if name == "b200" else "")
    torch.cuda.reset_peak_memory_stats()
