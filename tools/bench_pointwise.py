"""GB/s of the generic pointwise interpreter on typical short programs (residual add, scale, bias add) at Llama shapes."""
import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [root, os.path.join(root, "lightning-thunder_b200")]
import torch
from b200ex.pointwise import PwLaunch, PwNode, compile_chain

BF, F32 = torch.bfloat16, torch.float32
dev = "cuda:0"
R, H = 8192, 4096
s = (R, H)
progs = {
    "residual add bf16 (x + y)": ([PwNode("convert", "xf", F32, s, ("x",)), PwNode("convert", "yf", F32, s, ("y",)), PwNode("add", "a", F32, s, ("xf", "yf")),
                                   PwNode("convert", "o", BF, s, ("a",))], {"x": (BF, s), "y": (BF, s)}, ["o"], 3 * R * H * 2),
    "grad pre-scale bf16 (g / 8)": ([PwNode("convert", "xf", F32, s, ("x",)), PwNode("div", "a", F32, s, ("xf", 8.0)), PwNode("convert", "o", BF, s, ("a",))],
                                     {"x": (BF, s)}, ["o"], 2 * R * H * 2),
    "bias add + tanh bf16": ([PwNode("broadcast_in_dim", "bb", BF, s, ("b", (1,))), PwNode("convert", "xf", F32, s, ("x",)), PwNode("convert", "bf", F32, s, ("bb",)),
                              PwNode("add", "a", F32, s, ("xf", "bf")), PwNode("tanh", "t", F32, s, ("a",)), PwNode("convert", "o", BF, s, ("t",))],
                             {"x": (BF, s), "b": (BF, (H,))}, ["o"], 2 * R * H * 2),
}
import itertools
for (name, (nodes, meta, outs, nbytes)), use_jit in itertools.product(progs.items(), (True, False)):
    (spec,) = compile_chain(meta, nodes, outs)
    pw = PwLaunch(spec)
    pw.use_jit = use_jit
    name = name + (" [jit]" if use_jit else " [interp]")
    sets = [{k: torch.randn(*shape, device=dev).to(dt) for k, (dt, shape) in meta.items()} for _ in range(4)]
    for i in range(5):
        pw.run(sets[i % 4])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    n = 200
    for i in range(n):
        pw.run(sets[i % 4])
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / n
    print(f"{name:42s} {us:7.1f} us  {nbytes / us / 1e3:6.0f} GB/s  ({len(spec.code)} code words)")
