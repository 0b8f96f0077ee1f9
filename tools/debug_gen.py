"""In-kernel random leaves vs the materialised form, at the PwLaunch level."""
import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(root, "lightning-thunder_b200"), root]
import torch
from b200ex.pointwise import compile_chain, PwNode, PwLaunch
from b200ex import ops
BF = torch.bfloat16
for shape in ((64, 128), (512, 768), (8,), (3, 5, 16)):
    nodes = [PwNode("lt", "m", torch.bool, shape, ("u", 0.9)), PwNode("convert", "mf", BF, shape, ("m",)), PwNode("mul", "y", BF, shape, ("x", "mf")), PwNode("add", "o", BF, shape, ("y", "u"))]
    (spec,) = compile_chain({"x": (BF, shape), "u": (BF, shape)}, nodes, ["o"], generated={"u": (0.0, 1.0)})
    pw = PwLaunch(spec, {"u": (shape, BF, 0.0, 1.0, "seed", "offset")})
    x = torch.randn(shape, device="cuda", dtype=BF)
    env = {"x": x, "seed": 1234567, "offset": 12}
    a = pw.run(dict(env))["o"]
    b = pw.run(dict(env), _in_kernel_rng=False)["o"]
    u = ops.uniform_philox(shape, 0.0, 1.0, device="cuda", dtype=BF, seed=1234567, offset=12)
    ref = (x * (u < 0.9).to(BF)) + u
    torch.cuda.synchronize()
    print(shape, "jit launches", pw.jit_launches, "in-kernel == materialised:", torch.equal(a, b), " materialised == eager:", torch.equal(b, ref), " in-kernel == eager:", torch.equal(a, ref))
    nodes2 = [PwNode("add", "o", BF, shape, ("u", 0.0))]
    (spec2,) = compile_chain({"u": (BF, shape)}, nodes2, ["o"], generated={"u": (0.0, 1.0)})
    pw2 = PwLaunch(spec2, {"u": (shape, BF, 0.0, 1.0, "seed", "offset")})
    g = pw2.run(dict(env))["o"]
    torch.cuda.synchronize()
    badg = (g != u).flatten().nonzero().flatten()
    print("   generated-only program == kernel:", torch.equal(g, u), "mismatches", badg.numel(), "first", badg[:8].tolist(), "last", badg[-3:].tolist())
    if not torch.equal(a, b):
        bad = (a != b).flatten().nonzero().flatten()[:6].tolist()
        print("   first diffs", bad, [(a.flatten()[i].item(), b.flatten()[i].item()) for i in bad])
