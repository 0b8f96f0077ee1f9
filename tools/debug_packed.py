"""Show elements where the packed bf16x2 GELU backward differs from the fp32 form (two processes: the switch is read once)."""
import os, subprocess, sys, tempfile
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(root, "lightning-thunder_b200"), root]
import torch

if len(sys.argv) > 1:
    from b200ex import ops
    d = torch.load(sys.argv[1])
    x, b, g = (d[k].cuda() for k in ("x", "b", "g"))
    dx, db = ops.bias_act_bwd(x, b, g, "gelu")
    torch.save({"dx": dx.cpu()}, sys.argv[2])
    sys.exit(0)

g0 = torch.Generator().manual_seed(5)
x = torch.randn(8, 2048, generator=g0).to(torch.bfloat16)
b = torch.zeros(2048, dtype=torch.bfloat16)
g = torch.randn(8, 2048, generator=g0).to(torch.bfloat16)
tmp = tempfile.mkdtemp()
torch.save({"x": x, "b": b, "g": g}, f"{tmp}/in.pt")
outs = {}
for flag in ("1", "0"):
    env = dict(os.environ, B200K_BIAS_ACT_PACKED=flag)
    subprocess.run([sys.executable, __file__, f"{tmp}/in.pt", f"{tmp}/out{flag}.pt"], check=True, env=env)
    outs[flag] = torch.load(f"{tmp}/out{flag}.pt")["dx"]
from oracle import regions as O
edx, _ = O.bias_gelu_bwd(x, b, g)
diff = (outs["1"].view(torch.int16) != outs["0"].view(torch.int16))
print("differing:", diff.sum().item(), "of", diff.numel(), "; packed==oracle:", (outs["1"] == edx).float().mean().item(), " fp32form==oracle:", (outs["0"] == edx).float().mean().item())
idx = diff.flatten().nonzero().flatten()[:12]
for i in idx.tolist():
    print(f"x={x.flatten()[i].item():+.6f} g={g.flatten()[i].item():+.6f} packed={outs['1'].flatten()[i].item():+.6f} fp32form={outs['0'].flatten()[i].item():+.6f} oracle={edx.flatten()[i].item():+.6f}")
