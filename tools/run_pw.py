import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [root, os.path.join(root, "lightning-thunder_b200")]
import torch
from b200ex.pointwise import PwLaunch, PwNode, compile_chain
BF, F32 = torch.bfloat16, torch.float32
s = (8192, 4096)
nodes = [PwNode("convert", "xf", F32, s, ("x",)), PwNode("convert", "yf", F32, s, ("y",)), PwNode("add", "a", F32, s, ("xf", "yf")), PwNode("convert", "o", BF, s, ("a",))]
(spec,) = compile_chain({"x": (BF, s), "y": (BF, s)}, nodes, ["o"])
pw = PwLaunch(spec)
env = {"x": torch.randn(*s, device="cuda").to(BF), "y": torch.randn(*s, device="cuda").to(BF)}
for _ in range(4):
    pw.run(env)
torch.cuda.synchronize()
print("ok")
