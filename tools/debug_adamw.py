"""Mismatch statistics of b200ex.optim.AdamW against torch.optim.AdamW(fused=True), per dtype / step / tensor."""
import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(root, "lightning-thunder_b200"), root]
import torch
from b200ex.optim import AdamW
for dtype in (torch.bfloat16, torch.float16, torch.float32):
    torch.manual_seed(0)
    p0 = (torch.randn(1 << 16) * 0.5).to(dtype)
    ours, theirs = [torch.nn.Parameter(p0.clone().cuda())], [torch.nn.Parameter(p0.clone().cuda())]
    a = AdamW(ours, lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    b = torch.optim.AdamW(theirs, lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, fused=True)
    for step in range(1, 6):
        g = (torch.randn(1 << 16) * 0.1).to(dtype).cuda()
        ours[0].grad, theirs[0].grad = g.clone(), g.clone()
        a.step(); b.step(); torch.cuda.synchronize()
        out = []
        for name, x, y in (("p", ours[0].data, theirs[0].data), ("m", a.state[ours[0]]["exp_avg"], b.state[theirs[0]]["exp_avg"]), ("v", a.state[ours[0]]["exp_avg_sq"], b.state[theirs[0]]["exp_avg_sq"])):
            d = (x.float() - y.float()).abs()
            rel = (d / (y.float().abs() + 1e-30)).max().item()
            out.append(f"{name}: {(d != 0).float().mean().item() * 100:6.3f}% differ, max rel {rel:.2e}")
        print(str(dtype)[6:], "step", step, " | ".join(out))
