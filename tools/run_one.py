"""Run one workload's step a few times (short command for ncu).  usage: python tools/run_one.py [workload] [iters]"""
import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [root, os.path.join(root, "lightning-thunder_b200")]
import torch
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "rmsnorm"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 6
wl = bench.WORKLOADS[name]()
dev = torch.device("cuda", 0)
wl.alloc(dev)
for i in range(iters):
    wl.step(i)
torch.cuda.synchronize()
print("ok", name, iters)
