"""Short config-2 forward+backward for ncu captures (development aid)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
w = bench.build_workload(0, int(os.environ.get("B", 8)), dev)
for i in range(int(os.environ.get("ITERS", 2))):
    bench.step_device(w, i)
torch.cuda.synchronize()
print("ok", w["last"][1].mean().item())
