"""torch.profiler view of the end-to-end step (development aid): which device kernels / memsets sit beside ours."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
w = bench.build_workload(0, 8, dev)
gen, host, out_img, out_loss, h2d, d2h = bench.build_e2e(w, dev)
for _ in range(3):
    bench.step_e2e(gen, host, out_img, out_loss, dev)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        bench.step_e2e(gen, host, out_img, out_loss, dev)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
