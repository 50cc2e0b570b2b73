"""torch.profiler view of the end-to-end step (development aid): which device kernels / memsets sit beside ours."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
w = bench.build_workload(list(range(8)), dev)
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
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=30, max_name_column_width=60))
import time
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    bench.step_e2e(gen, host, out_img, out_loss, dev)
t1 = time.perf_counter()          # CPU-side enqueue time only (no sync inside the step)
torch.cuda.synchronize()
t2 = time.perf_counter()
print("cpu enqueue ms/step %.2f   wall ms/step %.2f" % ((t1 - t0) * 100, (t2 - t0) * 100))
evs = [e for e in prof.events()]
syncs = [e for e in evs if e.name in ("cudaStreamSynchronize", "cudaDeviceSynchronize", "cudaEventSynchronize")]
for s in syncs:
    rng = s.time_range
    enc = [e for e in evs if e is not s and e.time_range.start <= rng.start and e.time_range.end >= rng.end
           and e.device_type == s.device_type]
    enc.sort(key=lambda e: e.time_range.end - e.time_range.start)
    print(s.name, "%.0f us" % (rng.end - rng.start), "<-", [e.name[:40] for e in enc[:6]])
