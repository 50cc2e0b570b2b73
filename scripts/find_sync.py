"""Locate host<->device synchronisation points in the end-to-end step (development aid)."""
import os, sys, warnings, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
dev = torch.device("cuda:0"); torch.cuda.set_device(0)
w = bench.build_workload(list(range(8)), dev)
gen, host, out_img, out_loss, h2d, d2h = bench.build_e2e(w, dev)
for _ in range(2): bench.step_e2e(gen, host, out_img, out_loss, dev)
torch.cuda.synchronize()
def showwarning(message, category, filename, lineno, file=None, line=None):
    print("WARN:", message)
    traceback.print_stack(limit=7)
warnings.showwarning = showwarning
warnings.simplefilter("always")
torch.cuda.set_sync_debug_mode("warn")
bench.step_e2e(gen, host, out_img, out_loss, dev)
torch.cuda.set_sync_debug_mode("default")
torch.cuda.synchronize()
