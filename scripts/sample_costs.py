"""Per-sample render cost of the bench batch (one sample per launch): kernel times and (point, part) pair counts, to
calibrate rendering.sample_cost_proxy (the host-side load-balancing estimate of the strong-scaling split).
    python scripts/sample_costs.py > gpurun_out/sample_costs.json"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    G = bench.GLOBAL_BATCH
    rows = []
    for i in range(G):
        w = bench.build_workload([i], dev, G)
        for k in range(3):
            bench.step_device(w, k)
        torch.cuda.synchronize()
        timer = bench.KernelTimer()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for k in range(5):
            bench.step_device(w, 100 + k, timer)
        t1.record()
        torch.cuda.synchronize()
        (vc, ac), (vf, af) = bench.mean_valid_parts(w)
        n = w["n"]
        valid = int(w["last"][5].numel())
        rows.append({"sample": i, "ms": t0.elapsed_time(t1) / 5, "kernels": timer.totals(),
                     "pairs_coarse": vc * n * bench.NC, "pairs_fine": vf * n * bench.NF,
                     "active_coarse": ac * n * bench.NC, "active_fine": af * n * bench.NF})
        del w
    print(json.dumps(rows))


if __name__ == "__main__":
    main()
