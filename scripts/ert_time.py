"""Early ray termination on an OPAQUE synthetic scene at config-2 size (B=8, 128x128, 48+48): kernel times and the share
of decoder evaluations skipped, ert_eps = 0 vs 1e-6.   python scripts/ert_time.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from enarf_gan_b200 import ops

dev = torch.device("cuda", 0)
w = bench.build_workload(list(range(8)), dev)
w["tri_plane"][:, 96:] += 6.0
w["decoder"][2] = (w["decoder"][2][0], w["decoder"][2][1] + torch.tensor([0, 0, 0, 30.0], device=dev))
scene = ops.SceneBuffers(w["pose"], w["cscale"], w["canon"], w["tri_plane"], w["decoder"])
dmin, dmax, direction, _ = ops.depth_range(w["img"], w["inv_K"], w["pose"])


def timed(fn, n=10):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n, out


for eps in (0.0, 1e-6):
    cnt = torch.zeros(8, device=dev, dtype=torch.int32)
    always = None if eps > 0 else None
    tc, wb = timed(lambda: ops.render_coarse(scene, dmin, dmax, direction, 48, opaque_rays=cnt[:1], ert_eps=eps, ert_hint=None,
                                             ert_stats=cnt[1:3] if eps else None))
    bins = ops.fine_sample(wb, 48, seed=1)
    tf, out = timed(lambda: ops._render_fine_fwd(scene, dmin, dmax, direction, bins, want_points=True, want_feat=True,
                                                 want_pairs=True, ert_eps=eps, ert_hint=None, ert_stats=cnt[3:5] if eps else None))
    c = cnt.tolist()
    print(f"ert_eps {eps:g}: coarse {tc:.3f} ms, fine fwd {tf:.3f} ms" + (
        f"; coarse skipped {c[2] / max(1, c[1] + c[2]):.1%}, fine skipped {c[4] / max(1, c[3] + c[4]):.1%}" if eps else ""))
