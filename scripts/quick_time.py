"""Ad-hoc forward timing of config 2 (B=8, 128^2, 48+48) -- development aid, not the bench."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from enarf_gan_b200 import ops, rendering, synthetic as S
from enarf_gan_b200 import hostmath

B, size, Nc, Nf = int(os.environ.get("B", 8)), 128, 48, 48
dev = torch.device("cuda:0")
inp = S.make_inputs(B, size, seed=1236)
tp = S.make_triplane(B, seed=1236).to(dev)
scene = hostmath.random_scene(inp, tp, seed=1236, device=dev)
img = rendering.pixel_grid(size, size, B, dev)
invK = inp["inv_intrinsics"].to(dev)
def step():
    return rendering.render_rays(scene, img, invK, Nc, Nf, rng="philox", seed=1, return_intermediate=True)
for _ in range(3): out = step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
dmin, dmax, direction = out["depth_min"], out["depth_max"], out["direction"]
n = 10
def t(fn):
    torch.cuda.synchronize(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / n
print("valid rays", out["validity"].float().mean().item(), "mask mean", out["mask"].mean().item())
print("total ms", t(step))
print("depth_range ms", t(lambda: ops.depth_range(img, invK, scene.pose)))
print("coarse ms", t(lambda: ops.render_coarse(scene, dmin, dmax, direction, Nc)))
wb = out["coarse_weights_blurred"]
print("sample ms", t(lambda: ops.fine_sample(wb, Nf, seed=1)))
print("fine ms", t(lambda: ops.render_fine(scene, dmin, dmax, direction, out["bins"])))
print("repack ms", t(lambda: ops.repack_planes(tp)))
pts = torch.randn(B, 3, 1 << 20, device=dev)
_, _, m = ops.density_query(scene, (out["direction"][:, :, :1 << 14] * 9).contiguous(), want_mask=True)
R = B * size * size
print("rays/s fwd", R / (t(step) * 1e-3))
