"""Short config-2 forward for ncu captures (development aid)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from enarf_gan_b200 import ops, rendering, synthetic as S, hostmath
B, size, Nc, Nf = int(os.environ.get("B", 8)), 128, 48, 48
dev = torch.device("cuda:0")
inp = S.make_inputs(B, size, seed=1236)
tp = S.make_triplane(B, seed=1236).to(dev)
scene = hostmath.random_scene(inp, tp, seed=1236, device=dev)
img = rendering.pixel_grid(size, size, B, dev)
invK = inp["inv_intrinsics"].to(dev)
for _ in range(int(os.environ.get("ITERS", 2))):
    out = rendering.render_rays(scene, img, invK, Nc, Nf, rng="philox", seed=1)
torch.cuda.synchronize()
print("ok", out["mask"].mean().item())
