"""SURVEY config 4 shapes (B=4, 512x512 rays per sample, 96+96 samples, bf16 feature planes): forward + backward once,
timed with CUDA events -- a size check of the fused kernels, not the bench (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from enarf_gan_b200 import ops, rendering, synthetic as S, hostmath
B, size, Nc, Nf = int(os.environ.get("B", 4)), int(os.environ.get("SIZE", 512)), 96, 96
dev = torch.device("cuda:0")
inp = S.make_inputs(B, size, seed=1238)
tp = S.make_triplane(B, seed=1238).to(dev).requires_grad_()
scene0 = hostmath.random_scene(inp, tp.detach(), seed=1238, device=dev)
dec = [(w.clone().requires_grad_(), b.clone().requires_grad_()) for w, b in zip(scene0.w, scene0.b)]
img = rendering.pixel_grid(size, size, B, dev)
invK = inp["inv_intrinsics"].to(dev)
for dtype in (torch.float32, torch.bfloat16):
    for it in range(2):
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        scene = ops.SceneBuffers(scene0.pose, scene0.cscale, scene0.canon_pose, tp, dec, plane_dtype=dtype)
        out = rendering.render_rays(scene, img, invK, Nc, Nf, rng="philox", seed=it, tri_plane=tp, decoder=dec)
        e[1].record()
        loss = (out["color"] ** 2).mean() + ((1 - out["mask"]) ** 2).mean()
        loss.backward()
        e[2].record()
        torch.cuda.synchronize()
    R = B * size * size
    print("planes %s: fwd %.1f ms, bwd %.1f ms, %.2f M rays/s fwd+bwd, mask mean %.4f, |d_tp| max %.3e, peak mem %.1f GB" % (
        str(dtype).split(".")[-1], e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), R / e[0].elapsed_time(e[2]) / 1e3,
        out["mask"].mean().item(), tp.grad.abs().max().item(), torch.cuda.max_memory_allocated() / 2**30))
    tp.grad = None
