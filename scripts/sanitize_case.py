"""A small pass over every kernel of the library for compute-sanitizer (memcheck / racecheck / initcheck / synccheck):
24 x 24 images, 2 samples -- fp32 and bf16 planes, with and without the feature / pair checkpoints, early ray
termination on an opaque scene (rounds), point query, stand-alone samplers, DSO producer, ray selection, photometric
loss.   compute-sanitizer --tool racecheck python scripts/sanitize_case.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import enarf_oracle as O
import golden_util as G
from enarf_gan_b200 import ops, rendering, triplane_sampler as TS

g, cfg, inp, tp, layers, z_rend = G.load("b2_s24")
sc = G.scene(cfg, inp, tp, layers, z_rend)
d = lambda t: t.detach().float().cuda()
img = O.whole_image_grid(cfg["size"], cfg["size"], cfg["B"]).cuda()


def scene(dtype=torch.float32, opaque=False):
    t = sc.tri_plane.clone()
    dec = [(w.detach().clone(), b.detach().clone()) for w, b in sc.eff]
    if opaque:
        t[:, 96:] += 6.0
        dec[2][1][3] += 30.0
    tp_g = d(t).requires_grad_()
    dec_g = [(d(w).requires_grad_(), d(b).requires_grad_()) for w, b in dec]
    return ops.SceneBuffers(d(sc.pose), d(sc.cscale), d(sc.canonical_pose), tp_g, dec_g, dtype), tp_g, dec_g


def fwd_bwd(dtype, feat, pairs, ert=0.0, opaque=False):
    ops.SAVE_FEATURES, ops.PAIR_CKPT = feat, pairs
    gs, tp_g, dec_g = scene(dtype, opaque)
    out = rendering.render_rays(gs, img, sc.inv_intrinsics.cuda(), cfg["Nc"], cfg["Nf"], rng="philox", seed=1, tri_plane=tp_g,
                                decoder=dec_g, ert_eps=ert, bg_const=-1.0)
    if ert:      # second call: the coarse pass runs in rounds on the first call's hint
        out = rendering.render_rays(gs, img, sc.inv_intrinsics.cuda(), cfg["Nc"], cfg["Nf"], rng="philox", seed=2, tri_plane=tp_g,
                                    decoder=dec_g, ert_eps=ert, ert_prev=out["opaque_rays"])
    (out["color"].square().sum() + out["mask"].sum() + out["disparity"].sum() + out["fine_weights"].sum()).backward()
    torch.cuda.synchronize()
    return out


for dtype in (torch.float32, torch.bfloat16):
    for feat, pairs in ((True, True), (True, False), (False, False)):
        fwd_bwd(dtype, feat, pairs)
fwd_bwd(torch.float32, True, True, ert=1e-6, opaque=True)
fwd_bwd(torch.float32, False, False, ert=1e-6, opaque=True)
print("render fwd/bwd ok")
gs, _, _ = scene()
pts = torch.randn(2, 3, 3000, device="cuda") * 3
ops.density_query(gs, pts, want_color=True, want_mask=True)
# stand-alone sampler (generic + channels-last fast path), forward and backward
for C in (8, 32):
    x = torch.randn(1, 3 * C, 64, 64, device="cuda", requires_grad=True)
    grid = (torch.rand(1, 500, 1, 3, device="cuda") * 2.4 - 1.2).requires_grad_()
    TS.triplane_sampler(x, grid).sum().backward()
# DSO producer, ray selection, photometric loss
canon = torch.randn(3, 64, 64, 32, device="cuda", requires_grad=True)
flow = (torch.randn(2, 6, 64, 64, device="cuda") * 3).requires_grad_()
for dt in (torch.float32, torch.bfloat16):
    ops.warp_planes(canon, flow, dt).float().sum().backward()
mask = (torch.rand(2, 80, 72, device="cuda") > 0.9).float()
idx, homo = ops.select_rays(mask, 777, seed=3, radius=5)
rc = torch.randn(2, 3, 777, device="cuda", requires_grad=True)
rm = torch.rand(2, 777, device="cuda", requires_grad=True)
lc, lm = ops.photometric_loss(idx, rc, rm, torch.randn(2, 3, 80, 72, device="cuda"), torch.rand(2, 80, 72, device="cuda"), "mae")
(lc + lm).backward()
torch.cuda.synchronize()
print("sanitize_case: all kernels ran")
