import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, ROOT + "/oracle", ROOT + "/tests"]
import numpy as np, torch
import enarf_oracle as O, golden_util as G
from enarf_gan_b200 import ops

def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()

name = "b1_s32"
g, cfg, inp, tp, layers, z_rend = G.load(name)
B, Nc, Nf = cfg["B"], cfg["Nc"], cfg["Nf"]
img = O.whole_image_grid(cfg["size"], cfg["size"], B)
rng = {"mode": "given", "idx": torch.from_numpy(g["idx"].astype(np.int64)), "jitter": torch.from_numpy(g["jitter"])}
for dtype in (torch.float64, torch.float32):
  for term in ("color", "mask", "disparity", "fine_weights"):
    sc = G.scene(cfg, inp, tp, layers, z_rend, dtype)
    sc.tri_plane = sc.tri_plane.clone().requires_grad_()
    sc.eff = [(w.detach().clone().requires_grad_(), b.detach().clone().requires_grad_()) for w, b in sc.eff]
    ref = O.render(sc, img, Nc, Nf, rng)
    sel = ref["validity"][0]
    n = ref["fine_weights"].shape[1]
    gen = torch.Generator().manual_seed(3)
    pick = {"color": ref["color"][:, :, sel], "mask": ref["mask"][:, sel], "disparity": ref["disparity"][:, sel],
            "fine_weights": ref["fine_weights"]}[term]
    wgt = torch.randn(pick.shape, generator=gen)
    (pick * wgt).sum().backward()
    d = lambda t: t.detach().float().cuda()
    tp_g = d(sc.tri_plane).requires_grad_()
    dec = [(d(w).requires_grad_(), d(b).requires_grad_()) for w, b in sc.eff]
    gs = ops.SceneBuffers(d(sc.pose), d(sc.cscale), d(sc.canonical_pose), tp_g, dec)
    out = ops.render_fine(gs, d(ref["depth_min"]), d(ref["depth_max"]), d(ref["direction_all"][:, :, sel]).contiguous(),
                          d(ref["bins"]), tri_plane=tp_g, decoder=dec)
    o = {"color": out[0], "mask": out[1], "disparity": out[2], "fine_weights": out[3]}[term]
    (o * wgt.cuda()).sum().backward()
    gt = sc.tri_plane.grad
    msg = [f"{dtype} {term:13s} feat {rel(tp_g.grad[:, :96], gt[:, :96]):.2e} prob {rel(tp_g.grad[:, 96:], gt[:, 96:]):.2e}"]
    for i, ((w, b), (wr, br)) in enumerate(zip(dec, sc.eff)):
        msg.append(f"dW{i+1} {rel(w.grad, wr.grad):.2e} db{i+1} {rel(b.grad, br.grad):.2e}")
    print(" ".join(msg))
    if term == "color":
        diff = (tp_g.grad.cpu().double() - gt.double()).abs()
        k = diff.flatten().topk(5)
        for v, i in zip(k.values, k.indices):
            idx = np.unravel_index(i.item(), diff.shape)
            print("   worst", idx, "got", tp_g.grad.cpu()[idx].item(), "want", gt[idx].item())
