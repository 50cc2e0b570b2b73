"""DSO ("denarf") path on the GPU, through the C ABI: the deformation-field tri-plane producer
(``enarf_warp_planes_fwd/bwd``, reference models/narf.py:39-58) against the oracle and the reference fixtures, the
plane hand-off into the fused fine pass (channels-last planes in, channels-last gradient out, probability planes
shared by the batch), and the ``DSONARFGenerator`` mirror against the reference's own output."""
import numpy as np
import pytest
import torch

import enarf_oracle as O
import golden_util as G
from enarf_gan_b200 import synthetic as S

pytestmark = pytest.mark.gpu
TOL = 1e-4


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def to_nchw(feat):
    """(B,3,H,W,32) channels-last -> (B,96,H,W), the reference's channel order."""
    B, _, H, W, C = feat.shape
    return feat.permute(0, 1, 4, 2, 3).reshape(B, 3 * C, H, W)


def canonical_cl(canon):
    from enarf_gan_b200 import ops
    return ops.repack_planes(canon.cuda())[0]


@pytest.mark.parametrize("name", G.DSO_CASES)
def test_warp_forward_vs_oracle_and_reference(name):
    from enarf_gan_b200 import ops
    g, cfg, inp, canon, flow, layers = G.load_dso(name)
    feat = ops.warp_planes(canonical_cl(canon), flow.cuda())
    got = torch.cat([to_nchw(feat).cpu(), canon[:, 96:].expand(cfg["B"], -1, -1, -1)], dim=1)
    assert rel(got.reshape(-1)[::G.DSO_SUBSAMPLE], g["warped_sub"]) < 1e-5          # the reference's own closure
    with torch.no_grad():
        ref = O.warp_triplane(canon, flow)
    assert rel(got, ref) < 1e-5
    # bf16 planes (config 4): 1e-2
    feat16 = ops.warp_planes(canonical_cl(canon), flow.cuda(), torch.bfloat16)
    assert feat16.dtype == torch.bfloat16
    assert rel(to_nchw(feat16.float()).cpu(), ref[:, :96]) < 1e-2


def test_warp_far_outside_and_zero_flow():
    """Zero flow is the identity; a flow that leaves the plane by more than a texel gives exact zeros (zero padding)."""
    from enarf_gan_b200 import ops
    canon = S.make_triplane(1, num_part=23, seed=7)
    cl = canonical_cl(canon)
    zero = torch.zeros(1, 6, 256, 256, device="cuda")
    assert rel(to_nchw(ops.warp_planes(cl, zero)).cpu(), canon[:, :96]) < 1e-6
    far = torch.full((1, 6, 256, 256), 1.0e6, device="cuda")
    far[:, 1::2] = -400.0
    assert ops.warp_planes(cl, far).abs().max().item() == 0.0


@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
def test_warp_backward_vs_oracle_autograd(gdtype):
    """d(canonical planes) and d(flow) against autograd through the oracle's elementwise restatement."""
    from enarf_gan_b200 import ops
    B = 2
    canon = S.make_triplane(1, num_part=23, seed=11, smooth=1)
    flow = S.make_flow(B, seed=11)
    up = torch.randn(B, 96, 256, 256, generator=torch.Generator().manual_seed(3))
    c_ref = canon.clone().requires_grad_()
    f_ref = flow.clone().requires_grad_()
    (O.warp_triplane(c_ref, f_ref)[:, :96] * up).sum().backward()

    cl = canonical_cl(canon).requires_grad_()
    f_gpu = flow.cuda().requires_grad_()
    out = ops.warp_planes(cl, f_gpu, gdtype)
    up_cl = up.reshape(B, 3, 32, 256, 256).permute(0, 1, 3, 4, 2).contiguous().cuda()
    (out.float() * up_cl).sum().backward() if gdtype == torch.float32 else out.backward(up_cl.to(gdtype))
    tol = TOL if gdtype == torch.float32 else 1e-2
    assert rel(to_nchw(cl.grad[None]).cpu(), c_ref.grad[:, :96]) < tol, "d canonical planes"
    assert rel(f_gpu.grad, f_ref.grad) < tol, "d flow"


def test_repack_canonical_gradient_is_the_transpose():
    from enarf_gan_b200 import ops
    tp = S.make_triplane(1, num_part=23, seed=2).cuda().requires_grad_()
    up = torch.randn(3, 256, 256, 32, device="cuda")
    (ops.RepackCanonical.apply(tp) * up).sum().backward()
    assert rel(tp.grad[:, :96], to_nchw(up[None])) == 0.0
    assert tp.grad[:, 96:].abs().max().item() == 0.0


def _dso_generator(cfg, canon, flow, g, plane_dtype=torch.float32):
    from enarf_gan_b200.generator import DSONARFGenerator

    class Flow(torch.nn.Module):
        def __init__(self, f):
            super().__init__()
            self.f = torch.nn.Parameter(f)

        def forward(self, z, bone_length, truncation_psi=1):
            return self.f[:z.shape[0]]

    nerf_params = dict(Nc=cfg["Nc"], Nf=cfg["Nf"], origin_location="center_fixed", coordinate_scale=3,
                       deformation_field=True, no_ray_direction=True, time_conditional=True, pose_conditional=True,
                       render_bs=16384)
    config = dict(use_triplane=True, ray_batchsize=cfg["ray_bs"], nerf_params=nerf_params)
    gen = DSONARFGenerator(config, cfg["size"], num_bone=24, parent_id=S.SMPL_PARENTS, num_bone_param=23,
                           flow_generator=Flow(flow), plane_dtype=plane_dtype)
    gen.register_canonical_pose(S.canonical_pose())
    sd = {"nerf.tri_plane": canon}
    for i in range(3):
        sd[f"nerf.mlp.layers.{i}.conv.weight"] = torch.from_numpy(g[f"mlp{i}_weight"])[None, :, :, None]
        sd[f"nerf.mlp.layers.{i}.conv.modulation.weight"] = torch.from_numpy(g[f"mlp{i}_mod_weight"])
        sd[f"nerf.mlp.layers.{i}.conv.modulation.bias"] = torch.from_numpy(g[f"mlp{i}_mod_bias"])
        sd[f"nerf.mlp.layers.{i}.bias"] = torch.from_numpy(g[f"mlp{i}_bias"])[None, :, None]
    missing, unexpected = gen.load_state_dict(sd, strict=False)
    assert not unexpected
    return gen.cuda()


@pytest.mark.parametrize("name", G.DSO_CASES)
def test_dso_generator_matches_reference_golden(name):
    """DSONARFGenerator.forward with the reference's parameters, ray ids and fine-pass draws -> its colour / mask."""
    g, cfg, inp, canon, flow, layers = G.load_dso(name)
    gen = _dso_generator(cfg, canon, flow, g).eval()
    ray_idx = torch.from_numpy(g["ray_idx"]).long().cuda()
    gen.ray_sampler = lambda mask, ray_bs: (ray_idx, G.dso_image_coord(ray_idx.cpu(), cfg["size"])[:, None].cuda())
    gen.nerf.rng_mode = "given"
    gen.nerf.rng_given = (torch.from_numpy(g["idx"].astype(np.int64)).cuda(), torch.from_numpy(g["jitter"]).cuda())
    c = lambda k: inp[k].cuda()
    fg = S.make_foreground_mask(cfg["B"], cfg["size"], seed=cfg["seed"]).cuda()
    with torch.no_grad():
        color, mask, grid = gen(c("pose_to_camera"), None, fg, torch.from_numpy(g["frame_time"]).cuda(),
                                c("bone_length"), c("inv_intrinsics"))
    tol = 1e-4 if cfg["B"] == 1 else 3e-3
    assert torch.equal(grid, ray_idx)
    assert rel(color, g["color"]) < tol
    assert rel(mask, g["mask"]) < tol
    assert rel(gen.nerf.buffers_tensors["fine_weights"][:, 0], g["fine_weights"]) < tol


def test_mask_based_sampler_contract():
    """ray_sampler.py:7-39: ray_bs distinct pixels per image, all inside the dilated mask when it has enough pixels,
    pixel-centre homogeneous coordinates."""
    from enarf_gan_b200.generator import mask_based_sampler
    size, ray_bs = 256, 512
    fg = torch.zeros(2, size, size, device="cuda")
    fg[:, 10:20, 30:40] = 1                     # dilation by 64 -> rows [0, 84), columns [0, 104)
    idx, homo = mask_based_sampler(fg, ray_bs)
    assert idx.shape == (2, ray_bs) and homo.shape == (2, 1, 3, ray_bs)
    x, y = idx % size, idx // size
    assert (x < 104).all() and (y < 84).all()
    assert all(len(set(r.tolist())) == ray_bs for r in idx)
    assert torch.equal(homo[:, 0, 0], x.float() + 0.5) and torch.equal(homo[:, 0, 1], y.float() + 0.5)
    assert (homo[:, 0, 2] == 1).all()


@pytest.mark.parametrize("name,plane_dtype,tol", [("dso_b2_s48", torch.float32, TOL), ("dso_b1_s64", torch.float32, TOL),
                                                  ("dso_b2_s48", torch.bfloat16, 1e-2)])
def test_dso_step_gradients(name, plane_dtype, tol):
    """One DSO step (warp -> sparse render -> loss -> backward) through the mirror: gradients of the canonical
    tri-plane parameter (features through the warp, probabilities summed over the batch), of the flow and of the
    decoder parameters against autograd through the oracle."""
    g, cfg, inp, canon, flow, layers = G.load_dso(name)
    B, Nc, Nf = cfg["B"], cfg["Nc"], cfg["Nf"]
    z = G.dso_latent(g["frame_time"], inp["pose_to_camera"])
    rng = {"mode": "given", "idx": torch.from_numpy(g["idx"].astype(np.int64)), "jitter": torch.from_numpy(g["jitter"])}
    img = G.dso_image_coord(g["ray_idx"], cfg["size"])
    # oracle
    for L in layers:
        for k in ("weight", "mod_weight"):
            L[k].requires_grad_()
    c_ref = canon.clone().requires_grad_()
    f_ref = flow.clone().requires_grad_()
    warped = O.warp_triplane(c_ref, f_ref)
    if plane_dtype == torch.bfloat16:
        # bf16 planes move the point the gradient is taken at (activation kinks flip: a few % of the max-norm), which
        # says nothing about the kernels: the oracle renders the same bf16-rounded feature planes (straight-through),
        # so what is compared is the backward arithmetic incl. the bf16 rounding of the plane gradient
        feat = warped[:, :96]
        warped = torch.cat([feat + (feat.bfloat16().float() - feat).detach(), warped[:, 96:]], dim=1)
    sc = O.Scene(inp["pose_to_camera"], inp["bone_length"], inp["inv_intrinsics"], warped,
                 layers, z, S.canonical_pose(), S.SMPL_PARENTS, origin_location="center_fixed", coordinate_scale=3.0)
    ref = O.render(sc, img, Nc, Nf, rng)
    up = torch.randn(ref["color"].shape, generator=torch.Generator().manual_seed(9))
    loss = (ref["color"] * up).sum() + ref["mask"].square().sum()
    loss.backward()
    # mirror
    gen = _dso_generator(cfg, canon, flow, g, plane_dtype).train()
    ray_idx = torch.from_numpy(g["ray_idx"]).long().cuda()
    gen.ray_sampler = lambda mask, ray_bs: (ray_idx, img[:, None].cuda())
    gen.nerf.rng_mode = "given"
    gen.nerf.rng_given = (rng["idx"].cuda(), rng["jitter"].cuda())
    c = lambda k: inp[k].cuda()
    fg = S.make_foreground_mask(B, cfg["size"], seed=cfg["seed"]).cuda()
    color, mask, _ = gen(c("pose_to_camera"), None, fg, torch.from_numpy(g["frame_time"]).cuda(), c("bone_length"),
                         c("inv_intrinsics"), background=0.0)
    loss_g = (color * up.cuda()).sum() + mask.square().sum()
    loss_g.backward()
    assert rel(loss_g, loss) < tol
    gt = gen.nerf.tri_plane.grad
    assert rel(gt[:, :96], c_ref.grad[:, :96]) < tol, "canonical feature planes"
    assert rel(gt[:, 96:], c_ref.grad[:, 96:]) < tol, "probability planes (summed over the batch)"
    assert rel(gen.nerf.flow_generator.f.grad, f_ref.grad) < tol, "flow"
    for i, L in enumerate(layers):
        conv = gen.nerf.mlp.layers[i].conv
        assert rel(conv.weight.grad[0, :, :, 0], L["weight"].grad) < max(tol, 1e-3), f"decoder weight {i}"
        assert rel(conv.modulation.weight.grad, L["mod_weight"].grad) < max(tol, 1e-3), f"modulation weight {i}"


def test_dso_render_entire_img_agrees_with_sparse_forward():
    """DSONARFGenerator.render_entire_img (chunked, batch forced to 1, planes warped once) against forward() on all
    pixels of the same frame: same scene, independent fine-pass draws -> Monte-Carlo tolerance; rays that miss every part
    are exactly zero / background in both."""
    g, cfg, inp, canon, flow, layers = G.load_dso("dso_b1_s64")
    size = cfg["size"]
    gen = _dso_generator(dict(cfg, ray_bs=size * size), canon, flow, g).eval()
    gen.nerf.rng_mode = "philox"
    gen.nerf.config["render_bs"] = 1000                               # 5 chunks, the last one ragged
    c = lambda k: inp[k].cuda()
    t = torch.from_numpy(g["frame_time"]).cuda()
    ray_idx = torch.arange(size * size, device="cuda")[None]
    gen.ray_sampler = lambda mask, ray_bs: (ray_idx, G.dso_image_coord(ray_idx.cpu(), size)[:, None].cuda())
    with torch.no_grad():
        color, mask, _ = gen(c("pose_to_camera"), None, None, t, c("bone_length"), c("inv_intrinsics"), background=0.0)
        img_c, img_m, img_d = gen.render_entire_img(c("pose_to_camera"), c("inv_intrinsics"), t, c("bone_length"),
                                                    render_size=size)
    assert img_c.shape == (3, size, size) and img_m.shape == (size, size) and img_d.shape == (size, size)
    m = mask.reshape(size, size)
    assert img_m.max() > 0.5
    assert (img_m - m).abs().mean() < 0.02 and (img_c - color.reshape(3, size, size)).abs().mean() < 0.02
    assert not gen.nerf.tri_plane.requires_grad or img_c.grad_fn is None        # no_grad=True: nothing tracked
