"""The reference-facing Python API on the GPU: generator mirror vs the reference goldens, the
``triplane_sampler`` autograd Function vs F.grid_sample, render_entire_img chunk invariance."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import golden_util as G
from enarf_gan_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


class _Planes(torch.nn.Module):
    def __init__(self, tp):
        super().__init__()
        self.tp = torch.nn.Parameter(tp)

    def forward(self, z, bone_length, truncation_psi=1):
        return self.tp[:z.shape[0]]


def make_generator(cfg, tp, g, constant=False, render_bs=16384):
    from enarf_gan_b200.generator import TriNARFGenerator
    nerf_params = dict(Nc=cfg["Nc"], Nf=cfg["Nf"], origin_location=cfg["origin"], coordinate_scale=3,
                       constant_triplane=constant, render_bs=render_bs)
    config = dict(z_dim=256, background_ratio=0.7, nerf_params=nerf_params)
    gen = TriNARFGenerator(config, cfg["size"], num_bone=24, num_bone_param=23, parent_id=S.SMPL_PARENTS,
                           black_background=True, tri_plane_gen=None if constant else _Planes(tp))
    gen.register_canonical_pose(S.canonical_pose())
    sd = {}
    for i in range(3):
        sd[f"nerf.mlp.layers.{i}.conv.weight"] = torch.from_numpy(g[f"mlp{i}_weight"])[None, :, :, None]
        sd[f"nerf.mlp.layers.{i}.conv.modulation.weight"] = torch.from_numpy(g[f"mlp{i}_mod_weight"])
        sd[f"nerf.mlp.layers.{i}.conv.modulation.bias"] = torch.from_numpy(g[f"mlp{i}_mod_bias"])
        sd[f"nerf.mlp.layers.{i}.bias"] = torch.from_numpy(g[f"mlp{i}_bias"])[None, :, None]
    missing, unexpected = gen.load_state_dict(sd, strict=False)
    assert not unexpected
    return gen.cuda()


@pytest.mark.parametrize("name", ["b1_s32", "b2_s24", "b1_s32_center_fixed", "config1_b1_s64"])
def test_generator_matches_reference_golden(name):
    """TriNARFGenerator.forward with the reference's parameters, inputs and RNG draws -> the reference's
    image / mask / disparity (1e-4 for B == 1; the reference's own batched-gather error bar for B > 1)."""
    g, cfg, inp, tp, layers, z_rend = G.load(name)
    gen = make_generator(cfg, tp, g).eval()
    gen.nerf.rng_mode = "given"
    gen.nerf.rng_given = (torch.from_numpy(g["idx"].astype(np.int64)).cuda(), torch.from_numpy(g["jitter"]).cuda())
    z = S.make_latent(cfg["B"], seed=cfg["seed"]).cuda()
    c = lambda k: inp[k].cuda()
    with torch.no_grad():
        color, mask, disp = gen(c("pose_to_camera"), c("pose_to_world"), c("bone_length"), z, c("inv_intrinsics"),
                                return_disparity=True)
    tol = 1e-4 if cfg["B"] == 1 else 3e-3
    assert rel(color, g["color"]) < tol
    assert rel(mask, g["mask"]) < tol
    assert rel(disp, g["disparity"]) < tol
    fw = gen.nerf.buffers_tensors["fine_weights"]
    assert fw.shape == (cfg["B"], 1, int(g["n_rays"]), cfg["Nf"] - 1)
    if "fine_weights" in g:
        assert rel(fw[:, 0], g["fine_weights"]) < tol
        assert rel(gen.nerf.buffers_tensors["fine_depth"][:, 0], g["fine_depth"]) < 1e-5


def test_generator_training_step_gradients_flow():
    """forward + backward through the public API: gradients reach the tri-plane producer and every decoder
    parameter (modulation, weight, bias)."""
    g, cfg, inp, tp, layers, z_rend = G.load("b2_s24")
    gen = make_generator(cfg, tp, g).train()
    gen.nerf.rng_mode = "philox"
    z = S.make_latent(cfg["B"], seed=3).cuda()
    c = lambda k: inp[k].cuda()
    color, mask, fw, fd = gen(c("pose_to_camera"), c("pose_to_world"), c("bone_length"), z, c("inv_intrinsics"))
    assert color.shape == (2, 3, 24, 24) and mask.shape == (2, 24, 24)
    loss = (color ** 2).mean() + ((1 - mask) ** 2).mean()
    loss.backward()
    assert gen.nerf.tri_plane_gen.tp.grad.abs().sum() > 0
    for name, p in gen.nerf.mlp.named_parameters():
        if "noise" in name:
            continue
        assert p.grad is not None and torch.isfinite(p.grad).all() and p.grad.abs().sum() > 0, name


def test_render_entire_img_is_chunk_invariant():
    """ray_batchsize (render_bs) chunking must not change the image (philox draws are keyed per call, so use
    the coarse-deterministic quantities: validity / mask support) and must match the one-shot render."""
    g, cfg, inp, tp, layers, z_rend = G.load("b1_s32")
    outs = []
    for bs in (1024, 256, 100):
        gen = make_generator(cfg, tp, g, render_bs=bs).eval()
        gen.nerf.rng_mode = "philox"
        gen.nerf.rng_seed = 11
        z = S.make_latent(1, seed=cfg["seed"]).cuda()
        z_nerf, z_rend_ = z[:, :512], z[:, 512:]
        c = lambda k: inp[k].cuda()
        color, mask, disp = gen.nerf.render_entire_img(c("pose_to_camera"), c("inv_intrinsics"), z_nerf, z_rend_,
                                                       c("bone_length"), render_size=32, Nc=48, Nf=48)
        assert color.shape == (3, 32, 32) and mask.shape == (32, 32)
        outs.append((color, mask))
    # different chunking -> different random draws, same scene: images agree to Monte-Carlo noise
    for color, mask in outs[1:]:
        assert (mask - outs[0][1]).abs().mean() < 0.02
        assert (color - outs[0][0]).abs().mean() < 0.02


@pytest.mark.parametrize("mode,padding,align", [("bilinear", "zeros", False), ("bilinear", "border", False),
                                                ("bilinear", "reflection", True), ("bilinear", "zeros", True),
                                                ("nearest", "zeros", False)])
@pytest.mark.parametrize("C", [5, 32])      # 32 = channels-last fast path (bilinear), 5 = generic NCHW kernels
def test_triplane_sampler_matches_grid_sample(mode, padding, align, C):
    """The stand-alone op (forward and both gradients) against three F.grid_sample calls."""
    from enarf_gan_b200.triplane_sampler import triplane_sampler
    torch.manual_seed(0)
    N, C, H, W, n = 2, C, 17, 23, 1000
    inp = torch.randn(N, 3 * C, H, W, device="cuda", requires_grad=True)
    grid = (torch.rand(N, n, 1, 3, device="cuda") * 2.6 - 1.3).requires_grad_()
    out = triplane_sampler(inp, grid, mode, padding, align)
    ref = 0
    for k in range(3):
        g2 = torch.stack([grid[..., k], grid[..., (k + 1) % 3]], dim=-1)
        ref = ref + F.grid_sample(inp[:, k * C:(k + 1) * C], g2, mode=mode, padding_mode=padding, align_corners=align)
    assert rel(out, ref) < 1e-5
    w = torch.randn_like(ref)
    gi, gg = torch.autograd.grad((out * w).sum(), [inp, grid])
    ri, rg = torch.autograd.grad((ref * w).sum(), [inp, grid])
    assert rel(gi, ri) < 1e-5
    if mode == "bilinear":
        assert rel(gg, rg) < 1e-4


def test_sampling_front_ends_match_oracle():
    """sample_triplane_part_prob / sample_weighted_feature_v2 (reference signatures) against the oracle."""
    import enarf_oracle as O
    from enarf_gan_b200 import sampling
    torch.manual_seed(1)
    B, P, n = 2, 3, 500
    tp = torch.randn(B, (32 + P) * 3, 64, 64)
    canon = torch.rand(B, P, 3, n) * 2.2 - 1.1
    valid = (canon.abs() < 1).all(dim=2)
    feat, weight = O.gather_features(tp, canon, valid)
    masked = canon * valid[:, :, None] + 2 * ~valid[:, :, None]
    w = sampling.sample_triplane_part_prob(tp[:, 96:].reshape(B * P, 3, 64, 64).cuda(), masked.cuda(), valid.cuda())
    assert rel(w.cpu() * valid, weight) < 1e-5
    f = sampling.sample_weighted_feature_v2(32, tp[:, :96].cuda(), masked.cuda(), w, valid.cuda())
    assert rel(f, feat) < 1e-5


def test_density_grid_matches_oracle_query():
    """The volume create_mesh feeds to marching cubes (mesh_rendering.py:50-71): grid layout, centre offset, coordinate
    scale and chunking, against the oracle's point query on the same grid."""
    import enarf_oracle as O
    g, cfg, inp, tp, layers, z_rend = G.load("b1_s32")
    gen = make_generator(cfg, tp, g).eval()
    nerf = gen.nerf
    sc = G.scene(cfg, inp, tp, layers, z_rend)
    c = lambda k: inp[k].cuda()
    pose, bl = nerf.transform_pose(c("pose_to_camera"), c("bone_length"))
    z = S.make_latent(1, seed=cfg["seed"]).cuda()
    mi = {"z": z[:, :512], "z_rend": z[:, 512:], "bone_length": bl, "tri_plane_feature": tp.cuda(), "truncation_psi": 1}
    center = pose[:, 0, :3, 3:].clone()                              # (1,3,1): around the first part
    before = pose.clone()
    vol = nerf.density_grid(pose, center, voxel_size=0.1, model_input=mi, chunk=3000)
    assert torch.equal(pose, before), "the caller's pose must not be scaled in place"
    assert vol.shape == (21, 21, 21)
    bins = torch.arange(-10, 11) / 10
    pts = (torch.stack(torch.meshgrid(bins, bins, bins, indexing="ij")).reshape(1, 3, -1) + center.cpu()) * 3.0
    with torch.no_grad():
        den = sc.query(pts)[0]
    assert den.abs().max() > 0
    assert rel(vol.reshape(-1), den.reshape(-1)) < 1e-4


def test_render_entire_img_ray_blocks_assemble_the_image():
    """B < G sharding: the ray blocks ``shard_rays`` hands to 3 ranks, rendered one after the other here,
    concatenate to the one-shot image (same scene; fine-pass draws differ per call -> Monte-Carlo tolerance; rays that
    miss every part box are exactly zero in both)."""
    from enarf_gan_b200 import rendering
    g, cfg, inp, tp, layers, z_rend = G.load("b1_s32")
    gen = make_generator(cfg, tp, g, render_bs=4096).eval()
    nerf = gen.nerf
    nerf.rng_mode = "philox"
    z = S.make_latent(1, seed=cfg["seed"]).cuda()
    c = lambda k: inp[k].cuda()
    pose, bl = nerf.transform_pose(c("pose_to_camera"), c("bone_length"))
    mi = {"z": z[:, :512], "z_rend": z[:, 512:], "bone_length": bl, "tri_plane_feature": tp.cuda(), "truncation_psi": 1}
    full_c, full_m, _ = rendering.render_entire_img(nerf, pose, c("inv_intrinsics"), render_size=32, Nc=48, Nf=48,
                                                    model_input=mi)
    n, world = 32 * 32, 3
    blocks = [rendering.render_entire_img(nerf, pose, c("inv_intrinsics"), render_size=32, Nc=48, Nf=48, model_input=mi,
                                          ray_range=rendering.shard_rays(n, r, world)) for r in range(world)]
    col = torch.cat([b[0] for b in blocks], dim=-1).reshape(3, 32, 32)
    msk = torch.cat([b[1] for b in blocks], dim=-1).reshape(32, 32)
    from enarf_gan_b200 import ops
    img = rendering.pixel_grid(32, 32, 1, pose.device)
    miss = ~ops.depth_range(img, c("inv_intrinsics"), rendering._scaled(nerf, pose))[3].reshape(32, 32)
    assert miss.any() and (msk[miss] == 0).all() and (full_m[miss] == 0).all()
    assert (msk - full_m).abs().mean() < 0.02 and (col - full_c).abs().mean() < 0.02
    # single process: the sharded entry point is the plain one
    one_c, one_m, _ = rendering.render_entire_img_sharded(nerf, pose, c("inv_intrinsics"), render_size=32, Nc=48, Nf=48,
                                                          model_input=mi)
    assert one_c.shape == (3, 32, 32) and (one_m[miss] == 0).all() and (one_m - full_m).abs().mean() < 0.02
