"""Host side of the DSO mirror on CPU (no kernels involved): ray selection against the reference's own draw stored in
the fixtures, the time / pose latents, the photometric loss, and the constructor's parameter layout."""
import numpy as np
import pytest
import torch

import golden_util as G
from enarf_gan_b200 import synthetic as S
from enarf_gan_b200.generator import DSONARFGenerator, PhotometricLoss, mask_based_sampler


@pytest.mark.parametrize("name", G.DSO_CASES)
def test_mask_based_sampler_reproduces_the_reference_draw(name):
    """Same generator state as the reference run (make_golden_dso.py seeds seed+1 right before the forward, and the
    sampler's uniform_ is the first draw): the separable dilation + topk must select the reference's ray ids."""
    g, cfg, *_ = G.load_dso(name)
    fg = S.make_foreground_mask(cfg["B"], cfg["size"], seed=cfg["seed"])
    torch.manual_seed(cfg["seed"] + 1)
    idx, homo = mask_based_sampler(fg, cfg["ray_bs"])
    assert sorted(idx[0].tolist()) == sorted(g["ray_idx"][0].tolist())
    assert torch.equal(idx, torch.from_numpy(g["ray_idx"]).long())
    assert torch.equal(homo[:, 0], G.dso_image_coord(g["ray_idx"], cfg["size"]))


def _generator(Nf=48):
    nerf_params = dict(Nc=48, Nf=Nf, origin_location="center_fixed", coordinate_scale=3, deformation_field=True,
                       no_ray_direction=True, time_conditional=True, pose_conditional=True, render_bs=16384)
    config = dict(use_triplane=True, ray_batchsize=64, nerf_params=nerf_params)
    gen = DSONARFGenerator(config, 32, num_bone=24, parent_id=S.SMPL_PARENTS, num_bone_param=23)
    gen.register_canonical_pose(S.canonical_pose())
    return gen


def test_latents_and_parameter_layout():
    gen = _generator()
    inp = S.make_inputs(2, 32, seed=3)
    t = torch.tensor([0.25, 0.8])
    z1, z2 = gen.get_latents(t, inp["pose_to_camera"])
    assert z1 is z2 and z1.shape == (2, 20 + 23 * 9)
    assert torch.allclose(z1, G.dso_latent(t.numpy(), inp["pose_to_camera"]), atol=1e-6)
    # the reference's parameter names / shapes (models/narf.py:38, libraries/NeRF/net.py:10-27) so its checkpoints load
    sd = gen.state_dict()
    assert sd["nerf.tri_plane"].shape == (1, 32 * 3 + 23 * 3, 256, 256)
    assert sd["nerf.mlp.layers.0.conv.modulation.weight"].shape == (32, 227)
    assert sd["nerf.mlp.layers.2.conv.weight"].shape == (1, 4, 64, 1)
    assert gen.nerf.canonical_pose.shape == (23, 4, 4)


def test_unsupported_switches_fail_loudly():
    with pytest.raises(NotImplementedError):
        DSONARFGenerator(dict(use_triplane=False, ray_batchsize=64, nerf_params={}), 32, num_bone=24,
                         parent_id=S.SMPL_PARENTS)
    nerf_params = dict(Nc=48, Nf=48, origin_location="center_fixed", coordinate_scale=3, no_ray_direction=False,
                       time_conditional=True, pose_conditional=False, render_bs=16384)
    with pytest.raises(NotImplementedError):          # view-dependent colour is outside the fused path
        DSONARFGenerator(dict(use_triplane=True, ray_batchsize=64, nerf_params=nerf_params), 32, num_bone=24,
                         parent_id=S.SMPL_PARENTS)


@pytest.mark.parametrize("kind", ["mse", "mae"])
def test_photometric_loss_restates_the_reference(kind):
    """libraries/NeRF/loss.py:14-48 written out with explicit indexing."""
    torch.manual_seed(0)
    B, size, n = 2, 8, 20
    color, mask = torch.rand(B, 3, size, size), torch.rand(B, size, size)
    grid = torch.stack([torch.randperm(size * size)[:n] for _ in range(B)])
    sparse_c, sparse_m = torch.rand(B, 3, n), torch.rand(B, n)
    lc, lm = PhotometricLoss(dict(nerf_loss_type=kind, color_coef=2.0, mask_coef=0.5))(grid, sparse_c, sparse_m, color, mask)
    ref_c = torch.stack([color[b].reshape(3, -1)[:, grid[b]] for b in range(B)])
    ref_m = torch.stack([mask[b].reshape(-1)[grid[b]] for b in range(B)])
    want_c = ((ref_c - sparse_c) ** 2).mean() if kind == "mse" else (ref_c - sparse_c).abs().clamp_min(0.01).mean()
    assert torch.allclose(lc, want_c * 2.0) and torch.allclose(lm, ((ref_m - sparse_m) ** 2).mean() * 0.5)
