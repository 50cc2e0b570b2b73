"""Edge cases of the path on the GPU, through the C ABI: empty and ragged ray sets, rays that miss everything, the
per-ray sample limits, and the error behaviour of the boundary (codes + messages instead of CUDA faults)."""
import numpy as np
import pytest
import torch

import enarf_oracle as O
import golden_util as G

pytestmark = pytest.mark.gpu
TOL = 1e-4


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def gpu_scene(sc, plane_dtype=torch.float32):
    from enarf_gan_b200 import ops
    d = lambda t: t.cuda()
    return ops.SceneBuffers(d(sc.pose), d(sc.cscale), d(sc.canonical_pose), d(sc.tri_plane),
                            [(d(w), d(b)) for w, b in sc.eff], plane_dtype)


@pytest.fixture(scope="module")
def b2():
    g, cfg, inp, tp, layers, z_rend = G.load("b2_s24")
    sc = G.scene(cfg, inp, tp, layers, z_rend)
    return dict(cfg=cfg, sc=sc, gs=gpu_scene(sc))


@pytest.mark.parametrize("n", [1, 7, 33, 250])
def test_ragged_ray_counts_match_the_oracle(b2, n):
    """Ray counts that are not a multiple of a tile (16 rays) or a warp, down to a single ray per sample: the in-kernel
    Philox draws are keyed by the launch ray index, so the oracle reproduces the bins bit-exactly and the render to 1e-4."""
    from enarf_gan_b200 import rendering
    cfg, sc = b2["cfg"], b2["sc"]
    B, s = cfg["B"], cfg["size"]
    full = O.whole_image_grid(s, s, B)
    pick = torch.linspace(0, s * s - 1, n).long()          # spread over the image: hits and misses
    img = full[:, :, pick].contiguous()
    seed = 991
    with torch.no_grad():
        ref = O.render(sc, img, cfg["Nc"], cfg["Nf"], {"mode": "philox", "seed": seed, "ray_ids": torch.arange(B * n)})
    out = rendering.render_rays(b2["gs"], img.cuda(), sc.inv_intrinsics.cuda(), cfg["Nc"], cfg["Nf"], rng="philox",
                                seed=seed, return_intermediate=True)
    assert torch.equal(out["validity"].cpu(), ref["validity"])
    assert torch.equal(out["bins"].cpu(), ref["bins"])
    for k in ("color", "mask", "disparity", "fine_weights"):
        assert rel(out[k], ref[k]) < TOL or ref[k].abs().max() == 0, k


def test_rays_that_miss_every_part(b2):
    """Pixels far outside the image: no probe falls inside a part box.  B > 1 keeps the rays (exact zeros out);
    B == 1 compacts them away and takes the reference's empty branch (rendering.py:351-354)."""
    from enarf_gan_b200 import ops, rendering
    cfg, sc = b2["cfg"], b2["sc"]
    B, n = cfg["B"], 40
    img = torch.stack([torch.full((n,), -5000.0), torch.linspace(-9000, -8000, n), torch.ones(n)])[None].repeat(B, 1, 1)
    out = rendering.render_rays(b2["gs"], img.cuda(), sc.inv_intrinsics.cuda(), cfg["Nc"], cfg["Nf"], rng="philox")
    assert not out["validity"].any()
    for k in ("color", "mask", "disparity", "fine_weights"):
        assert out[k].abs().max().item() == 0.0, k
    g1 = ops.SceneBuffers(b2["gs"].pose[:1], b2["gs"].cscale[:1], b2["gs"].canon_pose, b2["gs"].tri_plane[:1],
                          [(w[:1], b) for w, b in zip(b2["gs"].w, b2["gs"].b)])
    one = rendering.render_rays(g1, img[:1].cuda(), sc.inv_intrinsics[:1].cuda(), cfg["Nc"], cfg["Nf"], rng="philox")
    assert one["fine_weights"] is None and one["color"].shape == (1, 3, n) and one["color"].abs().max().item() == 0.0


def test_maximum_samples_per_ray(b2):
    """Nc = Nf = 512 (the documented limit of the fused passes): finite, sorted, and the compositing identities hold."""
    from enarf_gan_b200 import rendering
    cfg, sc = b2["cfg"], b2["sc"]
    s = cfg["size"]
    img = O.whole_image_grid(s, s, cfg["B"])[:, :, ::5].contiguous()
    out = rendering.render_rays(b2["gs"], img.cuda(), sc.inv_intrinsics.cuda(), 512, 512, rng="philox", seed=3,
                                return_intermediate=True)
    assert out["bins"].shape[-1] == 512 and (out["bins"][..., 1:] >= out["bins"][..., :-1]).all()
    for k in ("color", "mask", "disparity", "fine_weights"):
        assert torch.isfinite(out[k]).all(), k
    assert rel(out["fine_weights"].sum(-1), out["mask"]) < 1e-4
    assert out["mask"].max() <= 1 + 1e-5 and out["mask"].max() > 0.1


def test_limits_and_bad_arguments_are_reported_not_faulted(b2):
    from enarf_gan_b200 import ops, rendering
    cfg, sc, gs = b2["cfg"], b2["sc"], b2["gs"]
    B = cfg["B"]
    img = O.whole_image_grid(cfg["size"], cfg["size"], B).cuda()
    dmin, dmax, direction, _ = ops.depth_range(img, sc.inv_intrinsics.cuda(), gs.pose)
    with pytest.raises(RuntimeError, match="libenarf_b200 error"):       # 513 samples per ray: over the limit
        ops.render_coarse(gs, dmin, dmax, direction, 513)
    with pytest.raises(RuntimeError, match="libenarf_b200 error"):       # a single coarse sample: no interval
        ops.render_coarse(gs, dmin, dmax, direction, 1)
    with pytest.raises(ValueError, match="contiguous"):                  # strided tensors are refused, not copied silently
        ops.render_coarse(gs, dmin[:, ::2], dmax, direction, 48)
    with pytest.raises(ValueError, match="CUDA tensors"):                # host tensors: there is no CPU path
        ops.fine_sample(torch.rand(4, 48), 48)
    with pytest.raises(TypeError):                                       # fp64 is not part of the boundary
        ops.render_coarse(gs, dmin.double(), dmax, direction, 48)
    with pytest.raises(RuntimeError, match="multiple of 4"):             # warp producer: W % 4 != 0
        ops.warp_planes(torch.zeros(3, 8, 6, 32, device="cuda"), torch.zeros(1, 6, 8, 6, device="cuda"))
    # the device is still healthy afterwards
    wb = ops.render_coarse(gs, dmin, dmax, direction, cfg["Nc"])
    assert torch.isfinite(wb).all()


def test_more_than_32_parts_is_refused():
    from enarf_gan_b200 import ops
    B, P = 1, 33
    pose = torch.eye(4, device="cuda").repeat(B, P, 1, 1)
    with pytest.raises((RuntimeError, AssertionError)):
        gs = ops.SceneBuffers(pose, torch.ones(B, P, device="cuda"), pose[0], torch.zeros(B, (32 + P) * 3, 16, 16, device="cuda"),
                              [(torch.zeros(B, 64, 32, device="cuda"), torch.zeros(64, device="cuda")),
                               (torch.zeros(B, 64, 64, device="cuda"), torch.zeros(64, device="cuda")),
                               (torch.zeros(B, 4, 64, device="cuda"), torch.zeros(4, device="cuda"))])
        d = torch.ones(B, 8, device="cuda")
        ops.render_coarse(gs, d, d * 2, torch.ones(B, 3, 8, device="cuda"), 48)


def test_large_point_query_is_tiling_invariant_and_repeatable(b2):
    """Regression test for the tile_compact race fixed in round 2 (a warp that entered the compaction late could read
    an already advanced list length: corrupt active list -> wrong densities / a faulting texel offset, seen only at
    large point counts).  The point query over 2 x 2^19 points must give bit-identical masks and densities on every
    repetition and whatever the tiling (one call vs sixteen slices: a point's result depends on that point only)."""
    from enarf_gan_b200 import ops
    gs = b2["gs"]
    B, N = b2["cfg"]["B"], 1 << 19
    g = torch.Generator().manual_seed(77)
    centre = gs.pose[:, :, :3, 3].mean(dim=1)                                   # (B,3): the body's centre in camera space
    pts = (centre[:, :, None].cpu() + 1.2 * torch.randn(B, 3, N, generator=g)).cuda().contiguous()
    den0, _, m0 = ops.density_query(gs, pts, want_color=False, want_mask=True)
    assert (m0 != 0).float().mean().item() > 0.05, "the cloud must hit the parts"
    for _ in range(5):
        den, _, m = ops.density_query(gs, pts, want_color=False, want_mask=True)
        assert torch.equal(m, m0) and torch.equal(den, den0)
    step = N // 16
    for k in range(16):
        sl = pts[:, :, k * step:(k + 1) * step].contiguous()
        den, _, m = ops.density_query(gs, sl, want_color=False, want_mask=True)
        assert torch.equal(m, m0[:, k * step:(k + 1) * step]) and torch.equal(den, den0[:, :, k * step:(k + 1) * step])
