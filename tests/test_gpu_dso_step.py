"""DSO step kernels around the renderer (csrc/dso.cu): ray selection against the reference's
mask_based_sampler computation (max_pool2d + noise + topk, libraries/NeRF/ray_sampler.py:7-39) and the fused photometric
loss against libraries/NeRF/loss.py:14-48 restated in torch, forward and gradients."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _mask(B, h, w, seed):
    g = torch.Generator().manual_seed(seed)
    m = torch.zeros(B, h, w)
    for b in range(B):     # a few blobs
        for _ in range(3):
            cy, cx = torch.randint(0, h, (1,), generator=g).item(), torch.randint(0, w, (1,), generator=g).item()
            m[b, max(0, cy - 9):cy + 9, max(0, cx - 6):cx + 6] = 1.0
    return m.cuda()


@pytest.mark.parametrize("B,h,w,k", [(2, 96, 80, 500), (1, 200, 200, 40000), (3, 64, 64, 64 * 64), (2, 256, 256, 4096)])
def test_ray_selection_matches_maxpool_noise_topk(B, h, w, k):
    from enarf_gan_b200 import ops
    mask = _mask(B, h, w, 3)
    noise = torch.rand(B, h, w, generator=torch.Generator().manual_seed(4)).cuda()
    idx, homo = ops.select_rays(mask, k, noise=noise, radius=64)
    # the reference's computation
    dil = F.max_pool2d(mask[:, None], 129, stride=1, padding=64)[:, 0]
    score = (dil + noise).reshape(B, -1)
    ref = torch.topk(score, k, dim=1, sorted=False)[1]
    assert torch.equal(idx.sort(dim=1)[0], ref.sort(dim=1)[0]), "the selected set must be torch.topk's"
    assert torch.equal(idx, idx.sort(dim=1)[0]), "ascending pixel order"
    assert torch.equal(homo[:, 0], (idx % w).float() + 0.5) and torch.equal(homo[:, 1], (idx // w).float() + 0.5)
    assert torch.all(homo[:, 2] == 1)


def test_ray_selection_with_ties_and_philox():
    from enarf_gan_b200 import ops
    mask = _mask(2, 64, 64, 5)
    # all-equal noise: every dilated pixel ties; exactly k come out, lowest pixel ids first within the tie
    idx, _ = ops.select_rays(mask, 100, noise=torch.zeros(2, 64, 64).cuda(), radius=2)
    dil = F.max_pool2d(mask[:, None], 5, stride=1, padding=2)[:, 0].reshape(2, -1)
    for b in range(2):
        on = dil[b].nonzero()[:, 0]
        assert torch.equal(idx[b], on[:100]) if on.numel() >= 100 else set(on.tolist()) <= set(idx[b].tolist())
    a, _ = ops.select_rays(mask, 300, seed=1)
    b_, _ = ops.select_rays(mask, 300, seed=1)
    c, _ = ops.select_rays(mask, 300, seed=2)
    assert torch.equal(a, b_) and not torch.equal(a, c)
    assert all(len(set(r.tolist())) == 300 for r in a)


@pytest.mark.parametrize("kind", ["mse", "mae"])
@pytest.mark.parametrize("with_mask", [True, False])
def test_photometric_loss_matches_torch(kind, with_mask):
    from enarf_gan_b200.generator import PhotometricLoss
    B, size, k = 3, 48, 700
    g = torch.Generator().manual_seed(8)
    gt_c = torch.randn(B, 3, size, size, generator=g).cuda()
    gt_m = torch.rand(B, size, size, generator=g).cuda()
    idx = torch.stack([torch.randperm(size * size, generator=g)[:k] for _ in range(B)]).cuda()
    # rendered values near the ground truth so that the MAE dead zone (|diff| < 0.01) is exercised
    rc0 = (torch.gather(gt_c.reshape(B, 3, -1), 2, idx[:, None].expand(-1, 3, -1)) + 0.02 * torch.randn(B, 3, k, generator=g).cuda())
    rm0 = torch.rand(B, k, generator=g).cuda()
    cfg = dict(nerf_loss_type=kind, color_coef=0.7, mask_coef=1.3)
    out = []
    for fused in (True, False):
        rc, rm = rc0.clone().requires_grad_(), rm0.clone().requires_grad_()
        if fused:
            lc, lm = PhotometricLoss(cfg)(idx, rc, rm, gt_c, gt_m if with_mask else None)
        else:     # loss.py:14-48 in torch
            sc = torch.gather(gt_c.reshape(B, 3, -1), 2, idx[:, None].expand(-1, 3, -1))
            lc = (F.mse_loss(sc, rc) if kind == "mse" else torch.clamp_min((sc - rc).abs(), 0.01).mean()) * 0.7
            lm = F.mse_loss(torch.gather(gt_m.reshape(B, -1), 1, idx), rm) * 1.3 if with_mask else 0
        (lc * 2.0 + lm * 3.0).backward()
        out.append((float(lc), float(lm), rc.grad.clone(), rm.grad.clone() if rm.grad is not None else None))
    (lc1, lm1, gc1, gm1), (lc2, lm2, gc2, gm2) = out
    assert abs(lc1 - lc2) < 1e-5 * max(1, abs(lc2)) and abs(lm1 - lm2) < 1e-5 * max(1, abs(lm2))
    assert torch.allclose(gc1, gc2, rtol=1e-4, atol=1e-8)
    if with_mask:
        assert torch.allclose(gm1, gm2, rtol=1e-4, atol=1e-8)
