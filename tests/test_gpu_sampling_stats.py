"""Distribution test of the in-kernel PHILOX fine sampler: the reference draws its fine-pass bins with
``torch.multinomial(weights, Nf, replacement=True)`` (libraries/NeRF/rendering.py:192-194).  The kernel's draws cannot
reproduce that stream, so they are tested to follow the same categorical law: chi-square goodness of fit of the
kernel's bin counts against w / sum(w), with torch.multinomial put through the identical test as the control; the
jitter is tested for uniformity."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _chi2(counts, prob):
    exp = prob * counts.sum(dim=1, keepdim=True)
    return ((counts - exp) ** 2 / exp).sum(dim=1)


def test_philox_bins_follow_the_multinomial_law():
    from enarf_gan_b200 import ops
    R, Nc, Nf, K = 64, 48, 48, 400
    g = torch.Generator(device="cuda").manual_seed(7)
    # blurred-weight-like rows: a floor of 0.01 (rendering.py:190) plus a few bumps
    w = 0.01 + torch.rand(R, Nc, device="cuda", generator=g) ** 6
    prob = (w / w.sum(dim=1, keepdim=True)).double()
    ours = torch.zeros(R, Nc, device="cuda", dtype=torch.float64)
    ctrl = torch.zeros_like(ours)
    frac = []
    for k in range(K):
        bins, idx = ops.fine_sample(w, Nf, seed=1000 + k, want_idx=True)
        ours.scatter_add_(1, idx.long(), torch.ones(R, Nf, device="cuda", dtype=torch.float64))
        ctrl.scatter_add_(1, torch.multinomial(w, Nf, replacement=True, generator=g),
                          torch.ones(R, Nf, device="cuda", dtype=torch.float64))
        frac.append(bins * Nc - torch.floor(bins * Nc))          # the jitter inside the bin
        assert torch.all(bins[:, 1:] >= bins[:, :-1]), "bins must come out sorted"
        # bins = (idx + jitter) / Nc (rendering.py:195-196): each sorted bin lies in its index's cell (up to rounding)
        assert ((bins * Nc - idx.sort(dim=1)[0].float() - 0.5).abs() <= 0.5 + 1e-3).all()
    dof = R * (Nc - 1)
    sd = (2 * dof) ** 0.5
    s_ours, s_ctrl = _chi2(ours, prob).sum().item(), _chi2(ctrl, prob).sum().item()
    print(f"\n  chi2 over {R} rays x {K * Nf} draws: kernel {s_ours:.0f}, torch.multinomial {s_ctrl:.0f}, "
          f"expected {dof} +- {sd:.0f}")
    assert abs(s_ours - dof) < 5 * sd, (s_ours, dof, sd)
    assert abs(s_ctrl - dof) < 5 * sd, (s_ctrl, dof, sd)        # the control passes the same test
    # per-ray statistic: no single ray far out in the tail (chi2(47): mean 47, sd 9.7)
    assert _chi2(ours, prob).max().item() < 47 + 6 * 9.7
    # jitter ~ U[0,1): 20-cell chi-square over all draws
    f = torch.cat(frac).reshape(-1)
    cells = torch.histc(f, bins=20, min=0.0, max=1.0).double()
    c2 = ((cells - f.numel() / 20) ** 2 / (f.numel() / 20)).sum().item()
    print(f"  jitter uniformity chi2(19) = {c2:.1f}")
    assert c2 < 19 + 6 * (2 * 19) ** 0.5


def test_philox_draws_differ_between_rays_and_seeds():
    from enarf_gan_b200 import ops
    w = torch.full((8, 48), 1.0, device="cuda")
    a = ops.fine_sample(w, 48, seed=1)
    b = ops.fine_sample(w, 48, seed=2)
    assert not torch.equal(a, b)
    assert not torch.equal(a[0], a[1])
    assert torch.equal(a, ops.fine_sample(w, 48, seed=1)), "same seed, same draws"
