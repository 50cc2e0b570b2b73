"""The GAN training step (enarf_gan_b200.train.train_step) on one GPU with the real kernels: the step
runs, every generator / discriminator parameter moves, losses stay finite -- with the stand-in networks of
scripts/train_gan.py and, when the reference archive is staged, with the reference's own StyleGAN2 discriminator."""
import importlib.util
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _script():
    spec = importlib.util.spec_from_file_location("train_gan", os.path.join(ROOT, "scripts", "train_gan.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _step(gen, dis, B, iters=2, n_accum=1):
    from enarf_gan_b200 import train
    T = _script()
    dev = torch.device("cuda", 0)
    cfg = train.StepConfig(z_total=256 * 3, n_accum_step=n_accum, adv_loss_type="ce", tri_plane_reg_coef=1e-3, r1_every=2)
    go = torch.optim.Adam(gen.parameters(), lr=1e-3, betas=(0.0, 0.99))
    do = torch.optim.Adam(dis.parameters(), lr=2e-3, betas=(0.0, 0.99))
    d = T.synthetic_batch(0, B, dev)
    before = [p.detach().clone() for p in list(gen.parameters()) + list(dis.parameters())]
    logs = []
    for it in range(1, 1 + iters):
        fake, log = train.train_step(it, B, gen, dis, go, do, cfg, d["pose_to_camera"], d["pose_to_world"], d["bone_length"],
                                     d["inv_intrinsics"], d["bone_mask"], d["real_img"])
        logs.append(log)
        assert fake.shape == (B, 3, T.SIZE, T.SIZE) and torch.isfinite(fake).all()
    after = list(gen.parameters()) + list(dis.parameters())
    moved = [not torch.equal(a, b) for a, b in zip(after, before)]
    names = [n for n, _ in list(gen.named_parameters()) + list(dis.named_parameters())]
    n_gen = len(list(gen.parameters()))
    still = [n for n, m in zip(names[:n_gen], moved[:n_gen]) if not m and "noise" not in n]
    assert not still, f"generator parameters that never moved: {still}"
    # (a saturated discriminator can hand its last bias an exactly-zero gradient: ask for the bulk)
    assert sum(moved[n_gen:]) >= 0.9 * len(moved[n_gen:]), "the discriminator must train too"
    for log in logs:
        assert all(torch.isfinite(v) for v in log.values()), log
    assert "r1_reg" in logs[1]
    return logs


def test_training_step_with_stand_in_networks():
    T = _script()
    gen, dis = T.build(torch.device("cuda", 0))
    _step(gen, dis, B=2)
    _step(gen, dis, B=4, n_accum=2)


def test_training_step_with_the_reference_discriminator():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import ref_harness as H
    if not H.available():
        pytest.skip("reference archive not staged")
    H.install()
    from libraries.custom_stylegan2.net import Discriminator
    T = _script()
    gen, _ = T.build(torch.device("cuda", 0))
    dis = Discriminator(H._EasyDict(minibatch_std=False), size=T.SIZE).cuda().train()
    _step(gen, dis, B=2)


def test_channels_last_producer_hand_off_matches_nchw():
    """SURVEY 8f N1, GAN half: a producer that emits (channels-last feature planes, probability planes) renders the
    same image and receives the same parameter gradients as the NCHW producer with the same parameters -- without the
    repack / unpack kernels."""
    T = _script()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    gen_a, _ = T.build(dev, channels_last=False)
    gen_b, _ = T.build(dev, channels_last=True)
    gen_b.load_state_dict(gen_a.state_dict())
    d = T.synthetic_batch(0, 2, dev)
    z = torch.randn(2, 768, device=dev)
    outs = []
    for gen in (gen_a, gen_b):
        gen.nerf.rng_seed = 5
        img, mask, _, _ = gen(d["pose_to_camera"], d["pose_to_world"], d["bone_length"], z, d["inv_intrinsics"])
        ((img ** 2).mean() + ((1 - mask) ** 2).mean()).backward()
        outs.append((img.detach(), mask.detach(), {k: p.grad.clone() for k, p in gen.named_parameters() if p.grad is not None}))
    (ia, ma, ga), (ib, mb, gb) = outs
    rel = lambda a, b: ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()
    assert rel(ib, ia) < 1e-5 and rel(mb, ma) < 1e-5
    assert set(ga) == set(gb)
    for k in ga:
        assert rel(gb[k], ga[k]) < 1e-4, k
