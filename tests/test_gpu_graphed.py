"""CUDA-graph capture of the generator step (enarf_gan_b200.graphed.GraphedRenderStep): a replay computes what the eager
step computes for the same inputs and the same device-side Philox seed, draws afresh on the next replay, and follows
new inputs."""
import pytest
import torch

import golden_util as G
from enarf_gan_b200 import synthetic as S
from test_gpu_api import make_generator

pytestmark = pytest.mark.gpu


def _loss(color, mask):
    return (color ** 2).mean() + ((1 - mask) ** 2).mean()


def test_graph_replay_matches_eager_and_redraws():
    from enarf_gan_b200.graphed import GraphedRenderStep
    g, cfg, inp, tp, layers, z_rend = G.load("b2_s24")
    gen = make_generator(cfg, tp, g).train()
    gen.nerf.rng_mode = "philox"
    gen.nerf.rng_seed_dev = torch.zeros(1, device="cuda", dtype=torch.int64)
    z = S.make_latent(cfg["B"], seed=3).cuda()
    args = [inp[k].cuda() for k in ("pose_to_camera", "pose_to_world", "bone_length")] + [z, inp["inv_intrinsics"].cuda()]
    step = GraphedRenderStep(gen, _loss, *args)
    # replay with a known seed
    gen.nerf.rng_seed_dev.fill_(1000)
    color, mask, loss = step(*args)
    c1, l1 = color.clone(), loss.clone()
    grads1 = {k: p.grad.clone() for k, p in gen.named_parameters() if p.grad is not None}
    assert int(gen.nerf.rng_seed_dev) == 1001, "the seed advances inside the graph"
    # the eager step at the same seed
    gen.nerf.rng_seed_dev.fill_(1000)
    gen.zero_grad(set_to_none=True)
    ce, me, _, _ = gen(*args)
    le = _loss(ce, me)
    le.backward()
    assert torch.equal(ce, c1), "same seed, same image"
    assert abs(le.item() - l1.item()) < 1e-6
    rel = lambda a, b: ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()
    for k, p in gen.named_parameters():
        if p.grad is not None:
            assert rel(grads1[k], p.grad) < 1e-4, k
    # the next replay draws other samples, and a different latent changes the decoder
    color2, _, _ = step(*args)
    assert not torch.equal(color2, c1)
    args2 = list(args)
    args2[3] = S.make_latent(cfg["B"], seed=4)          # a HOST tensor: copied in by the step
    gen.nerf.rng_seed_dev.fill_(1000)
    color3, _, _ = step(*args2)
    assert not torch.equal(color3, c1)
