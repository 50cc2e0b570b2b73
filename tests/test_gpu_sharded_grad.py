"""B < G with gradients on real GPUs: one image's rays split over 2 ranks (NCCL), the owner's tri-plane and
decoder weights broadcast, every rank's partial gradients reduced onto the owner -- the owner's gradients equal the
one-GPU render_entire_img ones up to the order of the atomic accumulation.  Needs >= 2 GPUs (skipped otherwise; run
with ``gpurun --gpus 2``)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu
SIZE, NC, NF = 64, 48, 48


def _generator(device):
    for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import golden_util as G
    from test_gpu_api import make_generator
    g, cfg, inp, tp, layers, z_rend = G.load("config1_b1_s64")
    gen = make_generator(cfg, tp, g, render_bs=1 << 20).to(device).train()
    gen.nerf.rng_mode = "given"            # the same fine-pass draws on every layout: compare gradients, not noise
    return gen, cfg, inp, g


def _loss(color, mask, disp):
    g = torch.Generator().manual_seed(1)
    wc = torch.randn(color.shape, generator=g).to(color.device)
    wm = torch.randn(mask.shape, generator=g).to(color.device)
    return (color * wc).sum() + (mask * wm).sum() + (disp * wm).sum() * 0.1


def _draws(gen, n_valid, lo=0, hi=None):
    g = torch.Generator().manual_seed(7)
    idx = torch.randint(0, NC, (n_valid, NF), generator=g)
    jit = torch.rand(n_valid, NF, generator=g)
    return idx[lo:hi].cuda(), jit[lo:hi].cuda()


def _render(gen, cfg, inp, sharded, group=None):
    """Full-image render with gradients; the fine-pass draws are tied to the RAY (not to the rank's block) so that the
    sharded and the one-GPU render sample the same points."""
    from enarf_gan_b200 import ops, rendering, synthetic as S
    dev = next(gen.parameters()).device
    c = lambda k: inp[k].to(dev)
    z = S.make_latent(1, seed=cfg["seed"]).to(dev)
    z_nerf, z_rend = z[:, :512], z[:, 512:]
    nerf = gen.nerf
    # valid rays of the whole image, to cut the per-ray draws for this rank's block
    pose, bl = nerf.transform_pose(c("pose_to_camera"), c("bone_length"))
    ps = rendering._scaled(nerf, pose)
    valid = ops.depth_range(rendering.pixel_grid(SIZE, SIZE, 1, dev), c("inv_intrinsics"), ps)[3][0]
    n = SIZE * SIZE
    lo, hi = (0, n)
    if sharded:
        lo, hi = rendering.shard_rays(n, dist.get_rank(), dist.get_world_size())
    before = int(valid[:lo].sum())
    mine = int(valid[lo:hi].sum())
    nerf.rng_given = _draws(gen, int(valid.sum()), before, before + mine)
    if sharded:
        return nerf.render_entire_img_sharded(c("pose_to_camera"), c("inv_intrinsics"), z_nerf, z_rend, c("bone_length"),
                                              render_size=SIZE, Nc=NC, Nf=NF, no_grad=False, group=group)
    return nerf.render_entire_img(c("pose_to_camera"), c("inv_intrinsics"), z_nerf, z_rend, c("bone_length"),
                                  render_size=SIZE, Nc=NC, Nf=NF, no_grad=False)


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        gen, cfg, inp, g = _generator(torch.device("cuda", rank))
        gen.nerf.sync_near_far = False
        color, mask, disp = _render(gen, cfg, inp, sharded=True)
        _loss(color, mask, disp).backward()
        grads = {k: (p.grad.detach().cpu() if p.grad is not None else None) for k, p in gen.named_parameters()}
        torch.save({"color": color.detach().cpu(), "mask": mask.detach().cpu(), "grads": grads}, os.path.join(out_dir, f"s{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_ray_sharded_gradients_match_one_gpu(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (torch.load(os.path.join(tmp_path, f"s{r}.pt")) for r in range(2))
    gen, cfg, inp, g = _generator(torch.device("cuda", 0))
    color, mask, disp = _render(gen, cfg, inp, sharded=False)
    _loss(color, mask, disp).backward()
    rel = lambda a, b: ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()
    assert rel(r0["color"], color.detach().cpu()) < 1e-5 and rel(r1["mask"], mask.detach().cpu()) < 1e-5
    for k, p in gen.named_parameters():
        if p.grad is None:
            continue
        e = rel(r0["grads"][k], p.grad.detach().cpu())
        print(f"  {k}: owner vs one GPU {e:.2e}")
        assert e < 1e-4, k
        other = r1["grads"][k]
        assert other is None or other.abs().sum() == 0, f"non-owner got a gradient for {k}"
