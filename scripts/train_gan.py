#!/usr/bin/env python
"""BASELINE configs[4]: the full ENARF-GAN training step -- generator render through the fused kernels + discriminator,
NCCL gradient all-reduce -- one process per GPU, synthetic data.

    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 scripts/train_gan.py \
        --global-batch 64 --iters 20
    python scripts/train_gan.py --global-batch 8 --iters 10          # one GPU

Prints one JSON line on rank 0: s/iter, rendered rays/s, and the share of the step spent waiting for the gradient
all-reduce (measured as the difference to the same step run under DDP's no_sync).

The two networks AROUND the renderer are outside this repository's scope (StyleGAN2-ADA tri-plane
generator, StyleGAN2 discriminator), so stand-ins with the same interfaces and comparable parameter counts are used:
``PlaneProducer`` (latent + bone-length code -> (B,168,256,256) tri-planes) and ``ConvDiscriminator`` (residual
stride-2 conv stack, ``forward(img, ddp, world_size)``).  The step itself -- ``enarf_gan_b200.train.train_step`` -- is
the one a reference user would call.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch import nn  # noqa: E402

SIZE, NC, NF = 128, 48, 48


class PlaneProducer(nn.Module):
    """Stand-in for the StyleGAN2-ADA tri-plane generator: called as ``tri_plane_gen(z, c, truncation_psi=)`` with c the
    bone-length positional encoding (narf.py:286-289).  A learnable base plane set plus a per-sample low-rank term."""

    def __init__(self, z_dim=512, c_dim=192, channels=168, res=256, rank=8, channels_last=False):
        super().__init__()
        # channels_last: emit the renderer's own layout -- (B,3,H,W,32) feature planes + (B,72,H,W) probability planes
        #  -- instead of the reference's NCHW (B,168,H,W); the parameters are the same either way
        self.channels_last = channels_last
        self.base = nn.Parameter(torch.randn(1, channels, res, res) * 0.5)
        self.to_code = nn.Sequential(nn.Linear(z_dim + c_dim, 256), nn.LeakyReLU(0.2), nn.Linear(256, channels * rank))
        self.basis = nn.Parameter(torch.randn(rank, res, res) * 0.1)
        self.rank, self.channels = rank, channels
        with torch.no_grad():
            self.base[:, 96:] *= 4.0          # probability channels ~ N(0, 2^2) like the synthetic fixtures

    def forward(self, z, c, truncation_psi=1):
        code = self.to_code(torch.cat([z, c], dim=1)).reshape(-1, self.channels, self.rank)
        if not self.channels_last:
            return self.base + torch.einsum("bcr,rhw->bchw", code, self.basis)
        B, H, W = z.shape[0], self.base.shape[2], self.base.shape[3]
        feat = self.base[:, :96].reshape(1, 3, 32, H, W).permute(0, 1, 3, 4, 2) \
            + torch.einsum("bkcr,rhw->bkhwc", code[:, :96].reshape(B, 3, 32, self.rank), self.basis)
        prob = self.base[:, 96:] + torch.einsum("bcr,rhw->bchw", code[:, 96:], self.basis)
        return feat.contiguous(), prob


class ConvDiscriminator(nn.Module):
    """Stand-in for the StyleGAN2 discriminator (custom_stylegan2/net.py:609-676): 128 -> 4 residual stride-2 blocks."""

    def __init__(self, size=128, width=64):
        super().__init__()
        ch = [width, width * 2, width * 4, width * 8, width * 8, width * 8]
        self.stem = nn.Conv2d(3, ch[0], 1)
        blocks = []
        for i in range(5):
            blocks.append(nn.ModuleDict(dict(a=nn.Conv2d(ch[i], ch[i], 3, padding=1), b=nn.Conv2d(ch[i], ch[i + 1], 3, 2, 1),
                                             skip=nn.Conv2d(ch[i], ch[i + 1], 1, 2, bias=False))))
        self.blocks = nn.ModuleList(blocks)
        self.head = nn.Sequential(nn.Flatten(), nn.Linear(ch[5] * (size // 32) ** 2, ch[5]), nn.LeakyReLU(0.2), nn.Linear(ch[5], 1))

    def forward(self, img, ddp=False, world_size=1):
        x = nn.functional.leaky_relu(self.stem(img), 0.2)
        for blk in self.blocks:
            y = nn.functional.leaky_relu(blk["a"](x), 0.2)
            y = nn.functional.leaky_relu(blk["b"](y), 0.2)
            x = (y + blk["skip"](x)) * 2 ** -0.5
        return self.head(x)


def build(device, ert_eps=0.0, channels_last=False):
    from enarf_gan_b200 import synthetic as S
    from enarf_gan_b200.generator import TriNARFGenerator
    cfg = dict(z_dim=256, background_ratio=0.3,
               nerf_params=dict(Nc=NC, Nf=NF, origin_location="center+head", coordinate_scale=3, render_bs=16384, ert_eps=ert_eps))
    gen = TriNARFGenerator(cfg, SIZE, num_bone=24, num_bone_param=23, parent_id=S.SMPL_PARENTS, black_background=True,
                           tri_plane_gen=PlaneProducer(channels_last=channels_last))
    gen.register_canonical_pose(S.canonical_pose())
    gen.nerf.rng_mode = "philox"
    return gen.to(device).train(), ConvDiscriminator(SIZE).to(device).train()


def synthetic_batch(first, B, device):
    from enarf_gan_b200 import synthetic as S
    per = [S.make_inputs(1, SIZE, seed=4321 + 17 * (first + i)) for i in range(B)]
    inp = {k: torch.cat([q[k] for q in per], dim=0).to(device) for k in per[0]}
    g = torch.Generator().manual_seed(99 + first)
    inp["bone_mask"] = (torch.rand(B, SIZE, SIZE, generator=g) > 0.8).float().to(device)
    inp["real_img"] = torch.randn(B, 3, SIZE, SIZE, generator=g).clamp(-1, 1).to(device)
    return inp


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--global-batch", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--n-accum", type=int, default=1)
    ap.add_argument("--channels-last", action="store_true", help="producer emits the renderer's plane layout (no repack / unpack)")
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    from enarf_gan_b200 import rendering, train
    lo, hi = rendering.shard_samples(a.global_batch, rank, world)
    B = hi - lo
    gen, dis = build(device, channels_last=a.channels_last)
    gen_ddp, dis_ddp = train.wrap_ddp(gen, device), train.wrap_ddp(dis, device)
    cfg = train.StepConfig(z_total=256 * 3, n_accum_step=a.n_accum, adv_loss_type="ce", tri_plane_reg_coef=0.0)
    go = torch.optim.Adam(gen_ddp.parameters(), lr=1e-4, betas=(0.0, 0.99))
    do = torch.optim.Adam(dis_ddp.parameters(), lr=2e-4, betas=(0.0, 0.99))
    d = synthetic_batch(lo, B, device)

    def step(it, g_=gen_ddp, d_=dis_ddp):
        return train.train_step(it, B, g_, d_, go, do, cfg, d["pose_to_camera"], d["pose_to_world"], d["bone_length"],
                                d["inv_intrinsics"], d["bone_mask"], d["real_img"], ddp=world > 1, world_size=world)

    def timed(fn, iters):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for it in range(iters):
            _, log = fn(1 + it)            # 1..: the lazy R1 runs on iteration 16
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item(), log

    for it in range(a.warmup):
        step(1 + it)
    ms, log = timed(step, a.iters)
    ms_nosync = None
    if world > 1:      # the same step without the gradient all-reduce (parameters drift apart: timing only, done last)
        def step_nosync(it):
            with gen_ddp.no_sync(), dis_ddp.no_sync():
                return step(it)
        ms_nosync, _ = timed(step_nosync, a.iters)
    if rank == 0:
        n_par = lambda m: sum(p.numel() for p in m.parameters())
        out = {"workload": f"BASELINE configs[4]: ENARF-GAN training step (generator render {SIZE}x{SIZE}, {NC}+{NF} samples, P=24 "
                           f"+ discriminator, Adam), global batch {a.global_batch} over {world} GPU(s), stand-in producer / "
                           f"discriminator networks", "n_gpus": world, "global_batch": a.global_batch, "samples_per_gpu": B,
               "s_per_iter": ms * 1e-3, "ms_per_iter": ms, "rays_per_s": a.global_batch * SIZE * SIZE / (ms * 1e-3),
               "generator_params_M": n_par(gen) / 1e6, "discriminator_params_M": n_par(dis) / 1e6,
               "losses": {k: float(v) for k, v in log.items()}}
        if ms_nosync is not None:
            out["ms_per_iter_without_allreduce"] = ms_nosync
            out["allreduce_share_of_step"] = max(0.0, (ms - ms_nosync) / ms)
            out["nccl_ranks"] = dist.get_world_size()
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
