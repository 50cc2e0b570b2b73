"""The GAN training step around the renderer (BASELINE configs[4]), one process per GPU.

``train_step`` mirrors ``train_ENARF_GAN.py:102-170`` of the reference -- generator update over ``n_accum_step``
micro-batches (adversarial + bone-guided mask loss + optional tri-plane regulariser), discriminator update on the
detached fakes and a real batch, lazy R1 every 16 iterations -- with what the reference leaves out for multi-GPU use:

* ``wrap_ddp`` puts generator and discriminator under ``DistributedDataParallel`` (NCCL over NVLink on GPUs, gloo in
  the CPU tests): the ONLY exchange of the training step is the bucketed gradient all-reduce of G and D, overlapped
  with the backward by DDP's hooks; the render itself is sample-sharded and needs nothing but the 2-float batch-global
  near/far (``rendering.global_near_far``).  The reference's own DDP branch is dead code (train_ENARF_GAN.py:203-206);
* micro-batches but the last run under ``no_sync`` so that gradients cross the wire once per optimiser step;
* no ``torch.cuda.empty_cache()`` stalls (train_ENARF_GAN.py:138,158,168), latents drawn on the device.

The discriminator is a black box here (``dis(img, ddp, world_size)`` as in custom_stylegan2/net.py:655); the reference's
StyleGAN2 discriminator, or any module with that call, is injected by the caller.
"""
import contextlib
from dataclasses import dataclass
from typing import Optional

import torch
from torch import nn

from . import losses


@dataclass
class StepConfig:
    """The values the reference's train_step reads from its global ``config`` (configs/enarfgan_train/*/config.yml)."""
    z_total: int                      # width of the latent the generator's forward splits (z_dim * 3 or * 4)
    n_accum_step: int = 1
    adv_loss_type: str = "ce"
    bone_guided_coef: float = 1.0     # config.loss.bone_guided_coef
    tri_plane_reg_coef: float = 0.0   # config.loss.tri_plane_reg_coef
    r1_loss_coef: float = 10.0
    r1_every: int = 16


def wrap_ddp(module: nn.Module, device: Optional[torch.device] = None, **kw) -> nn.Module:
    """DistributedDataParallel when a process group with more than one rank is up, the bare module otherwise."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return module
    if not any(p.requires_grad for p in module.parameters()):
        return module
    ids = [device.index] if device is not None and device.type == "cuda" else None
    # buffers (canonical pose ...) are constants set identically on every rank: no per-forward broadcast
    kw.setdefault("broadcast_buffers", False)
    kw.setdefault("gradient_as_bucket_view", True)
    return nn.parallel.DistributedDataParallel(module, device_ids=ids, **kw)


def _unwrap(m):
    return m.module if isinstance(m, nn.parallel.DistributedDataParallel) else m


def _no_sync(m, active):
    return m.no_sync() if active and isinstance(m, nn.parallel.DistributedDataParallel) else contextlib.nullcontext()


def generator_loss(gen, dis, fake_img, fake_low_res_mask, bone_mask, cfg: StepConfig, ddp, world_size):
    """``loss`` of train_ENARF_GAN.py:79-99 (without its optimiser zeroing, which train_step does once)."""
    background_ratio = _unwrap(gen).background_ratio
    loss_bone = losses.nerf_patch_loss(fake_low_res_mask, bone_mask, background_ratio) * cfg.bone_guided_coef
    # the bare module: no discriminator gradient is produced here, so DDP's reducer must not wait for one
    loss_adv = losses.adv_loss_gen(_unwrap(dis)(fake_img, ddp, world_size), cfg.adv_loss_type, tmp=1)
    total = loss_adv + loss_bone
    if cfg.tri_plane_reg_coef > 0:
        planes = _unwrap(gen).nerf.buffers_tensors["tri_plane_feature"]
        if isinstance(planes, (tuple, list)):      # handed over as (channels-last features, probabilities)
            sq = sum(p.float().square().sum() for p in planes) / sum(p.numel() for p in planes)
        else:
            sq = planes.square().mean()
        total = total + sq * cfg.tri_plane_reg_coef
    return total, {"adv_loss_gen": loss_adv.detach(), "bone_loss": loss_bone.detach()}


def train_step(it, batchsize, gen, dis, gen_optimizer, dis_optimizer, cfg: StepConfig, pose_to_camera, pose_to_world,
               bone_length, inv_intrinsic, bone_mask, real_img, ddp=False, world_size=1, generator=None):
    """One training iteration on this rank's ``batchsize`` samples.  -> (fake images (B,3,s,s) detached, dict of losses).
    ``gen`` / ``dis`` may be DDP-wrapped (``wrap_ddp``); every tensor argument lives on this rank's device."""
    device = pose_to_camera.device
    forward_bs = batchsize // cfg.n_accum_step
    fakes, log = [], {}
    # ---- generator ---------------------------------------------------------------------------------------
    dis.requires_grad_(False)
    gen_optimizer.zero_grad(set_to_none=True)
    starts = list(range(0, batchsize, forward_bs))
    for i in starts:
        sl = slice(i, i + forward_bs)
        z = torch.randn(forward_bs, cfg.z_total, device=device, generator=generator)
        with _no_sync(gen, i != starts[-1]):          # gradients cross the wire once, on the last micro-batch
            fake_img, fake_mask, _, _ = gen(pose_to_camera[sl], pose_to_world[sl], bone_length[sl], z, inv_intrinsic[sl])
            loss_gen, log = generator_loss(gen, dis, fake_img, fake_mask, bone_mask[sl], cfg, ddp, world_size)
            loss_gen.backward()
        fakes.append(fake_img.detach())
    fake_img = torch.cat(fakes)
    gen_optimizer.step()
    # ---- discriminator -----------------------------------------------------------------------------------
    dis.requires_grad_(True)
    dis_optimizer.zero_grad(set_to_none=True)
    loss_dis = losses.adv_loss_dis(dis(real_img, ddp, world_size), dis(fake_img, ddp, world_size), cfg.adv_loss_type)
    loss_dis.backward()
    dis_optimizer.step()
    log["adv_loss_dis"] = loss_dis.detach()
    if it % cfg.r1_every == 0:                        # lazy R1 (train_ENARF_GAN.py:155-169)
        dis_optimizer.zero_grad(set_to_none=True)
        real = real_img.detach().requires_grad_(True)
        dis_real = dis(real, ddp, world_size)
        r1 = losses.d_r1_loss(dis_real, real)
        (0.5 * r1 * cfg.r1_every * cfg.r1_loss_coef + 0 * dis_real[0]).sum().backward()   # 0 * logit keeps every
        dis_optimizer.step()                                                              # parameter in the graph
        log["r1_reg"] = r1.detach()
    return fake_img, log
