"""Losses of the GAN training step, mirroring ``models/loss.py:5-30`` and ``libraries/gan/loss.py`` of the reference
(same names, arguments and values).  They are O(B * size^2) element-wise work on the rendered mask /
discriminator logits and stay in torch."""
import torch
import torch.nn.functional as F


def push_to_background(fake_mask, background_ratio=0.3):
    """Mean square of the ``background_ratio`` smallest mask values of the whole batch (models/loss.py:5-13): the
    generator is asked to leave that share of the pixels empty."""
    if not background_ratio > 0:
        return 0
    flat = fake_mask.reshape(-1)
    k = int(flat.numel() * background_ratio)
    smallest = torch.topk(flat, k=k, largest=False, sorted=False)[0]
    return smallest.square().mean()


def nerf_bone_loss(fake_mask, bone_mask):
    """(1 - mask)^2 averaged over the pixels the 2-D bone rasterisation covers (models/loss.py:16-24); the bone mask is
    max-pooled down to the rendered resolution when it is larger."""
    assert fake_mask.ndim == bone_mask.ndim
    if fake_mask.shape[-1] != bone_mask.shape[-1]:
        rate = bone_mask.shape[-1] // fake_mask.shape[-1]
        bone_mask = F.max_pool2d(bone_mask[:, None], rate, rate, 0)[:, 0]
    on = bone_mask > 0.5
    return ((1 - fake_mask).square() * on).sum() / on.sum()


def nerf_patch_loss(fake_mask, bone_mask, background_ratio=0.3, coef=10):
    """models/loss.py:27-30."""
    return (push_to_background(fake_mask, background_ratio) + nerf_bone_loss(fake_mask, bone_mask)) * coef


def adv_loss_dis(real, fake, adv_loss_type, tmp=1.0):
    """libraries/gan/loss.py: hinge or non-saturating ("ce") discriminator loss."""
    if adv_loss_type == "hinge":
        return F.relu(1 - real).mean() + F.relu(1 + fake).mean()
    if adv_loss_type == "ce":
        return F.softplus(-real * tmp).mean() + F.softplus(fake * tmp).mean()
    raise ValueError(f"{adv_loss_type} is not supported")


def adv_loss_gen(fake, adv_loss_type, tmp=1.0):
    if adv_loss_type == "hinge":
        return -fake.mean()
    if adv_loss_type == "ce":
        return F.softplus(-fake * tmp).mean()
    raise ValueError(f"{adv_loss_type} is not supported")


def d_r1_loss(real_pred, real_img):
    """R1 penalty: E |d D(x) / d x|^2 over the real batch."""
    (grad,) = torch.autograd.grad(outputs=real_pred.sum(), inputs=real_img, create_graph=True)
    return grad.square().reshape(grad.shape[0], -1).sum(1).mean()
