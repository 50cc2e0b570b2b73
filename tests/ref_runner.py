"""Run the UNMODIFIED reference generator (``tests/golden/ref_harness.py``) on one synthetic case -- on the CPU in the
authoring container, on CUDA on the GPU box -- recording everything a parity test needs: outputs, the fine pass's
RNG draws, the part-validity bits the reference itself computed, and autograd gradients of a fixed linear loss with
respect to the tri-plane and every decoder parameter.  TEST INFRASTRUCTURE (used by ``tests/golden/make_golden_grad.py``
and the ``-m gpu`` full-size parity tests)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.dirname(HERE), os.path.join(HERE, "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)

import ref_harness as H  # noqa: E402
from enarf_gan_b200 import synthetic as S  # noqa: E402


def available():
    return H.available()


def loss_weights(B, n_all, n, Nf, seed=5):
    """Fixed random cotangents: loss = <color, w_c> + <mask, w_m> + <disparity, w_d> + <fine_weights, w_f>.  Colour, mask
    and disparity are the full (zero-filled for B == 1) ``n_all`` rays; fine_weights keeps the ``n`` rays the render
    kept (rendering.py:107-110)."""
    g = torch.Generator().manual_seed(seed)
    return dict(color=torch.randn(B, 3, n_all, generator=g), mask=torch.randn(B, n_all, generator=g),
                disparity=torch.randn(B, n_all, generator=g) * 3, fine_weights=torch.randn(B, n, Nf - 1, generator=g))


def linear_loss(color, mask, disparity, fine_weights, lw):
    dev = color.device
    B = color.shape[0]
    return ((color.reshape(B, 3, -1) * lw["color"].to(dev)).sum() + (mask.reshape(B, -1) * lw["mask"].to(dev)).sum()
            + (disparity.reshape(B, -1) * lw["disparity"].to(dev)).sum()
            + (fine_weights.reshape(lw["fine_weights"].shape) * lw["fine_weights"].to(dev)).sum())


def decoder_state(gen):
    """The reference's decoder parameters under the names our mirror's ``load_state_dict`` expects."""
    sd = {}
    for i, layer in enumerate(gen.nerf.mlp.layers):
        sd[f"nerf.mlp.layers.{i}.conv.weight"] = layer.conv.weight.detach().clone()
        sd[f"nerf.mlp.layers.{i}.conv.modulation.weight"] = layer.conv.modulation.weight.detach().clone()
        sd[f"nerf.mlp.layers.{i}.conv.modulation.bias"] = layer.conv.modulation.bias.detach().clone()
        sd[f"nerf.mlp.layers.{i}.bias"] = layer.bias.detach().clone()
    return sd


def run(B, size, Nc, Nf, seed, origin="center+head", device="cpu", grad=True, keep_masks=True, loss_seed=5,
        inputs=None, dtype=torch.float32, allow_tf32=False,
        force_draws=None):
    """``allow_tf32``: torch's cuDNN convolutions default to TF32 on Ampere+ (``torch.backends.cudnn.allow_tf32`` is True
    out of the box, in the reference's torch 1.10 as in 2.11), so ON A GPU the reference's grouped ``conv1d`` decoder
    (custom_stylegan2/net.py:250) runs at ~1e-3 relative precision "as shipped".  Parity is adjudicated against the
    same code with that switch off (true fp32); ``allow_tf32=True`` reproduces the stock behaviour for the record.
    ``force_draws=(idx, jitter)`` pins the fine pass's multinomial indices and jitter (e.g. to those of an fp64 run, so
    that the fp32 and the fp64 evaluation of the reference can be compared sample for sample).
    ``dtype=torch.float64`` runs the same unmodified code with torch's default dtype switched to double (so every
    tensor the reference creates itself is fp64 too): the "fp64-cast reference" of SURVEY 8c(2), free of the fp32
    x-remap noise of its batched gather (H1) -- the adjudicator when fp32 reference and fp32 kernels disagree at 1e-4."""
    if inputs is None:          # generated under the fp32 default: the same numbers whatever ``dtype`` the run uses
        P = 24 if origin == "center+head" else 23
        inputs = (S.make_inputs(B, size, seed=seed), S.make_triplane(B, num_part=P, seed=seed), S.make_latent(B, seed=seed))
    # the random-init generator too is drawn under the fp32 default (an fp64 default would draw different weights)
    H.set_device("cpu")
    gen = H.build_generator(size, S.canonical_pose(), S.SMPL_PARENTS, seed=seed, Nc=Nc, Nf=Nf, origin_location=origin)
    with torch.no_grad():
        for layer in gen.nerf.mlp.layers:       # non-trivial biases so the bias path is exercised
            layer.bias.normal_(0, 0.3)
    if dtype == torch.float64:
        torch.set_default_dtype(torch.float64)
    keep = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = bool(allow_tf32)     # the stock default is True
    torch.backends.cuda.matmul.allow_tf32 = False          # the stock default (torch >= 1.12)
    try:
        return _run(gen, B, size, Nc, Nf, seed, origin, device, grad, keep_masks, loss_seed, inputs, dtype, force_draws)
    finally:
        torch.set_default_dtype(torch.float32)
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = keep


def _run(gen, B, size, Nc, Nf, seed, origin, device, grad, keep_masks, loss_seed, inputs, dtype, force_draws=None):
    """-> dict with the reference's outputs (on ``device``):
    color (B,3,size,size), mask (B,size,size), disparity (B,n_all) [already x coordinate_scale], fine_weights
    (B,n,Nf-1), fine_depth (B,n,Nf), validity (B,n_all) bool, idx (B*n,Nf) / jitter (B*n,Nf) = the fine pass's draws,
    valid_parts (B,P,n*Nf) bool (reference's own in_cube & |canon|<1 at the fine samples), decoder state dict, and with
    ``grad``: d_tri_plane (B,C,256,256), param_grads {state-dict name: grad}, loss."""
    H.set_device(device)
    dev = torch.device(device)
    P = gen.nerf.num_bone
    inp, tp, z = inputs
    assert tp.shape[1] == (32 + P) * 3
    gen = gen.to(dev).to(dtype).eval()          # eval: temporal_state keeps the intermediates; arithmetic unchanged
    nerf = gen.nerf
    tp_d = tp.detach().to(dev).to(dtype).clone().requires_grad_(grad)
    orig_forward = nerf._forward

    def forward_with_planes(*a, **k):           # models/narf.py:187-193 takes the planes from model_input when given
        a = list(a)
        a[8]["tri_plane_feature"] = tp_d
        return orig_forward(*a, **k)

    nerf._forward = forward_with_planes
    draws = {}
    real_multinomial = torch.multinomial

    def recording_multinomial(w, n, replacement=False):
        idx = real_multinomial(w, n, replacement=replacement)
        if force_draws is not None:
            idx = force_draws[0].to(idx.device).reshape(idx.shape)
        draws["idx"] = idx.clone()
        draws["weights"] = w.detach().clone()
        draws["state"] = torch.cuda.get_rng_state(dev) if dev.type == "cuda" else torch.get_rng_state()
        return idx

    torch.manual_seed(seed + 1)
    torch.multinomial = recording_multinomial
    H.FORCED_JITTER = None if force_draws is None else force_draws[1].to(dtype)
    c = lambda k: inp[k].to(dev).to(dtype)
    try:
        with torch.set_grad_enabled(grad):
            color, mask, disparity = gen(c("pose_to_camera"), c("pose_to_world"), c("bone_length"), z.to(dev).to(dtype),
                                         c("inv_intrinsics"), return_disparity=True)
    finally:
        torch.multinomial = real_multinomial
        H.FORCED_JITTER = None
    fw = nerf.buffers_tensors["fine_weights"][:, 0]
    fd = nerf.buffers_tensors["fine_depth"][:, 0]
    n = fw.shape[1]
    # the jitter is the next draw after the multinomial (rendering.py:192-194): replay it from the recorded state
    if dev.type == "cuda":
        keep = torch.cuda.get_rng_state(dev)
        torch.cuda.set_rng_state(draws["state"], dev)
        jitter = torch.empty(B, 1, n, Nf, device=dev).uniform_().reshape(B * n, Nf)
        torch.cuda.set_rng_state(keep, dev)
    else:
        keep = torch.get_rng_state()
        torch.set_rng_state(draws["state"])
        jitter = torch.empty(B, 1, n, Nf).uniform_().reshape(B * n, Nf)
        torch.set_rng_state(keep)
    if force_draws is not None:
        jitter = force_draws[1].to(dev).reshape(B * n, Nf)
    ts = nerf.temporal_state
    out = dict(color=color.detach(), mask=mask.detach(), disparity=disparity.detach(), fine_weights=fw.detach(),
               fine_depth=fd.detach(), idx=draws["idx"], jitter=jitter, coarse_weights_blurred=draws["weights"],
               n_rays=n, P=P, decoder=decoder_state(gen), inputs=(inp, tp, z))
    if keep_masks and "in_cube" in ts:
        out["valid_parts"] = (ts["in_cube"] & (ts["canonical_fine_points"].abs() < 1).all(dim=2)).detach()
        out["fine_points"] = ts["fine_points"].detach()         # (B,3,n*Nf), scaled camera frame
    if grad:
        lw = loss_weights(B, size * size, n, Nf, loss_seed)
        loss = linear_loss(color, mask, disparity, nerf.buffers_tensors["fine_weights"], lw)
        loss.backward()
        out["loss"] = loss.detach()
        out["loss_weights"] = lw
        out["d_tri_plane"] = tp_d.grad.detach()
        grads = {}
        for i, layer in enumerate(gen.nerf.mlp.layers):
            grads[f"nerf.mlp.layers.{i}.conv.weight"] = layer.conv.weight.grad.detach()
            grads[f"nerf.mlp.layers.{i}.conv.modulation.weight"] = layer.conv.modulation.weight.grad.detach()
            grads[f"nerf.mlp.layers.{i}.conv.modulation.bias"] = layer.conv.modulation.bias.grad.detach()
            grads[f"nerf.mlp.layers.{i}.bias"] = layer.bias.grad.detach()
        out["param_grads"] = grads
    return out
