"""Generate the golden fixtures by running the UNMODIFIED reference on CPU (authoring container only).

    python tests/golden/make_golden.py

Writes ``tests/golden/case_*.npz``.  Inputs are regenerated from seeds by
``enarf_gan_b200.synthetic`` (a 44 MB tri-plane cannot be committed); a checksum of every
regenerated input is stored so generator drift is detected.  The random-init decoder parameters,
the reference's RNG draws for the fine pass and the reference's outputs / intermediates are stored.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_harness as H  # noqa: E402
from enarf_gan_b200 import synthetic as S  # noqa: E402

CASES = {
    # name: (B, image size, Nc, Nf, seed, origin_location, keep intermediates)
    "b1_s32": (1, 32, 48, 48, 1301, "center+head", True),
    "b2_s24": (2, 24, 48, 48, 1302, "center+head", True),
    "b1_s32_center_fixed": (1, 32, 48, 64, 1303, "center_fixed", True),
    "config1_b1_s64": (1, 64, 48, 48, 1235, "center+head", False),
}


def decoder_params(gen):
    out = {}
    for i, layer in enumerate(gen.nerf.mlp.layers):
        out[f"mlp{i}_mod_weight"] = layer.conv.modulation.weight.detach().numpy()
        out[f"mlp{i}_mod_bias"] = layer.conv.modulation.bias.detach().numpy()
        out[f"mlp{i}_weight"] = layer.conv.weight.detach().numpy()[0, :, :, 0]
        out[f"mlp{i}_bias"] = layer.bias.detach().numpy()[0, :, 0]
    return out


def checksum(t):
    t = t.double()
    return np.array([t.sum().item(), t.abs().sum().item(), t.flatten()[::9973].sum().item()])


def run_case(name, B, size, Nc, Nf, seed, origin, keep):
    gen = H.build_generator(size, S.canonical_pose(), S.SMPL_PARENTS, seed=seed, Nc=Nc, Nf=Nf,
                            origin_location=origin)
    P = gen.nerf.num_bone
    inp = S.make_inputs(B, size, seed=seed)
    tp = S.make_triplane(B, num_part=P, seed=seed)
    z = S.make_latent(B, seed=seed)
    with torch.no_grad():
        for layer in gen.nerf.mlp.layers:       # non-trivial biases so the bias path is exercised
            layer.bias.normal_(0, 0.3)
    gen.eval()
    nerf = gen.nerf

    # the reference takes the tri-plane from model_input["tri_plane_feature"] when given
    # (models/narf.py:187-193); constant_triplane=True only supports one shared plane.
    orig_forward = nerf._forward

    def forward_with_planes(*a, **k):
        a = list(a)
        a[8]["tri_plane_feature"] = tp
        return orig_forward(*a, **k)

    nerf._forward = forward_with_planes

    draws = {}
    real_multinomial = torch.multinomial

    def recording_multinomial(w, n, replacement=False):
        idx = real_multinomial(w, n, replacement=replacement)
        draws["idx"] = idx.clone()
        draws["weights"] = w.clone()
        draws["state"] = torch.get_rng_state()
        return idx

    torch.manual_seed(seed + 1)
    torch.multinomial = recording_multinomial
    try:
        with torch.no_grad():
            color, mask, disparity = gen(inp["pose_to_camera"], inp["pose_to_world"], inp["bone_length"], z,
                                         inp["inv_intrinsics"], return_disparity=True)
    finally:
        torch.multinomial = real_multinomial
    fw = nerf.buffers_tensors["fine_weights"]
    fd = nerf.buffers_tensors["fine_depth"]
    n = fw.shape[2]
    torch.set_rng_state(draws["state"])
    jitter = torch.empty(B, 1, n, Nf).uniform_().reshape(B * n, Nf)
    ts = nerf.temporal_state

    g = dict(decoder_params(gen))
    g.update(meta=np.array([B, size, Nc, Nf, seed, P]), origin=np.array(origin),
             cs_pose=checksum(inp["pose_to_camera"]), cs_triplane=checksum(tp), cs_z=checksum(z),
             color=color.numpy(), mask=mask.numpy(), disparity=disparity.numpy(),
             n_rays=np.array(n))
    # disparity above already carries the generator's x coordinate_scale (generator.py:114)
    if keep:
        valid = ts["in_cube"] & (ts["canonical_fine_points"].abs() < 1).all(dim=2)
        g.update(fine_weights=fw[:, 0].numpy(), fine_depth=fd[:, 0].numpy(),
                 idx=draws["idx"].numpy().astype(np.int16), jitter=jitter.numpy(),
                 coarse_weights_blurred=draws["weights"].numpy(),
                 coarse_density=ts["coarse_density"][:, 0].numpy(),
                 coarse_depth=ts["coarse_depth"][:, 0].numpy(),
                 fine_points=ts["fine_points"].numpy(),
                 fine_density=ts["fine_density"][:, 0].numpy(),
                 fine_valid_bits=np.packbits(valid.numpy()),
                 fine_valid_shape=np.array(valid.shape),
                 fine_part_weight_sum=ts["weight"].mul(valid).sum(dim=1).numpy())
    else:
        g.update(fine_weights_sum=fw.sum(dim=-1)[:, 0].numpy(), fine_depth_first=fd[:, 0, :, 0].numpy(),
                 idx=draws["idx"].numpy().astype(np.int16), jitter=jitter.numpy())
    path = os.path.join(HERE, f"case_{name}.npz")
    np.savez_compressed(path, **g)
    print(name, "rays", n, "mask sum", float(mask.sum()), os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    only = sys.argv[1:]
    for name, cfg in CASES.items():
        if not only or name in only:
            run_case(name, *cfg)
