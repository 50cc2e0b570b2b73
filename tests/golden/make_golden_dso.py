"""Golden fixtures of the DSO ("denarf") path, made by running the UNMODIFIED reference on CPU (authoring
container only):

    python tests/golden/make_golden_dso.py

* ``TriPlaneNARF``'s deformation-field producer (models/narf.py:39-58): the closure ``tri_plane_gen`` is called as the
  reference defines it; only the StyleGAN2-ADA flow head (``prepare_stylegan2``, out of the hot path's scope) is
  replaced by a module returning a seeded synthetic flow.  A warped tri-plane is 43 MB per sample, so the fixture keeps
  a strided subsample plus checksums of it.
* ``DSONARFGenerator.forward`` (models/generator.py:219-258): mask-based ray selection, latents from time / pose,
  sparse render.  The ray ids the reference drew and its fine-pass RNG draws are stored.
Inputs are regenerated from seeds by ``enarf_gan_b200.synthetic`` (checksummed against drift).
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
from make_golden import checksum, decoder_params  # noqa: E402

CASES = {
    # name: (B, image size, rays per image, Nc, Nf, seed)
    "dso_b1_s64": (1, 64, 768, 48, 64, 1401),
    "dso_b2_s48": (2, 48, 320, 48, 48, 1402),
}
SUBSAMPLE = 4999        # stride of the stored subsample of the warped tri-plane


class _Flow(torch.nn.Module):
    """Stand-in for the StyleGAN2-ADA flow head: returns the flow it was given."""
    flow = None

    def forward(self, z, *a, **k):
        return self.flow


def frame_times(B, seed):
    return torch.from_numpy(np.random.RandomState(seed + 3).uniform(0, 1, size=B).astype(np.float32))


def build(size, ray_bs, Nc, Nf, seed):
    H.install()
    from models.generator import DSONARFGenerator
    from models.narf import TriPlaneNARF
    TriPlaneNARF.prepare_stylegan2 = lambda self, c: _Flow()
    cfg = H.gen_config(Nc=Nc, Nf=Nf, origin_location="center_fixed", constant_triplane=False)
    cfg.nerf_params.deformation_field = True
    cfg.nerf_params.no_ray_direction = True
    cfg.nerf_params.time_conditional = True
    cfg.nerf_params.pose_conditional = True
    cfg.ray_batchsize = ray_bs
    torch.manual_seed(seed)
    gen = DSONARFGenerator(cfg, size, num_bone=24, parent_id=S.SMPL_PARENTS, num_bone_param=23)
    gen.register_canonical_pose(S.canonical_pose())
    return gen


def run_case(name, B, size, ray_bs, Nc, Nf, seed):
    gen = build(size, ray_bs, Nc, Nf, seed)
    nerf = gen.nerf
    P = nerf.num_bone
    inp = S.make_inputs(B, size, seed=seed)
    canon = S.make_triplane(1, num_part=P, seed=seed)
    flow = S.make_flow(B, seed=seed)
    fg = S.make_foreground_mask(B, size, seed=seed)
    t = frame_times(B, seed)
    with torch.no_grad():
        nerf.tri_plane.copy_(canon)
        for layer in nerf.mlp.layers:
            layer.bias.normal_(0, 0.3)
    nerf.flow_generator.flow = flow
    gen.eval()

    with torch.no_grad():
        warped = nerf.tri_plane_gen(torch.zeros(B, 1))          # the reference's closure, narf.py:43-57
    flat = warped.reshape(-1)

    draws = {}
    real_multinomial = torch.multinomial

    def recording_multinomial(w, n, replacement=False):
        idx = real_multinomial(w, n, replacement=replacement)
        draws["idx"] = idx.clone()
        draws["state"] = torch.get_rng_state()
        return idx

    torch.manual_seed(seed + 1)
    torch.multinomial = recording_multinomial
    try:
        with torch.no_grad():
            color, mask, ray_idx = gen(inp["pose_to_camera"], None, fg, t, inp["bone_length"], inp["inv_intrinsics"])
    finally:
        torch.multinomial = real_multinomial
    n = nerf.buffers_tensors["fine_weights"].shape[2]
    torch.set_rng_state(draws["state"])
    jitter = torch.empty(B, 1, n, Nf).uniform_().reshape(B * n, Nf)

    g = dict(decoder_params(gen))
    g.update(meta=np.array([B, size, ray_bs, Nc, Nf, seed, P]),
             cs_pose=checksum(inp["pose_to_camera"]), cs_canon=checksum(canon), cs_flow=checksum(flow),
             warped_sub=flat[::SUBSAMPLE].numpy(), warped_cs=checksum(warped),
             frame_time=t.numpy(), ray_idx=ray_idx.numpy().astype(np.int32),
             color=color.numpy(), mask=mask.numpy(), n_rays=np.array(n),
             idx=draws["idx"].numpy().astype(np.int16), jitter=jitter.numpy(),
             fine_weights=nerf.buffers_tensors["fine_weights"][:, 0].numpy())
    path = os.path.join(HERE, f"case_{name}.npz")
    np.savez_compressed(path, **g)
    print(name, "valid rays", n, "mask sum", float(mask.sum()), os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    only = sys.argv[1:]
    for name, cfg in CASES.items():
        if not only or name in only:
            run_case(name, *cfg)
