"""Rebuild oracle inputs from a committed golden fixture (tests/golden/case_*.npz)."""
import os

import numpy as np
import torch

import enarf_oracle as O
from enarf_gan_b200 import synthetic as S

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["b1_s32", "b2_s24", "b1_s32_center_fixed", "config1_b1_s64"]


def checksum(t):
    t = t.double()
    return np.array([t.sum().item(), t.abs().sum().item(), t.flatten()[::9973].sum().item()])


def load(name):
    g = dict(np.load(os.path.join(GOLDEN_DIR, f"case_{name}.npz")))
    B, size, Nc, Nf, seed, P = [int(v) for v in g["meta"]]
    origin = str(g["origin"])
    inp = S.make_inputs(B, size, seed=seed)
    tp = S.make_triplane(B, num_part=P, seed=seed)
    z = S.make_latent(B, seed=seed)
    # generator drift check: the regenerated inputs must be the ones the reference saw
    np.testing.assert_allclose(checksum(inp["pose_to_camera"]), g["cs_pose"], rtol=1e-12)
    np.testing.assert_allclose(checksum(tp), g["cs_triplane"], rtol=1e-12)
    np.testing.assert_allclose(checksum(z), g["cs_z"], rtol=1e-12)
    layers = [{"mod_weight": torch.from_numpy(g[f"mlp{i}_mod_weight"]),
               "mod_bias": torch.from_numpy(g[f"mlp{i}_mod_bias"]),
               "weight": torch.from_numpy(g[f"mlp{i}_weight"]),
               "bias": torch.from_numpy(g[f"mlp{i}_bias"])} for i in range(3)]
    z_rend = z[:, 512:768]          # generator.py:79 split [2*256 | 256]
    cfg = dict(B=B, size=size, Nc=Nc, Nf=Nf, seed=seed, P=P, origin=origin)
    return g, cfg, inp, tp, layers, z_rend


def scene(cfg, inp, tp, layers, z_rend, dtype=torch.float32):
    return O.Scene(inp["pose_to_camera"], inp["bone_length"], inp["inv_intrinsics"], tp, layers, z_rend,
                   S.canonical_pose(), S.SMPL_PARENTS, origin_location=cfg["origin"], coordinate_scale=3.0,
                   dtype=dtype)


# ---- DSO ("denarf") fixtures: tests/golden/make_golden_dso.py ------------------------------------------
DSO_CASES = ["dso_b1_s64", "dso_b2_s48"]
DSO_SUBSAMPLE = 4999


def load_dso(name):
    """-> golden dict, cfg, inputs, canonical tri-plane (1,C,256,256), flow (B,6,256,256), decoder layers."""
    g = dict(np.load(os.path.join(GOLDEN_DIR, f"case_{name}.npz")))
    B, size, ray_bs, Nc, Nf, seed, P = [int(v) for v in g["meta"]]
    inp = S.make_inputs(B, size, seed=seed)
    canon = S.make_triplane(1, num_part=P, seed=seed)
    flow = S.make_flow(B, seed=seed)
    np.testing.assert_allclose(checksum(inp["pose_to_camera"]), g["cs_pose"], rtol=1e-12)
    np.testing.assert_allclose(checksum(canon), g["cs_canon"], rtol=1e-12)
    np.testing.assert_allclose(checksum(flow), g["cs_flow"], rtol=1e-9)
    layers = [{"mod_weight": torch.from_numpy(g[f"mlp{i}_mod_weight"]),
               "mod_bias": torch.from_numpy(g[f"mlp{i}_mod_bias"]),
               "weight": torch.from_numpy(g[f"mlp{i}_weight"]),
               "bias": torch.from_numpy(g[f"mlp{i}_bias"])} for i in range(3)]
    cfg = dict(B=B, size=size, ray_bs=ray_bs, Nc=Nc, Nf=Nf, seed=seed, P=P, origin="center_fixed")
    return g, cfg, inp, canon, flow, layers


def dso_latent(frame_time, pose_to_camera):
    """DSONARFGenerator.get_latents restated (models/generator.py:184-217): 10-frequency encoding of the frame time
    (cos then sin) followed by every joint rotation relative to the root's."""
    t = torch.as_tensor(frame_time)
    x = t[:, None] * 2 ** torch.arange(10) * np.pi
    rel = torch.matmul(pose_to_camera[:, :1, :3, :3].permute(0, 1, 3, 2), pose_to_camera[:, 1:, :3, :3])
    return torch.cat([torch.cos(x), torch.sin(x), rel.reshape(rel.shape[0], -1)], dim=1)


def dso_image_coord(ray_idx, size):
    """Pixel centres of the selected rays (ray_sampler.py:30-38) -> (B,3,n)."""
    idx = torch.as_tensor(ray_idx).long()
    x = (idx % size).float() + 0.5
    y = torch.div(idx, size, rounding_mode="floor").float() + 0.5
    return torch.stack([x, y, torch.ones_like(x)], dim=1)
