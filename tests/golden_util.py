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
