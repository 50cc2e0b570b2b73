"""Small host-side (torch) pieces of the path that stay outside the kernels because they are O(B*P) work:
joint frames -> part frames, canonical part frames, per-sample modulated decoder weights.
Mirrors libraries/NARF/pose_utils.py:129-148, models/narf.py:84-120,165 and
libraries/custom_stylegan2/net.py:161-174,229-247 of the reference."""
import math

import numpy as np
import torch


_PARENT_INDEX = {}   # (parent ids, device) -> index tensor; built once so the render path has no host sync


def _parent_index(parent_id, device):
    key = (tuple(int(p) for p in np.asarray(parent_id)[1:]), str(device))
    if key not in _PARENT_INDEX:
        _PARENT_INDEX[key] = torch.tensor(key[0], dtype=torch.long, device=device)
    return _PARENT_INDEX[key]


def part_frames(pose_to_camera, bone_length, origin_location, parent_id):
    """(B,J,4,4),(B,J-1,1) -> part frames (B,P,4,4), bone length (B,P,1)   [transform_pose]."""
    par = _parent_index(parent_id, pose_to_camera.device)
    mid = (pose_to_camera[:, 1:, :, 3:] + pose_to_camera[:, par, :, 3:]) / 2
    if origin_location == "center":
        return torch.cat([pose_to_camera[:, 1:, :, :3], mid], dim=-1), bone_length
    if origin_location == "center_fixed":
        return torch.cat([pose_to_camera[:, par, :, :3], mid], dim=-1), bone_length
    if origin_location == "center+head":
        body = torch.cat([pose_to_camera[:, par, :, :3], mid], dim=-1)
        frames = torch.cat([body, pose_to_camera[:, 15:16]], dim=1)
        return frames, torch.cat([bone_length, torch.ones_like(bone_length[:, :1])], dim=1)
    raise ValueError(f"unsupported origin_location {origin_location!r}")


def canonical_part_frames(pose, parent_id, origin_location):
    """numpy (J,4,4) canonical joints -> (canonical part frames (P,4,4), canonical bone length (P,)) float32
    [register_canonical_pose]."""
    pose = np.asarray(pose, dtype=np.float64)
    par = np.asarray(parent_id)[1:]
    joints = pose[:, :3, 3]
    length = np.linalg.norm(joints[1:] - joints[par], axis=1)
    mid = (pose[1:, :, 3:] + pose[par, :, 3:]) / 2
    if origin_location == "center":
        frames = np.concatenate([pose[1:, :, :3], mid], axis=-1)
    elif origin_location == "center_fixed":
        frames = np.concatenate([pose[par, :, :3], mid], axis=-1)
    elif origin_location == "center+head":
        frames = np.concatenate([np.concatenate([pose[par, :, :3], mid], axis=-1), pose[15:16]], axis=0)
        length = np.concatenate([length, np.ones(1)])
    else:
        raise ValueError(f"unsupported origin_location {origin_location!r}")
    return torch.tensor(frames, dtype=torch.float32), torch.tensor(length, dtype=torch.float32)


def scaled_pose(frames, coordinate_scale):
    out = frames.clone()
    out[:, :, :3, 3] = out[:, :, :3, 3] * coordinate_scale
    return out


def modulated_weight(weight, mod_weight, mod_bias, style):
    """weight (out,in), EqualLinear(style_dim -> in, bias_init 1) parameters, style (B,style_dim)
    -> per-sample normalised weight (B,out,in)."""
    s = torch.nn.functional.linear(style, mod_weight * (1.0 / math.sqrt(mod_weight.shape[1])), mod_bias)
    w = (1.0 / math.sqrt(weight.shape[1])) * weight[None] * s[:, None, :]
    return torch.nn.functional.normalize(w, dim=-1)


def bone_length_encoding(bone_length, num_frequency, num_bone):
    """The conditioning vector the reference hands to its StyleGAN2-ADA tri-plane / flow generator:
    ``multi_part_positional_encoding(bone_length, L, num_bone)[:, :, 0]`` (libraries/NeRF/utils.py:46-71, called at
    models/narf.py:286-289).  bone_length (B,num_bone,1) -> (B, num_bone*2L): per part
    [sin(l 2^0 pi) .. sin(l 2^(L-1) pi), cos(l 2^0 pi) .. cos(l 2^(L-1) pi)], zeroed for parts with |l| > 1."""
    B = bone_length.shape[0]
    l = bone_length.reshape(B, num_bone, 1)
    ang = l * 2 ** torch.arange(num_frequency, device=l.device) * math.pi           # (B,P,L): (l * 2^i) * pi
    enc = torch.cat([torch.sin(ang), torch.cos(ang)], dim=2)
    enc = enc * (1 - (l.abs() > 1).to(enc.dtype))
    return enc.reshape(B, num_bone * 2 * num_frequency)


def random_scene(inp, tri_plane, seed=0, device="cuda", origin_location="center+head", coordinate_scale=3.0,
                 plane_dtype=torch.float32):
    """A SceneBuffers with random-init decoder weights for synthetic benchmarks / smoke runs."""
    from . import ops, synthetic as S
    g = torch.Generator().manual_seed(seed)
    B = tri_plane.shape[0]
    frames, bl = part_frames(inp["pose_to_camera"], inp["bone_length"], origin_location, S.SMPL_PARENTS)
    pose = scaled_pose(frames, coordinate_scale)
    cframes, clen = canonical_part_frames(S.canonical_pose(), S.SMPL_PARENTS, origin_location)
    cscale = (clen[None, :, None] / bl / coordinate_scale)[:, :, 0]
    z = torch.randn(B, 256, generator=g)
    decoder = []
    for cin, cout in ((32, 64), (64, 64), (64, 4)):
        w = modulated_weight(torch.randn(cout, cin, generator=g), torch.randn(cin, 256, generator=g),
                             torch.ones(cin), z)
        decoder.append((w.to(device), (0.3 * torch.randn(cout, generator=g)).to(device)))
    return ops.SceneBuffers(pose.to(device), cscale.to(device), cframes.to(device), tri_plane.to(device), decoder,
                            plane_dtype)
