"""Deterministic synthetic inputs for the renderer hot path.

Nothing here touches the reference or a dataset: a fixed SMPL-like 24-joint skeleton, random
axis-angle poses run through forward kinematics, a random camera, pinhole intrinsics and random
tri-planes.  The same generator feeds the oracle, the parity tests, ``smoke()`` and ``bench.py``.

Reference shapes being reproduced (not code): ``pose_to_camera (B,24,4,4)``, ``bone_length (B,23,1)``,
``inv_intrinsics (B,3,3)`` as consumed by ``models/generator.py:56-58``; the canonical pose
``(24,4,4)`` consumed by ``models/narf.py:84-120``; the SMPL parent table of ``ENARF_GAN_demo.py:27-28``.
"""
import math

import numpy as np
import torch

SMPL_PARENTS = np.array([-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13,
                         14, 16, 17, 18, 19, 20, 21])

# Rest (T-pose) joints in metres, hips centred.  x: subject's left, y: up, z: forward.
REST_JOINTS = np.array([
    [0.00, 0.00, 0.00],    # 0 pelvis
    [0.07, -0.09, 0.00],   # 1 l_hip
    [-0.07, -0.09, 0.00],  # 2 r_hip
    [0.00, 0.11, -0.02],   # 3 spine1
    [0.10, -0.47, 0.00],   # 4 l_knee
    [-0.10, -0.47, 0.00],  # 5 r_knee
    [0.00, 0.25, 0.00],    # 6 spine2
    [0.09, -0.87, -0.03],  # 7 l_ankle
    [-0.09, -0.87, -0.03],  # 8 r_ankle
    [0.00, 0.30, 0.02],    # 9 spine3
    [0.11, -0.93, 0.09],   # 10 l_foot
    [-0.11, -0.93, 0.09],  # 11 r_foot
    [0.00, 0.51, -0.01],   # 12 neck
    [0.08, 0.42, 0.00],    # 13 l_collar
    [-0.08, 0.42, 0.00],   # 14 r_collar
    [0.00, 0.60, 0.04],    # 15 head
    [0.19, 0.45, -0.01],   # 16 l_shoulder
    [-0.19, 0.45, -0.01],  # 17 r_shoulder
    [0.45, 0.44, -0.03],   # 18 l_elbow
    [-0.45, 0.44, -0.03],  # 19 r_elbow
    [0.70, 0.45, -0.03],   # 20 l_wrist
    [-0.70, 0.45, -0.03],  # 21 r_wrist
    [0.79, 0.44, -0.04],   # 22 l_hand
    [-0.79, 0.44, -0.04],  # 23 r_hand
], dtype=np.float64)


def canonical_pose() -> np.ndarray:
    """(24,4,4) canonical joint frames: identity rotations at the rest joints."""
    pose = np.tile(np.eye(4), (24, 1, 1))
    pose[:, :3, 3] = REST_JOINTS
    return pose.astype(np.float32)


def _rodrigues(aa: np.ndarray) -> np.ndarray:
    """(...,3) axis-angle -> (...,3,3) rotation matrices (float64)."""
    theta = np.linalg.norm(aa, axis=-1, keepdims=True)
    k = aa / np.maximum(theta, 1e-12)
    kx, ky, kz = k[..., 0], k[..., 1], k[..., 2]
    zero = np.zeros_like(kx)
    K = np.stack([zero, -kz, ky, kz, zero, -kx, -ky, kx, zero], axis=-1).reshape(aa.shape[:-1] + (3, 3))
    s = np.sin(theta)[..., None]
    c = np.cos(theta)[..., None]
    return np.eye(3) + s * K + (1 - c) * (K @ K)


def forward_kinematics(aa: np.ndarray) -> np.ndarray:
    """(B,24,3) per-joint axis-angle -> (B,24,4,4) joint-to-world frames."""
    B = aa.shape[0]
    R_local = _rodrigues(aa)
    out = np.zeros((B, 24, 4, 4))
    out[:, :, 3, 3] = 1
    for j in range(24):
        p = SMPL_PARENTS[j]
        if p < 0:
            out[:, j, :3, :3] = R_local[:, j]
            out[:, j, :3, 3] = REST_JOINTS[j]
        else:
            off = REST_JOINTS[j] - REST_JOINTS[p]
            out[:, j, :3, :3] = out[:, p, :3, :3] @ R_local[:, j]
            out[:, j, :3, 3] = out[:, p, :3, 3] + np.einsum("bij,j->bi", out[:, p, :3, :3], off)
    return out


def make_inputs(B: int, size: int, seed: int = 1234, pose_sigma: float = 0.35, span: float = 0.7):
    """Synthetic pose / camera batch.

    Returns a dict of float32 torch CPU tensors: ``pose_to_camera (B,24,4,4)``,
    ``pose_to_world (B,24,4,4)``, ``bone_length (B,23,1)``, ``intrinsics (B,3,3)``,
    ``inv_intrinsics (B,3,3)``.  The camera looks down +z; the body is rotated by a random yaw and
    pushed to a root depth in [2.5, 4] m; the focal length makes a 1.7 m body span ``span`` of the
    image (f = span * size * z_root / 1.7).
    """
    rng = np.random.RandomState(seed)
    aa = rng.normal(0.0, pose_sigma, size=(B, 24, 3))
    p2w = forward_kinematics(aa)
    yaw = rng.uniform(0, 2 * math.pi, size=B)
    depth = rng.uniform(2.5, 4.0, size=B)
    cam = np.tile(np.eye(4), (B, 1, 1))
    c, s = np.cos(yaw), np.sin(yaw)
    # yaw about the vertical axis, then flip y so image-y points down (pinhole convention)
    Ry = np.stack([c, 0 * c, s, 0 * c, 1 + 0 * c, 0 * c, -s, 0 * c, c], axis=-1).reshape(B, 3, 3)
    flip = np.diag([1.0, -1.0, 1.0])
    cam[:, :3, :3] = flip @ Ry
    cam[:, :3, 3] = np.stack([rng.uniform(-0.1, 0.1, B), rng.uniform(-0.1, 0.1, B) + 0.15, depth], axis=-1)
    p2c = cam[:, None] @ p2w
    joints = p2w[:, :, :3, 3]
    bone_length = np.linalg.norm(joints[:, 1:] - joints[:, SMPL_PARENTS[1:]], axis=-1)[..., None]
    K = np.zeros((B, 3, 3))
    f = span * size * depth / 1.7
    K[:, 0, 0] = f
    K[:, 1, 1] = f
    K[:, 0, 2] = size / 2
    K[:, 1, 2] = size / 2
    K[:, 2, 2] = 1
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return {
        "pose_to_camera": t(p2c), "pose_to_world": t(p2w), "bone_length": t(bone_length),
        "intrinsics": t(K), "inv_intrinsics": t(np.linalg.inv(K)),
    }


def make_triplane(B: int, num_part: int = 24, seed: int = 1234, res: int = 256, feat_dim: int = 32,
                  feat_sigma: float = 0.5, prob_sigma: float = 2.0, smooth: int = 0) -> torch.Tensor:
    """Random tri-plane ``(B,(feat_dim+num_part)*3,res,res)`` fp32 in the reference's NCHW channel
    order (feature ``32*k+c``, probability ``96+3*p+k``; ``models/narf.py:239``, ``sampling.py:28``).
    ``smooth`` > 0 box-blurs the planes that many times (keeps gradients well conditioned)."""
    g = torch.Generator().manual_seed(seed)
    ch = (feat_dim + num_part) * 3
    tp = torch.randn(B, ch, res, res, generator=g)
    tp[:, :feat_dim * 3] *= feat_sigma
    tp[:, feat_dim * 3:] *= prob_sigma
    for _ in range(smooth):
        tp = torch.nn.functional.avg_pool2d(tp, 3, 1, 1)
    return tp.contiguous()


def make_latent(B: int, z_dim: int = 256, n_chunks: int = 3, seed: int = 1234) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed + 77)
    return torch.randn(B, z_dim * n_chunks, generator=g)


def make_flow(B: int, seed: int = 1234, res: int = 256, coarse: int = 8, amplitude: float = 4.0) -> torch.Tensor:
    """Smooth synthetic deformation field ``(B,6,res,res)`` in texels, the shape the DSO flow head emits
    (``models/narf.py:45-46``: channel 2k / 2k+1 = x / y offset of plane k).  Low-resolution Gaussian noise
    upsampled bicubically: offsets of a few texels, large enough near the borders to leave the plane."""
    g = torch.Generator().manual_seed(seed + 991)
    low = torch.randn(B, 6, coarse, coarse, generator=g) * amplitude
    return torch.nn.functional.interpolate(low, size=(res, res), mode="bicubic", align_corners=False).contiguous()


def make_foreground_mask(B: int, size: int, seed: int = 1234) -> torch.Tensor:
    """``(B,size,size)`` 0/1 foreground mask: a centred ellipse per sample with a seeded radius."""
    rng = np.random.RandomState(seed + 5)
    yy, xx = np.meshgrid(np.arange(size) + 0.5, np.arange(size) + 0.5, indexing="ij")
    out = np.zeros((B, size, size), dtype=np.float32)
    for b in range(B):
        rx, ry = rng.uniform(0.15, 0.3) * size, rng.uniform(0.3, 0.45) * size
        out[b] = (((xx - size / 2) / rx) ** 2 + ((yy - size / 2) / ry) ** 2 <= 1).astype(np.float32)
    return torch.from_numpy(out)
