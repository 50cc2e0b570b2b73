"""Public generator API, mirroring ``models/generator.py:14-140`` of the reference (``TriNARFGenerator``).

Signatures, latent split, return tuples and the ``nerf.buffers_tensors`` side effects are the reference's;
the render itself is ``TriPlaneNARF`` over the fused kernels.  The StyleGAN2 networks that the reference
instantiates around the renderer (tri-plane generator, 2-D background generator) are out of this path's scope
(SURVEY.md section 2, rows 9/12/13): they are accepted as injected ``nn.Module``s.
"""
from typing import Optional

import numpy as np
import torch
from torch import nn

from .narf import TriPlaneNARF, _cfg
from .rendering import pixel_grid


def whole_image_grid_ray_sampler(render_size: int, patch_size: int, batchsize: int, device="cuda"):
    """(grid (B,p,p,2) in [-1,1], homo_img (B,1,3,p^2)) -- libraries/NeRF/ray_sampler.py:42-67."""
    homo = pixel_grid(render_size, patch_size, batchsize, device)
    grid = (homo[:, :2] / (render_size / 2) - 1).permute(0, 2, 1).reshape(batchsize, patch_size, patch_size, 2)
    return grid, homo[:, None]


class TriNARFGenerator(nn.Module):
    def __init__(self, config, size, num_bone=1, parent_id=None, num_bone_param=None, black_background=False,
                 tri_plane_gen: Optional[nn.Module] = None, background_generator: Optional[nn.Module] = None,
                 plane_dtype=torch.float32):
        super().__init__()
        self.config = config
        self.size = size
        self.num_bone = num_bone
        self.ray_sampler = whole_image_grid_ray_sampler
        self.background_ratio = _cfg(config, "background_ratio")
        self.black_background = black_background
        z_dim = _cfg(config, "z_dim")
        nerf_params = _cfg(config, "nerf_params")
        if _cfg(nerf_params, "origin_location") == "center+head":
            num_bone_param = num_bone
        self.nerf = TriPlaneNARF(nerf_params, z_dim=[z_dim * 2, z_dim], num_bone=num_bone, bone_length=True,
                                 parent=parent_id, num_bone_param=num_bone_param, tri_plane_gen=tri_plane_gen,
                                 plane_dtype=plane_dtype)
        if not black_background:
            if background_generator is None:
                raise ValueError("black_background=False needs an injected background_generator module "
                                 "(the StyleGAN2 background network is outside the renderer hot path)")
            self.background_generator = background_generator

    def register_canonical_pose(self, pose: np.ndarray):
        self.nerf.register_canonical_pose(pose)

    def normalized_inv_intrinsics(self, intrinsics: torch.Tensor):
        normalized = torch.cat([intrinsics[:2] / self.size, intrinsics.new_tensor([[0, 0, 1]])], dim=0)
        return torch.linalg.inv(normalized)

    def _split(self, z):
        if not self.black_background:
            d = z.shape[1] // 4
            return torch.split(z, [d * 2, d, d], dim=1)
        d = z.shape[1] // 3
        return tuple(torch.split(z, [d * 2, d], dim=1)) + (None,)

    def forward(self, pose_to_camera, pose_to_world, bone_length, z=None, inv_intrinsics=None,
                return_intermediate=False, truncation_psi=1, black_bg_if_possible=False, return_disparity=False,
                return_bg=False):
        """Same contract as models/generator.py:56-118.  ``pose_to_world`` is accepted and unused, as there."""
        assert self.num_bone == 1 or (bone_length is not None and pose_to_camera is not None)
        batchsize = pose_to_camera.shape[0]
        device = pose_to_camera.device
        _, homo_img = self.ray_sampler(self.size, self.size, batchsize, device)
        z_nerf, z_rend, z_bg = self._split(z)
        if isinstance(inv_intrinsics, np.ndarray):
            inv_intrinsics = torch.from_numpy(inv_intrinsics)
        inv_intrinsics = inv_intrinsics.float().to(device)
        nparams = _cfg(self.config, "nerf_params")
        # a constant (black) background is composited inside the fine pass's epilogue; an image background, or a caller
        # who wants the bare foreground back (return_bg), goes through the torch expression below
        const_bg = (self.black_background or black_bg_if_possible) and not return_bg
        out = self.nerf(batchsize, homo_img, pose_to_camera, inv_intrinsics, z_nerf, z_rend, bone_length,
                        Nc=_cfg(nparams, "Nc"), Nf=_cfg(nparams, "Nf"), return_intermediate=return_intermediate,
                        truncation_psi=truncation_psi, return_disparity=return_disparity,
                        background=-1.0 if const_bg else 0.0)
        fg_color, fg_mask = out[:2]
        fine_weights = self.nerf.buffers_tensors.get("fine_weights")
        fine_depth = self.nerf.buffers_tensors.get("fine_depth")
        fg_color = fg_color.reshape(batchsize, 3, self.size, self.size)
        fg_mask = fg_mask.reshape(batchsize, self.size, self.size)
        if not self.black_background and not black_bg_if_possible:
            n_latent = self.background_generator.n_latent
            bg_color, _ = self.background_generator([z_bg, z_rend], inject_index=n_latent - 4)
        else:
            bg_color = -1
        rendered_color = fg_color if const_bg else fg_color + (1 - fg_mask[:, None]) * bg_color      # generator.py:107
        if return_intermediate:
            fine_points, fine_density = out[-1]
            return rendered_color, fg_mask, fine_points, fine_density
        if return_disparity:
            return rendered_color, fg_mask, out[2] * _cfg(nparams, "coordinate_scale")
        if return_bg:
            return fg_color, fg_mask, bg_color
        return rendered_color, fg_mask, fine_weights, fine_depth


def mask_based_sampler(mask: torch.Tensor, ray_batchsize: int, generator: Optional[torch.Generator] = None):
    """DSO ray selection (libraries/NeRF/ray_sampler.py:7-39): dilate the foreground mask with a 129 x 129 max-pool,
    add U(0,1) noise and keep the ``ray_batchsize`` largest pixels of every image.
    mask (B,h,w) -> (ray_idx (B,ray_bs) int64 flat pixel ids, homo_img (B,1,3,ray_bs) pixel centres)."""
    batchsize, h, w = mask.shape
    pad_size = 64
    if mask.is_cuda:
        # on the device: separable dilation + noise, exact radix select and the pixel centres in two small kernels
        # (csrc/dso.cu).  The selected SET is torch.topk's for the same noise; rays come out in ascending pixel order.
        from . import ops
        noise = torch.empty(batchsize, h, w, device=mask.device).uniform_(generator=generator)
        ray_idx, homo = ops.select_rays(mask, ray_batchsize, noise=noise, radius=pad_size)
        return ray_idx, homo[:, None]
    # CPU tensors (host-side tests): the same computation in torch.  A square max-pool is separable: two 1-D passes give the identical dilation at 2*129 instead of 129^2 taps
    m = torch.nn.functional.max_pool2d(mask.float()[:, None], (pad_size * 2 + 1, 1), stride=1, padding=(pad_size, 0))
    m = torch.nn.functional.max_pool2d(m, (1, pad_size * 2 + 1), stride=1, padding=(0, pad_size))[:, 0]
    m = m.reshape(batchsize, h * w)
    m = m + torch.empty_like(m).uniform_(generator=generator)
    ray_idx = torch.topk(m, ray_batchsize, dim=1, sorted=False)[1]
    x = ray_idx % w
    y = torch.div(ray_idx, w, rounding_mode="floor")
    rays = (torch.stack([x, y], dim=2) + 0.5).permute(0, 2, 1)
    homo_img = torch.cat([rays, torch.ones(batchsize, 1, ray_batchsize, device=mask.device)], dim=1)
    return ray_idx, homo_img.reshape(batchsize, 1, 3, -1)


class DSONARFGenerator(nn.Module):
    """Dynamic-scene-overfitting generator (models/generator.py:143-278) for the tri-plane ("denarf") model:
    time / pose conditioned latents -> deformation-field tri-plane -> sparse render of ``ray_batchsize`` rays per
    image chosen by ``mask_based_sampler``.  ``flow_generator`` is the injected flow head (the reference builds a
    StyleGAN2-ADA network there, outside this path)."""

    def __init__(self, config, size, num_bone=1, parent_id=None, num_bone_param=None,
                 flow_generator: Optional[nn.Module] = None, tri_plane_gen: Optional[nn.Module] = None,
                 plane_dtype=torch.float32):
        super().__init__()
        self.config = config
        self.size = size
        self.num_bone = num_bone
        self.ray_sampler = mask_based_sampler
        if not _cfg(config, "use_triplane"):
            raise NotImplementedError("MLPNARF (use_triplane: False) is outside the tri-plane hot path")
        nerf_params = _cfg(config, "nerf_params")
        self.time_conditional = _cfg(nerf_params, "time_conditional")
        self.pose_conditional = _cfg(nerf_params, "pose_conditional")
        z_dim = (20 if self.time_conditional else 0) + ((num_bone - 1) * 9 if self.pose_conditional else 0)
        if _cfg(nerf_params, "origin_location") == "center+head":
            num_bone_param = num_bone
        view_dependent = not _cfg(nerf_params, "no_ray_direction")
        self.nerf = TriPlaneNARF(nerf_params, z_dim=z_dim, num_bone=num_bone, bone_length=True, parent=parent_id,
                                 num_bone_param=num_bone_param, view_dependent=view_dependent,
                                 tri_plane_gen=tri_plane_gen, plane_dtype=plane_dtype, flow_generator=flow_generator)

    def register_canonical_pose(self, pose: np.ndarray):
        self.nerf.register_canonical_pose(pose)

    @staticmethod
    def positional_encoding(x: torch.Tensor, num_frequency: int) -> torch.Tensor:
        """(B,) -> (B, 2*num_frequency): cos then sin of x * 2^i * pi."""
        x = x[:, None] * 2 ** torch.arange(num_frequency, device=x.device) * np.pi
        return torch.cat([torch.cos(x), torch.sin(x)], dim=1)

    @staticmethod
    def pose_encoding(pose: torch.Tensor):
        """(B,num_bone,4,4) -> (B,(num_bone-1)*9): every joint rotation relative to the root's."""
        R = pose[..., :3, :3]
        relative = R[:, :1].transpose(-1, -2) @ R[:, 1:]            # root^T . joint
        return relative.flatten(1)

    def get_latents(self, frame_time: torch.Tensor, pose_to_camera: torch.Tensor):
        """(z for the tri-plane / flow head, z for the decoder modulation): the same conditioning vector twice
        (models/generator.py:204-217)."""
        parts = ([self.positional_encoding(frame_time, num_frequency=10)] if self.time_conditional else []) \
            + ([self.pose_encoding(pose_to_camera)] if self.pose_conditional else [])
        if not parts:
            raise AssertionError("DSONARFGenerator needs time_conditional or pose_conditional")
        latent = torch.cat(parts, dim=1)
        return latent, latent

    def forward(self, pose_to_camera, camera_pose, mask, frame_time, bone_length, inv_intrinsics,
                background: Optional[float] = None):
        """-> (rendered_color (B,3,ray_bs), rendered_mask (B,ray_bs), ray_idx (B,ray_bs))."""
        assert bone_length is not None and pose_to_camera is not None
        assert isinstance(inv_intrinsics, torch.Tensor)
        batchsize = pose_to_camera.shape[0]
        grid, img_coord = self.ray_sampler(mask, _cfg(self.config, "ray_batchsize"))
        z1, z2 = self.get_latents(frame_time, pose_to_camera)
        nparams = _cfg(self.config, "nerf_params")
        if background is None:
            background = -1
        # composited in the fine pass's epilogue: color + background * (1 - mask), models/generator.py:255-257
        color, rendered_mask = self.nerf(batchsize, img_coord, pose_to_camera, inv_intrinsics, z1, z2, bone_length,
                                         Nc=_cfg(nparams, "Nc"), Nf=_cfg(nparams, "Nf"), return_intermediate=False,
                                         camera_pose=camera_pose, background=float(background))
        return color, rendered_mask, grid

    def render_entire_img(self, pose_to_camera, inv_intrinsics, frame_time, bone_length, camera_pose=None,
                          render_size=128, semantic_map=False, use_normalized_intrinsics=False, no_grad=True,
                          bbox=None):
        z1, z2 = self.get_latents(frame_time, pose_to_camera)
        nparams = _cfg(self.config, "nerf_params")
        return self.nerf.render_entire_img(pose_to_camera, inv_intrinsics, z1, z2, bone_length, camera_pose,
                                           render_size, _cfg(nparams, "Nc"), _cfg(nparams, "Nf"), semantic_map,
                                           use_normalized_intrinsics, no_grad=no_grad, bbox=bbox)


class PhotometricLoss:
    """Sparse photometric loss of the DSO step (libraries/NeRF/loss.py:5-48): gather the ground-truth pixels at the
    sampled ray ids, then MSE or truncated MAE on colour and MSE on the mask."""

    def __init__(self, config):
        self.config = config

    def __call__(self, grid, sparse_color, sparse_mask, color, mask=None):
        batchsize, _, size, _ = color.shape
        kind = _cfg(self.config, "nerf_loss_type")
        if sparse_color.is_cuda and kind in ("mse", "mae"):
            # fused gather + loss + gradients (csrc/dso.cu: enarf_photometric_loss)
            from . import ops
            lc, lm = ops.photometric_loss(grid, sparse_color, sparse_mask if mask is not None else None, color, mask, kind,
                                          _cfg(self.config, "color_coef"), _cfg(self.config, "mask_coef"))
            return lc, (lm if mask is not None else 0)
        sampled_color = torch.gather(color.reshape(batchsize, 3, size * size), 2, grid[:, None].expand(-1, 3, -1))
        sampled_mask = torch.gather(mask.reshape(batchsize, size * size), 1, grid) if mask is not None else None
        kind = _cfg(self.config, "nerf_loss_type")
        if kind == "mse":
            loss_color = torch.nn.functional.mse_loss(sampled_color, sparse_color)
        elif kind == "mae":
            loss_color = torch.clamp_min(torch.abs(sampled_color - sparse_color), 0.01).mean()
        else:
            raise ValueError(kind)
        loss_color = loss_color * _cfg(self.config, "color_coef")
        loss_mask = torch.nn.functional.mse_loss(sampled_mask, sparse_mask) * _cfg(self.config, "mask_coef") \
            if sampled_mask is not None else 0
        return loss_color, loss_mask
