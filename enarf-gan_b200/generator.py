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
        out = self.nerf(batchsize, homo_img, pose_to_camera, inv_intrinsics, z_nerf, z_rend, bone_length,
                        Nc=_cfg(nparams, "Nc"), Nf=_cfg(nparams, "Nf"), return_intermediate=return_intermediate,
                        truncation_psi=truncation_psi, return_disparity=return_disparity)
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
        rendered_color = fg_color + (1 - fg_mask[:, None]) * bg_color
        if return_intermediate:
            fine_points, fine_density = out[-1]
            return rendered_color, fg_mask, fine_points, fine_density
        if return_disparity:
            return rendered_color, fg_mask, out[2] * _cfg(nparams, "coordinate_scale")
        if return_bg:
            return fg_color, fg_mask, bg_color
        return rendered_color, fg_mask, fine_weights, fine_depth
