"""Host-side mirror of the reference's radiance-field model for the tri-plane path.

``TriPlaneNARF`` keeps the reference's constructor, method names, parameter/buffer names (so a reference
checkpoint's ``nerf.*`` keys load unchanged) and the ``buffers_tensors`` / ``temporal_state`` side effects
(models/narf.py:17-290, libraries/NARF/base.py:11-83, libraries/NeRF/base.py:11-151); every heavy operation
is delegated to the fused kernels through ``rendering`` / ``ops``.  The decoder layers mirror
``StyledMLP`` -> ``StyledConv(conv_1d=True)`` -> ``ModulatedConv1d`` + ``EqualLinear``
(libraries/NeRF/net.py:10-27, libraries/custom_stylegan2/net.py:128-179,194-254,270-320) but only hold the
parameters and compute the per-sample effective weights; the convolution itself runs inside the kernels.
"""
import math
from typing import Dict, Optional

import numpy as np
import torch
from torch import nn

from . import hostmath, ops, rendering
from ._abi import on_device_of

_UNSUPPORTED = ("selector_mlp", "no_selector", "clamp_mask", "multiply_density_with_triplane_wieght",
                "constant_trimask")


class EqualLinear(nn.Module):
    """Parameters of the reference's EqualLinear (weight ~ N(0,1)/lr_mul, runtime scale 1/sqrt(in))."""

    def __init__(self, in_dim, out_dim, bias_init=0.0):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(out_dim, in_dim))
        self.bias = nn.Parameter(torch.full((out_dim,), float(bias_init)))
        self.scale = 1.0 / math.sqrt(in_dim)

    def forward(self, x):
        return torch.nn.functional.linear(x, self.weight * self.scale, self.bias)


class ModulatedConv1d(nn.Module):
    """Holds ``weight (1,out,in,1)`` and ``modulation``; ``effective_weight(style)`` is what the reference
    feeds to its grouped conv1d: normalize_in(weight * style / sqrt(in))."""

    def __init__(self, in_channel, out_channel, style_dim):
        super().__init__()
        self.in_channel, self.out_channel = in_channel, out_channel
        self.scale = 1.0 / math.sqrt(in_channel)
        self.weight = nn.Parameter(torch.randn(1, out_channel, in_channel, 1))
        self.modulation = EqualLinear(style_dim, in_channel, bias_init=1.0)

    def effective_weight(self, style):
        s = self.modulation(style)                                     # (B,in)
        w = self.scale * self.weight[0, :, :, 0][None] * s[:, None, :]  # (B,out,in)
        return torch.nn.functional.normalize(w, dim=-1)


class _Noise(nn.Module):  # present in the reference's StyledConv even when unused; kept for checkpoint-key parity
    def __init__(self):
        super().__init__()
        # never used by the 1-D decoder (net.py:311: noise only for conv_2d), hence frozen: DDP's reducer would
        # otherwise wait for a gradient that never comes
        self.weight = nn.Parameter(torch.zeros(1), requires_grad=False)


class StyledConv1d(nn.Module):
    def __init__(self, in_channel, out_channel, style_dim):
        super().__init__()
        self.conv = ModulatedConv1d(in_channel, out_channel, style_dim)
        self.bias = nn.Parameter(torch.zeros(1, out_channel, 1))
        self.noise = _Noise()


class StyledMLP(nn.Module):
    def __init__(self, in_dim, hidden_dim, out_dim, style_dim=512, num_layers=3):
        super().__init__()
        dims = [in_dim] + [hidden_dim] * (num_layers - 1) + [out_dim]
        self.layers = nn.ModuleList([StyledConv1d(dims[i], dims[i + 1], style_dim) for i in range(num_layers)])
        self.hidden_dim = hidden_dim

    def effective(self, style):
        """[(W_eff (B,out,in), bias (out,)), ...] -- autograd-tracked."""
        return [(l.conv.effective_weight(style), l.bias.reshape(-1)) for l in self.layers]


def _cfg(config, key, default=None):
    if isinstance(config, dict):
        return config.get(key, default)
    return getattr(config, key, default)


class TriPlaneNARF(nn.Module):
    tri_plane_based = True

    def __init__(self, config, z_dim=256, num_bone=1, bone_length=True, parent=None, num_bone_param=None,
                 view_dependent: bool = False, tri_plane_gen: Optional[nn.Module] = None,
                 plane_dtype=torch.float32, flow_generator: Optional[nn.Module] = None):
        super().__init__()
        assert bone_length and parent is not None
        if view_dependent:
            raise NotImplementedError("view-dependent colour is outside the fused path (SURVEY section 8a, a10)")
        for key in _UNSUPPORTED:
            if _cfg(config, key, False):
                raise NotImplementedError(f"nerf_params.{key} is not supported by the fused renderer")
        self.config = config
        self.origin_location = _cfg(config, "origin_location")
        self.coordinate_scale = _cfg(config, "coordinate_scale", 1)
        if self.origin_location not in ("center", "center_fixed", "center+head"):
            raise ValueError(f"origin_location {self.origin_location!r}")       # NeRF/base.py:23
        self.z_dim, self.z2_dim = (z_dim if isinstance(z_dim, (list, tuple)) else (z_dim, z_dim))
        self.w_dim, self.feat_dim = 512, 32
        self.view_dependent = False
        self.num_bone = num_bone - 1 if self.origin_location in ("center", "center_fixed") else num_bone
        self.num_frequency_for_other = _cfg(config, "num_frequency_for_other", 4)    # NeRF/base.py:29
        self.parent_id = np.asarray(parent)
        self.plane_dtype = plane_dtype
        self.rng_mode = "torch"      # "torch": the reference's multinomial/uniform_ stream; "philox": in-kernel
        self.rng_seed = 0
        # early ray termination of the fine pass (north_star step (e)): 0 = off = the reference's exact semantics;
        # e.g. 1e-6 stops a ray once its transmittance is below that (outputs move by < 1e-6), see rendering.render_rays
        self.ert_eps = float(_cfg(config, "ert_eps", 0.0) or 0.0)
        self.temporal_state = {}
        self.buffers_tensors = {}
        self.deformation_field = False
        if _cfg(config, "constant_triplane", False):
            self.tri_plane = nn.Parameter(torch.zeros(1, 32 * 3 + self.num_bone * 3, 256, 256))
        elif _cfg(config, "deformation_field", False):
            # DSO "denarf" (narf.py:39-58): one canonical tri-plane, warped per sample by a 6-channel flow field.
            # The flow head itself is the StyleGAN2-ADA generator (out of scope): pass it as ``flow_generator``
            # (called like the reference's: flow_generator(z, c, truncation_psi=...) -> (B,6,256,256) with c the
            # positional encoding of the bone lengths, ``encoded_bone_length``) or hand the flow
            # over as model_input["flow"].
            self.deformation_field = True
            self.tri_plane = nn.Parameter(torch.zeros(1, 32 * 3 + self.num_bone * 3, 256, 256))
            if flow_generator is not None:
                self.flow_generator = flow_generator
        elif tri_plane_gen is not None:
            self.tri_plane_gen = tri_plane_gen     # e.g. the StyleGAN2-ADA generator (out of scope here)
        self.mlp = StyledMLP(32, 64, 4, style_dim=self.z2_dim)

    # ---- canonical pose (models/narf.py:84-120) -------------------------------------------------------
    def register_canonical_pose(self, pose: np.ndarray) -> None:
        pose = np.asarray(pose)
        par = self.parent_id[1:]
        self.register_buffer("canonical_joints", torch.tensor(pose[1:, :3, 3], dtype=torch.float32))
        self.register_buffer("canonical_parent_joints", torch.tensor(pose[par, :3, 3], dtype=torch.float32))
        frames, length = hostmath.canonical_part_frames(pose, self.parent_id, self.origin_location)
        self.register_buffer("canonical_bone_length", length)
        self.register_buffer("canonical_pose", frames)

    def transform_pose(self, pose_to_camera, bone_length):
        return hostmath.part_frames(pose_to_camera, bone_length, self.origin_location, self.parent_id)

    # ---- tri-plane ------------------------------------------------------------------------------------
    def encoded_bone_length(self, bone_length):
        """c of the StyleGAN2-ADA generators (c_dim = 2 L num_bone, narf.py:79-82,286-289)."""
        return hostmath.bone_length_encoding(bone_length, self.num_frequency_for_other, self.num_bone)

    def compute_flow(self, z, bone_length, truncation_psi=1):
        if not hasattr(self, "flow_generator"):
            raise RuntimeError("deformation_field needs a flow_generator module or model_input['flow']")
        return self.flow_generator(z, self.encoded_bone_length(bone_length), truncation_psi=truncation_psi)

    def warped_planes(self, flow):
        """flow (B,6,H,W) -> (feature planes (B,3,H,W,32) channels-last in ``plane_dtype``, probability planes
        (1,3P,H,W)): the deformation-field producer (narf.py:43-57) emitting the renderer's own layout."""
        canon = ops.RepackCanonical.apply(self.tri_plane)
        return ops.warp_planes(canon, flow, self.plane_dtype), self.tri_plane[:, 32 * 3:]

    def compute_tri_plane_feature(self, z, bone_length, truncation_psi=1):
        if self.deformation_field:
            # the reference's NCHW result (narf.py:55-57), for callers that want the tensor itself; the render path
            # (build_scene) hands the channels-last planes over without building this
            B = z.shape[0]
            feat = ops.warp_planes(ops.RepackCanonical.apply(self.tri_plane),
                                   self.compute_flow(z, bone_length, truncation_psi))
            H, W = feat.shape[2:4]
            warped = feat.permute(0, 1, 4, 2, 3).reshape(B, 32 * 3, H, W)
            return torch.cat([warped, self.tri_plane[:, 32 * 3:].expand(B, -1, -1, -1)], dim=1)
        if hasattr(self, "tri_plane"):
            return self.tri_plane.expand(z.shape[0], -1, -1, -1)
        if not hasattr(self, "tri_plane_gen"):
            raise RuntimeError("no tri-plane source: pass constant_triplane=True, a tri_plane_gen module or "
                               "model_input['tri_plane_feature']")
        return self.tri_plane_gen(z, self.encoded_bone_length(bone_length), truncation_psi=truncation_psi)

    def build_scene(self, pose_scaled, model_input: Dict):
        """SceneBuffers for one render call + the autograd-tracked tensors behind it."""
        bone_length, z, z_rend = model_input["bone_length"], model_input["z"], model_input["z_rend"]
        tri_plane = model_input.get("tri_plane_feature")
        decoder = model_input.get("decoder")        # handed over by rendering.render_entire_img_sharded (owner's weights)
        if decoder is None:
            decoder = self.mlp.effective(z_rend)
        cscale = (self.canonical_bone_length[None, :, None] / bone_length / self.coordinate_scale)[:, :, 0]
        if self.deformation_field and tri_plane is None:
            planes = model_input.get("planes")
            if planes is None:
                flow = model_input.get("flow")
                if flow is None:
                    flow = self.compute_flow(z, bone_length, model_input.get("truncation_psi", 1))
                planes = self.warped_planes(flow)
            scene = ops.SceneBuffers(pose_scaled, cscale, self.canonical_pose, None, decoder,
                                     feat_planes=planes[0], prob_planes=planes[1])
            return scene, planes, decoder
        if tri_plane is None:
            tri_plane = model_input.get("planes")
        if tri_plane is None:
            tri_plane = self.compute_tri_plane_feature(z, bone_length, model_input.get("truncation_psi", 1))
        if isinstance(tri_plane, (tuple, list)):
            # producer hand-off on the GAN path: a tri-plane generator whose last layer already writes the
            # renderer's layout -- (feature planes (B,3,H,W,32) channels-last fp32 / bf16, probability planes (B,3P,H,W))
            # -- so that neither the NCHW -> channels-last repack nor the gradient unpack runs, and the gradients come
            # back in the same two layouts
            feat, prob = tri_plane
            self.buffers_tensors["tri_plane_feature"] = (feat, prob)
            scene = ops.SceneBuffers(pose_scaled, cscale, self.canonical_pose, None, decoder,
                                     feat_planes=feat.contiguous(), prob_planes=prob.to(torch.float32).contiguous())
            return scene, (feat, prob), decoder
        tri_plane = tri_plane.to(torch.float32).contiguous()      # materialises an expanded constant plane
        self.buffers_tensors["tri_plane_feature"] = tri_plane      # narf.py:194
        if not self.training:
            self.temporal_state["tri_plane_feature"] = tri_plane
        scene = ops.SceneBuffers(pose_scaled, cscale, self.canonical_pose, tri_plane, decoder, self.plane_dtype)
        return scene, tri_plane, decoder

    # ---- point query (models/narf.py:176-211) ---------------------------------------------------------
    @on_device_of(1)
    def calc_density_and_color_from_camera_coord_v2(self, position, pose_to_camera, ray_direction, model_input):
        """position (B,3,n) in the scaled camera frame -> density (B,1,n), color (B,3,n).  Forward only."""
        scene, _, _ = self.build_scene(pose_to_camera, model_input)
        with torch.no_grad():
            den, col, _ = ops.density_query(scene, position.to(torch.float32).contiguous())
        return den, col

    @on_device_of(1)
    def density_grid(self, pose_to_camera, center, voxel_size=0.003, model_input=None, chunk=1 << 24):
        """The density volume ``create_mesh`` feeds to marching cubes (libraries/NARF/mesh_rendering.py:50-71):
        (2/voxel_size + 1)^3 points on a regular grid around ``center`` (1,3,1), in the scaled camera frame, queried
        through ``enarf_density_query``.  pose_to_camera (1,P,4,4) are part frames (``transform_pose`` applied, not yet
        scaled; unlike the reference the caller's tensor is not modified in place).  The scene is built once, the
        grid is generated on the device chunk by chunk and only the density volume is kept.
        -> (S,S,S) fp32 on the device, S = 2*int(1/voxel_size)+1, indexed [x][y][z] like the reference's."""
        cube = int(1 / voxel_size)
        S = 2 * cube + 1
        dev = pose_to_camera.device
        scene, _, _ = self.build_scene(rendering._scaled(self, pose_to_camera), dict(model_input or {}))
        bins = torch.arange(-cube, cube + 1, device=dev, dtype=torch.float32) / cube
        c = center.to(dev, torch.float32).reshape(3)
        out = torch.empty(S * S * S, device=dev)
        with torch.no_grad():
            for start in range(0, S * S * S, chunk):
                idx = torch.arange(start, min(start + chunk, S * S * S), device=dev)
                ix, iy, iz = torch.div(idx, S * S, rounding_mode="floor"), torch.div(idx, S, rounding_mode="floor") % S, idx % S
                pts = torch.stack([bins[ix] + c[0], bins[iy] + c[1], bins[iz] + c[2]], dim=0)[None]
                pts = (pts * self.coordinate_scale).contiguous()
                den, _, _ = ops.density_query(scene, pts, want_color=False)
                out[start:start + idx.numel()] = den.reshape(-1)
        return out.reshape(S, S, S)

    # ---- NARFBase.forward / render_entire_img (libraries/NARF/base.py:26-63) --------------------------
    def forward(self, batchsize, sampled_img_coord, pose_to_camera, inv_intrinsics, z, z_rend, bone_length,
                render_scale=1, Nc=64, Nf=128, return_intermediate=False, truncation_psi=1,
                camera_pose: Optional[torch.Tensor] = None, return_disparity=False, background: float = 0.0):
        """``background``: a constant the colour is composited over inside the fine pass (rgb + (1 - mask) * background)
        instead of three element-wise passes in the caller; 0 = the reference's plain foreground colour."""
        model_input = {"z": z, "z_rend": z_rend, "bone_length": bone_length, "truncation_psi": truncation_psi}
        pose_to_camera, model_input["bone_length"] = self.transform_pose(pose_to_camera, bone_length)
        self._bg_const = float(background)
        try:
            out = rendering.render(self, sampled_img_coord, pose_to_camera, inv_intrinsics, render_scale=render_scale,
                                   Nc=Nc, Nf=Nf, return_intermediate=return_intermediate, camera_pose=camera_pose,
                                   model_input=model_input)
        finally:
            self._bg_const = 0.0
        if return_intermediate:
            color, mask, _, inter = out
            return color, mask, inter
        color, mask, disparity = out
        return (color, mask, disparity) if return_disparity else (color, mask)

    def render_entire_img(self, pose_to_camera, inv_intrinsics, z, z_rend, bone_length, camera_pose=None,
                          render_size=128, Nc=64, Nf=128, semantic_map=False, use_normalized_intrinsics=False,
                          no_grad=True, truncation_psi=1, bbox=None):
        model_input = {"z": z, "z_rend": z_rend, "bone_length": bone_length, "truncation_psi": truncation_psi}
        pose_to_camera, model_input["bone_length"] = self.transform_pose(pose_to_camera, bone_length)
        with torch.set_grad_enabled(not no_grad):
            # NARF/base.py:59-60 hoists the plane generation out of the chunk loop.  It hands the generator the
            # UNtransformed bone_length there, which only works where transform_pose leaves it alone ("center",
            # "center_fixed"; with "center+head" its positional encoding cannot even be reshaped to 24 parts): the
            # transformed lengths are the same numbers in those cases and also work for "center+head".
            bl = model_input["bone_length"]
            if self.deformation_field:      # in the hand-off layout
                model_input["planes"] = self.warped_planes(self.compute_flow(z[:1], bl[:1], truncation_psi))
            else:
                model_input["tri_plane_feature"] = self.compute_tri_plane_feature(z, bl)
        return rendering.render_entire_img(self, pose_to_camera, inv_intrinsics, camera_pose, render_size, Nc, Nf,
                                           semantic_map, use_normalized_intrinsics, no_grad, model_input, bbox=bbox)

    def render_entire_img_sharded(self, pose_to_camera, inv_intrinsics, z, z_rend, bone_length, render_size=128, Nc=64,
                                  Nf=128, no_grad=True, truncation_psi=1, bbox=None, group=None, owner=0):
        """``render_entire_img`` with the one image's rays split over the ranks of ``group`` (fewer samples than GPUs);
        with ``no_grad=False`` the gradients of every rank's ray block are summed onto ``owner``'s parameters
        (rendering.render_entire_img_sharded)."""
        model_input = {"z": z, "z_rend": z_rend, "bone_length": bone_length, "truncation_psi": truncation_psi}
        pose_to_camera, model_input["bone_length"] = self.transform_pose(pose_to_camera, bone_length)
        return rendering.render_entire_img_sharded(self, pose_to_camera, inv_intrinsics, render_size=render_size, Nc=Nc,
                                                   Nf=Nf, model_input=model_input, group=group, no_grad=no_grad,
                                                   owner=owner, bbox=bbox)
