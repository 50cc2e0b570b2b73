"""Drop-in for ``libraries/triplane/sampling.py`` (reference lines 9-127): the three gather front-ends with
their reference signatures, for callers that use them directly.  The render loop itself does not go
through these (it uses the fused kernels); they exist so that code importing
``sample_feature / sample_triplane_part_prob / sample_weighted_feature_v2`` keeps working.  All of them run
on the stand-alone tri-plane sampler kernel, per sample, i.e. without the reference's "planes side by side"
trick and its fp32 x-remap noise."""
from typing import Optional

import torch

from .triplane_sampler import triplane_sampler


def sample_feature(tri_plane_features, position, reduction: str = "sum", clamp_mask=False,
                   batch_idx: Optional[torch.Tensor] = None):
    """tri_plane_features (B, C*3, h, w), position (B,3,n) in [-1,1] -> (B,C,n) ("sum") or (B,1,n) ("prod":
    product over planes of sigmoid(sample), C must be 1)."""
    if batch_idx is not None:
        raise NotImplementedError("batch_idx is an artefact of the reference's side-by-side layout; pass the "
                                  "batched planes instead")
    grid = position.permute(0, 2, 1).contiguous()[:, :, None, :]                 # (B,n,1,3)
    if reduction == "sum":
        return triplane_sampler(tri_plane_features, grid)[:, :, :, 0]
    if reduction != "prod":
        raise ValueError(reduction)
    B, C3, h, w = tri_plane_features.shape
    assert C3 == 3, "prod reduction samples one probability channel per plane"
    # channel c of plane k is kept only when c == k, so output channel k holds plane k's sample alone
    eye = torch.eye(3, device=tri_plane_features.device, dtype=tri_plane_features.dtype)
    spread = (tri_plane_features[:, :, None] * eye[None, :, :, None, None]).reshape(B, 9, h, w)
    feature = triplane_sampler(spread, grid)[:, :, :, 0]                         # (B,3,n)
    if clamp_mask:
        feature = (feature.detach().clamp(-2, 5) - feature.detach()) + feature
    return torch.sigmoid(feature).prod(dim=1, keepdim=True)


def sample_triplane_part_prob(tri_plane_weights, position, position_validity, mode="prod", clamp_mask=False):
    """tri_plane_weights (B*P,3,h,w), position (B,P,3,n) -> weight (B,P,n)   [sampling.py:54-76]."""
    bs, n_bone, _, n = position.shape
    position = position.reshape(bs * n_bone, 3, n)
    if mode == "prod":
        return sample_feature(tri_plane_weights, position, clamp_mask=clamp_mask, reduction="prod").reshape(bs, n_bone, n)
    if mode == "sum":
        raise NotImplementedError("mode='sum' needs 3-channel-per-plane weights; unused by the reference's callers")
    return torch.ones(bs, n_bone, n, device=position.device) / n_bone


def sample_weighted_feature_v2(feat_dim, tri_plane_features, position, weight, position_validity,
                               clamp_mask: bool = False):
    """tri_plane_features (B,3*feat_dim,h,w), position (B,P,3,n) (invalid parts at coordinate 2), weight and
    position_validity (B,P,n) -> (B,feat_dim,n) = sum_p validity * weight * sample   [sampling.py:79-127]."""
    B, P, _, n = position.shape
    pos = position.permute(0, 2, 1, 3).reshape(B, 3, P * n)
    value = sample_feature(tri_plane_features, pos, clamp_mask=clamp_mask).reshape(B, feat_dim, P, n)
    return (value * (weight * position_validity)[:, None]).sum(dim=2)
