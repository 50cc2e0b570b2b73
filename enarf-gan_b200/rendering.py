"""Host side of the render loop: the drop-in for ``libraries/NeRF/rendering.py`` of the reference.

``render`` / ``render_entire_img`` keep the reference's signatures and side effects
(rendering.py:227-359, 362-427: ``model.buffers_tensors["fine_weights"|"fine_depth"|...]``); the work is
done by four kernel launches (depth range, fused coarse pass, fine sampling, fused fine pass) instead of
~60 eager torch ops.  ``render_rays`` is the same pipeline on explicit tensors.
"""
from typing import Dict, Optional

import torch

from . import ops

NEAR_PLANE = 0.3   # rendering.py:249
FAR_PLANE = 5.0    # rendering.py:138


def pixel_grid(render_size, patch_size, batchsize, device):
    """whole_image_grid_ray_sampler's homogeneous pixel centres (ray_sampler.py:42-67) -> (B,3,patch^2)."""
    a = torch.arange(patch_size, device=device, dtype=torch.float32)
    c = render_size * (a + 0.5) / patch_size
    x = c[None, :].expand(patch_size, patch_size).reshape(-1)
    y = c[:, None].expand(patch_size, patch_size).reshape(-1)
    return torch.stack([x, y, torch.ones_like(x)], dim=0)[None].expand(batchsize, 3, -1).contiguous()


def render_rays(scene: ops.SceneBuffers, image_coord, inv_intrinsics, Nc, Nf, rng="philox", seed=0,
                idx=None, jitter=None, compact_b1=True, tri_plane=None, decoder=None, nearfar=None,
                generator=None, return_intermediate=False):
    """image_coord (B,3,n) homogeneous pixels -> dict(color (B,3,n), mask (B,n), disparity (B,n),
    fine_weights (B,n',Nf-1), fine_depth (B,n',Nf), validity (B,n)); n' = valid rays when B == 1
    (rendering.py:107-110), else n.

    rng: "philox" (in-kernel draws, counter = launch ray index), "torch" (torch.multinomial + uniform_ on
    the device generator, the reference's own order, rendering.py:192-194) or "given" (idx, jitter).
    """
    B, _, n_all = image_coord.shape
    dev = image_coord.device
    inv_K = inv_intrinsics.to(torch.float32)
    if inv_K.ndim == 2:
        inv_K = inv_K[None].expand(B, 3, 3)
    inv_K = inv_K.contiguous()
    image_coord = image_coord.to(torch.float32).contiguous()
    dmin, dmax, direction, validity = ops.depth_range(image_coord, inv_K, scene.pose, nearfar)
    out = {"validity": validity}
    sel = None
    if B == 1 and compact_b1:                       # data-dependent shape: the reference syncs here too
        sel = validity[0].nonzero()[:, 0]
        dmin, dmax = dmin[:, sel].contiguous(), dmax[:, sel].contiguous()
        direction = direction[:, :, sel].contiguous()
    n = dmin.shape[1]
    if n == 0:                                      # rendering.py:351-354
        out.update(color=torch.zeros(B, 3, n_all, device=dev), mask=torch.zeros(B, n_all, device=dev),
                   disparity=torch.zeros(B, n_all, device=dev), fine_weights=None, fine_depth=None)
        return out
    with torch.no_grad():
        wb = ops.render_coarse(scene, dmin, dmax, direction, Nc)
        if rng == "philox":
            bins = ops.fine_sample(wb, Nf, seed=seed)
        elif rng == "torch":
            idx = torch.multinomial(wb.reshape(B * n, Nc), Nf, replacement=True, generator=generator)
            jitter = torch.empty(B, 1, n, Nf, device=dev).uniform_(generator=generator)
            bins = ops.fine_sample(wb, Nf, idx=idx.reshape(B, n, Nf), jitter=jitter.reshape(B, n, Nf))
        elif rng == "given":
            bins = ops.fine_sample(wb, Nf, idx=idx.reshape(B, n, Nf), jitter=jitter.reshape(B, n, Nf))
        else:
            raise ValueError(f"unknown rng mode {rng!r}")
    color, mask, disp, fw, fd = ops.render_fine(scene, dmin, dmax, direction, bins, tri_plane, decoder)
    if sel is not None:                             # rendering.py:337-350
        full_c = torch.zeros(B, 3, n_all, device=dev).index_copy(2, sel, color)
        full_m = torch.zeros(B, n_all, device=dev).index_copy(1, sel, mask)
        full_d = torch.zeros(B, n_all, device=dev).index_copy(1, sel, disp)
        color, mask, disp = full_c, full_m, full_d
    out.update(color=color, mask=mask, disparity=disp, fine_weights=fw, fine_depth=fd)
    if return_intermediate:
        out.update(coarse_weights_blurred=wb, bins=bins, depth_min=dmin, depth_max=dmax, direction=direction)
    return out
