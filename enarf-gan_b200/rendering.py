"""Host side of the render loop: the drop-in for ``libraries/NeRF/rendering.py`` of the reference.

``render`` / ``render_entire_img`` keep the reference's signatures and side effects
(rendering.py:227-359, 362-427: ``model.buffers_tensors["fine_weights"|"fine_depth"|...]``); the work is
done by four kernel launches (depth range, fused coarse pass, fine sampling, fused fine pass) instead of
~60 eager torch ops.  ``render_rays`` is the same pipeline on explicit tensors.
"""
from typing import Dict, Optional

import torch

from . import ops
from ._abi import on_device_of

NEAR_PLANE = 0.3   # rendering.py:249
FAR_PLANE = 5.0    # rendering.py:138


def pixel_grid(render_size, patch_size, batchsize, device):
    """whole_image_grid_ray_sampler's homogeneous pixel centres (ray_sampler.py:42-67) -> (B,3,patch^2)."""
    a = torch.arange(patch_size, device=device, dtype=torch.float32)
    c = render_size * (a + 0.5) / patch_size
    x = c[None, :].expand(patch_size, patch_size).reshape(-1)
    y = c[:, None].expand(patch_size, patch_size).reshape(-1)
    return torch.stack([x, y, torch.ones_like(x)], dim=0)[None].expand(batchsize, 3, -1).contiguous()


@on_device_of(1)
def render_rays(scene: ops.SceneBuffers, image_coord, inv_intrinsics, Nc, Nf, rng="philox", seed=0, seed_dev=None,
                idx=None, jitter=None, compact_b1=True, tri_plane=None, decoder=None, nearfar=None,
                generator=None, return_intermediate=False, want_points=False, ert_eps=0.0, ert_prev=None,
                bg_const=0.0):
    """image_coord (B,3,n) homogeneous pixels -> dict(color (B,3,n), mask (B,n), disparity (B,n),
    fine_weights (B,n',Nf-1), fine_depth (B,n',Nf), validity (B,n)); n' = valid rays when B == 1
    (rendering.py:107-110), else n.

    ert_eps > 0: early ray termination of the fine pass (samples behind the depth where the ray's transmittance has
    fallen below ert_eps are neither evaluated nor differentiated; outputs move by < ert_eps), and of the coarse pass.
    Rounds (extra barriers) only run where rays are opaque: the fine pass asks this call's coarse pass, the coarse pass
    asks ``ert_prev`` = out["opaque_rays"] of the previous call on this scene (None: assume translucent).
    out["ert_stats"] is an int32 (4,) on the device: fine samples evaluated, skipped, coarse evaluated, skipped.

    bg_const: colour = rgb + (1 - mask) * bg_const, composited in the fine pass's epilogue (rays that are not rendered
    come out as bg_const).

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
        out.update(color=torch.full((B, 3, n_all), float(bg_const), device=dev), mask=torch.zeros(B, n_all, device=dev),
                   disparity=torch.zeros(B, n_all, device=dev), fine_weights=None, fine_depth=None)
        return out
    hint = stats = None
    if ert_eps > 0:
        counters = torch.zeros(6, device=dev, dtype=torch.int32)
        hint, stats, prev = counters[:1], counters[1:5], (ert_prev if ert_prev is not None else counters[5:])
        out["opaque_rays"] = hint
    with torch.no_grad():
        wb = ops.render_coarse(scene, dmin, dmax, direction, Nc, opaque_rays=hint, ert_eps=ert_eps,
                               ert_hint=prev if ert_eps > 0 else None, ert_stats=stats[2:] if ert_eps > 0 else None)
        if rng == "philox":
            bins = ops.fine_sample(wb, Nf, seed=seed, seed_dev=seed_dev)
        elif rng == "torch":
            idx = torch.multinomial(wb.reshape(B * n, Nc), Nf, replacement=True, generator=generator)
            jitter = torch.empty(B, 1, n, Nf, device=dev).uniform_(generator=generator)
            bins = ops.fine_sample(wb, Nf, idx=idx.reshape(B, n, Nf), jitter=jitter.reshape(B, n, Nf))
        elif rng == "given":
            bins = ops.fine_sample(wb, Nf, idx=idx.reshape(B, n, Nf), jitter=jitter.reshape(B, n, Nf))
        else:
            raise ValueError(f"unknown rng mode {rng!r}")
    color, mask, disp, fw, fd, pts = ops.render_fine(scene, dmin, dmax, direction, bins, tri_plane, decoder,
                                                     want_points=want_points, bg_const=bg_const, ert_eps=ert_eps, ert_hint=hint,
                                                     ert_stats=stats[:2] if ert_eps > 0 else None)
    if ert_eps > 0:
        out["ert_stats"] = stats
        if pts is not None:
            pts = torch.cat([pts[..., :3], pts[..., 3:].clamp_min(0)], dim=-1)    # dropped samples are marked sigma = -1
    if sel is not None:                             # rendering.py:337-350
        full_c = torch.full((B, 3, n_all), float(bg_const), device=dev).index_copy(2, sel, color)
        full_m = torch.zeros(B, n_all, device=dev).index_copy(1, sel, mask)
        full_d = torch.zeros(B, n_all, device=dev).index_copy(1, sel, disp)
        color, mask, disp = full_c, full_m, full_d
    out.update(color=color, mask=mask, disparity=disp, fine_weights=fw, fine_depth=fd, points=pts)
    if return_intermediate:
        out.update(coarse_weights_blurred=wb, bins=bins, depth_min=dmin, depth_max=dmax, direction=direction)
    return out


# ------------------------------------------------------------------------------------------------------
# Drop-in signatures of the reference module (libraries/NeRF/rendering.py:227-427)
# ------------------------------------------------------------------------------------------------------
def _scaled(model, pose_to_camera):
    if model.coordinate_scale != 1:                      # rendering.py:258-260
        pose_to_camera = pose_to_camera.clone()
        pose_to_camera[:, :, :3, 3] *= model.coordinate_scale
    return pose_to_camera


@on_device_of(2)
def render(model, image_coord, pose_to_camera, inv_intrinsics, render_scale: float = 1, Nc: int = 64,
           Nf: int = 128, semantic_map: bool = False, return_intermediate: bool = False,
           camera_pose: Optional[torch.Tensor] = None, model_input: Dict = {}, _scene=None):
    """Same contract as the reference's ``render``: image_coord (B,1,3,n), pose_to_camera (B,P,4,4) part
    frames -> (color (B,3,n), mask (B,n), disparity (B,n)[, (fine_points, fine_density)]); stores
    ``model.buffers_tensors["fine_weights"|"fine_depth"|"fine_density"]`` ((B,1,n',Nf-1), (B,1,n',Nf),
    (B,1,n'*Nf)) like rendering.py:287,323-324."""
    if render_scale != 1:
        raise NotImplementedError("render_scale != 1 is never used by the reference's callers")
    if semantic_map:
        raise NotImplementedError("semantic map rendering is disabled in the reference too (rendering.py:299)")
    if pose_to_camera.requires_grad:
        raise NotImplementedError("Currently pose should not be differentiable")     # rendering.py:216-217
    B, _, _, n_all = image_coord.shape
    pose_scaled = _scaled(model, pose_to_camera)
    if not hasattr(model, "buffers_tensors"):
        model.buffers_tensors = {}
    if _scene is None:
        _scene = model.build_scene(pose_scaled, model_input)
    scene, tri_plane, decoder = _scene
    planes_grad = any(t.requires_grad for t in tri_plane) if isinstance(tri_plane, (tuple, list)) \
        else tri_plane.requires_grad
    track = torch.is_grad_enabled() and (planes_grad or any(w.requires_grad for w, _ in decoder))
    # sample-sharded batch (one process per GPU): near / far are batch-global in the reference (rendering.py:15-17), so
    # the ranks exchange two floats (switch off with model.sync_near_far = False), unless model.batch_near_far is set
    nearfar = getattr(model, "batch_near_far", None)     # a caller that knows the whole batch's range hands it over
    if nearfar is None and getattr(model, "sync_near_far", True):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            nearfar = global_near_far(scene.pose)
    out = render_rays(scene, image_coord.reshape(B, 3, n_all), inv_intrinsics, Nc, Nf, nearfar=nearfar,
                      rng=getattr(model, "rng_mode", "torch"), seed=getattr(model, "rng_seed", 0),
                      seed_dev=getattr(model, "rng_seed_dev", None),
                      idx=getattr(model, "rng_given", (None, None))[0],
                      jitter=getattr(model, "rng_given", (None, None))[1],
                      tri_plane=tri_plane if track else None, decoder=decoder if track else None,
                      return_intermediate=return_intermediate or not model.training,
                      # per-sample outputs only where somebody reads them (evaluation, return_intermediate): a training
                      # step then skips the rays' last samples, whose density the reference computes and drops
                      want_points=return_intermediate or not model.training,
                      ert_eps=float(getattr(model, "ert_eps", 0.0)), ert_prev=getattr(model, "_ert_prev", None),
                      bg_const=float(getattr(model, "_bg_const", 0.0)))
    if "ert_stats" in out:
        model.buffers_tensors["ert_stats"] = out["ert_stats"]
        model._ert_prev = out["opaque_rays"]        # the next call's coarse pass decides on it whether to run in rounds
    if getattr(model, "rng_mode", "torch") == "philox":
        if getattr(model, "rng_seed_dev", None) is not None:
            model.rng_seed_dev.add_(1)                          # on the device: also advances inside a replayed CUDA graph
        else:
            model.rng_seed = getattr(model, "rng_seed", 0) + 1  # fresh draws on the next call
    if out["fine_weights"] is not None:
        n = out["fine_weights"].shape[1]
        model.buffers_tensors["fine_weights"] = out["fine_weights"][:, None]
        model.buffers_tensors["fine_depth"] = out["fine_depth"][:, None]
        # (a training step without return_intermediate does not materialise the per-sample densities: None)
        model.buffers_tensors["fine_density"] = (out["points"][..., 3].reshape(B, 1, n * Nf) if out["points"] is not None
                                                 else None)
        if not model.training:
            model.temporal_state.update(fine_depth=out["fine_depth"][:, None],
                                        fine_density=out["points"][..., :-1, 3][:, None],
                                        coarse_weights=out["coarse_weights_blurred"].reshape(B * n, Nc))
    if return_intermediate:
        inter = None
        if out["fine_weights"] is not None:
            u = out["bins"][:, None]                                                      # (B,1,n,Nf)
            start = (out["depth_min"][:, None] * out["direction"])[..., None]
            end = (out["depth_max"][:, None] * out["direction"])[..., None]
            fine_points = (start * (1 - u) + end * u).reshape(B, 3, -1)
            inter = (fine_points, model.buffers_tensors["fine_density"])
        return out["color"], out["mask"], out["disparity"], inter
    return out["color"], out["mask"], out["disparity"]


@on_device_of(1)
def render_entire_img(model, pose_to_camera, inv_intrinsics, camera_pose: Optional[torch.Tensor] = None,
                      render_size: int = 128, Nc: int = 64, Nf: int = 128, semantic_map: bool = False,
                      use_normalized_intrinsics: bool = False, no_grad: bool = True, model_input: Dict = {},
                      bbox=None, ray_range=None):
    """Chunked full-image render with the reference's ``render_bs`` semantics (rendering.py:362-427): batch
    forced to 1, pixel centres (optionally bbox-cropped / normalised), chunks of ``model.config.render_bs``
    rays rendered independently and concatenated.  The scene (repacked planes, decoder weights) is built once."""
    cfg = model.config
    ray_batchsize = cfg["render_bs"] if isinstance(cfg, dict) else cfg.render_bs
    if bbox is not None:
        width, height = bbox[2] - bbox[0], bbox[3] - bbox[1]
        x_off, y_off = bbox[0], bbox[1]
    else:
        width, height, x_off, y_off = render_size, render_size, 0, 0
    dev = pose_to_camera.device
    ar = torch.arange(width * height, device=dev)
    x = (ar % width + 0.5 + x_off).float()
    y = (torch.div(ar, width, rounding_mode="floor") + 0.5 + y_off).float()
    if use_normalized_intrinsics:
        x, y = x / render_size, y / render_size
    img_coord = torch.stack([x, y, torch.ones_like(x)], dim=0)[None, None]                # (1,1,3,n)
    if ray_range is not None:          # this rank's block of the image's rays (render_entire_img_sharded)
        img_coord = img_coord[..., ray_range[0]:ray_range[1]]
    pose1 = pose_to_camera[:1]
    mi = dict(model_input)
    for key in ("z", "z_rend", "bone_length", "tri_plane_feature"):
        if mi.get(key) is not None:
            mi[key] = mi[key][:1]
    colors, masks, disps = [], [], []
    with torch.set_grad_enabled(not no_grad):
        scene = model.build_scene(_scaled(model, pose1), mi)
        for i in range(0, img_coord.shape[-1], ray_batchsize):
            c, m, d = render(model, img_coord[:, :, :, i:i + ray_batchsize], pose1, inv_intrinsics, Nc=Nc, Nf=Nf,
                             camera_pose=camera_pose, model_input=mi, _scene=scene)
            colors.append(c)
            masks.append(m)
            disps.append(d)
        color, mask, disp = torch.cat(colors, dim=2), torch.cat(masks, dim=1), torch.cat(disps, dim=1)
    if ray_range is not None:
        return color[0], mask[0], disp[0]                                                 # (3,n_local), (n_local,) x2
    return color.reshape(3, height, width), mask.reshape(height, width), disp.reshape(height, width)


# ------------------------------------------------------------------------------------------------------
# Multi-GPU: sample-major ray sharding.  No data-path collective; the only cross-sample
# quantity of the reference is the batch-global near/far (rendering.py:15-17), exchanged as two floats.
# ------------------------------------------------------------------------------------------------------
def shard_samples(batch: int, rank: int, world: int):
    """Contiguous block of samples rendered by ``rank`` (sizes differ by at most one)."""
    base, extra = divmod(batch, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def balanced_shards(costs, world: int):
    """Deal the batch's samples to ``world`` ranks so that the ranks' summed ``costs`` are even: longest-processing-time
    first under the constraint that no rank gets more than ceil(B / world) samples (every GPU holds its samples'
    planes).  ``costs``: one positive number per sample, e.g. ``sample_cost_proxy``.  -> list of index lists.
    A sample's render cost varies with how much of the image the body covers and how many part boxes overlap; contiguous
    blocks (``shard_samples``) leave the slowest of 8 ranks ~15 % behind the mean on the benchmark batch."""
    costs = [float(c) for c in costs]
    cap = -(-len(costs) // world)
    order = sorted(range(len(costs)), key=lambda i: -costs[i])
    load, out = [0.0] * world, [[] for _ in range(world)]
    for i in order:
        r = min((r for r in range(world) if len(out[r]) < cap), key=lambda r: load[r])
        out[r].append(i)
        load[r] += costs[i]
    return [sorted(b) for b in out]


def sample_cost_proxy(pose_scaled, cscale=None):
    """A cheap host-side estimate of a sample's render cost from its part frames alone (B,P,4,4): the number of
    (point, part) pairs grows with the image area every part box covers, ~ sum over parts of 1 / depth^2."""
    tz = pose_scaled[:, :, 2, 3].double().clamp_min(1e-3)
    return (1.0 / tz ** 2).sum(dim=1).tolist()


def global_near_far(pose_local, near0: float = NEAR_PLANE, far0: float = FAR_PLANE, group=None):
    """near/far over the WHOLE batch from this rank's shard of the (scaled) part frames (B_local,P,4,4).

    near = max(min t_z - sqrt3, near0), far = max(max t_z + sqrt3, far0) (rendering.py:15-17); the min / max
    over ranks is ONE all_reduce(MAX) of the pair (-min, max) (NCCL on GPUs, gloo in the CPU tests).  Returns the
    2-element tensor ``ops.depth_range(..., nearfar=)`` expects; with no process group it is the local value.
    """
    import torch.distributed as dist
    tz = pose_local[:, :, 2, 3].to(torch.float32)
    ext = torch.stack([-tz.min(), tz.max()])           # one MAX reduction carries both extremes
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(ext, op=dist.ReduceOp.MAX, group=group)
    sqrt3 = 3 ** 0.5
    near = torch.clamp_min(-ext[0] - sqrt3, near0)
    far = torch.clamp_min(ext[1] + sqrt3, far0)
    return torch.stack([near, far])


def shard_rays(n_rays: int, rank: int, world: int):
    """Contiguous block of one image's rays rendered by ``rank`` when there are fewer samples than GPUs
    (``render_entire_img`` has batch 1; SURVEY 8e "B < G")."""
    return shard_samples(n_rays, rank, world)


def gather_rays(local: torch.Tensor, n_rays: int, group=None):
    """All-gather the per-rank ray blocks ``(..., n_local)`` into ``(..., n_rays)`` on every rank.  Blocks differ by
    at most one ray, so they are padded to a common length for one ``all_gather`` and trimmed afterwards."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    longest = -(-n_rays // world)
    pad = longest - local.shape[-1]
    send = torch.nn.functional.pad(local, (0, pad)) if pad else local.contiguous()
    parts = [torch.empty_like(send) for _ in range(world)]
    dist.all_gather(parts, send.contiguous(), group=group)
    out = []
    for r, part in enumerate(parts):
        lo, hi = shard_rays(n_rays, r, world)
        out.append(part[..., :hi - lo])
    return torch.cat(out, dim=-1)


class _FromOwner(torch.autograd.Function):
    """The owner rank's tensor on every rank (broadcast); in the backward the ranks' partial gradients are summed back
    onto the owner (reduce), the other ranks get zeros.  Every rank passes a tensor of the same shape and dtype."""

    @staticmethod
    def forward(ctx, t, owner, group):
        import torch.distributed as dist
        ctx.owner, ctx.group = owner, group
        out = t.detach().clone().contiguous()
        dist.broadcast(out, src=owner, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        import torch.distributed as dist
        g = g.contiguous().clone()
        dist.reduce(g, dst=ctx.owner, op=dist.ReduceOp.SUM, group=ctx.group)
        if dist.get_rank(ctx.group) != ctx.owner:
            g.zero_()
        return g, None, None


class _GatherRays(torch.autograd.Function):
    """``gather_rays`` with a gradient: every rank evaluates the same loss on the assembled tensor, so the gradient of
    its own block is simply its slice of the incoming gradient (the sum over ranks happens where the blocks' shared
    inputs are reduced, ``_FromOwner``)."""

    @staticmethod
    def forward(ctx, local, n_rays, group):
        import torch.distributed as dist
        ctx.range = shard_rays(n_rays, dist.get_rank(group), dist.get_world_size(group))
        return gather_rays(local.detach(), n_rays, group)

    @staticmethod
    def backward(ctx, g):
        lo, hi = ctx.range
        return g[..., lo:hi].contiguous(), None, None


def from_owner(t, owner=0, group=None):
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return t
    return _FromOwner.apply(t, owner, group)


def render_entire_img_sharded(model, pose_to_camera, inv_intrinsics, render_size: int = 128, Nc: int = 64, Nf: int = 128,
                              model_input: Dict = {}, group=None, no_grad: bool = True, owner: int = 0, **kw):
    """``render_entire_img`` with the image's rays split across the ranks of ``group`` -- fewer samples than GPUs
    (SURVEY 8e "B < G"; the reference renders the one image chunk by chunk on one GPU, rendering.py:362-427).
    Rank r renders rays ``shard_rays(n, r, world)``; one all_gather per output assembles the image on every rank.

    ``no_grad=True``: every rank holds the model and the planes (replicated parameters, or
    ``model_input["tri_plane_feature"]`` given on every rank).
    ``no_grad=False`` (training on one image): the OWNER's tri-plane and per-sample decoder weights are broadcast
    (44 MB + 26 KB over NVLink), every rank renders and differentiates its ray block, and in the backward the partial
    plane / decoder gradients are summed back onto the owner (one ``reduce`` each), where they flow on into whatever
    produced them (tri-plane generator, modulation); the other ranks' parameters receive no gradient.  Every rank
    must evaluate the same loss on the returned image."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    bbox = kw.get("bbox")
    width, height = (bbox[2] - bbox[0], bbox[3] - bbox[1]) if bbox is not None else (render_size, render_size)
    n = width * height
    mi = dict(model_input)
    if not no_grad and world > 1:
        with torch.set_grad_enabled(True):
            tp = mi.get("tri_plane_feature")
            if tp is None:
                tp = model.compute_tri_plane_feature(mi["z"][:1], mi["bone_length"][:1], mi.get("truncation_psi", 1))
            mi["tri_plane_feature"] = from_owner(tp[:1].to(torch.float32).contiguous(), owner, group)
            mi["decoder"] = [(from_owner(w, owner, group), from_owner(b, owner, group))
                             for w, b in model.mlp.effective(mi["z_rend"][:1])]
    color, mask, disp = render_entire_img(model, pose_to_camera, inv_intrinsics, render_size=render_size, Nc=Nc, Nf=Nf,
                                          model_input=mi, ray_range=shard_rays(n, rank, world), no_grad=no_grad, **kw)
    if world > 1 and not no_grad:
        color, mask, disp = (_GatherRays.apply(t, n, group) for t in (color, mask, disp))
    else:
        color, mask, disp = gather_rays(color, n, group), gather_rays(mask, n, group), gather_rays(disp, n, group)
    return color.reshape(3, height, width), mask.reshape(height, width), disp.reshape(height, width)
