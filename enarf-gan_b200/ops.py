"""Torch-facing operators over the C ABI: one function per exported entry point plus the autograd
Function of the fused fine pass.  Tensors in, tensors out; every buffer is allocated here by torch's
caching allocator and handed to the library as a raw pointer on torch's current stream."""
import ctypes

import torch

from . import _abi
from ._abi import check, lib, ptr, stream_ptr

FEAT_DIM = 32
HIDDEN = 64
# Checkpoint the gathered features (B,n,Nf,32) in the forward so the backward skips its gather; costs 128 B per
# fine sample of device memory.  Set False to recompute instead.
SAVE_FEATURES = True


class SceneBuffers:
    """Owns the device tensors one render call needs and the enarf_scene_t that points at them.

    pose (B,P,4,4) is the part frames with scaled translation (rendering.py:258-260 applied), cscale
    (B,P) = canonical_bone_length / bone_length / coordinate_scale (narf.py:165), canon_pose (P,4,4),
    tri_plane (B,(32+P)*3,H,W) NCHW as the reference produces it, decoder = [(W_eff (B,out,in), bias), ...].
    """

    def __init__(self, pose, cscale, canon_pose, tri_plane, decoder, plane_dtype=torch.float32, feat_planes=None,
                 prob_planes=None):
        B, P = pose.shape[:2]
        f32 = lambda t: t.detach().to(torch.float32).contiguous()
        self.pose, self.cscale, self.canon_pose = f32(pose), f32(cscale), f32(canon_pose)
        if tri_plane is None:
            # producer hand-off: channels-last feature planes (B,3,H,W,32) as warp_planes emits them
            # and NCHW probability planes (B or 1, 3P, H, W); a batch of 1 is shared by every sample (stride 0)
            assert feat_planes is not None and prob_planes is not None
            assert feat_planes.shape[0] == B and feat_planes.shape[1] == 3 and feat_planes.shape[4] == FEAT_DIM
            assert prob_planes.shape[0] in (1, B) and prob_planes.shape[1] == 3 * P
            assert prob_planes.dtype == torch.float32 and prob_planes[0].is_contiguous()
            H, W = feat_planes.shape[2:4]
            assert tuple(prob_planes.shape[-2:]) == (H, W)
            plane_dtype = feat_planes.dtype
            self.tri_plane = None
        else:
            H, W = tri_plane.shape[-2:]
            assert tri_plane.shape[1] == (FEAT_DIM + P) * 3, "tri-plane must have (32+P)*3 channels"
            self.tri_plane = tri_plane if (tri_plane.dtype == torch.float32 and tri_plane.is_contiguous()) \
                else tri_plane.detach().to(torch.float32).contiguous()
        self.plane_dtype = plane_dtype
        self.B, self.P, self.H, self.W = B, P, H, W
        (w1, b1), (w2, b2), (w3, b3) = decoder
        assert w1.shape[1:] == (HIDDEN, FEAT_DIM) and w2.shape[1:] == (HIDDEN, HIDDEN) and w3.shape[1:] == (4, HIDDEN)
        self.w = [f32(w1), f32(w2), f32(w3)]
        self.b = [f32(b1).reshape(-1), f32(b2).reshape(-1), f32(b3).reshape(-1)]
        if feat_planes is None:
            feat_planes = repack_planes(self.tri_plane.detach(), plane_dtype)
        self.feat_planes = feat_planes.detach()
        assert self.feat_planes.is_contiguous()
        if self.tri_plane is None:
            self.prob_view = prob_planes.detach()
            prob_stride = 0 if prob_planes.shape[0] == 1 else prob_planes.stride(0)
        else:
            self.prob_view = self.tri_plane.detach()[:, FEAT_DIM * 3:]
            prob_stride = self.tri_plane.stride(0)
        s = _abi.Scene()
        s.B, s.P, s.H, s.W = B, P, H, W
        s.pose, s.cscale, s.canon_pose = ptr(self.pose), ptr(self.cscale), ptr(self.canon_pose)
        s.feat_planes = ptr(self.feat_planes, None)
        s.feat_dtype = _abi.DTYPE_F32 if plane_dtype == torch.float32 else _abi.DTYPE_BF16
        s.prob_planes = self.prob_view.data_ptr()
        s.prob_batch_stride = prob_stride
        s.w1, s.w2, s.w3 = (ptr(t) for t in self.w)
        s.b1, s.b2, s.b3 = (ptr(t) for t in self.b)
        self.c = s

    @property
    def ref(self):
        return ctypes.byref(self.c)


def repack_planes(tri_plane, dtype=torch.float32):
    """(B,C,H,W) fp32 NCHW -> channels-last feature planes (B,3,H,W,32)."""
    B, C, H, W = tri_plane.shape
    out = torch.empty(B, 3, H, W, FEAT_DIM, device=tri_plane.device, dtype=dtype)
    code = _abi.DTYPE_F32 if dtype == torch.float32 else _abi.DTYPE_BF16
    check(lib().enarf_repack_planes(ptr(tri_plane), B, C, H, W, ptr(out, None), code, stream_ptr()))
    return out


def unpack_plane_grads(d_feat, d_tri_plane, accumulate=True):
    """Adds (or, with accumulate=False, stores) d_feat (B,3,H,W,32) into channels [0,96) of the NCHW gradient."""
    B, C, H, W = d_tri_plane.shape
    check(lib().enarf_unpack_plane_grads(ptr(d_feat), B, C, H, W, int(accumulate), ptr(d_tri_plane), stream_ptr()))
    return d_tri_plane


class WarpPlanes(torch.autograd.Function):
    """Deformation-field producer (models/narf.py:39-58) emitting the renderer's layout.  canon_feat (3,H,W,32) fp32
    channels-last, flow (B,6,H,W) fp32 -> (B,3,H,W,32) in ``dtype``.  The incoming gradient is consumed in the
    output's dtype (autograd casts it), d_canon_feat is accumulated in fp32."""

    @staticmethod
    def forward(ctx, canon_feat, flow, dtype):
        _, H, W, _ = canon_feat.shape
        B = flow.shape[0]
        assert flow.shape == (B, 6, H, W) and canon_feat.shape == (3, H, W, FEAT_DIM)
        canon_feat, flow = canon_feat.contiguous(), flow.to(torch.float32).contiguous()
        out = torch.empty(B, 3, H, W, FEAT_DIM, device=flow.device, dtype=dtype)
        check(lib().enarf_warp_planes_fwd(ptr(canon_feat), ptr(flow), B, H, W, ptr(out, None), _dtype_code(dtype),
                                          stream_ptr()))
        ctx.save_for_backward(canon_feat, flow)
        return out

    @staticmethod
    def backward(ctx, g):
        canon_feat, flow = ctx.saved_tensors
        B, _, H, W = flow.shape
        g = g.contiguous()
        d_canon = torch.zeros_like(canon_feat) if ctx.needs_input_grad[0] else None
        d_flow = torch.empty_like(flow) if ctx.needs_input_grad[1] else None
        if d_canon is not None or d_flow is not None:
            check(lib().enarf_warp_planes_bwd(ptr(g, None), _dtype_code(g.dtype), ptr(canon_feat), ptr(flow), B, H, W,
                                              ptr(d_canon, allow_none=True), ptr(d_flow, allow_none=True), stream_ptr()))
        return d_canon, d_flow, None


def _dtype_code(dtype):
    if dtype == torch.float32:
        return _abi.DTYPE_F32
    if dtype == torch.bfloat16:
        return _abi.DTYPE_BF16
    raise TypeError(f"plane dtype must be float32 or bfloat16, got {dtype}")


def warp_planes(canon_feat, flow, dtype=torch.float32):
    """canon_feat (3,H,W,32) channels-last fp32, flow (B,6,H,W) -> warped feature planes (B,3,H,W,32)."""
    return WarpPlanes.apply(canon_feat, flow, dtype)


class RepackCanonical(torch.autograd.Function):
    """(1,C,H,W) NCHW parameter -> (3,H,W,32) channels-last feature planes, with the transposed gradient."""

    @staticmethod
    def forward(ctx, tri_plane):
        ctx.shape = tri_plane.shape
        return repack_planes(tri_plane.detach().to(torch.float32).contiguous())[0]

    @staticmethod
    def backward(ctx, g):
        d = torch.zeros(ctx.shape, device=g.device)
        return unpack_plane_grads(g.contiguous().float()[None], d, accumulate=False)


def near_far(pose, near0=0.3, far0=5.0):
    """Batch-global near/far as a 2-element device tensor (no host sync)."""
    B, P = pose.shape[:2]
    out = torch.empty(2, device=pose.device, dtype=torch.float32)
    check(lib().enarf_near_far(ptr(pose), B, P, near0, far0, ptr(out), stream_ptr()))
    return out


def depth_range(image_coord, inv_K, pose, nearfar=None):
    """image_coord (B,3,n), inv_K (B,3,3), pose (B,P,4,4) -> dmin, dmax (B,n), dir (B,3,n), validity (B,n) bool."""
    B, _, n = image_coord.shape
    P = pose.shape[1]
    if nearfar is None:
        nearfar = near_far(pose)
    dev = pose.device
    dmin = torch.empty(B, n, device=dev)
    dmax = torch.empty(B, n, device=dev)
    direction = torch.empty(B, 3, n, device=dev)
    valid = torch.empty(B, n, device=dev, dtype=torch.uint8)
    check(lib().enarf_depth_range(ptr(image_coord), ptr(inv_K), ptr(pose), ptr(nearfar), B, P, n, ptr(dmin), ptr(dmax),
                                  ptr(direction), ptr(valid, torch.uint8), stream_ptr()))
    return dmin, dmax, direction, valid.bool()


def render_coarse(scene, dmin, dmax, direction, Nc, want_density=False, opaque_rays=None, ert_eps=0.0, ert_hint=None,
                  ert_stats=None):
    """-> blurred coarse weights (B,n,Nc) [, density (B,n,Nc)].  ``opaque_rays``: a zeroed int32 tensor of one element
    that receives the number of rays the coarse pass found opaque (the early-termination hint of the fine pass).
    ert_eps / ert_hint / ert_stats: early termination of the coarse samples (see render_rays)."""
    B, n = dmin.shape
    wb = torch.empty(B, n, Nc, device=dmin.device)
    den = torch.empty(B, n, Nc, device=dmin.device) if want_density else None
    check(lib().enarf_render_coarse(scene.ref, ptr(dmin), ptr(dmax), ptr(direction), n, Nc, ptr(wb),
                                    ptr(den, allow_none=True), ptr(opaque_rays, torch.int32, True), float(ert_eps),
                                    ptr(ert_hint, torch.int32, True), ptr(ert_stats, torch.int32, True), stream_ptr()))
    return (wb, den) if want_density else wb


def fine_sample(weights, Nf, idx=None, jitter=None, seed=0, ray_id_offset=0, want_idx=False, seed_dev=None):
    """weights (...,Nc) -> sorted bins (...,Nf).  idx+jitter given -> parity mode, else Philox with key
    ``seed`` (+ ``seed_dev``: a one-element int64 device tensor, for renders replayed from a CUDA graph)."""
    Nc = weights.shape[-1]
    R = weights.numel() // Nc
    bins = torch.empty(*weights.shape[:-1], Nf, device=weights.device)
    idx_out = torch.empty(*weights.shape[:-1], Nf, device=weights.device, dtype=torch.int32) if want_idx else None
    if idx is not None:
        idx = idx.to(torch.int32).contiguous()
        jitter = jitter.to(torch.float32).contiguous()
        mode = _abi.RNG_GIVEN
    else:
        mode = _abi.RNG_PHILOX
    check(lib().enarf_fine_sample(ptr(weights), R, Nc, Nf, mode, ptr(idx, torch.int32, True), ptr(jitter, allow_none=True),
                                  seed, ptr(seed_dev, torch.int64, True), ray_id_offset, ptr(bins),
                                  ptr(idx_out, torch.int32, True), stream_ptr()))
    return (bins, idx_out) if want_idx else bins


def density_query(scene, points, want_color=True, want_mask=False):
    """points (B,3,N) -> density (B,1,N), color (B,3,N) | None, part mask (B,N) int32 | None."""
    B, _, N = points.shape
    dev = points.device
    den = torch.empty(B, 1, N, device=dev)
    col = torch.empty(B, 3, N, device=dev) if want_color else None
    msk = torch.empty(B, N, device=dev, dtype=torch.int32) if want_mask else None
    check(lib().enarf_density_query(scene.ref, ptr(points), N, ptr(den), ptr(col, allow_none=True),
                                    ptr(msk, torch.int32, True), stream_ptr()))
    return den, col, msk


PAIR_CKPT = True              # the forward also keeps per-(point, part) records / features for the backward
PAIR_CKPT_PER_POINT = 2.0     # capacity in pairs per launched sample point (tiles that do not fit are recomputed)
PAIR_CKPT_MAX = 48 << 20      # pairs (208 B each)


class PairCheckpoint:
    """Caller-owned buffers of enarf_pair_ckpt_t (include/enarf_b200.h) for one fine pass."""

    def __init__(self, B, n, Nf, device, skip_last=False):
        tiles = lib().enarf_fine_tiles_per_sample(n, Nf)
        self.skip_last = bool(skip_last)
        self.capacity = int(max(1, min(PAIR_CKPT_MAX, PAIR_CKPT_PER_POINT * B * n * Nf)))
        self.rec = torch.empty(self.capacity, 20, device=device)
        self.feat = torch.empty(self.capacity, FEAT_DIM, device=device)
        self.tile_base = torch.empty(max(1, B * tiles), device=device, dtype=torch.int32)
        self.cursor = torch.zeros(1, device=device, dtype=torch.int32)
        self.count = torch.empty(B, n, Nf, device=device, dtype=torch.uint8)     # valid parts per sample
        self.c = _abi.PairCkpt(self.rec.data_ptr(), self.feat.data_ptr(), self.tile_base.data_ptr(),
                               self.cursor.data_ptr(), self.capacity, int(self.skip_last), 0, self.count.data_ptr())

    @property
    def ref(self):
        return ctypes.byref(self.c)

    def tensors(self):
        return self.rec, self.feat, self.tile_base, self.cursor, self.count

    @classmethod
    def from_tensors(cls, rec, feat, tile_base, cursor, count, skip_last=False):
        self = cls.__new__(cls)
        self.rec, self.feat, self.tile_base, self.cursor, self.capacity = rec, feat, tile_base, cursor, rec.shape[0]
        self.count = count
        self.skip_last = bool(skip_last)
        self.c = _abi.PairCkpt(rec.data_ptr(), feat.data_ptr(), tile_base.data_ptr(), cursor.data_ptr(), self.capacity,
                               int(self.skip_last), 0, count.data_ptr())
        return self


def _render_fine_fwd(scene, dmin, dmax, direction, bins, want_points=False, want_feat=False, want_pairs=False,
                     ert_eps=0.0, ert_hint=None, ert_stats=None, bg_const=0.0, skip_last=False):
    """-> color, mask, disp, fine_weights, fine_depth [, points (B,n,Nf,4)] [, feat (B,n,Nf,32) checkpoint: only
    the rows of samples inside a part are written] [, PairCheckpoint]."""
    B, n, Nf = bins.shape
    dev = bins.device
    color = torch.empty(B, 3, n, device=dev)
    mask = torch.empty(B, n, device=dev)
    disp = torch.empty(B, n, device=dev)
    fw = torch.empty(B, n, Nf - 1, device=dev)
    fd = torch.empty(B, n, Nf, device=dev)
    pts = torch.empty(B, n, Nf, 4, device=dev) if (want_points or ert_eps > 0) else None
    feat = torch.empty(B, n, Nf, FEAT_DIM, device=dev) if want_feat else None
    # skip_last (with the pair checkpoint only): the last sample of every ray -- evaluated and then dropped by the
    # reference, rendering.py:306-308 -- is left out; its row of `points` is 0
    pairs = PairCheckpoint(B, n, Nf, dev, skip_last=skip_last) if want_pairs else None
    check(lib().enarf_render_fine_fwd(scene.ref, ptr(dmin), ptr(dmax), ptr(direction), ptr(bins), n, Nf, ptr(color),
                                      ptr(mask), ptr(disp), ptr(fw), ptr(fd), ptr(pts, allow_none=True),
                                      ptr(feat, allow_none=True), pairs.ref if pairs else None, float(ert_eps),
                                      ptr(ert_hint, torch.int32, True), ptr(ert_stats, torch.int32, True), float(bg_const),
                                      stream_ptr()))
    out = (color, mask, disp, fw, fd)
    if want_points:
        out += (pts,)
    if want_feat:
        out += (feat,)
    if want_pairs:
        out += (pairs,)
    return out


def _fine_forward(ctx, scene, dmin, dmax, direction, bins, ert):
    """Forward of both autograd Functions: run the fused pass with the checkpoints a backward wants and stash them.
    ``ert`` = (ert_eps, ert_hint, ert_stats[, bg_const[, lean]]): early ray termination, the constant background, and
    ``lean`` = the caller does not read the per-sample outputs (training): the last sample of every ray is not evaluated."""
    pairs_on = SAVE_FEATURES and PAIR_CKPT
    lean = bool(ert[4]) if len(ert) > 4 else False
    out = _render_fine_fwd(scene, dmin, dmax, direction, bins, want_points=True, want_feat=SAVE_FEATURES,
                           want_pairs=pairs_on, ert_eps=ert[0], ert_hint=ert[1], ert_stats=ert[2],
                           bg_const=ert[3] if len(ert) > 3 else 0.0, skip_last=lean and pairs_on)
    ctx.skip_last = lean and pairs_on
    ctx.scene = scene
    ctx.ert_eps = float(ert[0])
    ctx.bg_const = float(ert[3]) if len(ert) > 3 else 0.0
    saved = [dmin, dmax, direction, bins, out[5]]
    if SAVE_FEATURES:
        saved.append(out[6])
    if pairs_on:
        saved.extend(out[7].tensors())
    ctx.has_feat, ctx.has_pairs = SAVE_FEATURES, pairs_on
    ctx.save_for_backward(*saved)
    out = out[:6]
    ctx.mark_non_differentiable(out[4], out[5])
    return out


def _fine_backward(ctx, grads, d_feat, d_prob, d_prob_stride):
    """Backward of both: launches enarf_render_fine_bwd into the caller's zero-initialised plane-gradient buffers and
    returns (dW1, dW2, dW3, db1, db2, db3) shaped like the decoder tensors the forward was given."""
    scene = ctx.scene
    dmin, dmax, direction, bins, pts, *rest = ctx.saved_tensors
    feat = rest.pop(0) if ctx.has_feat else None
    pairs = PairCheckpoint.from_tensors(*rest, skip_last=ctx.skip_last) if ctx.has_pairs else None
    _, n, Nf = bins.shape
    g = [None if t is None else t.contiguous().float() for t in grads]
    dw = [torch.zeros_like(w) for w in scene.w]
    db = [torch.zeros_like(b) for b in scene.b]
    check(lib().enarf_render_fine_bwd(
        scene.ref, ptr(dmin), ptr(dmax), ptr(direction), ptr(bins), ptr(pts), ptr(feat, allow_none=True),
        pairs.ref if pairs else None, n, Nf,
        ptr(g[0], allow_none=True), ptr(g[1], allow_none=True), ptr(g[2], allow_none=True),
        ptr(g[3], allow_none=True), ptr(d_feat), d_prob.data_ptr(), d_prob_stride,
        ptr(dw[0]), ptr(dw[1]), ptr(dw[2]), ptr(db[0]), ptr(db[1]), ptr(db[2]), ctx.ert_eps, ctx.bg_const,
        stream_ptr()))
    return (dw[0], dw[1], dw[2], db[0].reshape(ctx_shape(scene, 0)), db[1].reshape(ctx_shape(scene, 1)),
            db[2].reshape(ctx_shape(scene, 2)))


class RenderFine(torch.autograd.Function):
    """Fused fine pass.  Differentiable inputs: tri_plane (NCHW, features and probabilities), the three
    W_eff and the three biases; everything else (pose, rays, bins) carries no gradient, as in the
    reference (rendering.py:216-217)."""

    @staticmethod
    def forward(ctx, scene, dmin, dmax, direction, bins, tri_plane, w1, w2, w3, b1, b2, b3, ert=(0.0, None, None)):
        ctx.tp_shape = tri_plane.shape
        return _fine_forward(ctx, scene, dmin, dmax, direction, bins, ert)

    @staticmethod
    def backward(ctx, g_color, g_mask, g_disp, g_fw, _g_fd, _g_pts):
        scene = ctx.scene
        B, dev = scene.B, ctx.saved_tensors[0].device
        need_tp = ctx.needs_input_grad[5]
        d_feat = torch.zeros(B, 3, scene.H, scene.W, FEAT_DIM, device=dev)
        if need_tp:
            # the feature channels of d_tp are overwritten by unpack_plane_grads: only the probability channels
            # (scattered into directly) and the channels-last scratch need zero-filling.  Per sample: each slice is
            # contiguous, so the fill is the vectorised one (the strided view of the whole batch falls back to torch's
            # generic element-wise fill: 78 us instead of ~25 us for 151 MB)
            d_tp = torch.empty(ctx.tp_shape, device=dev)
            for b in range(d_tp.shape[0]):
                d_tp[b, FEAT_DIM * 3:].zero_()
            d_prob, stride = d_tp[:, FEAT_DIM * 3:], d_tp.stride(0)
        else:
            # only the decoder wants gradients: the kernel still needs somewhere to scatter d(probability), but not
            # a full NCHW tri-plane gradient
            d_tp = None
            d_prob = torch.zeros(B, ctx.tp_shape[1] - FEAT_DIM * 3, scene.H, scene.W, device=dev)
            stride = d_prob.stride(0)
        dec = _fine_backward(ctx, (g_color, g_mask, g_disp, g_fw), d_feat, d_prob, stride)
        if need_tp:
            unpack_plane_grads(d_feat, d_tp, accumulate=False)
        return (None, None, None, None, None, d_tp) + dec + (None,)


class RenderFinePlanes(torch.autograd.Function):
    """Fused fine pass on handed-over planes: feat_planes (B,3,H,W,32) channels-last (fp32 / bf16) and prob_planes
    (B or 1, 3P, H, W) NCHW.  The gradients come back in the same layouts -- d_feat exactly as the kernel
    accumulates it (no unpack), d_prob summed over the batch when the planes are shared."""

    @staticmethod
    def forward(ctx, scene, dmin, dmax, direction, bins, feat_planes, prob_planes, w1, w2, w3, b1, b2, b3,
                ert=(0.0, None, None)):
        ctx.prob_shape = prob_planes.shape
        return _fine_forward(ctx, scene, dmin, dmax, direction, bins, ert)

    @staticmethod
    def backward(ctx, g_color, g_mask, g_disp, g_fw, _g_fd, _g_pts):
        scene = ctx.scene
        dev = ctx.saved_tensors[0].device
        d_feat = torch.zeros(scene.B, 3, scene.H, scene.W, FEAT_DIM, device=dev)
        d_prob = torch.zeros(ctx.prob_shape, device=dev)
        stride = 0 if ctx.prob_shape[0] == 1 else d_prob.stride(0)
        dec = _fine_backward(ctx, (g_color, g_mask, g_disp, g_fw), d_feat, d_prob, stride)
        return (None, None, None, None, None, d_feat if ctx.needs_input_grad[5] else None,
                d_prob if ctx.needs_input_grad[6] else None) + dec + (None,)


def ctx_shape(scene, i):
    return scene.b_shapes[i] if hasattr(scene, "b_shapes") else scene.b[i].shape


def render_fine(scene, dmin, dmax, direction, bins, tri_plane=None, decoder=None, want_points=None, ert_eps=0.0,
                ert_hint=None, ert_stats=None, bg_const=0.0):
    """-> color (B,3,n), mask (B,n), disparity (B,n), fine_weights (B,n,Nf-1), fine_depth (B,n,Nf)
    [, points (B,n,Nf,4) = per-sample (r,g,b,sigma) when ``want_points`` is given (None/False/True)].
    Pass the autograd-tracked ``tri_plane`` and ``decoder`` tensors to get gradients.  ``bg_const``: colour comes out
    composited over a constant background, rgb + (1 - mask) * bg_const."""
    if tri_plane is None or not torch.is_grad_enabled():
        out = _render_fine_fwd(scene, dmin, dmax, direction, bins, want_points=bool(want_points), ert_eps=ert_eps,
                               ert_hint=ert_hint, ert_stats=ert_stats, bg_const=bg_const)
        if ert_eps > 0 and not want_points:
            out = out[:5]                       # the point records were only needed to mark the dropped samples
        return out if want_points is None or want_points else out + (None,)
    (w1, b1), (w2, b2), (w3, b3) = decoder
    scene.b_shapes = [b1.shape, b2.shape, b3.shape]
    lean = not want_points        # nobody reads the per-sample outputs: the rays' last samples need not be evaluated
    if isinstance(tri_plane, (tuple, list)):        # (feat_planes channels-last, prob_planes NCHW)
        out = RenderFinePlanes.apply(scene, dmin, dmax, direction, bins, tri_plane[0], tri_plane[1], w1, w2, w3, b1, b2, b3,
                                     (ert_eps, ert_hint, ert_stats, bg_const, lean))
        return out[:5] if want_points is None else (out if want_points else out[:5] + (None,))
    out = RenderFine.apply(scene, dmin, dmax, direction, bins, tri_plane, w1, w2, w3, b1, b2, b3,
                           (ert_eps, ert_hint, ert_stats, bg_const, lean))
    return out[:5] if want_points is None else (out if want_points else out[:5] + (None,))


# ---- DSO step around the renderer ------------------------------------------------------------------
def select_rays(mask, ray_batchsize, noise=None, seed=0, radius=64):
    """mask_based_sampler (libraries/NeRF/ray_sampler.py:7-39) on the device: dilate the (B,h,w) foreground mask by a
    (2*radius+1)^2 max-pool, add U[0,1) noise (``noise`` (B,h,w) given, else in-kernel Philox keyed by ``seed``) and keep
    the ``ray_batchsize`` best pixels of every image.  -> (ray_idx (B,k) int64 flat pixel ids in ascending order,
    homo_img (B,3,k) pixel centres)."""
    B, h, w = mask.shape
    m = mask.to(torch.float32).contiguous()
    scratch, score = torch.empty_like(m), torch.empty_like(m)
    if noise is not None:
        noise = noise.to(torch.float32).reshape(B, h, w).contiguous()
    check(lib().enarf_ray_scores(ptr(m), B, h, w, int(radius), ptr(noise, allow_none=True), int(seed), ptr(scratch), ptr(score),
                                 stream_ptr()))
    idx = torch.empty(B, ray_batchsize, device=mask.device, dtype=torch.int64)
    homo = torch.empty(B, 3, ray_batchsize, device=mask.device)
    check(lib().enarf_select_rays(ptr(score), B, h * w, w, int(ray_batchsize), ptr(idx, torch.int64), ptr(homo), stream_ptr()))
    return idx, homo


class _PhotometricLoss(torch.autograd.Function):
    """(colour loss, mask loss) of libraries/NeRF/loss.py:14-48 with the gather of the ground truth fused in; the
    gradients with respect to the rendered colour / mask are produced by the same launch and scaled in the backward."""

    @staticmethod
    def forward(ctx, r_color, r_mask, gt_color, gt_mask, ray_idx, loss_type, color_coef, mask_coef):
        B, _, k = r_color.shape
        npix = gt_color.shape[-1] * gt_color.shape[-2] if gt_color.ndim == 4 else gt_color.shape[-1]
        rc = r_color.detach().to(torch.float32).contiguous()
        rm = r_mask.detach().to(torch.float32).contiguous() if r_mask is not None else None
        gc = gt_color.detach().to(torch.float32).reshape(B, 3, npix).contiguous()
        gm = gt_mask.detach().to(torch.float32).reshape(B, npix).contiguous() if gt_mask is not None else None
        loss = torch.zeros(2, device=rc.device)
        d_color = torch.empty_like(rc)
        d_mask = torch.empty_like(rm) if gm is not None else None
        check(lib().enarf_photometric_loss(ptr(rc), ptr(rm, allow_none=True), ptr(gc), ptr(gm, allow_none=True),
                                           ptr(ray_idx.contiguous(), torch.int64), B, k, npix, int(loss_type), float(color_coef),
                                           float(mask_coef), ptr(loss), ptr(d_color), ptr(d_mask, allow_none=True), stream_ptr()))
        ctx.save_for_backward(d_color, d_mask if d_mask is not None else torch.empty(0, device=rc.device))
        ctx.has_mask = d_mask is not None
        return loss[0], loss[1]

    @staticmethod
    def backward(ctx, g_c, g_m):
        d_color, d_mask = ctx.saved_tensors
        return d_color * g_c, (d_mask * g_m if ctx.has_mask else None), None, None, None, None, None, None


def photometric_loss(ray_idx, rendered_color, rendered_mask, gt_color, gt_mask, loss_type="mse", color_coef=1.0, mask_coef=1.0):
    """-> (loss_color, loss_mask) as 0-d tensors (loss_mask is 0 when ``gt_mask`` is None)."""
    code = {"mse": 0, "mae": 1}[loss_type]
    return _PhotometricLoss.apply(rendered_color, rendered_mask, gt_color, gt_mask, ray_idx, code, color_coef, mask_coef)
