"""CPU ORACLE for the articulated tri-plane renderer hot path.  TEST INFRASTRUCTURE ONLY.

This file is the checker, never the product: only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it.  The product path
(``enarf-gan_b200/``) must not, and fails loudly when its CUDA library is missing.

It restates the reference algorithm (nogu-atsu/ENARF-GAN @ca46d03) with plain torch-CPU elementwise
ops so that (i) every operation has one defined rounding (no matmul / grid_sample / cumsum whose
accumulation order is a library detail), (ii) it runs in fp32 or fp64, (iii) autograd yields the
gradients.  Each function cites the reference file:line it follows.

Pinning: the reference ships no tests or golden vectors, so this oracle is
pinned by fixtures produced by running the reference's own Python code on CPU in the authoring
container (``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``) and re-checked by
``tests/test_oracle_vs_golden.py``.

ARITHMETIC CONTRACT for the bit-exact quantities (ray validity, part-selection masks, sample bin
indices): fp32, round-to-nearest, every product and sum rounded separately (no FMA contraction),
sums associated left to right exactly as written below.  The CUDA kernels use __fmul_rn/__fadd_rn
in the same order.
"""
import math

import torch

SQRT3 = 3 ** 0.5
N_PROBE = 32          # rendering.py:18
LARGE = 1e3           # rendering.py:54
NEAR_DEFAULT = 0.3    # rendering.py:249
FAR_DEFAULT = 5.0     # rendering.py:138 (far_plane default)


# --------------------------------------------------------------------------------------------
# a2: joint frames -> part frames                       libraries/NARF/pose_utils.py:129-148
# --------------------------------------------------------------------------------------------
def transform_pose(pose_to_camera, bone_length, origin_location, parent_id):
    """(B,24,4,4),(B,23,1) -> (B,P,4,4),(B,P,1).  Rotation of the parent (or own for "center"),
    translation = midpoint of joint and parent; "center+head" appends joint 15 with length 1."""
    par = torch.as_tensor(parent_id[1:], dtype=torch.long)
    child_t = pose_to_camera[:, 1:, :, 3:]
    parent_t = pose_to_camera[:, par, :, 3:]
    mid = (child_t + parent_t) / 2
    if origin_location == "center":
        out = torch.cat([pose_to_camera[:, 1:, :, :3], mid], dim=-1)
    elif origin_location == "center_fixed":
        out = torch.cat([pose_to_camera[:, par, :, :3], mid], dim=-1)
    elif origin_location == "center+head":
        body = torch.cat([pose_to_camera[:, par, :, :3], mid], dim=-1)
        out = torch.cat([body, pose_to_camera[:, 15:16]], dim=1)
        bone_length = torch.cat([bone_length, torch.ones_like(bone_length[:, :1])], dim=1)
    else:
        raise ValueError(origin_location)
    return out, bone_length


def canonical_parts(canonical_pose, parent_id, origin_location):
    """models/narf.py:84-120 -> (canonical part frames (P,4,4), canonical bone length (P,))."""
    pose = torch.as_tensor(canonical_pose, dtype=torch.float64)
    par = torch.as_tensor(parent_id[1:], dtype=torch.long)
    coord = pose[:, :3, 3]
    length = (coord[1:] - coord[par]).norm(dim=1)
    mid = (pose[1:, :, 3:] + pose[par, :, 3:]) / 2
    if origin_location == "center":
        parts = torch.cat([pose[1:, :, :3], mid], dim=-1)
    elif origin_location == "center_fixed":
        parts = torch.cat([pose[par, :, :3], mid], dim=-1)
    elif origin_location == "center+head":
        parts = torch.cat([torch.cat([pose[par, :, :3], mid], dim=-1), pose[15:16]], dim=0)
        length = torch.cat([length, torch.ones(1, dtype=length.dtype)])
    else:
        raise ValueError(origin_location)
    return parts.float(), length.float()


def scale_pose(pose, coordinate_scale):
    """rendering.py:258-260: translations times coordinate_scale."""
    pose = pose.clone()
    pose[:, :, :3, 3] = pose[:, :, :3, 3] * coordinate_scale
    return pose


# --------------------------------------------------------------------------------------------
# a3: depth range                                        libraries/NeRF/rendering.py:10-79
# --------------------------------------------------------------------------------------------
def near_far(pose, near_plane=NEAR_DEFAULT, far_plane=FAR_DEFAULT):
    """Batch-global near/far (rendering.py:15-17).  Returns two 0-d tensors of pose.dtype."""
    tz = pose[:, :, 2, 3]
    near = torch.clamp_min(tz.min() - SQRT3, near_plane)
    far = torch.clamp_min(tz.max() + SQRT3, far_plane)
    return near, far


def linspace_contract(start, end, steps, dtype):
    """torch.linspace's CPU formula written out: step=(end-start)/(steps-1); first half counts up
    from start, second half counts down from end (each product / sum rounded separately)."""
    start = torch.as_tensor(start, dtype=dtype)
    end = torch.as_tensor(end, dtype=dtype)
    step = (end - start) / torch.tensor(float(steps - 1), dtype=dtype)
    i = torch.arange(steps, dtype=dtype)
    up = start + step * i
    down = end - step * (float(steps - 1) - i)
    return torch.where(i < steps // 2, up, down)


def to_local(points, pose):
    """local = R^T (x - t) per part (models/narf.py:158-162, NeRF/utils.py:13-32).
    points (B,3,N), pose (B,P,4,4) -> (B,P,3,N).  Contract order: ((R0i*d0 + R1i*d1) + R2i*d2)."""
    R = pose[:, :, :3, :3]
    t = pose[:, :, :3, 3]
    d = points[:, None] - t[..., None]                      # (B,P,3,N)
    out = []
    for i in range(3):
        li = (R[:, :, 0, i, None] * d[:, :, 0] + R[:, :, 1, i, None] * d[:, :, 1]) + R[:, :, 2, i, None] * d[:, :, 2]
        out.append(li)
    return torch.stack(out, dim=2)


def depth_range(image_coord, pose, inv_intrinsics, near, far):
    """image_coord (B,3,n) homogeneous pixels, pose (B,P,4,4) already scaled, inv_intrinsics (B,3,3).
    Returns depth_min, depth_max (B,n), ray_direction (B,3,n), validity (B,n) bool."""
    dt = pose.dtype
    K = inv_intrinsics
    if K.ndim == 2:
        K = K[None].expand(image_coord.shape[0], 3, 3)
    u = image_coord
    direction = torch.stack(
        [(K[:, i, 0, None] * u[:, 0] + K[:, i, 1, None] * u[:, 1]) + K[:, i, 2, None] * u[:, 2] for i in range(3)],
        dim=1)                                              # (B,3,n)
    probes = linspace_contract(near, far, N_PROBE, dt)      # (32,)
    B, _, n = direction.shape
    pts = (direction[..., None] * probes).reshape(B, 3, n * N_PROBE)
    local = to_local(pts, pose)                             # (B,P,3,n*32)
    inside = (local.abs() <= 1).all(dim=2).reshape(B, -1, n, N_PROBE)
    big = torch.tensor(LARGE, dtype=dt)
    dmin = torch.where(inside, probes, big).amin(dim=3).amin(dim=1)      # (B,n)
    dmax = torch.where(inside, probes, -big).amax(dim=3).amax(dim=1)
    validity = dmin != big
    dmin = torch.where(validity, dmin, torch.as_tensor(near, dtype=dt))
    dmax = torch.where(dmax != -big, dmax, torch.as_tensor(far, dtype=dt))
    dmin = torch.maximum(dmin, torch.as_tensor(near, dtype=dt))
    return dmin, dmax, direction, validity


# --------------------------------------------------------------------------------------------
# a4: coarse sampling                                    libraries/NeRF/rendering.py:119-133
# --------------------------------------------------------------------------------------------
def lerp_contract(a, b, t):
    """a*(1-t) + b*t with each op rounded separately (rendering.py:126-129,198-200)."""
    return a * (1 - t) + b * t


def coarse_sample(dmin, dmax, direction, Nc):
    """dmin,dmax (B,n); direction (B,3,n) -> coarse_depth (B,n,Nc+1), coarse_points (B,3,n,Nc),
    start,end (B,3,n)."""
    dt = dmin.dtype
    start = dmin[:, None] * direction
    end = dmax[:, None] * direction
    bins = linspace_contract(0.0, 1.0, Nc + 1, dt)
    depth = lerp_contract(dmin[..., None], dmax[..., None], bins)
    raw = lerp_contract(start[..., None], end[..., None], bins)           # (B,3,n,Nc+1)
    points = (raw[..., 1:] + raw[..., :-1]) / 2
    return depth, points, start, end


# --------------------------------------------------------------------------------------------
# a6: part transform + validity                          models/narf.py:147-174,198-201,237
# --------------------------------------------------------------------------------------------
def canonical_scale(canonical_bone_length, bone_length, coordinate_scale):
    """(P,),(B,P,1) -> (B,P): L_c / L / coordinate_scale (narf.py:165)."""
    return (canonical_bone_length[None, :, None] / bone_length / coordinate_scale)[:, :, 0]


def part_transform(points, pose, cscale, canonical_pose):
    """points (B,3,N), pose (B,P,4,4) scaled, cscale (B,P), canonical_pose (P,4,4)
    -> local (B,P,3,N), canon (B,P,3,N), valid (B,P,N) bool.
    valid = all|local|<=1 and all|canon|<1 (NeRF/utils.py:39-43, narf.py:200-201)."""
    local = to_local(points, pose)
    s = local * cscale[:, :, None, None]
    Rc = canonical_pose[:, :3, :3].to(points.dtype)
    tc = canonical_pose[:, :3, 3].to(points.dtype)
    canon = torch.stack(
        [((Rc[None, :, i, 0, None] * s[:, :, 0] + Rc[None, :, i, 1, None] * s[:, :, 1])
          + Rc[None, :, i, 2, None] * s[:, :, 2]) + tc[None, :, i, None] for i in range(3)], dim=2)
    valid = (local.abs() <= 1).all(dim=2) & (canon.abs() < 1).all(dim=2)
    return local, canon, valid


# --------------------------------------------------------------------------------------------
# a7/a8/a9: bilinear tri-plane gather    libraries/triplane/sampling.py:9-127,
#                                        cuda_extension/TriplaneSampler_kernel.cu:13-92
# --------------------------------------------------------------------------------------------
def bilinear_zeros(plane, x, y):
    """grid_sample(bilinear, zeros padding, align_corners=False) written out.
    plane (C,H,W); x,y (N,) in [-1,1] (x -> width, y -> height) -> (C,N)."""
    C, H, W = plane.shape
    ix = ((x + 1) * W - 1) / 2
    iy = ((y + 1) * H - 1) / 2
    x0 = torch.floor(ix)
    y0 = torch.floor(iy)
    out = 0
    flat = plane.reshape(C, H * W)
    for dy in (0, 1):
        for dx in (0, 1):
            xi = x0 + dx
            yi = y0 + dy
            wx = (ix - x0) if dx else (x0 + 1 - ix)
            wy = (iy - y0) if dy else (y0 + 1 - iy)
            ok = (xi >= 0) & (xi <= W - 1) & (yi >= 0) & (yi <= H - 1)
            idx = (yi.clamp(0, H - 1) * W + xi.clamp(0, W - 1)).long()
            out = out + flat[:, idx] * (wx * wy * ok.to(plane.dtype))
    return out


PLANE_AXES = ((0, 1), (1, 2), (2, 0))   # plane k samples (x=p[k], y=p[(k+1)%3]); sampling.py:30


def triplane_sample_sum(tri_plane, position):
    """The raw ``triplane_sampler`` op (bilinear/zeros/align_corners=False): tri_plane (3C,H,W),
    position (3,N) -> (C,N) summed over the three planes (TriplaneSampler_kernel.cu:36-74)."""
    C = tri_plane.shape[0] // 3
    out = 0
    for k, (ax, ay) in enumerate(PLANE_AXES):
        out = out + bilinear_zeros(tri_plane[k * C:(k + 1) * C], position[ax], position[ay])
    return out


def warp_triplane(tri_plane, flow, feat_dim=32):
    """Deformation-field tri-plane producer of the DSO model (models/narf.py:43-57).
    tri_plane (1,(feat_dim+P)*3,S,S) canonical planes; flow (B,6,S,S) = the flow head's output, channel 2k / 2k+1
    the x / y offset (in texels) of plane k.  grid = (pixel centre + flow) / (S/2) - 1 (narf.py:49; the reference
    writes the 128 of S = 256), every feature plane is resampled there (bilinear, zeros, align_corners=False), the
    probability planes are passed through unchanged.  -> (B,(feat_dim+P)*3,S,S)."""
    B = flow.shape[0]
    S = tri_plane.shape[-1]
    dt = tri_plane.dtype
    ar = torch.arange(S, dtype=dt) + 0.5
    gx = ar[None, :].expand(S, S).reshape(-1)
    gy = ar[:, None].expand(S, S).reshape(-1)
    fl = flow.to(dt).reshape(B, 3, 2, S * S)
    planes = []
    for b in range(B):
        for k in range(3):
            x = (gx + fl[b, k, 0]) / (S / 2) - 1
            y = (gy + fl[b, k, 1]) / (S / 2) - 1
            planes.append(bilinear_zeros(tri_plane[0, k * feat_dim:(k + 1) * feat_dim], x, y))
    warped = torch.stack(planes).reshape(B, 3 * feat_dim, S, S)
    return torch.cat([warped, tri_plane[:, 3 * feat_dim:].expand(B, -1, -1, -1)], dim=1)


def gather_features(tri_plane, canon, valid, feat_dim=32):
    """tri_plane (B,(feat_dim+P)*3,H,W); canon (B,P,3,N); valid (B,P,N)
    -> feature (B,feat_dim,N), part weight (B,P,N) (zero where invalid).
    weight = prod_k sigmoid(prob plane sample) (sampling.py:48,60-64); feature = sum over valid
    parts of weight * sum_k bilinear(feature plane k) (sampling.py:110-123).  Invalid parts are at
    coordinate 2 in the reference (narf.py:237) and contribute exactly zero feature."""
    B, P, _, N = canon.shape
    feat = torch.zeros(B, feat_dim, N, dtype=canon.dtype)
    weight = torch.zeros(B, P, N, dtype=canon.dtype)
    for b in range(B):
        fplanes = tri_plane[b, :feat_dim * 3]
        for p in range(P):
            sel = valid[b, p].nonzero()[:, 0]
            if sel.numel() == 0:
                continue
            pos = canon[b, p][:, sel]
            w = 1
            for k, (ax, ay) in enumerate(PLANE_AXES):
                ch = feat_dim * 3 + 3 * p + k
                w = w * torch.sigmoid(bilinear_zeros(tri_plane[b, ch:ch + 1], pos[ax], pos[ay])[0])
            f = triplane_sample_sum(fplanes, pos)
            feat[b] = feat[b].index_add(1, sel, f * w)
            weight[b, p] = weight[b, p].index_add(0, sel, w)
    return feat, weight


# --------------------------------------------------------------------------------------------
# a10: decoder                       libraries/custom_stylegan2/net.py:128-179,194-254,270-320
#                                    libraries/triplane/triplane_nerf.py:32-48, NeRF/activation.py
# --------------------------------------------------------------------------------------------
class _LeakyBackwardReLU(torch.autograd.Function):
    """MyReLU (NeRF/activation.py:5-16): relu forward; backward g*[x>=0] + 0.1*g*[x<0]*[g<0]."""

    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return x.clamp_min(0)

    @staticmethod
    def backward(ctx, g):
        x, = ctx.saved_tensors
        return g * (x >= 0) + g * (x < 0) * (g < 0) * 0.1


def effective_weights(layers, z_rend):
    """Per-sample modulated + normalised decoder weights.
    ``layers``: list of dicts {mod_weight (in,style), mod_bias (in,), weight (out,in), bias (out,)}.
    s = z W_mod^T/sqrt(style) + b_mod (EqualLinear, net.py:161-174); W = normalize_in(W*s/sqrt(in))
    (net.py:213,236,242-243).  Returns [(W_eff (B,out,in), bias (out,)), ...]."""
    out = []
    for L in layers:
        style_dim = L["mod_weight"].shape[1]
        in_ch = L["weight"].shape[1]
        s = torch.nn.functional.linear(z_rend, L["mod_weight"] * (1 / math.sqrt(style_dim)), L["mod_bias"])
        w = (1 / math.sqrt(in_ch)) * L["weight"][None] * s[:, None, :]
        w = w / w.norm(dim=-1, keepdim=True).clamp_min(1e-12)
        out.append((w, L["bias"]))
    return out


def decoder(feature, eff):
    """feature (B,32,N) -> color (B,3,N), raw density (B,1,N).  Every layer, the last included, is
    conv + bias -> LeakyReLU(0.2) * sqrt(2) (net.py:311,317-318); rgb = tanh (triplane_nerf.py:46)."""
    h = feature
    for w, b in eff:
        h = torch.einsum("boi,bin->bon", w, h) + b[None, :, None]
        h = torch.nn.functional.leaky_relu(h, 0.2) * (2 ** 0.5)
    return torch.tanh(h[:, :3]), h[:, 3:4]


def density_color(points, pose, cscale, canonical_pose, tri_plane, eff, feat_dim=32):
    """calc_density_and_color_from_camera_coord_v2 (models/narf.py:176-211, 213-275).
    Returns density (B,1,N), color (B,3,N), valid (B,P,N), part weight (B,P,N)."""
    _, canon, valid = part_transform(points, pose, cscale, canonical_pose)
    feat, weight = gather_features(tri_plane, canon, valid, feat_dim)
    color, raw = decoder(feat, eff)
    density = _LeakyBackwardReLU.apply(raw) * 10                              # narf.py:274
    density = density * valid.any(dim=1, keepdim=True).to(density.dtype)     # narf.py:204
    return density, color, valid, weight


# --------------------------------------------------------------------------------------------
# a5: coarse weights -> fine bins                        libraries/NeRF/rendering.py:174-200
# --------------------------------------------------------------------------------------------
def seq_cumsum(x):
    """Inclusive prefix sum along the last dim, strictly sequential in x.dtype."""
    out = [x[..., 0]]
    for j in range(1, x.shape[-1]):
        out.append(out[-1] + x[..., j])
    return torch.stack(out, dim=-1)


def ray_weights(density, depth, render_scale=1.0):
    """density (...,N), depth (...,N+1) -> alpha-compositing weights (...,N)
    (rendering.py:178-184 and 313-321): dd = sigma*delta; T = exp(-(cumsum(dd)-dd))."""
    delta = depth[..., 1:] - depth[..., :-1]
    dd = density * delta * render_scale
    T = torch.exp(-(seq_cumsum(dd) - dd))
    return T * (1 - torch.exp(-dd))


def blur_weights(w):
    """(max(w[j-1],w[j]) + max(w[j],w[j+1]))/2 + 0.01 with zero padding (rendering.py:187-190)."""
    p = torch.nn.functional.pad(w, (1, 1))
    return (torch.maximum(p[..., :-2], p[..., 1:-1]) + torch.maximum(p[..., 1:-1], p[..., 2:])) / 2 + 0.01


_PHILOX_M0, _PHILOX_M1 = 0xD2511F53, 0xCD9E8D57
_PHILOX_W0, _PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32(c0, c1, c2, c3, k0, k1, rounds=10):
    """Philox4x32-10 (Salmon et al. 2011) on int64 tensors holding uint32 values."""
    M = 0xFFFFFFFF
    c0, c1, c2, c3 = [torch.as_tensor(c, dtype=torch.int64) & M for c in (c0, c1, c2, c3)]

    def mulhilo(a, b):
        # 32x32 -> hi, lo without overflowing int64: split b into 16-bit halves
        lo = (a * (b & 0xFFFF) + ((a * (b >> 16)) & 0xFFFF) * 65536) & M
        t = a * (b & 0xFFFF)
        u = a * (b >> 16)
        hi = ((u + (t >> 16)) >> 16) & M
        return hi, lo

    for _ in range(rounds):
        h0, l0 = mulhilo(c0, _PHILOX_M0)
        h1, l1 = mulhilo(c2, _PHILOX_M1)
        c0, c1, c2, c3 = (h1 ^ c1 ^ k0) & M, l1, (h0 ^ c3 ^ k1) & M, l0
        k0 = (k0 + _PHILOX_W0) & M
        k1 = (k1 + _PHILOX_W1) & M
    return c0, c1, c2, c3


def philox_uniforms(ray_ids, Nf, seed):
    """Two U[0,1) fp32 draws per (ray, sample): counter=(ray, sample, 0, 0), key=(seed lo, seed hi);
    u = (word >> 8) * 2^-24.  Returns (u_select, u_jitter), each (R,Nf) float32."""
    r = torch.as_tensor(ray_ids, dtype=torch.int64)[:, None].expand(-1, Nf)
    s = torch.arange(Nf, dtype=torch.int64)[None].expand(r.shape[0], -1)
    z = torch.zeros_like(r)
    o0, o1, _, _ = philox4x32(r, s, z, z, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    f = lambda o: ((o >> 8).to(torch.float32) * (1.0 / 16777216.0))
    return f(o0), f(o1)


def invert_cdf(wb, u_select):
    """wb (R,Nc) blurred weights, u_select (R,Nf) -> bin index (R,Nf) int64.
    c = sequential fp32 prefix sum; target = u*c[-1]; idx = #{j < Nc-1 : c_j <= target}."""
    c = seq_cumsum(wb)
    target = u_select * c[:, -1:]
    return (c[:, None, :-1] <= target[:, :, None]).sum(dim=-1)


def fine_bins(idx, jitter, Nc):
    """idx (R,Nf) int, jitter (R,Nf) U[0,1) -> sorted bins (R,Nf): idx/Nc + jitter/Nc, ascending
    (rendering.py:192-197)."""
    b = idx.to(jitter.dtype) / Nc + jitter / Nc
    return torch.sort(b, dim=-1)[0]


def fine_sample(bins, dmin, dmax, start, end):
    """bins (B,n,Nf) -> fine_depth (B,n,Nf), fine_points (B,3,n,Nf) (rendering.py:198-200)."""
    depth = lerp_contract(dmin[..., None], dmax[..., None], bins)
    pts = lerp_contract(start[..., None], end[..., None], bins[:, None])
    return depth, pts


# --------------------------------------------------------------------------------------------
# a11: compositing                                       libraries/NeRF/rendering.py:307-335
# --------------------------------------------------------------------------------------------
def composite(density, color, fine_depth, render_scale=1.0):
    """density (B,n,Nf), color (B,3,n,Nf), fine_depth (B,n,Nf).  The last sample only closes the
    last interval.  Returns color (B,3,n), mask (B,n), disparity (B,n), weights (B,n,Nf-1)."""
    w = ray_weights(density[..., :-1], fine_depth, render_scale)
    rgb = (w[:, None] * color[..., :-1]).sum(dim=-1)
    mask = w.sum(dim=-1)
    disp = (w * 1 / fine_depth[..., :-1]).sum(dim=-1)
    return rgb, mask, disp, w


# --------------------------------------------------------------------------------------------
# a1: pixel grid                                         libraries/NeRF/ray_sampler.py:42-67
# --------------------------------------------------------------------------------------------
def whole_image_grid(render_size, patch_size, B, dtype=torch.float32):
    """Homogeneous pixel-centre coordinates (B,3,patch_size^2), row-major (x fastest)."""
    a = torch.arange(patch_size, dtype=dtype)
    c = render_size * (a + 0.5) / patch_size
    x = c[None, :].expand(patch_size, patch_size).reshape(-1)
    y = c[:, None].expand(patch_size, patch_size).reshape(-1)
    return torch.stack([x, y, torch.ones_like(x)], dim=0)[None].expand(B, 3, -1).contiguous()


# --------------------------------------------------------------------------------------------
# the whole path                                         libraries/NeRF/rendering.py:227-359
# --------------------------------------------------------------------------------------------
class Scene:
    """Everything constant across one render call, in one dtype."""

    def __init__(self, pose_to_camera, bone_length, inv_intrinsics, tri_plane, layers, z_rend,
                 canonical_pose, parent_id, origin_location="center+head", coordinate_scale=3.0,
                 dtype=torch.float32):
        c = lambda t: torch.as_tensor(t).to(dtype)
        self.dtype = dtype
        pose, bl = transform_pose(c(pose_to_camera), c(bone_length), origin_location, parent_id)
        self.pose = scale_pose(pose, coordinate_scale)
        cparts, clen = canonical_parts(canonical_pose, parent_id, origin_location)
        self.canonical_pose = cparts.to(dtype)
        self.cscale = canonical_scale(clen.to(dtype), bl, coordinate_scale)
        self.inv_intrinsics = c(inv_intrinsics)
        self.tri_plane = tri_plane if tri_plane.dtype == dtype else tri_plane.to(dtype)
        self.layers = [{k: (v if v.dtype == dtype else v.to(dtype)) for k, v in L.items()} for L in layers]
        self.z_rend = c(z_rend)
        self.eff = effective_weights(self.layers, self.z_rend)
        self.near, self.far = near_far(self.pose)

    def query(self, points):
        return density_color(points, self.pose, self.cscale, self.canonical_pose, self.tri_plane, self.eff)


def render(scene, image_coord, Nc, Nf, rng, compact_b1=True):
    """rendering.render restated.  ``rng`` is one of
    {"mode":"given","idx":(R,Nf) int,"jitter":(R,Nf)} | {"mode":"philox","seed":int,"ray_ids":(R,)} |
    {"mode":"torch"} (torch.multinomial + uniform_ on the global generator, the reference's order).
    Rays are flattened batch-major (R = B*n, or the valid rays only when B == 1 and compact_b1).
    Returns a dict with color (B,3,n), mask (B,n), disparity (B,n), fine_weights (B,n',Nf-1),
    fine_depth (B,n',Nf), validity (B,n) and the intermediates used by the stage-wise tests."""
    dt = scene.dtype
    image_coord = image_coord.to(dt)
    B, _, n_all = image_coord.shape
    dmin, dmax, direction, validity = depth_range(image_coord, scene.pose, scene.inv_intrinsics,
                                                  scene.near, scene.far)
    out = {"validity": validity, "depth_min_all": dmin, "depth_max_all": dmax, "direction_all": direction}
    compact = B == 1 and compact_b1                               # rendering.py:107-110
    if compact:
        sel = validity[0]
        dmin, dmax, direction = dmin[:, sel], dmax[:, sel], direction[:, :, sel]
    n = dmin.shape[-1]
    zeros = lambda *s: torch.zeros(*s, dtype=dt)
    if n == 0:                                                    # rendering.py:351-354
        out.update(color=zeros(B, 3, n_all), mask=zeros(B, n_all), disparity=zeros(B, n_all),
                   fine_weights=None, fine_depth=None)
        return out
    cdepth, cpts, start, end = coarse_sample(dmin, dmax, direction, Nc)
    with torch.no_grad():
        cden, _, cvalid, _ = scene.query(cpts.reshape(B, 3, n * Nc))
        cden = cden.reshape(B, n, Nc)
        wb = blur_weights(ray_weights(cden, cdepth).reshape(B * n, Nc))
    if rng["mode"] == "given":
        idx, jitter = rng["idx"], rng["jitter"].to(dt)
    elif rng["mode"] == "philox":
        us, uj = philox_uniforms(rng["ray_ids"], Nf, rng["seed"])
        idx, jitter = invert_cdf(wb.float(), us), uj.to(dt)
    elif rng["mode"] == "torch":
        idx = torch.multinomial(wb.float(), Nf, replacement=True)
        jitter = torch.empty(B, 1, n, Nf).uniform_().reshape(B * n, Nf).to(dt)
    else:
        raise ValueError(rng["mode"])
    bins = fine_bins(idx, jitter, Nc).reshape(B, n, Nf)
    fdepth, fpts = fine_sample(bins, dmin, dmax, start, end)
    fden, fcol, fvalid, fweight = scene.query(fpts.reshape(B, 3, n * Nf))
    fden = fden.reshape(B, n, Nf)
    fcol = fcol.reshape(B, 3, n, Nf)
    rgb, mask, disp, w = composite(fden, fcol, fdepth)
    if compact:                                                   # rendering.py:337-350
        full = lambda t, *s: zeros(*s).index_copy(len(s) - 1, sel.nonzero()[:, 0], t)
        rgb, mask, disp = full(rgb, B, 3, n_all), full(mask, B, n_all), full(disp, B, n_all)
    out.update(color=rgb, mask=mask, disparity=disp, fine_weights=w, fine_depth=fdepth,
               coarse_depth=cdepth, coarse_points=cpts, coarse_density=cden, coarse_valid=cvalid,
               coarse_weights_blurred=wb, idx=idx, jitter=jitter, bins=bins, fine_points=fpts,
               fine_density=fden, fine_color=fcol, fine_valid=fvalid, fine_part_weight=fweight,
               depth_min=dmin, depth_max=dmax)
    return out
