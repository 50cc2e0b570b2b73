"""Full-size parity at BASELINE configs[1..2] per-sample size (128 x 128 rays, 48+48 samples, P = 24) against
  (1) the UNMODIFIED reference running its eager ``grid_sample`` path ON THE SAME B200 (tests/golden/ref_harness.py
      imports it from oracle/_ref/reference_py.zip; its own CUDA extension is loaded as ``triplane_sampler_cuda``), and
  (2) the CPU oracle,
forward AND gradients (tri-plane feature / probability channels, every decoder parameter through the modulation
chain), with the reference's own fine-pass draws fed to the kernels.  B = 1 cases are the adjudicators (the
reference's B > 1 gather is itself only ~1e-3 accurate); the B = 2 case is held to 3e-3 against the reference and to
1e-4 against the oracle.  Part masks are compared bit for bit with the masks the reference computed on the GPU."""
import numpy as np
import pytest
import torch

import enarf_oracle as O
import ref_runner as R
from enarf_gan_b200 import synthetic as S
from parity import check_param_grads, outlier_stats, rel_max, report

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not R.available(), reason="reference archive not staged")]
NC = NF = 48


class _Planes(torch.nn.Module):
    def __init__(self, tp):
        super().__init__()
        self.tp = torch.nn.Parameter(tp)

    def forward(self, z, bone_length, truncation_psi=1):
        return self.tp[:z.shape[0]]


def _mirror(size, tp, state):
    from enarf_gan_b200.generator import TriNARFGenerator
    cfg = dict(z_dim=256, background_ratio=0.7,
               nerf_params=dict(Nc=NC, Nf=NF, origin_location="center+head", coordinate_scale=3, render_bs=16384))
    gen = TriNARFGenerator(cfg, size, num_bone=24, num_bone_param=23, parent_id=S.SMPL_PARENTS, black_background=True,
                           tri_plane_gen=_Planes(tp.clone()))
    gen.register_canonical_pose(S.canonical_pose())
    missing, unexpected = gen.load_state_dict({k: v.cpu() for k, v in state.items()}, strict=False)
    assert not unexpected
    return gen.cuda().eval()


def _oracle(ref, B, size):
    """CPU oracle on the same inputs / draws, with autograd; returns outputs + gradients."""
    inp, tp, z = ref["inputs"]
    st = ref["decoder"]
    layers = [dict(mod_weight=st[f"nerf.mlp.layers.{i}.conv.modulation.weight"].cpu().clone().requires_grad_(),
                   mod_bias=st[f"nerf.mlp.layers.{i}.conv.modulation.bias"].cpu().clone().requires_grad_(),
                   weight=st[f"nerf.mlp.layers.{i}.conv.weight"].cpu()[0, :, :, 0].clone().requires_grad_(),
                   bias=st[f"nerf.mlp.layers.{i}.bias"].cpu()[0, :, 0].clone().requires_grad_()) for i in range(3)]
    tpo = tp.detach().clone().requires_grad_()
    sc = O.Scene(inp["pose_to_camera"], inp["bone_length"], inp["inv_intrinsics"], tpo, layers, z[:, 512:768],
                 S.canonical_pose(), S.SMPL_PARENTS, origin_location="center+head", coordinate_scale=3.0)
    rng = {"mode": "given", "idx": ref["idx"].cpu(), "jitter": ref["jitter"].cpu()}
    out = O.render(sc, O.whole_image_grid(size, size, B), NC, NF, rng)
    color = out["color"] + (1 - out["mask"][:, None]) * -1.0
    loss = R.linear_loss(color, out["mask"], out["disparity"] * 3.0, out["fine_weights"], ref["loss_weights"])
    loss.backward()
    grads = {}
    for i, L in enumerate(layers):
        grads[f"nerf.mlp.layers.{i}.conv.weight"] = L["weight"].grad[None, :, :, None]
        grads[f"nerf.mlp.layers.{i}.conv.modulation.weight"] = L["mod_weight"].grad
        grads[f"nerf.mlp.layers.{i}.conv.modulation.bias"] = L["mod_bias"].grad
        grads[f"nerf.mlp.layers.{i}.bias"] = L["bias"].grad[None, :, None]
    return dict(color=color, mask=out["mask"], disparity=out["disparity"] * 3.0, fine_weights=out["fine_weights"],
                fine_depth=out["fine_depth"], validity=out["validity"], d_tri_plane=tpo.grad, param_grads=grads)


def _ours(ref, B, size):
    """The kernels through the public API (TriNARFGenerator mirror), fed the reference's parameters, inputs and draws.
    -> generator, outputs / gradients, and for every rendered ray its pixel (B == 1 renders only the valid rays)."""
    from enarf_gan_b200 import ops, rendering
    inp, tp, z = ref["inputs"]
    gen = _mirror(size, tp.detach().float(), {k: v.float() for k, v in ref["decoder"].items()})
    gen.nerf.rng_mode = "given"
    gen.nerf.rng_given = (ref["idx"], ref["jitter"].float())
    c = lambda k: inp[k].cuda()
    color, mask, disp = gen(c("pose_to_camera"), c("pose_to_world"), c("bone_length"), z.cuda(), c("inv_intrinsics"),
                            return_disparity=True)
    fw = gen.nerf.buffers_tensors["fine_weights"]
    fd = gen.nerf.buffers_tensors["fine_depth"]
    assert fw.shape == (B, 1, ref["n_rays"], NF - 1), "ray validity differs from the reference on the GPU"
    R.linear_loss(color, mask, disp, fw, ref["loss_weights"]).backward()
    scene, _, _ = gen.nerf.build_scene(*_scene_args(gen, inp, z))
    valid = ops.depth_range(rendering.pixel_grid(size, size, B, "cuda"), c("inv_intrinsics"), scene.pose)[3]
    pix = valid[0].nonzero()[:, 0][None] if B == 1 else torch.arange(size * size, device="cuda")[None].expand(B, -1)
    return gen, scene, dict(color=color, mask=mask, disparity=disp, fine_weights=fw[:, 0], fine_depth=fd[:, 0], pix=pix,
                            d_tri_plane=gen.nerf.tri_plane_gen.tp.grad,
                            param_grads={k: p.grad for k, p in gen.named_parameters() if k in ref["param_grads"]})


def _mask_flips(scene, ref):
    """Part-selection bits of the kernels at the reference's own fine points vs the bits the reference computed.
    -> (#flipped bits, #bits, (B,n) bool: rays that contain a flipped sample)."""
    from enarf_gan_b200 import ops
    pts = ref["fine_points"].float().contiguous()
    _, _, m = ops.density_query(scene, pts, want_color=False, want_mask=True)
    bits = torch.stack([(m >> p) & 1 for p in range(24)], dim=1).bool()
    diff = bits != ref["valid_parts"]
    B = diff.shape[0]
    return int(diff.sum()), diff.numel(), diff.any(dim=1).reshape(B, ref["n_rays"], NF).any(dim=2)


def _compare(tag, ours, other, B, tol, tol_grad, fails, drop_rays=None, robust=False, own=None):
    """Prints every figure first and collects the violated bars in ``fails`` (one GPU run shows the whole picture).
    ``drop_rays`` (B,n) bool: rays left out of the forward comparison (boundary flips against an fp64 run)."""
    print(f"  -- vs {tag} (forward bar {tol:g}, gradient bar {tol_grad:g})")
    pix = ours["pix"]
    keep = torch.ones_like(pix, dtype=torch.bool) if drop_rays is None else ~drop_rays
    for k in ("color", "mask", "disparity", "fine_weights"):
        a, b = ours[k], other[k]
        if k == "fine_weights":
            a, b = a[keep], b.to(a.device)[keep]
        else:       # image-shaped: pick the rendered (and kept) rays' pixels
            ch = 3 if k == "color" else 1
            sel = lambda t: torch.cat([t.reshape(B, ch, -1)[i][:, pix[i][keep[i]]] for i in range(B)], dim=1)
            a, b = sel(a), sel(b.to(a.device))
        m_, e_ = report(k, a.reshape(-1), b.reshape(-1))
        if robust:
            # Against an fp64 run every fp32 evaluation carries the rounding of its sample positions (x - t cancels
            # ~30x at these depths) through white-noise planes and the decoder; how much depends on the scene (measured
            # 2.5e-5 on the B = 2 case, 6e-4 on seed 2201).  The bar: 1e-4, or no further from the fp64 result than the
            # reference's own fp32 evaluation of the same samples (``own``, x 1.5 for the max over ~7000 rays).
            bar = max(tol, 1.5 * own[k]) if own and own[k] == own[k] else tol
            print(f"      bar {bar:.2e} (reference's own fp32 distance {own[k] if own else float('nan'):.2e})")
            if not m_ < bar:
                fails.append((tag, k, "rel_max", m_, "bar", bar))
            continue
        if not m_ < tol:
            fails.append((tag, k, "rel_max", m_))
        if not e_ < 100 * tol:                  # every element above 1 % of the max: within 100 x the bar
            fails.append((tag, k, "rel_elem", e_))
    d = rel_max(ours["fine_depth"], other["fine_depth"])
    if not d < 1e-5:
        fails.append((tag, "fine_depth", d))
    for which, sl in (("feature", slice(0, 96)), ("probability", slice(96, None))):
        a, b = ours["d_tri_plane"][:, sl].cpu(), other["d_tri_plane"][:, sl].cpu()
        nz = (b != 0) | (a != 0)
        frac, q999, mx = outlier_stats(a[nz], b[nz], tol_grad)
        print(f"  d {which} planes: nnz {int(nz.sum())}  outliers(>{tol_grad:g}) {frac:.2e}  q99.9 {q999:.2e}  max {mx:.2e}")
        # derivative kinks (parity.outlier_stats): 99.9 % of the touched texels inside the bar; a texel fed by a single
        # flipped point can be off by a large share of its own value, so the rest is only capped (1e-1 of the max-norm)
        if not (q999 < tol_grad and frac < 1e-3 and mx < 1e-1):
            fails.append((tag, f"d {which} planes", frac, q999, mx))
    try:
        check_param_grads(f"B={B} {tag}", ours["param_grads"], other["param_grads"], tol_grad)
    except AssertionError as e:
        fails.append((tag, "decoder parameter gradients", str(e)[:200]))


# (B, seed, run the fp64-cast reference too).  What was measured on the B200 and the bars that follow from it:
#  * part-selection masks vs the reference's own bits on the same GPU: 0 flips of 8.2 M / 9.0 M / 37.7 M bits;
#  * forward vs the fp64-cast reference (its arithmetic without the fp32 x-remap noise of its gather, H1):
#    the adjudicator, bar 1e-4, on the rays whose samples are classified the same way in fp32 and fp64 (a sample within
#    ~1e-7 of a part boundary is classified differently by ANY fp32 implementation, the reference included): rays with a
#    detected mask flip are left out and at most 1 ray in 2000 may exceed the bar (measured: B = 2 all 32 768 rays
#    inside 2.5e-5; B = 1 seed 2201: see the printed per-ray figures);
#  * forward vs the fp32 reference on the GPU with cuDNN's TF32 convolutions off: its own fp32 noise, measured 2-4e-4
#    (B = 1) -- bar 1e-3 (3e-3 for B > 1); vs the CPU oracle: two fp32 implementations, measured 0.6-1.4e-4, bar 3e-4;
#  * gradients: 99.9 % of the touched texels / the bulk of every parameter gradient inside 1e-4 of the max-norm against
#    the fp32 runs (derivative kinks make the rest, see tests/parity.py); the fp64 run takes different sides of more
#    kinks (measured q99.9 2-3e-4): bar 1e-3;
#  * reported only: the reference exactly as shipped on a GPU, where its grouped conv1d decoder runs in TF32
#    (torch.backends.cudnn.allow_tf32 defaults to True).
@pytest.mark.parametrize("B,seed,with_fp64", [(1, 2201, True), (1, 2202, False), (2, 2203, True)])
def test_render_and_gradients_match_reference_on_gpu(B, seed, with_fp64):
    size = 128
    fails = []
    ref = R.run(B, size, NC, NF, seed, device="cuda", grad=True)
    gen, scene, ours = _ours(ref, B, size)
    flips, total, _ = _mask_flips(scene, ref)
    print(f"\n  B={B} seed={seed}: {ref['n_rays']} rays, part-mask bits vs reference-on-GPU: {flips} flips of {total}")
    # bit-exact up to boundary points: the reference's cuBLAS / fused kernels contract multiply-adds differently from
    # the separately rounded contract arithmetic; a flip needs |coordinate| within ~1e-7 of 1
    if flips > max(2, total // 10_000_000):
        fails.append(("part masks", flips, total))
    g_bar = 1e-4 if B == 1 else 1e-3           # B > 1: the reference's batched gather is only ~1e-3 accurate (H1)
    _compare("reference-on-GPU fp32 (cuDNN TF32 off)", ours, ref, B, 1e-3 if B == 1 else 3e-3, g_bar, fails)
    _compare("oracle (CPU fp32)", ours, _oracle(ref, B, size), B, 3e-4, 1e-4, fails)
    del ref, ours, gen, scene
    torch.cuda.empty_cache()
    if with_fp64:
        ref64 = R.run(B, size, NC, NF, seed, device="cuda", grad=True, dtype=torch.float64)
        _, scene64, ours64 = _ours(ref64, B, size)
        flips, total, bad = _mask_flips(scene64, ref64)
        print(f"  part-mask bits vs the fp64-cast reference: {flips} flips of {total} on {int(bad.sum())} ray(s), left out below")
        if flips > max(4, total // 2_000_000):
            fails.append(("part masks vs fp64", flips, total))
        # the reference's OWN fp32 evaluation of the same samples (fp64 run's draws pinned): how far fp32 arithmetic sits
        # from the fp64 result on this scene.  The kernels must be inside 1e-4 or no further than the reference itself.
        ref32f = R.run(B, size, NC, NF, seed, device="cuda", grad=False, force_draws=(ref64["idx"], ref64["jitter"]))
        own = {}
        for k in ("color", "mask", "disparity", "fine_weights"):
            own[k] = rel_max(ref32f[k].reshape(-1), ref64[k].reshape(-1)) if ref32f[k].shape == ref64[k].shape else float("nan")
        print("  reference fp32 vs its fp64 self, same draws: " + ", ".join(f"{k} {v:.2e}" for k, v in own.items()))
        _compare("reference-on-GPU fp64-cast", ours64, ref64, B, 1e-4, 1e-3, fails, drop_rays=bad, robust=True, own=own)
        del ref64, ours64, scene64, ref32f
        torch.cuda.empty_cache()
        shipped = R.run(B, size, NC, NF, seed, device="cuda", grad=True, allow_tf32=True)
        _, _, ours_s = _ours(shipped, B, size)
        _compare("reference-on-GPU as shipped (TF32 conv1d) -- reported only", ours_s, shipped, B, 1e-4, 1e-4, [])
    assert not fails, fails


def _scene_args(gen, inp, z):
    from enarf_gan_b200 import rendering
    pose, bl = gen.nerf.transform_pose(inp["pose_to_camera"].cuda(), inp["bone_length"].cuda())
    mi = {"z": z[:, :512].cuda(), "z_rend": z[:, 512:768].cuda(), "bone_length": bl, "truncation_psi": 1}
    return rendering._scaled(gen.nerf, pose), mi
