"""Early ray termination of the fine pass (north_star step (e)).  The reference composites every sample
(libraries/NeRF/rendering.py:307-335); with ``ert_eps`` > 0 the kernels stop evaluating a ray once its exact prefix
transmittance is below ert_eps.  Dropped-tail bound: the weights of the dropped samples sum to less than ert_eps, so
colour (|c| <= 1), mask and every fine_weight move by < ert_eps and disparity by < ert_eps / near.

Checked here on a synthetic OPAQUE scene (probability planes >> 0, density bias > 0: most fine samples sit behind
the first surface): outputs and gradients agree with the exhaustive render within 1e-4 while a large share of the
decoder evaluations is skipped; on the ordinary (translucent) fixture the result is bit-identical and nothing is
skipped; forward and backward agree on the dropped set with and without the pair checkpoint."""
import numpy as np
import pytest
import torch

import enarf_oracle as O
import golden_util as G
from parity import rel_max

pytestmark = pytest.mark.gpu
EPS = 1e-6


def _scene(opaque, grad=False):
    from enarf_gan_b200 import ops
    g, cfg, inp, tp, layers, z_rend = G.load("b2_s24")
    sc = G.scene(cfg, inp, tp, layers, z_rend)
    tp = sc.tri_plane.clone()
    dec = [(w.detach().clone(), b.detach().clone()) for w, b in sc.eff]
    if opaque:
        tp[:, 96:] += 6.0            # every part claims its points: part weight ~ 1
        dec[2][1][3] += 30.0         # density bias: sigma ~ 10 * sqrt2 * (o + 30) > 400 inside the body (a surface)
    d = lambda t: t.detach().float().cuda()
    tp_g = d(tp).requires_grad_(grad)
    dec_g = [(d(w).requires_grad_(grad), d(b).requires_grad_(grad)) for w, b in dec]
    gs = ops.SceneBuffers(d(sc.pose), d(sc.cscale), d(sc.canonical_pose), tp_g, dec_g)
    img = O.whole_image_grid(cfg["size"], cfg["size"], cfg["B"]).cuda()
    return cfg, sc, gs, img, tp_g, dec_g


def _render(opaque, ert_eps, grad=False, seed=3):
    from enarf_gan_b200 import rendering
    cfg, sc, gs, img, tp_g, dec_g = _scene(opaque, grad)
    out = rendering.render_rays(gs, img, sc.inv_intrinsics.cuda(), cfg["Nc"], cfg["Nf"], rng="philox", seed=seed,
                                tri_plane=tp_g if grad else None, decoder=dec_g if grad else None, ert_eps=ert_eps,
                                want_points=True)
    return out, tp_g, dec_g


def test_opaque_scene_matches_the_exhaustive_render_and_skips_work():
    full, _, _ = _render(True, 0.0)
    fast, _, _ = _render(True, EPS)
    ev, sk = [int(v) for v in fast["ert_stats"].tolist()[:2]]
    frac = sk / max(1, ev + sk)
    print(f"\n  opaque scene: {ev} fine samples evaluated, {sk} skipped ({100 * frac:.1f} % of the decoder evaluations)")
    # hierarchical sampling already concentrates the fine samples at the first surface: what is left to skip are the
    # samples the 0.01 floor of the blurred weights spreads behind it
    assert frac > 0.2, "an opaque scene should retire rays early"
    for k in ("color", "mask", "disparity", "fine_weights"):
        err = (fast[k] - full[k]).abs().max().item()
        print(f"  {k}: max |diff| {err:.2e}  (rel_max {rel_max(fast[k], full[k]):.2e})")
        assert rel_max(fast[k], full[k]) < 1e-4, k
    assert (fast["color"] - full["color"]).abs().max().item() < 10 * EPS
    assert (fast["mask"] - full["mask"]).abs().max().item() < 10 * EPS
    assert torch.equal(fast["fine_depth"], full["fine_depth"])
    # dropped samples are exposed with density 0 and weight 0, never negative
    assert fast["points"][..., 3].min().item() >= 0
    assert torch.equal(fast["validity"], full["validity"])


@pytest.mark.parametrize("pair_ckpt", [True, False])
def test_opaque_scene_gradients(pair_ckpt, monkeypatch):
    from enarf_gan_b200 import ops
    monkeypatch.setattr(ops, "PAIR_CKPT", pair_ckpt)
    grads = []
    for eps in (0.0, EPS):
        out, tp_g, dec_g = _render(True, eps, grad=True)
        g = torch.Generator().manual_seed(9)
        wc = torch.randn(out["color"].shape, generator=g).cuda()
        wm = torch.randn(out["mask"].shape, generator=g).cuda()
        wf = torch.randn(out["fine_weights"].shape, generator=g).cuda()
        ((out["color"] * wc).sum() + (out["mask"] * wm).sum() + (out["fine_weights"] * wf).sum()).backward()
        grads.append((tp_g.grad.clone(), [w.grad.clone() for w, _ in dec_g], [b.grad.clone() for _, b in dec_g]))
    (tp0, w0, b0), (tp1, w1, b1) = grads
    print(f"\n  d feature planes {rel_max(tp1[:, :96], tp0[:, :96]):.2e}  d probability planes {rel_max(tp1[:, 96:], tp0[:, 96:]):.2e}")
    assert rel_max(tp1[:, :96], tp0[:, :96]) < 1e-4
    assert rel_max(tp1[:, 96:], tp0[:, 96:]) < 1e-4
    for i in range(3):
        assert rel_max(w1[i], w0[i]) < 1e-4, f"dW{i + 1}"
        assert rel_max(b1[i], b0[i]) < 1e-3, f"db{i + 1}"


def test_translucent_scene_is_bit_identical_and_skips_nothing():
    """The ordinary fixture: the coarse pass reports (almost) no opaque rays, the fine pass runs in one round and the
    group-major pair order changes no per-sample result."""
    full, _, _ = _render(False, 0.0)
    fast, _, _ = _render(False, EPS)
    for k in ("color", "mask", "disparity", "fine_weights", "fine_depth"):
        assert torch.equal(fast[k], full[k]), k
    assert int(fast["ert_stats"][1]) == 0


def test_rounds_forced_on_a_translucent_scene_change_nothing():
    """ert_hint = NULL runs every tile in rounds; no ray of this scene reaches T < 1e-6 ... unless one does: compare."""
    from enarf_gan_b200 import ops
    cfg, sc, gs, img, _, _ = _scene(False)
    dmin, dmax, direction, _ = ops.depth_range(img, sc.inv_intrinsics.cuda(), gs.pose)
    wb = ops.render_coarse(gs, dmin, dmax, direction, cfg["Nc"])
    bins = ops.fine_sample(wb, cfg["Nf"], seed=5)
    a = ops.render_fine(gs, dmin, dmax, direction, bins, want_points=True)
    stats = torch.zeros(2, device="cuda", dtype=torch.int32)
    b = ops.render_fine(gs, dmin, dmax, direction, bins, want_points=True, ert_eps=EPS, ert_hint=None, ert_stats=stats)
    for x, y in zip(a[:4], b[:4]):
        assert rel_max(y, x) < 1e-5
    print(f"\n  forced rounds: evaluated {int(stats[0])}, skipped {int(stats[1])}")


def test_coarse_pass_terminates_too():
    """The stratified coarse samples are where an opaque scene wastes most: about half of them lie behind the surface."""
    from enarf_gan_b200 import ops
    cfg, sc, gs, img, _, _ = _scene(True)
    dmin, dmax, direction, _ = ops.depth_range(img, sc.inv_intrinsics.cuda(), gs.pose)
    full, den_full = ops.render_coarse(gs, dmin, dmax, direction, cfg["Nc"], want_density=True)
    stats = torch.zeros(2, device="cuda", dtype=torch.int32)
    opaque = torch.zeros(1, device="cuda", dtype=torch.int32)
    fast, den = ops.render_coarse(gs, dmin, dmax, direction, cfg["Nc"], want_density=True, opaque_rays=opaque, ert_eps=EPS,
                                  ert_hint=None, ert_stats=stats)
    ev, sk = [int(v) for v in stats.tolist()]
    print(f"\n  coarse pass: {ev} samples evaluated, {sk} skipped ({100 * sk / max(1, ev + sk):.1f} %), "
          f"{int(opaque)} of {dmin.numel()} rays opaque")
    assert sk / max(1, ev + sk) > 0.25
    assert (fast - full).abs().max().item() < 10 * EPS          # blurred weights: a dropped sample keeps the 0.01 floor
    assert den.min().item() >= 0
    # a hint below the threshold: one round, nothing skipped, identical weights
    stats.zero_()
    none = ops.render_coarse(gs, dmin, dmax, direction, cfg["Nc"], ert_eps=EPS, ert_hint=torch.zeros(1, device="cuda", dtype=torch.int32),
                             ert_stats=stats)
    assert torch.equal(none, full) and int(stats[1]) == 0


def test_model_level_switch_feeds_the_hint_forward():
    """nerf.ert_eps on the public API: the second call's coarse pass runs in rounds on the first call's opaque count."""
    from test_gpu_api import make_generator
    from enarf_gan_b200 import synthetic as S
    g, cfg, inp, tp, layers, z_rend = G.load("b2_s24")
    tp = tp.clone()
    tp[:, 96:] += 6.0
    gen = make_generator(cfg, tp, g).eval()
    with torch.no_grad():
        gen.nerf.mlp.layers[2].bias[0, 3, 0] += 30.0
    gen.nerf.rng_mode = "philox"
    gen.nerf.ert_eps = EPS
    z = S.make_latent(cfg["B"], seed=cfg["seed"]).cuda()
    c = lambda k: inp[k].cuda()
    with torch.no_grad():
        for _ in range(2):
            gen(c("pose_to_camera"), c("pose_to_world"), c("bone_length"), z, c("inv_intrinsics"))
            st = gen.nerf.buffers_tensors["ert_stats"].tolist()
    print(f"\n  second call: fine {st[0]} / skipped {st[1]}, coarse {st[2]} / skipped {st[3]}")
    assert st[3] > 0 and st[1] > 0
    assert gen.nerf.buffers_tensors["fine_density"].min().item() >= 0
