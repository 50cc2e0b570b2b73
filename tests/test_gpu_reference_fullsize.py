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
    """The kernels through the public API (TriNARFGenerator mirror), fed the reference's parameters, inputs and draws."""
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
    return gen, dict(color=color, mask=mask, disparity=disp, fine_weights=fw[:, 0], fine_depth=fd[:, 0],
                     d_tri_plane=gen.nerf.tri_plane_gen.tp.grad,
                     param_grads={k: p.grad for k, p in gen.named_parameters() if k in ref["param_grads"]})


def _compare(tag, ours, other, B, tol):
    print(f"  -- vs {tag} (bar {tol:g})")
    for k in ("color", "mask", "disparity", "fine_weights"):
        m_, e_ = report(k, ours[k].reshape(B, -1), other[k].reshape(B, -1))
        assert m_ < tol, (tag, k, m_)
        assert e_ < 50 * tol, (tag, k, "element-wise", e_)     # every element above 1 % of the max: within 50 x the bar
    assert rel_max(ours["fine_depth"], other["fine_depth"]) < 1e-5
    for which, sl in (("feature", slice(0, 96)), ("probability", slice(96, None))):
        a, b = ours["d_tri_plane"][:, sl].cpu(), other["d_tri_plane"][:, sl].cpu()
        nz = (b != 0) | (a != 0)
        frac, q999, mx = outlier_stats(a[nz], b[nz], tol)
        print(f"  d {which} planes: nnz {int(nz.sum())}  outliers(>{tol:g}) {frac:.2e}  q99.9 {q999:.2e}  max {mx:.2e}")
        # derivative kinks (parity.outlier_stats): 99.9 % of the touched texels inside the bar, the rest capped
        assert q999 < tol and frac < 1e-3 and mx < 3e-2, (tag, which, frac, q999, mx)
    check_param_grads(f"B={B} {tag}", ours["param_grads"], other["param_grads"], tol)


# (B, seed, run the fp64-cast reference too).  Bars: 1e-4 against the fp64-cast reference (the
# reference's arithmetic without its fp32 x-remap noise, H1) and against the oracle; against the reference as shipped
# (fp32) 1e-3 for B == 1 and 3e-3 for B > 1 -- its own distance from its fp64 self, measured here in this container at
# 3.3e-4 (colour, 48 x 48 image) where the oracle sits at 7.7e-5 from the fp64 run.
@pytest.mark.parametrize("B,seed,with_fp64", [(1, 2201, True), (1, 2202, False), (2, 2203, True)])
def test_render_and_gradients_match_reference_on_gpu(B, seed, with_fp64):
    size = 128
    ref = R.run(B, size, NC, NF, seed, device="cuda", grad=True)
    gen, ours = _ours(ref, B, size)
    inp, tp, z = ref["inputs"]

    # ---- part-selection masks: the reference's own bits, computed on this GPU, at the reference's own fine points ----
    from enarf_gan_b200 import ops
    scene, _, _ = gen.nerf.build_scene(*_scene_args(gen, inp, z))
    _, _, m = ops.density_query(scene, ref["fine_points"].contiguous(), want_color=False, want_mask=True)
    bits = torch.stack([(m >> p) & 1 for p in range(24)], dim=1).bool()
    flips = int((bits != ref["valid_parts"]).sum())
    total = bits.numel()
    print(f"\n  B={B} seed={seed}: {ref['n_rays']} rays, part-mask bits vs reference-on-GPU: {flips} flips of {total}")
    # bit-exact up to boundary points: the reference's cuBLAS / fused kernels contract multiply-adds differently from
    # the separately rounded contract arithmetic; a flip needs |coordinate| within ~1e-7 of 1
    assert flips <= max(2, total // 10_000_000), f"{flips} mask flips"

    _compare("reference-on-GPU fp32 (as shipped)", ours, ref, B, 1e-3 if B == 1 else 3e-3)
    _compare("oracle (CPU fp32)", ours, _oracle(ref, B, size), B, 1e-4)
    del ref, ours, gen
    torch.cuda.empty_cache()
    if with_fp64:
        ref64 = R.run(B, size, NC, NF, seed, device="cuda", grad=True, dtype=torch.float64)
        _, ours64 = _ours(ref64, B, size)
        _compare("reference-on-GPU fp64-cast", ours64, ref64, B, 1e-4)


def _scene_args(gen, inp, z):
    from enarf_gan_b200 import rendering
    pose, bl = gen.nerf.transform_pose(inp["pose_to_camera"].cuda(), inp["bone_length"].cuda())
    mi = {"z": z[:, :512].cuda(), "z_rend": z[:, 512:768].cuda(), "bone_length": bl, "truncation_psi": 1}
    return rendering._scaled(gen.nerf, pose), mi
