"""The conditioning vector of the tri-plane / flow generators: our ``bone_length_encoding`` against the reference's
``multi_part_positional_encoding`` (libraries/NeRF/utils.py:46-71) and its use in ``compute_tri_plane_feature``."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import ref_harness as H  # noqa: E402
from enarf_gan_b200 import hostmath, synthetic as S  # noqa: E402


def test_encoding_layout_and_mask():
    bl = torch.tensor([[[0.25], [0.5], [1.5]]])                  # the third part is out of range -> zeros
    enc = hostmath.bone_length_encoding(bl, 2, 3)
    assert enc.shape == (1, 12)
    import math
    exp0 = [math.sin(0.25 * math.pi), math.sin(0.5 * math.pi), math.cos(0.25 * math.pi), math.cos(0.5 * math.pi)]
    assert torch.allclose(enc[0, :4], torch.tensor(exp0), atol=1e-6)
    assert torch.all(enc[0, 8:] == 0)


@pytest.mark.skipif(not H.available(), reason="reference not present")
def test_encoding_matches_reference():
    H.install()
    from libraries.NeRF.utils import multi_part_positional_encoding
    g = torch.Generator().manual_seed(3)
    bl = torch.rand(5, 24, 1, generator=g) * 1.3             # some parts beyond 1
    ref = multi_part_positional_encoding(bl, 4, num_bone=24)[:, :, 0]
    assert torch.equal(hostmath.bone_length_encoding(bl, 4, 24), ref)


def test_generator_is_called_with_the_encoding():
    """An injected tri-plane generator receives c = encoding of the transformed bone lengths, shape (B, 2*4*24)."""
    from enarf_gan_b200.narf import TriPlaneNARF
    seen = {}

    class Gen(torch.nn.Module):
        c_dim = 2 * 4 * 24

        def forward(self, z, c, truncation_psi=1):
            seen["c"] = c
            return torch.zeros(z.shape[0], (32 + 24) * 3, 256, 256)

    nerf = TriPlaneNARF(dict(origin_location="center+head", coordinate_scale=3, Nc=8, Nf=8, render_bs=64), z_dim=[512, 256],
                        num_bone=24, parent=S.SMPL_PARENTS, num_bone_param=24, tri_plane_gen=Gen())
    inp = S.make_inputs(2, 16, seed=1)
    _, bl = nerf.transform_pose(inp["pose_to_camera"], inp["bone_length"])
    nerf.compute_tri_plane_feature(torch.zeros(2, 512), bl)
    assert seen["c"].shape == (2, Gen.c_dim)
    assert torch.equal(seen["c"], hostmath.bone_length_encoding(bl, 4, 24))
