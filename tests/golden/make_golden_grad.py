"""Gradient fixtures: autograd through the UNMODIFIED reference on CPU (authoring container only).

    python tests/golden/make_golden_grad.py

For every case of ``make_golden.py`` the reference generator is run again with the same seeds (its outputs and RNG
draws are checked to be bit-identical to the forward fixture) with autograd on, and the gradients of the fixed linear
loss of ``tests/ref_runner.py`` are stored in ``tests/golden/case_<name>_grad.npz``:
  * every decoder parameter's gradient (conv.weight, conv.modulation.weight / bias, bias of the three layers) in full;
  * the tri-plane gradient as per-channel sums and |.|-sums plus a fixed subsample of its non-zero entries (the largest
    4096 in magnitude and 60 000 pseudo-randomly chosen ones; the full tensor is 44 MB per sample).
These pin the hand-written pieces a restatement can get wrong silently: MyReLU's backward (NeRF/activation.py:12-16),
the modulation / normalisation chain (custom_stylegan2/net.py:229-250) and the "coarse pass carries no gradient" rule.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import ref_runner as R  # noqa: E402
from make_golden import CASES  # noqa: E402

N_TOP, N_RAND = 4096, 60000


def subsample(grad, seed):
    flat = grad.reshape(-1)
    nz = flat.nonzero()[:, 0]
    top = flat.abs().topk(min(N_TOP, nz.numel()))[1]
    g = torch.Generator().manual_seed(seed)
    pick = nz[torch.randperm(nz.numel(), generator=g)[:N_RAND]]
    idx = torch.unique(torch.cat([top, pick]))
    return idx.numpy().astype(np.int64), flat[idx].numpy(), int(nz.numel())


def run_case(name, B, size, Nc, Nf, seed, origin, keep):
    fwd = np.load(os.path.join(HERE, f"case_{name}.npz"))
    o = R.run(B, size, Nc, Nf, seed, origin, device="cpu", grad=True)
    # same run as the forward fixture: identical draws and image
    assert np.array_equal(o["idx"].numpy().astype(np.int16), fwd["idx"]), "fine-pass draws differ from the forward fixture"
    assert np.array_equal(o["jitter"].numpy(), fwd["jitter"])
    assert np.array_equal(o["color"].numpy(), fwd["color"]), "image differs from the forward fixture"
    d_tp = o["d_tri_plane"]
    idx, val, nnz = subsample(d_tp, seed)
    g = {"meta": np.array([B, size, Nc, Nf, seed, o["P"]]), "loss": o["loss"].numpy(), "loss_seed": np.array(5),
         "n_rays": np.array(o["n_rays"]),
         "d_tp_channel_sum": d_tp.double().sum(dim=(2, 3)).numpy(), "d_tp_channel_abs": d_tp.double().abs().sum(dim=(2, 3)).numpy(),
         "d_tp_max": d_tp.abs().amax(dim=(2, 3)).numpy(), "d_tp_idx": idx, "d_tp_val": val, "d_tp_nnz": np.array(nnz)}
    for k, v in o["param_grads"].items():
        g["grad:" + k] = v.numpy()
    path = os.path.join(HERE, f"case_{name}_grad.npz")
    np.savez_compressed(path, **g)
    print(name, "loss", float(o["loss"]), "nnz", nnz, "stored", idx.size, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    only = sys.argv[1:]
    for name, cfg in CASES.items():
        if not only or name in only:
            run_case(name, *cfg)
