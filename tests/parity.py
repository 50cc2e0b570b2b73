"""Error metrics shared by the parity tests.

``rel_max``  = max|a - b| / max|b|                       (max-norm relative error: the 1e-4 / 1e-2 bars of north_star)
``rel_elem`` = max_i |a_i - b_i| / max(|b_i|, floor)     (element-wise relative error; ``floor`` = floor_frac * max|b|
               is the absolute floor below which an element is judged on absolute error -- a sum of ~50 fp32 terms of
               magnitude max|b| cannot be relatively exact where it cancels to ~0)
Both are computed in float64 on the CPU."""
import torch

FLOOR_FRAC = 1e-2


def _pair(a, b):
    return torch.as_tensor(a).detach().double().cpu(), torch.as_tensor(b).detach().double().cpu()


def rel_max(a, b):
    a, b = _pair(a, b)
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def rel_elem(a, b, floor_frac=FLOOR_FRAC):
    a, b = _pair(a, b)
    floor = floor_frac * b.abs().max().clamp_min(1e-30)
    return ((a - b).abs() / b.abs().clamp_min(floor)).max().item()


def report(name, a, b, floor_frac=FLOOR_FRAC):
    """-> (rel_max, rel_elem) and a printed line (shown with ``pytest -s`` / on failure)."""
    m, e = rel_max(a, b), rel_elem(a, b, floor_frac)
    print(f"  {name:34s} rel_max {m:.3e}   rel_elem(floor {floor_frac:g}*max) {e:.3e}")
    return m, e


def outlier_stats(a, b, tol):
    """For gradients of a piecewise-linear network: -> (fraction of elements with |a-b| > tol*max|b|, the 99.9th
    percentile of |a-b|/max|b|, rel_max).  Two fp32 evaluations that round a LeakyReLU / MyReLU pre-activation to
    different sides of 0 (|pre-activation| ~ 1e-7 of its scale; a handful per million activations) differ by the whole
    slope jump on every gradient element that point feeds, so the max-norm cannot be the only criterion."""
    a, b = _pair(a, b)
    e = (a - b).abs().reshape(-1) / b.abs().max().clamp_min(1e-30)
    k = max(1, int(round(0.999 * e.numel())))
    return (e > tol).double().mean().item(), e.kthvalue(k)[0].item(), e.max().item()


def check_param_grads(name, grads, ref_grads, tol, verbose=True):
    """Decoder parameter gradients against the reference's autograd.  A derivative-kink flip at hidden unit o of one
    point (see parity.outlier_stats) lands in ROW o of that layer's conv.weight gradient and in element o of its bias
    gradient (measured on config1_b1_s64: errors 7e-3 / 2e-3 of the max-norm in rows 31 and 22 of layer 0 with bias
    errors at the same two units, every other row <= 4e-4 -- the signature of two flipped units, not of a formula
    difference; a flipped layer-2 unit additionally spreads over every layer-1 row at the 1e-4 level).  So: the median
    element within ``tol`` of the max-norm, at least 90 % of the rows / units within 10 x ``tol``, everything within
    2e-2."""
    for key, mine in grads.items():
        ref = ref_grads[key].detach().double().cpu()
        mine = mine.detach().double().cpu().reshape(ref.shape)
        e = (mine - ref).abs() / ref.abs().max().clamp_min(1e-30)
        if key.endswith("conv.weight"):
            rows = e.reshape(ref.shape[1], -1).amax(dim=1)        # (1,out,in,1)
        elif key.endswith(".bias") and "modulation" not in key:
            rows = e.reshape(-1)
        else:
            rows = None
        if verbose:
            print(f"  {name} {key}: max {e.max():.2e} median {e.median():.2e}"
                  + ("" if rows is None else f" rows>10tol {int((rows > 10 * tol).sum())}/{rows.numel()}"))
        assert e.max() < 2e-2, (key, e.max().item())
        # (the modulation gradients sum a flipped row into every one of their few elements: 10 x tol on their median)
        assert e.median() < (tol if rows is not None else 10 * tol), (key, e.median().item())
        if rows is not None:
            assert (rows > 10 * tol).double().mean() <= 0.1, (key, rows.topk(4))
