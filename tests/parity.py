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
    """Decoder parameter gradients against another implementation's autograd.  A derivative-kink flip at hidden unit o
    of one point (see outlier_stats) lands in ROW o of that layer's conv.weight gradient and in element o of its bias
    gradient (measured on config1_b1_s64 against the reference: errors 7e-3 / 2e-3 of the max-norm in rows 31 and 22 of
    layer 0 with bias errors at the same two units, every other row <= 4e-4 -- the signature of two flipped units, not
    of a formula difference; a flipped layer-2 unit additionally spreads over every layer-1 row at the 1e-4 level, and
    the number of flips grows with the number of activations: ~45 M at 128 x 128).  So the bar is statistical: the
    median element within ``tol`` of the max-norm, 90 % of the elements within 10 x ``tol``, everything within 3e-2.
    The modulation gradients (32 / 64 elements) sum every row, flipped ones included, into each element: 10 x looser."""
    for key, mine in grads.items():
        ref = ref_grads[key].detach().double().cpu()
        mine = mine.detach().double().cpu().reshape(ref.shape)
        e = ((mine - ref).abs() / ref.abs().max().clamp_min(1e-30)).reshape(-1)
        q90 = e.kthvalue(max(1, int(round(0.9 * e.numel()))))[0].item()
        if verbose:
            print(f"  {name} {key}: median {e.median():.2e}  q90 {q90:.2e}  max {e.max():.2e}")
        t = 10 * tol if "modulation" in key else tol
        if e.numel() < 32:          # db3 has 4 elements: its "90th percentile" is its maximum
            q90 = 0.0
        assert e.median() < t and q90 < 10 * t and e.max() < 3e-2, (key, e.median().item(), q90, e.max().item())
