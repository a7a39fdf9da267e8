"""Host plan of the GPU resampler (vits_b200/resample.py) against scipy.signal: the filter, the offsets and the kernel's
index arithmetic (restated in numpy) reproduce scipy.signal.resample_poly.  The reference's own resampler (librosa's
default, vits_wrap.py:186-194) is not in this image and not pinned by the reference: see the module docstring."""
import numpy as np
import pytest
import scipy.signal as ss

from vits_b200 import resample as R

RATES = [(16000, 8000), (16000, 24000), (int(16000 / 1.1), 16000), (int(16000 / 0.9), 16000), (16000, 22050), (8000, 16000), (16000, 44100)]


def kernel_restatement(p, x):
    """y[n] = sum_k taps[(n + n_pre_remove) * down - k * up] * x[k]  (csrc/resample.cu), float64."""
    n_out = p.n_out(len(x))
    y = np.zeros(n_out)
    for n in range(n_out):
        t = (n + p.n_pre_remove) * p.down
        k_hi = min(t // p.up, len(x) - 1)
        first = t - (len(p.taps) - 1)
        k_lo = 0 if first <= 0 else (first + p.up - 1) // p.up
        ks = np.arange(k_lo, k_hi + 1)
        y[n] = np.sum(p.taps[t - ks * p.up] * x[ks])
    return y


@pytest.mark.parametrize("orig,target", RATES)
def test_plan_matches_scipy(orig, target):
    p = R.plan(orig, target)
    mr = max(p.up, p.down)
    h = ss.firwin(20 * mr + 1, 1.0 / mr, window=("kaiser", 5.0)) * p.up
    pad = len(p.taps) - len(h)
    assert pad == p.down - (10 * mr) % p.down and not p.taps[:pad].any()
    assert np.abs(p.taps[pad:] - h).max() <= 1e-14
    for n_in in (1, 2, 57, 403):
        x = np.random.default_rng(n_in).standard_normal(n_in)
        y = ss.resample_poly(x, p.up, p.down)
        assert p.n_out(n_in) == len(y)
        assert np.abs(kernel_restatement(p, x) - y).max() <= 1e-12


def test_identity_and_errors():
    p = R.plan(16000, 16000)
    assert (p.up, p.down) == (1, 1) and p.n_out(123) == 123
    with pytest.raises(ValueError):
        R.plan(0, 16000)
