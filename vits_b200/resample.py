"""Host plan of the GPU resampler (`vitsb_resample_poly`, csrc/resample.cu): the `pitch` / `sampling_rate` step of
`VITSWrap.speaking` (vits_wrap.py:186-194).

The reference calls `librosa.resample(wav, orig_sr=..., target_sr=...)` with librosa's default resampler (soxr_hq today,
kaiser_best before 0.10; requirements.txt does not pin it and neither is in this image), so there is no reference output
to reproduce bit for bit.  What is built is the published polyphase method of `scipy.signal.resample_poly` - a
Kaiser-windowed (beta 5) sinc low-pass of 20 * max(up, down) + 1 taps applied as up-sample / FIR / down-sample - restated
here with numpy only; `tests/` compares plan and output with scipy itself.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np


def firwin_lowpass_kaiser(numtaps: int, cutoff: float, beta: float = 5.0) -> np.ndarray:
    """scipy.signal.firwin(numtaps, cutoff, window=('kaiser', beta)) for one low-pass band: windowed ideal response,
    scaled to unit gain at DC."""
    m = np.arange(numtaps, dtype=np.float64) - 0.5 * (numtaps - 1)
    h = cutoff * np.sinc(cutoff * m)
    h *= np.kaiser(numtaps, beta)
    return h / h.sum()


def _upfirdn_len(len_h: int, n_in: int, up: int, down: int) -> int:
    return (((n_in - 1) * up + len_h) - 1) // down + 1


@dataclass(frozen=True)
class ResamplePlan:
    up: int
    down: int
    taps: np.ndarray          # float64, zero-padded and scaled by `up`
    n_pre_remove: int

    def n_out(self, n_in: int) -> int:
        n = n_in * self.up
        return n // self.down + (1 if n % self.down else 0)


def plan(orig_sr: int, target_sr: int) -> ResamplePlan:
    """The filter and offsets scipy.signal.resample_poly(x, target_sr, orig_sr) uses (padtype 'constant', default window)."""
    orig_sr, target_sr = int(orig_sr), int(target_sr)
    if orig_sr < 1 or target_sr < 1:
        raise ValueError("sampling rates must be positive integers")
    g = math.gcd(orig_sr, target_sr)
    up, down = target_sr // g, orig_sr // g
    if up == 1 and down == 1:
        return ResamplePlan(1, 1, np.ones(1), 0)
    max_rate = max(up, down)
    half_len = 10 * max_rate
    h = firwin_lowpass_kaiser(2 * half_len + 1, 1.0 / max_rate) * up
    n_pre_pad = down - half_len % down
    n_pre_remove = (half_len + n_pre_pad) // down
    # scipy appends zeros until the filtered length covers n_out + n_pre_remove for the given n_in; trailing zeros do not
    # change any output sample, and the kernel simply stops at the last tap, so they are not materialised here
    taps = np.concatenate((np.zeros(n_pre_pad), h))
    return ResamplePlan(up, down, taps, n_pre_remove)
