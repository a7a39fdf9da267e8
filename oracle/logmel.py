"""Log-mel spectrogram for the parity metric (log-mel L1).

TEST INFRASTRUCTURE.  Follows `mel_spectrogram_torch`
(mel_processing.py:92-119): reflect pad (n_fft-hop)/2, `torch.stft`
center=False with a hann window of win_length, magnitude
sqrt(re^2+im^2+1e-6), mel projection, log(clamp(.,1e-5)) (mel_processing.py:26-32).
The mel basis is `librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax)`
(mel_processing.py:107; librosa is third party and absent here): restated
below from librosa's published algorithm — Slaney mel scale (linear below
1 kHz, log above), triangular filters, `norm='slaney'` area normalisation.
"""
from __future__ import annotations

import numpy as np
import torch


def _hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-10) / min_log_hz) / logstep, mels)


def _mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * m
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), freqs)


def mel_basis(sr: int, n_fft: int, n_mels: int, fmin: float = 0.0, fmax=None) -> np.ndarray:
    if fmax is None:
        fmax = sr / 2.0
    fftfreqs = np.linspace(0, sr / 2.0, 1 + n_fft // 2)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    weights = np.zeros((n_mels, 1 + n_fft // 2))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, None]
    return weights.astype(np.float32)


def log_mel(y: torch.Tensor, n_fft=1024, num_mels=80, sampling_rate=16000, hop_size=192,
            win_size=768, fmin=0.0, fmax=None) -> torch.Tensor:
    """y: [B, t] float32 in [-1,1] -> [B, num_mels, frames]."""
    y = y.float()
    pad = int((n_fft - hop_size) / 2)
    y = torch.nn.functional.pad(y.unsqueeze(1), (pad, pad), mode="reflect").squeeze(1)
    spec = torch.stft(y, n_fft, hop_length=hop_size, win_length=win_size,
                      window=torch.hann_window(win_size), center=False, normalized=False,
                      onesided=True, return_complex=True)
    spec = torch.sqrt(spec.real.pow(2) + spec.imag.pow(2) + 1e-6)
    mel = torch.from_numpy(mel_basis(sampling_rate, n_fft, num_mels, fmin, fmax))
    spec = torch.matmul(mel, spec)
    return torch.log(torch.clamp(spec, min=1e-5))


def log_mel_l1(a: torch.Tensor, b: torch.Tensor, **kw) -> float:
    """Mean absolute difference of log-mels of two waveforms [B,t]."""
    return (log_mel(a, **kw) - log_mel(b, **kw)).abs().mean().item()
