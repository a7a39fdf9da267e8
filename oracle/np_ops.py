"""Explicit-loop numpy conv primitives (TEST INFRASTRUCTURE).

Independent restatement of the two ATen ops whose semantics the hot path
depends on — `F.conv1d` (zero padding, dilation; models.py:288, modules.py:115,229-241)
and `F.conv_transpose1d` (weight [Cin,Cout,k], stride u, padding (k-u)//2;
models.py:289-293) — used to pin index conventions on small cases."""
import numpy as np


def conv1d(x, w, b=None, dilation=1, padding=0):
    B, Cin, T = x.shape
    Cout, _, k = w.shape
    To = T + 2 * padding - dilation * (k - 1)
    y = np.zeros((B, Cout, To), dtype=np.float64)
    for t in range(To):
        for j in range(k):
            s = t + j * dilation - padding
            if 0 <= s < T:
                y[:, :, t] += x[:, :, s].astype(np.float64) @ w[:, :, j].T.astype(np.float64)
    if b is not None:
        y += b[None, :, None]
    return y.astype(np.float32)


def conv_transpose1d(x, w, b=None, stride=1, padding=0):
    B, Cin, T = x.shape
    _, Cout, k = w.shape
    To = (T - 1) * stride - 2 * padding + k
    y = np.zeros((B, Cout, To), dtype=np.float64)
    for s in range(T):
        for j in range(k):
            t = s * stride + j - padding
            if 0 <= t < To:
                y[:, :, t] += x[:, :, s].astype(np.float64) @ w[:, :, j].astype(np.float64)
    if b is not None:
        y += b[None, :, None]
    return y.astype(np.float32)
