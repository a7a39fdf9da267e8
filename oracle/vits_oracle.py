"""Functional CPU restatement of emotional-vits' flow-reverse + Generator.

TEST INFRASTRUCTURE: the checker for the CUDA path and the timed CPU
baseline.  Never imported by ``vits_b200``.

The reference is pure PyTorch; its arithmetic lives in ATen (third party,
``requirements.txt:5`` pins torch==1.13, this image has 2.11).  This file
restates the *composition* of those ops on a plain ``state_dict`` with
``torch.nn.functional`` calls, no ``nn.Module`` and no import of the
reference.  Parity pin: ``oracle/gen_golden.py`` runs the reference's own
modules (imported read-only from /root/reference in the build container) on
the inputs of ``oracle/fixtures.py`` and commits the outputs under
``tests/golden/``; ``tests/test_oracle.py`` checks this file against them.
The reference has no tests or golden vectors of its own.

All paths are relative to /root/reference/emotional-vits/.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1  # modules.py:11


from vits_b200.config import HotPathConfig  # noqa: E402,F401  (the product owns the schema; the oracle follows it)


# ---------------------------------------------------------------------------
# weight-norm folding (models.py:467-474; torch.nn.utils.weight_norm, dim=0)
# ---------------------------------------------------------------------------
def fold_weight_norm(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Return a state_dict where every `X.weight_g`/`X.weight_v` pair is
    replaced by `X.weight = g * v / ||v||`, the norm taken over every dim but 0
    (dim 0 is Cout for Conv1d/Linear and Cin for ConvTranspose1d)."""
    if not any(k.endswith(".weight_v") for k in sd):
        return sd
    out: Dict[str, torch.Tensor] = {}
    for k, v in sd.items():
        if k.endswith(".weight_g"):
            continue
        if k.endswith(".weight_v"):
            base = k[: -len("_v")]
            g = sd[base + "_g"]
            dims = tuple(range(1, v.dim()))
            norm = v.float().pow(2).sum(dim=dims, keepdim=True).sqrt()
            out[base] = (g.float() * v.float() / norm).to(v.dtype)
        else:
            out[k] = v
    return out


def _get_padding(kernel_size: int, dilation: int = 1) -> int:
    return int((kernel_size * dilation - dilation) / 2)  # commons.py:13-14


# ---------------------------------------------------------------------------
# WN (modules.py:130-182)
# ---------------------------------------------------------------------------
def wn(sd, prefix: str, x: torch.Tensor, g: torch.Tensor, cfg: HotPathConfig,
       dilation_rate: int = 1, x_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`WN.infer` (modules.py:157-182) when x_mask is None, else the masked
    `WN.forward` (modules.py:130-155)."""
    H, k, nl = cfg.hidden_channels, cfg.flow_kernel_size, cfg.flow_n_layers
    G = F.linear(g, sd[prefix + "cond_layer.weight"], sd[prefix + "cond_layer.bias"])
    output = torch.zeros_like(x)
    for i in range(nl):
        dil = dilation_rate ** i
        pad = int((k * dil - dil) / 2)
        x_in = F.conv1d(x, sd[f"{prefix}in_layers.{i}.weight"], sd[f"{prefix}in_layers.{i}.bias"],
                        dilation=dil, padding=pad)
        acts = x_in + G[:, i * 2 * H:(i + 1) * 2 * H].unsqueeze(-1)
        acts = torch.tanh(acts[:, :H]) * torch.sigmoid(acts[:, H:])
        rs = F.conv1d(acts, sd[f"{prefix}res_skip_layers.{i}.weight"], sd[f"{prefix}res_skip_layers.{i}.bias"])
        if i < nl - 1:
            x = x + rs[:, :H]
            if x_mask is not None:
                x = x * x_mask
            output = output + rs[:, H:]
        else:
            output = output + rs
    if x_mask is not None:
        output = output * x_mask
    return output


# ---------------------------------------------------------------------------
# ResidualCouplingLayer reverse, mean_only (modules.py:341-375)
# ---------------------------------------------------------------------------
def coupling_reverse(sd, prefix: str, x: torch.Tensor, g: torch.Tensor, cfg: HotPathConfig,
                     dilation_rate: int = 1, x_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    half = cfg.inter_channels // 2
    x0, x1 = x[:, :half], x[:, half:]
    h = F.conv1d(x0, sd[prefix + "pre.weight"], sd[prefix + "pre.bias"])
    if x_mask is not None:
        h = h * x_mask
    h = wn(sd, prefix + "enc.", h, g, cfg, dilation_rate, x_mask)
    m = F.conv1d(h, sd[prefix + "post.weight"], sd[prefix + "post.bias"])
    if x_mask is not None:
        m = m * x_mask
    x1 = x1 - m            # mean_only: logs == 0 (modules.py:349-350,370-372)
    if x_mask is not None:
        x1 = x1 * x_mask
    return torch.cat([x0, x1], 1)


def flow_reverse(sd, z_p: torch.Tensor, g: torch.Tensor, cfg: HotPathConfig,
                 x_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`ResidualCouplingBlock.infer(reverse=True)` (models.py:228-235) or, with
    x_mask [B,1,T], `.forward(x, x_mask, g, reverse=True)` (models.py:219-226).
    Module list is [RCL0, Flip, RCL2, Flip, ...]; reversed: Flip first."""
    sd = fold_weight_norm(sd)
    x = z_p
    for i in reversed(range(cfg.n_flows)):
        x = torch.flip(x, [1])                                   # modules.py:278-289
        x = coupling_reverse(sd, f"flow.flows.{2 * i}.", x, g, cfg, cfg.dilation_rate[i], x_mask)
    return x


# ---------------------------------------------------------------------------
# ResBlock2 (modules.py:250-260) and Generator (models.py:306-318)
# ---------------------------------------------------------------------------
def resblock2(sd, prefix: str, x: torch.Tensor, g: torch.Tensor, k: int, dils: Sequence[int]) -> torch.Tensor:
    for d_idx, d in enumerate(dils):
        xt = F.leaky_relu(x, LRELU_SLOPE)
        xt = F.conv1d(xt, sd[f"{prefix}convs1.{d_idx}.weight"], sd[f"{prefix}convs1.{d_idx}.bias"],
                      dilation=d, padding=_get_padding(k, d))
        gs = F.linear(g, sd[f"{prefix}conds.{d_idx}.weight"], sd[f"{prefix}conds.{d_idx}.bias"])
        xa, xb = torch.chunk(xt, 2, dim=1)
        sa, sb = torch.chunk(gs, 2, dim=1)
        xt = torch.tanh(xa + sa.unsqueeze(-1)) * torch.sigmoid(xb + sb.unsqueeze(-1))
        xt = F.conv1d(xt, sd[f"{prefix}convs2.{d_idx}.weight"], sd[f"{prefix}convs2.{d_idx}.bias"],
                      dilation=1, padding=_get_padding(k, 1))
        x = xt + x
    return x


def generator(sd, z: torch.Tensor, g: torch.Tensor, cfg: HotPathConfig,
              taps: Optional[Dict[str, torch.Tensor]] = None) -> torch.Tensor:
    """`Generator.forward` (models.py:306-318).  `taps`, if given, receives
    per-stage activations for layer-level parity tests."""
    sd = fold_weight_norm(sd)
    nk = len(cfg.resblock_kernel_sizes)
    x = F.conv1d(z, sd["dec.conv_pre.weight"], sd["dec.conv_pre.bias"], padding=3)
    if taps is not None:
        taps["conv_pre"] = x
    for i, (u, k) in enumerate(zip(cfg.upsample_rates, cfg.upsample_kernel_sizes)):
        x = F.leaky_relu(x, LRELU_SLOPE)
        x = F.conv_transpose1d(x, sd[f"dec.ups.{i}.weight"], sd[f"dec.ups.{i}.bias"],
                               stride=u, padding=(k - u) // 2)
        if taps is not None:
            taps[f"ups.{i}"] = x
        xs = None
        for j in range(nk):
            r = resblock2(sd, f"dec.resblocks.{i * nk + j}.", x, g,
                          cfg.resblock_kernel_sizes[j], cfg.resblock_dilation_sizes[j])
            xs = r if xs is None else xs + r
        x = xs / nk
        if taps is not None:
            taps[f"mrf.{i}"] = x
    x = F.leaky_relu(x)                       # default slope 0.01 (models.py:315)
    x = F.conv1d(x, sd["dec.conv_post.weight"], None, padding=3)   # bias=False (models.py:301)
    return torch.tanh(x)


# ---------------------------------------------------------------------------
# seam: infer_p2 (models.py:568-575) and the tail of inference() (models.py:529-535)
# ---------------------------------------------------------------------------
def length_regulate(attn, m_p, s_p, noise):
    m = torch.matmul(attn, m_p.transpose(1, 2)).transpose(1, 2)   # models.py:570
    s = torch.matmul(attn, s_p.transpose(1, 2)).transpose(1, 2)   # models.py:571
    return m + noise * s                                          # models.py:572


def infer_p2(sd, attn, m_p, s_p, g, noise, cfg: HotPathConfig) -> torch.Tensor:
    z_p = length_regulate(attn, m_p, s_p, noise)
    z = flow_reverse(sd, z_p, g, cfg)
    return generator(sd, z, g, cfg)


def sequence_mask(lengths: torch.Tensor, max_length: int) -> torch.Tensor:
    ar = torch.arange(max_length, dtype=lengths.dtype)            # commons.py:120-124
    return ar.unsqueeze(0) < lengths.unsqueeze(1)


def synthesize_padded(sd, z_p, g, lengths, cfg: HotPathConfig):
    """Batched reference semantics of `inference()` (models.py:533-534):
    masked reverse flow, then UNMASKED decoder on z*mask."""
    mask = sequence_mask(lengths, z_p.size(2)).unsqueeze(1).to(z_p.dtype)
    z = flow_reverse(sd, z_p, g, cfg, x_mask=mask)
    return z, generator(sd, z * mask, g, cfg)


def synthesize_ragged(sd, z_p, g, lengths, cfg: HotPathConfig):
    """Per-utterance semantics: row b equals B=1 `flow.infer` + `dec`
    (models.py:573-574) on its own length; output zero beyond it."""
    B, _, T = z_p.shape
    up = cfg.total_upsample
    z = torch.zeros_like(z_p)
    o = torch.zeros(B, 1, T * up, dtype=z_p.dtype)
    for b in range(B):
        L = int(lengths[b])
        zb = flow_reverse(sd, z_p[b:b + 1, :, :L], g[b:b + 1], cfg)
        z[b:b + 1, :, :L] = zb
        o[b:b + 1, :, :L * up] = generator(sd, zb, g[b:b + 1], cfg)
    return z, o
