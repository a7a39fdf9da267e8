"""Functional CPU restatement of `SynthesizerTrn.infer_p1` (text encoder + duration predictor).

TEST INFRASTRUCTURE: the checker for `vitsb_infer_p1`; never imported by ``vits_b200``.
Composition of ATen ops on a plain state_dict, each function citing the reference lines it follows (paths relative
to /root/reference/emotional-vits/).  Pinned by `oracle/gen_golden.py` (reference modules run in the build container,
outputs committed as tests/golden/base_p1_*.npz) and `tests/test_oracle.py`.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from vits_b200.config import HotPathConfig


def gen_sin_table(max_len: int, d_model: int) -> torch.Tensor:
    """commons.py:176-190."""
    pe = torch.zeros(max_len, d_model)
    position = torch.arange(0, max_len, dtype=torch.float32).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32) * -(math.log(10000.0) / d_model))
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    return pe.unsqueeze(0)


def _ln(x, gamma, beta, eps=1e-5):
    """modules.LayerNorm (modules.py:33-44): layer norm over the channel dim of [B, C, T]."""
    return F.layer_norm(x.transpose(1, -1), (x.size(1),), gamma, beta, eps).transpose(1, -1)


def mha(sd, p: str, x, n_heads: int, attn_mask=None):
    """MultiHeadAttention.forward(x, x, attn_mask) (attentions.py:78-102)."""
    q = F.conv1d(x, sd[p + "conv_q.weight"], sd[p + "conv_q.bias"])
    k = F.conv1d(x, sd[p + "conv_k.weight"], sd[p + "conv_k.bias"])
    v = F.conv1d(x, sd[p + "conv_v.weight"], sd[p + "conv_v.bias"])
    b, d, t = k.shape
    kc = d // n_heads
    q = q.view(b, n_heads, kc, t).transpose(2, 3)
    k = k.view(b, n_heads, kc, t).transpose(2, 3)
    v = v.view(b, n_heads, kc, t).transpose(2, 3)
    scores = torch.matmul(q / math.sqrt(kc), k.transpose(-2, -1))
    if attn_mask is not None:
        scores = scores.masked_fill(attn_mask == 0, -1e4)
    o = torch.matmul(F.softmax(scores, dim=-1), v)
    o = o.transpose(2, 3).contiguous().view(b, d, t)
    return F.conv1d(o, sd[p + "conv_o.weight"], sd[p + "conv_o.bias"])


def ffn2(sd, p: str, x, g, k: int, x_mask=None):
    """FFN2.infer / .forward (attentions.py:148-166)."""
    x = F.conv1d(x, sd[p + "conv_1.weight"], sd[p + "conv_1.bias"], padding=k // 2)
    c = F.linear(g, sd[p + "cond.weight"], sd[p + "cond.bias"])
    xa, xb = torch.chunk(x, 2, dim=1)
    sa, sb = torch.chunk(c, 2, dim=1)
    x = torch.tanh(xa + sa.unsqueeze(-1)) * torch.sigmoid(xb + sb.unsqueeze(-1))
    if x_mask is not None:
        x = x * x_mask
    x = F.conv1d(x, sd[p + "conv_2.weight"], sd[p + "conv_2.bias"], padding=k // 2)
    return x if x_mask is None else x * x_mask


def encoder(sd, x, g, cfg: HotPathConfig, x_mask=None):
    """attentions.Encoder.infer (attentions.py:48-54) / .forward (:34-46) with FFN2 layers."""
    attn_mask = None
    if x_mask is not None:
        attn_mask = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
        x = x * x_mask
    for i in range(cfg.n_layers):
        e = "enc_p.encoder."
        y = mha(sd, f"{e}attn_layers.{i}.", x, cfg.n_heads, attn_mask)
        x = _ln(x + y, sd[f"{e}norm_layers_1.{i}.gamma"], sd[f"{e}norm_layers_1.{i}.beta"])
        y = ffn2(sd, f"{e}ffn_layers.{i}.", x, g, cfg.kernel_size, x_mask)
        x = _ln(x + y, sd[f"{e}norm_layers_2.{i}.gamma"], sd[f"{e}norm_layers_2.{i}.beta"])
    return x if x_mask is None else x * x_mask


def text_encoder(sd, text, emo, g, cfg: HotPathConfig, x_lengths: Optional[torch.Tensor] = None):
    """TextEncoder.infer (models.py:180-189); with x_lengths the masked .forward (:166-178).
    text [B, Tx, text_channels], emo [B, 1024], g [B, gin] -> x [B, H, Tx], m, logs [B, inter, Tx]."""
    H = cfg.hidden_channels
    x = F.linear(text, sd["enc_p.emb.0.weight"], sd["enc_p.emb.0.bias"])
    x = F.layer_norm(x, (H,), sd["enc_p.emb.1.weight"], sd["enc_p.emb.1.bias"], 1e-5)
    x = x + F.linear(emo, sd["enc_p.emo_proj.weight"], sd["enc_p.emo_proj.bias"]).unsqueeze(1)
    T = x.size(1)
    pe = gen_sin_table(max(T, 256 + 128), H)[:, :T]            # models.py:150-154,160-164 (same table values either way)
    x = x * math.sqrt(H) + pe * sd["enc_p.alpha"]
    x = x.transpose(1, -1)
    x_mask = None
    if x_lengths is not None:
        x_mask = (torch.arange(T).unsqueeze(0) < x_lengths.unsqueeze(1)).unsqueeze(1).to(x.dtype)
    x = encoder(sd, x, g, cfg, x_mask)
    stats = F.conv1d(x, sd["enc_p.proj.weight"], sd["enc_p.proj.bias"])
    if x_mask is not None:
        stats = stats * x_mask
    m, logs = torch.split(stats, cfg.inter_channels, dim=1)
    return x, m, logs, x_mask


def duration_predictor(sd, x, g, cfg: HotPathConfig, x_mask=None):
    """DurationPredictor.infer (models.py:59-67) / .forward (:48-57); act_func ReLU, dropout off."""
    k = cfg.kernel_size_d
    x = F.conv1d(x, sd["dp.pre.weight"], sd["dp.pre.bias"]) + F.linear(g, sd["dp.cond1.weight"], sd["dp.cond1.bias"]).unsqueeze(-1)
    x = F.conv1d(x if x_mask is None else x * x_mask, sd["dp.conv_1.weight"], sd["dp.conv_1.bias"], padding=k // 2)
    x = _ln(F.relu(x), sd["dp.norm_1.gamma"], sd["dp.norm_1.beta"])
    x = x + F.linear(g, sd["dp.cond2.weight"], sd["dp.cond2.bias"]).unsqueeze(-1)
    x = F.conv1d(x if x_mask is None else x * x_mask, sd["dp.conv_2.weight"], sd["dp.conv_2.bias"], padding=k // 2)
    x = _ln(F.relu(x), sd["dp.norm_2.gamma"], sd["dp.norm_2.beta"])
    x = F.conv1d(x if x_mask is None else x * x_mask, sd["dp.proj.weight"], sd["dp.proj.bias"])
    return x if x_mask is None else x * x_mask


def infer_p1(sd, text, emo, sid, cfg: HotPathConfig, x_lengths: Optional[torch.Tensor] = None):
    """SynthesizerTrn.infer_p1 (models.py:558-566): returns m_p, s_p = exp(logs_p), logw, g.  With x_lengths: the
    batched head of inference() (models.py:518-522), masked."""
    g = F.embedding(sid, sd["emb_g.weight"])                  # models.py:561
    x, m_p, logs_p, x_mask = text_encoder(sd, text, emo, g, cfg, x_lengths)
    s_p = torch.exp(logs_p)
    logw = duration_predictor(sd, x, g, cfg, x_mask)
    return m_p, s_p, logw, g


def durations_to_path(logw, length_scale: float = 1.0):
    """EmoVITS.infer's glue (infer.py:169-176, commons.py:143-155): w_ceil, y_length and the 0/1 path, B = 1."""
    w_ceil = torch.ceil(torch.exp(logw) * length_scale)
    y_length = int(torch.clamp_min(torch.sum(w_ceil), 1).item())
    t_x = logw.size(-1)
    cum = torch.cumsum(w_ceil, -1).view(t_x)
    path = (torch.arange(y_length).unsqueeze(0) < cum.unsqueeze(1)).float()      # [t_x, t_y]
    path = path - F.pad(path, (0, 0, 1, 0))[:-1]
    return w_ceil, y_length, path.transpose(0, 1).unsqueeze(0)
