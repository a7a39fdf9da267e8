"""Deterministic synthetic weights and inputs for the hot path.

TEST INFRASTRUCTURE.  Generates the `dec.*` / `flow.*` state_dict layout of
SURVEY.md Appendix A (models.py:282-304, modules.py:93-128,223-248,314-339)
straight from the config, without the reference, so that fixtures can be
re-created on the GPU box where /root/reference does not exist.

Why not the reference's default init: `flow.*.post` is zero-initialised
(modules.py:338-339) which makes the reverse flow an identity, and every
`std=0.01` init (commons.py:8-10) gives a waveform of std 0.01; parity tests
on such weights are vacuous.  Here every hot weight is drawn at fan-in scale.
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import torch

from vits_b200.config import HotPathConfig


def hot_param_specs(cfg: HotPathConfig) -> List[Tuple[str, Tuple[int, ...], str, int]]:
    """[(key, shape, kind, fan_in)] for folded (no weight_g/v) hot keys, in a
    fixed order.  kind in {conv, convT, linear, bias}; weight-normed kinds are
    suffixed '+wn' (modules.py:110,117,127,228,236,244; models.py:291)."""
    specs = []
    C0, gin = cfg.upsample_initial_channel, cfg.gin_channels
    specs.append(("dec.conv_pre.weight", (C0, cfg.inter_channels, 7), "conv", cfg.inter_channels * 7))
    specs.append(("dec.conv_pre.bias", (C0,), "bias", 0))
    ch = C0
    for i, (u, k) in enumerate(zip(cfg.upsample_rates, cfg.upsample_kernel_sizes)):
        cin, cout = C0 // (2 ** i), C0 // (2 ** (i + 1))
        specs.append((f"dec.ups.{i}.weight", (cin, cout, k), "convT+wn", max(1, cin * k // u)))
        specs.append((f"dec.ups.{i}.bias", (cout,), "bias", 0))
        ch = cout
        for j, kk in enumerate(cfg.resblock_kernel_sizes):
            r = i * len(cfg.resblock_kernel_sizes) + j
            inter = (ch // 16) * 16                                # modules.py:226
            for d in range(len(cfg.resblock_dilation_sizes[j])):
                p = f"dec.resblocks.{r}."
                specs.append((p + f"convs1.{d}.weight", (inter, ch, kk), "conv+wn", ch * kk))
                specs.append((p + f"convs1.{d}.bias", (inter,), "bias", 0))
                specs.append((p + f"convs2.{d}.weight", (ch, inter // 2, kk), "conv+wn", (inter // 2) * kk))
                specs.append((p + f"convs2.{d}.bias", (ch,), "bias", 0))
                specs.append((p + f"conds.{d}.weight", (inter, gin), "linear+wn", gin))
                specs.append((p + f"conds.{d}.bias", (inter,), "bias", 0))
    specs.append(("dec.conv_post.weight", (1, ch, 7), "conv", ch * 7))
    H, half, k, nl = cfg.hidden_channels, cfg.inter_channels // 2, cfg.flow_kernel_size, cfg.flow_n_layers
    for f in range(cfg.n_flows):
        p = f"flow.flows.{2 * f}."
        specs.append((p + "pre.weight", (H, half, 1), "conv", half))
        specs.append((p + "pre.bias", (H,), "bias", 0))
        for l in range(nl):
            specs.append((p + f"enc.in_layers.{l}.weight", (2 * H, H, k), "conv+wn", H * k))
            specs.append((p + f"enc.in_layers.{l}.bias", (2 * H,), "bias", 0))
            rs = 2 * H if l < nl - 1 else H                       # modules.py:121-124
            specs.append((p + f"enc.res_skip_layers.{l}.weight", (rs, H, 1), "conv+wn", H))
            specs.append((p + f"enc.res_skip_layers.{l}.bias", (rs,), "bias", 0))
        specs.append((p + "enc.cond_layer.weight", (2 * H * nl, gin), "linear+wn", gin))
        specs.append((p + "enc.cond_layer.bias", (2 * H * nl,), "bias", 0))
        specs.append((p + "post.weight", (half, H, 1), "conv", H))
        specs.append((p + "post.bias", (half,), "bias", 0))
    return specs


def make_state_dict(cfg: HotPathConfig, seed: int = 1234, weight_norm: bool = False,
                    gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """fp32 hot-path state_dict.  With weight_norm=True the weight-normed
    tensors are emitted as `weight_g`/`weight_v` whose fold equals the
    weight_norm=False tensors (up to fp32 rounding)."""
    gen = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    for key, shape, kind, fan_in in hot_param_specs(cfg):
        if kind == "bias":
            sd[key] = (torch.rand(shape, generator=gen) * 0.2 - 0.1)
            continue
        # conditioning on a unit-variance g adds to pre-gate activations: keep it O(0.5)
        scale = gain / math.sqrt(fan_in)
        if kind.startswith("linear"):
            scale *= 0.5
        w = torch.randn(shape, generator=gen) * scale
        if weight_norm and kind.endswith("+wn"):
            dims = tuple(range(1, w.dim()))
            row_scale = torch.rand((shape[0],) + (1,) * (w.dim() - 1), generator=gen) + 0.5
            sd[key + "_g"] = w.pow(2).sum(dim=dims, keepdim=True).sqrt()
            sd[key + "_v"] = w * row_scale
        else:
            if kind.endswith("+wn"):
                # consume the same random numbers so both variants share every other tensor
                torch.rand((shape[0],) + (1,) * (w.dim() - 1), generator=gen)
            sd[key] = w
    return sd


def make_lengths(B: int, lo: int, hi: int, seed: int) -> List[int]:
    gen = torch.Generator().manual_seed(seed)
    return [int(x) for x in torch.randint(lo, hi + 1, (B,), generator=gen)]


def make_latents(cfg: HotPathConfig, lengths: List[int], seed: int = 1234, noise_scale: float = 0.707):
    """Synthetic `z_p`-side inputs per SURVEY.md 8(d): returns dict with
    z_p [B,192,Tmax] (already length-regulated: m_p + noise*s_p on valid
    frames, zero beyond), g [B,gin], lengths."""
    gen = torch.Generator().manual_seed(seed)
    B, Tm, C = len(lengths), max(lengths), cfg.inter_channels
    z_p = torch.zeros(B, C, Tm)
    for b, L in enumerate(lengths):
        m = torch.randn(C, L, generator=gen)
        s = torch.rand(C, L, generator=gen) * 0.7 + 0.3
        n = torch.randn(C, L, generator=gen) * noise_scale
        z_p[b, :, :L] = m + n * s
    g = torch.randn(B, cfg.gin_channels, generator=gen)
    return {"z_p": z_p, "g": g, "lengths": torch.tensor(lengths, dtype=torch.int64)}


def make_p2_inputs(cfg: HotPathConfig, T: int, seed: int = 1234, noise_scale: float = 0.707):
    """Inputs of `infer_p2(attn, m_p, s_p, g, noise)` (models.py:569; shapes per
    export.py:171-180): Tx = ceil(T/7) phonemes with integer durations summing
    to T, hard monotonic attn as built by commons.infer_path (commons.py:143-155)."""
    gen = torch.Generator().manual_seed(seed)
    C = cfg.inter_channels
    Tx = max(1, (T + 6) // 7)
    dur = torch.full((Tx,), T // Tx, dtype=torch.int64)
    dur[: T - int(dur.sum())] += 1
    perm = torch.randperm(Tx, generator=gen)
    dur = dur[perm]
    cum = torch.cumsum(dur, 0)
    t = torch.arange(T).unsqueeze(1)
    attn = ((t < cum.unsqueeze(0)) & (t >= (cum - dur).unsqueeze(0))).float().unsqueeze(0)  # [1,T,Tx]
    m_p = torch.randn(1, C, Tx, generator=gen)
    s_p = torch.rand(1, C, Tx, generator=gen) * 0.7 + 0.3
    noise = torch.randn(1, C, T, generator=gen) * noise_scale
    g = torch.randn(1, cfg.gin_channels, generator=gen)
    return {"attn": attn, "m_p": m_p, "s_p": s_p, "g": g, "noise": noise}


def p1_param_specs(cfg: HotPathConfig) -> List[Tuple[str, Tuple[int, ...], str, int]]:
    """[(key, shape, kind, fan_in)] of the `enc_p.*`, `dp.*`, `emb_g.*` keys (models.py:20-47, 117-164;
    attentions.py:12-33, 58-77, 128-146; none of them weight-normed)."""
    H, F_, tc, gin, k = cfg.hidden_channels, cfg.filter_channels, cfg.text_channels, cfg.gin_channels, cfg.kernel_size
    sp = [("enc_p.alpha", (), "scalar", 0),
          ("enc_p.emb.0.weight", (H, tc), "linear", tc), ("enc_p.emb.0.bias", (H,), "bias", 0),
          ("enc_p.emb.1.weight", (H,), "gamma", 0), ("enc_p.emb.1.bias", (H,), "bias", 0),
          ("enc_p.emo_proj.weight", (H, cfg.emo_channels), "linear", cfg.emo_channels), ("enc_p.emo_proj.bias", (H,), "bias", 0)]
    for i in range(cfg.n_layers):
        a = f"enc_p.encoder.attn_layers.{i}."
        for n in ("q", "k", "v", "o"):
            sp += [(a + f"conv_{n}.weight", (H, H, 1), "conv", H), (a + f"conv_{n}.bias", (H,), "bias", 0)]
        sp += [(f"enc_p.encoder.norm_layers_1.{i}.gamma", (H,), "gamma", 0), (f"enc_p.encoder.norm_layers_1.{i}.beta", (H,), "bias", 0)]
        f = f"enc_p.encoder.ffn_layers.{i}."
        sp += [(f + "conv_1.weight", (2 * F_, H, k), "conv", H * k), (f + "conv_1.bias", (2 * F_,), "bias", 0),
               (f + "conv_2.weight", (H, F_, k), "conv", F_ * k), (f + "conv_2.bias", (H,), "bias", 0),
               (f + "cond.weight", (2 * F_, gin), "linear", gin), (f + "cond.bias", (2 * F_,), "bias", 0)]
        sp += [(f"enc_p.encoder.norm_layers_2.{i}.gamma", (H,), "gamma", 0), (f"enc_p.encoder.norm_layers_2.{i}.beta", (H,), "bias", 0)]
    sp += [("enc_p.proj.weight", (2 * cfg.inter_channels, H, 1), "conv", H), ("enc_p.proj.bias", (2 * cfg.inter_channels,), "bias", 0)]
    D, kd = cfg.hidden_size_d, cfg.kernel_size_d
    sp += [("dp.pre.weight", (D, H, 1), "conv", H), ("dp.pre.bias", (D,), "bias", 0),
           ("dp.conv_1.weight", (D, D, kd), "conv", D * kd), ("dp.conv_1.bias", (D,), "bias", 0),
           ("dp.norm_1.gamma", (D,), "gamma", 0), ("dp.norm_1.beta", (D,), "bias", 0),
           ("dp.conv_2.weight", (D, D, kd), "conv", D * kd), ("dp.conv_2.bias", (D,), "bias", 0),
           ("dp.norm_2.gamma", (D,), "gamma", 0), ("dp.norm_2.beta", (D,), "bias", 0),
           ("dp.proj.weight", (1, D, 1), "conv", D), ("dp.proj.bias", (1,), "bias", 0),
           ("dp.cond1.weight", (D, gin), "linear", gin), ("dp.cond1.bias", (D,), "bias", 0),
           ("dp.cond2.weight", (D, gin), "linear", gin), ("dp.cond2.bias", (D,), "bias", 0),
           ("emb_g.weight", (cfg.n_speakers, gin), "embed", 0)]
    return sp


def make_p1_state_dict(cfg: HotPathConfig, seed: int = 4321) -> Dict[str, torch.Tensor]:
    """fp32 `enc_p.*` / `dp.*` / `emb_g.*` state_dict at fan-in scale (LayerNorm gains around 1, biases +-0.1,
    emb_g ~ N(0, 1) as nn.Embedding's default); the duration head is biased so that durations land around 3-12 frames."""
    gen = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    for key, shape, kind, fan_in in p1_param_specs(cfg):
        if kind == "scalar":
            sd[key] = torch.tensor(1.0) + 0.1 * torch.randn((), generator=gen)
        elif kind == "bias":
            sd[key] = torch.rand(shape, generator=gen) * 0.2 - 0.1
        elif kind == "gamma":
            sd[key] = 1.0 + 0.2 * (torch.rand(shape, generator=gen) - 0.5)
        elif kind == "embed":
            sd[key] = torch.randn(shape, generator=gen)
        else:
            scale = 1.0 / math.sqrt(fan_in)
            if kind == "linear" and fan_in == cfg.gin_channels:
                scale *= 0.5
            sd[key] = torch.randn(shape, generator=gen) * scale
    # log-scale head of the prior: s_p = exp(logs_p) around 0.3 .. 1.1, the range SURVEY 8(d) states for the path's inputs
    ic = cfg.inter_channels
    sd["enc_p.proj.weight"][ic:] *= 0.25
    sd["enc_p.proj.bias"][ic:] -= 0.5
    sd["dp.proj.bias"] = torch.tensor([1.6])
    sd["dp.proj.weight"] = sd["dp.proj.weight"] * 0.5
    return sd


def make_p1_inputs(cfg: HotPathConfig, lengths: List[int], seed: int = 99):
    """Inputs of infer_p1 / the batched encoder head: text features [B, Tx, text_channels] (zero beyond each row's
    length), emo [B, 1024], sid [B]."""
    gen = torch.Generator().manual_seed(seed)
    B, Tm = len(lengths), max(lengths)
    text = torch.zeros(B, Tm, cfg.text_channels)
    for b, L in enumerate(lengths):
        text[b, :L] = torch.randn(L, cfg.text_channels, generator=gen)
    emo = torch.randn(B, cfg.emo_channels, generator=gen)
    sid = torch.tensor([(b * 37 + 5) % cfg.n_speakers for b in range(B)], dtype=torch.long)
    return {"text": text, "emo": emo, "sid": sid, "lengths": torch.tensor(lengths, dtype=torch.int64)}
