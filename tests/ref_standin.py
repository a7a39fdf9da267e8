"""Stand-in for the reference's `SynthesizerTrn` with exactly the attributes `vits_b200.install` touches.

TEST INFRASTRUCTURE.  /root/reference does not exist on the GPU box, so the drop-in is exercised against an
`nn.Module` tree that has the reference's attribute names, parameter names and call conventions
(emotional-vits/models.py:192-235 ResidualCouplingBlock, :282-318 Generator, :517-575 SynthesizerTrn.inference /
infer_p2; modules.py:93-128 WN, :223-260 ResBlock2, :314-375 ResidualCouplingLayer) and whose forward methods run the
CPU oracle on the module's own state_dict.  `tests/test_install_cpu.py` checks, in the build container where the
reference is importable, that the real `SynthesizerTrn` exposes the same attributes and state_dict keys.
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch.nn.utils import weight_norm

from oracle import vits_oracle as O


class Flip(nn.Module):
    pass


class WN(nn.Module):
    def __init__(self, hidden, k, dilation_rate, n_layers, gin):
        super().__init__()
        self.in_layers, self.res_skip_layers = nn.ModuleList(), nn.ModuleList()
        self.cond_layer = weight_norm(nn.Linear(gin, 2 * hidden * n_layers))
        for i in range(n_layers):
            d = dilation_rate ** i
            self.in_layers.append(weight_norm(nn.Conv1d(hidden, 2 * hidden, k, dilation=d, padding=(k * d - d) // 2)))
            self.res_skip_layers.append(weight_norm(nn.Conv1d(hidden, 2 * hidden if i < n_layers - 1 else hidden, 1)))


class ResidualCouplingLayer(nn.Module):
    def __init__(self, channels, hidden, k, dilation_rate, n_layers, gin):
        super().__init__()
        self.pre = nn.Conv1d(channels // 2, hidden, 1)
        self.enc = WN(hidden, k, dilation_rate, n_layers, gin)
        self.post = nn.Conv1d(hidden, channels // 2, 1)


class ResidualCouplingBlock(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.channels, self.hidden_channels = cfg.inter_channels, cfg.hidden_channels
        self.kernel_size, self.dilation_rate = cfg.flow_kernel_size, list(cfg.dilation_rate)
        self.n_layers, self.n_flows, self.gin_channels = cfg.flow_n_layers, cfg.n_flows, cfg.gin_channels
        self._cfg = cfg
        self.flows = nn.ModuleList()
        for i in range(cfg.n_flows):
            self.flows.append(ResidualCouplingLayer(cfg.inter_channels, cfg.hidden_channels, cfg.flow_kernel_size,
                                                    cfg.dilation_rate[i], cfg.flow_n_layers, cfg.gin_channels))
            self.flows.append(Flip())

    def _sd(self):
        return {"flow." + k: v.float() for k, v in self.state_dict().items()}

    def forward(self, x, x_mask, g=None, reverse=False):
        if not reverse:
            return "reference-forward-direction"          # sentinel: install must leave this path alone
        return O.flow_reverse(self._sd(), x.float(), g.float(), self._cfg, x_mask=x_mask.float()).to(x.dtype)

    def infer(self, x, g, reverse=True):
        return O.flow_reverse(self._sd(), x.float(), g.float(), self._cfg).to(x.dtype)


class ResBlock2(nn.Module):
    def __init__(self, ch, k, dils, gin):
        super().__init__()
        inter = (ch // 16) * 16
        self.convs1 = nn.ModuleList([weight_norm(nn.Conv1d(ch, inter, k, dilation=d, padding=(k * d - d) // 2)) for d in dils])
        self.convs2 = nn.ModuleList([weight_norm(nn.Conv1d(inter // 2, ch, k, padding=(k - 1) // 2)) for _ in dils])
        self.conds = nn.ModuleList([weight_norm(nn.Linear(gin, inter)) for _ in dils])


class Generator(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self._cfg = cfg
        self.num_kernels, self.num_upsamples = len(cfg.resblock_kernel_sizes), len(cfg.upsample_rates)
        C0 = cfg.upsample_initial_channel
        self.conv_pre = nn.Conv1d(cfg.inter_channels, C0, 7, 1, padding=3)
        self.ups, self.resblocks = nn.ModuleList(), nn.ModuleList()
        ch = C0
        for i, (u, k) in enumerate(zip(cfg.upsample_rates, cfg.upsample_kernel_sizes)):
            self.ups.append(weight_norm(nn.ConvTranspose1d(C0 // 2 ** i, C0 // 2 ** (i + 1), k, u, padding=(k - u) // 2)))
            ch = C0 // 2 ** (i + 1)
            for k2, d in zip(cfg.resblock_kernel_sizes, cfg.resblock_dilation_sizes):
                self.resblocks.append(ResBlock2(ch, k2, d, cfg.gin_channels))
        self.conv_post = nn.Conv1d(ch, 1, 7, 1, padding=3, bias=False)
        self.infer = self.forward

    def forward(self, x, g):
        sd = {"dec." + k: v.float() for k, v in self.state_dict().items()}
        return O.generator(sd, x.float(), g.float(), self._cfg).to(x.dtype)


class SynthesizerStandIn(nn.Module):
    """`dec`, `flow`, `infer_p2` and the batched tail of `inference` (models.py:529-535) of the reference model."""

    def __init__(self, cfg):
        super().__init__()
        self.dec = Generator(cfg)
        self.flow = ResidualCouplingBlock(cfg)

    def remove_weight_norm(self):
        def _rm(m):
            try:
                torch.nn.utils.remove_weight_norm(m)
            except ValueError:
                return
        self.apply(_rm)

    @torch.no_grad()
    def infer_p2(self, attn, m_p, s_p, g, noise):
        z_p = O.length_regulate(attn, m_p, s_p, noise)
        z = self.flow.infer(z_p, g=g, reverse=True)
        return self.dec(z, g=g)

    @torch.no_grad()
    def inference_tail(self, z_p, y_mask, g, max_len=None):
        z = self.flow(z_p, y_mask, g=g, reverse=True)                     # models.py:533
        return self.dec((z * y_mask)[:, :, :max_len], g=g), z             # models.py:534


def attach_param_tree(model: nn.Module, sd: dict):
    """Register the tensors of a flat state_dict (e.g. the enc_p.* / dp.* / emb_g.* keys of the reference model,
    models.py:117-164, 20-47, 465) as parameters under the same dotted names, so that model.state_dict() carries them."""
    for key, val in sd.items():
        parts = key.split(".")
        mod = model
        for p_ in parts[:-1]:
            if isinstance(mod, nn.ModuleDict):
                if p_ not in mod:
                    mod[p_] = nn.ModuleDict()
                mod = mod[p_]
            else:
                if not hasattr(mod, p_):
                    setattr(mod, p_, nn.ModuleDict())
                mod = getattr(mod, p_)
        mod.register_parameter(parts[-1], nn.Parameter(val.clone(), requires_grad=False))
    return model


def load_fixture(model: nn.Module, sd: dict, weight_normed: bool):
    """Strict load of the fixture state_dict (weight_g / weight_v form when the model still carries weight norm)."""
    own = model.state_dict()
    missing = set(own) ^ set(sd)
    assert not missing, sorted(missing)[:8]
    model.load_state_dict({k: v.reshape(own[k].shape) for k, v in sd.items()}, strict=True)
    return model
