"""Drop-in installation behind the reference's own `SynthesizerTrn`.

    model = export.load_model(ckpt, hps)          # reference, unchanged (export.py:22-61)
    vits_b200.install(model, precision="fp16")    # after load_state_dict, before or after remove_weight_norm
    wav = model.infer_p2(attn, m_p, s_p, g, noise)    # same signature, same layout, now on libvitsb200.so

What is rebound (signatures unchanged; paths relative to emotional-vits/):
  model.infer_p2                      models.py:568-575
  model.flow.infer / .forward(reverse=True)   models.py:219-235   (forward(reverse=False) keeps the reference code)
  model.dec.forward / .infer          models.py:304-318
  model.infer_p1                      models.py:558-566   (when the state_dict carries enc_p.* / dp.* / emb_g.*)
`model.infer` (models.py:537-556) and `model.inference` (models.py:517-535) call
`self.flow`/`self.dec` and therefore pick the new path up without being touched.
The text encoder and duration predictor run on the library too (fp32) when their weights are present; `model.infer`'s own
noise sampling and length regulation stay reference PyTorch (they call the rebound `flow` / `dec`).
"""
from __future__ import annotations

import types
from typing import Optional

import torch

from .engine import FlowDecoder


def config_from_model(model) -> dict:
    """Read the hot-path hyper-parameters off the constructed reference modules
    (models.py:192-217, 282-304; modules.py:223-248)."""
    flow, dec = model.flow, model.dec
    nk = dec.num_kernels
    rb_type = type(dec.resblocks[0]).__name__
    if rb_type != "ResBlock2":
        raise RuntimeError(f"vits_b200 supports resblock '2' only, model has {rb_type} "
                           "(ResBlock1 is shape-inconsistent in the reference, modules.py:185-220)")
    return dict(
        inter_channels=flow.channels,
        hidden_channels=flow.hidden_channels,
        gin_channels=flow.gin_channels,
        n_flows=flow.n_flows,
        flow_kernel_size=flow.kernel_size,
        flow_n_layers=flow.n_layers,
        dilation_rate=list(flow.dilation_rate),
        resblock="2",
        resblock_kernel_sizes=[dec.resblocks[j].convs1[0].kernel_size[0] for j in range(nk)],
        resblock_dilation_sizes=[[c.dilation[0] for c in dec.resblocks[j].convs1] for j in range(nk)],
        upsample_rates=[u.stride[0] for u in dec.ups],
        upsample_kernel_sizes=[u.kernel_size[0] for u in dec.ups],
        upsample_initial_channel=dec.conv_pre.out_channels,
    )


def install(model, precision: str = "fp16", device: Optional[torch.device] = None) -> FlowDecoder:
    """Pack `model.flow` / `model.dec` weights into a FlowDecoder and rebind the
    reference entry points to it.  Call again after `load_state_dict` or a dtype /
    device change to re-pack.  Returns the engine (also stored as `model._vitsb`)."""
    if device is None:
        device = next(model.parameters()).device
    if torch.device(device).type != "cuda":
        raise RuntimeError("vits_b200.install: move the model to a CUDA device first (no CPU fallback)")
    eng = FlowDecoder(config_from_model(model), model.state_dict(), device=device, precision=precision)
    model._vitsb = eng
    ref_flow_forward = model.flow.forward

    def _dtype():
        return next(model.dec.parameters()).dtype

    def flow_infer(self, x, g, reverse=True):
        if not reverse:
            raise NotImplementedError("vits_b200 accelerates the reverse flow only")
        return eng.flow_reverse(x, g).to(x.dtype)

    def flow_forward(self, x, x_mask, g=None, reverse=False):
        if not reverse:
            return ref_flow_forward(x, x_mask, g=g, reverse=False)      # training direction: reference code
        lengths = x_mask.reshape(x_mask.size(0), -1).sum(-1).round().to(torch.int64)
        return eng.flow_reverse(x, g, lengths.tolist()).to(x.dtype)

    def dec_forward(self, x, g):
        return eng.decode(x, g).to(x.dtype)

    def infer_p2(self, attn, m_p, s_p, g, noise):
        return eng.infer_p2(attn, m_p, s_p, g, noise).to(_dtype())

    def infer_p1(self, x, emo, sid):
        # models.py:558-566 asserts B == 1; outputs in the model's dtype, as the reference returns them
        assert x.size(0) == 1
        dt = _dtype()
        return tuple(t.to(dt) for t in eng.infer_p1(x, emo, sid))

    if eng.get_option("plan.p1") == 1:                  # the checkpoint carried enc_p.* / dp.* / emb_g.*
        model.infer_p1 = types.MethodType(infer_p1, model)
    model.flow.infer = types.MethodType(flow_infer, model.flow)
    model.flow.forward = types.MethodType(flow_forward, model.flow)
    model.dec.forward = types.MethodType(dec_forward, model.dec)
    model.dec.infer = model.dec.forward
    model.infer_p2 = types.MethodType(infer_p2, model)
    return eng
