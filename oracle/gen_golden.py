"""Generate tests/golden/*.npz by running the UNMODIFIED reference modules.

Run in the build container only (needs /root/reference, read-only):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.gen_golden

Inputs and weights come from `oracle/fixtures.py` (seeded, reference-free), so
the GPU box can rebuild the same inputs and compare against the committed
outputs.  The reference model is built exactly as export.py:32-38 does
(config via json, because utils.py needs soundfile), the fixture state_dict
is loaded over its `dec.*`/`flow.*` keys (every hot key must match in name and
shape), and outputs come from `SynthesizerTrn.infer_p2` (models.py:568-575),
`flow.infer`/`dec` (models.py:573-574) and the tail of `inference`
(models.py:533-534).
"""
from __future__ import annotations

import json
import os
import sys
import warnings

import numpy as np
import torch

REF = os.environ.get("VITS_REF_PATH", "/root/reference/emotional-vits")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def build_reference(cfgj, sd):
    sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    from models import SynthesizerTrn  # noqa: the reference, imported read-only
    m = SynthesizerTrn(cfgj["data"]["text_channels"], cfgj["data"]["filter_length"] // 2 + 1,
                       cfgj["train"]["segment_size"] // cfgj["data"]["hop_length"],
                       n_speakers=cfgj["data"]["n_speakers"], **cfgj["model"]).eval()
    ref_hot = {k: v for k, v in m.state_dict().items() if k.startswith(("dec.", "flow."))}
    assert set(ref_hot) == set(sd), sorted(set(ref_hot) ^ set(sd))[:5]
    for k in sd:
        assert tuple(ref_hot[k].shape) == tuple(sd[k].shape), k
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all(not k.startswith(("dec.", "flow.")) for k in missing)
    return m


def main():
    from oracle import fixtures as FX
    from vits_b200.config import HotPathConfig
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    torch.set_grad_enabled(False)

    # ---------------- base.json ----------------
    cfgj = json.load(open(os.path.join(REF, "configs", "base.json")))
    cfg = HotPathConfig.from_json(cfgj)
    sd_wn = FX.make_state_dict(cfg, seed=1234, weight_norm=True)
    m = build_reference(cfgj, sd_wn)      # weight-normed load (checkpoint-on-disk form)
    import commons  # reference commons.py (sequence_mask); importable once REF is on sys.path
    # G1: infer_p2, B=1, T=33 (odd, not a tile multiple), weight-normed modules
    inp = FX.make_p2_inputs(cfg, 33, seed=1234)
    o_wn = m.infer_p2(inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"])
    m.remove_weight_norm()                # models.py:467-474, as infer.py:65 does
    o = m.infer_p2(inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"])
    assert (o - o_wn).abs().max() < 1e-5
    z_p = torch.matmul(inp["attn"], inp["m_p"].transpose(1, 2)).transpose(1, 2) + inp["noise"] * \
        torch.matmul(inp["attn"], inp["s_p"].transpose(1, 2)).transpose(1, 2)
    z = m.flow.infer(z_p, g=inp["g"], reverse=True)
    taps = {}
    hooks = [m.dec.conv_pre.register_forward_hook(lambda mod, i, out: taps.__setitem__("conv_pre", out))]
    for i in range(4):
        hooks.append(m.dec.ups[i].register_forward_hook(
            lambda mod, inp_, out, i=i: taps.__setitem__(f"ups.{i}", out)))
    m.dec(z, g=inp["g"])
    for h in hooks:
        h.remove()
    np.savez_compressed(os.path.join(OUT, "base_p2_T33.npz"), o=o.numpy(), z=z.numpy(), z_p=z_p.numpy(),
                        **{"tap_" + k: v[0, :8].numpy() for k, v in taps.items()})
    # G2: tiny lengths around the halo sizes (T=1,2,5)
    tiny = {}
    for T in (1, 2, 5):
        lat = FX.make_latents(cfg, [T], seed=100 + T)
        zz = m.flow.infer(lat["z_p"], g=lat["g"], reverse=True)
        tiny[f"z_T{T}"] = zz.numpy()
        tiny[f"o_T{T}"] = m.dec(zz, g=lat["g"]).numpy()
    np.savez_compressed(os.path.join(OUT, "base_tiny.npz"), **tiny)
    # G3: batched, padded semantics of inference() (models.py:533-534) and per-row B=1 (ragged)
    lens = [20, 33, 7]
    lat = FX.make_latents(cfg, lens, seed=77)
    mask = commons.sequence_mask(lat["lengths"], max(lens)).unsqueeze(1).float()
    zb = m.flow(lat["z_p"], mask, g=lat["g"], reverse=True)
    ob = m.dec((zb * mask)[:, :, :None], g=lat["g"])
    rag = {}
    for b, L in enumerate(lens):
        zr = m.flow.infer(lat["z_p"][b:b + 1, :, :L], g=lat["g"][b:b + 1], reverse=True)
        rag[f"z_row{b}"] = zr.numpy()
        rag[f"o_row{b}"] = m.dec(zr, g=lat["g"][b:b + 1]).numpy()
    np.savez_compressed(os.path.join(OUT, "base_batch3.npz"), z_padded=zb.numpy(), o_padded=ob.numpy(),
                        lengths=np.array(lens), **rag)

    # ---------------- infer_p1: text encoder + duration predictor ----------------
    sd1 = FX.make_p1_state_dict(cfg, seed=4321)
    ref_p1 = {k: v for k, v in m.state_dict().items() if k.startswith(("enc_p.", "dp.", "emb_g."))}
    assert set(ref_p1) == set(sd1), sorted(set(ref_p1) ^ set(sd1))[:6]
    for k in sd1:
        assert tuple(ref_p1[k].shape) == tuple(sd1[k].shape), k
    missing, unexpected = m.load_state_dict(sd1, strict=False)
    assert not unexpected and all(not k.startswith(("enc_p.", "dp.", "emb_g.")) for k in missing)
    p1 = {}
    for Tx, seed in ((60, 99), (1, 3), (5, 4), (131, 7)):
        inp1 = FX.make_p1_inputs(cfg, [Tx], seed=seed)
        m_p, s_p, logw, g1 = m.infer_p1(inp1["text"], inp1["emo"], inp1["sid"])      # models.py:558-566
        p1.update({f"m_p_T{Tx}": m_p.numpy(), f"s_p_T{Tx}": s_p.numpy(), f"logw_T{Tx}": logw.numpy(), f"g_T{Tx}": g1.numpy()})
    # batched, masked head of inference() (models.py:518-522)
    lens1 = [23, 60, 7]
    inpb = FX.make_p1_inputs(cfg, lens1, seed=11)
    gb = m.emb_g(inpb["sid"])
    xb, mb, logsb, xmask = m.enc_p(inpb["text"], inpb["lengths"], inpb["emo"], g=gb)
    logwb = m.dp(xb, xmask, g=gb)
    p1.update({"m_p_batch": mb.numpy(), "logs_p_batch": logsb.numpy(), "logw_batch": logwb.numpy(), "lengths_batch": np.array(lens1)})
    np.savez_compressed(os.path.join(OUT, "base_p1.npz"), **p1)

    # ---------------- sat/configs/adapt.json (second shape set) ----------------
    cfga = json.load(open(os.path.join(REF, "sat", "configs", "adapt.json")))
    cfga["data"].setdefault("hop_length", 96)
    cfga["train"].setdefault("segment_size", 9216)
    cfgA = HotPathConfig.from_json(cfga)
    sdA = FX.make_state_dict(cfgA, seed=4321, weight_norm=False)
    sys.path.insert(0, REF)
    from models import SynthesizerTrn
    mA = SynthesizerTrn(cfga["data"]["text_channels"], cfga["data"]["filter_length"] // 2 + 1,
                        cfga["train"]["segment_size"] // cfga["data"]["hop_length"],
                        n_speakers=cfga["data"]["n_speakers"], **cfga["model"]).eval()
    mA.remove_weight_norm()
    missing, unexpected = mA.load_state_dict(sdA, strict=False)
    assert not unexpected and all(not k.startswith(("dec.", "flow.")) for k in missing), (missing[:4], unexpected[:4])
    latA = FX.make_latents(cfgA, [19], seed=5, noise_scale=0.382)
    zA = mA.flow.infer(latA["z_p"], g=latA["g"], reverse=True)
    oA = mA.dec(zA, g=latA["g"])
    np.savez_compressed(os.path.join(OUT, "adapt_T19.npz"), z=zA.numpy(), o=oA.numpy())
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
