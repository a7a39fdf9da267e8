"""GPU: the drop-in itself.  `vits_b200.install(model)` on a module tree with the reference's attribute names
(tests/ref_standin.py; the real SynthesizerTrn when VITS_REF_PATH exists) must leave signatures, dtypes and layouts
untouched and route infer_p2 / flow.infer / flow.forward(reverse=True) / dec.forward through libvitsb200.so."""
import os

import numpy as np
import pytest
import torch

from oracle import fixtures as FX
from tests.ref_standin import SynthesizerStandIn, load_fixture

pytestmark = pytest.mark.gpu

TOL = {"fp32": (1e-5, 2e-5), "fp16": (2e-3, 4e-3)}     # (waveform, z)


def _model(base_cfg, weight_normed, half=False):
    sd = FX.make_state_dict(base_cfg, seed=1234, weight_norm=True)
    m = load_fixture(SynthesizerStandIn(base_cfg).eval(), sd, True)
    if not weight_normed:
        m.remove_weight_norm()                       # infer.py:65
    m = m.half() if half else m                      # infer.py:66
    return m.cuda()


@pytest.mark.parametrize("prec", ["fp32", "fp16"])
@pytest.mark.parametrize("weight_normed", [True, False])
def test_install_rebinds_reference_entry_points(base_cfg, golden_dir, prec, weight_normed):
    import vits_b200
    m = _model(base_cfg, weight_normed)
    eng = vits_b200.install(m, precision=prec)
    assert m._vitsb is eng
    tw, tz = TOL[prec]
    gold = np.load(os.path.join(golden_dir, "base_p2_T33.npz"))
    inp = {k: v.cuda() for k, v in FX.make_p2_inputs(base_cfg, 33, seed=1234).items()}
    # SynthesizerTrn.infer_p2(attn, m_p, s_p, g, noise)                      models.py:568-575, infer.py:179
    o = m.infer_p2(inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"])
    assert o.is_cuda and o.dtype == torch.float32 and tuple(o.shape) == (1, 1, 33 * 192)
    assert np.abs(o.cpu().numpy() - gold["o"]).max() <= tw
    # flow.infer(z_p, g=g, reverse=True) and dec(z, g=g), the calls of SynthesizerTrn.infer    models.py:554-555
    z = m.flow.infer(torch.from_numpy(gold["z_p"]).cuda(), g=inp["g"], reverse=True)
    assert np.abs(z.cpu().numpy() - gold["z"]).max() <= tz
    o2 = m.dec(torch.from_numpy(gold["z"]).cuda(), g=inp["g"])
    assert np.abs(o2.cpu().numpy() - gold["o"]).max() <= tw
    assert m.dec.infer(torch.from_numpy(gold["z"]).cuda(), g=inp["g"]).shape == o2.shape      # models.py:304 alias
    # batched tail of inference(): self.flow(z_p, y_mask, g=g, reverse=True); self.dec((z * y_mask)[:, :, :max_len], g=g)
    gb = np.load(os.path.join(golden_dir, "base_batch3.npz"))
    lens = [int(x) for x in gb["lengths"]]
    lat = FX.make_latents(base_cfg, lens, seed=77)
    y_mask = (torch.arange(max(lens))[None, :] < torch.tensor(lens)[:, None]).unsqueeze(1).float().cuda()
    op, zm = m.inference_tail(lat["z_p"].cuda(), y_mask, lat["g"].cuda())
    assert np.abs(zm.cpu().numpy() - gb["z_padded"]).max() <= tz
    assert np.abs(op.cpu().numpy() - gb["o_padded"]).max() <= tw
    # the training direction keeps the reference code
    assert m.flow(lat["z_p"].cuda(), y_mask, g=lat["g"].cuda(), reverse=False) == "reference-forward-direction"
    with pytest.raises(NotImplementedError):
        m.flow.infer(lat["z_p"].cuda(), g=lat["g"].cuda(), reverse=False)


def test_install_on_half_model_returns_model_dtype(base_cfg, golden_dir):
    """infer.py:66 runs the model in fp16: inputs arrive as half tensors, outputs must come back as half on the
    same device (the caller applies .float(), infer.py:182)."""
    import vits_b200
    m = _model(base_cfg, weight_normed=False, half=True)
    vits_b200.install(m, precision="fp16")
    gold = np.load(os.path.join(golden_dir, "base_p2_T33.npz"))
    inp = {k: v.cuda().half() for k, v in FX.make_p2_inputs(base_cfg, 33, seed=1234).items()}
    o = m.infer_p2(inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"])
    assert o.dtype == torch.float16 and o.is_cuda
    # half-rounded weights and inputs: the reference's own fp16 noise floor is ~3e-3 here; this checks plumbing, not parity
    assert np.abs(o.float().cpu().numpy() - gold["o"]).max() <= 2e-2
    z = m.flow.infer(torch.from_numpy(gold["z_p"]).cuda().half(), g=inp["g"], reverse=True)
    assert z.dtype == torch.float16


def test_reinstall_after_load_state_dict_repacks(base_cfg):
    import vits_b200
    m = _model(base_cfg, weight_normed=False)
    vits_b200.install(m, precision="fp32")
    lat = FX.make_latents(base_cfg, [40], seed=9)
    w1 = m.dec(lat["z_p"].cuda(), g=lat["g"].cuda())
    sd2 = FX.make_state_dict(base_cfg, seed=99)
    own = m.state_dict()
    m.load_state_dict({k: v.reshape(own[k].shape) for k, v in sd2.items()})
    vits_b200.install(m, precision="fp32")
    w2 = m.dec(lat["z_p"].cuda(), g=lat["g"].cuda())
    from vits_b200 import FlowDecoder
    ref = FlowDecoder(base_cfg, sd2, device="cuda:0", precision="fp32").decode(lat["z_p"].cuda(), lat["g"].cuda())
    assert torch.equal(w2, ref) and not torch.equal(w1, w2)


def test_install_rebinds_infer_p1_when_the_checkpoint_has_it(base_cfg, golden_dir):
    """models.py:558-566: with enc_p.* / dp.* / emb_g.* in the state_dict, install() also routes infer_p1 through the library."""
    import vits_b200
    from tests.ref_standin import attach_param_tree
    m = _model(base_cfg, weight_normed=False).cpu()
    attach_param_tree(m, FX.make_p1_state_dict(base_cfg, seed=4321))
    m = m.cuda()
    assert "infer_p1" not in m.__dict__
    eng = vits_b200.install(m, precision="fp16")
    assert eng.get_option("plan.p1") == 1 and "infer_p1" in m.__dict__
    gold = np.load(os.path.join(golden_dir, "base_p1.npz"))
    inp = FX.make_p1_inputs(base_cfg, [60], seed=99)
    m_p, s_p, logw, g = m.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"].cuda())
    assert m_p.dtype == torch.float32 and m_p.is_cuda
    assert np.abs(m_p.cpu().numpy() - gold["m_p_T60"]).max() <= 2e-4
    assert np.abs(logw.cpu().numpy() - gold["logw_T60"]).max() <= 2e-4
    m2 = _model(base_cfg, weight_normed=False)                      # no encoder weights: infer_p1 is left alone
    vits_b200.install(m2, precision="fp16")
    assert "infer_p1" not in m2.__dict__
