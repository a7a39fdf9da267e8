"""GPU: infer_p1 (text encoder + duration predictor) through the C ABI against the reference-generated
goldens (tests/golden/base_p1.npz, made by oracle/gen_golden.py from the reference's own SynthesizerTrn.infer_p1) and the
CPU oracle (oracle/p1_oracle.py).  fp32 arithmetic on the device; tolerance 2e-4 absolute on m_p / logw (values of
O(1)), 2e-4 relative on s_p = exp(logs_p)."""
import os

import numpy as np
import pytest
import torch

from oracle import fixtures as FX
from oracle import p1_oracle as P
from oracle import vits_oracle as O

pytestmark = pytest.mark.gpu
TOL = 2e-4

_eng = {}


def engine(base_cfg, base_sd, prec="fp16"):
    from vits_b200 import FlowDecoder
    if prec not in _eng:
        sd = dict(base_sd)
        sd.update(FX.make_p1_state_dict(base_cfg, seed=4321))
        _eng[prec] = FlowDecoder(base_cfg, sd, device="cuda:0", precision=prec)
    return _eng[prec]


def test_infer_p1_golden(base_cfg, base_sd, golden_dir):
    gold = np.load(os.path.join(golden_dir, "base_p1.npz"))
    eng = engine(base_cfg, base_sd)
    for Tx, seed in ((60, 99), (1, 3), (5, 4), (131, 7)):
        inp = FX.make_p1_inputs(base_cfg, [Tx], seed=seed)
        m_p, s_p, logw, g = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"])
        assert m_p.shape == (1, 192, Tx) and s_p.shape == (1, 192, Tx) and logw.shape == (1, 1, Tx) and g.shape == (1, 1024)
        assert np.array_equal(g.cpu().numpy(), gold[f"g_T{Tx}"])
        assert np.abs(m_p.cpu().numpy() - gold[f"m_p_T{Tx}"]).max() <= TOL
        assert np.abs(s_p.cpu().numpy() / gold[f"s_p_T{Tx}"] - 1).max() <= TOL
        assert np.abs(logw.cpu().numpy() - gold[f"logw_T{Tx}"]).max() <= TOL
        # the quantity that matters downstream: integer durations (infer.py:169-171)
        w_ref = np.ceil(np.exp(gold[f"logw_T{Tx}"]))
        w = np.ceil(np.exp(logw.cpu().numpy()))
        near = np.abs(np.exp(gold[f"logw_T{Tx}"]) - np.round(np.exp(gold[f"logw_T{Tx}"]))) < 1e-3      # within tolerance of an integer
        assert np.array_equal(w[~near], w_ref[~near])


def test_infer_p1_ragged_batch_equals_per_row_calls(base_cfg, base_sd):
    eng = engine(base_cfg, base_sd)
    sd1 = FX.make_p1_state_dict(base_cfg, seed=4321)
    lens = [23, 60, 7, 1]
    inp = FX.make_p1_inputs(base_cfg, lens, seed=11)
    m_p, s_p, logw, g = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"], lens)
    for _ in range(3):                 # direct issue, graph capture, replay: identical; new inputs flow through the replay
        m2, s2, l2, _ = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"], lens)
        assert torch.equal(m2, m_p) and torch.equal(s2, s_p) and torch.equal(l2, logw)
    inp_b = FX.make_p1_inputs(base_cfg, lens, seed=12)
    m3, _, l3, _ = eng.infer_p1(inp_b["text"].cuda(), inp_b["emo"].cuda(), inp_b["sid"], lens)
    with torch.no_grad():
        mr3, _, lr3, _ = P.infer_p1(sd1, inp_b["text"][1:2, :60], inp_b["emo"][1:2], inp_b["sid"][1:2], base_cfg)
    assert (m3[1:2, :, :60].cpu() - mr3).abs().max().item() <= TOL and (l3[1:2, :, :60].cpu() - lr3).abs().max().item() <= TOL
    for b, L in enumerate(lens):
        with torch.no_grad():
            mr, sr, lr, gr = P.infer_p1(sd1, inp["text"][b:b + 1, :L], inp["emo"][b:b + 1], inp["sid"][b:b + 1], base_cfg)
        assert (m_p[b:b + 1, :, :L].cpu() - mr).abs().max().item() <= TOL
        assert (s_p[b:b + 1, :, :L].cpu() / sr - 1).abs().max().item() <= TOL
        assert (logw[b:b + 1, :, :L].cpu() - lr).abs().max().item() <= TOL
        assert torch.equal(g[b:b + 1].cpu(), gr)
        assert m_p[b, :, L:].abs().sum().item() == 0 and s_p[b, :, L:].abs().sum().item() == 0 and logw[b, :, L:].abs().sum().item() == 0
        # bit-identical to the B = 1 call on that row
        m1, s1, l1, _ = eng.infer_p1(inp["text"][b:b + 1, :L].contiguous().cuda(), inp["emo"][b:b + 1].cuda(), inp["sid"][b:b + 1])
        assert torch.equal(m1[0], m_p[b, :, :L]) and torch.equal(l1[0], logw[b, :, :L]) and torch.equal(s1[0], s_p[b, :, :L])


@pytest.mark.parametrize("prec,tol", [("fp32", 5e-4), ("fp16", 2e-3)])
def test_infer_chain_p1_glue_p2_vs_oracle(base_cfg, base_sd, prec, tol):
    """EmoVITS.infer's chain (infer.py:166-182): infer_p1 -> w_ceil / path glue -> infer_p2, against the CPU oracle chain
    on the same noise."""
    eng = engine(base_cfg, base_sd, prec)
    sd1 = FX.make_p1_state_dict(base_cfg, seed=4321)
    inp = FX.make_p1_inputs(base_cfg, [40], seed=21)
    with torch.no_grad():
        mr, sr, lr, gr = P.infer_p1(sd1, inp["text"], inp["emo"], inp["sid"], base_cfg)
        w_ref, y_ref, path_ref = P.durations_to_path(lr)
        noise = torch.randn(1, 192, y_ref, generator=torch.Generator().manual_seed(5)) * 0.707
        o_ref = O.infer_p2(base_sd, path_ref, mr, sr, gr, noise, base_cfg)
    m_p, s_p, logw, g = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"])
    w, y_len, path = P.durations_to_path(logw.cpu())                 # host glue, as the reference does it (infer.py:169-176)
    assert y_len == y_ref and torch.equal(w, w_ref)
    o = eng.infer_p2(path.cuda(), m_p, s_p, g, noise.cuda()).cpu()
    assert o.shape == o_ref.shape == (1, 1, y_ref * 192)
    assert (o - o_ref).abs().max().item() <= tol


def test_infer_p1_needs_its_weights(base_cfg, base_sd):
    from vits_b200 import FlowDecoder
    from vits_b200 import lib as L
    eng = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision="fp16")      # dec.* / flow.* only
    inp = FX.make_p1_inputs(base_cfg, [5], seed=1)
    with pytest.raises(L.VitsbError):
        eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"])
    full = engine(base_cfg, base_sd)
    with pytest.raises(L.VitsbError):
        full.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), [99999])          # sid out of range (infer.py:145)


@pytest.mark.parametrize("prec", ["fp32", "fp16"])
def test_device_side_durations_equal_the_host_glue(base_cfg, base_sd, prec):
    """vitsb_synthesize_durations (durations, cumsum, lengths and the gather on the device, lengths never
    on the host) == infer_p1 -> host glue (infer.py:169-176) -> infer_p2, and == the CPU oracle chain."""
    eng = engine(base_cfg, base_sd, prec)
    sd1 = FX.make_p1_state_dict(base_cfg, seed=4321)
    tol = {"fp32": 5e-4, "fp16": 2e-3}[prec]
    for Tx, seed, scale in ((40, 21, 1.0), (7, 22, 1.3), (1, 23, 1.0)):
        inp = FX.make_p1_inputs(base_cfg, [Tx], seed=seed)
        m_p, s_p, logw, g = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"])
        w, y_len, path = P.durations_to_path(logw.cpu(), scale)
        T_cap = y_len + 37
        noise = torch.randn(1, 192, T_cap, generator=torch.Generator().manual_seed(seed)) * 0.707
        ref = eng.infer_p2(path.cuda(), m_p, s_p, g, noise[:, :, :y_len].contiguous().cuda())           # existing path, host glue
        wav, yl, z = eng.synthesize_durations(m_p, s_p, logw, g, noise.cuda(), T_cap, length_scale=scale, return_z=True)
        assert int(yl.item()) == y_len
        assert wav.shape == (1, 1, T_cap * 192)
        assert wav[:, :, y_len * 192:].abs().sum().item() == 0 and z[:, :, y_len:].abs().sum().item() == 0
        assert (wav[:, :, :y_len * 192] - ref).abs().max().item() <= 1e-6
        # compact noise: the reference's flat pool viewed as [1, inter, y_length] (infer.py:172-174)
        pool = noise[:, :, :y_len].contiguous().view(-1)
        pool = torch.cat([pool, torch.zeros(192 * T_cap - pool.numel())])
        wav_c, _ = eng.synthesize_durations(m_p, s_p, logw, g, pool.cuda(), T_cap, length_scale=scale, compact_noise=True)
        assert torch.equal(wav_c, wav)
        # against the CPU oracle chain
        with torch.no_grad():
            mr, sr, lr, gr = P.infer_p1(sd1, inp["text"], inp["emo"], inp["sid"], base_cfg)
            wr, yr, pr = P.durations_to_path(lr, scale)
            assert yr == y_len
            o_ref = O.infer_p2(base_sd, pr, mr, sr, gr, noise[:, :, :yr], base_cfg)
        assert (wav[:, :, :yr * 192].cpu() - o_ref).abs().max().item() <= tol
    # ragged batch, truncation at T_cap, int16 output, end-to-end infer()
    lens = [23, 40, 7]
    inp = FX.make_p1_inputs(base_cfg, lens, seed=31)
    m_p, s_p, logw, g = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"], lens)
    T_cap = 150                                                      # shorter than the longest row: truncated
    noise = torch.randn(3, 192, T_cap, generator=torch.Generator().manual_seed(9)) * 0.707
    wav, yl = eng.synthesize_durations(m_p, s_p, logw, g, noise.cuda(), T_cap, x_lengths=lens)
    yl = yl.tolist()
    for b, L in enumerate(lens):
        _, y_b, path = P.durations_to_path(logw[b:b + 1, :, :L].cpu())
        assert yl[b] == min(y_b, T_cap)
        if y_b <= T_cap:
            one = eng.infer_p2(path.cuda(), m_p[b:b + 1, :, :L].contiguous(), s_p[b:b + 1, :, :L].contiguous(), g[b:b + 1],
                               noise[b:b + 1, :, :y_b].contiguous().cuda())
            assert (wav[b:b + 1, :, :y_b * 192] - one).abs().max().item() <= 1e-6
        assert wav[b, :, yl[b] * 192:].abs().sum().item() == 0
    wav2, yl2 = eng.infer(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"], noise.cuda(), T_cap, x_lengths=lens)
    assert torch.equal(wav2, wav) and yl2.tolist() == yl
    pcm, _ = eng.synthesize_durations(m_p, s_p, logw, g, noise.cuda(), T_cap, x_lengths=lens, pcm16=True, volume=0.8)
    assert np.array_equal(pcm.cpu().numpy(), np.clip(wav.cpu().numpy() * np.float32(0.8) * 32767, -32768, 32767).astype(np.int16))
    # repeated calls (graph capture + replay of the workspace-only launches) stay identical
    for _ in range(3):
        w3, _ = eng.synthesize_durations(m_p, s_p, logw, g, noise.cuda(), T_cap, x_lengths=lens)
        assert torch.equal(w3, wav)


def test_infer_p1_long_text_and_empty_rows(base_cfg, base_sd):
    """Tx = 300 takes the global-memory attention kernel (K / V of a head no longer fit shared memory); a row of zero
    phonemes stays empty through infer_p1 and the durations front end; Tx beyond the table limit is refused."""
    from vits_b200 import lib as L
    eng = engine(base_cfg, base_sd)
    sd1 = FX.make_p1_state_dict(base_cfg, seed=4321)
    inp = FX.make_p1_inputs(base_cfg, [300], seed=41)
    m_p, s_p, logw, g = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"])
    with torch.no_grad():
        mr, sr, lr, _ = P.infer_p1(sd1, inp["text"], inp["emo"], inp["sid"], base_cfg)
    assert (m_p.cpu() - mr).abs().max().item() <= TOL and (logw.cpu() - lr).abs().max().item() <= TOL
    assert (s_p.cpu() / sr - 1).abs().max().item() <= TOL
    # ragged batch with an empty row
    lens = [9, 0, 14]
    inp = FX.make_p1_inputs(base_cfg, [9, 1, 14], seed=42)
    m_p, s_p, logw, g = eng.infer_p1(inp["text"].cuda(), inp["emo"].cuda(), inp["sid"], lens)
    assert m_p[1].abs().sum().item() == 0 and logw[1].abs().sum().item() == 0
    T_cap = 160
    noise = torch.randn(3, 192, T_cap, generator=torch.Generator().manual_seed(2)) * 0.707
    wav, yl = eng.synthesize_durations(m_p, s_p, logw, g, noise.cuda(), T_cap, x_lengths=lens)
    yl = yl.tolist()
    assert yl[1] == 0 and wav[1].abs().sum().item() == 0 and torch.isfinite(wav).all()
    for b in (0, 2):
        L_ = lens[b]
        _, y_b, path = P.durations_to_path(logw[b:b + 1, :, :L_].cpu())
        assert yl[b] == y_b
        one = eng.infer_p2(path.cuda(), m_p[b:b + 1, :, :L_].contiguous(), s_p[b:b + 1, :, :L_].contiguous(), g[b:b + 1],
                           noise[b:b + 1, :, :y_b].contiguous().cuda())
        assert (wav[b:b + 1, :, :y_b * 192] - one).abs().max().item() <= 1e-6
    big = FX.make_p1_inputs(base_cfg, [1030], seed=43)
    with pytest.raises(L.VitsbError):
        eng.infer_p1(big["text"].cuda(), big["emo"].cuda(), big["sid"])
