"""CPU: the oracle restatement against the golden vectors produced by the
reference's own modules (oracle/gen_golden.py), plus an independent numpy
restatement of the conv primitives on small cases."""
import os

import numpy as np
import torch

from oracle import fixtures as FX
from oracle import vits_oracle as O

TOL = 2e-5  # fp32 op-order noise between two torch compositions of the same math


def test_p2_golden(base_cfg, base_sd, golden_dir):
    gold = np.load(os.path.join(golden_dir, "base_p2_T33.npz"))
    inp = FX.make_p2_inputs(base_cfg, 33, seed=1234)
    z_p = O.length_regulate(inp["attn"], inp["m_p"], inp["s_p"], inp["noise"])
    assert np.abs(z_p.numpy() - gold["z_p"]).max() < 1e-6
    z = O.flow_reverse(base_sd, z_p, inp["g"], base_cfg)
    assert np.abs(z.numpy() - gold["z"]).max() < TOL
    assert np.abs(z.numpy() - z_p.numpy()).max() > 0.5      # the flow is not an identity
    taps = {}
    o = O.generator(base_sd, z, inp["g"], base_cfg, taps)
    assert np.abs(o.numpy() - gold["o"]).max() < TOL
    assert o.std() > 0.1                                     # O(1)-scale signal, not std 0.01
    for k in ("conv_pre", "ups.0", "ups.1", "ups.2", "ups.3"):
        assert np.abs(taps[k][0, :8].numpy() - gold["tap_" + k]).max() < TOL * 5, k
    o2 = O.infer_p2(base_sd, inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"], base_cfg)
    assert np.abs(o2.numpy() - gold["o"]).max() < TOL


def test_weight_norm_fold_matches(base_cfg, base_sd):
    sd_wn = FX.make_state_dict(base_cfg, seed=1234, weight_norm=True)
    assert any(k.endswith("weight_g") for k in sd_wn)
    folded = O.fold_weight_norm(sd_wn)
    assert set(folded) == set(base_sd)
    assert len(sd_wn) == 463 and len(folded) == 315
    for k in base_sd:
        assert (folded[k] - base_sd[k]).abs().max() < 1e-6, k


def test_tiny_lengths(base_cfg, base_sd, golden_dir):
    gold = np.load(os.path.join(golden_dir, "base_tiny.npz"))
    for T in (1, 2, 5):
        lat = FX.make_latents(base_cfg, [T], seed=100 + T)
        z = O.flow_reverse(base_sd, lat["z_p"], lat["g"], base_cfg)
        o = O.generator(base_sd, z, lat["g"], base_cfg)
        assert o.shape == (1, 1, T * 192)
        assert np.abs(z.numpy() - gold[f"z_T{T}"]).max() < TOL
        assert np.abs(o.numpy() - gold[f"o_T{T}"]).max() < TOL


def test_batch_padded_and_ragged(base_cfg, base_sd, golden_dir):
    gold = np.load(os.path.join(golden_dir, "base_batch3.npz"))
    lens = [int(x) for x in gold["lengths"]]
    lat = FX.make_latents(base_cfg, lens, seed=77)
    z, o = O.synthesize_padded(base_sd, lat["z_p"], lat["g"], lat["lengths"], base_cfg)
    assert np.abs(z.numpy() - gold["z_padded"]).max() < TOL
    assert np.abs(o.numpy() - gold["o_padded"]).max() < TOL
    zr, orr = O.synthesize_ragged(base_sd, lat["z_p"], lat["g"], lat["lengths"], base_cfg)
    for b, L in enumerate(lens):
        assert np.abs(zr[b:b + 1, :, :L].numpy() - gold[f"z_row{b}"]).max() < TOL
        assert np.abs(orr[b:b + 1, :, :L * 192].numpy() - gold[f"o_row{b}"]).max() < TOL
        assert orr[b, :, L * 192:].abs().sum() == 0
    # the two batch semantics differ only in the tail of shorter rows
    assert (orr[1] - o[1]).abs().max() < TOL
    assert (orr[0, :, :20 * 192 - 3000] - o[0, :, :20 * 192 - 3000]).abs().max() < TOL
    assert (orr[0, :, :20 * 192] - o[0, :, :20 * 192]).abs().max() > 1e-2


def test_adapt_shape_set(adapt_cfg, golden_dir):
    gold = np.load(os.path.join(golden_dir, "adapt_T19.npz"))
    sd = FX.make_state_dict(adapt_cfg, seed=4321)
    lat = FX.make_latents(adapt_cfg, [19], seed=5, noise_scale=0.382)
    z = O.flow_reverse(sd, lat["z_p"], lat["g"], adapt_cfg)
    o = O.generator(sd, z, lat["g"], adapt_cfg)
    assert o.shape == (1, 1, 19 * 96)
    assert np.abs(z.numpy() - gold["z"]).max() < TOL
    assert np.abs(o.numpy() - gold["o"]).max() < TOL


def test_numpy_conv_primitives():
    """Independent explicit-loop numpy conv1d / conv_transpose1d against the
    ATen ops the reference calls (third-party arithmetic)."""
    from oracle import np_ops
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, 6, 23)).astype(np.float32)
    w = rng.standard_normal((4, 6, 5)).astype(np.float32)
    b = rng.standard_normal(4).astype(np.float32)
    for d in (1, 3):
        ref = torch.nn.functional.conv1d(torch.from_numpy(x), torch.from_numpy(w), torch.from_numpy(b),
                                         dilation=d, padding=2 * d).numpy()
        assert np.abs(np_ops.conv1d(x, w, b, dilation=d, padding=2 * d) - ref).max() < 1e-4
    for (k, u) in ((16, 8), (12, 6), (4, 2), (8, 4)):
        wt = rng.standard_normal((6, 3, k)).astype(np.float32)
        bt = rng.standard_normal(3).astype(np.float32)
        ref = torch.nn.functional.conv_transpose1d(torch.from_numpy(x), torch.from_numpy(wt), torch.from_numpy(bt),
                                                   stride=u, padding=(k - u) // 2).numpy()
        got = np_ops.conv_transpose1d(x, wt, bt, stride=u, padding=(k - u) // 2)
        assert got.shape == ref.shape == (2, 3, 23 * u)
        assert np.abs(got - ref).max() < 1e-4


def test_logmel_sane():
    from oracle.logmel import log_mel, mel_basis
    mb = mel_basis(16000, 1024, 80)
    assert mb.shape == (80, 513) and (mb >= 0).all()
    # slaney-normalised triangles: every filter integrates to ~1 bin-width-scaled area
    assert np.all(mb.sum(1) > 0)
    t = torch.arange(16000) / 16000.0
    y = 0.5 * torch.sin(2 * np.pi * 1000 * t).unsqueeze(0)
    lm = log_mel(y)
    assert lm.shape[1] == 80 and lm.shape[2] == 16000 // 192
    peak = lm.mean(-1)[0].argmax().item()
    centre = np.argmax(mb[:, round(1000 / (16000 / 1024))])
    assert abs(peak - centre) <= 1


def test_p1_oracle_matches_reference_golden(base_cfg, golden_dir):
    """oracle/p1_oracle.py against outputs of the reference's own infer_p1 (models.py:558-566) and
    the masked batched head of inference() (models.py:518-522), generated by oracle/gen_golden.py."""
    import numpy as np
    import torch
    from oracle import fixtures as FX
    from oracle import p1_oracle as P
    gold = np.load(os.path.join(golden_dir, "base_p1.npz"))
    sd = FX.make_p1_state_dict(base_cfg, seed=4321)
    with torch.no_grad():
        for Tx, seed in ((60, 99), (1, 3), (5, 4), (131, 7)):
            inp = FX.make_p1_inputs(base_cfg, [Tx], seed=seed)
            m_p, s_p, logw, g = P.infer_p1(sd, inp["text"], inp["emo"], inp["sid"], base_cfg)
            assert np.abs(m_p.numpy() - gold[f"m_p_T{Tx}"]).max() <= 2e-5
            assert np.abs(s_p.numpy() / gold[f"s_p_T{Tx}"] - 1).max() <= 2e-5
            assert np.abs(logw.numpy() - gold[f"logw_T{Tx}"]).max() <= 2e-5
            assert np.array_equal(g.numpy(), gold[f"g_T{Tx}"])
        lens = [int(x) for x in gold["lengths_batch"]]
        inp = FX.make_p1_inputs(base_cfg, lens, seed=11)
        m_p, s_p, logw, _ = P.infer_p1(sd, inp["text"], inp["emo"], inp["sid"], base_cfg, x_lengths=inp["lengths"])
        assert np.abs(m_p.numpy() - gold["m_p_batch"]).max() <= 2e-5
        assert np.abs(torch.log(s_p).numpy() - gold["logs_p_batch"]).max() <= 2e-5
        assert np.abs(logw.numpy() - gold["logw_batch"]).max() <= 2e-5
        # glue: durations -> path equals commons.infer_path semantics (one source position per frame, row sums 1)
        logw60 = torch.from_numpy(gold["logw_T60"])
        w_ceil, y_len, path = P.durations_to_path(logw60)
        assert path.shape == (1, y_len, 60) and torch.all(path.sum(-1) == 1) and torch.equal(path.sum(1).view(-1), w_ceil.view(-1))
