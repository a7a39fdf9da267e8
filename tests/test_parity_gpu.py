"""GPU parity tests: the CUDA path through the C ABI against (a) the committed
golden vectors produced by the reference's own modules and (b) the CPU oracle
on the same seeded inputs.

Tolerances (BASELINE.json north_star): waveform max-abs <= 2e-3 and log-mel L1
<= 1e-2 on the 16-bit tensor-core path, <= 1e-5 on the fp32-verification build.
The 16-bit default is fp16 operands (the reference's production dtype,
infer.py:66): bf16 operand rounding alone gives ~1.2e-2 max-abs on these
unit-scale fixtures (measured with a pure-PyTorch emulation), so the
bf16 variant is held to its own measured floor instead.
"""
import os

import numpy as np
import pytest
import torch

from oracle import fixtures as FX
from oracle import vits_oracle as O
from oracle.logmel import log_mel_l1

pytestmark = pytest.mark.gpu

TOL_WAV = {"fp32": 1e-5, "fp16": 2e-3, "bf16": 8e-3}
# bf16 operands at BASELINE sizes: the emulated floor of bf16 rounding alone is 1.2e-2 at T=417 (DESIGN.md section 3,
# BASELINE.md "precision deviation"); the bf16 variant is held to twice that, and to the north-star log-mel bound
TOL_WAV_BF16_LONG = 2.5e-2
# z is an intermediate, not a north-star quantity: |z| reaches ~5.4 on these fixtures, so 4e-3 is about one fp16 ulp of
# the largest values (the WN in_layers run a single fp16 term; the 3-term split everywhere gave 6e-5)
TOL_Z = {"fp32": 2e-5, "fp16": 4e-3, "bf16": 5e-3}
TOL_MEL = {"fp32": 1e-4, "fp16": 1e-2, "bf16": 1e-2}

_engines = {}


def engine(cfg, sd, prec, key="base"):
    from vits_b200 import FlowDecoder
    k = (key, prec)
    if k not in _engines:
        _engines[k] = FlowDecoder(cfg, sd, device="cuda:0", precision=prec)
    return _engines[k]


@pytest.mark.parametrize("prec", ["fp32", "fp16", "bf16"])
def test_infer_p2_golden(base_cfg, base_sd, golden_dir, prec):
    gold = np.load(os.path.join(golden_dir, "base_p2_T33.npz"))
    inp = FX.make_p2_inputs(base_cfg, 33, seed=1234)
    eng = engine(base_cfg, base_sd, prec)
    o = eng.infer_p2(*(inp[k].cuda() for k in ("attn", "m_p", "s_p", "g", "noise"))).cpu()
    assert o.shape == (1, 1, 33 * 192)
    ref = torch.from_numpy(gold["o"])
    assert (o - ref).abs().max().item() <= TOL_WAV[prec]
    assert log_mel_l1(o[:, 0], ref[:, 0]) <= TOL_MEL[prec]
    z = eng.flow_reverse(torch.from_numpy(gold["z_p"]).cuda(), inp["g"].cuda()).cpu()
    assert (z - torch.from_numpy(gold["z"])).abs().max().item() <= TOL_Z[prec]


@pytest.mark.parametrize("prec", ["fp32", "fp16"])
def test_tiny_lengths_golden(base_cfg, base_sd, golden_dir, prec):
    gold = np.load(os.path.join(golden_dir, "base_tiny.npz"))
    eng = engine(base_cfg, base_sd, prec)
    for T in (1, 2, 5):
        lat = FX.make_latents(base_cfg, [T], seed=100 + T)
        wav, z = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), return_z=True)
        assert (z.cpu().numpy() - gold[f"z_T{T}"]).__abs__().max() <= TOL_Z[prec]
        assert (wav.cpu().numpy() - gold[f"o_T{T}"]).__abs__().max() <= TOL_WAV[prec]


@pytest.mark.parametrize("prec", ["fp32", "fp16"])
def test_batch_golden_ragged_and_padded(base_cfg, base_sd, golden_dir, prec):
    gold = np.load(os.path.join(golden_dir, "base_batch3.npz"))
    lens = [int(x) for x in gold["lengths"]]
    lat = FX.make_latents(base_cfg, lens, seed=77)
    eng = engine(base_cfg, base_sd, prec)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    wav, z = eng.synthesize(zp, g, lens, ragged=True, return_z=True)
    wav, z = wav.cpu().numpy(), z.cpu().numpy()
    for b, L in enumerate(lens):      # row b == the reference's B=1 infer on its own length
        assert np.abs(z[b:b + 1, :, :L] - gold[f"z_row{b}"]).max() <= TOL_Z[prec]
        assert np.abs(wav[b:b + 1, :, :L * 192] - gold[f"o_row{b}"]).max() <= TOL_WAV[prec]
        assert np.abs(wav[b, :, L * 192:]).sum() == 0 and np.abs(z[b, :, L:]).sum() == 0
    # padded semantics == SynthesizerTrn.inference (masked flow, unmasked decoder)
    wavp = eng.synthesize(zp, g, lens, ragged=False).cpu().numpy()
    assert np.abs(wavp - gold["o_padded"]).max() <= TOL_WAV[prec]
    zm = eng.flow_reverse(zp, g, lens).cpu().numpy()
    assert np.abs(zm - gold["z_padded"]).max() <= TOL_Z[prec]
    # standalone decode() of the golden z equals the golden waveform
    od = eng.decode(torch.from_numpy(gold["z_padded"]).cuda(), g, lens, ragged=False).cpu().numpy()
    assert np.abs(od - gold["o_padded"]).max() <= TOL_WAV[prec]


@pytest.mark.parametrize("prec", ["fp32", "fp16"])
def test_adapt_shape_set_golden(adapt_cfg, golden_dir, prec):
    gold = np.load(os.path.join(golden_dir, "adapt_T19.npz"))
    sd = FX.make_state_dict(adapt_cfg, seed=4321)
    eng = engine(adapt_cfg, sd, prec, key="adapt")
    lat = FX.make_latents(adapt_cfg, [19], seed=5, noise_scale=0.382)
    wav, z = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), return_z=True)
    assert wav.shape == (1, 1, 19 * 96)
    assert np.abs(z.cpu().numpy() - gold["z"]).max() <= TOL_Z[prec]
    assert np.abs(wav.cpu().numpy() - gold["o"]).max() <= TOL_WAV[prec]


@pytest.mark.parametrize("gin", [1280, 72])
def test_conditioning_width_other_than_base_vs_oracle(gin):
    """The conditioning GEMV has a register-resident kernel for gin <= 1024 and a generic one above; neither the base
    (1024) nor the adapt (256) shape set reaches the generic kernel or a gin that leaves a K quarter partly empty."""
    from vits_b200 import FlowDecoder
    from vits_b200.config import HotPathConfig
    cfg = HotPathConfig(gin_channels=gin)
    sd = FX.make_state_dict(cfg, seed=77)
    lens = [23, 9, 31, 17, 5, 12, 30, 2, 19, 27, 8, 14, 25, 3, 21, 11, 16, 29, 7]        # 19 rows: a second batch pass of 16
    lat = FX.make_latents(cfg, lens, seed=3)
    eng = FlowDecoder(cfg, sd, device="cuda:0", precision="fp32")
    wav, z = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), lat["lengths"], return_z=True)
    wav, z = wav.cpu(), z.cpu()
    for b, n in enumerate(lens):
        with torch.no_grad():
            z_ref = O.flow_reverse(sd, lat["z_p"][b:b + 1, :, :n], lat["g"][b:b + 1], cfg)
            o_ref = O.generator(sd, z_ref, lat["g"][b:b + 1], cfg)
        assert (z[b, :, :n] - z_ref[0]).abs().max().item() <= TOL_Z["fp32"]
        assert (wav[b, :, :n * cfg.hop_length] - o_ref[0]).abs().max().item() <= TOL_WAV["fp32"]


def test_weight_norm_state_dict_packs_identically(base_cfg, base_sd):
    from vits_b200 import FlowDecoder
    sd_wn = FX.make_state_dict(base_cfg, seed=1234, weight_norm=True)
    a = engine(base_cfg, base_sd, "fp32")
    b = FlowDecoder(base_cfg, sd_wn, device="cuda:0", precision="fp32")
    lat = FX.make_latents(base_cfg, [40], seed=9)
    wa = a.synthesize(lat["z_p"].cuda(), lat["g"].cuda())
    wb = b.synthesize(lat["z_p"].cuda(), lat["g"].cuda())
    assert (wa - wb).abs().max().item() <= 1e-5


@pytest.mark.parametrize("prec", ["fp32", "fp16", "bf16"])
def test_c1_single_utterance_vs_oracle(base_cfg, base_sd, prec):
    """BASELINE config 1: one ~5 s utterance (T=417)."""
    lat = FX.make_latents(base_cfg, [417], seed=1234)
    with torch.no_grad():
        z_ref = O.flow_reverse(base_sd, lat["z_p"], lat["g"], base_cfg)
        o_ref = O.generator(base_sd, z_ref, lat["g"], base_cfg)
    eng = engine(base_cfg, base_sd, prec)
    wav = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda()).cpu()
    assert (wav - o_ref).abs().max().item() <= (TOL_WAV_BF16_LONG if prec == "bf16" else TOL_WAV[prec])
    assert log_mel_l1(wav[:, 0], o_ref[:, 0]) <= TOL_MEL[prec]


_c2_cache = {}


def _c2_oracle(base_cfg, base_sd):
    """CPU oracle on the full C2 batch, row by row (ragged semantics); ~95 s of audio, a few seconds of CPU."""
    if not _c2_cache:
        lens = FX.make_lengths(16, 84, 834, seed=1234)
        lat = FX.make_latents(base_cfg, lens, seed=1234)
        with torch.no_grad():
            z_ref, o_ref = O.synthesize_ragged(base_sd, lat["z_p"], lat["g"], lat["lengths"], base_cfg)
        _c2_cache.update(lens=lens, lat=lat, z=z_ref, o=o_ref)
    return _c2_cache


@pytest.mark.parametrize("prec", ["fp32", "fp16", "bf16"])
def test_c2_batch16_vs_oracle_every_row(base_cfg, base_sd, prec):
    """BASELINE config 2 at full size (16 mixed-length utterances, 1-10 s), compared DIRECTLY with the CPU oracle:
    waveform max-abs and log-mel L1 of every row, z of every row."""
    c = _c2_oracle(base_cfg, base_sd)
    lens, lat = c["lens"], c["lat"]
    eng = engine(base_cfg, base_sd, prec)
    wav, z = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), lens, ragged=True, return_z=True)
    wav, z = wav.cpu(), z.cpu()
    tol_w = TOL_WAV_BF16_LONG if prec == "bf16" else TOL_WAV[prec]
    tol_z = 3e-2 if prec == "bf16" else TOL_Z[prec]
    for b, L in enumerate(lens):
        assert (wav[b, :, :L * 192] - c["o"][b, :, :L * 192]).abs().max().item() <= tol_w, b
        assert log_mel_l1(wav[b, :, :L * 192], c["o"][b, :, :L * 192]) <= TOL_MEL[prec], b
        assert (z[b, :, :L] - c["z"][b, :, :L]).abs().max().item() <= tol_z, b
        assert wav[b, :, L * 192:].abs().sum().item() == 0 and z[b, :, L:].abs().sum().item() == 0


def test_c2_batch16_properties(base_cfg, base_sd):
    """C2 through size-independent properties: determinism, row independence (row b of the ragged batch == the B=1
    call on that row, bit for bit), zero beyond len."""
    lens = FX.make_lengths(16, 84, 834, seed=1234)
    lat = FX.make_latents(base_cfg, lens, seed=1234)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    e16 = engine(base_cfg, base_sd, "fp16")
    w1 = e16.synthesize(zp, g, lens, ragged=True)
    w2 = e16.synthesize(zp, g, lens, ragged=True)
    assert torch.equal(w1, w2)
    for b in (0, 7, 15):
        L = lens[b]
        wb = e16.synthesize(zp[b:b + 1, :, :L].contiguous(), g[b:b + 1])
        assert torch.equal(wb[0, 0], w1[b, 0, :L * 192])
        assert w1[b, 0, L * 192:].abs().sum().item() == 0
    assert not torch.isnan(w1).any()


@pytest.mark.parametrize("prec", ["fp16", "fp32"])
def test_graph_replay_is_bit_identical(base_cfg, base_sd, prec):
    """The engine issues a shape directly on its first call, captures the workspace-only launches into a CUDA
    graph on the second and replays it afterwards: all three must agree bit for bit, new input values and
    fresh input / output buffers must flow through the replay, and a change of lengths must not reuse it."""
    from vits_b200 import FlowDecoder
    eng = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision=prec)      # fresh engine: empty graph cache
    lens = [40, 97, 64]
    lat = FX.make_latents(base_cfg, lens, seed=77)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    w = [eng.synthesize(zp, g, lens).clone() for _ in range(4)]                 # direct, capture, replay, replay
    for x in w[1:]:
        assert torch.equal(w[0], x)
    lat2 = FX.make_latents(base_cfg, lens, seed=78)
    zp2, g2 = lat2["z_p"].cuda(), lat2["g"].cuda()
    keep = torch.empty(1 << 20, device="cuda")                                  # perturb the allocator: new buffer addresses
    w_replay = eng.synthesize(zp2, g2, lens)
    ref = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision=prec).synthesize(zp2, g2, lens)   # first call: direct
    assert torch.equal(w_replay, ref) and not torch.equal(w_replay, w[0])
    # different lengths with the SAME tile counts replay the same graph (the lengths are read from device memory) ...
    lens3 = [40, 90, 64]
    w3 = eng.synthesize(zp, g, lens3)
    ref3 = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision=prec).synthesize(zp, g, lens3)
    assert torch.equal(w3, ref3)
    assert w3[1, 0, 90 * 192:].abs().sum().item() == 0
    # ... and lengths that change a tile count take a different one
    lens4 = [40, 97, 12]
    w4 = [eng.synthesize(zp, g, lens4).clone() for _ in range(3)]
    ref4 = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision=prec).synthesize(zp, g, lens4)
    for x in w4:
        assert torch.equal(x, ref4)
    del keep


def test_reload_invalidates_cached_graphs(base_cfg, base_sd):
    """A second load_state_dict on the same handle must not replay graphs captured against the old weight pools."""
    from vits_b200 import FlowDecoder
    eng = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision="fp16")
    lens = [40, 97, 64]
    lat = FX.make_latents(base_cfg, lens, seed=77)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    w_old = [eng.synthesize(zp, g, lens).clone() for _ in range(3)]          # direct, capture, replay
    sd2 = FX.make_state_dict(base_cfg, seed=4242)
    eng.load_state_dict(sd2)
    w_new = [eng.synthesize(zp, g, lens).clone() for _ in range(3)]
    ref = FlowDecoder(base_cfg, sd2, device="cuda:0", precision="fp16").synthesize(zp, g, lens)
    for x in w_new:
        assert torch.equal(x, ref)
    assert not torch.equal(w_new[0], w_old[0])
    bad = dict(sd2)
    bad.pop("dec.conv_post.weight")
    from vits_b200 import lib as L
    with pytest.raises(L.VitsbError):
        eng.load_state_dict(bad)
    with pytest.raises(L.VitsbError):                   # a failed reload leaves the handle unloaded, not half-built
        eng.synthesize(zp, g, lens)


def test_two_threads_share_one_handle(base_cfg, base_sd):
    """include/vitsb200.h: entry points take the handle's mutex; concurrent calls need distinct workspaces."""
    import ctypes as C
    import threading
    eng = engine(base_cfg, base_sd, "fp16")
    lens = [33, 50]
    lats = [FX.make_latents(base_cfg, lens, seed=900 + i) for i in range(2)]
    refs = [eng.synthesize(l["z_p"].cuda(), l["g"].cuda(), lens).clone() for l in lats]
    outs, errs = [None, None], []

    def work(i):
        try:
            with torch.cuda.device(0):
                st = torch.cuda.Stream()
                zp, g = lats[i]["z_p"].cuda(), lats[i]["g"].cuda()
                ws = torch.empty(eng.workspace_bytes(2, max(lens)), dtype=torch.uint8, device="cuda")
                wav = torch.empty(2, 1, max(lens) * 192, device="cuda")
                torch.cuda.synchronize()
                arr = (C.c_int32 * 2)(*lens)
                for _ in range(20):
                    rc = eng.lib.vitsb_synthesize(eng._h, zp.data_ptr(), g.data_ptr(), arr, 2, max(lens), 1, None, wav.data_ptr(),
                                                  ws.data_ptr(), ws.numel(), C.c_void_p(st.cuda_stream))
                    assert rc == 0
                st.synchronize()
                outs[i] = wav
        except Exception as e:       # noqa
            errs.append(e)

    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    for i in range(2):
        assert torch.equal(outs[i], refs[i])


def test_row_chunking_small_workspace_and_large_batch(base_cfg, base_sd):
    """Memory plan (C4): a workspace smaller than the batch's layout, or more rows than the kernels' 512-row tile
    tables, runs the batch in row chunks - bit-identical to the single-pass result."""
    import ctypes as C
    eng = engine(base_cfg, base_sd, "fp16")
    lens = FX.make_lengths(12, 8, 60, seed=5)
    lat = FX.make_latents(base_cfg, lens, seed=5)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    ref = eng.synthesize(zp, g, lens)
    T = max(lens)
    small = eng.workspace_bytes(5, T)
    assert small < eng.workspace_bytes(12, T)
    ws = torch.empty(small, dtype=torch.uint8, device="cuda")
    wav = torch.empty_like(ref)
    arr = (C.c_int32 * 12)(*lens)
    rc = eng.lib.vitsb_synthesize(eng._h, zp.data_ptr(), g.data_ptr(), arr, 12, T, 1, None, wav.data_ptr(), ws.data_ptr(),
                                  ws.numel(), eng._stream())
    assert rc == 0
    assert torch.equal(wav, ref)
    tiny = torch.empty(eng.workspace_bytes(1, T) - 1024, dtype=torch.uint8, device="cuda")
    rc = eng.lib.vitsb_synthesize(eng._h, zp.data_ptr(), g.data_ptr(), arr, 12, T, 1, None, wav.data_ptr(), tiny.data_ptr(),
                                  tiny.numel(), eng._stream())
    assert rc == 5                                                       # VITSB_ERR_WORKSPACE
    # 520 rows > 512: chunked internally
    lens_b = [3 + (i % 5) for i in range(520)]
    latb = FX.make_latents(base_cfg, lens_b, seed=6)
    wb = eng.synthesize(latb["z_p"].cuda(), latb["g"].cuda(), lens_b)
    for b in (0, 511, 512, 519):
        L = lens_b[b]
        one = eng.synthesize(latb["z_p"][b:b + 1, :, :L].contiguous().cuda(), latb["g"][b:b + 1].cuda())
        assert torch.equal(one[0, 0], wb[b, 0, :L * 192])


@pytest.mark.parametrize("prec", ["fp32", "fp16"])
def test_c3_long_form_vs_oracle(base_cfg, base_sd, prec):
    """BASELINE config 3: 60 s single utterance (T=5000), compared directly with the CPU oracle (waveform max-abs and
    log-mel L1 over the whole minute), plus time-locality."""
    lat = FX.make_latents(base_cfg, [5000], seed=3)
    if "c3" not in _c2_cache:
        with torch.no_grad():
            zr = O.flow_reverse(base_sd, lat["z_p"], lat["g"], base_cfg)
            _c2_cache["c3"] = (zr, O.generator(base_sd, zr, lat["g"], base_cfg))
    z_ref, o_ref = _c2_cache["c3"]
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    eng = engine(base_cfg, base_sd, prec)
    w, z = eng.synthesize(zp, g, return_z=True)
    assert w.shape == (1, 1, 960000)
    assert (w.cpu() - o_ref).abs().max().item() <= TOL_WAV[prec]
    assert log_mel_l1(w.cpu()[:, 0], o_ref[:, 0]) <= TOL_MEL[prec]
    assert (z.cpu() - z_ref).abs().max().item() <= TOL_Z[prec]
    if prec == "fp32":
        # time-locality: the first 2 s do not depend on what follows them (receptive field << 1 s)
        ws = eng.synthesize(zp[:, :, :400].contiguous(), g)
        assert (ws[0, 0, :32000] - w[0, 0, :32000]).abs().max().item() <= 1e-5


def test_c5_distinct_speakers_per_row(base_cfg, base_sd):
    """BASELINE config 5: every row has its own g; rows must not leak conditioning."""
    lens = [50, 50, 50, 50]
    lat = FX.make_latents(base_cfg, lens, seed=11)
    zp = lat["z_p"][:1].repeat(4, 1, 1).cuda()          # same latent, 4 different speakers
    g = lat["g"].cuda()
    e = engine(base_cfg, base_sd, "fp16")
    w = e.synthesize(zp, g, lens)
    for b in range(4):
        wb = e.synthesize(zp[b:b + 1], g[b:b + 1])
        assert torch.equal(wb[0], w[b])
    assert (w[0] - w[1]).abs().max().item() > 1e-2


def test_error_paths(base_cfg, base_sd):
    from vits_b200 import lib as L
    e = engine(base_cfg, base_sd, "fp16")
    lat = FX.make_latents(base_cfg, [10], seed=1)
    with pytest.raises(ValueError):
        e.synthesize(lat["z_p"][:, :100].cuda(), lat["g"].cuda())
    with pytest.raises(L.VitsbError):
        e.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), [11])       # len > T_max
    bad = dict(base_sd)
    bad.pop("dec.conv_post.weight")
    from vits_b200 import FlowDecoder
    with pytest.raises(L.VitsbError):
        FlowDecoder(base_cfg, bad, device="cuda:0")


def test_host_pipeline_matches_blocking_calls(base_cfg, base_sd):
    """Serving loop (H2D / synthesis / D2H on three streams, depth 2) returns exactly what the blocking
    host call returns, request by request, including when a slot is reused."""
    eng = engine(base_cfg, base_sd, "fp16")
    lens = [50, 33, 64, 20]
    B, T = len(lens), max(lens)
    pipe = eng.host_pipeline(B, T)
    reqs = []
    for i in range(5):
        lat = FX.make_latents(base_cfg, lens, seed=500 + i)
        reqs.append((lat["z_p"].pin_memory(), lat["g"].pin_memory(),
                     torch.empty(B, 1, T * 192, dtype=torch.float32).pin_memory()))
    tickets = [pipe.submit(z, g, lens, out) for z, g, out in reqs]
    for t in tickets:
        pipe.wait(t)
    for z, g, out in reqs:
        ref = eng.synthesize_host(z, g, lens)
        assert torch.equal(out, ref)


@pytest.mark.parametrize("opts", [{"flow_fused": 0}, {"flow_fused": 0, "flow_in_split": 1}, {"rb_fused": 0}, {"tc_pair": 0},
                                  {"cuda_graph": 0, "pdl": 0}, {"branch_streams": 0}])
def test_alternative_launch_plans_agree(base_cfg, base_sd, opts):
    """The launch-plan options (layer-at-a-time flow instead of the fused coupling kernel, 3-term split in_layers,
    layer-at-a-time ResBlocks, no CTA pairs, no graphs + PDL) stay within the same tolerance of the CPU oracle and of
    the default plan on a ragged batch."""
    from vits_b200 import FlowDecoder
    lens = [97, 40, 64]
    lat = FX.make_latents(base_cfg, lens, seed=321)
    with torch.no_grad():
        z_ref, o_ref = O.synthesize_ragged(base_sd, lat["z_p"], lat["g"], torch.tensor(lens), base_cfg)
    alt = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision="fp16", options=opts)
    for k, v in opts.items():
        assert alt.get_option(k) == v
    deflt = engine(base_cfg, base_sd, "fp16")
    assert deflt.get_option("plan.flow_fused") == 1 and deflt.get_option("plan.rb_fused") == 1
    if "flow_fused" in opts:
        assert alt.get_option("plan.flow_fused") == 0
    w_alt, z_alt = alt.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), lens, ragged=True, return_z=True)
    w_alt2 = alt.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), lens, ragged=True)
    w_def, z_def = deflt.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), lens, ragged=True, return_z=True)
    w_alt, w_def = w_alt.cpu(), w_def.cpu()
    assert torch.equal(w_alt, w_alt2.cpu())
    assert (w_alt - o_ref).abs().max().item() <= TOL_WAV["fp16"]
    assert (w_def - o_ref).abs().max().item() <= TOL_WAV["fp16"]
    assert (w_alt - w_def).abs().max().item() <= TOL_WAV["fp16"]
    assert (z_alt.cpu() - z_ref).abs().max().item() <= TOL_Z["fp16"] and (z_def.cpu() - z_ref).abs().max().item() <= TOL_Z["fp16"]
    for b, L in enumerate(lens):
        assert w_alt[b, 0, L * 192:].abs().sum().item() == 0
    with pytest.raises(Exception):
        alt.set_option("no_such_option", 1)


def test_length_regulate_matches_reference_matmuls(base_cfg, base_sd):
    """the head of infer_p2 (models.py:570-572).  On a 0/1 duration path the gather kernel is
    bit-identical to the reference's two fp32 matmuls + noise FMA; on an arbitrary attn it is the same weighted sum."""
    eng = engine(base_cfg, base_sd, "fp32")
    for T, seed in ((33, 1234), (417, 7), (1, 3)):
        inp = FX.make_p2_inputs(base_cfg, T, seed=seed)
        ref = O.length_regulate(inp["attn"], inp["m_p"], inp["s_p"], inp["noise"])
        z = eng.length_regulate(inp["attn"].cuda(), inp["m_p"].cuda(), inp["s_p"].cuda(), inp["noise"].cuda()).cpu()
        assert torch.equal(z, ref)
    g = torch.Generator().manual_seed(5)
    B, Ty, Tx, C = 3, 70, 45, base_cfg.inter_channels
    attn = torch.rand(B, Ty, Tx, generator=g) * (torch.rand(B, Ty, Tx, generator=g) < 0.2)
    m_p, s_p, noise = (torch.randn(B, C, n, generator=g) for n in (Tx, Tx, Ty))
    ref = O.length_regulate(attn, m_p, s_p, noise)
    z = eng.length_regulate(attn.cuda(), m_p.cuda(), s_p.cuda(), noise.cuda()).cpu()
    assert (z - ref).abs().max().item() <= 1e-5
    with pytest.raises(ValueError):
        eng.length_regulate(attn.cuda(), m_p[:, :10].cuda(), s_p[:, :10].cuda(), noise.cuda())


def test_c4_batch64_rows_independent(base_cfg, base_sd):
    """BASELINE config 4 (batch 64-512 utterances, sharded by utterance): a 64-row ragged call (short rows so that
    the test stays fast) equals the per-row B=1 calls bit for bit, stays within tolerance of the fp32 build and
    stores zeros beyond every row's length; larger batches are the same launch plan with more tiles."""
    lens = FX.make_lengths(64, 8, 90, seed=64)
    lat = FX.make_latents(base_cfg, lens, seed=64)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    e16, e32 = engine(base_cfg, base_sd, "fp16"), engine(base_cfg, base_sd, "fp32")
    w16 = e16.synthesize(zp, g, lens, ragged=True)
    w32 = e32.synthesize(zp, g, lens, ragged=True)
    assert (w16 - w32).abs().max().item() <= TOL_WAV["fp16"]
    for b in (0, 17, 40, 63):
        L = lens[b]
        wb = e16.synthesize(zp[b:b + 1, :, :L].contiguous(), g[b:b + 1])
        assert torch.equal(wb[0, 0], w16[b, 0, :L * 192])
        assert w16[b, 0, L * 192:].abs().sum().item() == 0


@pytest.mark.parametrize("prec", ["fp16", "fp32"])
def test_pcm16_output_matches_reference_postprocessing(base_cfg, base_sd, prec):
    """int16 output written by the last kernel == np.clip(wav * volume * 32767, -32768, 32767)
    .astype(np.int16) (vits_wrap.py:195) applied to the same call's float waveform, bit for bit."""
    eng = engine(base_cfg, base_sd, prec)
    lens = [50, 21, 64]
    lat = FX.make_latents(base_cfg, lens, seed=909)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    wav = eng.synthesize(zp, g, lens).cpu().numpy()
    for volume in (1.0, 0.7, 3.0):                      # 3.0 clips
        pcm = eng.synthesize_pcm16(zp, g, lens, volume=volume).cpu().numpy()
        ref = np.clip(wav * volume * 32767, -32768, 32767).astype(np.int16)
        assert pcm.dtype == np.int16 and pcm.shape == ref.shape
        assert np.array_equal(pcm, ref)
    assert np.abs(pcm).max() == 32767


def test_pcm16_written_straight_into_pinned_host_memory(base_cfg, base_sd):
    """SURVEY 8(f) N2, "write to pinned host memory": the output pointer of the C ABI may be page-locked host memory
    (device-accessible under unified addressing); the last kernel then stores the int16 samples over PCIe itself and the
    request needs no device buffer and no separate D2H copy.  Same samples as the device-buffer call."""
    eng = engine(base_cfg, base_sd, "fp16")
    lens = [50, 21, 64]
    lat = FX.make_latents(base_cfg, lens, seed=909)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    ref = eng.synthesize_pcm16(zp, g, lens, volume=0.7).cpu()
    host = torch.zeros(ref.shape, dtype=torch.int16).pin_memory()
    out = eng.synthesize_pcm16(zp, g, lens, volume=0.7, out=host)
    torch.cuda.synchronize()
    assert out.data_ptr() == host.data_ptr() and not host.is_cuda
    assert torch.equal(host, ref)
    wav_host = torch.zeros(ref.shape, dtype=torch.float32).pin_memory()
    eng.synthesize(zp, g, lens, out=wav_host)
    torch.cuda.synchronize()
    assert torch.equal(wav_host, eng.synthesize(zp, g, lens).cpu())


def test_synthesize_segments_equals_sequential_calls(base_cfg, base_sd):
    """the serving loop's sequential per-segment calls (vits_wrap.py:177-183) as ragged batches -
    same waveforms bit for bit, original order, whatever the batch split."""
    eng = engine(base_cfg, base_sd, "fp16")
    lens = [33, 120, 7, 64, 90, 15, 1]
    segs = []
    for i, T in enumerate(lens):
        lat = FX.make_latents(base_cfg, [T], seed=700 + i)
        segs.append((lat["z_p"].cuda(), lat["g"].cuda()))
    ref = [eng.synthesize(z, g) for z, g in segs]
    for kw in ({}, {"max_batch": 3}, {"max_batch_frames": 200}):
        got = eng.synthesize_segments(segs, **kw)
        assert len(got) == len(ref)
        for a, b in zip(got, ref):
            assert a.shape == b.shape and torch.equal(a, b)
    with pytest.raises(ValueError):
        eng.synthesize_segments([(segs[0][0][:, :10], segs[0][1])])


def test_dynamic_batcher_across_threads(base_cfg, base_sd):
    """requests submitted from several threads are grouped into ragged batches by one dispatcher and every
    caller gets exactly its B = 1 result."""
    import threading
    from vits_b200.serving import DynamicBatcher
    eng = engine(base_cfg, base_sd, "fp16")
    lens = [33, 120, 7, 64, 90, 15, 1, 50, 77, 21, 40, 9]
    lats = [FX.make_latents(base_cfg, [T], seed=800 + i) for i, T in enumerate(lens)]
    refs = [eng.synthesize(l["z_p"].cuda(), l["g"].cuda()).clone() for l in lats]
    got = [None] * len(lens)
    with DynamicBatcher(eng, max_batch=5, max_wait_ms=50.0) as bt:
        def worker(idx):
            for i in idx:
                got[i] = bt.submit(lats[i]["z_p"], lats[i]["g"]).result(timeout=60)
        th = [threading.Thread(target=worker, args=(range(k, len(lens), 4),)) for k in range(4)]
        [t.start() for t in th]
        [t.join() for t in th]
        assert sum(bt.batches) == len(lens) and max(bt.batches) <= 5 and max(bt.batches) > 1      # really batched
        with pytest.raises(ValueError):
            bt.submit(lats[0]["z_p"][:, :10], lats[0]["g"])
    for a, b in zip(got, refs):
        assert a.shape == b.shape and torch.equal(a, b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_engines_on_two_devices_in_one_process(base_cfg, base_sd):
    """The dynamic shared-memory opt-in is per device: a second engine on another GPU of the same process must work."""
    from vits_b200 import FlowDecoder
    lens = [50, 33]
    lat = FX.make_latents(base_cfg, lens, seed=5)
    e0 = engine(base_cfg, base_sd, "fp16")
    w0 = e0.synthesize(lat["z_p"].cuda(0), lat["g"].cuda(0), lens)
    e1 = FlowDecoder(base_cfg, base_sd, device="cuda:1", precision="fp16")
    w1 = e1.synthesize(lat["z_p"].cuda(1), lat["g"].cuda(1), lens)
    assert w1.device.index == 1 and torch.equal(w0.cpu(), w1.cpu())
    z0 = e0.flow_reverse(lat["z_p"].cuda(0), lat["g"].cuda(0), lens)
    rc_dev = torch.empty_like(z0, device="cuda:1")
    import ctypes as C
    arr = (C.c_int32 * 2)(*lens)
    with torch.cuda.device(1):
        zp1, g1 = lat["z_p"].cuda(1).contiguous(), lat["g"].cuda(1).contiguous()
        rc = e1.lib.vitsb_flow_reverse_noalloc(e1._h, zp1.data_ptr(), g1.data_ptr(), arr, 2, max(lens), rc_dev.data_ptr(), e1._stream())
        torch.cuda.synchronize()
    assert rc == 0 and torch.equal(rc_dev.cpu(), z0.cpu())          # the static scratch array's address is per device


@pytest.mark.parametrize("lens", [[100], [336], [113, 37, 230], [500] * 3 + [17], [417] * 41])
def test_coupling_pairs_match_single_cta(base_cfg, base_sd, lens):
    """The coupling kernel in CTA-pair mode (cta_group::2: one MMA over both CTAs' tiles, half of every weight box per CTA)
    against the same kernel without pairs: one tile with a dead partner, exactly three tiles, odd / even tile counts across
    batch rows, and more pairs than SM pairs (41 x 4 tiles = 82 pairs on 74: two tiles per pair for some).  The per-row
    accumulation order is the same in both modes, so z is bit-identical; both stay within tolerance of the CPU oracle."""
    from vits_b200 import FlowDecoder
    lat = FX.make_latents(base_cfg, lens, seed=99)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    pair = engine(base_cfg, base_sd, "fp16")
    single = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision="fp16", options={"tc_pair": 0})
    assert pair.get_option("tc_pair") == 1 and single.get_option("tc_pair") == 0
    z1 = pair.flow_reverse(zp, g, lens)
    z2 = pair.flow_reverse(zp, g, lens)
    z0 = single.flow_reverse(zp, g, lens)
    torch.cuda.synchronize()
    assert torch.equal(z1, z2)
    assert torch.equal(z1, z0)
    if len(lens) <= 4:
        with torch.no_grad():
            for b, L in enumerate(lens):                        # per-utterance flow of the oracle (models.py:573)
                z_ref = O.flow_reverse(base_sd, lat["z_p"][b:b + 1, :, :L], lat["g"][b:b + 1], base_cfg)
                assert (z1[b:b + 1, :, :L].cpu() - z_ref).abs().max().item() <= TOL_Z["fp16"]
    for b, L in enumerate(lens):
        assert z1[b, :, L:].abs().sum().item() == 0
