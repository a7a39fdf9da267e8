"""GPU: vitsb_resample_poly (include/vitsb200.h; SURVEY 8(f) N2, vits_wrap.py:186-195) against scipy.signal.resample_poly
through the C ABI, for the rate pairs the serving path produces (`pitch` -> int(sr / pitch) -> sr, `sampling_rate`),
batched rows, and the fused int16 conversion against the numpy expression of vits_wrap.py:195."""
import numpy as np
import pytest
import scipy.signal as ss
import torch

from oracle import fixtures as FX
from vits_b200.config import HotPathConfig

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from vits_b200 import FlowDecoder
    cfg = HotPathConfig()
    return FlowDecoder(cfg, FX.make_state_dict(cfg, seed=1234, weight_norm=False), device="cuda:0", precision="fp16")


@pytest.mark.parametrize("orig,target", [(16000, 8000), (16000, 24000), (int(16000 / 1.1), 16000), (int(16000 / 0.9), 16000),
                                         (16000, 22050), (8000, 16000)])
def test_resample_matches_scipy(eng, orig, target):
    from vits_b200 import resample as R
    p = R.plan(orig, target)
    rng = np.random.default_rng(orig + target)
    for shape in ((1, 1, 80064), (3, 1, 4801), (2, 7)):
        t = np.arange(shape[-1]) / orig
        x = (0.5 * np.sin(2 * np.pi * 440.0 * t) + 0.1 * rng.standard_normal(shape)).astype(np.float32)
        y = eng.resample(torch.from_numpy(x), orig, target).cpu().numpy()
        ref = ss.resample_poly(x.astype(np.float64), p.up, p.down, axis=-1)
        assert y.shape == ref.shape and y.dtype == np.float32
        assert np.abs(y - ref).max() <= 2e-6                      # float32 rounding of a float64 sum
        pcm = eng.resample(torch.from_numpy(x), orig, target, pcm16=True, volume=0.8).cpu().numpy()
        want = np.clip(y * np.float32(0.8) * np.float32(32767), -32768, 32767).astype(np.int16)      # vits_wrap.py:195 on the fp32 result
        assert pcm.dtype == np.int16 and np.array_equal(pcm, want)


def test_resample_identity_and_after_synthesis(eng):
    cfg = HotPathConfig()
    lat = FX.make_latents(cfg, [40], seed=3)
    wav = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda())
    same = eng.resample(wav, 16000, 16000)
    assert torch.equal(same, wav)
    down = eng.resample(wav, 16000, 8000)
    assert down.shape == (1, 1, wav.shape[-1] // 2)
    ref = ss.resample_poly(wav.cpu().numpy().astype(np.float64), 1, 2, axis=-1)
    assert np.abs(down.cpu().numpy() - ref).max() <= 2e-6


@pytest.mark.parametrize("pitch,rate,tail", [(1.0, None, 0.0), (1.1, None, 0.25), (1.0, 8000, 0.1), (0.9, 24000, 0.0)])
def test_speaking_output_matches_reference_steps(eng, pitch, rate, tail):
    """VITSWrap.speaking's per-segment output steps (vits_wrap.py:186-197) with scipy's resample_poly standing in for
    librosa.resample: same order of operations, same int16 conversion, same tail padding."""
    rng = np.random.default_rng(7)
    sr = 16000
    wav = (0.4 * np.sin(2 * np.pi * 220.0 * np.arange(24000) / sr) + 0.05 * rng.standard_normal(24000)).astype(np.float32)
    volume = 0.9
    got = eng.speaking_output(torch.from_numpy(wav), volume=volume, pitch=pitch, sampling_rate=rate, tail_silence=tail).cpu().numpy()
    from vits_b200 import resample as R
    ref = wav.astype(np.float64)
    if pitch != 1.0:
        p = R.plan(int(sr / pitch), sr)
        ref = ss.resample_poly(ref.astype(np.float32).astype(np.float64), p.up, p.down)
    out_sr = sr if rate is None else rate
    if out_sr != sr:
        p = R.plan(sr, out_sr)
        ref = ss.resample_poly(ref.astype(np.float32).astype(np.float64), p.up, p.down)
    want = np.clip(ref.astype(np.float32) * np.float32(volume) * np.float32(32767), -32768, 32767).astype(np.int16)
    if tail > 0:
        want = np.pad(want, [0, int(tail * out_sr)])
    assert got.dtype == np.int16 and got.shape == want.shape
    assert np.abs(got.astype(np.int32) - want.astype(np.int32)).max() <= 1        # float32 rounding of the resampled value at an integer boundary
