"""Error (vs the CPU oracle, C2 batch, every row) and decoder time for each upsampler split mask."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import fixtures as FX
from oracle import vits_oracle as O
from oracle.logmel import log_mel_l1
from vits_b200 import FlowDecoder
from vits_b200.config import HotPathConfig
cfg = HotPathConfig(); sd = FX.make_state_dict(cfg, seed=1234)
lens = FX.make_lengths(16, 84, 834, seed=1234)
lat = FX.make_latents(cfg, lens, seed=1234)
torch.set_num_threads(os.cpu_count())
with torch.no_grad():
    z_ref, o_ref = O.synthesize_ragged(sd, lat["z_p"], lat["g"], lat["lengths"], cfg)
zp, g = lat["z_p"].cuda(), lat["g"].cuda()
masks = [int(m) for m in sys.argv[1:]] or [15, 7, 3, 1, 0, 12, 9]
for mask in masks:
    eng = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16", options={"ups_split_mask": mask})
    w = eng.synthesize(zp, g, lens).cpu()
    err = max((w[b, :, :L * 192] - o_ref[b, :, :L * 192]).abs().max().item() for b, L in enumerate(lens))
    rms = (w - o_ref).pow(2).mean().sqrt().item()
    mel = max(log_mel_l1(w[b, :, :L * 192], o_ref[b, :, :L * 192]) for b, L in enumerate(lens))
    for _ in range(3):
        eng.synthesize(zp, g, lens)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        eng.synthesize(zp, g, lens)
    b.record(); torch.cuda.synchronize()
    print(f"ups_split_mask={mask:2d}: C2 max-abs {err:.3e} rms {rms:.3e} log-mel L1 {mel:.3e}  step {a.elapsed_time(b) / 20:.3f} ms", flush=True)
    del eng
