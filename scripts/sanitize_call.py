"""Small ragged synthesis through every kernel family (for compute-sanitizer): fp16 + fp32 engines, pcm16, length_regulate."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import fixtures as FX
from vits_b200.config import HotPathConfig
from vits_b200 import FlowDecoder
cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234)
lens = [37, 5, 21]
lat = FX.make_latents(cfg, lens, seed=3)
z, g = lat["z_p"].cuda(), lat["g"].cuda()
for prec in ("fp16", "fp32"):
    eng = FlowDecoder(cfg, sd, device="cuda:0", precision=prec)
    for _ in range(3):                       # direct, graph capture, graph replay
        w = eng.synthesize(z, g, lens)
    p = eng.synthesize_pcm16(z, g, lens, volume=0.8)
    inp = FX.make_p2_inputs(cfg, 33, seed=1)
    o = eng.infer_p2(*(inp[k].cuda() for k in ("attn", "m_p", "s_p", "g", "noise")))
    torch.cuda.synchronize()
    print(prec, "ok", float(w.abs().max()), int(p.abs().max()), float(o.abs().max()))
