"""Print the fp16-path error against the fp32-verification build and the golden vector (GPU box).
Usage: python scripts/err_check.py   (VITSB_LIB selects the library)"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import fixtures as FX
from vits_b200.config import HotPathConfig
from oracle.logmel import log_mel_l1
from vits_b200 import FlowDecoder

cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234, weight_norm=False)
e16 = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16")
e32 = FlowDecoder(cfg, sd, device="cuda:0", precision="fp32")
gold = np.load(os.path.join("tests", "golden", "base_p2_T33.npz"))
inp = FX.make_p2_inputs(cfg, 33, seed=1234)
o = e16.infer_p2(*(inp[k].cuda() for k in ("attn", "m_p", "s_p", "g", "noise"))).cpu()
ref = torch.from_numpy(gold["o"])
print("golden T=33 : max-abs %.3e  log-mel L1 %.3e" % ((o - ref).abs().max().item(), log_mel_l1(o[:, 0], ref[:, 0])))
for T in (417, 834, 5000):
    lat = FX.make_latents(cfg, [T], seed=5)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    w16, w32 = e16.synthesize(zp, g).cpu(), e32.synthesize(zp, g).cpu()
    print("T=%d vs fp32 build: max-abs %.3e  log-mel L1 %.3e" % (T, (w16 - w32).abs().max().item(), log_mel_l1(w16[:, 0], w32[:, 0])))
lens = FX.make_lengths(16, 84, 834, seed=1234)
lat = FX.make_latents(cfg, lens, seed=1234)
zp, g = lat["z_p"].cuda(), lat["g"].cuda()
w16, w32 = e16.synthesize(zp, g, lens).cpu(), e32.synthesize(zp, g, lens).cpu()
print("C2 batch vs fp32 build: max-abs %.3e" % (w16 - w32).abs().max().item())
