import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import fixtures as FX
from vits_b200 import FlowDecoder
from vits_b200.config import HotPathConfig
cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234); sd.update(FX.make_p1_state_dict(cfg, seed=4321))
eng = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16")
inp = FX.make_p1_inputs(cfg, [84], seed=99)
text, emo, sid = inp["text"].cuda(), inp["emo"].cuda(), inp["sid"]
for _ in range(3):
    out = eng.infer_p1(text, emo, sid)
torch.cuda.synchronize()
print("ok", float(out[2].mean()))
