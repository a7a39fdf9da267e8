"""clock64 timeline of the fused coupling kernel (debug library): VITSB_LIB=vits_b200/libvitsb200_dbg.so
VITSB_FF_TRACE=<launches> python scripts/gpu_flow_trace.py [c2|one]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import fixtures as FX  # noqa: E402
from vits_b200 import FlowDecoder  # noqa: E402
from vits_b200.config import HotPathConfig  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "c2"
cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234)
eng = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16", options={"cuda_graph": 0})
lens = FX.make_lengths(16, 84, 834, seed=1234) if mode == "c2" else [417]
lat = FX.make_latents(cfg, lens, seed=1234)
z = eng.flow_reverse(lat["z_p"].cuda(), lat["g"].cuda(), lens)
torch.cuda.synchronize()
print("done", mode, float(z.abs().max()))
