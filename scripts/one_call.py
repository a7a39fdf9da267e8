"""One synthesize call of the C2 batch (for ncu captures)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import fixtures as FX
from vits_b200.config import HotPathConfig
from vits_b200 import FlowDecoder
cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234)
eng = FlowDecoder(cfg, sd, device="cuda:0", precision=sys.argv[1] if len(sys.argv) > 1 else "fp16",
                  options={"cuda_graph": 0, **{k: int(v) for k, v in (kv.split("=") for kv in os.environ.get("VITSB_OPTS", "").split(",") if kv)}})
# (cuda_graph 0: plain launches, every kernel visible to ncu; VITSB_OPTS="rb_nslab=16,pdl=0": extra per-handle options for A/B runs)
lens = FX.make_lengths(16, 84, 834, seed=1234)
lat = FX.make_latents(cfg, lens, seed=1234)
z, g = lat["z_p"].cuda(), lat["g"].cuda()
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    w = eng.synthesize(z, g, lens)
torch.cuda.synchronize()
print("ok", float(w.abs().max()))
