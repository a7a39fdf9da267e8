import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import fixtures as FX
from vits_b200 import FlowDecoder
from vits_b200.config import HotPathConfig
cfg = HotPathConfig(); sd = FX.make_state_dict(cfg, seed=1234)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def bench(eng, zp, g, lens, n=30):
    for _ in range(5): eng.synthesize(zp, g, lens)
    torch.cuda.synchronize(); tot = 0.0
    for _ in range(n):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); eng.synthesize(zp, g, lens); b.record(); torch.cuda.synchronize(); tot += a.elapsed_time(b)
    return tot / n
outs = {}
for opts in ({"branch_streams": 0}, {"branch_streams": 1}, {"branch_streams": 0}, {"branch_streams": 1}, {"branch_streams": 1, "cuda_graph": 0}):
    eng = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16", options=opts)
    res = []
    for name, lens in (("C1", [417]), ("C3", [5000]), ("C2", FX.make_lengths(16, 84, 834, seed=1234))):
        lat = FX.make_latents(cfg, lens, seed=3)
        zp, g = lat["z_p"].cuda(), lat["g"].cuda()
        ll = lens if len(lens) > 1 else None
        w = eng.synthesize(zp, g, ll).clone()
        if name in outs: assert torch.equal(outs[name], w), name
        outs[name] = w
        res.append(bench(eng, zp, g, ll))
    print(opts, "C1 %.3f ms  C3 %.3f ms  C2 %.3f ms" % tuple(res), "ws C2 %.2f GB" % (eng.workspace_bytes(16, 812) / 2**30), flush=True)
