"""GPU bring-up of the fused coupling kernel: z of the reverse flow against the CPU oracle and against the
layer-at-a-time plan, plus flow-only timing.  Usage: python scripts/gpu_flow_bringup.py"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import fixtures as FX  # noqa: E402
from oracle import vits_oracle as O  # noqa: E402
from vits_b200 import FlowDecoder  # noqa: E402
from vits_b200.config import HotPathConfig  # noqa: E402

cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234)
fused = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16")
plain = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16", options={"flow_fused": 0})
print("plan fused:", fused.get_option("plan.flow_fused"), "plain:", plain.get_option("plan.flow_fused"), flush=True)
ok = True
for lens in ([33], [112], [113], [417], [97, 40, 64], [300, 5, 224, 1]):
    lat = FX.make_latents(cfg, lens, seed=7 + len(lens))
    with torch.no_grad():
        z_ref, _ = O.synthesize_ragged(sd, lat["z_p"], lat["g"], lat["lengths"], cfg) if len(lens) > 1 else \
            (O.flow_reverse(sd, lat["z_p"], lat["g"], cfg), None)
    zp, g = lat["z_p"].cuda(), lat["g"].cuda()
    zf = fused.flow_reverse(zp, g, lens if len(lens) > 1 else None).cpu()
    zl = plain.flow_reverse(zp, g, lens if len(lens) > 1 else None).cpu()
    ef, el = (zf - z_ref).abs().max().item(), (zl - z_ref).abs().max().item()
    bad = ef > 4e-3 or not torch.isfinite(zf).all()
    ok &= not bad
    print(f"lens={lens}: fused err {ef:.3e}  layerwise err {el:.3e}  {'FAIL' if bad else 'ok'}", flush=True)
    if bad:
        d = (zf - z_ref).abs()
        b, c, t = [int(x) for x in torch.nonzero(d == d.max())[0]]
        print("   worst at b,c,t =", b, c, t, " per-row-of-time max:", d[b].max(0).values[:160:8].tolist())
# flow-only timing on the C2 batch
lens = FX.make_lengths(16, 84, 834, seed=1234)
lat = FX.make_latents(cfg, lens, seed=1234)
zp, g = lat["z_p"].cuda(), lat["g"].cuda()
for name, e in (("fused", fused), ("layerwise", plain)):
    for _ in range(3):
        e.flow_reverse(zp, g, lens)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        e.flow_reverse(zp, g, lens)
    b.record()
    torch.cuda.synchronize()
    print(f"flow_reverse C2 ({name}): {a.elapsed_time(b) / 20:.3f} ms per call (includes pack, cond gemv, unpack)", flush=True)
print("BRINGUP", "OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
