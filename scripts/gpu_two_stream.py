"""Experiment: two C2 batches in flight on two streams (two handles, two workspaces) vs one stream back to back.
Does overlapping consecutive steps fill the SMs the flow (79 of 148) and the partial last waves leave idle?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import fixtures as FX
from vits_b200.config import HotPathConfig
from vits_b200 import FlowDecoder

cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234)
lens = FX.make_lengths(16, 84, 834, seed=1234)
audio = sum(lens) * cfg.hop_length / cfg.sampling_rate
engs = [FlowDecoder(cfg, sd, device="cuda:0", precision="fp16") for _ in range(2)]
lats = [FX.make_latents(cfg, lens, seed=1234 + i) for i in range(2)]
zs = [l["z_p"].cuda() for l in lats]
gs = [l["g"].cuda() for l in lats]
outs = [torch.empty(16, 1, max(lens) * cfg.hop_length, device="cuda") for _ in range(2)]
streams = [torch.cuda.Stream(), torch.cuda.Stream()]
N = 40


def run(two):
    for i in range(4):                                   # warm-up (graphs)
        with torch.cuda.stream(streams[i % 2 if two else 0]):
            engs[i % 2].synthesize(zs[i % 2], gs[i % 2], lens, out=outs[i % 2])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for st in streams:
        st.wait_event(e0)
    for i in range(N):
        with torch.cuda.stream(streams[i % 2 if two else 0]):
            engs[i % 2].synthesize(zs[i % 2], gs[i % 2], lens, out=outs[i % 2])
    for st in streams:
        torch.cuda.current_stream().wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / N
    return ms


for rep in range(2):
    for two in (False, True):
        ms = run(two)
        print(f"{'two streams' if two else 'one stream '}: {ms:.3f} ms per batch, {audio / ms * 1e3:.0f} audio-s/s", flush=True)
