"""Single-utterance latency of the pieces of EmoVITS.infer on the library (B = 1, ~5 s of audio)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import fixtures as FX
from oracle import p1_oracle as P
from vits_b200 import FlowDecoder
from vits_b200.config import HotPathConfig
cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234); sd.update(FX.make_p1_state_dict(cfg, seed=4321))
eng = FlowDecoder(cfg, sd, device="cuda:0", precision="fp16")
Tx = 84
inp = FX.make_p1_inputs(cfg, [Tx], seed=99)
text, emo, sid = inp["text"].cuda(), inp["emo"].cuda(), inp["sid"]
m_p, s_p, logw, g = eng.infer_p1(text, emo, sid)
w, y_len, path = P.durations_to_path(logw.cpu())
T_cap = 768
noise = (torch.randn(1, 192, T_cap) * 0.707).cuda()
print("Tx", Tx, "y_len", y_len, "audio_s", y_len * 192 / 16000)

def timeit(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3

print("infer_p1                       : %.3f ms" % timeit(lambda: eng.infer_p1(text, emo, sid)))
print("synthesize_durations (p2)      : %.3f ms" % timeit(lambda: eng.synthesize_durations(m_p, s_p, logw, g, noise, T_cap)))
print("infer (p1 + durations + p2)    : %.3f ms" % timeit(lambda: eng.infer(text, emo, sid, noise, T_cap)))
def host_glue():
    a, b, c, d = eng.infer_p1(text, emo, sid)
    ww, yl, pth = P.durations_to_path(c.cpu())
    return eng.infer_p2(pth.cuda(), a, b, d, noise[:, :, :yl].contiguous())
print("infer_p1 + HOST glue + infer_p2: %.3f ms" % timeit(host_glue))
zp = torch.randn(1, 192, y_len).cuda()
print("synthesize (flow+dec only)     : %.3f ms" % timeit(lambda: eng.synthesize(zp, g)))
