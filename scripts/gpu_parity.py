"""Engine parity against the CPU oracle for one precision; prints per-case errors."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from oracle import fixtures as FX  # noqa: E402
from oracle import vits_oracle as O  # noqa: E402
from oracle.logmel import log_mel_l1  # noqa: E402
from vits_b200 import FlowDecoder  # noqa: E402
from vits_b200.config import HotPathConfig  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
cfg = HotPathConfig()
sd = FX.make_state_dict(cfg, seed=1234)
eng = FlowDecoder(cfg, sd, device="cuda:0", precision=prec)
print(f"[{prec}] flop/frame {eng.flop_per_frame:.0f} launches {eng.launches_per_call(1, 100)}")
torch.set_grad_enabled(False)
for lens in ([33], [1], [5], [130], [20, 33, 7], [417]):
    lat = FX.make_latents(cfg, lens, seed=77)
    z_ref, o_ref = O.synthesize_ragged(sd, lat["z_p"], lat["g"], lat["lengths"], cfg)
    wav, z = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), lens, ragged=True, return_z=True)
    torch.cuda.synchronize()
    wav, z = wav.cpu(), z.cpu()
    zf = eng.flow_reverse(lat["z_p"].cuda(), lat["g"].cuda(), lens).cpu()
    od = eng.decode(z_ref.cuda(), lat["g"].cuda(), lens, ragged=True).cpu()
    lm = log_mel_l1(wav[:, 0], o_ref[:, 0]) if max(lens) * 192 >= 2048 else float("nan")
    print(f"[{prec}] lens {lens}: z err {(z - z_ref).abs().max():.3e} (flow-only {(zf - z_ref).abs().max():.3e}) "
          f"wav err {(wav - o_ref).abs().max():.3e} dec-only {(od - o_ref).abs().max():.3e} logmel L1 {lm:.3e} "
          f"nan {int(torch.isnan(wav).sum())}")
    if len(lens) > 1:
        zp_ref, op_ref = O.synthesize_padded(sd, lat["z_p"], lat["g"], lat["lengths"], cfg)
        wavp = eng.synthesize(lat["z_p"].cuda(), lat["g"].cuda(), lens, ragged=False).cpu()
        print(f"[{prec}] lens {lens} padded semantics: wav err {(wavp - op_ref).abs().max():.3e}")
