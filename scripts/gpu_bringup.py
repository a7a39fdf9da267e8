"""GPU bring-up report: selftest matrix of the tcgen05 conv against the fp32
CUDA-core conv (each config in its own process so a trap cannot poison the
rest), then engine parity against the CPU oracle.  Writes gpurun_out/bringup.log."""
import ctypes as C
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = [
    # (prec, Cin, N, k, dil, L, B, gate)
    (0, 64, 64, 1, 1, 256, 1, 0),
    (0, 64, 128, 3, 1, 300, 2, 0),
    (0, 256, 256, 3, 1, 300, 2, 0),
    (0, 256, 256, 11, 5, 700, 3, 1),
    (0, 256, 512, 5, 1, 417, 2, 1),
    (0, 192, 512, 7, 1, 417, 1, 0),
    (0, 512, 2048, 3, 1, 200, 1, 0),
    (0, 256, 96, 1, 1, 417, 2, 0),
    (0, 128, 128, 7, 3, 1000, 2, 1),
    (0, 64, 128, 11, 1, 1000, 2, 0),
    (0, 32, 64, 7, 1, 1000, 2, 0),
    (0, 32, 32, 11, 5, 2000, 2, 1),
    (0, 16, 32, 11, 1, 2000, 2, 0),
    (0, 64, 64, 3, 3, 1000, 1, 1),
    (1, 256, 256, 3, 1, 300, 2, 0),
    (1, 128, 128, 7, 3, 1000, 2, 1),
]


def one(cfg):
    from vits_b200 import lib as L
    lib = L.load_library()
    me, sc = C.c_double(-1), C.c_double(-1)
    rc = lib.vitsb_selftest_conv(0, *cfg, 1234, C.byref(me), C.byref(sc))
    print(json.dumps({"cfg": cfg, "rc": rc, "max_err": me.value, "ref_scale": sc.value}))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        one(tuple(json.loads(sys.argv[2])))
        return
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    log = open(os.path.join(ROOT, "gpurun_out", "bringup.log"), "w")

    def say(*a):
        s = " ".join(str(x) for x in a)
        print(s, flush=True)
        log.write(s + "\n")
        log.flush()

    say("== selftest: tcgen05 conv vs fp32 conv ==")
    for cfg in CONFIGS:
        t = time.time()
        try:
            r = subprocess.run([sys.executable, __file__, "--one", json.dumps(cfg)], capture_output=True, text=True, timeout=120)
            out = (r.stdout.strip().splitlines() or ["<no output>"])[-1]
            say(out, "exit", r.returncode, "%.1fs" % (time.time() - t), (r.stderr.strip()[-300:] if r.returncode else ""))
        except subprocess.TimeoutExpired:
            say(json.dumps({"cfg": cfg}), "TIMEOUT")
    say("== engine parity vs oracle ==")
    for prec in ("fp32", "fp16", "bf16"):
        try:
            r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "gpu_parity.py"), prec], capture_output=True,
                               text=True, timeout=600)
            say(r.stdout.strip()[-3000:], "exit", r.returncode, r.stderr.strip()[-1500:] if r.returncode else "")
        except subprocess.TimeoutExpired:
            say(prec, "TIMEOUT")


if __name__ == "__main__":
    main()
