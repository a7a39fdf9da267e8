"""Experiment: slab-variant conv (VITSB_TC_MODE=1,2,3) selftests, then bench per working mode."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = [(0, 64, 64, 1, 1, 256, 1, 0), (0, 64, 128, 3, 1, 300, 2, 0), (0, 256, 256, 3, 1, 300, 2, 0), (0, 256, 256, 11, 5, 700, 3, 1),
       (0, 128, 128, 7, 3, 1000, 2, 1), (0, 32, 32, 11, 5, 2000, 2, 1), (0, 16, 32, 11, 1, 2000, 2, 0), (0, 192, 512, 7, 1, 417, 1, 0)]
modes = [int(m) for m in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 2, 3]
good = []
for m in modes:
    env = dict(os.environ, VITSB_TC_MODE=str(m))
    ok = True
    for c in CFG:
        try:
            r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "gpu_bringup.py"), "--one", json.dumps(c)],
                               capture_output=True, text=True, timeout=120, env=env)
            line = (r.stdout.strip().splitlines() or ["<none>"])[-1]
            print("mode", m, line, "exit", r.returncode, r.stderr.strip()[-200:] if r.returncode else "", flush=True)
            try:
                d = json.loads(line)
                if not (d["rc"] == 0 and 0 <= d["max_err"] < 2e-5):
                    ok = False
            except Exception:
                ok = False
        except subprocess.TimeoutExpired:
            print("mode", m, c, "TIMEOUT", flush=True)
            ok = False
    if ok:
        good.append(m)
print("working modes:", good, flush=True)
for m in good:
    env = dict(os.environ, VITSB_TC_MODE=str(m))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "10", "--no-cpu-baseline", "--profile-out",
                        os.path.join(ROOT, "gpurun_out", f"profile_mode{m}.csv")], capture_output=True, text=True, env=env, timeout=600)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        print("mode", m, "audio-s/s", d["value"], "ms/step", d["ms_per_step"], "tflops", d["roofline"]["achieved"], flush=True)
    except Exception:
        print("mode", m, "bench failed", r.stdout[-500:], r.stderr[-500:], flush=True)
    print(open(os.path.join(ROOT, "gpurun_out", f"profile_mode{m}.csv")).read() if os.path.exists(os.path.join(ROOT, "gpurun_out", f"profile_mode{m}.csv")) else "")
