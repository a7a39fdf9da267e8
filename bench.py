#!/usr/bin/env python
"""bench.py — generated audio-seconds per second of the flow-reverse + Generator hot path.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--precision fp16|bf16|fp32]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A step is one `synthesize` pass (reverse flow + decoder)
over one batch of synthetic latents.  Workload at N=1 is BASELINE.json
configs[1]: 16 mixed-length utterances (1-10 s, masked/ragged), emotional-vits
default config, random-init weights at fan-in scale, explicitly injected noise.
For N>1 the global batch is 16*N utterances sharded by utterance, one process
per GPU, the same length set on every rank (so per-GPU work is identical at every
N), no collective on the synthesis path (weak scaling).

Prints ONE JSON line (rank 0).  `--impl reference` times the CPU oracle port of
the reference path on the host cores instead (see oracle/vits_oracle.py).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SR, HOP = 16000, 192
FRAME_SEC = HOP / SR
PER_GPU_BATCH = 16
LEN_LO, LEN_HI = 84, 834          # 1-10 s in 12 ms frames
C2_WORKLOAD = "C2: emotional-vits base config, 16 mixed-length utterances per GPU (1-10 s, ragged/masked), flow-reverse + Generator"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-side-legs", action="store_true", help="skip the sustained / varying-lengths / C1 / C3 / C4 side records")
    ap.add_argument("--option", action="append", default=[], metavar="KEY=INT",
                    help="library option for an A/B run (vitsb_set_option), e.g. --option tc_pair=0; recorded in config.options")
    ap.add_argument("--profile-out", default=None, help="write the per-launch-family table to this file")
    ap.add_argument("--batch-per-gpu", type=int, default=PER_GPU_BATCH,
                    help="utterances per GPU and step (default 16 = BASELINE config C2; 64-512 are C4 sweep points, not the headline)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "tflops_burst": d["bf16_tflops"],
                "hbm_gbs": d["hbm_gbs"], "source": "measured"}
    return {"tflops_sustained": 1400.0, "tflops_burst": 1590.0, "hbm_gbs": 6650.0, "source": "fallback"}


def global_lengths(n_gpus: int):
    """Weak scaling with IDENTICAL per-GPU work: every rank gets the same 16-length set of config C2 (seed 1234), so
    the per-N efficiency the driver computes isolates scaling from workload (latents and speakers differ per rank)."""
    from oracle import fixtures as FX
    base = FX.make_lengths(PER_GPU_BATCH, LEN_LO, LEN_HI, seed=1234)
    return base * n_gpus


def lpt_shard(lengths, n):
    """Longest-processing-time-first assignment of utterances to ranks."""
    bins = [[] for _ in range(n)]
    load = [0] * n
    for i in sorted(range(len(lengths)), key=lambda i: -lengths[i]):
        r = min(range(n), key=lambda r: (load[r], len(bins[r])))
        bins[r].append(i)
        load[r] += lengths[i]
    return bins


def reduce_job(audio_s, frames, ms, world, device):
    """Whole-job aggregation for N ranks: audio and frames SUM over ranks, time MAX over ranks."""
    import torch
    import torch.distributed as dist
    tot = torch.tensor([audio_s, frames], dtype=torch.float64, device=device)
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return tot[0].item(), int(round(tot[1].item())), t.item()


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def mark(self):
        return time.time()

    def stop(self, t0=None, t1=None):
        if self.proc:
            self.proc.terminate()
        rows = [r for ts, r in self.rows if (t0 is None or ts >= t0) and (t1 is None or ts <= t1 + 0.15)]
        if not rows:
            rows = [r for _, r in self.rows[-3:]]
        sm = [float(r[0]) for r in rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------
# CPU leg: the oracle port of the reference's own PyTorch path, on host cores
# ---------------------------------------------------------------------------
def cpu_oracle_run(lengths, budget_s: float, min_utts: int = 1):
    """Time per-utterance `flow.infer` + `dec` (models.py:573-574) of the oracle
    port on as many utterances of the batch as fit in `budget_s`."""
    import torch
    from oracle import fixtures as FX
    from oracle import vits_oracle as O
    from vits_b200.config import HotPathConfig
    cfg = HotPathConfig()
    sd = FX.make_state_dict(cfg, seed=1234)
    torch.set_num_threads(os.cpu_count() or 1)
    lat = FX.make_latents(cfg, lengths, seed=1234)
    order = sorted(range(len(lengths)), key=lambda i: lengths[i])
    done_frames, used, n = 0, 0.0, 0
    with torch.no_grad():
        O.generator(sd, O.flow_reverse(sd, lat["z_p"][:1, :, :32], lat["g"][:1], cfg), lat["g"][:1], cfg)  # warm-up
        for i in order:
            L = lengths[i]
            t = time.perf_counter()
            z = O.flow_reverse(sd, lat["z_p"][i:i + 1, :, :L], lat["g"][i:i + 1], cfg)
            O.generator(sd, z, lat["g"][i:i + 1], cfg)
            used += time.perf_counter() - t
            done_frames += L
            n += 1
            if used >= budget_s and n >= min_utts:
                break
    return {"audio_s": done_frames * FRAME_SEC, "seconds": used, "utts": n, "frames": done_frames,
            "threads": torch.get_num_threads()}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lengths = global_lengths(1)
    total_steps = args.steps + args.warmup
    budget = max(1.0, min(15.0, 150.0 / max(1, total_steps)))
    rec = []
    for s in range(total_steps):
        r = cpu_oracle_run(lengths, budget)
        if s >= args.warmup:
            rec.append(r)
    audio = sum(r["audio_s"] for r in rec)
    secs = sum(r["seconds"] for r in rec)
    val = audio / secs
    sample = f"{rec[0]['utts']} shortest utterance(s) of the 16-utterance C2 batch per step ({rec[0]['audio_s']:.1f} s audio), " \
             f"per-utterance fp32 torch-CPU oracle port of flow.infer+dec"
    out = {"impl": "reference", "metric": "audio-sec/sec (RTF^-1)", "value": val, "unit": "audio-s/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(1, len(rec)),
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": C2_WORKLOAD, "global_batch": 16, "valid_audio_s_per_step": rec[0]["audio_s"],
                      "parallelism": "host cores of rank 0", "precision": "fp32",
                      "sample": "all utterances of the batch that fit the per-step CPU budget, per utterance at B = 1"},
           "cpu_baseline": {"value": val, "unit": "audio-s/s", "cores": rec[0]["threads"], "kind": "port", "sample": sample},
           "e2e": {"value": val, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


# ---------------------------------------------------------------------------
# GPU leg
# ---------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from oracle import fixtures as FX            # input/weight generator only (never on the timed path)
    from vits_b200 import FlowDecoder
    from vits_b200.config import HotPathConfig

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    cfg = HotPathConfig()
    sd = FX.make_state_dict(cfg, seed=1234)
    sd.update(FX.make_p1_state_dict(cfg, seed=4321))       # text encoder + duration predictor, for the end-to-end infer() side leg
    lib_options = {kv.split("=", 1)[0]: int(kv.split("=", 1)[1]) for kv in args.option}
    eng = FlowDecoder(cfg, sd, device=dev, precision=args.precision, options=lib_options or None)
    all_len = global_lengths(world)
    mine = list(range(rank * PER_GPU_BATCH, (rank + 1) * PER_GPU_BATCH))          # identical length set on every rank
    lens = [all_len[i] for i in mine]
    lat = FX.make_latents(cfg, lens, seed=1234 + rank)                           # distinct latents and speakers (g) per rank and row
    B, T = len(lens), max(lens)
    up = eng.total_upsample
    z_host = lat["z_p"].pin_memory()
    g_host = lat["g"].pin_memory()
    z_dev, g_dev = z_host.to(dev), g_host.to(dev)
    wav_dev = torch.empty(B, 1, T * up, dtype=torch.float32, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)       # > 126 MB L2
    audio_s = sum(lens) * FRAME_SEC
    frames = sum(lens)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        eng.synthesize(z_dev, g_dev, lens, ragged=True, out=wav_dev)

    def timed(fn, steps, warmup, flush_l2=True, sample=True):
        sampler = ClockSampler(local)
        if rank == 0 and sample:
            sampler.start()                                  # started before warm-up: nvidia-smi needs ~0.5 s to attach
        for _ in range(warmup):
            fn()
        barrier()
        evs = []
        t0 = sampler.mark()
        for _ in range(steps):
            if flush_l2:
                flush.fill_(1)                               # evict L2 between timed iterations (untimed)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            evs.append((a, b))
        barrier()
        t1 = sampler.mark()
        clocks = sampler.stop(t0, t1) if (rank == 0 and sample) else None
        ms = sum(a.elapsed_time(b) for a, b in evs)
        return reduce_job(0.0, 0, ms, world, dev)[2], clocks

    def timed_pipeline(steps, warmup, pcm16):
        """End-to-end through the serving API (FlowDecoder.host_pipeline, depth 2): every step copies its latents
        from pinned host memory and its result back to pinned host memory inside the timed region; the copies of
        neighbouring steps overlap the synthesis.  pcm16: the result is the int16 PCM the serving path returns
        (vits_wrap.py:195), written by the last kernel.  Timed as ONE region with CUDA events on the compute stream
        (first H2D ordered after the start event, end event ordered after the last D2H).  No L2 flush here: the
        inputs come from the host every step and a step's activation traffic (GBs) is far beyond the 126 MB L2."""
        pipe = eng.host_pipeline(B, T, pcm16=pcm16)
        dt = torch.int16 if pcm16 else torch.float32
        outs = [torch.empty(B, 1, T * up, dtype=dt).pin_memory() for _ in range(2)]
        prev = None
        for i in range(warmup):
            t = pipe.submit(z_host, g_host, lens, outs[i % 2])
            if prev is not None:
                pipe.wait(prev)
            prev = t
        pipe.wait(prev)
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        prev = None
        for i in range(steps):
            t = pipe.submit(z_host, g_host, lens, outs[i % 2], after=ev0 if i == 0 else None)
            if prev is not None:
                pipe.wait(prev)
            prev = t
        torch.cuda.current_stream().wait_event(pipe.join_event(prev))
        ev1.record()
        pipe.wait(prev)
        barrier()
        return reduce_job(0.0, 0, ev0.elapsed_time(ev1), world, dev)[2], outs[0].numel() * outs[0].element_size()

    W = max(3, args.warmup)
    ms_res, clocks = timed(step_resident, args.steps, W)
    ms_e2e, d2h_bytes = timed_pipeline(args.steps, W, pcm16=True)
    ms_e2e_f32, d2h_f32 = timed_pipeline(args.steps, W, pcm16=False)
    audio_all, frames_all, _ = reduce_job(audio_s, frames, 0.0, world, dev)

    # ---- side legs (rank-local, N = 1 only): sustained clocks, varying lengths, the other BASELINE configs ----
    side = {}
    if world == 1 and not args.no_side_legs:
        # (a) sustained: >= 3 s of back-to-back steps, no flush, clocks sampled throughout (the power-capped regime)
        n_sus = max(args.steps, int(3000.0 / max(ms_res / args.steps, 0.1)))
        ms_sus, clk_sus = timed(step_resident, n_sus, 3, flush_l2=False)
        side["sustained"] = {"steps": n_sus, "seconds": ms_sus * 1e-3, "ms_per_step": ms_sus / n_sus,
                             "value": audio_s * n_sus / (ms_sus * 1e-3), "clocks": clk_sus}
        # (b) a fresh length draw every step (same distribution): a server never sees the same 16 lengths twice, so this
        # leg cannot be served by replaying one CUDA graph per length vector
        draws = [FX.make_lengths(B, LEN_LO, LEN_HI, seed=5000 + i) for i in range(args.steps + W)]
        Tv = max(max(d) for d in draws)
        latv = FX.make_latents(cfg, [Tv] * B, seed=77)
        zv, gv = latv["z_p"].to(dev), latv["g"].to(dev)
        wv = torch.empty(B, 1, Tv * up, dtype=torch.float32, device=dev)
        it = iter(draws)
        fr = [0]

        def step_vary():
            d = next(it)
            fr[0] += sum(d)
            eng.synthesize(zv, gv, d, ragged=True, out=wv)
        for _ in range(W):
            step_vary()
        fr[0] = 0
        ms_v, _ = timed(step_vary, args.steps, 0, sample=False)
        side["vary_lengths"] = {"value": fr[0] * FRAME_SEC / (ms_v * 1e-3), "ms_per_step": ms_v / args.steps,
                                "note": "fresh U{84..834} length draw per step, T_max fixed; lengths are device data, graphs are keyed on tile counts"}
        # (c) BASELINE configs C1 (5 s, B=1), C3 (60 s, B=1): single-stream latency; C4: 64-512 utterances on this GPU
        def latency(Tn, reps):
            l1 = FX.make_latents(cfg, [Tn], seed=3)
            z1, g1 = l1["z_p"].to(dev), l1["g"].to(dev)
            o1 = torch.empty(1, 1, Tn * up, dtype=torch.float32, device=dev)
            f = lambda: eng.synthesize(z1, g1, None, ragged=True, out=o1)
            ms, _ = timed(f, reps, 3, sample=False)
            return {"frames": Tn, "audio_s": Tn * FRAME_SEC, "latency_ms": ms / reps, "value": Tn * FRAME_SEC * reps / (ms * 1e-3)}
        # (d) the whole of EmoVITS.infer for one utterance (infer.py:136-184): text -> infer_p1 -> device-side durations ->
        # flow -> decoder -> int16 PCM, no host synchronisation inside
        p1in = FX.make_p1_inputs(cfg, [84], seed=99)
        tx, em, sid1 = p1in["text"].to(dev), p1in["emo"].to(dev), p1in["sid"]
        Tc = 768
        nz = (torch.randn(1, cfg.inter_channels, Tc, generator=torch.Generator().manual_seed(1)) * 0.707).to(dev)
        _, yl1 = eng.infer(tx, em, sid1, nz, Tc, pcm16=True)
        yl1 = int(yl1.item())
        ms_p1, _ = timed(lambda: eng.infer_p1(tx, em, sid1), 20, 3, sample=False)
        ms_inf, _ = timed(lambda: eng.infer(tx, em, sid1, nz, Tc, pcm16=True), 20, 3, sample=False)
        side["infer_text_to_pcm_B1"] = {"phonemes": 84, "frames": yl1, "audio_s": yl1 * FRAME_SEC, "infer_p1_ms": ms_p1 / 20,
                                        "latency_ms": ms_inf / 20, "value": yl1 * FRAME_SEC * 20 / (ms_inf * 1e-3),
                                        "note": "vitsb_infer_p1 (fp32) + vitsb_synthesize_durations; device-timed, L2 flushed between calls"}
        side["C1_latency"] = latency(417, 20)
        side["C3_latency"] = latency(5000, 10)
        c4 = {}
        for Bn in (64, 128, 256, 512):
            ln = sorted(FX.make_lengths(Bn, LEN_LO, LEN_HI, seed=1234 + Bn), reverse=True)   # length-sorted rows: chunks of similar length
            Tn = max(ln)
            zb = torch.randn(Bn, cfg.inter_channels, Tn, device=dev)
            for r_, L_ in enumerate(ln):
                zb[r_, :, L_:] = 0
            gb = torch.randn(Bn, cfg.gin_channels, device=dev)
            ob = torch.empty(Bn, 1, Tn * up, dtype=torch.int16, device=dev)
            f = lambda: eng.synthesize_pcm16(zb, gb, ln, ragged=True, out=ob)
            ms, _ = timed(f, 3, 2, sample=False)
            c4[str(Bn)] = {"audio_s": sum(ln) * FRAME_SEC, "ms_per_step": ms / 3, "value": sum(ln) * FRAME_SEC * 3 / (ms * 1e-3),
                           "workspace_gb": eng.workspace_bytes(Bn, Tn) / 2 ** 30}
            del zb, gb, ob
            eng._ws = None
            torch.cuda.empty_cache()
        side["C4_batch_sweep_1gpu"] = c4

    # per-launch-family device time (separate, untimed pass): which share of the step the tensor-core launches are
    eng.profile(True)
    for _ in range(3):
        step_resident()
    table = eng.profile_table()
    eng.profile(False)
    conv_rows = [r for r in table if r[3] > 0]
    conv_ms = sum(r[1] for r in conv_rows) / 3
    all_ms = sum(r[1] for r in table) / 3
    conv_flops = sum(r[3] for r in conv_rows) / 3
    conv_launches = sum(r[2] for r in conv_rows) // 3
    gen_rows = [r for r in conv_rows if not r[0].startswith("flow")]
    gen_ms, gen_flops = sum(r[1] for r in gen_rows) / 3, sum(r[3] for r in gen_rows) / 3
    pk = peaks()
    share = conv_ms / all_ms if all_ms else 0.0
    step_ms = ms_res / args.steps
    # tensor-core launches' achieved rate INSIDE the timed pass: their algorithmic FLOP over their share of the
    # driver-timed step (share from the per-launch event pass above; recomputable from ms_per_step)
    achieved_tf = conv_flops / (step_ms * share * 1e-3) / 1e12 if share > 0 else 0.0
    gen_share = gen_ms / all_ms if all_ms else 0.0
    gen_tf = gen_flops / (step_ms * gen_share * 1e-3) / 1e12 if gen_share > 0 else 0.0

    if rank == 0:
        value = audio_all * args.steps / (ms_res * 1e-3)
        e2e = audio_all * args.steps / (ms_e2e * 1e-3)
        traffic = None
        tf = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tf):
            traffic = json.load(open(tf)).get("dram_bytes_per_launch")
        out = {
            "metric": "audio-sec/sec (RTF^-1)", "value": value, "unit": "audio-s/s", "n_gpus": world,
            "steps": args.steps, "warmup": W, "ms_per_step": step_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp16": "f16", "bf16": "bf16", "fp32": "f32"}[args.precision] + " operands, f32 accumulate",
            "data": "synthetic",
            "config": {"workload": C2_WORKLOAD if PER_GPU_BATCH == 16 else
                                   f"C4 sweep point: emotional-vits base config, {PER_GPU_BATCH} mixed-length utterances per GPU (1-10 s, ragged/masked), flow-reverse + Generator",
                       "note": "at N>1 this is C5-shaped: every row of every rank has its own speaker embedding g" if world > 1 else "",
                       "global_batch": PER_GPU_BATCH * world, "valid_audio_s_per_step": audio_all,
                       "parallelism": f"utterance-sharded replicas x{world}, the same 16-length set on every rank, no collective",
                       "l2": "256 MiB flush between timed steps; per-step activation traffic >> 126 MB L2",
                       "precision": args.precision, **({"options": lib_options} if lib_options else {})},
            "e2e": {"value": e2e, "unit": "audio-s/s", "h2d_bytes_per_step": int(z_host.numel() * 4 + g_host.numel() * 4 + 4 * B) * world,
                    "d2h_bytes_per_step": int(d2h_bytes) * world, "ms_per_step": ms_e2e / args.steps,
                    "api": "FlowDecoder.host_pipeline(pcm16=True) depth 2: per-step H2D of the latents and D2H of the int16 PCM the serving "
                           "path returns (vits_wrap.py:195, written by the last kernel) overlap the neighbouring steps' synthesis",
                    "float32_result": {"value": audio_all * args.steps / (ms_e2e_f32 * 1e-3), "d2h_bytes_per_step": int(d2h_f32) * world}},
            "gpu_launches": eng.launches_per_call(B, T) * args.steps * world,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "rb_fused_kernel + flow_fused_kernel + conv_tc_kernel (all tcgen05 launches of a step; rb_fused is ~80 % of it)",
                         "achieved": achieved_tf, "peak": pk["tflops_burst"], "unit": "TFLOP/s",
                         "frac": achieved_tf / pk["tflops_burst"],
                         "peak_source": pk["source"] + " (burst cuBLAS bf16: the timed region is < 1 s at full clocks)",
                         "frac_of_sustained_peak": achieved_tf / pk["tflops_sustained"], "sustained_peak": pk["tflops_sustained"],
                         "formula": "achieved = algorithmic FLOP of the tensor-core launches per step / (ms_per_step * kernel_share_of_step)",
                         "traffic": traffic, "launches_per_step": conv_launches,
                         "avg_launch_ms": step_ms * share / max(1, conv_launches),
                         "flop_per_launch": conv_flops / max(1, conv_launches),
                         "kernel_share_of_step": share,
                         "generator_convs": {"tflops": gen_tf, "frac_burst": gen_tf / pk["tflops_burst"],
                                             "frac_sustained": gen_tf / pk["tflops_sustained"], "share_of_step": gen_share},
                         "step_level_tflops": eng.flop_per_frame * frames_all * args.steps / (ms_res * 1e-3) / 1e12},
        }
        if "sustained" in side:
            sus = side["sustained"]
            sus_tf = eng.flop_per_frame * frames * 1e-12 / (sus["ms_per_step"] * 1e-3)
            sus["step_level_tflops"] = sus_tf
            sus["frac_of_sustained_peak"] = sus_tf / pk["tflops_sustained"]
        if side:
            out["configs"] = side
        if args.profile_out:
            with open(args.profile_out, "w") as f:
                f.write(f"# per-launch-family device time, mean of 3 steps, B={B} T_max={T} precision={args.precision}\n")
                f.write("family,ms_per_step,launches_per_step,algorithmic_gflop_per_step,tflops\n")
                for name, ms, n, fl in sorted(table, key=lambda r: -r[1]):
                    f.write(f"{name},{ms / 3:.4f},{n // 3},{fl / 3 / 1e9:.2f},{(fl / (ms * 1e-3) / 1e12) if ms > 0 else 0:.1f}\n")
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_oracle_run(lens, 12.0)
            out["cpu_baseline"] = {"value": r["audio_s"] / r["seconds"], "unit": "audio-s/s", "cores": r["threads"], "kind": "port",
                                   "sample": f"{r['utts']} shortest utterance(s) of the same batch ({r['audio_s']:.1f} s audio, "
                                             f"{r['seconds']:.1f} s CPU), per-utterance fp32 torch-CPU oracle port"}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    global PER_GPU_BATCH
    args = parse()
    PER_GPU_BATCH = args.batch_per_gpu
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
