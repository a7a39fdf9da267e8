#!/bin/bash
# per-kernel A/B of differently built libraries with ncu launch durations (each kernel timed alone at boost clocks: far
# less run-to-run noise than the bench's family table).  LIBS="a.so b.so" K=regex bash scripts/gpu_ab_ncu.sh
mkdir -p gpurun_out
LIBS=${LIBS:-"libvitsb200_base.so libvitsb200.so"}
# an entry may carry per-handle options for scripts/one_call.py: libvitsb200.so@rb_nslab=16,pdl=0
for ent in $LIBS; do
  lib=${ent%%@*}; opts=""; [ "$ent" != "$lib" ] && opts=${ent#*@}
  VITSB_OPTS=$opts VITSB_LIB=$PWD/vits_b200/$lib timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"${K:-rb_fused|conv_tc|flow_fused}" \
    --csv --log-file gpurun_out/abn_$ent.csv python scripts/one_call.py fp16 ${CALLS:-3} > gpurun_out/abn_$ent.log 2>&1 || tail -3 gpurun_out/abn_$ent.log
done
python - <<PY
import csv, collections
libs = "$LIBS".split()
tab = {}
for lib in libs:
    rows = [r for r in csv.reader(l for l in open(f"gpurun_out/abn_{lib}.csv") if l.startswith('"'))]
    h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    per = collections.defaultdict(list)
    for r in rows[1:]:
        per[r[ki].split("(")[0].replace(", 1>", ">").replace(", 0>", ">")[-40:]].append(float(r[vi].replace(",", "")))
    tab[lib] = per
names = list(tab[libs[0]].keys())
tot = {l: 0.0 for l in libs}
for n in names:
    line = f"{n:<42}"
    for l in libs:
        v = tab[l].get(n, [])
        ncall = ${CALLS:-3}
        v = v[len(v) // ncall:]                      # drop the first call (cold)
        m = sum(v) / max(1, len(v)) / 1000.0         # us per launch
        tot[l] += sum(v) / 1000.0 / (ncall - 1)
        line += f" {m:9.2f} x{len(v) // (ncall - 1):<3}"
    print(line)
print("sum us per call:", {l: round(t, 1) for l, t in tot.items()})
PY
