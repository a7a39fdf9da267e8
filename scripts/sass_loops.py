"""Instruction mix of every loop of one kernel in the built library: python scripts/sass_loops.py <mangled-name-substring>"""
import re, subprocess, sys
pat = sys.argv[1]
lib = "vits_b200/libvitsb200.so"
names = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = re.findall(r"Function : (\S+)", names)
fn = [f for f in funcs if pat in f][0]
sass = subprocess.run(["cuobjdump", "-sass", "-fun", fn, lib], capture_output=True, text=True).stdout
ins = []
for l in sass.split("\n"):
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
print(fn, len(ins), "instructions")
addr = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", t)
    if not m: continue
    tgt = int(m.group(1), 16)
    if tgt >= a or tgt not in addr: continue
    body = [x for _, x in ins[addr[tgt]:i + 1]]
    if len(body) < 60: continue
    c = lambda k: sum(1 for x in body if k in x)
    print(f"loop {tgt:#x}-{a:#x}: {len(body)} instr | LDTM {c('LDTM')} MUFU {c('MUFU')} STS {c('STS')} LDS {c('LDS')} LDG {c('LDG')} "
          f"FADD {c('FADD')} FMUL {c('FMUL')} FFMA {c('FFMA')} FMNMX {c('FMNMX')} F2FP {c('F2FP')} HADD2 {c('HADD2')} PRMT {c('PRMT')} SEL {c('SEL')} "
          f"IMAD {c('IMAD')} IADD {c('IADD')} LOP {c('LOP3')} ISETP {c('ISETP')} BAR {c('BAR.')} SYNCS {c('SYNCS')} MOV {c('MOV')} SHF {c('SHF')}")
