"""Bucket the SASS of an ncu report into loop regions and print each region's share of executed
warp instructions, issue cycles (FFMA2/FMUL2/FADD2 count 2) and stall samples.
usage: python scripts/ncu_regions.py report.ncu-rep [min_ffma2_run]
Regions: maximal address ranges between backward-branch targets; a region is labelled `scan`
when >= 50 % of its instructions are packed FP32."""
import csv, io, re, subprocess, sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
ins = []
for r in rows:
    try:
        a = int(r["Address"], 16)
    except ValueError:
        continue
    src = r["Source"].strip()
    op = src.split()[1] if src.startswith("@") else src.split()[0]
    ins.append((a, op.split(".")[0], src, int(r["Instructions Executed"]), int(r["# Samples"])))
base = ins[0][0]
tot_i = sum(x[3] for x in ins)
tot_c = sum(x[3] * (2 if x[1] in ("FFMA2", "FMUL2", "FADD2") else 1) for x in ins)
tot_s = sum(x[4] for x in ins)
# split wherever the executed count changes by > 2x relative to the previous instruction (basic-block-ish), then merge by similarity
blocks = []
cur = [ins[0]]
for x in ins[1:]:
    p = cur[-1]
    if (x[3] > 1.5 * p[3] or p[3] > 1.5 * x[3]) and max(x[3], p[3]) > tot_i * 1e-5:
        blocks.append(cur); cur = [x]
    else:
        cur.append(x)
blocks.append(cur)
print(f"total warp inst {tot_i:.3e}, issue cycles {tot_c:.3e}, samples {tot_s}")
print(f"{'addr':>8} {'n':>5} {'exec/inst':>11} {'inst%':>6} {'cyc%':>6} {'smp%':>6} {'pk%':>5}  first")
for b in blocks:
    bi = sum(x[3] for x in b)
    bc = sum(x[3] * (2 if x[1] in ("FFMA2", "FMUL2", "FADD2") else 1) for x in b)
    bs = sum(x[4] for x in b)
    if bc < tot_c * 0.004:
        continue
    pk = sum(1 for x in b if x[1] in ("FFMA2", "FMUL2", "FADD2")) / len(b)
    print(f"{b[0][0]-base:8x} {len(b):5d} {bi/len(b):11.3e} {100*bi/tot_i:6.2f} {100*bc/tot_c:6.2f} {100*bs/tot_s:6.2f} {100*pk:5.0f}  {b[0][2][:60]}")
