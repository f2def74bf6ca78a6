"""ncu source page joined with nvdisasm line info: share of instructions / issue cycles / stall samples per source line.
usage: cuobjdump -xelf all libsrt_cuda.so && nvdisasm --print-line-info srt_backend.sm_100a.cubin > gpurun_out/all_dis.txt
       python scripts/ncu_phases.py report.ncu-rep <mangled kernel name>"""
import csv, io, re, subprocess, sys, collections
rep = sys.argv[1]; kern = sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
dl = open('gpurun_out/all_dis.txt').read().splitlines()
s = next(i for i,l in enumerate(dl) if l.startswith('.text.'+kern))
cur=None; li=[]
for l in dl[s+1:]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur=(m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/', l)
    if m: li.append(cur)
    if l.startswith('.text.') and len(li)>100: break
agg = collections.defaultdict(lambda:[0,0,0])
i=0
for r in rows:
    try: int(r["Address"],16)
    except ValueError: continue
    src = li[i] if i < len(li) else None; i+=1
    ex=int(r["Instructions Executed"]); sm=int(r["# Samples"])
    op=r["Source"].split()[0] if not r["Source"].strip().startswith('@') else r["Source"].split()[1]
    cyc = ex*(2 if op.split('.')[0] in ("FFMA2","FMUL2","FADD2") else 1)
    a=agg[src]; a[0]+=ex; a[1]+=cyc; a[2]+=sm
T=[sum(v[k] for v in agg.values()) for k in range(3)]
byfile=collections.defaultdict(lambda:[0,0,0])
for (src),v in agg.items():
    f = src[0] if src else None
    for k in range(3): byfile[f][k]+=v[k]
for f,v in sorted(byfile.items(), key=lambda x:-x[1][1]): print(f"{str(f):32s} inst {100*v[0]/T[0]:5.1f}% cyc {100*v[1]/T[1]:5.1f}% smp {100*v[2]/T[2]:5.1f}%")
print("--- lines >= 0.7% cycles")
for src,v in sorted(agg.items(), key=lambda x:-x[1][1]):
    if v[1] < 0.007*T[1]: break
    print(f"{str(src):44s} inst {100*v[0]/T[0]:5.1f}% cyc {100*v[1]/T[1]:5.1f}% smp {100*v[2]/T[2]:5.1f}%")
