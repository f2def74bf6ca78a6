"""ncu source page joined with nvdisasm line info: stall reasons per instruction (pass the reasons to list as extra arguments).
usage: cuobjdump -xelf all libsrt_cuda.so && nvdisasm --print-line-info srt_backend.sm_100a.cubin > gpurun_out/all_dis.txt
       python scripts/ncu_stalls.py report.ncu-rep <mangled kernel name> [stall_long_sb ...]"""
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
    if m: cur=(m.group(1).split('/')[-1].replace('srt_','').replace('.cuh','').replace('.cu',''), int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/', l)
    if m: li.append(cur)
    if l.startswith('.text.') and len(li)>100: break
keys = ["stall_long_sb","stall_no_inst","stall_wait","stall_short_sb","stall_membar","stall_lg","stall_branch_resolving","stall_math","stall_mio","stall_sleep","stall_not_selected","stall_selected","stall_dispatch"]
tot = collections.Counter()
per = collections.defaultdict(collections.Counter)
i=0
for r in rows:
    try: int(r["Address"],16)
    except ValueError: continue
    src = li[i] if i < len(li) else None; i+=1
    for k in keys:
        v = int(r[k] or 0); tot[k]+=v; per[k][(src, r["Source"].strip()[:40])]+=v
T = sum(tot.values())
print("total samples", T)
for k,v in tot.most_common(): print(f"{k:26s} {v:8d} {100*v/T:5.1f}%")
for k in sys.argv[3:]:
    print("==", k)
    for (src, ins), v in per[k].most_common(14): print(f"   {v:7d} {100*v/tot[k]:5.1f}%  {src}  {ins}")
