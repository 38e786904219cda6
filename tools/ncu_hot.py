#!/usr/bin/env python3
"""Hot spots of one kernel from an .ncu-rep source page (SASS view): instructions executed per SASS line, grouped
in address order into runs, plus the top lines.  usage: python tools/ncu_hot.py rep kernel_name [topN]"""
import csv, subprocess, sys, io
rep, kn = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kn, "--launch-count", "1"] , capture_output=True, text=True).stdout
lines = raw.splitlines()
# find header
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[1:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"): break
    data.append((r[ix["Source"]].strip(), int(r[ix["Instructions Executed"]]), int(r[ix["# Samples"]]), int(r[ix["Thread Instructions Executed"]])))
tot = sum(d[1] for d in data); tots = sum(d[2] for d in data)
print(f"# {kn}: {len(data)} SASS lines, {tot} warp-instructions, {tots} samples")
# cumulative profile by position: print segments of 25 lines
seg = 25
for i in range(0, len(data), seg):
    chunk = data[i:i + seg]
    w = sum(d[1] for d in chunk); s = sum(d[2] for d in chunk); th = sum(d[3] for d in chunk)
    print(f"lines {i:4d}-{i+len(chunk)-1:4d}: inst {100*w/tot:5.1f}%  samples {100*s/max(tots,1):5.1f}%  avg_threads {th/max(w,1):4.1f}  first: {chunk[0][0][:60]}")
print("# top lines by samples")
for j in sorted(range(len(data)), key=lambda j: -data[j][2])[:top]:
    d = data[j]
    print(f"{j:4d} inst {100*d[1]/tot:4.1f}% samp {100*d[2]/max(tots,1):4.1f}% thr {d[3]/max(d[1],1):4.1f}  {d[0][:90]}")
