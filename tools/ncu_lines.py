#!/usr/bin/env python3
"""Per-source-line instruction / stall-sample counts of one kernel: the SASS page of an `ncu --set full` capture joined
(by instruction order) with the line table nvdisasm prints for the object file the kernel came from.
usage: ncu_lines.py capture.ncu-rep kernel_regex object.o [min_pct] [launch_index]"""
import csv, io, re, subprocess, sys, tempfile, os, glob
rep, kern, obj = sys.argv[1], sys.argv[2], sys.argv[3]
minpct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
skip = sys.argv[5] if len(sys.argv) > 5 else "0"
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
kname = rows[0][1]
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hdr = rows[hi[0]]; ix = {n: i for i, n in enumerate(hdr)}
body = rows[hi[0] + 1: hi[1] if len(hi) > 1 else None]
mangled = None
with tempfile.TemporaryDirectory() as td:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=td, capture_output=True)
    cubin = glob.glob(td + "/*.cubin")[0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
# split per function, pick the one whose demangled name matches
funcs, cur, line = {}, None, None
for l in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        cur = m.group(1); funcs[cur] = []; line = None; continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        line = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m and cur:
        funcs[cur].append((line, m.group(2).strip()))
base = re.sub(r"\(.*", "", kname).split("::")[-1].split("<")[0]
body = [r for r in body if len(r) > 5 and r[ix["Source"]].strip()]
cands = [f for f in funcs if base in f and abs(len(funcs[f]) - len(body)) <= 1]
if not cands:
    print("no function of", len(body), "instructions matching", base, {f: len(v) for f, v in funcs.items() if base in f}); sys.exit(1)
fn = funcs[cands[0]]
ci, si = ix["Instructions Executed"], ix["# Samples"]
tot = sum(int(r[ci]) for r in body); stot = sum(int(r[si]) for r in body)
agg = {}
for (ln, _), r in zip(fn, body):
    a = agg.setdefault(ln, [0, 0]); a[0] += int(r[ci]); a[1] += int(r[si])
print(f"{kname[:100]}\ntotal warp instructions {tot}, stall samples {stot}")
src = {}
for ln, (c, s) in sorted(agg.items(), key=lambda kv: (kv[0] or ("", 0))):
    if 100.0 * c / tot >= minpct or 100.0 * s / max(stot, 1) >= minpct:
        text = ""
        if ln:
            p = [q for q in glob.glob(os.path.dirname(os.path.abspath(obj)) + "/../**/" + ln[0], recursive=True)]
            if p:
                src.setdefault(p[0], open(p[0]).read().splitlines())
                text = src[p[0]][ln[1] - 1].strip()[:100] if ln[1] - 1 < len(src[p[0]]) else ""
        print(f"{(ln[0] + ':' + str(ln[1])) if ln else '?':>22} inst {100.0*c/tot:5.1f}%  stall {100.0*s/max(stot,1):5.1f}%  {text}")
