#!/usr/bin/env python3
"""Per-launch table from an `ncu --metrics ... --csv --log-file launches.csv` pass.  usage: launch_table.py launches.csv"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; ik = h.index("Kernel Name"); im = h.index("Metric Name"); iv = h.index("Metric Value"); ii = h.index("ID")
d = {}
for r in rows[1:]:
    d.setdefault((int(r[ii]), r[ik].split('(')[0]), {})[r[im]] = float(r[iv].replace(',', ''))
tot = sum(v.get("gpu__time_duration.sum", 0) for v in d.values())
print(f"{'id':>3} {'kernel':42s} {'us':>9} {'share':>6} {'rd MB':>8} {'wr MB':>8} {'Mwarp-inst':>10}")
for (i, k), v in sorted(d.items()):
    t = v.get("gpu__time_duration.sum", 0)
    print(f"{i:3d} {k:42s} {t/1e3:9.1f} {t/tot:6.3f} {v.get('dram__bytes_read.sum',0)/1e6:8.1f} {v.get('dram__bytes_write.sum',0)/1e6:8.1f} {v.get('smsp__inst_executed.sum',0)/1e6:10.2f}")
print(f"total {tot/1e3:.1f} us")
