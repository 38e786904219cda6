#!/usr/bin/env python3
"""Per-kernel-name aggregate of an `ncu --set full` capture (run here, no GPU needed) as JSON: launches, summed
duration, DRAM bytes per step (read + write), duration-weighted issue-slot / pipe utilisation.  bench.py reads the
committed result (profiles/ncu_latest.json) for roofline.traffic.
usage: python tools/ncu_to_json.py gpurun_out/prof.ncu-rep frames_per_step > profiles/ncu_latest.json"""
import csv, io, json, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}


def val(r, name):
    v = float(r[ix[name]].replace(",", ""))
    return v * SCALE.get(units[ix[name]], 1)


agg = {}
for r in rows[2:]:
    name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").split("<")[0].strip()
    a = agg.setdefault(name, {"launches": 0, "time_us": 0.0, "dram_bytes": 0.0, "_issue": 0.0, "_alu": 0.0, "_lsu": 0.0,
                              "_dram": 0.0, "warp_inst": 0.0, "bank_conflicts": 0.0, "_bar": 0.0, "_lsb": 0.0, "_occ": 0.0})
    t = val(r, "gpu__time_duration.sum")
    a["launches"] += 1; a["time_us"] += t
    a["dram_bytes"] += val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
    a["_issue"] += t * val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
    a["_alu"] += t * val(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active")
    a["_lsu"] += t * val(r, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active")
    a["_dram"] += t * val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")
    a["warp_inst"] += val(r, "smsp__inst_executed.sum")
    a["bank_conflicts"] += val(r, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
    a["_bar"] += t * val(r, "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio")
    a["_lsb"] += t * val(r, "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio")
    a["_occ"] += t * val(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
out = {"source": sys.argv[1].split("/")[-1], "frames_per_step": int(sys.argv[2]) if len(sys.argv) > 2 else None, "kernels": {}}
for k, a in agg.items():
    t = a["time_us"]
    out["kernels"][k] = {"launches": a["launches"], "time_us": round(t, 1), "dram_bytes": int(a["dram_bytes"]),
                         "issue_active_pct": round(a["_issue"] / t, 1), "pipe_alu_pct": round(a["_alu"] / t, 1),
                         "pipe_lsu_pct": round(a["_lsu"] / t, 1), "dram_throughput_pct": round(a["_dram"] / t, 1),
                         "warp_inst": int(a["warp_inst"]), "smem_bank_conflicts": int(a["bank_conflicts"]),
                         "stall_barrier_per_issue": round(a["_bar"] / t, 2), "stall_long_scoreboard_per_issue": round(a["_lsb"] / t, 2),
                         "warps_active_pct": round(a["_occ"] / t, 1)}
print(json.dumps(out, indent=1))
