#!/usr/bin/env python3
"""bench.py -- stereo frames/s of the ORB stereo front-end (extract L + extract R + ComputeStereoMatches)
at 752x480, 1000 keypoints, 8 levels, scale 1.2, FAST 20/7 (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--frames F] [--impl reference]

A step = one pass of the whole hot path over one batch of F synthetic stereo frames.
  value   : device-resident inputs, CUDA-event timed on the launching stream, max over ranks.
  e2e     : the same through the C ABI with HOST (pinned) buffers: H2D of the images and D2H of
            keypoints/descriptors/uRight/depth inside the timed region (gforb_stereo_frames_batch).
  roofline: dominant kernel (by measured per-stage CUDA-event time) vs the measured HBM copy peak.
  cpu_baseline: the CPU oracle (a port of the reference algorithm; the reference itself cannot be built here)
            on the box's host cores, bounded sample, rank 0 at N=1 only.
Multi-GPU: frames are independent -> each rank runs its own F frames (weak scaling), no collective on the
data path; torch.distributed only provides the barrier and the max-over-ranks of the device time.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

W, H, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH = 752, 480, 1000, 1.2, 8, 20, 7
BF = 47.90639384423901
B = BF / 435.2046959714599
METRIC = "stereo frames/s ORB extract+match @752x480x1000kp"
UNIT = "stereo_frames/s"
BASE_FRAMES = 16           # distinct generated stereo frames; the batch cycles them with shifts + fresh noise


def make_frames(n, seed0=0):
    """n distinct synthetic rectified stereo pairs (uint8 [n,H,W] x2)."""
    from synth import synth_stereo
    nb = min(n, BASE_FRAMES)
    try:
        from concurrent.futures import ProcessPoolExecutor
        with ProcessPoolExecutor(max_workers=min(nb, os.cpu_count() or 1)) as ex:
            base = list(ex.map(synth_stereo, range(seed0, seed0 + nb)))
    except Exception:
        base = [synth_stereo(seed0 + i) for i in range(nb)]
    left = np.empty((n, H, W), np.uint8)
    right = np.empty((n, H, W), np.uint8)
    rng = np.random.default_rng(99 + seed0)
    for i in range(n):
        l, r = base[i % nb]
        shift = 16 * (i // nb)
        if shift:
            l, r = np.roll(l, shift, axis=1), np.roll(r, shift, axis=1)
            nz = rng.integers(-1, 2, (2, H, W), dtype=np.int16)
            l = np.clip(l.astype(np.int16) + nz[0], 0, 255).astype(np.uint8)
            r = np.clip(r.astype(np.int16) + nz[1], 0, 255).astype(np.uint8)
        left[i], right[i] = l, r
    return left, right


def algorithmic_bytes(level_sizes, n_cand, n_kp):
    """Per EYE image, SURVEY.md 8(d) (FAST survivors are 4 B packed here, not 8)."""
    P = sum(w * h for w, h in level_sizes)
    Ppad = sum((w + 38) * (h + 38) for w, h in level_sizes)
    return {
        "pyramid": W * H + Ppad,
        "fast": P + 4 * n_cand,
        "quadtree": 4 * n_cand + 4 * n_kp,
        "blur": 2 * P,
        "describe": (961 + 1369 + 60) * n_kp,
    }


class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_baseline(left, right, threads, target_s=15.0):
    """Oracle port on the host cores, 'all cores' shape: every worker runs whole stereo frames."""
    from oracle import pyoracle as O
    n = min(len(left), 8)
    imgs = np.empty((2 * n, H, W), np.uint8)
    imgs[0::2], imgs[1::2] = left[:n], right[:n]
    secs, _ = O.bench_stereo(imgs, max(threads, 2), threads, 0, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, BF, B)
    per_frame = secs / max(threads, 2) * threads            # thread-seconds per frame
    nframes = int(max(threads * 2, min(4000, target_s * threads / max(per_frame, 1e-4))))
    secs, chk = O.bench_stereo(imgs, nframes, threads, 0, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, BF, B)
    return nframes / secs, nframes, secs, chk


def cpu_reference_shaped(left, right, nframes=48):
    """The reference's own shape (src/Frame.cc:268-282): one frame at a time, one thread per eye, serial match."""
    from oracle import pyoracle as O
    n = min(len(left), 8)
    imgs = np.empty((2 * n, H, W), np.uint8)
    imgs[0::2], imgs[1::2] = left[:n], right[:n]
    secs, _ = O.bench_stereo(imgs, nframes, 2, 1, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, BF, B)
    return nframes / secs


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm's CPU path (oracle port; OpenCV C++ is not in the image so the
    reference's own sources cannot be compiled) on all host threads, same config/metric."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    left, right = make_frames(8)
    t0 = time.time()
    vals, total_frames = [], 0
    # bounded sample per step so the whole run ends within minutes
    per_step_s = max(2.0, min(15.0, 150.0 / max(args.steps + args.warmup, 1)))
    for i in range(args.warmup + args.steps):
        fps, nframes, secs, _ = cpu_baseline(left, right, threads, target_s=per_step_s)
        if i >= args.warmup:
            vals.append((nframes, secs)); total_frames += nframes
    tot_s = sum(s for _, s in vals)
    value = total_frames / tot_s
    sample = f"{total_frames} synthetic 752x480 stereo frames over {args.steps} steps, {threads} threads"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * tot_s / max(args.steps, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[1]: 752x480 stereo pair, 1000 kp, 8 levels, 1.2, FAST 20/7, classic stereo match",
                       "frames_per_step": total_frames // max(args.steps, 1)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.time() - t0}
    emit(line)


def bind_to_gpu_numa_node(index):
    """Pin this rank (and the pinned host buffers it is about to allocate: first touch) to the NUMA node its GPU hangs
    off, so that the end-to-end arm's H2D / D2H traffic does not cross the socket interconnect.  Best effort: returns
    the node or None when the topology cannot be read (single-socket boxes report -1)."""
    try:
        q = subprocess.run(["nvidia-smi", "-i", str(index), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                           capture_output=True, text=True, timeout=20).stdout.strip().lower()
        bus = q[-12:] if len(q) >= 12 else q                       # sysfs uses a 4-digit domain: 0000:1b:00.0
        node = int(Path(f"/sys/bus/pci/devices/{bus}/numa_node").read_text())
        if node < 0:
            return None
        cpus = set()
        for part in Path(f"/sys/devices/system/node/node{node}/cpulist").read_text().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        try:                                                        # prefer that node's memory for what we allocate next
            import ctypes
            numa = ctypes.CDLL("libnuma.so.1")
            if numa.numa_available() >= 0:
                numa.numa_set_preferred(node)
        except OSError:
            pass
        return node
    except Exception:
        return None


_RESULT_FD = None


def quiet_stdout():
    """Library chatter (NCCL prints its version banner on stdout when NCCL_DEBUG is set in the environment) must not
    share stdout with the ONE JSON line: fd 1 is pointed at stderr for the run, the result goes to the saved fd."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_RESULT_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frames", type=int, default=256, help="stereo frames per step per GPU")
    ap.add_argument("--impl", default="gforb", choices=["gforb", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--chunk", type=int, default=0, help="frames per pipeline chunk of the end-to-end arm")
    args = ap.parse_args()
    quiet_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == "gforb" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return 0

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; gforb has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    gf = importlib.import_module("gf-orb-slam2_b200")
    from sharding import aggregate_throughput

    F = args.frames
    left, right = make_frames(F, seed0=1000 * rank)
    full_affinity = os.sched_getaffinity(0)
    numa_node = bind_to_gpu_numa_node(local_rank)     # before the pinned buffers are allocated
    sb = gf.StereoBatch(W, H, F, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)
    stream = torch.cuda.Stream()        # non-default: handle 0 (legacy default) would mean "library's own stream"
    torch.cuda.set_stream(stream)
    sb.set_stream(stream.cuda_stream)
    out = sb.alloc_outputs(F)                       # pinned
    pin_l = gf.PinnedArray((F, H, W), np.uint8); pin_r = gf.PinnedArray((F, H, W), np.uint8)
    pin_l.array[:] = left; pin_r.array[:] = right

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident arm ----------------
    sb.upload(pin_l.array, pin_r.array)
    torch.cuda.synchronize()
    for _ in range(args.warmup):
        sb.run(F, BF, B)
    sb.sync()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        sb.run(F, BF, B)
    e1.record(stream)
    barrier()
    sb.sync()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    launches_value = sb.last_launches() * args.steps          # kernels inside the timed region of `value`
    clocks = sampler.stop() if rank == 0 else None
    value = aggregate_throughput(F, args.steps, ms, world, dist)          # all ranks' frames / slowest rank's time
    ms_total = world * F * args.steps / value * 1e3

    # ---------------- per-stage times (same stream, CUDA events between the stages) ----------------
    sb.stage_timing(True)
    acc = {}
    for _ in range(max(3, min(args.steps, 10))):
        sb.run(F, BF, B)
        for k, v in sb.stage_times().items():
            acc.setdefault(k, []).append(v)
    sb.stage_timing(False)
    stage_ms = {k: float(np.mean(v)) for k, v in acc.items()}

    # ---------------- end-to-end arm: host buffers in, host buffers out ----------------
    sb.set_chunk(args.chunk)
    for _ in range(2):
        sb.process(pin_l.array, pin_r.array, out, BF, B)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        sb.process(pin_l.array, pin_r.array, out, BF, B)
    e1.record(stream)
    barrier()
    ms2 = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    launches_e2e = sb.last_launches() * args.steps
    e2e_blocking = aggregate_throughput(F, args.steps, ms2, world, dist)

    # ---------------- end-to-end arm, streaming form: two batch handles used alternately ----------------
    # gforb_stereo_frames_batch_async: step i is submitted on handle i % 2 after that handle's previous step has
    # completed, so the exposed tail of one step (last kernels + downloads) overlaps the uploads of the next.  Every
    # step still moves its own inputs host->device and its own results device->host inside the timed region; all
    # results are on the host when the clock stops.
    sb2 = gf.StereoBatch(W, H, F, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)
    sb2.set_stream(torch.cuda.Stream().cuda_stream)
    sb2.set_chunk(args.chunk)
    out2 = sb2.alloc_outputs(F)
    hs, outs = (sb, sb2), (out, out2)
    for i in range(4):
        hs[i % 2].sync(); hs[i % 2].submit(pin_l.array, pin_r.array, outs[i % 2], BF, B)
    sb.sync(); sb2.sync()
    barrier()
    t_e2e = time.perf_counter()
    e0.record(stream)
    for i in range(args.steps):
        hs[i % 2].sync(); hs[i % 2].submit(pin_l.array, pin_r.array, outs[i % 2], BF, B)
    sb2.sync(); sb.sync()              # sb last: its stream carries e1
    e1.record(stream)
    barrier()
    wall_ms = (time.perf_counter() - t_e2e) * 1e3
    ms3 = torch.tensor([max(e0.elapsed_time(e1), 0.0)], device="cuda")
    e2e_value = aggregate_throughput(F, args.steps, ms3, world, dist)
    assert np.array_equal(out["counts"], out2["counts"]) and np.array_equal(out["depth"], out2["depth"])
    K = sb.K
    h2d = 2 * F * W * H
    d2h = 2 * F * 4 + 2 * F * K * 28 + 2 * F * K * 32 + 2 * F * K * 4

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---------------- roofline of the dominant kernel ----------------
    counts = out["counts"][: 2 * F]
    n_kp = float(np.mean(counts))
    ex = gf.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)
    ex(left[0])
    sizes = [ex.level_geometry(l)[1:3] for l in range(NLEVELS)]
    n_cand = float(sum(len(ex.debug_candidates(l)) for l in range(NLEVELS)))
    per_img = algorithmic_bytes(sizes, n_cand, n_kp)
    per_launch = {k: v * 2 * F for k, v in per_img.items()}
    per_launch["stereo"] = (60 * 2 * n_kp + 352 * n_kp + 8 * n_kp) * F
    peaks_file = ROOT / "MEASURED_PEAKS.json"
    if peaks_file.exists():
        peak, peak_src = float(json.loads(peaks_file.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    dom = max(stage_ms, key=stage_ms.get)
    stages = {}
    for k, t in stage_ms.items():
        gbs = per_launch[k] / (t * 1e-3) / 1e9 if t > 0 else 0.0
        stages[k] = {"ms": round(t, 4), "share": round(t / sum(stage_ms.values()), 4),
                     "alg_bytes": int(per_launch[k]), "gbs": round(gbs, 1), "frac_hbm": round(gbs / peak, 4)}
    # DRAM traffic and issue-slot utilisation per stage from the committed ncu capture of this same command
    # (profiles/ncu_latest.json, written by tools/ncu_to_json.py; only meaningful at the captured batch size)
    prof_file = ROOT / "profiles" / "ncu_latest.json"
    if prof_file.exists():
        prof = json.loads(prof_file.read_text())
        if prof.get("frames_per_step") == F:
            kmap = {"pyramid": ("pyr_level0_fast_kernel", "pyr_resize_fast_kernel"), "fast": ("fast_cells_kernel",),
                    "quadtree": ("octree_kernel",), "blur": ("blur_kernel",), "describe": ("describe_kernel",),
                    "stereo": ("stereo_kernel",)}
            for k, names in kmap.items():
                ks = [prof["kernels"][n] for n in names if n in prof["kernels"]]
                if k in stages and len(ks) == len(names):
                    tt = sum(x["time_us"] for x in ks)
                    stages[k]["ncu_dram_bytes"] = int(sum(x["dram_bytes"] for x in ks))
                    stages[k]["ncu_issue_active_pct"] = round(sum(x["issue_active_pct"] * x["time_us"] for x in ks) / tt, 1)
                    stages[k]["ncu_pipe_alu_pct"] = round(sum(x["pipe_alu_pct"] * x["time_us"] for x in ks) / tt, 1)
                    stages[k]["ncu_pipe_lsu_pct"] = round(sum(x["pipe_lsu_pct"] * x["time_us"] for x in ks) / tt, 1)
    ach = stages[dom]["gbs"]
    roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": stages[dom].get("ncu_dram_bytes"), "issue_active_pct": stages[dom].get("ncu_issue_active_pct"),
                "int_pipe_alu_pct": stages[dom].get("ncu_pipe_alu_pct"),
                "peak_source": peak_src, "launches_in_stage": NLEVELS if dom == "pyramid" else 1,
                "note": "FAST/quadtree/describe/stereo are integer/LSU/latency-bound (SURVEY.md 8d): their ceiling is the "
                        "warp-instruction issue rate (issue_active_pct = ncu smsp__issue_active; int_pipe_alu_pct = share of the INT ALU "
                        "pipe's 64 lanes/clk/SM, measured in profiles/int_peaks_r1.json, that the kernel keeps busy), "
                        "not HBM; the HBM fraction is still reported for every stage under 'stages'; traffic = "
                        "dram__bytes_read+write per launch from profiles/ncu_latest.json"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[1]: 752x480 stereo pair, 1000 kp, 8 levels, 1.2, FAST 20/7, classic stereo match",
                       "frames_per_step_per_gpu": F, "parallelism": f"frame-sharded x{world}, no collective",
                       "l2": "inputs+intermediates per step (%d MB) exceed the 126 MB L2" % ((2 * F * (W * H + 3 * 1344493)) >> 20),
                       "mean_keypoints_per_image": round(n_kp, 1), "fast_candidates_per_image": n_cand},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "gpu_launches": launches_e2e,
                    "api": "gforb_stereo_frames_batch_async on two batch handles used alternately + gforb_batch_sync",
                    "blocking_single_handle": e2e_blocking, "numa_node_rank0": numa_node, "wall_ms_per_step": wall_ms / args.steps,
                    "pipeline": "chunks of %s frames: H2D / kernels / D2H of successive chunks overlap (2 copy + 4 kernel streams)" % (args.chunk or "8,24,32..,16,8 (auto)")},
            "gpu_launches": launches_value,
            "clocks": clocks, "roofline": roofline, "stages": stages}
    if not args.no_cpu_baseline and world == 1:
        os.sched_setaffinity(0, full_affinity)          # the CPU baseline gets every host core back
        threads = os.cpu_count() or 1
        fps, nframes, secs, _ = cpu_baseline(left, right, threads)
        line["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": threads, "kind": "port",
                                "reference_shaped_2_threads": round(cpu_reference_shaped(left, right), 1),
                                "sample": f"{nframes} synthetic 752x480 stereo frames (same generator), oracle port, "
                                          f"{threads} threads each running whole frames, {secs:.1f} s"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
