#!/usr/bin/env python3
"""bench.py -- stereo frames/s of the ORB stereo front-end (extract L + extract R + ComputeStereoMatches).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config C] [--frames F] [--impl reference]

A step = one pass of the whole hot path over one batch of F stereo frames; every step runs a DIFFERENT batch of
distinct seeded frames (tools/synthgen.cu, seed = 1234 + global frame number), generated on the device before the clock
starts.
  --config 1      (default) BASELINE.json configs[1]: 752x480, 1000 kp, 8 levels, 1.2, FAST 20/7, classic stereo match
           1euroc the same on the bundled EuRoC pair (tests/data/EuRoC_{l,r}.pgm = the reference's test/EuRoC_*.png)
           2tum   configs[2]: the bundled TUM pair (512x512 as shipped)
           2vga   configs[2]: synthetic 640x480 (the size BASELINE.json quotes for TUM)
           2hd    configs[2]: synthetic 1280x720, 2000 kp, 256 stereo frames per step
           3      configs[3]: the brute-force Hamming matcher at --match-n x --match-n descriptors (default 64k x 64k; its own
                  metric, G pairs/s; roofline.bound = "tensor": the kernel is a tcgen05 int8 contraction)
           4      configs[4]: 65,536 distinct synthetic 752x480 stereo frames in ONE pass, sharded over the ranks
                  (strong scaling: --steps is derived, 65,536 / (N * F))
  value   : device-resident inputs, CUDA-event timed on the launching stream, max over ranks.
  e2e     : the same through the C ABI with HOST (pinned) buffers: H2D of the images and D2H of
            keypoints/descriptors/uRight/depth inside the timed region (gforb_stereo_frames_batch_async).
  roofline: dominant kernel (by measured per-stage CUDA-event time) vs the measured HBM copy peak, with the INT-issue
            roofline beside it; every stage carries both fractions under `stages`.
  cpu_baseline / --impl reference: the reference's own code (oracle/_ref: its ORBextractor.cc compiled unmodified +
            Frame::ComputeStereoMatches lifted verbatim; kind "reference") on the box's host cores, bounded sample;
            the oracle port (kind "port") only when oracle/_ref is absent.
Multi-GPU: frames are independent -> each rank runs its own frames, no collective on the data path;
torch.distributed only provides the barrier and the max-over-ranks of the device time.
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
sys.path.insert(0, str(ROOT / "tools"))

SCALE, NLEVELS, INI_TH, MIN_TH = 1.2, 8, 20, 7
BF = 47.90639384423901
B = BF / 435.2046959714599
UNIT = "stereo_frames/s"
TAIL = "8 levels, 1.2, FAST 20/7, classic stereo match"
CONFIGS = {
    "1": dict(w=752, h=480, nfeat=1000, source="synthetic",
              workload="configs[1]: 752x480 stereo pair, 1000 kp, " + TAIL),
    "1euroc": dict(w=752, h=480, nfeat=1000, source="EuRoC",
                   workload="configs[1] on the bundled test/EuRoC_l.png + EuRoC_r.png (one pair, replicated over the batch), 1000 kp, " + TAIL),
    "2tum": dict(w=512, h=512, nfeat=1000, source="TUM",
                 workload="configs[2]: bundled test/TUM_l.png + TUM_r.png (512x512 as shipped, one pair replicated over the batch), 1000 kp, " + TAIL),
    "2vga": dict(w=640, h=480, nfeat=1000, source="synthetic",
                 workload="configs[2]: synthetic 640x480 stereo frames (the size BASELINE.json quotes for TUM), 1000 kp, " + TAIL),
    "2hd": dict(w=1280, h=720, nfeat=2000, source="synthetic",
                workload="configs[2]: synthetic 1280x720 x 256 stereo frames, 2000 kp, " + TAIL),
    "4": dict(w=752, h=480, nfeat=1000, source="synthetic", total_frames=65536,
              workload="configs[4]: 65,536 distinct synthetic 752x480 stereo frames in one pass, frame-sharded, 1000 kp, " + TAIL),
}
INT_PEAK_GOPS_FALLBACK = 18400.0      # IADD3 / IMAD / VABSDIFF4 thread-ops per second, profiles/int_peaks_r1.json
STAGE_CLASS = {"pyramid": "hbm", "fast": "hbm", "blur": "hbm", "describe": "int_issue", "stereo": "int_issue",
               "quadtree": "latency"}      # SURVEY.md 8(d) "Which roofline"


# the default configuration's constants and frame source, for the side benches under tools/
W, H, NFEAT = CONFIGS["1"]["w"], CONFIGS["1"]["h"], CONFIGS["1"]["nfeat"]


def make_frames(n, seed0=0):
    """n distinct seeded synthetic 752x480 stereo frames on the host (left, right uint8 [n, H, W]) -- the bytes the
    device generator writes for the same seeds"""
    import synthgen
    return synthgen.host_frames(seed0, n, W, H)


def metric_name(cfg):
    return f"stereo frames/s ORB extract+match @{cfg['w']}x{cfg['h']}x{cfg['nfeat']}kp"


def bundled_pair(name):
    def pgm(path):
        raw = Path(path).read_bytes()
        parts = raw.split(None, 4)
        assert parts[0] == b"P5" and parts[3] == b"255"
        w, h = int(parts[1]), int(parts[2])
        return np.frombuffer(raw[len(raw) - w * h:], np.uint8).reshape(h, w)
    return pgm(ROOT / "tests" / "data" / f"{name}_l.pgm"), pgm(ROOT / "tests" / "data" / f"{name}_r.pgm")


def host_sample(cfg, n, seed0=0):
    """n stereo frames of the config's workload on the host (the CPU arm's input): the same bytes the device generator
    writes for the same seeds, or the bundled pair."""
    if cfg["source"] == "synthetic":
        import synthgen
        return synthgen.host_frames(seed0, n, cfg["w"], cfg["h"])
    l, r = bundled_pair(cfg["source"])
    return np.repeat(l[None], n, 0), np.repeat(r[None], n, 0)


def algorithmic_bytes(W, H, level_sizes, n_cand, n_kp):
    """Per EYE image, SURVEY.md 8(d) (FAST survivors are 4 B packed here, not 8)."""
    P = sum(w * h for w, h in level_sizes)
    # the batch path writes each level's ROI plus the 3-px frame the blur reads (k_pyramid.cu v3), not the 19-px frame of 8(d)
    Ppad = sum((w + 6) * (h + 6) for w, h in level_sizes)
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
                                          "--format=csv,noheader,nounits", "-lms", "50"],
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


def cpu_model():
    try:
        for line in Path("/proc/cpuinfo").read_text().splitlines():
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def cpu_arm():
    """(module, kind, flags): the reference's own code when oracle/_ref was built (it travels prebuilt), else the port."""
    from oracle import pyref
    if (ROOT / "oracle" / "_ref" / "libgforb_ref.so").exists():
        return pyref.load("ref"), "reference", ("oracle/_ref/libgforb_ref.so: the reference's src/ORBextractor.cc unmodified + Frame::ComputeStereoMatches "
                                                "lifted verbatim, g++ -O3 -march=x86-64-v3 -ffp-contract=off (the prebuilt library has to run on whatever "
                                                "host the snapshot lands on; the -march=native build timed within 1 % of it where both ran)")
    from oracle import pyoracle
    return pyoracle, "port", "oracle/_build/libgforb_oracle.so: CPU restatement, g++ -O3 -march=x86-64-v3 -ffp-contract=off"


def cpu_baseline(cfg, left, right, threads, target_s=15.0):
    """The CPU arm on the host cores, 'all cores' shape: every worker runs whole stereo frames."""
    O, kind, flags = cpu_arm()
    n = min(len(left), 16)
    imgs = np.empty((2 * n, cfg["h"], cfg["w"]), np.uint8)
    imgs[0::2], imgs[1::2] = left[:n], right[:n]
    a = (cfg["nfeat"], SCALE, NLEVELS, INI_TH, MIN_TH, BF, B)
    secs, _ = O.bench_stereo(imgs, max(threads, 2), threads, 0, *a)
    per_frame = secs / max(threads, 2) * threads            # thread-seconds per frame
    nframes = int(max(threads * 2, min(8000, target_s * threads / max(per_frame, 1e-4))))
    secs, chk = O.bench_stereo(imgs, nframes, threads, 0, *a)
    return nframes / secs, nframes, secs, kind, flags, n


def cpu_reference_shaped(cfg, left, right, nframes=48):
    """The reference's own shape (src/Frame.cc:268-282): one frame at a time, one thread per eye, serial match."""
    O, _, _ = cpu_arm()
    n = min(len(left), 16)
    imgs = np.empty((2 * n, cfg["h"], cfg["w"]), np.uint8)
    imgs[0::2], imgs[1::2] = left[:n], right[:n]
    secs, _ = O.bench_stereo(imgs, nframes, 2, 1, cfg["nfeat"], SCALE, NLEVELS, INI_TH, MIN_TH, BF, B)
    return nframes / secs


def run_reference(args, cfg, rank, world):
    """--impl reference: the reference's CPU implementation of the path (oracle/_ref, see cpu_arm) on all host threads,
    same config / metric; each step is a bounded sample of the workload."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    left, right = host_sample(cfg, 16)
    t0 = time.time()
    vals, total_frames = [], 0
    per_step_s = max(2.0, min(15.0, 150.0 / max(args.steps + args.warmup, 1)))
    kind = flags = None
    for i in range(args.warmup + args.steps):
        fps, nframes, secs, kind, flags, ndistinct = cpu_baseline(cfg, left, right, threads, target_s=per_step_s)
        if i >= args.warmup:
            vals.append((nframes, secs)); total_frames += nframes
    tot_s = sum(s for _, s in vals)
    value = total_frames / tot_s
    sample = (f"{total_frames} stereo frames over {args.steps} steps ({ndistinct} distinct {cfg['w']}x{cfg['h']} frames of the workload, "
              f"cycled), {threads} threads each running whole frames")
    line = {"impl": "reference", "metric": metric_name(cfg), "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * tot_s / max(args.steps, 1), "higher_is_better": True,
            "scaling": "strong" if "total_frames" in cfg else "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic" if cfg["source"] == "synthetic" else "bundled",
            "config": {"workload": cfg["workload"], "frames_per_step": total_frames // max(args.steps, 1)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample,
                             "cpu_model": cpu_model(), "build": flags},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.time() - t0}
    emit(line)


def run_config3(args):
    """--config 3: BASELINE.json configs[3], the brute-force 256-bit Hamming matcher (best / second-best + ratio test) at
    N x N descriptors (default 64k x 64k, the sweep's largest; the whole sweep is tools/bench_match.py).  A step = one
    gforb_match_bruteforce call; value = device-resident descriptors, e2e = host descriptors in, host results out.  The
    kernel is a tensor-core contraction (k_match_tc.cu): roofline.bound = "tensor", 2 * 256 int8 ops per pair against
    twice the measured dense bf16 rate.  Replicas only: the matcher does not shard (one rank runs; others exit)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n = args.match_n
    rng = np.random.default_rng(1234)
    A = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    Bd = A[rng.permutation(n)] ^ np.packbits(rng.random((n, 256)) < 0.1, axis=1)        # SURVEY.md 8(d) config 3 input
    metric = f"G descriptor pairs/s brute-force Hamming {n}x{n}"
    if args.impl == "reference":
        from oracle import pyoracle as O
        threads = os.cpu_count() or 1
        rows = min(n, max(256, int(2.0e10 / n / 8)))                  # bounded sample: a slice of the query set against all of B
        vals = []
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter(); O.match_bruteforce(A[:rows], Bd, 0.6, 50, nthreads=threads); dt = time.perf_counter() - t0
            if i >= args.warmup:
                vals.append(dt)
        v = rows * n / float(np.mean(vals)) / 1e9
        emit({"impl": "reference", "metric": metric, "value": v, "unit": "Gpairs/s", "n_gpus": args.gpus, "steps": args.steps,
              "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(vals)), "higher_is_better": True, "scaling": "weak",
              "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": {"workload": f"configs[3]: {n} x {n} descriptors, ratio 0.6, TH_LOW 50"},
              "cpu_baseline": {"value": v, "unit": "Gpairs/s", "cores": threads, "kind": "port", "cpu_model": cpu_model(),
                               "sample": f"{rows} query descriptors against all {n} train descriptors per step (the reference's nearest-neighbour loop, src/ORBmatcher.cc:562-576, over all train descriptors)"},
              "e2e": {"value": v, "unit": "Gpairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})
        return 0
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; gforb has no CPU fallback (use --impl reference for the CPU arm)")
    gf = importlib.import_module("gf-orb-slam2_b200")
    m = gf.ORBmatcher(0.6)
    dA, dB = gf.DeviceArray.from_host(A), gf.DeviceArray.from_host(Bd)
    outs = [gf.DeviceArray((n,), np.int32) for _ in range(4)]
    sampler = ClockSampler(0)
    for _ in range(max(args.warmup, 3)):
        m.BruteForceDevice(dA, n, dB, n, 50, *outs)
    m.sync()
    sampler.start()
    ms = []
    for _ in range(args.steps):
        m.BruteForceDevice(dA, n, dB, n, 50, *outs)
        ms.append(m.sync())                                            # CUDA events around the call's kernels, on its stream
    clocks = sampler.stop()
    t = float(np.mean(ms))
    e = []
    for i in range(min(args.steps, 10) + 1):
        t0 = time.perf_counter(); r = m.BruteForce(A, Bd, 50); dt = time.perf_counter() - t0
        if i:
            e.append(dt)
    peaks_file = ROOT / "MEASURED_PEAKS.json"
    bf16 = float(json.loads(peaks_file.read_text())["bf16_tflops"]) if peaks_file.exists() else 1600.0
    tops = 2.0 * 256 * n * n / (t * 1e-3) / 1e12
    tc = n >= 2048 and not os.environ.get("GFORB_NO_TC")
    line = {"metric": metric, "value": n * n / (t * 1e-3) / 1e9, "unit": "Gpairs/s", "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": t, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "s8 x s8 -> s32" if tc else "u32 popc",
            "data": "synthetic", "config": {"workload": f"configs[3]: {n} x {n} descriptors, ratio 0.6, TH_LOW 50", "name": "3",
                                            "l2": "the expanded train set (%d MB) is re-streamed by every query block" % ((n * 256) >> 20),
                                            "matched_frac": float((r["match"] >= 0).mean())},
            "e2e": {"value": n * n / float(np.mean(e)) / 1e9, "unit": "Gpairs/s", "h2d_bytes_per_step": 2 * n * 32, "d2h_bytes_per_step": 4 * n * 4,
                    "api": "gforb_match_bruteforce (pageable host arrays in, host results out)"},
            "gpu_launches": (5 if tc else 2) * args.steps, "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "match_tc_kernel" if tc else "match_scan_kernel", "achieved": tops, "peak": 2.0 * bf16,
                         "unit": "TOP/s (int8)", "frac": tops / (2.0 * bf16), "traffic": None,
                         "peak_source": "2 x MEASURED_PEAKS.json bf16_tflops (dense int8 runs at twice the dense bf16 rate)",
                         "note": "algorithmic ops = 2 * 256 per descriptor pair; the kernel is bound by its epilogue's ALU pipe "
                                 "(best / second-best scan off TMEM), see DESIGN.md section 8"}}
    if not args.no_cpu_baseline:
        from oracle import pyoracle as O
        threads = os.cpu_count() or 1
        rows = min(n, max(256, int(2.0e10 / n / 8)))
        t0 = time.perf_counter(); ref = O.match_bruteforce(A[:rows], Bd, 0.6, 50, nthreads=threads); dt = time.perf_counter() - t0
        ok = all(np.array_equal(r[k][:rows], ref[k]) for k in ref)
        line["cpu_baseline"] = {"value": rows * n / dt / 1e9, "unit": "Gpairs/s", "cores": threads, "kind": "port", "cpu_model": cpu_model(),
                                "sample": f"{rows} of the {n} query descriptors against all train descriptors, {dt:.1f} s; GPU == CPU on them: {ok}"}
    emit(line)
    return 0


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
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="1", choices=sorted(CONFIGS) + ["3"])
    ap.add_argument("--match-n", type=int, default=65536, help="--config 3: descriptors a side")
    ap.add_argument("--frames", type=int, default=256, help="stereo frames per step per GPU")
    ap.add_argument("--pool", type=int, default=8, help="distinct batches kept resident and cycled by the steps")
    ap.add_argument("--impl", default="gforb", choices=["gforb", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--chunk", type=int, default=0, help="frames per pipeline chunk of the end-to-end arm")
    args = ap.parse_args()
    quiet_stdout()
    if args.config == "3":
        return run_config3(args)
    cfg = CONFIGS[args.config]
    W, H, NFEAT = cfg["w"], cfg["h"], cfg["nfeat"]

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, cfg, rank, world)
        return 0
    args.warmup = max(args.warmup, 3)

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
    from sharding import aggregate_throughput, frame_shard

    F = args.frames
    strong = "total_frames" in cfg
    if strong:                                      # configs[4]: one pass over this rank's shard of the 65,536 frames
        lo, hi = frame_shard(cfg["total_frames"], rank, world)
        NB = (hi - lo) // F
        args.steps = NB
        seed_base = lo
    else:
        NB = max(1, args.pool) if cfg["source"] == "synthetic" else 1
        seed_base = rank * NB * F
    full_affinity = os.sched_getaffinity(0)
    numa_node = bind_to_gpu_numa_node(local_rank)     # before the pinned buffers are allocated
    sb = gf.StereoBatch(W, H, F, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)
    stream = torch.cuda.Stream()        # non-default: handle 0 (legacy default) would mean "library's own stream"
    torch.cuda.set_stream(stream)
    sb.set_stream(stream.cuda_stream)
    out = sb.alloc_outputs(F)                       # pinned

    # ---------------- the frames: NB distinct batches resident in HBM, gforb_batch's layout (image 2f = left, 2f+1 = right)
    pool = torch.empty((NB, 2 * F, H, W), dtype=torch.uint8, device="cuda")
    if cfg["source"] == "synthetic":
        import synthgen
        for bidx in range(NB):
            synthgen.device_batch_input(seed_base + bidx * F, F, W, H, pool[bidx].data_ptr(), stream.cuda_stream)
    else:
        l, r = bundled_pair(cfg["source"])
        pool[0, 0::2] = torch.from_numpy(l.copy()).cuda()
        pool[0, 1::2] = torch.from_numpy(r.copy()).cuda()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_step(i):
        sb.set_device_input(pool[i % NB].data_ptr())
        sb.run(F, BF, B)

    # ---------------- device-resident arm ----------------
    for i in range(args.warmup):
        run_step(i)
    sb.sync()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(args.steps):
        run_step(i)
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
    for i in range(max(3, min(args.steps, 10))):
        run_step(i)
        for k, v in sb.stage_times().items():
            acc.setdefault(k, []).append(v)
    sb.stage_timing(False)
    stage_ms = {k: float(np.mean(v)) for k, v in acc.items()}
    sb.set_device_input(None)

    # ---------------- end-to-end arm: host buffers in, host buffers out ----------------
    NBH = min(NB, 8)                                 # pinned host copies of the first NBH batches, cycled by the steps
    pin_l = gf.PinnedArray((NBH, F, H, W), np.uint8); pin_r = gf.PinnedArray((NBH, F, H, W), np.uint8)
    for bidx in range(NBH):
        pin_l.array[bidx] = pool[bidx, 0::2].cpu().numpy(); pin_r.array[bidx] = pool[bidx, 1::2].cpu().numpy()
    frame0 = (pin_l.array[0, 0].copy(), pin_r.array[0, 0].copy())
    del pool
    torch.cuda.empty_cache()
    sb.set_chunk(args.chunk)
    for i in range(2):
        sb.process(pin_l.array[i % NBH], pin_r.array[i % NBH], out, BF, B)
    e2e_steps = min(args.steps, 64)
    barrier()
    e0.record(stream)
    for i in range(e2e_steps):
        sb.process(pin_l.array[i % NBH], pin_r.array[i % NBH], out, BF, B)
    e1.record(stream)
    barrier()
    ms2 = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    launches_e2e = sb.last_launches() * e2e_steps
    e2e_blocking = aggregate_throughput(F, e2e_steps, ms2, world, dist)

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
        hs[i % 2].sync(); hs[i % 2].submit(pin_l.array[i % NBH], pin_r.array[i % NBH], outs[i % 2], BF, B)
    sb.sync(); sb2.sync()
    barrier()
    t_e2e = time.perf_counter()
    e0.record(stream)
    for i in range(e2e_steps):
        hs[i % 2].sync(); hs[i % 2].submit(pin_l.array[i % NBH], pin_r.array[i % NBH], outs[i % 2], BF, B)
    sb2.sync(); sb.sync()              # sb last: its stream carries e1
    e1.record(stream)
    barrier()
    wall_ms = (time.perf_counter() - t_e2e) * 1e3
    ms3 = torch.tensor([max(e0.elapsed_time(e1), 0.0)], device="cuda")
    e2e_value = aggregate_throughput(F, e2e_steps, ms3, world, dist)
    # ---------------- the copy-only ceiling of that arm, measured on THIS box in THIS run: the same calls with the
    # kernels switched off (gforb_batch_debug_copy_only) -- same chunks, streams, events and bytes, all ranks together
    sb.debug_copy_only(True); sb2.debug_copy_only(True)
    for i in range(2):
        hs[i % 2].sync(); hs[i % 2].submit(pin_l.array[i % NBH], pin_r.array[i % NBH], outs[i % 2], BF, B)
    sb.sync(); sb2.sync()
    barrier()
    e0.record(stream)
    for i in range(e2e_steps):
        hs[i % 2].sync(); hs[i % 2].submit(pin_l.array[i % NBH], pin_r.array[i % NBH], outs[i % 2], BF, B)
    sb2.sync(); sb.sync()
    e1.record(stream)
    barrier()
    ms4 = torch.tensor([max(e0.elapsed_time(e1), 1e-3)], device="cuda")
    ceiling_live = aggregate_throughput(F, e2e_steps, ms4, world, dist)
    sb.debug_copy_only(False); sb2.debug_copy_only(False)
    hs[0].submit(pin_l.array[0], pin_r.array[0], outs[0], BF, B); hs[1].submit(pin_l.array[0], pin_r.array[0], outs[1], BF, B)
    sb.sync(); sb2.sync()                      # real results back in the output buffers (counts feed the roofline below)
    K = sb.K
    h2d = 2 * F * W * H
    d2h = 2 * F * 4 + 2 * F * K * 28 + 2 * F * K * 32 + 2 * F * K * 4

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---------------- roofline of the dominant kernel ----------------
    counts = np.concatenate([out["counts"][: 2 * F], out2["counts"][: 2 * F]])
    n_kp = float(np.mean(counts))
    ex = gf.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)
    n_cand = 0.0
    for img in frame0:
        ex(img)
        n_cand += 0.5 * float(sum(len(ex.debug_candidates(l)) for l in range(NLEVELS)))
    sizes = [ex.level_geometry(l)[1:3] for l in range(NLEVELS)]
    per_img = algorithmic_bytes(W, H, sizes, n_cand, n_kp)
    per_launch = {k: v * 2 * F for k, v in per_img.items()}
    per_launch["stereo"] = (60 * 2 * n_kp + 352 * n_kp + 8 * n_kp) * F
    peaks_file = ROOT / "MEASURED_PEAKS.json"
    if peaks_file.exists():
        peak, peak_src = float(json.loads(peaks_file.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    int_peak, int_src = INT_PEAK_GOPS_FALLBACK, "fallback constant"
    ip = ROOT / "profiles" / "int_peaks_r1.json"
    if ip.exists():
        int_peak = float(json.loads(ip.read_text())["rates"]["IADD3"]["gops"])
        int_src = "profiles/int_peaks_r1.json (IADD3 / IMAD / VABSDIFF4 / IDP4A thread-ops per second, microbenchmarked on this pool's B200)"
    # instruction-issue ceiling: 4 schedulers x 32 lanes per SM and clock; the INT32 ALU pipe alone takes half of it
    # (64 lanes/clk/SM: the microbenchmarked IADD3 / IMAD / VABSDIFF4 rate in profiles/int_peaks_r1.json)
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    issue_peak = sms * 128 * float((clocks or {}).get("sm_max_mhz") or 1965.0) * 1e6 / 1e9
    dom = max(stage_ms, key=stage_ms.get)
    stages = {}
    for k, t in stage_ms.items():
        gbs = per_launch[k] / (t * 1e-3) / 1e9 if t > 0 else 0.0
        stages[k] = {"ms": round(t, 4), "share": round(t / sum(stage_ms.values()), 4), "class": STAGE_CLASS[k],
                     "alg_bytes": int(per_launch[k]), "gbs": round(gbs, 1), "frac_hbm": round(gbs / peak, 4)}
    # DRAM traffic, instruction counts and issue-slot utilisation per stage from the committed ncu capture of this same
    # command (profiles/ncu_latest.json, written by tools/ncu_to_json.py; only meaningful at the captured config / batch)
    prof_file = ROOT / "profiles" / "ncu_latest.json"
    if prof_file.exists():
        prof = json.loads(prof_file.read_text())
        if prof.get("frames_per_step") == F and prof.get("config", "1") == args.config:
            kmap = {"pyramid": ("pyr_level0_fast_kernel", "pyr_resize_fast_kernel", "pyr_cascade_kernel"), "fast": ("fast_cells_kernel",),
                    "quadtree": ("octree_kernel", "octree_redo_kernel"), "blur": ("blur_kernel",), "describe": ("describe_kernel",),
                    "stereo": ("stereo_kernel",)}
            for k, names in kmap.items():
                ks = [prof["kernels"][n] for n in names if n in prof["kernels"]]
                if k in stages and ks:
                    tt = sum(x["time_us"] for x in ks)
                    stages[k]["ncu_dram_bytes"] = int(sum(x["dram_bytes"] for x in ks))
                    stages[k]["ncu_issue_active_pct"] = round(sum(x["issue_active_pct"] * x["time_us"] for x in ks) / tt, 1)
                    stages[k]["ncu_pipe_alu_pct"] = round(sum(x["pipe_alu_pct"] * x["time_us"] for x in ks) / tt, 1)
                    stages[k]["ncu_pipe_lsu_pct"] = round(sum(x["pipe_lsu_pct"] * x["time_us"] for x in ks) / tt, 1)
                    if all("warp_inst" in x for x in ks):
                        # INT-issue roofline: thread-instructions per launch (ncu smsp__inst_executed x 32) over the
                        # LIVE stage time, against the microbenchmarked INT32 thread-op rate
                        tops = 32.0 * sum(x["warp_inst"] for x in ks)
                        gops = tops / (stage_ms[k] * 1e-3) / 1e9
                        stages[k]["int_issue"] = {"thread_inst_per_launch": int(tops), "gops": round(gops, 1),
                                                  "frac": round(gops / issue_peak, 4),
                                                  "alu_pipe_frac": round(stages[k]["ncu_pipe_alu_pct"] / 100.0, 4)}
    ach = stages[dom]["gbs"]
    roofline = {"bound": STAGE_CLASS[dom], "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": stages[dom].get("ncu_dram_bytes"), "issue_active_pct": stages[dom].get("ncu_issue_active_pct"),
                "int_pipe_alu_pct": stages[dom].get("ncu_pipe_alu_pct"),
                "int_issue": dict(stages[dom].get("int_issue", {}), peak=round(issue_peak, 1), unit="G thread-inst/s",
                                  peak_source="%d SMs x 4 schedulers x 32 lanes x max SM clock (issue slots)" % sms,
                                  alu_pipe_peak=int_peak, alu_pipe_peak_source=int_src) if "int_issue" in stages[dom] else None,
                "peak_source": peak_src, "launches_in_stage": "see stages",
                "note": "bound = the stage's class in SURVEY.md 8(d) (pyramid / FAST / blur: hbm; descriptor / stereo: int_issue; quadtree: "
                        "latency); achieved / peak / frac are algorithmic bytes over the live CUDA-event time against the measured HBM "
                        "copy rate.  FAST is classed hbm there but measures instruction-issue bound (ncu: issue slots 80 % busy, DRAM 7 %), so "
                        "its int_issue figure (thread-instructions from the committed ncu capture over the live time, against the "
                        "issue-slot ceiling; alu_pipe_frac = share of the INT32 ALU pipe, whose own ceiling is half of that) is the meaningful one; both fractions are under `stages` for every stage; traffic = "
                        "dram__bytes_read+write per launch from profiles/ncu_latest.json"}

    ceiling = ceiling_live
    line = {"metric": metric_name(cfg), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic" if cfg["source"] == "synthetic" else "bundled",
            "config": {"workload": cfg["workload"], "name": args.config,
                       "frames_per_step_per_gpu": F, "parallelism": f"frame-sharded x{world}, no collective",
                       "frames": ("%d distinct seeded frames per GPU (tools/synthgen.cu, seed 1234 + frame), %d resident batches cycled by the steps"
                                  % (NB * F, NB)) if cfg["source"] == "synthetic" else "one bundled pair replicated over the batch",
                       "frames_total": world * F * args.steps,
                       "l2": "inputs+intermediates per step (%d MB) exceed the 126 MB L2; consecutive steps read different frames"
                             % ((2 * F * (W * H + 3 * per_img["pyramid"])) >> 20),
                       "mean_keypoints_per_image": round(n_kp, 1), "fast_candidates_per_image": n_cand},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "gpu_launches": launches_e2e,
                    "api": "gforb_stereo_frames_batch_async on two batch handles used alternately + gforb_batch_sync",
                    "host_batches": NBH,
                    "copy_only_ceiling": ceiling, "link_frac": (e2e_value / ceiling) if ceiling else None,
                    "ceiling_source": "measured in this run on this box: the same gforb_stereo_frames_batch_async calls with the kernels "
                                      "switched off (gforb_batch_debug_copy_only), all ranks together; profiles/h2d_ceiling_r2.json "
                                      "holds an independent torch-free copy loop (tools/h2d_ceiling.py) from another box",
                    "blocking_single_handle": e2e_blocking, "numa_node_rank0": numa_node, "wall_ms_per_step": wall_ms / e2e_steps,
                    "pipeline": "chunks of %s frames: H2D / kernels / D2H of successive chunks overlap (2 copy + 4 kernel streams)" % (args.chunk or "8,24,32..,16,8 (auto)")},
            "gpu_launches": launches_value,
            "clocks": clocks, "roofline": roofline, "stages": stages}
    if not args.no_cpu_baseline and world == 1:
        os.sched_setaffinity(0, full_affinity)          # the CPU baseline gets every host core back
        threads = os.cpu_count() or 1
        hl, hr = host_sample(cfg, 16, seed_base)
        fps, nframes, secs, kind, flags, nd = cpu_baseline(cfg, hl, hr, threads)
        line["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": threads, "kind": kind, "cpu_model": cpu_model(), "build": flags,
                                "reference_shaped_2_threads": round(cpu_reference_shaped(cfg, hl, hr), 1),
                                "sample": f"{nframes} stereo frames ({nd} distinct frames of the same workload -- the bytes the GPU arm's first "
                                          f"frames hold -- cycled), {threads} threads each running whole frames, {secs:.1f} s"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
