#!/usr/bin/env python3
"""Copy-only ceiling of the end-to-end arm.

Every rank, concurrently, moves exactly what one bench.py step moves -- F stereo frames of 752x480 bytes host->device and
the step's results device->host (counts, keypoints, descriptors, uRight, depth: bench.py's d2h formula) -- from / to
pinned host memory with cudaMemcpyAsync on two streams, with NO kernels.  The rate it reaches, in stereo frames/s, is
what the box's host memory + PCIe fabric lets the end-to-end arm reach at that rank count; bench.py divides its e2e
value by it (`e2e.link_frac`).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/h2d_ceiling.py --out profiles/h2d_ceiling_r2.json          (N = 1: plain `python tools/h2d_ceiling.py`)

Variants timed (each: 3 warm-up + `--steps` timed steps, CUDA events on the copy streams, max over ranks):
  whole      one H2D copy per eye (2 x 92 MB) + one D2H copy per result array
  chunked    the same bytes cut the way gforb_stereo_frames_batch cuts them (8-frame head chunk, then 24, 32, ...)
  wc         `whole` with the input staging buffers allocated cudaHostAllocWriteCombined
  h2d_only / d2h_only   one direction alone (shows whether the two directions share a limit)
The output file accumulates one record per world size.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

W, H, K = 752, 480, 1032          # K: per-image keypoint capacity of the batch outputs (gf.StereoBatch.K at 1000 features)


def chk(r):
    err = r[0]
    if int(err) != 0:
        raise RuntimeError(f"CUDA error {err}")
    return r[1] if len(r) == 2 else r[1:]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--keys", type=int, default=K)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    from cuda.bindings import runtime as rt
    import torch.distributed as dist
    chk(rt.cudaSetDevice(local))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo")
    F, Kk = args.frames, args.keys
    h2d_bytes = 2 * F * W * H
    d2h_parts = [2 * F * 4, 2 * F * Kk * 28, 2 * F * Kk * 32, F * Kk * 4, F * Kk * 4]
    d2h_bytes = sum(d2h_parts)

    def host(nbytes, flags=0):
        p = chk(rt.cudaHostAlloc(nbytes, flags))
        (np.ctypeslib.as_array((__import__("ctypes").c_uint8 * nbytes).from_address(int(p))))[:] = 1    # first touch
        return p

    h_in = host(h2d_bytes)
    h_in_wc = host(h2d_bytes, rt.cudaHostAllocWriteCombined)
    h_out = host(d2h_bytes)
    d_in = chk(rt.cudaMalloc(h2d_bytes)); d_out = chk(rt.cudaMalloc(d2h_bytes))
    s_up = chk(rt.cudaStreamCreate()); s_dn = chk(rt.cudaStreamCreate())
    e = [chk(rt.cudaEventCreate()) for _ in range(4)]
    H2D, D2H = rt.cudaMemcpyKind.cudaMemcpyHostToDevice, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost

    def chunks():            # the frame chunking of the end-to-end arm (bench.py `pipeline`): 8, 24, 32, ..., 16, 8
        if F <= 56:
            return [F]
        mid = F - 56
        return [8, 24] + [32] * (mid // 32) + ([mid % 32] if mid % 32 else []) + [16, 8]

    def step(variant):
        src = h_in_wc if variant == "wc" else h_in
        up = variant != "d2h_only"; dn = variant != "h2d_only"
        if variant == "chunked":
            f0 = 0
            for c in chunks():
                for eye in range(2):
                    off = (eye * F + f0) * W * H
                    chk(rt.cudaMemcpyAsync(int(d_in) + off, int(src) + off, c * W * H, H2D, s_up))
                o = 0
                for part in d2h_parts:
                    a = part * f0 // F; b = part * (f0 + c) // F
                    chk(rt.cudaMemcpyAsync(int(h_out) + o + a, int(d_out) + o + a, b - a, D2H, s_dn))
                    o += part
                f0 += c
            return
        if up:
            for eye in range(2):
                off = eye * F * W * H
                chk(rt.cudaMemcpyAsync(int(d_in) + off, int(src) + off, F * W * H, H2D, s_up))
        if dn:
            o = 0
            for part in d2h_parts:
                chk(rt.cudaMemcpyAsync(int(h_out) + o, int(d_out) + o, part, D2H, s_dn))
                o += part

    def barrier():
        chk(rt.cudaDeviceSynchronize())
        if world > 1:
            dist.barrier()

    def timed(variant):
        for _ in range(3):
            step(variant)
        barrier()
        chk(rt.cudaEventRecord(e[0], s_up)); chk(rt.cudaEventRecord(e[2], s_dn))
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step(variant)
        chk(rt.cudaEventRecord(e[1], s_up)); chk(rt.cudaEventRecord(e[3], s_dn))
        chk(rt.cudaDeviceSynchronize())
        wall = (time.perf_counter() - t0) * 1e3
        ms_up = chk(rt.cudaEventElapsedTime(e[0], e[1])); ms_dn = chk(rt.cudaEventElapsedTime(e[2], e[3]))
        ms = max(ms_up, ms_dn)
        if world > 1:
            import torch
            t = torch.tensor([ms, ms_up, ms_dn, wall], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, ms_up, ms_dn, wall = t.tolist()
        barrier()
        per = ms / args.steps
        up_b = 0 if variant == "d2h_only" else h2d_bytes
        dn_b = 0 if variant == "h2d_only" else d2h_bytes
        return {"ms_per_step": per, "stereo_frames_per_s": world * F / (per * 1e-3),
                "h2d_gbs_per_rank": up_b / (ms_up / args.steps * 1e-3) / 1e9 if up_b else 0.0,
                "d2h_gbs_per_rank": dn_b / (ms_dn / args.steps * 1e-3) / 1e9 if dn_b else 0.0,
                "h2d_gbs_total": world * up_b / (ms_up / args.steps * 1e-3) / 1e9 if up_b else 0.0,
                "wall_ms_per_step": wall / args.steps}

    res = {v: timed(v) for v in ("whole", "chunked", "wc", "h2d_only", "d2h_only")}
    if rank == 0:
        try:
            aff = len(os.sched_getaffinity(0))
        except Exception:
            aff = None
        rec = {"n_gpus": world, "frames_per_step_per_gpu": F, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
               "steps": args.steps, "host_cpus": os.cpu_count(), "affinity_cpus": aff, "variants": res,
               "ceiling_stereo_frames_per_s": max(res["whole"]["stereo_frames_per_s"], res["chunked"]["stereo_frames_per_s"],
                                                  res["wc"]["stereo_frames_per_s"])}
        print(json.dumps(rec))
        if args.out:
            p = Path(args.out)
            allr = json.loads(p.read_text()) if p.exists() else {}
            allr[str(world)] = rec
            p.parent.mkdir(parents=True, exist_ok=True)
            p.write_text(json.dumps(allr, indent=1) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
