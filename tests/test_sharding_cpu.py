"""N>1 host logic on CPU: world-size-2 gloo run of the frame sharding + max-over-ranks timing reduction that
bench.py uses (no collective on the data path)."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total_frames, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sharding import frame_shard, aggregate_throughput
    lo, hi = frame_shard(total_frames, rank, world)
    ms = torch.tensor([10.0 * (rank + 1)])                       # pretend device time: slowest rank decides
    value = aggregate_throughput(hi - lo, steps=3, ms_local=ms, world=world, dist=dist)
    owned = torch.zeros(total_frames, dtype=torch.int32)
    owned[lo:hi] = 1
    dist.all_reduce(owned)                                        # test-only: every frame owned exactly once
    if rank == 0:
        out.put((value, owned.tolist()))
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [16, 17, 3])
def test_world2_frame_sharding(total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    value, owned = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert owned == [1] * total
    # weak-scaling aggregate: every rank's frames x steps over the MAX device time (20 ms here)
    per_rank = [len(range(*__import__("sharding").frame_shard(total, r, 2))) for r in range(2)]
    assert value == pytest.approx(sum(per_rank) * 3 / 0.020)
