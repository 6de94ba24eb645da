"""world_size-2 gloo test (CPU) of the multi-GPU host logic: contiguous clip shards cover the batch with
no overlap, the timing reduction is a max over ranks and the optional state gather is rank ordered."""
import os
import socket
import sys
from pathlib import Path

import torch
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_clips, out):
    sys.path.insert(0, str(ROOT))
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from yoho_b200 import dist as yd
    import torch.distributed as dist
    r, _, w = yd.init_process_group("gloo")
    lo, hi = yd.shard_range(n_clips, r, w)
    states = torch.arange(lo, hi, dtype=torch.float32).view(-1, 1, 1).expand(-1, 2, 3).contiguous()
    yd.barrier()
    elapsed = yd.max_over_ranks(10.0 + r)
    total = yd.sum_over_ranks(hi - lo)
    gathered = yd.gather_states(states) if (hi - lo) * w == n_clips else None
    out.put((r, lo, hi, elapsed, total, None if gathered is None else gathered[:, 0, 0].tolist()))
    dist.destroy_process_group()


def test_two_rank_sharding_and_reductions():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, n = _free_port(), 10
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (r0, lo0, hi0, e0, t0, g0), (r1, lo1, hi1, e1, t1, g1) = res
    assert (lo0, hi0, lo1, hi1) == (0, 5, 5, 10)
    assert e0 == e1 == 11.0                   # max over ranks, identical on both
    assert t0 == t1 == float(n)
    assert g0 == g1 == [float(i) for i in range(n)]
