"""The N > 1 path on CPU: two ranks over gloo (world_size 2) plan their disjoint window shards of one range exactly as
two torchrun ranks with vgen_cuda_set_shard(rank, 2) would, exchange them, and check that the union covers the range once.
Hit merging (concatenate per-rank ascending lists, sort) is checked the same way with oracle-derived hits."""
import os
import sys

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, start, end, hint, early, q):
    sys.path.insert(0, ROOT)
    import vgen_b200 as vg
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = vg.plan_windows(start, end, world, rank, sm_count=148, keys_per_launch_hint=hint, can_end_early=early)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    dist.barrier()
    if rank == 0:
        q.put(gathered)
    dist.destroy_process_group()


def _run(start, end, hint, early, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() + hash((start, end, hint, early))) % 300
    procs = [ctx.Process(target=_worker, args=(r, world, port, start, end, hint, early, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


@pytest.mark.parametrize("start,n,hint,early", [
    (2**64, 5 * 77594624 + 12345, 0, False),       # long sparse scan: full windows only, round-robin
    (2**64, 3_000_000, 524288, True),              # hit-limited dense scan: ramp levels first
    (1, 1000, 0, False),                           # tiny range: one small window on rank 0, nothing on rank 1
    (0, 2**20, 524288, False),                     # range starting at the invalid scalar 0
])
def test_two_ranks_cover_range_disjointly(start, n, hint, early):
    end = start + n - 1
    per_rank = _run(start, end, hint, early)
    lo = max(start, 1)
    wins = sorted(w for r in per_rank for w in r)
    pos = lo
    for first, cnt in wins:
        assert first == pos and cnt > 0        # contiguous, no overlap, no gap
        pos += cnt
    assert pos == end + 1
    if n > 10 * 77594624 // 2 and not early:
        assert abs(len(per_rank[0]) - len(per_rank[1])) <= 1      # balanced round-robin


def test_rank_union_equals_single_rank_plan():
    start, end = 2**65, 2**65 + 9 * 77594624
    one = __import__("vgen_b200").plan_windows(start, end, 1, 0)
    two = _run(start, end, 0, False)
    assert sum(c for _, c in one) == sum(c for r in two for _, c in r) == end - start + 1
