"""world_size-2 gloo test of the N>1 host logic (stream partition, no data-path collective, max-over-ranks
timing, gather of the result rows).  Runs on CPU: each rank plays its GPU with the CPU checker as the worker
(tests may use the oracle), and the concatenation must equal the single-process result bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_exactly():
    from simplefe_b200.shard import shard_range
    for S in (0, 1, 5, 127, 1024):
        for W in (1, 2, 4, 8):
            r = [shard_range(S, k, W) for k in range(W)]
            assert r[0][0] == 0 and r[-1][1] == S
            assert all(r[k][1] == r[k + 1][0] for k in range(W - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    assert [shard_range(1024, k, 8) for k in (0, 7)] == [(0, 128), (896, 1024)]
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _worker(rank, world, port, S, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    import fixtures as F
    from oracle import oracle as O
    from simplefe_b200.shard import gather_rows, max_over_ranks, shard_range
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = dict(F.CONFIGS["cfg5_small"], S=S, N=30011)
        lo, hi = shard_range(S, rank, world)
        iq = F.make_input(cfg, streams=range(lo, hi))          # stream s uses seed + s whatever rank owns it
        if hi > lo:
            y, _ = O.chain_run(iq, cfg["in_fmt"], cfg["blkconv"], cfg["stage_a"], cfg["stage_b"])
        else:
            y = np.zeros((0, 0), np.complex64)
        full = gather_rows(np.ascontiguousarray(y), S)
        t = max_over_ranks(1.0 + rank)                          # the slowest rank defines the step
        q.put((rank, lo, hi, full.shape, full.tobytes(), t))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("S", [5, 1])
def test_two_ranks_equal_one_process(S):
    import torch.multiprocessing as mp
    import fixtures as F
    from oracle import oracle as O
    O.build()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, S, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cfg = dict(F.CONFIGS["cfg5_small"], S=S, N=30011)
    ref, _ = O.chain_run(F.make_input(cfg), cfg["in_fmt"], cfg["blkconv"], cfg["stage_a"], cfg["stage_b"])
    assert [(r[1], r[2]) for r in res] == ([(0, 3), (3, 5)] if S == 5 else [(0, 1), (1, 1)])
    for rank, lo, hi, shape, raw, t in res:
        got = np.frombuffer(raw, np.complex64).reshape(shape)
        assert got.shape == ref.shape and np.array_equal(got, ref)
        assert t == 2.0
