"""Multi-GPU path on the CPU: streams are independent, shards are byte-balanced, there is no
exchange step.  Two `gloo` ranks each take their share of the same descriptor table from the library's own
partitioner (infl_batch_partition, host only), decode it with the CPU oracle (test infrastructure standing in
for the GPU here), and the ranks agree on: disjoint shares that cover the batch, balance, the aggregate byte
count bench.py reports as `value`, and the max-over-ranks time."""
import os
import sys
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def _batch():
    from defl_b200 import synth
    specs = synth.config5_specs(6 << 20, first_seed=77, lo=4096, hi=1 << 20)
    return synth.make_streams(specs, threads=2)


def test_partition_is_balanced_and_complete():
    import defl_b200 as d
    rng = np.random.default_rng(5)
    for n, parts in ((1, 2), (7, 8), (1000, 2), (1000, 8), (5, 1)):
        src = rng.integers(1, 1 << 22, size=n)
        dst = src * rng.integers(1, 12, size=n)
        part, load = d.batch_partition(src, dst, parts)
        assert part.min() >= 0 and part.max() < parts
        w = (src + dst).astype(np.uint64)
        for g in range(parts):
            assert int(w[part == g].sum()) == int(load[g])
        assert int(load.sum()) == int(w.sum())
        # greedy longest-processing-time: no shard exceeds the mean by more than the largest stream
        assert int(load.max()) <= int(w.sum()) // parts + int(w.max()) + parts
    part, load = d.batch_partition([], [], 4)
    assert len(part) == 0 and int(load.sum()) == 0


def _rank_main(rank, world, port, q):
    import torch
    import torch.distributed as dist
    import time
    import defl_b200 as d
    import oracle_lib as ol
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        streams = _batch()                                   # same seeds on every rank: the same table
        src = [len(z) for _, z in streams]
        dst = [len(r) for r, _ in streams]
        part, load = d.batch_partition(src, dst, world)
        mine = [i for i in range(len(streams)) if part[i] == rank]
        t0 = time.perf_counter()
        produced, adlers = 0, {}
        for i in mine:
            raw, z = streams[i]
            code, out, r = ol.port_inflate(z, len(raw), 1)
            assert code == 0 and out == raw
            produced += len(out)
            adlers[i] = r.adler
        dt = time.perf_counter() - t0
        # what bench.py does with its per-rank numbers: SUM of units, MAX of time
        t = torch.tensor([float(produced), dt, float(len(mine))], dtype=torch.float64)
        tot = t.clone(); dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        shares = [None] * world
        dist.all_gather_object(shares, (mine, adlers))
        q.put((rank, int(tot[0].item()), float(mx[1].item()), int(tot[2].item()), shares, [int(x) for x in load]))
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_a_batch_without_exchange():
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    streams = _batch()
    total = sum(len(r) for r, _ in streams)
    for rank, produced, tmax, count, shares, load in got:
        assert produced == total and count == len(streams) and tmax > 0
        a, b = set(shares[0][0]), set(shares[1][0])
        assert not (a & b) and a | b == set(range(len(streams)))
        w = [len(z) + len(r) for r, z in streams]
        assert abs(load[0] - load[1]) <= max(w) + 2
        for mine, adlers in shares:
            for i in mine:
                assert adlers[i] == (zlib.adler32(streams[i][0]) & 0xFFFFFFFF)
    assert got[0][1:4:2] == got[1][1:4:2]
