"""CPU, world_size 2 over gloo: the frame-sharded N>1 path (no data-path collective) — shards are
disjoint and cover all frames, per-rank results equal the single-process results, and the timing
reduction is max-over-ranks / sum of frames."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from orbslam3_cpp_fork_b200.sharding import reduce_time_and_count, round_robin, shard_range


def test_shard_partition():
    for n in (0, 1, 7, 100, 100001):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                seen += list(shard_range(n, r, world)) if n < 1000 else [len(shard_range(n, r, world))]
            if n < 1000:
                assert seen == list(range(n))
            else:
                assert sum(seen) == n and max(seen) - min(seen) <= 1
            rr = sorted(k for r in range(world) for k in (round_robin(n, r, world) if n < 1000 else []))
            assert n >= 1000 or rr == list(range(n))


def _worker(rank, world, port, n_frames, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import synth
    e = op.OracleExtractor(300, 1.2, 8, 20, 7)
    mine = {}
    for k in shard_range(n_frames, rank, world):
        _, kp, d = e.extract(synth.synth_image(200, 160, 500 + k))
        mine[k] = (len(kp), int(d.astype(np.int64).sum()))
    t, c = reduce_time_and_count(1.0 + rank, len(mine))
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        merged = {}
        for g in gathered:
            assert not (set(g) & set(merged)), "shards overlap"
            merged.update(g)
        out.put((t, c, merged))
    dist.destroy_process_group()


def test_two_rank_frame_sharding():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    n_frames = 5
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_frames, out)) for r in range(2)]
    [p.start() for p in procs]
    t, c, merged = out.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert t == 2.0 and c == n_frames            # max over ranks, sum of frames
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import synth
    e = op.OracleExtractor(300, 1.2, 8, 20, 7)
    for k in range(n_frames):
        _, kp, d = e.extract(synth.synth_image(200, 160, 500 + k))
        assert merged[k] == (len(kp), int(d.astype(np.int64).sum()))
