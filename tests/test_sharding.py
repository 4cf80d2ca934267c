"""Host logic of the multi-GPU path on CPU: shard bounds, and a world_size-2 gloo run that checks
the shards tile the rows exactly, that per-shard scoring reassembles to the single-rank result
(oracle as the scorer: no GPU here), and the max-over-ranks / all-reduce helpers."""
import os
import socket
import sys

import numpy as np
import pytest

from mlapi_b200 import sharding


def test_shard_bounds_tile_rows_exactly():
    for n in (0, 1, 2, 7, 100, 100_000_003):
        for g in (1, 2, 3, 4, 8):
            b = sharding.shard_bounds(n, g)
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(g - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_bounds(10, 0)
    with pytest.raises(ValueError):
        sharding.my_rows(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import synth
    from oracle import lr_oracle as orc

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    n = 1001
    X, W, b = synth.make_config("cfg1_d32_binary", n)
    lo, hi = sharding.my_rows(n, rank, world)
    part = torch.from_numpy(orc.predict_proba(X[lo:hi], W, b))
    sizes = [h - l for l, h in sharding.shard_bounds(n, world)]
    gathered = [torch.empty((s, 2), dtype=torch.float64) for s in sizes]
    dist.all_gather(gathered, part) if len(set(sizes)) == 1 else dist.all_gather_object(gathered, part)
    full = torch.cat([torch.as_tensor(g) for g in gathered]).numpy()
    same = bool(np.array_equal(full, orc.predict_proba(X, W, b)))      # no cross-row reduction
    t = sharding.max_over_ranks(1.0 + rank)
    g = torch.tensor([1.0 + rank, 10.0])
    sharding.allreduce_sum(g)
    q.put((rank, same, t, g.tolist()))
    dist.destroy_process_group()


def test_two_rank_gloo_shards_reassemble():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, same, t, g in results:
        assert same, "row shards did not reassemble bit-for-bit"
        assert t == 2.0                      # max over ranks
        assert g == [3.0, 20.0]              # sum over ranks
