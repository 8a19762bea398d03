"""Host logic of the sharded ladder on CPU, world_size 2, gloo: shard layout,
tile ownership, handle exchange and table split/merge.  No GPU work."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from analyzer_b200 import sharded
    handle = bytes([rank + 1]) * 64
    got = sharded.exchange_handles(handle)
    ok = got == [bytes([r + 1]) * 64 for r in range(world)]
    n_players = 1001
    rows = sharded.shard_rows(n_players, rank, world)
    total = torch.tensor([rows])
    dist.all_reduce(total)
    ok = ok and int(total) == n_players
    table = torch.arange(n_players * 2, dtype=torch.float64).reshape(n_players, 2)
    mine = sharded.split_table(table, rank, world)
    ok = ok and mine.shape[0] == rows and all(int(mine[i, 0]) // 2 % world == rank for i in range(rows))
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    ok = ok and torch.equal(sharded.merge_tables(gathered), table)
    tiles = sharded.tiles_of(rank, world, 1000, 16)
    all_tiles = [None] * world
    dist.all_gather_object(all_tiles, tiles)
    flat = sorted(t for ts in all_tiles for t in ts)
    ok = ok and flat[0][0] == 0 and flat[-1][1] == 1000 and all(a[1] == b[0] for a, b in zip(flat, flat[1:]))
    results[rank] = ok
    dist.destroy_process_group()


def test_sharded_host_logic_world2():
    world = 2
    with mp.Manager() as mgr:
        results = mgr.dict()
        mp.spawn(_worker, args=(world, 29531, results), nprocs=world, join=True)
        assert dict(results) == {0: True, 1: True}


def test_layout_math():
    from analyzer_b200 import sharded
    for world in (1, 2, 4, 8):
        for u in (1, 7, 8, 1000):
            assert sum(sharded.shard_rows(u, r, world) for r in range(world)) == u
            for p in range(u):
                assert sharded.local_row(p, world) < sharded.shard_rows(u, sharded.owner(p, world), world)
    with pytest.raises(ValueError):
        sharded.ShardedLadder(None, 10, 0, 3)
