"""Host logic of the sharded ladder on CPU, world_size 2, gloo: shard layout,
block ownership, handle exchange, table split/merge, arena layout.  No GPU work."""
import ctypes as C
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
    # every match of a stream lands on exactly one rank, in order, and the slices tile the stream
    n = 1000
    stream = torch.arange(n * 6, dtype=torch.int32).reshape(n, 6)
    part = sharded.take_slice(stream, rank, world)
    ok = ok and part.shape[0] == sharded.local_matches(n, rank, world)
    ok = ok and all(int(part[lm, 0]) // 6 == sharded.global_index(lm, rank, world) for lm in range(part.shape[0]))
    parts = [None] * world
    dist.all_gather_object(parts, part)
    back = torch.empty_like(stream)
    for r, pt in enumerate(parts):
        for lm in range(pt.shape[0]):
            back[sharded.global_index(lm, r, world)] = pt[lm]
    ok = ok and torch.equal(back, stream)
    # all ranks agree on every rank's arena layout
    lay = [sharded.arena_layout(n_players, n, 6, g, world) for g in range(world)]
    lays = [None] * world
    dist.all_gather_object(lays, lay)
    ok = ok and all(l == lay for l in lays)
    results[rank] = ok
    dist.destroy_process_group()


def test_sharded_host_logic_world2():
    world = 2
    with mp.Manager() as mgr:
        results = mgr.dict()
        mp.spawn(_worker, args=(world, 29531, results), nprocs=world, join=True)
        assert dict(results) == {0: True, 1: True}


def test_layout_math():
    from analyzer_b200 import _lib, sharded
    lib = _lib.lib()
    for world in (1, 2, 4, 8, 16):
        for u in (1, 7, 8, 1000):
            assert sum(sharded.shard_rows(u, r, world) for r in range(world)) == u
            for p in range(u):
                assert sharded.local_row(p, world) < sharded.shard_rows(u, sharded.owner(p, world), world)
        for n in (0, 1, 31, 32, 33, 64 * world - 1, 64 * world, 64 * world + 5, 1000, 12345):
            counts = [sharded.local_matches(n, r, world) for r in range(world)]
            assert sum(counts) == n
            assert counts == [lib.ts_shard_local_matches(n, r, world) for r in range(world)]
            seen = [0] * world
            for m in range(n):
                r = sharded.match_rank(m, world)
                assert sharded.local_index(m, world) == seen[r]           # local order is stream order
                assert sharded.global_index(seen[r], r, world) == m
                seen[r] += 1
            assert seen == counts
    offs, size = sharded.arena_layout(1000, 5000, 6, 1, 4)
    assert all(o % 256 == 0 for o in offs) and size % 256 == 0 and offs == sorted(offs)
    with pytest.raises(ValueError):
        sharded.ShardedLadder(None, 10, 100, 6, 0, 3)


def test_shard_entry_points_refuse_bad_views():
    """Argument checks of the C ABI that need no GPU."""
    from analyzer_b200 import _lib
    lib = _lib.lib()
    v = _lib.TsShardView()
    v.rank, v.n_shards = 0, 3
    assert lib.ts_shard_link(None, C.byref(v), 10, 6, 10, 1, 0, None, 0, None) == -1
    assert b"power of two" in lib.ts_last_error()
    assert lib.ts_shard_rate(None, C.byref(v), 10, 6, 10, 1, 0, None, 0, 0, None) == -1
    assert lib.ts_shard_local_matches(-1, 0, 2) < 0
    assert lib.ts_shard_link_workspace_bytes(1 << 22, 1 << 20) > (1 << 20) * 16 + (1 << 22) * 8
