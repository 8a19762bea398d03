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


def test_link_exchange_sizes_and_argument_checks():
    """The link exchange's host-side sizes agree between Python and the C ABI, and its entry points
    refuse views that do not fit the window before they touch a GPU."""
    from analyzer_b200 import _lib, sharded
    lib = _lib.lib()
    for n, world in ((1, 1), (31, 2), (32 * 8 + 1, 8), (100_000_000, 8), (500_000_000, 16)):
        nb = sharded.link_blocks_capacity(n, world)
        blocks = (n + 31) // 32
        assert nb % 8 == 0 and nb * world >= blocks
        # scratch of ts_shard_send: chunk sums per owner, totals, offsets of the own blocks
        need = lib.ts_shard_exchange_workspace_bytes(n, world)
        assert need >= world * nb * 4 and need < world * nb * 4 + world * (blocks // 4096 + 2) * 4 + 4096
        # the staging-free link workspace only holds the sort's buffers
        cap = sharded.owned_capacity(n, 6, world)
        assert cap <= max(1, n * 6) and (world > 1 or cap == n * 6)
        assert lib.ts_shard_link_workspace_bytes(0, min(cap, n * 6)) < lib.ts_shard_link_workspace_bytes(n * 6, min(cap, n * 6)) \
            or n * 6 < 4096
    assert lib.ts_shard_exchange_workspace_bytes(-1, 2) < 0 and lib.ts_shard_exchange_workspace_bytes(10, 17) < 0
    # arenas: the exchange adds its four pieces at the end; without it they shrink to place holders
    with_x, size_x = sharded.arena_layout(1000, 5000, 6, 1, 4)
    without, size_0 = sharded.arena_layout(1000, 5000, 6, 1, 4, exchange=False)
    assert with_x[:10] == without[:10] and len(with_x) == 14 and size_x > size_0
    v = _lib.TsShardView()
    v.rank, v.n_shards = 0, 2
    assert lib.ts_shard_count(None, C.byref(v), 100, 6, 10, 1, None) == -1            # no context
    ctx_stub = C.c_void_p(1)              # never dereferenced: the view is checked first
    v.link_blocks_cap, v.in_cap = 8, 600
    assert lib.ts_shard_count(ctx_stub, C.byref(v), 32 * 2 * 8 + 1, 6, 10, 1, None) == -1   # 9 blocks per rank > 8
    assert b"link_blocks_cap" in lib.ts_last_error()
    v.link_blocks_cap = 12                # not a multiple of 8
    assert lib.ts_shard_send(ctx_stub, C.byref(v), 100, 6, 10, 1, None, 0, None) == -1
    v.link_blocks_cap = 16
    assert lib.ts_shard_send(ctx_stub, C.byref(v), 100, 6, 10, 1, None, 0, None) == -1      # arrays not mapped
    assert b"not mapped" in lib.ts_last_error()
    v.n_shards = 3
    assert lib.ts_shard_count(ctx_stub, C.byref(v), 100, 6, 10, 1, None) == -1
    assert b"power of two" in lib.ts_last_error()
