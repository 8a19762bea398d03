"""One ladder over the GPUs of a node: hash-sharded player table in NVLink peer
memory (BASELINE.json configs[4]), one process per GPU.

Player p lives on rank p % world at local row p // world.  Every rank rates
the warp tiles t with t % world == rank of the same match stream; a match
waits for the versions of its players' rows wherever they live, so there is
no inter-GPU barrier and no collective while rating -- torch.distributed is
used only to exchange the 64-byte IPC handles once and to separate the three
phases (load / rate / read back).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from . import _lib as L

ROW_BYTES = 32


def owner(pid, world):
    return pid % world


def local_row(pid, world):
    return pid // world


def shard_rows(n_players, rank, world):
    """Number of players p < n_players with p % world == rank."""
    return max(0, (n_players - rank + world - 1) // world)


def split_table(table, rank, world):
    """Rows of a full [U, 2] table owned by `rank` (a copy, contiguous)."""
    return table.reshape(-1, 2)[rank::world].contiguous()


def merge_tables(shards):
    """Inverse of split_table over all ranks."""
    world = len(shards)
    total = sum(s.shape[0] for s in shards)
    out = shards[0].new_empty((total, 2))
    for r, s in enumerate(shards):
        out[r::world] = s
    return out


def tiles_of(rank, world, n_matches, tile):
    """Match tiles (start, stop) a rank rates: t % world == rank."""
    n_tiles = (n_matches + tile - 1) // tile
    return [(t * tile, min(n_matches, (t + 1) * tile)) for t in range(rank, n_tiles, world)]


def exchange_handles(handle, group=None):
    """All ranks' IPC handles (bytes), indexed by rank."""
    world = dist.get_world_size(group)
    out = [None] * world
    dist.all_gather_object(out, bytes(handle), group=group)
    return out


class ShardedLadder(object):
    def __init__(self, rater, n_players, rank, world, group=None):
        if world & (world - 1) or not 1 <= world <= 16:
            raise ValueError("world size must be a power of two <= 16")
        self.rater, self.n_players, self.rank, self.world, self.group = rater, n_players, rank, world, group
        self.lib = L.lib()
        self.device = rater.device.index or 0
        self.rows = shard_rows(n_players, rank, world)
        own = C.c_void_p()
        handle = (C.c_char * 64)()
        L.check(self.lib.ts_shard_alloc(self.device, max(self.rows, 1) * ROW_BYTES, C.byref(own), handle))
        self.own = own
        self.ptrs = (C.c_void_p * world)()
        self.ptrs[rank] = own.value
        self.opened = []
        handles = exchange_handles(handle.raw, group)
        for r, h in enumerate(handles):
            if r == rank:
                continue
            p = C.c_void_p()
            buf = C.create_string_buffer(h, 64)
            L.check(self.lib.ts_shard_open(self.device, buf, C.byref(p)))
            self.ptrs[r] = p.value
            self.opened.append(p)

    def _phase(self, phase, table_shard, ids=None, meta=None, pred=None, mode=1, lanes=0):
        n, p = (ids.shape if ids is not None else (0, 6))
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        out = L.TsOutputs()
        with torch.cuda.device(self.device):
            L.check(self.lib.ts_rate_stream_sharded(
                self.rater._ctx, phase, ptr(table_shard), self.n_players, self.rank, self.world, self.ptrs, mode,
                ptr(ids), ptr(meta), ptr(pred), n, p, C.byref(out), lanes,
                C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    def link(self, ids, meta, mode_cols=(-1, 0), data_group=None):
        """Distributed linking: every rank links the players it owns, then one
        element-wise MIN all-reduce assembles pred [N, P] on every rank."""
        n, p = ids.shape
        pred = torch.empty_like(ids)
        nbytes = L.check(self.lib.ts_link_workspace_bytes(n, p))
        ws = self.rater._workspace(nbytes)
        cols = list(mode_cols) + [-1] * (16 - len(mode_cols))
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        with torch.cuda.device(self.device):
            L.check(self.lib.ts_link_pred_sharded(
                self.rater._ctx, ptr(ids), ptr(meta), n, p, self.n_players, (C.c_int8 * 16)(*cols), 0, None,
                ptr(pred), self.rank, self.world, ptr(ws), ws.numel(),
                C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        if self.world > 1:
            dist.all_reduce(pred, op=dist.ReduceOp.MIN, group=data_group if data_group is not None else self.group)
        return pred

    def _barrier(self):
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def rate(self, table_shard, ids, meta, pred, mode=1, lanes=0):
        """table_shard: float64 [rows, 2] on this rank's GPU, updated in place.
        ids / meta / pred: the full arrays of the ladder on this GPU."""
        assert table_shard.is_cuda and table_shard.dtype == torch.float64 and table_shard.shape == (self.rows, 2)
        self._phase(0, table_shard)
        self._barrier()                       # every shard is loaded before anyone polls it
        self._phase(1, table_shard, ids, meta, pred, mode, lanes)
        self._barrier()                       # every rank is done before rows are read back
        self._phase(2, table_shard)
        torch.cuda.synchronize(self.device)

    def close(self):
        for p in self.opened:
            self.lib.ts_shard_close(p)
        self.opened = []
        if self.own:
            self.lib.ts_shard_free(self.own)
            self.own = C.c_void_p()
