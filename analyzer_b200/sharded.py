"""One ladder over the GPUs of a node (BASELINE.json configs[4]),
one process per GPU, rows pushed to where they are read (include/tsrate.h,
ts_shard_view).

Layout: blocks of 32 consecutive matches go round-robin to the ranks, so rank g
holds ids / meta / succ of its own 1/G of the stream only.  Player p is owned by
rank p % world (local row p // world): the owner's table shard holds the pair
before the first and after the last match.  Every rank has a mailbox of versioned
rows; a rated match stores each new row into the mailbox of the rank that rates
that player's next match, so a match polls only its own GPU's memory.  The links
(succ) are computed by the players' owners and stored straight into the slice of
the rank that needs them.  torch.distributed exchanges the 64-byte IPC handles
once and separates the phases (reset | link | rate |); there is no collective on
the data path.

emulate=True: one process plays all `world` ranks on one GPU (every pointer
local, one launch of the rating kernel for all ranks) -- the same kernels and
index arithmetic, for single-GPU boxes and tests.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from . import _lib as L

ROW_BYTES = 32
BLOCK = L.SHARD_BLOCK


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


def match_rank(m, world):
    """Rank that rates match m of a window."""
    return (m // BLOCK) % world


def local_index(m, world):
    """Index of match m in the slice of match_rank(m, world)."""
    return (m // BLOCK // world) * BLOCK + m % BLOCK


def global_index(lm, rank, world):
    """Inverse of (match_rank, local_index)."""
    return ((lm // BLOCK) * world + rank) * BLOCK + lm % BLOCK


def local_matches(n_matches, rank, world):
    """Matches of an n_matches window that live on `rank` (= ts_shard_local_matches)."""
    blocks = (n_matches + BLOCK - 1) // BLOCK
    if blocks <= rank:
        return 0
    mine = (blocks - rank + world - 1) // world
    last_is_mine = (blocks - 1) % world == rank
    return (mine - 1) * BLOCK + (n_matches - (blocks - 1) * BLOCK if last_is_mine else BLOCK)


def take_slice(full, rank, world):
    """Rows of a full [N, ...] stream tensor that live on `rank`, in local order."""
    n = full.shape[0]
    step = BLOCK * world * 32768                     # bounded index temporaries
    parts = []
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        m = torch.arange(lo, hi, device=full.device)
        parts.append(full[lo:hi][((m >> 5) % world) == rank])
    return torch.cat(parts) if parts else full[:0]


def exchange_handles(handle, group=None):
    """All ranks' IPC handles (bytes), indexed by rank."""
    world = dist.get_world_size(group)
    out = [None] * world
    dist.all_gather_object(out, bytes(handle), group=group)
    return out


def _align(x):
    return (x + 255) & ~255


def inbox_capacity(n_matches_cap, slots, world):
    """Link messages one sender may leave with one receiver (ts_shard_inbox_cap)."""
    return L.check(L.lib().ts_shard_inbox_cap(n_matches_cap, slots, world))


def ring_capacity(n_players, n_total, slots, world):
    """Row messages one sender may leave with one receiver over a whole ladder of n_total matches:
    its share of the rows that change GPU, plus the last rows going home, with room for skew."""
    if world == 1:
        return 1
    return int((n_total * slots + n_players) / (world * world) * 1.3) + (1 << 16)


def link_blocks_capacity(n_matches_cap, world):
    """Blocks of 32 matches per rank that the link exchange's count arrays are laid out for."""
    blocks = (n_matches_cap + BLOCK - 1) // BLOCK
    return max(8, ((blocks + world - 1) // world + 7) & ~7)


def owned_capacity(n_matches_cap, slots, world):
    """(player, slot) pairs one rank can be sent in a window: its share of the slots, with room
    for uneven ownership (all of them on one GPU)."""
    np_ = n_matches_cap * slots
    return max(1, np_) if world == 1 else min(max(1, np_), int(np_ / world * 1.25) + (1 << 16))


def arena_layout(n_players, n_matches_cap, slots, rank, world, n_total=None, exchange=True):
    """Byte offsets of (rows, table, ids, meta, succ, inbox, inbox counts, ring, ring finals, ring
    tails, link counts, received keys, received slot indices, received count) in rank's arena,
    and its size.  Every rank computes every rank's layout: peers' pointers are base + offset."""
    nl = local_matches(n_matches_cap, rank, world)
    ring = ring_capacity(n_players, n_total if n_total is not None else n_matches_cap, slots, world)
    owned = owned_capacity(n_matches_cap, slots, world) if exchange else 1
    counts = world * world * link_blocks_capacity(n_matches_cap, world) * 2 if exchange else 1
    sizes = (n_players * ROW_BYTES, max(1, shard_rows(n_players, rank, world)) * 16, max(1, nl) * slots * 4,
             max(1, nl) * 4, max(1, nl) * slots * 4, world * inbox_capacity(n_matches_cap, slots, world) * 8, 1024,
             world * ring * 32, 128, 256, counts, owned * 4, owned * 4, 256)
    offs, o = [], 0
    for sz in sizes:
        offs.append(o)
        o += _align(sz)
    return offs, o


class _Raw(object):
    """Device memory that torch did not allocate, for torch.as_tensor."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class ShardedLadder(object):
    def __init__(self, rater, n_players, n_matches_cap, slots, rank=0, world=1, group=None, emulate=False, n_total=None,
                 exchange=True):
        """n_matches_cap: the largest window; n_total: matches of the whole ladder when it is rated in
        several windows (the message rings are sized for the run; default: one window).
        exchange: links through ts_shard_count / ts_shard_send (every rank reads its own slice only);
        False: every rank reads the whole stream (ts_shard_link alone)."""
        if world & (world - 1) or not 1 <= world <= L.SHARD_MAX:
            raise ValueError("world size must be a power of two <= %d" % L.SHARD_MAX)
        if n_matches_cap * slots >= 0xFFFF0000:
            raise ValueError("a window holds fewer than 2^32 - 2^16 slot entries")
        self.rater, self.n_players, self.cap, self.slots = rater, n_players, n_matches_cap, slots
        self.rank, self.world, self.group, self.emulate = rank, world, group, emulate
        self.lib = L.lib()
        self.device = rater.device.index or 0
        dev = torch.device("cuda", self.device)
        self.n_total = n_total if n_total is not None else n_matches_cap
        self.exchange = bool(exchange) and world > 1
        self.layouts = [arena_layout(n_players, n_matches_cap, slots, g, world, self.n_total, self.exchange) for g in range(world)]
        self.bases = [0] * world
        self.opened, self.own = [], None
        self._keep = []
        if emulate:
            for g in range(world):
                t = torch.empty(self.layouts[g][1], dtype=torch.uint8, device=dev)
                self._keep.append(t)
                self.bases[g] = t.data_ptr()
            self.played = list(range(world))
        else:
            own = C.c_void_p()
            handle = (C.c_char * 64)()
            L.check(self.lib.ts_shard_alloc(self.device, self.layouts[rank][1], C.byref(own), handle))
            self.own = own
            self.bases[rank] = own.value
            if world > 1:
                for r, h in enumerate(exchange_handles(handle.raw, group)):
                    if r == rank:
                        continue
                    p = C.c_void_p()
                    L.check(self.lib.ts_shard_open(self.device, C.create_string_buffer(h, 64), C.byref(p)))
                    self.bases[r] = p.value
                    self.opened.append(p)
            self._keep.append(torch.as_tensor(_Raw(own.value, self.layouts[rank][1]), device=dev))
            self.played = [rank]
        # torch views of the arenas this process fills and reads
        self.ids, self.meta, self.table, self.succ = {}, {}, {}, {}
        for i, g in enumerate(self.played):
            arena = self._keep[i]
            offs, _ = self.layouts[g]
            nl = local_matches(n_matches_cap, g, world)
            rows = shard_rows(n_players, g, world)
            self.table[g] = arena[offs[1]:offs[1] + rows * 16].view(torch.float64).view(rows, 2)
            self.ids[g] = arena[offs[2]:offs[2] + nl * slots * 4].view(torch.int32).view(nl, slots)
            self.meta[g] = arena[offs[3]:offs[3] + nl * 4].view(torch.int32)
            self.succ[g] = arena[offs[4]:offs[4] + nl * slots * 4].view(torch.int32).view(nl, slots)
        self._ws = None
        self._gbuf = None
        self._flag = None
        self.epoch = 0

    def view(self, rank=None):
        v = L.TsShardView()
        v.rank = self.rank if rank is None else rank
        v.n_shards = self.world
        for g in range(self.world):
            offs, _ = self.layouts[g]
            v.rows[g], v.table[g] = self.bases[g] + offs[0], self.bases[g] + offs[1]
            v.ids[g], v.meta[g], v.succ[g] = self.bases[g] + offs[2], self.bases[g] + offs[3], self.bases[g] + offs[4]
            v.inbox[g], v.inbox_count[g] = self.bases[g] + offs[5], self.bases[g] + offs[6]
            v.ring[g], v.ring_final[g], v.ring_tail[g] = self.bases[g] + offs[7], self.bases[g] + offs[8], self.bases[g] + offs[9]
        v.inbox_cap = inbox_capacity(self.cap, self.slots, self.world)
        v.ring_cap = ring_capacity(self.n_players, self.n_total, self.slots, self.world)
        v.ring_epoch = self.epoch
        if self.exchange:
            for g in range(self.world):
                offs, _ = self.layouts[g]
                v.link_counts[g], v.in_keys[g], v.in_vals[g] = self.bases[g] + offs[10], self.bases[g] + offs[11], self.bases[g] + offs[12]
                v.in_count[g] = self.bases[g] + offs[13]
            v.in_cap = owned_capacity(self.cap, self.slots, self.world)
            v.link_blocks_cap = link_blocks_capacity(self.cap, self.world)
        return v

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def barrier(self, host=True):
        """All ranks' earlier work on their streams is done before any rank's later work starts.
        host=False with an NCCL group: a one-word all-reduce queued on the stream -- the GPUs wait for
        each other, the hosts do not (no synchronize, the next launches are queued behind it)."""
        if self.emulate or self.world == 1:
            if host:
                torch.cuda.synchronize(self.device)
            return
        if not host and dist.get_backend(self.group) == "nccl":
            if self._flag is None:
                self._flag = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", self.device))
            dist.all_reduce(self._flag, group=self.group)
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def set_stream(self, ids_full, meta_full):
        """Take this rank's blocks (all ranks' when emulating) out of a full stream on this GPU."""
        for g in self.played:
            nl = local_matches(ids_full.shape[0], g, self.world)
            self.ids[g][:nl].copy_(take_slice(ids_full, g, self.world))
            self.meta[g][:nl].copy_(take_slice(meta_full, g, self.world))

    def set_table(self, full):
        """Own shard(s) of a full [U, 2] table on this GPU."""
        for g in self.played:
            self.table[g].copy_(full.reshape(-1, 2)[g::self.world])

    def reset(self):
        self.epoch = self.epoch + 1 if self.epoch < 255 else 1      # the marker of this run's messages; 1 = rings cleared
        with torch.cuda.device(self.device):
            for g in self.played:
                v = self.view(g)
                L.check(self.lib.ts_shard_reset(self.rater._ctx, C.byref(v), self.n_players, self._stream()))

    def link(self, n_matches, mode=1, seq_base=0, owned_cap=None, gather=True):
        """succ slices and first rows of the window, for the players this rank owns."""
        if self.exchange:
            return self._link_exchange(n_matches, mode, seq_base)
        np_ = n_matches * self.slots
        if owned_cap is None:       # a rank owns ~1/world of the slots; skewed ladders may need more
            owned_cap = owned_capacity(n_matches, self.slots, self.world)
        self._workspace(L.check(self.lib.ts_shard_link_workspace_bytes(np_, owned_cap)))
        gathered = self._gather_stream(n_matches) if gather else None
        with torch.cuda.device(self.device):
            for g in self.played:
                v = self.view(g)
                if gathered is not None:         # every rank's slice from the local copy instead of over NVLink
                    all_ids, all_meta = gathered
                    for r in range(self.world):
                        v.ids[r], v.meta[r] = all_ids[r].data_ptr(), all_meta[r].data_ptr()
                L.check(self.lib.ts_shard_link(self.rater._ctx, C.byref(v), n_matches, self.slots, self.n_players, mode,
                                               seq_base, C.c_void_p(self._ws.data_ptr()), self._ws.numel(), self._stream()))

    def _workspace(self, nbytes):
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=torch.device("cuda", self.device))

    def _link_exchange(self, n_matches, mode, seq_base):
        """count | send | link: every rank reads its own slice, counts per block of 32 matches the
        live slots of every owner's players, and -- once all counts are known everywhere -- stores
        each (player, slot) pair at its stream-ordered place in the owner's memory; the owner sorts
        what it received.  A barrier across the ranks at every '|'."""
        in_cap = owned_capacity(self.cap, self.slots, self.world)
        self._workspace(max(L.check(self.lib.ts_shard_link_workspace_bytes(0, min(in_cap, n_matches * self.slots))),
                            L.check(self.lib.ts_shard_exchange_workspace_bytes(n_matches, self.world))))
        ws, nws = C.c_void_p(self._ws.data_ptr()), self._ws.numel()
        ctx = self.rater._ctx
        with torch.cuda.device(self.device):
            for g in self.played:
                v = self.view(g)
                L.check(self.lib.ts_shard_count(ctx, C.byref(v), n_matches, self.slots, self.n_players, mode, self._stream()))
            self.barrier(host=False)
            for g in self.played:
                v = self.view(g)
                L.check(self.lib.ts_shard_send(ctx, C.byref(v), n_matches, self.slots, self.n_players, mode, ws, nws, self._stream()))
            self.barrier(host=False)
            for g in self.played:
                v = self.view(g)
                L.check(self.lib.ts_shard_link(ctx, C.byref(v), n_matches, self.slots, self.n_players, mode, seq_base, ws, nws,
                                               self._stream()))

    def _gather_stream(self, n_matches):
        """The id stream is the one thing every rank must see whole (it keeps the slots of its own
        players): one NCCL all-gather of the slices -- a bulk transfer at link speed -- instead of
        every rank reading its peers' slices through NVLink loads (measured: 1.2 GB of peer loads take
        6 ms, 200 GB/s).  Only with an NCCL group; otherwise ts_shard_link reads the peers' memory."""
        if self.emulate or self.world == 1 or dist.get_backend(self.group) != "nccl":
            return None
        dev = torch.device("cuda", self.device)
        nl_max = local_matches(n_matches, 0, self.world)
        nl = local_matches(n_matches, self.rank, self.world)
        if self._gbuf is None or self._gbuf[0].shape[1] < nl_max:
            self._gbuf = None
            self._gbuf = (torch.empty((self.world, nl_max, self.slots), dtype=torch.int32, device=dev),
                          torch.empty((self.world, nl_max), dtype=torch.int32, device=dev),
                          torch.zeros((nl_max, self.slots), dtype=torch.int32, device=dev),
                          torch.zeros((nl_max,), dtype=torch.int32, device=dev))
        all_ids, all_meta, my_ids, my_meta = self._gbuf
        cap = all_ids.shape[1]
        my_ids[:nl].copy_(self.ids[self.rank][:nl])
        my_meta[:nl].copy_(self.meta[self.rank][:nl])
        dist.all_gather_into_tensor(all_ids.view(-1), my_ids.view(-1), group=self.group)
        dist.all_gather_into_tensor(all_meta.view(-1), my_meta.view(-1), group=self.group)
        return all_ids, all_meta

    def rate(self, n_matches, mode=1, seq_base=0, outputs=None, lanes=0):
        out = L.TsOutputs()
        if outputs:
            for k in ("quality", "out_queue", "status"):
                t = outputs.get(k)
                setattr(out, k, t.data_ptr() if t is not None else None)
        v = self.view()
        with torch.cuda.device(self.device):
            L.check(self.lib.ts_shard_rate(self.rater._ctx, C.byref(v), n_matches, self.slots, self.n_players, mode,
                                           seq_base, C.byref(out), lanes, 1 if self.emulate else 0, self._stream()))

    def status(self):
        with torch.cuda.device(self.device):
            return L.check(self.lib.ts_shard_status(self.rater._ctx, self._stream()))

    def check(self):
        st = self.status()
        if not self.emulate and self.world > 1:      # every rank sees every rank's flags: all raise together
            t = torch.tensor([st], dtype=torch.int32, device=torch.device("cuda", self.device))
            parts = [torch.zeros_like(t) for _ in range(self.world)]
            dist.all_gather(parts, t, group=self.group)
            for x in parts:
                st |= int(x)
        if st & L.SHARD_OVERFLOW:
            raise L.TsError("ts_shard_link: workspace or inbox too small for the slots this rank owns (skewed ladder: "
                            "pass a larger owned_cap / n_matches_cap)")
        if st & L.SHARD_DUPLICATE:
            raise L.TsError("sharded ladder: a player is listed twice in one match")
        if st & L.SHARD_STALLED:
            raise L.TsError("sharded ladder: a match's rows never arrived (links and stream do not belong together)")

    def run(self, n_matches, mode=1, seq_base=0, outputs=None, lanes=0, first=True):
        """One window: (reset |) link | rate |, tables updated in place in the owners' shards."""
        if first:
            self.reset()
        self.barrier()
        self.link(n_matches, mode, seq_base)
        self.check()                   # synchronises: duplicates / overflow must not reach the rating kernel
        self.barrier()
        self.rate(n_matches, mode, seq_base, outputs, lanes)
        self.barrier()
        self.check()

    def close(self):
        self.ids = self.meta = self.table = self.succ = {}
        self._keep = []
        self._gbuf = self._ws = None
        for p in self.opened:
            self.lib.ts_shard_close(p)
        self.opened = []
        if self.own:
            torch.cuda.synchronize(self.device)
            self.lib.ts_shard_free(self.own)
            self.own = None
