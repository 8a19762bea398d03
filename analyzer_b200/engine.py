"""Host side of the B200 rating path: environment, HBM player table, schedule
and rate calls over the C ABI (include/tsrate.h).  torch is used for device
memory and streams only; every kernel is in libtsrate.so.

Mirrors, in columnar form, what rater.py:30-37 (env), worker.py:176-192 (the
chronological per-match loop) and rater.py:69-169 (rate_match) do per object.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import _lib as L

N_MODES = 6
MODE_NAMES = ("casual", "ranked", "blitz", "br", "5v5_casual", "5v5_ranked")   # rater.py:70-82


# ---------------------------------------------------------------------------
# draw margin, the way the trueskill package computes it:
# ppf(y) = -sqrt(2) * erfcinv(2y), erfcinv = rational guess + 2 Newton steps
# ---------------------------------------------------------------------------
def _erfcinv(y):
    if y >= 2:
        return -100.0
    if y <= 0:
        return 100.0
    lower = y < 1
    if not lower:
        y = 2 - y
    t = math.sqrt(-2 * math.log(y / 2.0))
    x = -0.70711 * ((2.30753 + t * 0.27061) / (1.0 + t * (0.99229 + t * 0.04481)) - t)
    for _ in range(2):
        err = math.erfc(x) - y
        x += err / (1.12837916709551257 * math.exp(-(x ** 2)) - x * err)
    return x if lower else -x


def draw_margin(draw_probability, n_players, beta):
    """calc_draw_margin(draw_probability, size, env); exactly 0 at
    draw_probability == 0, where the kernel uses the analytic draw limit
    (with the reference's 50-digit backend ppf(0.5) is ~1e-34, not 0)."""
    if draw_probability == 0:
        return 0.0
    return -math.sqrt(2) * _erfcinv(draw_probability + 1.0) * math.sqrt(n_players) * beta


class Env(object):
    """rater.py:30-37 -- beta, tau, draw_probability of the TrueSkill environment."""

    def __init__(self, beta=10.0 / 30 * 3000, tau=1000 / 100.0, draw_probability=0.0,
                 mu=1500, sigma=1000, zero_draw="limit"):
        """zero_draw: what a draw (both winner flags equal) does at draw_probability = 0 -- "limit":
        the analytic limit v = -t, w = 1 (what the 50-digit backend was found to land on; unpinned by
        the reference); "raise": FloatingPointError for the batch, like the package's float backend."""
        assert zero_draw in ("limit", "raise")
        self.zero_draw = zero_draw
        self.mu, self.sigma = mu, sigma          # unused by the rating path (rater.py:121,132)
        self.beta, self.tau, self.draw_probability = float(beta), float(tau), float(draw_probability)

    def params(self):
        p = L.TsParams()
        p.beta, p.tau = self.beta, self.tau
        for n in range(L.MAX_SLOTS + 1):
            p.draw_margin[n] = draw_margin(self.draw_probability, n, self.beta) if n else 0.0
        p.draw_is_zero = (2 if self.zero_draw == "raise" else 1) if self.draw_probability == 0 else 0
        return p


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class PlayerTable(object):
    """HBM-resident player table: float64 [n_players, row_pairs, 2], NaN = NULL.

    layout "faithful": 8 pairs per row = one 128-byte line; col 0 shared
    (player.trueskill_mu/_sigma), cols 1..6 the queue ratings in MODE_NAMES
    order.  layout "ladder": one pair per row, no shared rating -- the single
    (mu, sigma) ladder of BASELINE.json.  layout "pair": shared + one queue.
    """

    def __init__(self, data, shared_col, mode_cols, seed=None):
        assert data.dtype == torch.float64 and data.dim() == 3 and data.shape[2] == 2
        assert data.is_cuda and data.is_contiguous()
        self.data, self.seed = data, seed
        self.shared_col = int(shared_col)
        self.mode_cols = list(mode_cols) + [-1] * (16 - len(mode_cols))
        if seed is not None:
            assert seed.dtype == torch.float64 and seed.shape == (data.shape[0], 2) and seed.is_cuda

    @property
    def n_players(self):
        return self.data.shape[0]

    @classmethod
    def faithful(cls, n_players, device, seed=None):
        data = torch.full((n_players, 8, 2), float("nan"), dtype=torch.float64, device=device)
        return cls(data, 0, [1, 2, 3, 4, 5, 6], seed)

    @classmethod
    def ladder(cls, data, mode=1, seed=None):
        cols = [-1] * N_MODES
        cols[mode] = 0
        return cls(data.view(-1, 1, 2), -1, cols, seed)

    @property
    def multi_queue(self):
        """Shared pair + more than one rated queue column (a batch that mixes game modes)."""
        return self.shared_col >= 0 and len(set(c for c in self.mode_cols if c >= 0)) > 1

    def mode_col_array(self):
        return (C.c_int8 * 16)(*self.mode_cols)

    def desc(self):
        t = L.TsTable()
        t.data = self.data.data_ptr()
        t.seed = self.seed.data_ptr() if self.seed is not None else None
        t.n_players = self.n_players
        t.row_pairs = self.data.shape[1]
        t.shared_col = self.shared_col
        for i in range(16):
            t.mode_col[i] = self.mode_cols[i]
        return t


class PackedIds(object):
    """The id column of a batch at the width it needs: ids [N, P] as a little-endian bit
    stream of `bits` bits each (all ones = empty slot), what ts_unpack_ids widens on the
    device.  `packed` is a uint8 numpy array or (pinned) torch tensor."""

    def __init__(self, packed, bits, n_matches, slots):
        self.packed, self.bits, self.shape = packed, int(bits), (int(n_matches), int(slots))

    @property
    def nbytes(self):
        return int(self.packed.numel() if isinstance(self.packed, torch.Tensor) else self.packed.size)


def id_bits(n_players):
    """Bits per packed id: ids 0 .. n_players - 1 and the all-ones value for an empty slot."""
    return max(1, int(n_players).bit_length())


def pack_ids(ids, n_players, pin=False):
    """Host-side packing of a uint32 [N, P] id column (EMPTY_SLOT allowed) for rate_batch."""
    ids = np.ascontiguousarray(ids).view(np.uint32)
    n, p = ids.shape
    bits = id_bits(n_players)
    flat = ids.reshape(-1)
    out = np.empty((flat.size * bits + 7) // 8, dtype=np.uint8)
    step = 1 << 22                                   # ids per piece: a multiple of 8, so pieces end on byte boundaries
    for lo in range(0, flat.size, step):
        x = flat[lo:lo + step].copy()
        x[x == L.EMPTY_SLOT] = (1 << bits) - 1
        b = np.unpackbits(x.view(np.uint8).reshape(-1, 4), axis=1, bitorder="little")[:, :bits]
        piece = np.packbits(b.reshape(-1), bitorder="little")
        o = lo * bits // 8
        out[o:o + piece.size] = piece
    packed = torch.from_numpy(out)
    return PackedIds(packed.pin_memory() if pin else packed, bits, n, p)


class Schedule(object):
    """Result of Rater.schedule: device tensors in tsrate.h's terms."""

    def __init__(self, n_matches, slots):
        self.n_matches, self.slots = n_matches, slots
        self.round = self.order = self.round_offsets = None
        self.ids_sorted = self.meta_sorted = self.info = None
        self._n_rounds = None

    @property
    def n_rounds(self):
        """Number of rounds; the first read waits for the schedule (ts_schedule itself does not)."""
        if self._n_rounds is None:
            info = self.info.cpu().numpy().view(np.uint32)
            if info[2]:
                raise L.TsError("ts_schedule: %d rounds need offsets_cap >= %d" % (info[0], info[0] + 2))
            self._n_rounds = int(info[0])
        return self._n_rounds


class Rater(object):
    """One context per device (ts_ctx)."""

    def __init__(self, env=None, device=0):
        if not torch.cuda.is_available():
            raise RuntimeError("analyzer_b200 needs a CUDA device (B200); there is no CPU fallback")
        self.env = env or Env()
        self.device = torch.device("cuda", device) if not isinstance(device, torch.device) else device
        self._lib = L.lib()
        self._ctx = C.c_void_p()
        prm = self.env.params()
        L.check(self._lib.ts_ctx_create(C.byref(prm), self.device.index or 0, C.byref(self._ctx)))
        self._ws = None
        self._vt = None
        self._copy_stream = None
        self._slot_rows = self._slot_succ = None
        self._slot_epoch = 0

    def close(self):
        if self._ctx:
            self._lib.ts_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def sm_count(self):
        return self._lib.ts_sm_count(self._ctx)

    def _workspace(self, nbytes):
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=self.device)
        return self._ws

    def schedule(self, ids, meta, n_players, mode_cols, offsets_cap=None):
        """ids: uint32-as-int32 [N, P] device tensor (torch has no uint32 math;
        the bits are what matter), meta: [N].  Returns a Schedule."""
        assert ids.is_cuda and ids.is_contiguous() and meta.is_contiguous()
        assert ids.element_size() == 4 and meta.element_size() == 4
        n, p = ids.shape
        sch = Schedule(n, p)
        dev = ids.device
        cap = max(3, int(offsets_cap if offsets_cap is not None else n + 2))
        sch.round = torch.empty(n, dtype=torch.int32, device=dev)
        sch.order = torch.empty(n, dtype=torch.int32, device=dev)
        sch.round_offsets = torch.empty(cap, dtype=torch.int32, device=dev)
        sch.ids_sorted = torch.empty_like(ids)
        sch.meta_sorted = torch.empty_like(meta)
        sch.info = torch.zeros(4, dtype=torch.int32, device=dev)
        nbytes = self._lib.ts_schedule_workspace_bytes(n, p)
        L.check(nbytes)
        ws = self._workspace(nbytes)
        cols = list(mode_cols) + [-1] * (16 - len(mode_cols))
        with torch.cuda.device(dev):
            # n_rounds_host = NULL: nothing here waits for the GPU; Schedule.n_rounds reads info when asked
            L.check(self._lib.ts_schedule(
                self._ctx, _ptr(ids), _ptr(meta), n, p, int(n_players), (C.c_int8 * 16)(*cols),
                _ptr(sch.round), _ptr(sch.order), _ptr(sch.round_offsets), cap,
                _ptr(sch.ids_sorted), _ptr(sch.meta_sorted), _ptr(sch.info),
                _ptr(ws), ws.numel(), None, _stream_ptr()))
        return sch

    def rate(self, table, sch, outputs=None, lanes_per_match=0):
        """Run every round of `sch` against `table` (in place).  outputs: dict
        with any of quality/out_shared/out_delta/out_queue/status device tensors."""
        out = L.TsOutputs()
        if outputs:
            for k in ("quality", "out_shared", "out_delta", "out_queue", "status"):
                t = outputs.get(k)
                setattr(out, k, t.data_ptr() if t is not None else None)
        desc = table.desc()
        with torch.cuda.device(table.data.device):
            L.check(self._lib.ts_rate(
                self._ctx, C.byref(desc), _ptr(sch.ids_sorted), _ptr(sch.meta_sorted), _ptr(sch.order),
                _ptr(sch.round_offsets), _ptr(sch.info), sch.n_matches, sch.slots, C.byref(out),
                int(lanes_per_match), _stream_ptr()))

    def link_pred(self, ids, meta, n_players, mode_cols, seq_base=0, last=None, out=None):
        """First half of the scheduler only: pred[N, P] (stream-wide number of the
        previous rated match of each slot's player) for the streaming rating
        kernel.  Windows: seq_base = matches in earlier windows, last = int32
        [n_players] device tensor of carried chain tails (updated in place)."""
        assert ids.is_cuda and ids.is_contiguous() and meta.is_contiguous()
        n, p = ids.shape
        pred = out if out is not None else torch.empty_like(ids)
        assert pred.shape == ids.shape and pred.is_contiguous()
        nbytes = L.check(self._lib.ts_link_workspace_bytes(n, p))
        ws = self._workspace(nbytes)
        cols = list(mode_cols) + [-1] * (16 - len(mode_cols))
        with torch.cuda.device(ids.device):
            L.check(self._lib.ts_link_pred(self._ctx, _ptr(ids), _ptr(meta), n, p, int(n_players),
                                           (C.c_int8 * 16)(*cols), int(seq_base), _ptr(last), _ptr(pred), _ptr(ws),
                                           ws.numel(), _stream_ptr()))
        return pred

    def link_pred_windows(self, ids, meta, n_players, mode_cols, window=None):
        """pred for a long stream, linked window by window with carried chain tails:
        same result as one link_pred call, but the scatter of every window stays
        inside L2 (a single pass over more than ~128 Mi slot entries read-modify-
        writes every sector) and the sort workspace is that of one window."""
        n, p = ids.shape
        if window is None:
            window = max(1, (60 << 20) // p)          # ~60 Mi slot entries per window
        if n <= window:
            return self.link_pred(ids, meta, n_players, mode_cols)
        pred = torch.empty_like(ids)
        last = torch.full((n_players,), -1, dtype=torch.int32, device=ids.device)   # TS_NONE
        for lo in range(0, n, window):
            hi = min(n, lo + window)
            self.link_pred(ids[lo:hi], meta[lo:hi], n_players, mode_cols, seq_base=lo, last=last, out=pred[lo:hi])
        return pred

    def rate_stream(self, table, ids, meta, pred, outputs=None, lanes_per_match=0, seq_base=0,
                    flags=L.STREAM_LOAD | L.STREAM_STORE):
        """Dataflow rating in stream order: single-pair ladders, or the shared pair
        plus any number of rated queue columns (mixed game modes)."""
        out = L.TsOutputs()
        if outputs:
            for k in ("quality", "out_shared", "out_delta", "out_queue", "status"):
                t = outputs.get(k)
                setattr(out, k, t.data_ptr() if t is not None else None)
        desc = table.desc()
        n, p = ids.shape
        nbytes = L.check(self._lib.ts_rate_stream_table_workspace_bytes(C.byref(desc)))
        if self._vt is None or self._vt.numel() < nbytes:
            assert flags & L.STREAM_LOAD, "the versioned rows of an earlier window are gone"
            self._vt = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(table.data.device):
            L.check(self._lib.ts_rate_stream(self._ctx, C.byref(desc), _ptr(ids), _ptr(meta), _ptr(pred), n, p,
                                             int(seq_base), int(flags), C.byref(out), int(lanes_per_match),
                                             _ptr(self._vt), self._vt.numel(), _stream_ptr()))

    def rate_slots(self, table, ids, meta, outputs=None, lanes_per_match=0, window=None):
        """Slot-indexed streaming form (ts_link_succ + ts_rate_slots) for single-pair ladders whose rows
        do not fit L2: every match reads its rows as one contiguous piece, only the writes scatter.
        Same bits as link_pred + rate_stream.  The ladder is rated in windows of `window` matches (about
        60 Mi slot entries by default: the link scatter stays L2-sized); the table carries the state from
        one window to the next, the rows buffer (32 bytes per slot of a window) is reused under a new
        epoch each time and never cleared."""
        n, p = ids.shape
        if window is None:
            window = max(1, (60 << 20) // p)
        window = min(window, max(n, 1))
        rows_bytes = L.check(self._lib.ts_slot_rows_bytes(window, p))
        if self._slot_rows is None or self._slot_rows.numel() < rows_bytes:
            self._slot_rows = None
            self._slot_rows = torch.zeros(max(rows_bytes, 256), dtype=torch.uint8, device=self.device)
        if self._slot_succ is None or self._slot_succ.numel() < window * p:
            self._slot_succ = None
            self._slot_succ = torch.empty(window * p, dtype=torch.int32, device=self.device)
        ws = self._workspace(L.check(self._lib.ts_link_workspace_bytes(window, p)))
        desc = table.desc()
        with torch.cuda.device(table.data.device):
            for lo in range(0, n, window):
                hi = min(n, lo + window)
                self._slot_epoch = self._slot_epoch + 1 if self._slot_epoch < 0xFFFFFFF0 else 1
                out = L.TsOutputs()
                if outputs:
                    for k in ("quality", "out_queue", "status"):
                        t = outputs.get(k)
                        setattr(out, k, t[lo:hi].data_ptr() if t is not None else None)
                a, b = ids[lo:hi], meta[lo:hi]
                L.check(self._lib.ts_link_succ(self._ctx, _ptr(a), _ptr(b), hi - lo, p, C.byref(desc), self._slot_epoch,
                                               _ptr(self._slot_succ), _ptr(self._slot_rows), _ptr(ws), ws.numel(), _stream_ptr()))
                L.check(self._lib.ts_rate_slots(self._ctx, C.byref(desc), _ptr(a), _ptr(b), _ptr(self._slot_succ), hi - lo, p,
                                                self._slot_epoch, C.byref(out), int(lanes_per_match), _ptr(self._slot_rows),
                                                _stream_ptr()))

    def rate_stream_windows(self, table, ids, meta, window, lanes_per_match=0):
        """The whole ladder in windows of `window` matches (worker.py:18-20's
        micro-batches): link + rate per window, chain tails and versioned rows
        carried from window to window.  Same result as one link + rate.
        (Linking window k+1 on a side stream while window k is rated was tried
        and gives nothing: the rating kernel holds the whole register file.)"""
        n, p = ids.shape
        last = torch.full((table.n_players,), -1, dtype=torch.int32, device=ids.device)   # TS_NONE
        for lo in range(0, max(n, 1), window):
            hi = min(n, lo + window)
            flags = (L.STREAM_LOAD if lo == 0 else 0) | (L.STREAM_STORE if hi >= n else 0)
            pred = self.link_pred(ids[lo:hi], meta[lo:hi], table.n_players, table.mode_cols, seq_base=lo, last=last)
            self.rate_stream(table, ids[lo:hi], meta[lo:hi], pred, None, lanes_per_match, seq_base=lo, flags=flags)

    def unpack_ids(self, packed, bits, out):
        """Packed id bytes (device uint8) -> the 32-bit ids the kernels read (out: int32 [n, p], contiguous)."""
        with torch.cuda.device(packed.device):
            L.check(self._lib.ts_unpack_ids(_ptr(packed), int(bits), out.numel(), _ptr(out), _stream_ptr()))
        return out

    def widen_meta(self, meta8, out=None):
        """uint8 meta column (device) -> the 32-bit words the kernels read."""
        n = meta8.numel()
        out = out if out is not None else torch.empty((n,), dtype=torch.int32, device=meta8.device)
        with torch.cuda.device(meta8.device):
            L.check(self._lib.ts_widen_meta(_ptr(meta8), n, _ptr(out), _stream_ptr()))
        return out

    def error_counts(self):
        """(matches that ended in TS_ST_FP_ERROR, in TS_ST_VALUE_ERROR) since the last
        reset_errors -- the two exceptions the trueskill package raises, kept apart on the
        device so that the batch can raise the right one without the status column."""
        if L.check(self._lib.ts_stream_stalled(self._ctx, _stream_ptr())):
            raise L.TsError("ts_rate_stream gave up: the links (pred) do not belong to the stream they were given with")
        if L.check(self._lib.ts_shard_status(self._ctx, _stream_ptr())) & L.SHARD_DUPLICATE:
            raise L.TsError("a player is listed twice in one match: not supported by the slot-indexed / sharded paths")
        fp, val = C.c_int64(0), C.c_int64(0)
        L.check(self._lib.ts_error_counts(self._ctx, C.byref(fp), C.byref(val), _stream_ptr()))
        return int(fp.value), int(val.value)

    def error_count(self):
        return sum(self.error_counts())

    def reset_errors(self):
        L.check(self._lib.ts_reset_errors(self._ctx, _stream_ptr()))

    def alloc_outputs(self, n, p, device=None):
        """Device buffers for every rate_match output (NaN = not written).  The four float64
        outputs are views of one flat buffer (key "_flat"): one fill, one copy back."""
        dev = device or self.device
        flat = torch.full((n * (1 + 5 * p) + 8,), float("nan"), dtype=torch.float64, device=dev)
        o = 0
        outs = {}
        for key, shape in (("quality", (n,)), ("out_shared", (n, p, 2)), ("out_delta", (n, p)), ("out_queue", (n, p, 2))):
            k = int(np.prod(shape))
            outs[key] = flat[o:o + k].view(shape)
            o = (o + k + 1) & ~1          # the pair outputs are stored as 16-byte double2
        outs["status"] = torch.zeros((n,), dtype=torch.uint8, device=dev)
        outs["_flat"] = flat
        return outs

    # ------------------------------------------------------------------
    def rate_batch(self, ids, meta, table, seed=None, shared_col=0, mode_cols=(1, 2, 3, 4, 5, 6),
                   want_outputs=True, lanes_per_match=0, table_out=None, path="auto", windows=None):
        """The reference-facing batch call with HOST buffers: numpy ids
        [N, P] uint32 (or a PackedIds: the same column at ceil(log2(U + 1)) bits per id, see
        pack_ids -- the id column is most of the PCIe volume), meta [N] uint32 or uint8 (only the low 8 bits are used; bytes
        quarter the column's PCIe volume), table [U, C, 2] float64 (NaN = NULL),
        optional seed [U, 2].  Copies in, schedules, rates, copies back.
        Returns dict(table=..., n_rounds=..., quality=..., out_*=..., status=...).
        Raises FloatingPointError / ValueError like the reference would for the
        batch (worker.py:195-197) -- the returned table is then not produced."""
        dev = self.device

        def host(a, as_i32=False):
            if a is None or isinstance(a, torch.Tensor):
                return a                      # torch CPU tensors (e.g. pinned) pass through
            a = np.ascontiguousarray(a)
            if as_i32 and a.dtype.itemsize == 1:
                return torch.from_numpy(a.view(np.uint8))      # meta shipped as bytes, widened on the device
            return torch.from_numpy(a.view(np.int32) if as_i32 else a)

        packed = ids if isinstance(ids, PackedIds) else None
        if packed is not None:
            ids_t = packed.packed if isinstance(packed.packed, torch.Tensor) else torch.from_numpy(packed.packed)
            n, p = packed.shape
        else:
            ids_t = host(ids, True)
            n, p = ids_t.shape
        meta_t, tab_t, seed_t = host(meta, True), host(table), host(seed)
        meta_bytes = meta_t.element_size() == 1
        # The streaming (dataflow) path wins whenever the versioned rows stay in L2 (small tables,
        # whatever the batch: 0.47 vs 0.58 ms for 500 matches) or chains are deep (several matches
        # per player: latency-bound, barriers hurt); a sparse batch over a table far larger than L2
        # is bandwidth-bound and cheaper round by round (16-byte rows, no versioned copy of the
        # whole table).  Measured crossovers, DESIGN.md 4.
        n_players = tab_t.shape[0]
        row_bytes = 32 if shared_col < 0 else 32 * (2 + len(set(c for c in mode_cols if c >= 0)))
        worth = n > 0 and p in (6, 10) and (n_players * row_bytes <= (64 << 20) or n * p >= 2 * n_players)
        stream = path == "stream" or (path == "auto" and shared_col < 0 and tab_t.shape[1] == 1 and worth)
        mode_cols = list(mode_cols)
        if path == "auto" and shared_col >= 0 and worth:
            # shared + queue layout: a batch that rates one queue column only gets the compact
            # 64-byte rows under one version; a batch that mixes game modes the ordinal-tagged rows
            # (a look at the meta column on the host, for batches of the worker's size only: a large
            # batch takes the ordinal-tagged rows, which serve one mode as well as many)
            bits = np.flatnonzero(np.bincount((meta_t.numpy() >> L.META_MODE_SHIFT) & 15, minlength=16)).tolist() \
                if n <= 65_536 else None
            if bits is not None:
                rated = [m for m in bits if m < len(mode_cols) and m < N_MODES and mode_cols[m] >= 0]
                if len(rated) == 1:
                    mode_cols = [c if i == rated[0] else -1 for i, c in enumerate(mode_cols)]
            stream = tab_t.shape[1] <= 16 and any(c >= 0 for c in mode_cols)
        if windows is None:
            # The GPU needs about as long for a window as PCIe needs to deliver it, and every
            # window costs ~0.27 ms of fixed GPU time (launches, look-back chains, ramp-up of the
            # rating kernel): T(W) ~ first copy + W * fixed + total work has its minimum near
            # 2.5M matches per window (scripts/e2e_windows.py, scripts/e2e_timeline.py)
            windows = max(1, min(8, n // 2_500_000)) if (stream and not want_outputs) else 1
            if windows >= 3:
                # growing parts (1 : 2 : ... : W): the GPU starts after a tenth of the copy instead of a
                # quarter, and the copies, faster than the GPU work, stay ahead (6.91 -> 6.58 ms at 10M matches)
                windows = list(range(1, windows + 1))
        # windows: a count (equal parts) or the parts themselves as fractions of the batch (a short first
        # window lets the GPU start early; the copies are faster than the GPU work, so it is not starved later)
        fractions = None
        if not isinstance(windows, int):
            fractions = [float(f) for f in windows]
            windows = len(fractions)
        self.reset_errors()
        if stream and windows > 1:
            # Pipelined ingest: the match stream is copied window by window on a side
            # stream while the previous window is linked and rated, so the PCIe copy
            # and the GPU work overlap (worker.py:18-20's micro-batches, in one call).
            # Windows are consecutive ranges of the stream; chain tails and the
            # versioned rows carry the state from one to the next.
            cur = torch.cuda.current_stream(dev)
            side = self._copy_stream or torch.cuda.Stream(dev)
            self._copy_stream = side
            ids_d = torch.empty((n, p), dtype=torch.int32, device=dev)
            ids_in = torch.empty((ids_t.numel(),), dtype=torch.uint8, device=dev) if packed is not None else ids_d
            meta_d = torch.empty((n,), dtype=torch.int32, device=dev)
            meta_in = torch.empty((n,), dtype=torch.uint8, device=dev) if meta_bytes else meta_d
            tab_d = tab_t.to(dev, non_blocking=True)
            seed_d = seed_t.to(dev, non_blocking=True) if seed_t is not None else None
            ptab = PlayerTable(tab_d, shared_col, list(mode_cols), seed_d)
            last = torch.full((ptab.n_players,), -1, dtype=torch.int32, device=dev)
            # windows start at multiples of 8 matches: packed id columns then split on byte boundaries
            if fractions is None:
                bounds = [((w * n) // windows) & ~7 for w in range(windows)] + [n]
            else:
                acc, bounds = 0.0, [0]
                for f in fractions[:-1]:
                    acc += f / sum(fractions)
                    bounds.append(max(bounds[-1], min(n, int(acc * n)) & ~7))
                bounds.append(n)
            cut = (lambda m: m * p * packed.bits // 8 if m < n else ids_t.numel()) if packed is not None else (lambda m: m)
            side.wait_stream(cur)
            ready = []
            with torch.cuda.stream(side):
                for w in range(windows):
                    lo, hi = bounds[w], bounds[w + 1]
                    ids_in[cut(lo):cut(hi)].copy_(ids_t[cut(lo):cut(hi)], non_blocking=True)
                    meta_in[lo:hi].copy_(meta_t[lo:hi], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(side)
                    ready.append(ev)
            for w in range(windows):
                lo, hi = bounds[w], bounds[w + 1]
                cur.wait_event(ready[w])
                if packed is not None:
                    self.unpack_ids(ids_in[cut(lo):cut(hi)], packed.bits, ids_d[lo:hi])
                if meta_bytes:
                    self.widen_meta(meta_in[lo:hi], meta_d[lo:hi])
                flags = (L.STREAM_LOAD if w == 0 else 0) | (L.STREAM_STORE if w == windows - 1 else 0)
                pred = self.link_pred(ids_d[lo:hi], meta_d[lo:hi], ptab.n_players, ptab.mode_cols, seq_base=lo, last=last)
                self.rate_stream(ptab, ids_d[lo:hi], meta_d[lo:hi], pred, None, lanes_per_match, seq_base=lo, flags=flags)
            outs = None
            n_rounds = None
        else:
            ids_d = ids_t.to(dev, non_blocking=True)
            if packed is not None:
                ids_d = self.unpack_ids(ids_d, packed.bits, torch.empty((n, p), dtype=torch.int32, device=dev))
            meta_d = meta_t.to(dev, non_blocking=True)
            if meta_bytes:
                meta_d = self.widen_meta(meta_d)
            tab_d = tab_t.to(dev, non_blocking=True)
            seed_d = seed_t.to(dev, non_blocking=True) if seed_t is not None else None
            ptab = PlayerTable(tab_d, shared_col, list(mode_cols), seed_d)
            outs = self.alloc_outputs(n, p) if want_outputs else None
            if stream:
                # single-pair ladders take the streaming (dataflow) path unless told otherwise
                pred = self.link_pred(ids_d, meta_d, ptab.n_players, ptab.mode_cols)
                self.rate_stream(ptab, ids_d, meta_d, pred, outs, lanes_per_match)
                n_rounds = None
            else:
                sch = self.schedule(ids_d, meta_d, ptab.n_players, ptab.mode_cols)
                self.rate(ptab, sch, outs, lanes_per_match)
                n_rounds = sch.n_rounds
        n_fp, n_val = self.error_counts()
        bad = n_fp + n_val
        result = dict(n_rounds=n_rounds, errors=bad)
        if outs is not None:
            flat_h = outs["_flat"].cpu().numpy()
            o = 0
            for k in ("quality", "out_shared", "out_delta", "out_queue"):
                size = outs[k].numel()
                result[k] = flat_h[o:o + size].reshape(tuple(outs[k].shape))
                o = (o + size + 1) & ~1
            result["status"] = outs["status"].cpu().numpy()
        if n_val:      # the reference raises on the first failing match of the batch; either kind fails it
            raise ValueError("sigma**2 should be greater than 0 / player without rating or seed / player id outside "
                             "the table (%d matches)" % n_val)
        if n_fp:
            raise FloatingPointError("%d matches outside the truncated-Gaussian domain" % n_fp)
        if table_out is not None:             # caller-provided (pinned) host tensor
            table_out.copy_(tab_d, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            result["table"] = table_out
        else:
            result["table"] = tab_d.cpu().numpy()
        return result


# ---------------------------------------------------------------------------
# synthetic ladders (BASELINE.json configs) generated in HBM
# ---------------------------------------------------------------------------
def make_gen(seed, n_players, slots=6, mode=1, p_draw=0.0, p_afk=0.0, zipf=False, zipf_q=1000.0):
    g = L.TsGen()
    g.seed, g.n_players, g.slots, g.mode = seed, n_players, slots, mode
    g.draw_thr = min(int(p_draw * 2 ** 32), 2 ** 32 - 1)
    g.afk_thr = min(int(p_afk * 2 ** 32), 2 ** 32 - 1)
    g.zipf = 1 if zipf else 0
    g.zipf_q = zipf_q
    g.zipf_a = zipf_q ** 0.2
    g.zipf_b = (n_players + zipf_q) ** 0.2 - zipf_q ** 0.2
    return g


def gen_matches(gen, first, count, device):
    ids = torch.empty((count, gen.slots), dtype=torch.int32, device=device)
    meta = torch.empty((count,), dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        L.check(L.lib().ts_gen_matches(C.byref(gen), first, count, _ptr(ids), _ptr(meta), _stream_ptr()))
    return ids, meta


def seed_table(vst_points, unknown_sigma, n_players, device, ranked=None, blitz=None, tier=None):
    """get_trueskill_seed (rater.py:42-62) for whole columns on the device.
    ranked / blitz: float64 device tensors with NaN = None; tier: int8.
    Returns (seed [U, 2], bad [U] uint8 -- 1 where the reference raises KeyError)."""
    seed = torch.empty((n_players, 2), dtype=torch.float64, device=device)
    bad = torch.empty((n_players,), dtype=torch.uint8, device=device)
    vst = (C.c_double * 31)(*[float(vst_points[t]) for t in range(-1, 30)])
    with torch.cuda.device(device):
        L.check(L.lib().ts_seed_table(_ptr(ranked), _ptr(blitz), _ptr(tier), n_players, float(unknown_sigma), vst,
                                      _ptr(seed), _ptr(bad), _stream_ptr()))
    return seed, bad


def gen_table(seed, first, count, device):
    t = torch.empty((count, 1, 2), dtype=torch.float64, device=device)
    with torch.cuda.device(device):
        L.check(L.lib().ts_gen_table(seed, first, count, _ptr(t), _stream_ptr()))
    return t
