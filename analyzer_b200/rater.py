"""Drop-in for the reference's `rater` module (rater.py) on the B200 path.

Same names and behaviour at the module boundary:

    rate_match(match)          rater.py:69-169   mutate one match's object graph
    get_trueskill_seed(player) rater.py:42-62
    vst_points                 rater.py:13-27
    env, TAU, UNKNOWN_PLAYER_SIGMA   rater.py:10-11,30-37

plus the batch forms the GPU wants:

    rate_matches(matches)      worker.py:176-192's loop in ONE call: columnar
                               ingest of the object graphs (shared Player
                               objects keep their identity, like the SQLAlchemy
                               session's identity map), GPU schedule + rate,
                               attribute write-back
    rate_batch(ids, meta, table, ...)   columns in, columns out (engine.Rater)

`worker.py` keeps working unchanged (`rater.rate_match(match)` per match), but
a worker that wants the throughput calls `rater.rate_matches(list_of_matches)`
once per batch instead of looping.

Everything numeric runs in libtsrate.so on the GPU.  There is no CPU fallback:
without the library or a CUDA device these functions raise.
"""
from __future__ import annotations

import ctypes as C
import logging
import os
import sys

import numpy as np

from . import _lib as L
from .engine import Env, Rater, MODE_NAMES

UNKNOWN_PLAYER_SIGMA = int(os.environ.get("UNKNOWN_PLAYER_SIGMA") or 500)   # rater.py:10
TAU = float(os.environ.get("TAU") or 1000 / 100.0)                          # rater.py:11


def _build_vst_points():
    """Tier (-1..29) -> average skill-tier points, piecewise linear in the
    tier with slopes 109 1/11, 50, 66 2/3, 133 1/3 and 200 (rater.py:13-27)."""
    pts = {-1: 1, 0: 1}
    anchor = 0
    for count, slope in ((11, 109 + 1 / 11), (4, 50), (9, 66 + 2 / 3), (3, 133 + 1 / 3), (2, 200)):
        base = pts[anchor] if anchor else 0
        for c in range(1, count + 1):
            pts[anchor + c] = base + slope * (c + 0.5)
        anchor += count
    return pts


vst_points = _build_vst_points()

env = Env(beta=10.0 / 30 * 3000, tau=TAU, draw_probability=0)   # rater.py:30-37

_MODE_INDEX = {name: i for i, name in enumerate(MODE_NAMES)}
_MODE_COLUMN = {name: "trueskill_" + name for name in MODE_NAMES}      # rater.py:70-82

logger = logging.getLogger("__name__")   # sic: the reference names its logger "__name__" (rater.py:177)
if not logger.handlers:
    logger.setLevel(logging.INFO)

    class _InfoFilter(logging.Filter):
        def filter(self, record):
            return record.levelno in (logging.DEBUG, logging.INFO)

    _h1 = logging.StreamHandler(sys.stdout)
    _h1.setLevel(logging.INFO)
    _h1.addFilter(_InfoFilter())
    logger.addHandler(_h1)
    _h2 = logging.StreamHandler(sys.stderr)
    _h2.setLevel(logging.WARNING)
    logger.addHandler(_h2)


def get_trueskill_seed(player):
    """First-touch (mu, sigma) of a player (rater.py:42-62)."""
    best = None
    for pts in (player.rank_points_ranked, player.rank_points_blitz):
        if pts is not None and pts != 0 and (best is None or pts > best):
            best = pts
    if best is not None:
        sigma = UNKNOWN_PLAYER_SIGMA * (2.0 / 3.0)
        return float(best) + sigma, sigma
    sigma = UNKNOWN_PLAYER_SIGMA
    return vst_points[player.skill_tier] + sigma, sigma   # KeyError outside -1..29, like the reference


_raters = {}


def _rater(device=0):
    r = _raters.get(device)
    if r is None or r.env is not env:
        r = Rater(env, device)
        _raters[device] = r
    return r


# ---------------------------------------------------------------------------
# columnar ingest of object graphs (replaces the per-match Python loop)
# ---------------------------------------------------------------------------
class _Batch(object):
    __slots__ = ("ids", "meta", "table", "seed", "players", "slot_parts", "matches", "index", "slots")


def ingest(matches):
    """Object graphs -> columns.  Every distinct Player object gets a row of the
    batch-local table (NaN = attribute is None).  Mirrors the reads of
    rater.py:70-130 without touching the objects.  Written for speed: this and
    _write_back are what bounds the object API (the GPU side is microseconds)."""
    b = _Batch()
    b.matches = matches = list(matches)
    index, players = {}, []
    n = len(matches)
    widest = 3
    for match in matches:
        for roster in match.rosters[:2]:
            k = len(roster.participants)
            if k > widest:
                widest = k
    if widest > 5:
        raise ValueError("rosters of more than 5 participants are not supported by the kernel")
    slots = 6 if widest <= 3 else 10
    half = slots // 2
    empty = L.EMPTY_SLOT
    ids = [empty] * (n * slots)                 # plain lists: element stores into numpy arrays are slow
    meta = [0] * n
    slot_parts = [None] * n
    need_seed = set()
    modes_seen = set()
    mode_index, unsupported, shift = _MODE_INDEX, L.MODE_UNSUPPORTED, L.META_MODE_SHIFT
    for m, match in enumerate(matches):
        mode = mode_index.get(match.game_mode, unsupported)
        bits = mode << shift
        if mode == unsupported:
            meta[m] = bits
            continue
        rosters = match.rosters
        if len(rosters) != 2:
            bits |= L.META_INVALID
        for part in match.participants:          # rater.py:95-100 (stops at the first AFK)
            if part.went_afk == 1:
                bits |= L.META_AFK
                break
        parts = []
        if not bits & (L.META_AFK | L.META_INVALID):
            modes_seen.add(mode)
            base = m * slots
            for t, roster in enumerate(rosters):
                if roster.winner:
                    bits |= (L.META_WIN_A << t)
                s = t * half
                for part in roster.participants:
                    player = part.player[0]
                    key = id(player)
                    row = index.get(key)
                    if row is None:
                        row = index[key] = len(players)
                        players.append(player)
                        if player.trueskill_mu is None:
                            need_seed.add(row)
                    ids[base + s] = row
                    parts.append((s, part, player))
                    s += 1
        meta[m] = bits
        slot_parts[m] = parts
    u = max(len(players), 1)
    table = np.full((u, 8, 2), np.nan)
    seed = np.full((u, 2), np.nan)
    # shared pair of every player; queue pairs only of the modes that are rated in this batch
    sh = [(float(p.trueskill_mu), float(p.trueskill_sigma)) if p.trueskill_mu is not None else (np.nan, np.nan)
          for p in players]
    if players:
        table[:len(players), 0] = sh
    for mi in sorted(modes_seen):
        mu_col, sg_col = _MODE_COLUMN[MODE_NAMES[mi]] + "_mu", _MODE_COLUMN[MODE_NAMES[mi]] + "_sigma"
        col = []
        for p in players:
            mu = getattr(p, mu_col, None)
            col.append((float(mu), float(getattr(p, sg_col))) if mu is not None else (np.nan, np.nan))
        if players:
            table[:len(players), 1 + mi] = col
    for row in need_seed:
        seed[row] = get_trueskill_seed(players[row])
    b.ids = np.array(ids, dtype=np.uint32).reshape(n, slots)
    b.meta = np.array(meta, dtype=np.uint32)
    b.table, b.seed = table, seed
    b.players, b.slot_parts, b.index, b.slots = players, slot_parts, index, slots
    return b


def _write_back(b, res):
    """Columns -> attributes, exactly the stores of rater.py:96-106,141,151-169."""
    status = res["status"].tolist()
    quality = res["quality"].tolist()
    o_sh, o_de, o_qu = res["out_shared"].tolist(), res["out_delta"].tolist(), res["out_queue"].tolist()
    meta = b.meta.tolist()
    info, error = logger.isEnabledFor(logging.INFO), logger.isEnabledFor(logging.ERROR)
    for m, match in enumerate(b.matches):
        st = status[m]
        if st == L.ST_SKIP_MODE:
            if info:
                logger.info("got unsupported game mode %s", match.game_mode)
            continue
        if st == L.ST_AFK or st == L.ST_INVALID:
            if error and (st == L.ST_INVALID or len(match.rosters) != 2):
                logger.error("got an invalid matchup %s", match.api_id)
            if info and meta[m] & L.META_AFK:
                logger.info("got an afk matchup %s", match.api_id)
            match.trueskill_quality = 0
            for part in match.participants:
                part.participant_items[0].any_afk = True
            continue
        if info:
            logger.info("got a valid matchup %s", match.api_id)
        col = _MODE_COLUMN[match.game_mode]
        mu_col, sg_col = col + "_mu", col + "_sigma"
        match.trueskill_quality = quality[m]
        for part in match.participants:
            part.participant_items[0].any_afk = False
        sh_m, de_m, qu_m = o_sh[m], o_de[m], o_qu[m]
        parts = b.slot_parts[m]
        seen = set()
        for s, part, player in parts:
            mu, sg = sh_m[s]
            # a player listed twice in one match (worker_test.py:130-131) has already
            # been updated when the reference reaches the second listing: delta 0
            key = id(player)
            if player.trueskill_mu is None:
                part.trueskill_delta = 0
            else:
                part.trueskill_delta = de_m[s] if key not in seen else 0.0
            seen.add(key)
            player.trueskill_mu = part.trueskill_mu = mu
            player.trueskill_sigma = part.trueskill_sigma = sg
        for s, part, player in parts:
            items = part.participant_items[0]
            mu, sg = qu_m[s]
            setattr(player, mu_col, mu)
            setattr(items, mu_col, mu)
            setattr(player, sg_col, sg)
            setattr(items, sg_col, sg)


# ---------------------------------------------------------------------------
# columnar ingest of the ROW STREAM (what the query of worker.py:176-191 yields,
# one row per participant) and the rows its single commit (worker.py:194) flushes
# -- no per-object, per-attribute Python anywhere
# ---------------------------------------------------------------------------
ROW_FIELDS = ("match", "game_mode", "n_rosters", "roster", "winner", "went_afk", "player", "skill_tier",
              "rank_points_ranked", "rank_points_blitz", "trueskill_mu", "trueskill_sigma", "queue")


class RowBatch(object):
    """Columns of a batch ingested from participant rows, plus what maps results back to rows."""
    __slots__ = ("ids", "meta", "table", "seed", "slots", "match_keys", "player_keys", "row_match", "row_slot",
                 "row_player", "n_rows")


_VST = np.array([vst_points[t] for t in range(-1, 30)])
_VST_C = (C.c_double * 31)(*[float(x) for x in _VST])


def _seed_columns(ranked, blitz, tier):
    """get_trueskill_seed (rater.py:42-62) for whole columns; NaN = None.  A tier outside -1..29
    with no rank points (the reference's KeyError) gives a NaN seed: such a player fails its
    first rated match with a ValueError status instead."""
    best = np.fmax(np.where(ranked == 0, np.nan, ranked), np.where(blitz == 0, np.nan, blitz))     # NaN only if both are
    has = ~np.isnan(best)
    t = np.asarray(tier, dtype=np.int64)
    ok = (t >= -1) & (t <= 29)
    s23 = UNKNOWN_PLAYER_SIGMA * (2.0 / 3.0)
    out = np.empty((len(t), 2))
    out[:, 0] = np.where(has, best + s23, np.where(ok, _VST[np.clip(t, -1, 29) + 1] + UNKNOWN_PLAYER_SIGMA, np.nan))
    out[:, 1] = np.where(has, s23, np.where(ok, float(UNKNOWN_PLAYER_SIGMA), np.nan))
    return out


def _first_appearance(keys):
    """(unique keys in order of first appearance, row -> index, first row of every key)"""
    uk, first, inv = np.unique(keys, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")
    rank = np.empty(len(uk), dtype=np.int64)
    rank[order] = np.arange(len(uk))
    return uk[order], rank[inv], first[order]


def ingest_rows(rows):
    """Participant rows -> columns, by the native host routine (ts_ingest_rows, csrc/ingest.cu); same
    contract and same result as ingest_rows_numpy below, whose docstring lists the columns."""
    lib = L.lib()
    col = lambda k, dt: np.ascontiguousarray(rows[k], dtype=dt)
    keep = [col("match", np.int64), col("player", np.int64), col("game_mode", np.int32), col("n_rosters", np.int32),
            col("roster", np.int32), col("skill_tier", np.int32), col("winner", np.uint8), col("went_afk", np.uint8),
            col("rank_points_ranked", np.float64), col("rank_points_blitz", np.float64), col("trueskill_mu", np.float64),
            col("trueskill_sigma", np.float64), col("queue", np.float64)]
    n = keep[0].shape[0]
    rc_ = L.TsRowColumns()
    rc_.n_rows = n
    for name, arr in zip(("match", "player", "game_mode", "n_rosters", "roster", "skill_tier", "winner", "went_afk",
                          "rank_points_ranked", "rank_points_blitz", "trueskill_mu", "trueskill_sigma", "queue"), keep):
        setattr(rc_, name, arr.ctypes.data)
    cap = max(n, 1)
    match_keys, player_keys = np.empty(cap, np.int64), np.empty(cap, np.int64)
    row_match, row_slot, row_player = np.empty(cap, np.int32), np.empty(cap, np.int32), np.empty(cap, np.int32)
    ids, meta = np.empty(cap * 10, np.uint32), np.empty(cap, np.uint32)
    table, seed = np.empty(cap * 16), np.empty(cap * 2)          # every used row is written in full by the routine
    nm, npl, slots = C.c_int64(0), C.c_int64(0), C.c_int32(6)
    ptr = lambda a: C.c_void_p(a.ctypes.data)
    L.check(lib.ts_ingest_rows(C.byref(rc_), float(UNKNOWN_PLAYER_SIGMA), _VST_C, ptr(match_keys), ptr(player_keys),
                               ptr(row_match), ptr(row_slot), ptr(row_player), ptr(ids), ptr(meta), ptr(table), ptr(seed),
                               C.byref(nm), C.byref(npl), C.byref(slots)))
    nm, npl, slots = nm.value, npl.value, slots.value
    b = RowBatch()
    b.n_rows, b.slots = n, slots
    b.match_keys, b.player_keys = match_keys[:nm], player_keys[:npl]
    b.row_match, b.row_slot, b.row_player = row_match[:n], row_slot[:n], row_player[:n]
    b.ids, b.meta = ids[:nm * slots].reshape(nm, slots), meta[:nm]
    u = max(npl, 1)
    if npl == 0:
        table[:16], seed[:2] = np.nan, np.nan
    b.table, b.seed = table[:u * 16].reshape(u, 8, 2), seed[:u * 2].reshape(u, 2)
    return b


def ingest_rows_numpy(rows):
    """Participant rows -> columns.  `rows`: a mapping (or numpy structured array) of equally long
    columns, one entry per participant in the order the query yields them (matches by created_at,
    worker.py:177):

        match               any integer key of the match (first appearance fixes the stream order)
        game_mode           0..5 in rater.py:70-82's order (casual, ranked, blitz, br, 5v5_casual,
                            5v5_ranked), anything else = unsupported
        n_rosters           len(match.rosters)                     (rater.py:91-93)
        roster              0 / 1: which of the two rosters the participant is on
        winner              that roster's .winner                   (rater.py:144)
        went_afk            participant.went_afk                    (rater.py:95-100)
        player              any integer key of the player (the identity the session's identity map shares)
        skill_tier, rank_points_ranked, rank_points_blitz          (rater.py:42-62; NaN = None)
        trueskill_mu, trueskill_sigma                              the player's shared pair (NaN = NULL)
        queue               float64 [rows, 6, 2]: the player's six trueskill_<mode>_mu/_sigma pairs

    Returns a RowBatch for Rater.rate_batch(b.ids, b.meta, b.table, seed=b.seed, shared_col=0,
    mode_cols=(1, ..., 6)); updates(b, result) turns the result into the rows to write."""
    col = rows.__getitem__
    match, player = np.asarray(col("match"), dtype=np.int64), np.asarray(col("player"), dtype=np.int64)
    n_rows = match.shape[0]
    b = RowBatch()
    b.n_rows = n_rows
    b.match_keys, row_match, first_row_of_match = _first_appearance(match)
    b.player_keys, row_player, first_row_of_player = _first_appearance(player)
    n, n_p = len(b.match_keys), len(b.player_keys)
    u = max(n_p, 1)
    # meta word per match (rater.py:70-106)
    mode = np.asarray(col("game_mode"), dtype=np.int64)[first_row_of_match]
    supported = (mode >= 0) & (mode <= 5)
    roster = np.asarray(col("roster"), dtype=np.int64)
    winner = np.asarray(col("winner"), dtype=bool)
    invalid = np.asarray(col("n_rosters"))[first_row_of_match] != 2
    afk = np.zeros(n, dtype=bool)
    afk[row_match[np.asarray(col("went_afk")) == 1]] = True
    live = supported & ~invalid & ~afk
    won = np.zeros((2, n), dtype=bool)                       # winner flags matter for rated matches only
    placed = (roster == 0) | (roster == 1)
    w_rows = np.flatnonzero(winner & placed)
    won[roster[w_rows], row_match[w_rows]] = True
    bits = np.where(supported, mode, L.MODE_UNSUPPORTED) << L.META_MODE_SHIFT
    bits = bits | np.where(supported & invalid, L.META_INVALID, 0) | np.where(supported & afk, L.META_AFK, 0)
    bits = bits | np.where(live & won[0], L.META_WIN_A, 0) | np.where(live & won[1], L.META_WIN_B, 0)
    b.meta = bits.astype(np.uint32)
    # slot of every row: position inside its (match, roster) group, in row order
    grp = row_match * 2 + np.where(placed, roster, 0)
    order = np.argsort(grp, kind="stable")
    g_sorted = grp[order]
    start = np.flatnonzero(np.r_[True, g_sorted[1:] != g_sorted[:-1]])
    pos = np.empty(n_rows, dtype=np.int64)
    pos[order] = np.arange(n_rows) - np.repeat(start, np.diff(np.r_[start, n_rows]))
    widest = int(pos[placed].max()) + 1 if placed.any() else 1
    if widest > 5:
        raise ValueError("rosters of more than 5 participants are not supported by the kernel")
    b.slots = slots = 6 if widest <= 3 else 10
    row_slot = np.where(placed, roster * (slots // 2) + pos, -1)
    ids = np.full((n, slots), L.EMPTY_SLOT, dtype=np.uint32)
    use = np.flatnonzero(placed & live[row_match])
    ids[row_match[use], row_slot[use]] = row_player[use]
    b.ids, b.row_match, b.row_slot, b.row_player = ids, row_match, row_slot, row_player
    # player table (pre-batch values: the same in every row of a player) and seeds
    table = np.full((u, 8, 2), np.nan)
    table[:n_p, 0, 0] = np.asarray(col("trueskill_mu"), dtype=np.float64)[first_row_of_player]
    table[:n_p, 0, 1] = np.asarray(col("trueskill_sigma"), dtype=np.float64)[first_row_of_player]
    table[:n_p, 1:7] = np.asarray(col("queue"), dtype=np.float64)[first_row_of_player]
    table[np.isnan(table[:, :, 0])] = np.nan                        # a NULL mu makes the pair NULL
    seed = np.full((u, 2), np.nan)
    seed[:n_p] = _seed_columns(np.asarray(col("rank_points_ranked"), dtype=np.float64)[first_row_of_player],
                               np.asarray(col("rank_points_blitz"), dtype=np.float64)[first_row_of_player],
                               np.asarray(col("skill_tier"))[first_row_of_player])
    b.table, b.seed = table, seed
    return b


def updates(b, res):
    """What the single commit of worker.py:194 flushes after rate_match ran over the batch, as
    arrays aligned with the input rows / matches / players of `b` (NaN or False mask = not written):

        match_quality [N], match_written [N]                         match.trueskill_quality (rater.py:103,141)
        part_mu, part_sigma, part_delta [rows], part_written [rows]   participant.trueskill_* (rater.py:150-157)
        items_mode [rows] (0..5), items_mu, items_sigma [rows]        participant_items <mode> pair (rater.py:167,169)
        items_any_afk [rows], items_written [rows]                    participant_items.any_afk (rater.py:99,105,110)
        player_table [U, 8, 2], player_touched [U]                    player.trueskill_* pairs (shared + six queues)
    """
    status = np.asarray(res["status"])
    n = status.shape[0]
    rated = status == L.ST_OK
    gated = (status == L.ST_AFK) | (status == L.ST_INVALID)
    out = {}
    out["match_quality"] = np.where(rated, np.asarray(res["quality"]), np.where(gated, 0.0, np.nan))
    out["match_written"] = rated | gated
    rm, rs = b.row_match, b.row_slot
    on = (rs >= 0) & rated[rm]
    sh, de, qu = np.asarray(res["out_shared"]), np.asarray(res["out_delta"]), np.asarray(res["out_queue"])
    safe_slot = np.where(rs >= 0, rs, 0)
    pick = lambda a: np.where(on, a[rm, safe_slot], np.nan)
    out["part_mu"], out["part_sigma"] = pick(sh[:, :, 0]), pick(sh[:, :, 1])
    out["part_delta"] = pick(de)
    out["part_written"] = on
    out["items_mode"] = ((b.meta >> L.META_MODE_SHIFT) & 15)[rm].astype(np.int64)
    out["items_mu"], out["items_sigma"] = pick(qu[:, :, 0]), pick(qu[:, :, 1])
    out["items_any_afk"] = gated[rm]
    out["items_written"] = (rated | gated)[rm]
    touched = np.zeros(b.table.shape[0], dtype=bool)
    touched[b.row_player[on]] = True
    out["player_table"] = np.asarray(res["table"])
    out["player_touched"] = touched
    return out


def rate_rows(rows, device=0):
    """ingest_rows + one GPU pass + updates: the row-stream form of worker.py:176-194."""
    b = ingest_rows(rows)
    res = _rater(device).rate_batch(b.ids, b.meta, b.table, seed=b.seed, shared_col=0,
                                    mode_cols=(1, 2, 3, 4, 5, 6), want_outputs=True)
    return b, updates(b, res)


def rate_matches(matches, device=0):
    """Rate a chronologically ordered batch of match object graphs in one GPU
    pass (the loop of worker.py:176-192).  Mutates the objects; returns None.
    Raises before any object is touched if the reference would have raised for
    the batch (FloatingPointError / ValueError / KeyError)."""
    b = ingest(matches)
    if not b.matches:
        return None
    res = _rater(device).rate_batch(b.ids, b.meta, b.table, seed=b.seed, shared_col=0,
                                    mode_cols=(1, 2, 3, 4, 5, 6), want_outputs=True)
    _write_back(b, res)
    return None


def rate_match(match):
    """Mutate a match structure by updating TrueSkill values (rater.py:69).
    Returns None like the reference (worker.py:192 discards the result)."""
    return rate_matches([match])


def rate_batch(ids, meta, table, seed=None, shared_col=0, mode_cols=(1, 2, 3, 4, 5, 6), device=0, **kw):
    """Columns in, columns out: see engine.Rater.rate_batch."""
    return _rater(device).rate_batch(ids, meta, table, seed=seed, shared_col=shared_col,
                                     mode_cols=mode_cols, **kw)
