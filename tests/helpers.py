"""Shared test helpers: synthetic ladders in columnar form and the slow
ground-truth run through the rater port + factor-graph shim (oracle O1)."""
from __future__ import annotations

import numpy as np

from oracle import rater_port as rp

EMPTY = 0xFFFFFFFF
N_COLS = 8          # row layout used by the faithful table: col 0 shared, 1..6 queues
SHARED_COL = 0
MODE_COLS = [1, 2, 3, 4, 5, 6]


def random_seed_fields(rng, n_players):
    """Static player attributes get_trueskill_seed() looks at (rater.py:42-62)."""
    tier = rng.integers(0, 30, n_players)
    ranked = np.where(rng.random(n_players) < 0.3, rng.integers(0, 3000, n_players), -1)
    blitz = np.where(rng.random(n_players) < 0.2, rng.integers(0, 3000, n_players), -1)
    return tier, ranked, blitz       # -1 stands for None


def seed_table(tier, ranked, blitz):
    out = np.empty((len(tier), 2))
    for i in range(len(tier)):
        r = None if ranked[i] < 0 else int(ranked[i])
        b = None if blitz[i] < 0 else int(blitz[i])
        out[i] = rp.seed_from_fields(r, b, int(tier[i]))
    return out


def players_from_table(table, tier, ranked, blitz):
    """Fake ORM players whose stored columns mirror the NaN-for-NULL table."""
    players = {}
    for i in range(table.shape[0]):
        kw = {}
        for mi, (_, col) in enumerate(rp.MODE_COLUMNS):
            mu, sg = table[i, MODE_COLS[mi]]
            if not np.isnan(mu):
                kw[col + "_mu"], kw[col + "_sigma"] = float(mu), float(sg)
        mu, sg = table[i, SHARED_COL]
        players[i] = rp.new_player(
            skill_tier=int(tier[i]),
            rank_points_ranked=None if ranked[i] < 0 else int(ranked[i]),
            rank_points_blitz=None if blitz[i] < 0 else int(blitz[i]),
            trueskill_mu=None if np.isnan(mu) else float(mu),
            trueskill_sigma=None if np.isnan(mu) else float(sg), **kw)
    return players


def table_from_players(players, n_players):
    t = np.full((n_players, N_COLS, 2), np.nan)
    for i in range(n_players):
        p = players[i]
        if p.trueskill_mu is not None:
            t[i, SHARED_COL] = (p.trueskill_mu, p.trueskill_sigma)
        for mi, (_, col) in enumerate(rp.MODE_COLUMNS):
            mu = getattr(p, col + "_mu")
            if mu is not None:
                t[i, MODE_COLS[mi]] = (mu, getattr(p, col + "_sigma"))
    return t


def run_objects(rate_match, ids, meta, table, tier, ranked, blitz):
    """Run `rate_match` (the port's, or the unmodified reference's) over the
    object graph built from the columns; return final table + per-match outputs."""
    n, p = ids.shape
    players = players_from_table(table, tier, ranked, blitz)
    matches = rp.matches_from_columns(ids, meta, players)
    quality = np.full(n, np.nan)
    o_sh = np.full((n, p, 2), np.nan)
    o_de = np.full((n, p), np.nan)
    o_qu = np.full((n, p, 2), np.nan)
    half = p // 2
    for m, match in enumerate(matches):
        rate_match(match)
        if match.trueskill_quality is not None:
            quality[m] = match.trueskill_quality
        mode = (int(meta[m]) >> 4) & 15
        if mode >= 6 or quality[m] == 0 or np.isnan(quality[m]):
            continue
        col = rp.MODE_COLUMNS[mode][1]
        for t, roster in enumerate(match.rosters[:2]):
            slots = [s for s in range(t * half, (t + 1) * half) if ids[m, s] != EMPTY]
            for s, part in zip(slots, roster.participants):
                o_sh[m, s] = (part.trueskill_mu, part.trueskill_sigma)
                o_de[m, s] = part.trueskill_delta
                items = part.participant_items[0]
                o_qu[m, s] = (getattr(items, col + "_mu"), getattr(items, col + "_sigma"))
    return dict(table=table_from_players(players, table.shape[0]), quality=quality,
                out_shared=o_sh, out_delta=o_de, out_queue=o_qu, matches=matches)


def small_ladder(rng, n_matches, n_players, slots=6, modes=(1,), p_draw=0.0, p_afk=0.0,
                 p_invalid=0.0, p_unsupported=0.0, ragged=False):
    ids = np.empty((n_matches, slots), dtype=np.uint32)
    meta = np.empty(n_matches, dtype=np.uint32)
    for m in range(n_matches):
        ids[m] = rng.choice(n_players, size=slots, replace=False)
        if ragged and rng.random() < 0.3:
            half = slots // 2
            ids[m, rng.integers(1, half)] = EMPTY
            if rng.random() < 0.5:
                ids[m, half + rng.integers(1, half)] = EMPTY
        bits = 1 if rng.random() < 0.5 else 2
        if rng.random() < p_draw:
            bits = 0 if rng.random() < 0.5 else 3
        if rng.random() < p_afk:
            bits |= 4
        if rng.random() < p_invalid:
            bits |= 8
        mode = int(rng.choice(modes))
        if rng.random() < p_unsupported:
            mode = 15
        meta[m] = bits | (mode << 4)
    return ids, meta


def rows_from_matches(matches):
    """The flat participant rows the query of worker.py:176-191 would yield for these object graphs
    (one row per participant, matches in order), as columns for analyzer_b200.rater.ingest_rows."""
    mode_index = {name: i for i, (name, _) in enumerate(rp.MODE_COLUMNS)}
    cols = {k: [] for k in ("match", "game_mode", "n_rosters", "roster", "winner", "went_afk", "player", "skill_tier",
                            "rank_points_ranked", "rank_points_blitz", "trueskill_mu", "trueskill_sigma", "queue")}
    parts = []
    nan = float("nan")
    for m, match in enumerate(matches):
        for t, roster in enumerate(match.rosters):
            for part in roster.participants:
                p = part.player[0]
                cols["match"].append(1000 + 7 * m)
                cols["game_mode"].append(mode_index.get(match.game_mode, 99))
                cols["n_rosters"].append(len(match.rosters))
                cols["roster"].append(t)
                cols["winner"].append(bool(roster.winner))
                cols["went_afk"].append(int(part.went_afk))
                cols["player"].append(id(p))
                cols["skill_tier"].append(p.skill_tier)
                cols["rank_points_ranked"].append(nan if p.rank_points_ranked is None else p.rank_points_ranked)
                cols["rank_points_blitz"].append(nan if p.rank_points_blitz is None else p.rank_points_blitz)
                cols["trueskill_mu"].append(nan if p.trueskill_mu is None else p.trueskill_mu)
                cols["trueskill_sigma"].append(nan if p.trueskill_sigma is None else p.trueskill_sigma)
                q = []
                for _, col in rp.MODE_COLUMNS:
                    mu = getattr(p, col + "_mu")
                    q.append((nan, nan) if mu is None else (mu, getattr(p, col + "_sigma")))
                cols["queue"].append(q)
                parts.append(part)
    out = {k: np.array(v) for k, v in cols.items()}
    out["queue"] = np.array(cols["queue"], dtype=np.float64).reshape(-1, 6, 2)
    return out, parts
