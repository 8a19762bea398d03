"""ORACLE -- test infrastructure, NOT product code.

CPU restatement of the reference's rating POLICY (/root/reference/rater.py),
kept separate from the arithmetic (oracle/shim/trueskill, restating the
third-party trueskill==0.4.4 package).  The reference file cannot travel to
the GPU box, so this port is what `bench.py --impl reference` and the
`cpu_baseline` leg time there; in this container it is validated bit-for-bit
against the unmodified rater.py (tests/test_oracle_port.py, and the golden
fixtures written by tests/golden/make_golden.py).

Citations (file:line under /root/reference):
    rater.py:10-11   UNKNOWN_PLAYER_SIGMA, TAU
    rater.py:13-27   vst_points table
    rater.py:30-37   environment
    rater.py:42-62   get_trueskill_seed
    rater.py:69-169  rate_match
    worker.py:176-192 chronological per-match loop

Parity status: the arithmetic underneath (shim) is pinned on the public
TrueSkill known-answer vectors only -- the reference holds no golden value
for rate()/quality() ("parity unpinned" by the reference's own tests).
"""
from __future__ import annotations

import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_SHIM = os.path.join(_HERE, "shim")
if _SHIM not in sys.path:
    sys.path.insert(0, _SHIM)

import mpmath  # noqa: E402
import trueskill  # noqa: E402  (the shim)

mpmath.mp.dps = 50  # rater.py:8

UNKNOWN_PLAYER_SIGMA = 500  # rater.py:10 default
TAU = 1000 / 100.0          # rater.py:11 default
BETA = 10.0 / 30 * 3000     # rater.py:34

# rater.py:70-82 -- game_mode -> column prefix on player / participant_items
MODE_COLUMNS = (
    ("casual", "trueskill_casual"),
    ("ranked", "trueskill_ranked"),
    ("blitz", "trueskill_blitz"),
    ("br", "trueskill_br"),
    ("5v5_casual", "trueskill_5v5_casual"),
    ("5v5_ranked", "trueskill_5v5_ranked"),
)
_MODE_TO_COLUMN = dict(MODE_COLUMNS)


def build_vst_points():
    """rater.py:13-27 -- tier (-1..29) -> average skill-tier points: tiers -1 and 0
    are worth 1 point; above that the table is piecewise linear in (c + 0.5), c
    counted from the last tier of the previous segment, with segment lengths
    11, 4, 9, 3, 2 and slopes 109 1/11, 50, 66 2/3, 133 1/3, 200."""
    pts = {-1: 1, 0: 1}
    segments = ((11, 109 + 1 / 11), (4, 50), (9, 66 + 2 / 3), (3, 133 + 1 / 3), (2, 200))
    top = 0                     # last tier of the previous segment
    for length, slope in segments:
        offset = pts[top] if top else 0
        for c in range(1, length + 1):
            pts[top + c] = offset + slope * (c + 0.5)
        top += length
    return pts


vst_points = build_vst_points()


def make_env(tau=TAU, draw_probability=0, beta=BETA, backend="mpmath"):
    """rater.py:30-37."""
    return trueskill.TrueSkill(backend=backend, mu=1500, sigma=1000, beta=beta,
                               tau=tau, draw_probability=draw_probability)


env = make_env()


def seed_from_fields(rank_points_ranked, rank_points_blitz, skill_tier,
                     unknown_sigma=UNKNOWN_PLAYER_SIGMA):
    """rater.py:42-62 on plain values: best non-zero rank points, else tier."""
    best = None
    for pts in (rank_points_ranked, rank_points_blitz):
        if pts is None or pts == 0:
            continue
        if best is None or pts > best:
            best = pts
    if best is not None:
        sigma = unknown_sigma * (2.0 / 3.0)
        return float(best) + sigma, sigma
    sigma = unknown_sigma
    return vst_points[skill_tier] + sigma, sigma


def get_trueskill_seed(player):
    return seed_from_fields(player.rank_points_ranked, player.rank_points_blitz, player.skill_tier)


def rate_match(match, env=env):
    """rater.py:69-169 restated.  Mutates the object graph in place."""
    column = _MODE_TO_COLUMN.get(match.game_mode)
    if column is None:                       # rater.py:83-85
        return
    gated = len(match.rosters) != 2          # rater.py:91-93
    for part in match.participants:          # rater.py:95-100
        part.participant_items[0].any_afk = False
        if part.went_afk == 1:
            gated = True
            break
    if gated:                                # rater.py:102-106
        match.trueskill_quality = 0
        for part in match.participants:
            part.participant_items[0].any_afk = True
        return

    mu_col, sigma_col = column + "_mu", column + "_sigma"
    shared_teams, queue_teams = [], []
    for roster in match.rosters:             # rater.py:108-136
        st, qt = [], []
        for part in roster.participants:
            player = part.player[0]
            if player.trueskill_mu is not None:
                s_mu, s_sigma = player.trueskill_mu, player.trueskill_sigma
            else:
                s_mu, s_sigma = get_trueskill_seed(player)
            st.append(env.create_rating(float(s_mu), float(s_sigma)))
            if getattr(player, mu_col) is not None:
                q_mu, q_sigma = getattr(player, mu_col), getattr(player, sigma_col)
            else:
                q_mu, q_sigma = s_mu, s_sigma
            qt.append(env.create_rating(float(q_mu), float(q_sigma)))
        shared_teams.append(st)
        queue_teams.append(qt)

    match.trueskill_quality = env.quality(queue_teams)          # rater.py:141
    ranks = [int(not r.winner) for r in match.rosters]          # rater.py:144

    new_shared = env.rate(shared_teams, ranks=ranks)            # rater.py:144-157
    for team, roster in zip(new_shared, match.rosters):
        for rating, part in zip(team, roster.participants):
            player = part.player[0]
            if player.trueskill_mu is not None:
                part.trueskill_delta = (float(rating.mu) - float(rating.sigma)) - \
                    (float(player.trueskill_mu) - float(player.trueskill_sigma))
            else:
                part.trueskill_delta = 0
            player.trueskill_mu = part.trueskill_mu = rating.mu
            player.trueskill_sigma = part.trueskill_sigma = rating.sigma

    new_queue = env.rate(queue_teams, ranks=[int(not r.winner) for r in match.rosters])  # rater.py:161-169
    for team, roster in zip(new_queue, match.rosters):
        for rating, part in zip(team, roster.participants):
            player = part.player[0]
            items = part.participant_items[0]
            setattr(player, mu_col, rating.mu)
            setattr(items, mu_col, rating.mu)
            setattr(player, sigma_col, rating.sigma)
            setattr(items, sigma_col, rating.sigma)


# --------------------------------------------------------------------------
# Fake ORM rows (same attribute surface as worker_test.py:6-63, plus the
# br / 5v5 columns the fakes there omit) and a columnar driver.
# --------------------------------------------------------------------------
class Obj(object):
    def __init__(self, **kw):
        self.__dict__.update(kw)


_ALL_COLS = [c for _, c in MODE_COLUMNS]


def new_player(skill_tier=None, rank_points_ranked=None, rank_points_blitz=None,
               trueskill_mu=None, trueskill_sigma=None, **queue):
    p = Obj(api_id="", skill_tier=skill_tier, rank_points_ranked=rank_points_ranked,
            rank_points_blitz=rank_points_blitz, trueskill_mu=trueskill_mu,
            trueskill_sigma=trueskill_sigma)
    for col in _ALL_COLS:
        setattr(p, col + "_mu", queue.get(col + "_mu"))
        setattr(p, col + "_sigma", queue.get(col + "_sigma"))
    return p


def new_participant(player, went_afk=0):
    items = Obj(api_id="", any_afk=False)
    for col in _ALL_COLS:
        setattr(items, col + "_mu", None)
        setattr(items, col + "_sigma", None)
    return Obj(api_id="", skill_tier=player.skill_tier, went_afk=went_afk,
               participant_items=[items], player=[player],
               trueskill_mu=None, trueskill_sigma=None, trueskill_delta=None)


def new_match(game_mode, rosters_spec, api_id=""):
    """rosters_spec: list of (winner, [participants])."""
    rosters = [Obj(api_id="", winner=w, participants=list(ps)) for w, ps in rosters_spec]
    return Obj(api_id=api_id, game_mode=game_mode, rosters=rosters,
               participants=[p for r in rosters for p in r.participants],
               trueskill_quality=None)


EMPTY = 0xFFFFFFFF
META_WIN_A, META_WIN_B, META_AFK, META_INVALID = 1, 2, 4, 8
MODE_SHIFT, MODE_MASK, MODE_UNSUPPORTED = 4, 0xF, 15
MODE_NAMES = [m for m, _ in MODE_COLUMNS]


def matches_from_columns(ids, meta, players):
    """Build the object graph a worker.py query would hand to rate_match
    (worker.py:176-192) from the columnar form used by the GPU path.
    `players` maps id -> player object and is shared by all matches, like the
    SQLAlchemy identity map."""
    n, p = ids.shape
    half = p // 2
    out = []
    for m in range(n):
        bits = int(meta[m])
        mode = (bits >> MODE_SHIFT) & MODE_MASK
        name = MODE_NAMES[mode] if mode < len(MODE_NAMES) else "unsupported"
        teams = []
        for t in range(2):
            parts = []
            for s in range(t * half, (t + 1) * half):
                pid = int(ids[m, s])
                if pid == EMPTY:
                    continue
                parts.append(new_participant(players[pid]))
            teams.append((bool(bits & (META_WIN_A << t)), parts))
        if bits & META_AFK and teams[0][1]:
            teams[0][1][0].went_afk = 1
        if bits & META_INVALID:
            teams.append((False, []))
        out.append(new_match(name, teams, api_id=str(m)))
    return out


def rate_ladder(ids, meta, ratings, env=env, mode_cols=None):
    """The single-ladder form of the path (BASELINE.json configs[1..4]): one
    (mu, sigma) per player and ONE env.rate per match -- rater.py:108-169 with
    only the queue-specific rating kept.  `ratings` maps player id -> (mu,
    sigma) floats and is updated in place, match after match in stream order
    (worker.py:192).  Gated matches are skipped (rater.py:83-85,102-106)."""
    n, p = ids.shape
    half = p // 2
    for m in range(n):
        bits = int(meta[m])
        mode = (bits >> MODE_SHIFT) & MODE_MASK
        if mode >= 6 or (mode_cols is not None and mode_cols[mode] < 0):
            continue
        if bits & (META_AFK | META_INVALID):
            continue
        teams, who = [], []
        for t in range(2):
            row = [int(x) for x in ids[m, t * half:(t + 1) * half] if x != EMPTY]
            who.append(row)
            teams.append([env.create_rating(float(ratings[i][0]), float(ratings[i][1])) for i in row])
        ranks = [int(not (bits & META_WIN_A)), int(not (bits & META_WIN_B))]
        new = env.rate(teams, ranks=ranks)
        for row, team in zip(who, new):
            for i, r in zip(row, team):
                ratings[i] = (r.mu, r.sigma)
    return ratings
