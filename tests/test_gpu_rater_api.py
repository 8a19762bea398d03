"""The drop-in `rater` module on the GPU: the reference's own four tests
(worker_test.py:66-189) restated against analyzer_b200.rater with the same
fake ORM classes' attribute surface, plus batch == sequential-port parity."""
import logging

import numpy as np
import pytest

from oracle import rater_port as rp
from tests import helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rater():
    from analyzer_b200 import rater as r
    r.logger.setLevel(logging.ERROR)
    return r


def _match(went_afk, shared, mode="ranked"):
    g_player = lambda: rp.new_player(skill_tier=15, trueskill_mu=2000 if shared else None,
                                     trueskill_sigma=100 if shared else None)
    part = lambda: rp.new_participant(g_player(), went_afk=went_afk)
    return rp.new_match(mode, [(True, [part()] * 3), (False, [part()] * 3)])   # worker_test.py:130-131


def test_rate_match(rater):
    match = _match(False, False)
    assert rater.rate_match(match) is None
    p0 = match.rosters[0].participants[0].player[0]
    p1 = match.rosters[1].participants[0].player[0]
    assert p0.trueskill_mu is not None
    assert p0.trueskill_ranked_mu is not None
    assert p0.trueskill_ranked_sigma < p0.trueskill_ranked_mu
    assert 500 < p0.trueskill_ranked_mu < 2500
    assert p0.trueskill_casual_mu is None
    assert p0.trueskill_mu > p1.trueskill_mu
    assert p0.trueskill_ranked_mu > p1.trueskill_ranked_mu
    # and the exact values the unmodified reference gives for this input
    ref = _match(False, False)
    rp.rate_match(ref)
    r0 = ref.rosters[0].participants[0].player[0]
    assert abs(p0.trueskill_mu - r0.trueskill_mu) < 1e-9 and abs(p0.trueskill_sigma - r0.trueskill_sigma) < 1e-9
    assert abs(p0.trueskill_ranked_mu - r0.trueskill_ranked_mu) < 1e-9
    assert abs(match.trueskill_quality - ref.trueskill_quality) < 1e-12
    part, rpart = match.rosters[0].participants[0], ref.rosters[0].participants[0]
    assert part.trueskill_delta == rpart.trueskill_delta == 0
    assert abs(part.participant_items[0].trueskill_ranked_mu - rpart.participant_items[0].trueskill_ranked_mu) < 1e-9
    assert part.participant_items[0].any_afk is False


def test_rate_match_returning(rater):
    match = _match(False, True)
    rater.rate_match(match)
    assert 1800 < match.rosters[0].participants[0].player[0].trueskill_ranked_mu < 2200


def test_rate_match_afk(rater):
    match = _match(True, False)
    rater.rate_match(match)
    assert match.rosters[0].participants[0].player[0].trueskill_mu is None
    assert match.rosters[0].participants[0].participant_items[0].any_afk == True   # noqa: E712
    assert match.trueskill_quality == 0


def test_unsupported_mode_is_untouched(rater):
    match = _match(False, True, mode="aral")
    rater.rate_match(match)
    assert match.trueskill_quality is None
    assert match.rosters[0].participants[0].player[0].trueskill_mu == 2000


def test_rate_matches_equals_sequential_port(rater):
    """worker.py:176-192: one batch, players shared between matches."""
    rng = np.random.default_rng(21)
    n_players = 60
    tier, ranked, blitz = H.random_seed_fields(rng, n_players)
    ids, meta = H.small_ladder(rng, 300, n_players, slots=10, modes=(0, 1, 2, 3, 4, 5), p_draw=0.1, p_afk=0.1,
                               p_invalid=0.03, p_unsupported=0.03, ragged=True)
    table = np.full((n_players, H.N_COLS, 2), np.nan)
    table[::5, H.SHARED_COL] = (1650.0, 180.0)
    want = H.run_objects(rp.rate_match, ids, meta, table, tier, ranked, blitz)
    got = H.run_objects(lambda m: None, ids, meta, table, tier, ranked, blitz)   # fresh, unrated graph
    players = got["matches"]
    rater.rate_matches(players)
    # collect from the mutated object graph
    pl = {}
    for match in players:
        for part in match.participants:
            pl[id(part.player[0])] = part.player[0]
    for mw, mg in zip(want["matches"], players):
        assert (mw.trueskill_quality is None) == (mg.trueskill_quality is None)
        if mw.trueskill_quality is not None:
            assert abs(mw.trueskill_quality - mg.trueskill_quality) < 1e-12
        for pw, pg in zip(mw.participants, mg.participants):
            assert pw.participant_items[0].any_afk == pg.participant_items[0].any_afk
            for attr in ("trueskill_mu", "trueskill_sigma", "trueskill_delta"):
                a, b = getattr(pw, attr), getattr(pg, attr)
                assert (a is None) == (b is None)
                if a is not None:
                    assert abs(a - b) < 2e-9
            for _, col in rp.MODE_COLUMNS:
                for suffix in ("_mu", "_sigma"):
                    a = getattr(pw.participant_items[0], col + suffix)
                    b = getattr(pg.participant_items[0], col + suffix)
                    assert (a is None) == (b is None)
                    if a is not None:
                        assert abs(a - b) < 1e-9
            for attr in ["trueskill_mu", "trueskill_sigma"] + [c + s for _, c in rp.MODE_COLUMNS for s in ("_mu", "_sigma")]:
                a, b = getattr(pw.player[0], attr), getattr(pg.player[0], attr)
                assert (a is None) == (b is None)
                if a is not None:
                    assert abs(a - b) < 1e-9


def test_seed_policy_on_device(rater):
    """ts_seed_table == get_trueskill_seed (rater.py:42-62), bit for bit, KeyError cases flagged."""
    import torch
    from analyzer_b200 import engine
    rng = np.random.default_rng(4)
    n = 5000
    tier = rng.integers(-3, 33, n).astype(np.int8)
    ranked = np.where(rng.random(n) < 0.4, rng.integers(0, 3000, n).astype(np.float64), np.nan)
    blitz = np.where(rng.random(n) < 0.3, rng.integers(0, 3000, n).astype(np.float64), np.nan)
    ranked[::17] = 0.0
    dev = torch.device("cuda", 0)
    seed, bad = engine.seed_table(rater.vst_points, rater.UNKNOWN_PLAYER_SIGMA, n, dev,
                                  torch.from_numpy(ranked).to(dev), torch.from_numpy(blitz).to(dev),
                                  torch.from_numpy(tier).to(dev))
    seed, bad = seed.cpu().numpy(), bad.cpu().numpy()
    for i in range(n):
        r = None if np.isnan(ranked[i]) else ranked[i]
        b = None if np.isnan(blitz[i]) else blitz[i]
        try:
            want = rp.seed_from_fields(r, b, int(tier[i]))
            assert bad[i] == 0 and tuple(seed[i]) == want, (i, seed[i], want)
        except KeyError:
            assert bad[i] == 1 and np.isnan(seed[i]).all()


def test_rate_rows_equals_rate_matches(rater):
    """worker.py:176-194 as a row stream: ingest_rows + one GPU pass + updates gives, row for row, what
    rate_matches writes into the object graphs of the same batch."""
    rng = np.random.default_rng(33)
    n_players = 70
    tier, ranked, blitz = H.random_seed_fields(rng, n_players)
    ids, meta = H.small_ladder(rng, 500, n_players, slots=10, modes=(0, 1, 2, 3, 4, 5), p_draw=0.1, p_afk=0.1,
                               p_invalid=0.0, p_unsupported=0.04, ragged=True)
    table = np.full((n_players, H.N_COLS, 2), np.nan)
    table[::4, H.SHARED_COL] = (1650.0, 180.0)
    table[::8, 2] = (1700.0, 150.0)
    matches = H.run_objects(lambda m: None, ids, meta, table, tier, ranked, blitz)["matches"]
    rows, parts = H.rows_from_matches(matches)           # the rows BEFORE rating
    b, up = rater.rate_rows(rows)
    rater.rate_matches(matches)                          # mutates the object graphs
    assert len(parts) == b.n_rows
    for m, match in enumerate(matches):
        q = match.trueskill_quality
        assert (q is None) == (not up["match_written"][m])
        if q is not None:
            assert abs(q - up["match_quality"][m]) < 1e-15
    for i, part in enumerate(parts):
        assert (part.trueskill_mu is None) == (not up["part_written"][i])
        if part.trueskill_mu is not None:
            assert part.trueskill_mu == up["part_mu"][i] and part.trueskill_sigma == up["part_sigma"][i]
            assert part.trueskill_delta == up["part_delta"][i]
        items = part.participant_items[0]
        if up["items_written"][i]:             # (unsupported game mode: the reference leaves the row alone)
            assert bool(items.any_afk) == bool(up["items_any_afk"][i])
        if up["part_written"][i]:
            col = rp.MODE_COLUMNS[int(up["items_mode"][i])][1]
            assert getattr(items, col + "_mu") == up["items_mu"][i] and getattr(items, col + "_sigma") == up["items_sigma"][i]
    row_of = {int(k): j for j, k in enumerate(b.player_keys)}
    seen = set()
    for part in parts:
        p = part.player[0]
        if id(p) in seen:
            continue
        seen.add(id(p))
        j = row_of[id(p)]
        if up["player_touched"][j]:
            assert p.trueskill_mu == up["player_table"][j, 0, 0] and p.trueskill_sigma == up["player_table"][j, 0, 1]
            for mi, (_, col) in enumerate(rp.MODE_COLUMNS):
                mu = getattr(p, col + "_mu")
                if mu is not None:
                    assert mu == up["player_table"][j, 1 + mi, 0]
