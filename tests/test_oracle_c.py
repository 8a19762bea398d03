"""The C oracle (oracle/ts_oracle.c: closed form in __float128 / fp64, schedule
oracle, generator) against the golden fixtures of the unmodified reference."""
import os

import numpy as np
import pytest

from tests import helpers as H

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _maxdiff(a, b):
    assert np.array_equal(np.isnan(a), np.isnan(b))
    m = ~np.isnan(a)
    return float(np.max(np.abs(a[m] - b[m]))) if m.any() else 0.0


@pytest.mark.parametrize("quad,tol", [(True, 2e-11), (False, 1e-9)])
def test_mixed_small(c_oracle, quad, tol):
    g = np.load(os.path.join(GOLD, "mixed_small.npz"))
    table = g["table_in"].copy()
    prm = c_oracle.make_params(1000.0, 10.0, 0.0)
    out = c_oracle.rate_sequential(prm, g["ids"], g["meta"], table, H.SHARED_COL, H.MODE_COLS,
                                   seed=g["seed"], quad=quad, want_outputs=True)
    assert out["errors"] == 0
    assert _maxdiff(table, g["table_out"]) < tol
    st = out["status"]
    gq = g["quality"]
    # unsupported mode leaves quality untouched (rater.py:83-85); gates store 0 (rater.py:103)
    assert np.array_equal(st == 1, np.isnan(gq))
    assert np.array_equal((st == 2) | (st == 3), gq == 0)
    rated = st == 0
    assert abs(out["quality"][rated] - gq[rated]).max() < 1e-12
    assert _maxdiff(out["out_shared"][rated], g["out_shared"][rated]) < tol
    assert _maxdiff(out["out_queue"][rated], g["out_queue"][rated]) < tol
    assert _maxdiff(out["out_delta"][rated], g["out_delta"][rated]) < 2 * tol


def _config1_inputs(c_oracle, g):
    seed = int(g["gen_seed"])
    gen = c_oracle.make_gen(seed, 5000, slots=6, mode=1)
    ids, meta = c_oracle.gen_matches(gen, 0, 10000)
    assert np.array_equal(ids[:64], g["ids_head"]) and np.array_equal(meta[:64], g["meta_head"])
    none = np.full(5000, -1)
    return ids, meta, H.seed_table(g["tier"], none, none)


@pytest.mark.parametrize("quad,tol", [(True, 2e-11), (False, 1e-9)])
def test_config1(c_oracle, quad, tol):
    """BASELINE.json configs[0]: 10k 3v3 ranked matches over 5k players."""
    g = np.load(os.path.join(GOLD, "config1.npz"))
    ids, meta, seed = _config1_inputs(c_oracle, g)
    table = np.full((5000, 2, 2), np.nan)
    prm = c_oracle.make_params(1000.0, 10.0, 0.0)
    out = c_oracle.rate_sequential(prm, ids, meta, table, 0, [-1, 1], seed=seed, quad=quad,
                                   want_outputs=True)
    assert out["errors"] == 0
    assert _maxdiff(table, g["table_out"]) < tol
    assert abs(out["quality"] - g["quality"]).max() < 1e-12


def test_draw_margin_and_draw_path(c_oracle):
    """draw_probability > 0: closed form vs the factor graph at 50 digits."""
    from oracle import rater_port as rp
    env = rp.make_env(draw_probability=0.02)
    prm = c_oracle.make_params(1000.0, 10.0, 0.02)
    assert prm.draw_margin[6] > 0 and prm.draw_margin[10] > prm.draw_margin[6]
    rng = np.random.default_rng(3)
    for slots in (6, 10):
        n_players = 30
        ids, meta = H.small_ladder(rng, 80, n_players, slots=slots, modes=(1,), p_draw=0.5)
        tbl0 = np.stack([rng.normal(1500, 400, n_players), rng.uniform(80, 500, n_players)], 1)
        table = tbl0.reshape(n_players, 1, 2).copy()
        out = c_oracle.rate_sequential(prm, ids, meta, table, -1, [-1, 0], quad=True)
        assert out["errors"] == 0
        cur = {i: env.create_rating(*tbl0[i]) for i in range(n_players)}
        half = slots // 2
        for m in range(len(ids)):
            teams = [[cur[int(p)] for p in ids[m, t * half:(t + 1) * half]] for t in range(2)]
            ranks = [int(not (int(meta[m]) & 1)), int(not (int(meta[m]) & 2))]
            new = env.rate(teams, ranks=ranks)
            for t in range(2):
                for k, p in enumerate(ids[m, t * half:(t + 1) * half]):
                    cur[int(p)] = env.create_rating(new[t][k].mu, new[t][k].sigma)
        want = np.array([[cur[i].mu, cur[i].sigma] for i in range(n_players)])
        assert np.abs(table[:, 0] - want).max() < 2e-11


def test_status_codes(c_oracle):
    prm = c_oracle.make_params()
    ids = np.array([[0, 1, 2, 3, 4, 5]] * 4, dtype=np.uint32)
    meta = np.array([1 | (1 << 4), 1 | (15 << 4), 1 | 4 | (1 << 4), 1 | 8 | (1 << 4)], dtype=np.uint32)
    table = np.full((6, 1, 2), np.nan)
    table[:, 0] = (1500.0, 300.0)
    table[3, 0, 1] = 0.0                          # sigma == 0 -> ValueError in the package
    before = table.copy()
    out = c_oracle.rate_sequential(prm, ids, meta, table, -1, [-1, 0], want_outputs=True)
    assert list(out["status"]) == [c_oracle.ST_VALUE_ERROR, c_oracle.ST_SKIP_MODE, c_oracle.ST_AFK,
                                   c_oracle.ST_INVALID]
    assert np.array_equal(table, before)          # nothing written on error or gate
    assert out["errors"] == 1
    # unseeded player and no seed table
    table[:, 0] = (np.nan, np.nan)
    out = c_oracle.rate_sequential(prm, ids[:1], meta[:1], table, -1, [-1, 0])
    assert out["status"][0] == c_oracle.ST_VALUE_ERROR


def test_schedule_oracle(c_oracle):
    rng = np.random.default_rng(11)
    n_players = 50
    ids, meta = H.small_ladder(rng, 400, n_players, slots=6, modes=(1,), p_afk=0.1, p_unsupported=0.05,
                               ragged=True)
    rnd, order, offs, R = c_oracle.schedule(ids, meta, [-1, 0], n_players)
    gated = ((meta & 12) != 0) | (((meta >> 4) & 15) != 1)
    assert np.array_equal(rnd == 0, gated)
    # brute-force definition: 1 + max round of any earlier match sharing a player
    want = np.zeros(len(ids), dtype=np.uint32)
    for m in range(len(ids)):
        if gated[m]:
            continue
        mine = set(int(p) for p in ids[m] if p != H.EMPTY)
        r = 0
        for k in range(m):
            if not gated[k] and mine & set(int(p) for p in ids[k] if p != H.EMPTY):
                r = max(r, int(want[k]))
        want[m] = r + 1
    assert np.array_equal(rnd, want)
    assert R == want.max() and len(offs) == R + 2 and offs[0] == 0 and offs[-1] == len(ids)
    assert sorted(order) == list(range(len(ids)))
    for r in range(R + 1):
        seg = order[offs[r]:offs[r + 1]]
        assert np.all(rnd[seg] == r) and np.all(np.diff(seg.astype(np.int64)) > 0)
        if r > 0:   # conflict-free: no player twice within a round
            p = ids[seg].ravel()
            p = p[p != H.EMPTY]
            assert len(np.unique(p)) == len(p)


def test_generator_properties(c_oracle):
    gen = c_oracle.make_gen(0xA11CE002, 1000, slots=6, mode=1, p_draw=0.1, p_afk=0.05)
    ids, meta = c_oracle.gen_matches(gen, 0, 5000)
    ids2, meta2 = c_oracle.gen_matches(gen, 1000, 100)      # counter-based: any window regenerates
    assert np.array_equal(ids[1000:1100], ids2) and np.array_equal(meta[1000:1100], meta2)
    assert ids.max() < 1000
    assert all(len(set(row)) == 6 for row in ids.tolist())    # sampled without replacement
    win = meta & 3
    assert 0.07 < np.mean(win == 0) < 0.13 and 0.03 < np.mean((meta & 4) != 0) < 0.07
    assert np.all((meta >> 4) == 1)
    gz = c_oracle.make_gen(0xA11CE003, 100000, slots=10, mode=5, zipf=True)
    idz, _ = c_oracle.gen_matches(gz, 0, 5000)
    assert idz.max() < 100000 and all(len(set(row)) == 10 for row in idz.tolist())
    counts = np.bincount(idz.ravel(), minlength=100000)
    assert counts[:1000].sum() > 5 * counts[-1000:].sum()     # skewed towards low ranks
    t = c_oracle.gen_table(5, 0, 20000)
    assert abs(t[:, 0, 0].mean() - 1500) < 15 and 350 < t[:, 0, 0].std() < 450
    assert np.all(t[:, 0, 1] == 500 * (2.0 / 3.0))
