"""GPU parity tests proper: the CUDA path, called through the C ABI
(libtsrate.so), against the oracles on the same seeded inputs.

Bars: schedule (round / order / offsets) bit-exact vs the sequential schedule
oracle; ratings, per-participant outputs and quality within 1e-9 absolute of
the sequential oracle (the tolerance BASELINE.json's north_star states); the
golden fixtures were produced by the unmodified reference rater.py."""
import os

import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-9


@pytest.fixture(scope="module")
def eng():
    from analyzer_b200 import engine
    return engine


@pytest.fixture(scope="module")
def rater(eng):
    return eng.Rater(eng.Env(), 0)


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


def _dev(a):
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _maxdiff(a, b):
    assert np.array_equal(np.isnan(a), np.isnan(b))
    m = ~np.isnan(a)
    return float(np.max(np.abs(a[m] - b[m]))) if m.any() else 0.0


def _check_schedule(c_oracle, sch, ids, meta, mode_cols, n_players):
    rnd, order, offs, R = c_oracle.schedule(ids, meta, mode_cols, n_players)
    assert sch.n_rounds == R
    assert np.array_equal(_u32(sch.round), rnd)
    assert np.array_equal(_u32(sch.order), order)
    assert np.array_equal(_u32(sch.round_offsets)[:R + 2], offs)
    assert np.array_equal(_u32(sch.ids_sorted), ids[order])
    assert np.array_equal(_u32(sch.meta_sorted), meta[order])
    info = _u32(sch.info)
    assert info[0] == R and info[1] == offs[1] and info[2] == 0
    return R


def test_golden_mixed_small(eng, rater, c_oracle):
    """Mixed 3v3/5v5 stream, 6 modes, ragged teams, AFK / invalid / unsupported,
    draws at draw_probability=0, shared + queue ratings with NULLs and seeds."""
    g = np.load(os.path.join(GOLD, "mixed_small.npz"))
    ids, meta = g["ids"], g["meta"]
    res = rater.rate_batch(ids, meta, g["table_in"], seed=g["seed"], shared_col=H.SHARED_COL,
                           mode_cols=H.MODE_COLS)
    assert _maxdiff(res["table"], g["table_out"]) < TOL
    st, gq = res["status"], g["quality"]
    assert np.array_equal(st == 1, np.isnan(gq))
    assert np.array_equal((st == 2) | (st == 3), gq == 0)
    rated = st == 0
    assert np.all(res["quality"][(st == 2) | (st == 3)] == 0)
    assert np.abs(res["quality"][rated] - gq[rated]).max() < 1e-12
    assert _maxdiff(res["out_shared"][rated], g["out_shared"][rated]) < TOL
    assert _maxdiff(res["out_queue"][rated], g["out_queue"][rated]) < TOL
    assert _maxdiff(res["out_delta"][rated], g["out_delta"][rated]) < 2 * TOL


@pytest.mark.parametrize("lanes", [1, 2, 4, 8, 16, 32])
def test_golden_config1_all_lane_widths(eng, rater, c_oracle, lanes):
    """BASELINE.json configs[0] (10k 3v3 ranked over 5k players, all unseeded)
    against the output of the unmodified reference, for every lanes-per-match
    variant of the rating kernel."""
    g = np.load(os.path.join(GOLD, "config1.npz"))
    gen = c_oracle.make_gen(int(g["gen_seed"]), 5000, slots=6, mode=1)
    ids, meta = c_oracle.gen_matches(gen, 0, 10000)
    none = np.full(5000, -1)
    seed = H.seed_table(g["tier"], none, none)
    table = np.full((5000, 2, 2), np.nan)
    res = rater.rate_batch(ids, meta, table, seed=seed, shared_col=0, mode_cols=[-1, 1],
                           lanes_per_match=lanes)
    assert res["errors"] == 0
    assert _maxdiff(res["table"], g["table_out"]) < TOL
    assert np.abs(res["quality"] - g["quality"]).max() < 1e-12


def test_device_generator_matches_host_twin(eng, c_oracle):
    dev = torch.device("cuda", 0)
    for kw in (dict(slots=6, mode=1, p_draw=0.1, p_afk=0.05), dict(slots=10, mode=5, zipf=True, p_draw=0.02)):
        n_players = 20000
        gd = eng.make_gen(0xA11CE003, n_players, **kw)
        gh = c_oracle.make_gen(0xA11CE003, n_players, **kw)
        ids_d, meta_d = eng.gen_matches(gd, 12345, 30000, dev)
        ids_h, meta_h = c_oracle.gen_matches(gh, 12345, 30000)
        assert np.array_equal(_u32(ids_d), ids_h) and np.array_equal(_u32(meta_d), meta_h)
    t = eng.gen_table(77, 1000, 50000, dev).cpu().numpy()
    assert np.array_equal(t, c_oracle.gen_table(77, 1000, 50000))


@pytest.mark.parametrize("slots,kw", [(6, dict()), (6, dict(p_afk=0.1)), (10, dict(zipf=True, p_draw=0.02))])
def test_schedule_bit_exact(eng, rater, c_oracle, slots, kw):
    dev = torch.device("cuda", 0)
    n, n_players = 300_000, 40_000
    gd = eng.make_gen(0xA11CE0B0 + slots, n_players, slots=slots, mode=1, **kw)
    ids_d, meta_d = eng.gen_matches(gd, 0, n, dev)
    cols = [-1, 0]
    sch = rater.schedule(ids_d, meta_d, n_players, cols)
    R = _check_schedule(c_oracle, sch, _u32(ids_d), _u32(meta_d), cols, n_players)
    assert R > 10


def test_schedule_edge_cases(eng, rater, c_oracle):
    cols = [-1, 0]
    # every match gated
    ids = np.arange(12, dtype=np.uint32).reshape(2, 6)
    meta = np.array([1 | 4 | (1 << 4), 1 | (15 << 4)], dtype=np.uint32)
    sch = rater.schedule(_dev(ids), _dev(meta), 12, cols)
    assert _check_schedule(c_oracle, sch, ids, meta, cols, 12) == 0
    # a single chain: every match shares player 0 -> one match per round
    n = 700
    ids = np.zeros((n, 6), dtype=np.uint32)
    ids[:, 1:] = np.arange(1, 1 + 5 * n, dtype=np.uint32).reshape(n, 5)
    meta = np.full(n, 1 | (1 << 4), dtype=np.uint32)
    sch = rater.schedule(_dev(ids), _dev(meta), 1 + 5 * n, cols)
    assert _check_schedule(c_oracle, sch, ids, meta, cols, 1 + 5 * n) == n
    # no conflicts at all -> one round; ragged teams; a player listed twice in a match
    ids = np.arange(6 * 50, dtype=np.uint32).reshape(50, 6)
    ids[3, 2] = H.EMPTY
    ids[7, 4] = ids[7, 0]
    meta = np.full(50, 2 | (1 << 4), dtype=np.uint32)
    sch = rater.schedule(_dev(ids), _dev(meta), 300, cols)
    assert _check_schedule(c_oracle, sch, ids, meta, cols, 300) == 1
    # duplicates across consecutive matches (chains must not self-link)
    ids = np.array([[0, 0, 0, 1, 1, 1], [0, 1, 2, 3, 4, 5], [0, 0, 2, 1, 1, 5]], dtype=np.uint32)
    meta = np.full(3, 1 | (1 << 4), dtype=np.uint32)
    sch = rater.schedule(_dev(ids), _dev(meta), 6, cols)
    assert _check_schedule(c_oracle, sch, ids, meta, cols, 6) == 3
    # too small an offsets array is reported, not overrun
    from analyzer_b200._lib import TsError
    # round_offsets too small: ts_schedule does not wait for the GPU, so the overflow is flagged on the
    # device -- reading n_rounds raises, and ts_rate on such a schedule rates nothing and counts an error
    sch = rater.schedule(_dev(ids), _dev(meta), 6, cols, offsets_cap=3)
    with pytest.raises(TsError):
        sch.n_rounds
    tab_d = _dev(np.tile(np.array([1500.0, 300.0]), (6, 1, 1)))
    before = tab_d.clone()
    rater.reset_errors()
    rater.rate(eng.PlayerTable(tab_d, -1, cols), sch)
    assert rater.error_counts() == (0, 1) and torch.equal(tab_d, before)
    rater.reset_errors()
    # other team sizes (2v2, 8v8): the run-time-slots form of the rounds kernel
    rng = np.random.default_rng(3)
    for slots in (4, 16):
        ids, meta = H.small_ladder(rng, 3000, 200, slots=slots, modes=(1,), p_afk=0.05, ragged=True)
        sch = rater.schedule(_dev(ids), _dev(meta), 200, cols)
        assert _check_schedule(c_oracle, sch, ids, meta, cols, 200) > 10
    # empty batch
    sch = rater.schedule(torch.empty((0, 6), dtype=torch.int32, device="cuda"),
                         torch.empty((0,), dtype=torch.int32, device="cuda"), 6, cols)
    assert sch.n_rounds == 0


@pytest.mark.parametrize("slots,p_draw,env_draw", [(6, 0.0, 0.0), (6, 0.05, 0.0), (10, 0.02, 0.02)])
def test_ladder_vs_float128_oracle(eng, c_oracle, slots, p_draw, env_draw):
    """Single-pair ladder (K=1) with deep chains: 200k matches over 3k players
    (~400 matches per player), uniform and Zipf, draws with and without margin."""
    dev = torch.device("cuda", 0)
    n, n_players, seed = 200_000, 3_000, 0xA11CE0C0 + slots
    r = eng.Rater(eng.Env(draw_probability=env_draw), 0)
    gd = eng.make_gen(seed, n_players, slots=slots, mode=1, p_draw=p_draw, zipf=(slots == 10))
    ids_d, meta_d = eng.gen_matches(gd, 0, n, dev)
    tab_d = eng.gen_table(seed, 0, n_players, dev)
    want = tab_d.cpu().numpy()
    ptab = eng.PlayerTable.ladder(tab_d, mode=1)
    sch = r.schedule(ids_d, meta_d, n_players, ptab.mode_cols)
    r.rate(ptab, sch)
    assert r.error_count() == 0
    prm = c_oracle.make_params(1000.0, 10.0, env_draw)
    out = c_oracle.rate_sequential(prm, _u32(ids_d), _u32(meta_d), want, -1, ptab.mode_cols[:6], quad=True)
    assert out["errors"] == 0
    assert np.abs(tab_d.cpu().numpy() - want).max() < TOL
    r.close()


def test_error_statuses_and_batch_failure(eng, rater, c_oracle):
    ids = np.array([[0, 1, 2, 3, 4, 5], [6, 7, 8, 9, 10, 11], [0, 1, 2, 6, 7, 8]], dtype=np.uint32)
    meta = np.full(3, 1 | (1 << 4), dtype=np.uint32)
    table = np.full((12, 1, 2), np.nan)
    table[:, 0] = (1500.0, 300.0)
    table[7, 0, 1] = 0.0                              # Rating(sigma=0) -> ValueError in the package
    with pytest.raises(ValueError):
        rater.rate_batch(ids, meta, table, shared_col=-1, mode_cols=[-1, 0])
    table[7, 0] = (np.nan, np.nan)                    # unseeded and no seed table
    with pytest.raises(ValueError):
        rater.rate_batch(ids, meta, table, shared_col=-1, mode_cols=[-1, 0])
    # the device-side statuses agree with the oracle's, match by match
    tab_d = _dev(table)
    ptab = eng.PlayerTable(tab_d, -1, [-1, 0])
    sch = rater.schedule(_dev(ids), _dev(meta), 12, ptab.mode_cols)
    outs = rater.alloc_outputs(3, 6)
    rater.reset_errors()
    rater.rate(ptab, sch, outs)
    want = table.copy()
    o = c_oracle.rate_sequential(c_oracle.make_params(), ids, meta, want, -1, [-1, 0], want_outputs=True)
    assert list(outs["status"].cpu().numpy()) == list(o["status"]) == [0, 5, 5]
    assert rater.error_count() == 2
    assert _maxdiff(tab_d.cpu().numpy(), want) < TOL
    rater.reset_errors()
    # far tails: a huge favourite winning / losing stays finite and inside the domain
    table = np.full((12, 1, 2), np.nan)
    table[:, 0] = (1500.0, 300.0)
    table[0:3, 0, 0] = 60000.0
    for bits in (1, 2):
        meta1 = np.array([bits | (1 << 4)], dtype=np.uint32)
        res = rater.rate_batch(ids[:1], meta1, table, shared_col=-1, mode_cols=[-1, 0])
        want = table.copy()
        o = c_oracle.rate_sequential(c_oracle.make_params(), ids[:1], meta1, want, -1, [-1, 0])
        assert o["errors"] == 0 and _maxdiff(res["table"], want) < 1e-7   # |mu| is 6e4 here


def test_full_size_properties(eng, rater, c_oracle):
    """Size-independent checks at a size the sequential oracle still finishes
    quickly (2M matches, 200k players): schedule bit-exact, every round
    conflict-free, per-player order preserved, ratings vs the fp64/float128 oracle."""
    dev = torch.device("cuda", 0)
    n, n_players, seed = 2_000_000, 200_000, 0xA11CE002
    gd = eng.make_gen(seed, n_players, slots=6, mode=1)
    ids_d, meta_d = eng.gen_matches(gd, 0, n, dev)
    tab_d = eng.gen_table(seed, 0, n_players, dev)
    want = tab_d.cpu().numpy()
    ptab = eng.PlayerTable.ladder(tab_d, mode=1)
    sch = rater.schedule(ids_d, meta_d, n_players, ptab.mode_cols)
    ids, meta = _u32(ids_d), _u32(meta_d)
    R = _check_schedule(c_oracle, sch, ids, meta, ptab.mode_cols, n_players)
    rnd, order, offs = _u32(sch.round), _u32(sch.order), _u32(sch.round_offsets)
    for r in (1, 2, R // 2, R):
        p = ids[order[offs[r]:offs[r + 1]]].ravel()
        assert len(np.unique(p)) == len(p)            # conflict-free
    # a player's matches get strictly increasing rounds in stream order
    flat_p = ids.ravel()
    flat_r = np.repeat(rnd, 6)
    o = np.argsort(flat_p, kind="stable")
    same = flat_p[o][1:] == flat_p[o][:-1]
    assert np.all(flat_r[o][1:][same] > flat_r[o][:-1][same])
    rater.rate(ptab, sch)
    assert rater.error_count() == 0
    out = c_oracle.rate_sequential(c_oracle.make_params(), ids, meta, want, -1, ptab.mode_cols[:6], quad=True)
    assert out["errors"] == 0
    assert np.abs(tab_d.cpu().numpy() - want).max() < TOL


def test_player_id_outside_the_table_is_refused(eng, rater):
    """A corrupt id (>= n_players) is never used as an index: its match gets TS_ST_VALUE_ERROR on
    every path, nobody else is touched through it, and the batch raises ValueError with or without
    the status column."""
    from analyzer_b200 import _lib
    rng = np.random.default_rng(5)
    n, u = 3000, 400
    ids, meta = H.small_ladder(rng, n, u, slots=6, modes=(1,))
    bad = ids.copy()
    bad[1234, 4] = u                      # exactly one past the table
    bad[2000, 0] = 0xFFFFFFF0
    tab = np.empty((u, 1, 2))
    tab[:, 0] = (1500.0, 300.0)
    finals = []
    for path in ("rounds", "stream"):
        tab_d = _dev(tab)
        ptab = eng.PlayerTable(tab_d, -1, [-1, 0])
        outs = rater.alloc_outputs(n, 6)
        rater.reset_errors()
        if path == "rounds":
            rater.rate(ptab, rater.schedule(_dev(bad), _dev(meta), u, ptab.mode_cols), outs)
        else:
            rater.rate_stream(ptab, _dev(bad), _dev(meta), rater.link_pred(_dev(bad), _dev(meta), u, ptab.mode_cols), outs)
        assert rater.error_counts() == (0, 2)
        st = outs["status"].cpu().numpy()
        assert st[1234] == st[2000] == _lib.ST_VALUE_ERROR and int((st != 0).sum()) == 2
        finals.append(tab_d.cpu().numpy())
    assert np.abs(finals[0] - finals[1]).max() < TOL
    # shared + queue layout, stream path
    tab2 = np.full((u, 2, 2), np.nan)
    tab2[:, 0] = (1500.0, 300.0)
    tab_d = _dev(tab2)
    ptab = eng.PlayerTable(tab_d, 0, [-1, 1])
    rater.reset_errors()
    rater.rate_stream(ptab, _dev(bad), _dev(meta), rater.link_pred(_dev(bad), _dev(meta), u, ptab.mode_cols))
    assert rater.error_counts() == (0, 2)
    rater.reset_errors()
    for want_outputs in (True, False):
        with pytest.raises(ValueError):
            rater.rate_batch(bad, meta, tab, shared_col=-1, mode_cols=[-1, 0], want_outputs=want_outputs)


def test_zero_probability_draw_switch(eng):
    """draw_probability = 0 and both winner flags equal: by default the analytic limit (mu of equal
    teams unchanged, sigma shrinks); with zero_draw="raise" the match fails like the float backend."""
    ids = np.array([[0, 1, 2, 3, 4, 5]], dtype=np.uint32)
    meta = np.array([3 | (1 << 4)], dtype=np.uint32)
    table = np.tile(np.array([2000.0, 100.0]), (6, 1, 1))
    r = eng.Rater(eng.Env(), 0)
    out = r.rate_batch(ids, meta, table, shared_col=-1, mode_cols=[-1, 0])
    assert np.abs(out["table"][:, 0, 0] - 2000.0).max() < 1e-9
    assert np.abs(out["table"][:, 0, 1] - 100.41498062126155).max() < 1e-9
    r.close()
    r = eng.Rater(eng.Env(zero_draw="raise"), 0)
    with pytest.raises(FloatingPointError):
        r.rate_batch(ids, meta, table, shared_col=-1, mode_cols=[-1, 0])
    meta[0] = 1 | (1 << 4)                     # a win is unaffected
    assert r.rate_batch(ids, meta, table, shared_col=-1, mode_cols=[-1, 0])["errors"] == 0
    r.close()
