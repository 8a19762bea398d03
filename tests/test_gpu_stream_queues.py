"""Streaming path for batches that mix game modes against the faithful table
(shared pair + six queue columns, rater.py:70-82): ts_link_pred + ts_rate_stream
through the C ABI, queue pairs versioned by ordinals.  Ratings and every
rate_match output are compared with the sequential float128 oracle and with
ts_schedule + ts_rate."""
import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


def _dev(a):
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _mixed_modes(rng, meta, modes, p_other=0.05):
    """Rewrite the mode bits of generated matches: a random rated mode each, a few
    matches of an unrated queue and of an unsupported mode."""
    meta = meta.copy()
    n = meta.shape[0]
    mode = rng.choice(np.array(modes, dtype=np.uint32), size=n)
    r = rng.random(n)
    mode = np.where(r < p_other, np.uint32(15), mode)
    return ((meta & ~np.uint32(0xF0)) | (mode.astype(np.uint32) << 4)).astype(np.uint32)


@pytest.fixture(scope="module")
def eng():
    from analyzer_b200 import engine
    return engine


def _faithful_inputs(rng, n_players, cols_used):
    table = np.full((n_players, 8, 2), np.nan)
    has_shared = rng.random(n_players) < 0.6
    table[has_shared, 0, 0] = rng.normal(1500, 400, has_shared.sum())
    table[has_shared, 0, 1] = rng.uniform(80, 600, has_shared.sum())
    for c in cols_used:
        has = rng.random(n_players) < 0.3
        table[has, c, 0] = rng.normal(1400, 400, has.sum())
        table[has, c, 1] = rng.uniform(80, 600, has.sum())
    seeds = np.stack([rng.normal(1700, 300, n_players), rng.uniform(300, 500, n_players)], 1)
    return table, seeds


@pytest.mark.parametrize("slots,lanes,modes", [(6, 0, (1, 2)), (6, 4, (0, 1, 2, 3)), (10, 0, (4, 5)), (10, 2, (0, 1, 2, 3, 4, 5)),
                                               (6, 32, (1, 2, 3)), (10, 16, (1, 5)), (6, 1, (0, 5))])
def test_stream_queues_vs_oracle_and_rounds(eng, c_oracle, slots, lanes, modes):
    """Mixed game modes in one chronological stream, deep chains (~240 slot entries per
    player), NULL shared / queue pairs, seeds, draws, AFK: table and all outputs."""
    dev = torch.device("cuda", 0)
    n, n_players, seed = 60_000, 1_500, 0xA11CE1A0 + slots + lanes
    rng = np.random.default_rng(seed)
    r = eng.Rater(eng.Env(), 0)
    gd = eng.make_gen(seed, n_players, slots=slots, mode=1, p_draw=0.05, p_afk=0.03, zipf=(slots == 10))
    ids_d, meta_gen = eng.gen_matches(gd, 0, n, dev)
    meta = _mixed_modes(rng, _u32(meta_gen), modes)
    meta_d = _dev(meta)
    mode_cols = [1, 2, 3, 4, 5, 6]
    table, seeds = _faithful_inputs(rng, n_players, [mode_cols[m] for m in modes])
    want = table.copy()
    out = c_oracle.rate_sequential(c_oracle.make_params(), _u32(ids_d), meta, want, 0, mode_cols,
                                   seed=seeds, quad=True, want_outputs=True)
    assert out["errors"] == 0
    tab_a, tab_b, seed_d = _dev(table), _dev(table), _dev(seeds)
    pa, pb = eng.PlayerTable(tab_a, 0, mode_cols, seed_d), eng.PlayerTable(tab_b, 0, mode_cols, seed_d)
    assert pa.multi_queue
    outs_a, outs_b = r.alloc_outputs(n, slots), r.alloc_outputs(n, slots)
    pred = r.link_pred(ids_d, meta_d, n_players, pa.mode_cols)
    r.rate_stream(pa, ids_d, meta_d, pred, outs_a, lanes)
    r.rate(pb, r.schedule(ids_d, meta_d, n_players, pb.mode_cols), outs_b, lanes)
    assert r.error_count() == 0
    for got in (tab_a.cpu().numpy(), tab_b.cpu().numpy()):
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert np.nanmax(np.abs(got - want)) < TOL
    st = outs_a["status"].cpu().numpy()
    assert np.array_equal(st, out["status"]) and np.array_equal(st, outs_b["status"].cpu().numpy())
    rated = st == 0
    assert rated.sum() > n // 2
    assert np.abs(outs_a["quality"].cpu().numpy()[rated] - out["quality"][rated]).max() < 1e-12
    for key, tol in (("out_shared", TOL), ("out_queue", TOL), ("out_delta", 2 * TOL)):
        g, w = outs_a[key].cpu().numpy()[rated], out[key][rated]
        assert np.array_equal(np.isnan(g), np.isnan(w)), key
        assert np.nanmax(np.abs(g - w)) < tol, key
    r.close()


@pytest.mark.parametrize("window", [1, 777, 50_000])
def test_queue_windows_equal_one_shot(eng, window):
    """Micro-batch windows (worker.py:18-20) over a mixed-mode stream: chain tails
    carried, rows and queue ordinals kept in the workspace."""
    dev = torch.device("cuda", 0)
    n, u = (3_000 if window == 1 else 120_000), 2_000
    rng = np.random.default_rng(window)
    r = eng.Rater(eng.Env(), 0)
    gen = eng.make_gen(0xA11CE1B7, u, slots=6, mode=1, p_afk=0.05)
    ids, meta_gen = eng.gen_matches(gen, 0, n, dev)
    meta = _dev(_mixed_modes(rng, _u32(meta_gen), (0, 1, 2)))
    table, seeds = _faithful_inputs(rng, u, [1, 2, 3])
    tab_a, tab_b, seed_d = _dev(table), _dev(table), _dev(seeds)
    cols = [1, 2, 3, 4, 5, 6]
    pa, pb = eng.PlayerTable(tab_a, 0, cols, seed_d), eng.PlayerTable(tab_b, 0, cols, seed_d)
    r.rate_stream_windows(pa, ids, meta, window)
    r.rate_stream(pb, ids, meta, r.link_pred(ids, meta, u, cols))
    assert r.error_count() == 0
    assert torch.equal(torch.nan_to_num(tab_a, nan=-1.0), torch.nan_to_num(tab_b, nan=-1.0))
    r.close()


def test_rate_batch_mixed_modes_streams(eng):
    """The host-buffer call picks the streaming path for a deep mixed-mode batch (no
    round count) and agrees with the round-based path; pipelined windows agree too."""
    r = eng.Rater(eng.Env(), 0)
    n, u = 200_000, 5_000
    rng = np.random.default_rng(99)
    gen = eng.make_gen(0xA11CE1C8, u, slots=6, mode=1, p_draw=0.02)
    ids_d, meta_gen = eng.gen_matches(gen, 0, n, torch.device("cuda", 0))
    ids = _u32(ids_d)
    meta = _mixed_modes(rng, _u32(meta_gen), (0, 1, 2, 4))
    table, seeds = _faithful_inputs(rng, u, [1, 2, 3, 5])
    a = r.rate_batch(ids, meta, table, seed=seeds)
    b = r.rate_batch(ids, meta, table, seed=seeds, path="rounds")
    c = r.rate_batch(ids, meta, table, seed=seeds, want_outputs=False, windows=4)
    assert a["n_rounds"] is None and b["n_rounds"] is not None and a["errors"] == b["errors"] == c["errors"] == 0
    for other in (b, c):
        assert np.array_equal(np.isnan(a["table"]), np.isnan(other["table"]))
        assert np.nanmax(np.abs(a["table"] - other["table"])) < TOL
    assert np.array_equal(a["status"], b["status"])
    rated = a["status"] == 0
    for key in ("quality", "out_shared", "out_queue", "out_delta"):
        g, w = a[key][rated], b[key][rated]
        assert np.array_equal(np.isnan(g), np.isnan(w)), key
        assert np.nanmax(np.abs(g - w)) < 2 * TOL, key
    r.close()


def test_two_modes_on_one_column_and_unrated_modes(eng, c_oracle):
    """Game modes 1 and 4 rate the same queue column, mode 3 is not rated; a player is
    listed twice in one match; ragged teams."""
    rng = np.random.default_rng(31)
    u, n = 400, 20_000
    ids, meta = H.small_ladder(rng, n, u, slots=10, modes=(0, 1, 2, 3, 4, 5), p_afk=0.05, p_unsupported=0.05, ragged=True)
    ids[7, 4] = ids[7, 0]                      # listed twice, same team: both slots write the same pair
    cols = [1, 2, 3, -1, 2, 6]
    table, seeds = _faithful_inputs(rng, u, [1, 2, 3, 6])
    want = table.copy()
    out = c_oracle.rate_sequential(c_oracle.make_params(), ids, meta, want, 0, cols, seed=seeds, quad=True, want_outputs=True)
    r = eng.Rater(eng.Env(), 0)
    tab, seed_d = _dev(table), _dev(seeds)
    pt = eng.PlayerTable(tab, 0, cols, seed_d)
    outs = r.alloc_outputs(n, 10)
    r.reset_errors()
    r.rate_stream(pt, _dev(ids), _dev(meta), r.link_pred(_dev(ids), _dev(meta), u, pt.mode_cols), outs)
    assert r.error_count() == out["errors"]
    got = tab.cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.nanmax(np.abs(got - want)) < TOL
    assert np.array_equal(outs["status"].cpu().numpy(), out["status"])
    r.reset_errors()
    r.close()


def test_more_than_1024_matches_of_one_queue_per_player(eng, c_oracle):
    """Queue ordinals are 10-bit tags: a player with thousands of matches in one queue wraps them."""
    rng = np.random.default_rng(37)
    u, n = 40, 30_000                         # ~4500 slot entries per player
    ids, meta = H.small_ladder(rng, n, u, slots=6, modes=(1, 2))
    cols = [1, 2, 3, 4, 5, 6]
    table, seeds = _faithful_inputs(rng, u, [2, 3])
    want = table.copy()
    out = c_oracle.rate_sequential(c_oracle.make_params(), ids, meta, want, 0, cols, seed=seeds, quad=True)
    assert out["errors"] == 0
    r = eng.Rater(eng.Env(), 0)
    tab, seed_d = _dev(table), _dev(seeds)
    pt = eng.PlayerTable(tab, 0, cols, seed_d)
    r.rate_stream(pt, _dev(ids), _dev(meta), r.link_pred(_dev(ids), _dev(meta), u, pt.mode_cols))
    assert r.error_count() == 0
    got = tab.cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.nanmax(np.abs(got - want)) < TOL
    r.close()


def test_queue_stream_empty_batch(eng):
    r = eng.Rater(eng.Env(), 0)
    dev = torch.device("cuda", 0)
    ptab = eng.PlayerTable.faithful(8, dev)
    e_ids = torch.zeros((0, 6), dtype=torch.int32, device=dev)
    e_meta = torch.zeros((0,), dtype=torch.int32, device=dev)
    r.rate_stream(ptab, e_ids, e_meta, r.link_pred(e_ids, e_meta, 8, ptab.mode_cols))
    assert r.error_count() == 0 and bool(torch.isnan(ptab.data).all())
    r.close()
