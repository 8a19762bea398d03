"""The streaming (dataflow) rating path: ts_link_pred + ts_rate_stream through
the C ABI, against the sequential oracles.  pred is integer work and must be
bit-exact; ratings within 1e-9 (north-star tolerance) and equal to what the
round-based path gives."""
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


def _pred_oracle(ids, meta, mode_cols):
    """Sequential definition: previous rated match of each slot's player."""
    n, p = ids.shape
    last = {}
    pred = np.full((n, p), 0xFFFFFFFF, dtype=np.uint32)
    live = np.zeros((n, p), dtype=bool)
    for m in range(n):
        bits = int(meta[m])
        mode = (bits >> 4) & 15
        if mode >= 6 or mode_cols[mode] < 0 or bits & 12:
            continue
        row = [int(x) for x in ids[m]]
        for s, pl in enumerate(row):
            if pl == H.EMPTY:
                continue
            live[m, s] = True
            pred[m, s] = last.get(pl, 0xFFFFFFFF)
        for pl in row:
            if pl != H.EMPTY:
                last[pl] = m
    return pred, live


@pytest.fixture(scope="module")
def eng():
    from analyzer_b200 import engine
    return engine


def test_link_pred_bit_exact(eng):
    rater = eng.Rater(eng.Env(), 0)
    rng = np.random.default_rng(5)
    ids, meta = H.small_ladder(rng, 20000, 900, slots=10, modes=(1,), p_afk=0.1, p_unsupported=0.05, ragged=True)
    ids[7, 4] = ids[7, 0]                      # a player listed twice in one match
    ids[8, 1] = ids[7, 0]
    cols = [-1, 0]
    pred = _u32(rater.link_pred(_dev(ids), _dev(meta), 900, cols))
    want, live = _pred_oracle(ids, meta, cols + [-1] * 14)
    assert np.array_equal(pred[live], want[live])
    rater.close()


@pytest.mark.parametrize("slots,lanes,p_draw,env_draw,zipf", [
    (6, 1, 0.0, 0.0, False), (6, 2, 0.05, 0.0, False), (6, 8, 0.0, 0.0, True),
    (10, 2, 0.02, 0.02, True), (10, 16, 0.02, 0.02, False), (6, 32, 0.0, 0.0, False)])
def test_stream_vs_oracle_and_rounds(eng, c_oracle, slots, lanes, p_draw, env_draw, zipf):
    """Deep chains (~400 matches per player): every dependency is exercised."""
    dev = torch.device("cuda", 0)
    n, n_players, seed = 200_000, 3_000, 0xA11CE0E0 + slots + lanes
    r = eng.Rater(eng.Env(draw_probability=env_draw), 0)
    gd = eng.make_gen(seed, n_players, slots=slots, mode=1, p_draw=p_draw, p_afk=0.03, zipf=zipf)
    ids_d, meta_d = eng.gen_matches(gd, 0, n, dev)
    tab_a = eng.gen_table(seed, 0, n_players, dev)
    tab_b = tab_a.clone()
    want = tab_a.cpu().numpy()
    pa, pb = eng.PlayerTable.ladder(tab_a, mode=1), eng.PlayerTable.ladder(tab_b, mode=1)
    outs_a, outs_b = r.alloc_outputs(n, slots), r.alloc_outputs(n, slots)
    pred = r.link_pred(ids_d, meta_d, n_players, pa.mode_cols)
    r.rate_stream(pa, ids_d, meta_d, pred, outs_a, lanes)
    sch = r.schedule(ids_d, meta_d, n_players, pb.mode_cols)
    r.rate(pb, sch, outs_b, lanes)
    assert r.error_count() == 0
    prm = c_oracle.make_params(1000.0, 10.0, env_draw)
    out = c_oracle.rate_sequential(prm, _u32(ids_d), _u32(meta_d), want, -1, pa.mode_cols[:6], quad=True,
                                   want_outputs=True)
    assert out["errors"] == 0
    got = tab_a.cpu().numpy()
    assert np.abs(got - want).max() < TOL
    assert np.abs(got - tab_b.cpu().numpy()).max() < TOL
    st = outs_a["status"].cpu().numpy()
    assert np.array_equal(st, out["status"]) and np.array_equal(st, outs_b["status"].cpu().numpy())
    rated = st == 0
    q = outs_a["quality"].cpu().numpy()
    assert np.abs(q[rated] - out["quality"][rated]).max() < 1e-12 and np.all(q[st == 2] == 0)
    oq, wq = outs_a["out_queue"].cpu().numpy()[rated], out["out_queue"][rated]
    assert np.array_equal(np.isnan(oq), np.isnan(wq))
    assert np.nanmax(np.abs(oq - wq)) < TOL
    r.close()


def test_stream_errors_do_not_stall_successors(eng, c_oracle):
    """A match the reference would raise on writes nothing but still releases
    the players' next matches (no deadlock), and is counted."""
    ids = np.array([[0, 1, 2, 3, 4, 5], [0, 6, 7, 8, 9, 10], [0, 1, 2, 6, 7, 8]], dtype=np.uint32)
    meta = np.full(3, 1 | (1 << 4), dtype=np.uint32)
    table = np.full((11, 1, 2), np.nan)
    table[:, 0] = (1500.0, 300.0)
    table[3, 0, 1] = 0.0                            # sigma == 0: ValueError for match 0
    r = eng.Rater(eng.Env(), 0)
    tab_d = _dev(table)
    ptab = eng.PlayerTable(tab_d, -1, [-1, 0])
    outs = r.alloc_outputs(3, 6)
    r.reset_errors()
    pred = r.link_pred(_dev(ids), _dev(meta), 11, ptab.mode_cols)
    r.rate_stream(ptab, _dev(ids), _dev(meta), pred, outs)
    want = table.copy()
    o = c_oracle.rate_sequential(c_oracle.make_params(), ids, meta, want, -1, [-1, 0], want_outputs=True)
    assert list(outs["status"].cpu().numpy()) == list(o["status"]) == [5, 0, 0]
    assert r.error_count() == 1
    got = tab_d.cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.nanmax(np.abs(got - want)) < TOL
    r.reset_errors()
    r.close()


def test_stream_rejects_table_without_rated_queue(eng):
    """A shared column needs at least one rated queue column next to it."""
    from analyzer_b200._lib import TsError
    r = eng.Rater(eng.Env(), 0)
    tab = torch.zeros((8, 2, 2), dtype=torch.float64, device="cuda")
    ptab = eng.PlayerTable(tab, 0, [-1, -1])
    ids = torch.zeros((1, 6), dtype=torch.int32, device="cuda")
    meta = torch.zeros((1,), dtype=torch.int32, device="cuda")
    with pytest.raises(TsError, match="at least one rated queue column"):
        r.rate_stream(ptab, ids, meta, ids.clone())
    r.close()


@pytest.mark.parametrize("slots,lanes,layout", [(6, 0, "pair"), (6, 4, "faithful"), (10, 0, "pair"), (10, 2, "faithful"),
                                                 (6, 1, "pair"), (6, 32, "pair")])
def test_stream_shared_plus_queue_vs_oracle_and_rounds(eng, c_oracle, slots, lanes, layout):
    """Reference-faithful semantics in stream order: shared + one queue pair per player,
    NULLs and seeds, quality / shared / delta / queue outputs -- against the sequential
    float128 oracle and against the round-based kernel."""
    dev = torch.device("cuda", 0)
    n, n_players, seed = 60_000, 1_500, 0xA11CE0FA + slots + lanes
    rng = np.random.default_rng(seed)
    r = eng.Rater(eng.Env(), 0)
    gd = eng.make_gen(seed, n_players, slots=slots, mode=2, p_draw=0.05, p_afk=0.03, zipf=(slots == 10))
    ids_d, meta_d = eng.gen_matches(gd, 0, n, dev)
    cols_n = 2 if layout == "pair" else 8
    shared_col, qcol = (0, 1) if layout == "pair" else (0, 3)
    mode_cols = [-1, -1, qcol]                   # only mode 2 (blitz) is rated
    table = np.full((n_players, cols_n, 2), np.nan)
    has_shared = rng.random(n_players) < 0.6
    table[has_shared, shared_col, 0] = rng.normal(1500, 400, has_shared.sum())
    table[has_shared, shared_col, 1] = rng.uniform(80, 600, has_shared.sum())
    has_queue = rng.random(n_players) < 0.3
    table[has_queue, qcol, 0] = rng.normal(1400, 400, has_queue.sum())
    table[has_queue, qcol, 1] = rng.uniform(80, 600, has_queue.sum())
    seeds = np.stack([rng.normal(1700, 300, n_players), rng.uniform(300, 500, n_players)], 1)
    want = table.copy()
    out = c_oracle.rate_sequential(c_oracle.make_params(), _u32(ids_d), _u32(meta_d), want, shared_col, mode_cols,
                                   seed=seeds, quad=True, want_outputs=True)
    assert out["errors"] == 0
    tab_a, tab_b, seed_d = _dev(table), _dev(table), _dev(seeds)
    pa, pb = eng.PlayerTable(tab_a, shared_col, mode_cols, seed_d), eng.PlayerTable(tab_b, shared_col, mode_cols, seed_d)
    outs_a, outs_b = r.alloc_outputs(n, slots), r.alloc_outputs(n, slots)
    pred = r.link_pred(ids_d, meta_d, n_players, pa.mode_cols)
    r.rate_stream(pa, ids_d, meta_d, pred, outs_a, lanes)
    r.rate(pb, r.schedule(ids_d, meta_d, n_players, pb.mode_cols), outs_b, lanes)
    assert r.error_count() == 0
    for got in (tab_a.cpu().numpy(), tab_b.cpu().numpy()):
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert np.nanmax(np.abs(got - want)) < TOL
    st = outs_a["status"].cpu().numpy()
    assert np.array_equal(st, out["status"])
    rated = st == 0
    assert np.abs(outs_a["quality"].cpu().numpy()[rated] - out["quality"][rated]).max() < 1e-12
    for key, tol in (("out_shared", TOL), ("out_queue", TOL), ("out_delta", 2 * TOL)):
        g, w = outs_a[key].cpu().numpy()[rated], out[key][rated]
        assert np.array_equal(np.isnan(g), np.isnan(w)), key
        assert np.nanmax(np.abs(g - w)) < tol, key
    r.close()


def test_large_index_range(eng):
    """300M slot entries (5v5, 30M matches, Zipf): index arithmetic beyond 2^28
    (a 64-bit overflow in the divide-by-slots multiplication once broke exactly
    this).  Checked on the device: every pred is an earlier match containing the
    same player, first appearances carry TS_NONE, and the streaming and
    round-based paths agree on the final table."""
    dev = torch.device("cuda", 0)
    n, u, slots = 30_000_000, 3_000_000, 10
    r = eng.Rater(eng.Env(draw_probability=0.02), 0)
    gen = eng.make_gen(0xA11CE003, u, slots=slots, mode=1, p_draw=0.02, zipf=True)
    ids, meta = eng.gen_matches(gen, 0, n, dev)
    cols = [-1, 0]
    pred = r.link_pred(ids, meta, u, cols)
    p64 = pred.to(torch.int64) & 0xFFFFFFFF
    has = p64 != 0xFFFFFFFF
    m_idx = torch.arange(n, device=dev).unsqueeze(1).expand(n, slots)
    assert bool((p64[has] < m_idx[has]).all())
    bad = 0
    for lo in range(0, n, 2_000_000):
        hi = min(n, lo + 2_000_000)
        rows = ids[p64[lo:hi].clamp(max=n - 1).reshape(-1)].reshape(hi - lo, slots, slots)
        same = (rows == ids[lo:hi].unsqueeze(2)).any(dim=2)
        bad += int((~same & has[lo:hi]).sum())
    assert bad == 0
    assert int((~has).sum()) == int(torch.unique(ids).numel())
    del p64, has, m_idx, rows, same
    tab_a = eng.gen_table(3, 0, u, dev)
    tab_b = tab_a.clone()
    pa, pb = eng.PlayerTable.ladder(tab_a, mode=1), eng.PlayerTable.ladder(tab_b, mode=1)
    r.rate_stream(pa, ids, meta, pred)
    sch = r.schedule(ids, meta, u, pb.mode_cols)
    r.rate(pb, sch)
    assert r.error_count() == 0
    assert float((tab_a - tab_b).abs().max()) < TOL
    # ... and not only with each other: a prefix of a stream is an exact check (the first k matches do
    # not depend on the rest), so the first 2M matches go through the same calls and against the
    # sequential oracle -- fp64 over all of them, __float128 over the first 300k
    from oracle import c_oracle as co
    prm = co.make_params(draw_probability=0.02)
    for k, quad in ((2_000_000, False), (300_000, True)):
        tab_p = eng.gen_table(3, 0, u, dev)
        want = tab_p.cpu().numpy()
        pp = eng.PlayerTable.ladder(tab_p, mode=1)
        r.rate_stream(pp, ids[:k], meta[:k], r.link_pred(ids[:k], meta[:k], u, cols))
        out = co.rate_sequential(prm, ids[:k].cpu().numpy().view(np.uint32), meta[:k].cpu().numpy().view(np.uint32), want,
                                 -1, cols, quad=quad)
        assert out["errors"] == 0 and r.error_count() == 0
        assert float(np.abs(tab_p.cpu().numpy() - want).max()) < TOL
    r.close()


def test_config2_full_size_against_the_oracles(eng):
    """BASELINE.json configs[1] at its full size (10M 3v3 matches, 1M players): the schedule bit for
    bit against the sequential schedule oracle, both GPU paths' final tables against the sequential
    fp64 oracle over the whole ladder, and against the __float128 one over the first million matches."""
    from oracle import c_oracle as co
    dev = torch.device("cuda", 0)
    n, u, seed = 10_000_000, 1_000_000, 0xA11CE002
    r = eng.Rater(eng.Env(), 0)
    ids_d, meta_d = eng.gen_matches(eng.make_gen(seed, u, slots=6, mode=1), 0, n, dev)
    tab_s = eng.gen_table(seed, 0, u, dev)
    tab_r = tab_s.clone()
    table0 = tab_s.cpu().numpy()
    ps, pr = eng.PlayerTable.ladder(tab_s, mode=1), eng.PlayerTable.ladder(tab_r, mode=1)
    r.rate_stream(ps, ids_d, meta_d, r.link_pred(ids_d, meta_d, u, ps.mode_cols))
    sch = r.schedule(ids_d, meta_d, u, pr.mode_cols)
    r.rate(pr, sch)
    assert r.error_count() == 0
    ids, meta = ids_d.cpu().numpy().view(np.uint32), meta_d.cpu().numpy().view(np.uint32)
    rnd, order, offs, R = co.schedule(ids, meta, pr.mode_cols, u)
    assert sch.n_rounds == R
    assert np.array_equal(sch.round.cpu().numpy().view(np.uint32), rnd)
    assert np.array_equal(sch.order.cpu().numpy().view(np.uint32), order)
    assert np.array_equal(sch.round_offsets.cpu().numpy().view(np.uint32)[:R + 2], offs)
    want = table0.copy()
    assert co.rate_sequential(co.make_params(), ids, meta, want, -1, pr.mode_cols[:6], quad=False)["errors"] == 0
    assert float(np.abs(tab_s.cpu().numpy() - want).max()) < TOL
    assert float(np.abs(tab_r.cpu().numpy() - want).max()) < TOL
    k = 1_000_000
    want = table0.copy()
    co.rate_sequential(co.make_params(), ids[:k], meta[:k], want, -1, pr.mode_cols[:6], quad=True)
    tab_p = torch.from_numpy(table0.copy()).to(dev)
    pp = eng.PlayerTable.ladder(tab_p, mode=1)
    r.rate_stream(pp, ids_d[:k], meta_d[:k], r.link_pred(ids_d[:k], meta_d[:k], u, pp.mode_cols))
    assert float(np.abs(tab_p.cpu().numpy() - want).max()) < TOL
    r.close()


@pytest.mark.parametrize("window", [1, 777, 50_000, 10**9])
def test_windows_equal_one_shot(eng, window):
    """Micro-batch windows with carried chain tails (worker.py:18-20) give the same
    table as one link + rate over the whole stream."""
    dev = torch.device("cuda", 0)
    n, u = (3_000 if window == 1 else 120_000), 2_000
    r = eng.Rater(eng.Env(), 0)
    gen = eng.make_gen(0xA11CE0F7, u, slots=6, mode=1, p_afk=0.05)
    ids, meta = eng.gen_matches(gen, 0, n, dev)
    tab_a = eng.gen_table(9, 0, u, dev)
    tab_b = tab_a.clone()
    pa, pb = eng.PlayerTable.ladder(tab_a, mode=1), eng.PlayerTable.ladder(tab_b, mode=1)
    r.rate_stream_windows(pa, ids, meta, window)
    pred = r.link_pred(ids, meta, u, pb.mode_cols)
    r.rate_stream(pb, ids, meta, pred)
    assert r.error_count() == 0
    assert torch.equal(tab_a, tab_b)
    r.close()


def test_rate_batch_pipelined_windows(eng):
    """Host-buffer call with the copy / compute pipeline over windows == one shot."""
    r = eng.Rater(eng.Env(), 0)
    n, u = 300_000, 20_000
    gen = eng.make_gen(0xA11CE0F8, u, slots=6, mode=1)
    ids_d, meta_d = eng.gen_matches(gen, 0, n, torch.device("cuda", 0))
    ids_h, meta_h = ids_d.cpu().pin_memory(), meta_d.cpu().pin_memory()
    tab_h = eng.gen_table(5, 0, u, torch.device("cuda", 0)).cpu().pin_memory()
    a = r.rate_batch(ids_h, meta_h, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, windows=5)
    b = r.rate_batch(ids_h, meta_h, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, windows=1)
    assert np.array_equal(a["table"], b["table"]) and a["errors"] == b["errors"] == 0
    # the meta column shipped as bytes and widened on the device: same result, both ingest forms
    meta8 = meta_h.to(torch.uint8).pin_memory()
    for windows in (5, 1):
        c = r.rate_batch(ids_h, meta8, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, windows=windows)
        assert np.array_equal(a["table"], c["table"]) and c["errors"] == 0
    d = r.rate_batch(ids_h.numpy().view(np.uint32), meta_h.numpy().astype(np.uint8), tab_h.numpy(), shared_col=-1,
                     mode_cols=[-1, 0], path="rounds")
    assert np.abs(a["table"] - d["table"]).max() < TOL
    # the id column shipped at the width it needs (15 bits for this table) and widened on the device
    packed = eng.pack_ids(ids_h.numpy().view(np.uint32), u, pin=True)
    assert packed.nbytes < ids_h.numel() * 4 * 0.7
    back = r.unpack_ids(packed.packed.cuda(), packed.bits, torch.empty_like(ids_h, device="cuda"))
    assert torch.equal(back.cpu(), ids_h)
    for windows in (5, 1, [1, 2, 3, 4], [0.5, 0.5]):          # equal parts, or growing ones given as fractions
        e = r.rate_batch(packed, meta8, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, windows=windows)
        assert np.array_equal(a["table"], e["table"]) and e["errors"] == 0
    # ragged records (empty slots) survive the packing
    ids_r = ids_h.numpy().view(np.uint32).copy()
    ids_r[::7, 2] = 0xFFFFFFFF
    f = r.rate_batch(ids_r, meta8, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, windows=1)
    g = r.rate_batch(eng.pack_ids(ids_r, u), meta8, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, windows=3)
    assert np.array_equal(f["table"], g["table"])
    r.close()


def test_watchdog_on_unsatisfiable_links(eng, monkeypatch):
    """pred from a different stream can never be satisfied: the kernel must give up
    (not hang the GPU) and the host must be told."""
    from analyzer_b200._lib import TsError
    monkeypatch.setenv("TSRATE_STREAM_MAX_SPINS", "4096")
    dev = torch.device("cuda", 0)
    r = eng.Rater(eng.Env(), 0)
    n, u = 4000, 500
    gen = eng.make_gen(11, u, slots=6, mode=1)
    ids, meta = eng.gen_matches(gen, 0, n, dev)
    tab = eng.gen_table(11, 0, u, dev)
    pt = eng.PlayerTable.ladder(tab, mode=1)
    pred = torch.full_like(ids, 123456)              # a match that is not in this stream
    r.rate_stream(pt, ids, meta, pred)
    with pytest.raises(TsError, match="gave up"):
        r.error_count()
    # the flag is sticky (a stall in one window of a pipelined batch must not be lost) ...
    with pytest.raises(TsError, match="gave up"):
        r.error_count()
    # ... until the errors are reset: a proper run afterwards is fine
    r.reset_errors()
    monkeypatch.delenv("TSRATE_STREAM_MAX_SPINS")
    tab2 = eng.gen_table(11, 0, u, dev)
    pt2 = eng.PlayerTable.ladder(tab2, mode=1)
    r.rate_stream(pt2, ids, meta, r.link_pred(ids, meta, u, pt2.mode_cols))
    assert r.error_count() == 0
    r.close()


def test_link_windows_and_sort_variants_bit_exact(eng, monkeypatch):
    """pred from windowed linking == one call; the 3-kernel radix passes (fallback for
    2^30 and more slot entries) == the single-sweep passes."""
    dev = torch.device("cuda", 0)
    r = eng.Rater(eng.Env(), 0)
    n, u = 400_000, 30_000
    gen = eng.make_gen(0xA11CE0F9, u, slots=10, mode=1, p_afk=0.05, zipf=True)
    ids, meta = eng.gen_matches(gen, 0, n, dev)
    cols = [-1, 0]
    one = r.link_pred(ids, meta, u, cols)
    win = r.link_pred_windows(ids, meta, u, cols, window=70_001)
    assert torch.equal(one, win)
    sch_a = r.schedule(ids, meta, u, cols)
    monkeypatch.setenv("TSRATE_ONESWEEP", "0")
    three = r.link_pred(ids, meta, u, cols)
    sch_b = r.schedule(ids, meta, u, cols)
    assert torch.equal(one, three)
    assert torch.equal(sch_a.order, sch_b.order) and torch.equal(sch_a.round, sch_b.round)
    # rounds by dataflow along the links (default) == level-synchronous frontier sweep
    monkeypatch.delenv("TSRATE_ONESWEEP")
    monkeypatch.setenv("TSRATE_SCHEDULE_FRONTIER", "1")
    sch_c = r.schedule(ids, meta, u, cols)
    assert sch_a.n_rounds == sch_c.n_rounds
    assert torch.equal(sch_a.order, sch_c.order) and torch.equal(sch_a.round, sch_c.round)
    k = sch_a.n_rounds + 2
    assert torch.equal(sch_a.round_offsets[:k], sch_c.round_offsets[:k])
    assert torch.equal(sch_a.info[:3], sch_c.info[:3])
    r.close()
