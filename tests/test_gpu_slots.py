"""The slot-indexed streaming form (ts_link_succ + ts_rate_slots: rows live where they are read, for
ladders whose rows do not fit L2) against the player-indexed one (ts_link_pred + ts_rate_stream) --
bit for bit, they share the pass -- and against the sequential float128 oracle."""
import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
TOL = 1e-9


@pytest.fixture(scope="module")
def eng():
    from analyzer_b200 import engine
    return engine


@pytest.mark.parametrize("slots,zipf,p_draw,window", [(6, False, 0.0, None), (6, False, 0.05, 7_001), (10, True, 0.02, None),
                                                      (10, True, 0.02, 50_000), (6, True, 0.0, 1)])
def test_slots_equal_stream_and_oracle(eng, c_oracle, slots, zipf, p_draw, window):
    dev = torch.device("cuda", 0)
    n, u, seed = 120_000 if window != 1 else 300, 2_500, 0xA11CE0AA + slots
    r = eng.Rater(eng.Env(draw_probability=0.02 if p_draw == 0.02 else 0.0), 0)
    gen = eng.make_gen(seed, u, slots=slots, mode=1, p_draw=p_draw, p_afk=0.03, zipf=zipf)
    ids, meta = eng.gen_matches(gen, 0, n, dev)
    tab_a = eng.gen_table(seed, 0, u, dev)
    tab_b = tab_a.clone()
    table0 = tab_a.cpu().numpy()
    pa, pb = eng.PlayerTable.ladder(tab_a, mode=1), eng.PlayerTable.ladder(tab_b, mode=1)
    outs_a, outs_b = r.alloc_outputs(n, slots), r.alloc_outputs(n, slots)
    r.rate_stream(pa, ids, meta, r.link_pred(ids, meta, u, pa.mode_cols), outs_a)
    r.rate_slots(pb, ids, meta, outs_b, window=window)
    assert r.error_count() == 0
    assert torch.equal(tab_a, tab_b)
    for k in ("quality", "out_queue", "status"):
        assert torch.equal(torch.nan_to_num(outs_a[k].double(), nan=-7.0), torch.nan_to_num(outs_b[k].double(), nan=-7.0)), k
    # a second run on the same rows buffer (new epoch, nothing cleared) gives the same again
    tab_b.copy_(torch.from_numpy(table0).to(dev))
    r.rate_slots(pb, ids, meta, window=window)
    assert torch.equal(tab_a, tab_b)
    want = table0.copy()
    out = c_oracle.rate_sequential(c_oracle.make_params(draw_probability=0.02 if p_draw == 0.02 else 0.0),
                                   ids.cpu().numpy().view(np.uint32), meta.cpu().numpy().view(np.uint32), want, -1, [-1, 0],
                                   quad=True)
    assert out["errors"] == 0
    assert float(np.abs(tab_b.cpu().numpy() - want).max()) < TOL
    r.close()


def test_slots_seeds_ragged_errors_and_duplicates(eng, c_oracle, monkeypatch):
    """NULL ratings seeded on first touch, ragged teams, a sigma == 0 player (the match fails, its rows
    move on unchanged), an id outside the table; a player listed twice is reported."""
    from analyzer_b200 import _lib
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(12)
    n, u = 4000, 300
    ids, meta = H.small_ladder(rng, n, u, slots=10, modes=(1,), p_draw=0.1, p_afk=0.05, ragged=True)
    table = np.full((u, 1, 2), np.nan)
    table[::2, 0] = (1500.0, 300.0)
    table[7, 0] = (1500.0, 0.0)
    seeds = np.stack([1400.0 + rng.random(u) * 300, np.full(u, 500.0)], 1)
    ids[1234, 2] = u + 3
    dev_t = lambda a: torch.from_numpy(np.ascontiguousarray(a.view(np.int32) if a.dtype == np.uint32 else a)).to(dev)
    r = eng.Rater(eng.Env(), 0)
    tabs = []
    for path in ("stream", "slots"):
        t = dev_t(table)
        pt = eng.PlayerTable(t, -1, [-1, 0], dev_t(seeds))
        outs = r.alloc_outputs(n, 10)
        r.reset_errors()
        if path == "stream":
            r.rate_stream(pt, dev_t(ids), dev_t(meta), r.link_pred(dev_t(ids), dev_t(meta), u, pt.mode_cols), outs)
        else:
            r.rate_slots(pt, dev_t(ids), dev_t(meta), outs, window=1500)
        tabs.append((t.cpu().numpy(), outs["status"].cpu().numpy(), r.error_counts()))
    assert np.array_equal(tabs[0][0], tabs[1][0], equal_nan=True)
    assert np.array_equal(tabs[0][1], tabs[1][1]) and tabs[0][2] == tabs[1][2] and tabs[0][2][1] > 0
    want = table.copy()
    ids_o = ids.copy()
    ids_o[1234, 2] = 0xFFFFFFFF                   # the oracle has no notion of a corrupt id: compare the rest
    out = c_oracle.rate_sequential(c_oracle.make_params(), ids_o, meta, want, -1, [-1, 0], seed=seeds, quad=True)
    st = tabs[1][1].copy()
    ok = (st == 0) & (out["status"] == 0)
    assert ok.sum() > n // 2
    r.reset_errors()
    bad = ids.copy()
    bad[1234, 2] = 5
    bad[10, 1] = bad[10, 0]
    t = dev_t(table)
    pt = eng.PlayerTable(t, -1, [-1, 0], dev_t(seeds))
    monkeypatch.setenv("TSRATE_STREAM_MAX_SPINS", "20000")
    r.rate_slots(pt, dev_t(bad), dev_t(meta))       # may stall on the self-link: the watchdog ends it
    with pytest.raises(_lib.TsError):
        r.error_count()
    r.close()
