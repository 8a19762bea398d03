"""Host-side logic that needs no GPU: seed policy, columnar ingest, draw margin,
and that libtsrate.so loads and exports every symbol include/tsrate.h declares."""
import ctypes
import os
import re

import numpy as np
import pytest

from analyzer_b200 import _lib as L
from analyzer_b200 import engine
from oracle import rater_port as rp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def rater_mod():
    from analyzer_b200 import rater
    return rater


def test_library_exports_every_declared_symbol():
    L.build()
    header = open(os.path.join(ROOT, "include", "tsrate.h")).read()
    declared = set(re.findall(r"\b(ts_[a-z_0-9]+)\s*\(", header))
    assert declared and declared == set(L.EXPORTS)
    lib = ctypes.CDLL(L.LIB_PATH)
    for name in declared:
        assert getattr(lib, name) is not None
    # struct sizes the ctypes mirrors must agree with (checked again on the GPU box by use)
    assert ctypes.sizeof(L.TsParams) == 8 * 2 + 8 * 17 + 8
    assert ctypes.sizeof(L.TsTable) == 8 + 8 + 8 + 4 + 4 + 16
    assert ctypes.sizeof(L.TsSchedInfo) == 16 and ctypes.sizeof(L.TsOutputs) == 40


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        engine.Rater(engine.Env(), 0)
    from analyzer_b200 import rater
    m = rp.new_match("ranked", [(True, [rp.new_participant(rp.new_player(skill_tier=3))]),
                                (False, [rp.new_participant(rp.new_player(skill_tier=4))])])
    with pytest.raises(RuntimeError):
        rater.rate_match(m)


def test_vst_points_and_seed_match_reference_policy(rater_mod):
    assert rater_mod.vst_points == rp.vst_points          # port is checked against rater.py:13-27 live
    assert rater_mod.env.beta == 1000.0 and rater_mod.env.tau == 10.0 and rater_mod.env.draw_probability == 0
    cases = [(None, None, 15), (2500, None, 0), (2500, 100, 0), (100, 2500, 0), (None, 2500, 0),
             (0, 0, 7), (0, 1200, -1), (3000.5, 2999, 29)]
    for ranked, blitz, tier in cases:
        p = rp.new_player(skill_tier=tier, rank_points_ranked=ranked, rank_points_blitz=blitz)
        assert rater_mod.get_trueskill_seed(p) == rp.seed_from_fields(ranked, blitz, tier)
    for ranked, blitz in [(2500, None), (2500, 100), (100, 2500), (None, 2500)]:   # worker_test.py:86-113
        mu, sigma = rater_mod.get_trueskill_seed(rp.new_player(skill_tier=0, rank_points_ranked=ranked,
                                                               rank_points_blitz=blitz))
        assert mu - sigma == 2500
    with pytest.raises(KeyError):
        rater_mod.get_trueskill_seed(rp.new_player(skill_tier=30))


def test_draw_margin_matches_package_formula(c_oracle):
    for p in (0.0, 0.02, 0.10):
        want = c_oracle.draw_margins(1000.0, p)     # 50-digit ppf through the shim
        for n in (2, 6, 10):
            got = engine.draw_margin(p, n, 1000.0)
            assert abs(got - want[n]) <= 1e-9 * max(1.0, abs(want[n]))
    prm = engine.Env(draw_probability=0.02).params()
    assert prm.draw_is_zero == 0 and prm.draw_margin[10] > prm.draw_margin[6] > 0
    assert engine.Env().params().draw_is_zero == 1


def test_ingest_columns(rater_mod):
    a, b, c = (rp.new_player(skill_tier=t) for t in (3, 4, 5))
    d = rp.new_player(trueskill_mu=2000, trueskill_sigma=100, trueskill_ranked_mu=1900, trueskill_ranked_sigma=90)
    mk = rp.new_participant
    m0 = rp.new_match("ranked", [(True, [mk(a), mk(b)]), (False, [mk(c), mk(d)])])
    m1 = rp.new_match("aral", [(True, [mk(a)]), (False, [mk(b)])])                       # unsupported
    m2 = rp.new_match("casual", [(False, [mk(a), mk(b, went_afk=1)]), (True, [mk(c)])])   # AFK
    m3 = rp.new_match("blitz", [(False, [mk(d)]), (False, [mk(a)]), (True, [mk(b)])])    # 3 rosters
    m4 = rp.new_match("5v5_ranked", [(False, [mk(d), mk(a), mk(b), mk(c)]), (True, [mk(a)])])
    bt = rater_mod.ingest([m0, m1, m2, m3, m4])
    assert bt.slots == 10 and bt.ids.shape == (5, 10)
    E = L.EMPTY_SLOT
    assert list(bt.ids[0]) == [0, 1, E, E, E, 2, 3, E, E, E]
    assert list(bt.ids[4]) == [3, 0, 1, 2, E, 0, E, E, E, E]
    assert np.all(bt.ids[1:4] == E)
    assert list(bt.meta) == [1 | (1 << 4), 15 << 4, 4 | (0 << 4), 8 | (2 << 4), 2 | (5 << 4)]
    assert bt.table.shape == (4, 8, 2)
    assert np.isnan(bt.table[0]).all() and list(bt.table[3, 0]) == [2000.0, 100.0]
    assert list(bt.table[3, 2]) == [1900.0, 90.0] and np.isnan(bt.table[3, 1]).all()
    assert tuple(bt.seed[0]) == rp.seed_from_fields(None, None, 3) and np.isnan(bt.seed[3]).all()
    with pytest.raises(ValueError):
        rater_mod.ingest([rp.new_match("ranked", [(True, [mk(a)] * 6), (False, [mk(b)])])])
    bad = rp.new_player(skill_tier=None)
    with pytest.raises(KeyError):
        rater_mod.ingest([rp.new_match("ranked", [(True, [mk(bad)]), (False, [mk(b)])])])


def test_pack_ids_bit_stream_round_trip():
    """engine.pack_ids: ids at ceil(log2(U + 1)) bits each, little-endian bit stream, all ones = empty slot."""
    from analyzer_b200 import engine
    rng = np.random.default_rng(3)
    for u in (2, 5, 255, 256, 1000, 1 << 20, 1_000_000, 12_500_000):
        bits = engine.id_bits(u)
        assert (1 << bits) - 1 >= u and (bits == 1 or (1 << (bits - 1)) - 1 < u)
        n, p = 1003, 6
        ids = rng.integers(0, u, size=(n, p)).astype(np.uint32)
        ids[rng.random((n, p)) < 0.05] = 0xFFFFFFFF
        ids[0, 0], ids[1, 1] = 0, u - 1
        pk = engine.pack_ids(ids, u)
        assert pk.bits == bits and pk.shape == (n, p) and pk.nbytes == (n * p * bits + 7) // 8
        stream = np.unpackbits(pk.packed.numpy(), bitorder="little")[:n * p * bits].reshape(n * p, bits)
        vals = (stream.astype(np.uint64) << np.arange(bits, dtype=np.uint64)).sum(axis=1).astype(np.uint32)
        vals[vals == (1 << bits) - 1] = 0xFFFFFFFF
        assert np.array_equal(vals.reshape(n, p), ids)


def test_ingest_rows_equals_object_ingest():
    """The row-stream ingest (numpy, no per-object Python) builds the same columns as the walk over
    the object graphs: ids through the player keys, meta, table and seeds row by row."""
    from analyzer_b200 import rater
    from tests import helpers as H
    rng = np.random.default_rng(9)
    n_players = 80
    tier, ranked, blitz = H.random_seed_fields(rng, n_players)
    ids, meta = H.small_ladder(rng, 400, n_players, slots=10, modes=(0, 1, 2, 3, 4, 5), p_draw=0.1, p_afk=0.1,
                               p_invalid=0.0, p_unsupported=0.05, ragged=True)
    table = np.full((n_players, H.N_COLS, 2), np.nan)
    table[::3, H.SHARED_COL] = (1650.0, 180.0)
    table[::6, 2] = (1700.0, 150.0)
    matches = H.run_objects(lambda m: None, ids, meta, table, tier, ranked, blitz)["matches"]
    a = rater.ingest(matches)
    rows, _ = H.rows_from_matches(matches)
    b = rater.ingest_rows(rows)                      # the native routine ...
    c = rater.ingest_rows_numpy(rows)                # ... and its numpy twin: the same columns
    assert b.slots == c.slots and np.array_equal(b.ids, c.ids) and np.array_equal(b.meta, c.meta)
    assert np.array_equal(b.table, c.table, equal_nan=True) and np.array_equal(b.seed, c.seed, equal_nan=True)
    assert np.array_equal(b.player_keys, c.player_keys) and np.array_equal(b.match_keys, c.match_keys)
    assert np.array_equal(b.row_match, c.row_match) and np.array_equal(b.row_slot, c.row_slot)
    assert np.array_equal(b.row_player, c.row_player)
    assert b.slots == a.slots and np.array_equal(a.meta, b.meta)
    key_of_a = np.array([id(p) for p in a.players] + [0], dtype=np.int64)
    key_of_b = np.r_[b.player_keys, 0]
    ka = key_of_a[np.where(a.ids == 0xFFFFFFFF, len(a.players), a.ids)]
    kb = key_of_b[np.where(b.ids == 0xFFFFFFFF, len(b.player_keys), b.ids)]
    assert np.array_equal(ka, kb)
    row_of = {int(k): i for i, k in enumerate(b.player_keys)}
    modes_seen = sorted(set(int(m >> 4) & 15 for m in a.meta if not m & 12 and (int(m >> 4) & 15) < 6))
    for i, p in enumerate(a.players):
        j = row_of[id(p)]
        for c in [0] + [1 + m for m in modes_seen]:            # the object ingest loads the rated queues only
            assert np.array_equal(a.table[i, c], b.table[j, c], equal_nan=True), (i, c)
        if np.isnan(a.table[i, 0, 0]):
            assert np.array_equal(a.seed[i], b.seed[j], equal_nan=True)
