"""Randomised small ladders through the C ABI against the sequential float128
oracle: both rating paths, both record widths, ragged teams, draws with and
without margin, AFK / invalid / unsupported matches, unseeded players with and
without seeds, very few players (deep chains) to very many (no chains)."""
import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _dev(a):
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("case", range(24))
def test_random_ladder(case, c_oracle):
    from analyzer_b200 import engine
    rng = np.random.default_rng(1000 + case)
    slots = int(rng.choice([6, 10]))
    n = int(rng.choice([1, 2, 17, 300, 2500]))
    n_players = int(rng.choice([slots, slots + 1, 40, 500, 20000]))
    env_draw = float(rng.choice([0.0, 0.02, 0.3]))
    ids, meta = H.small_ladder(rng, n, n_players, slots=slots, modes=(1,), p_draw=float(rng.choice([0.0, 0.2])),
                               p_afk=float(rng.choice([0.0, 0.1])), p_invalid=0.03, p_unsupported=0.03,
                               ragged=bool(rng.integers(0, 2)))
    table = np.empty((n_players, 1, 2))
    table[:, 0, 0] = rng.normal(1500, 500, n_players)
    table[:, 0, 1] = rng.uniform(30, 900, n_players)
    unseeded = rng.random(n_players) < 0.3
    table[unseeded, 0] = np.nan
    seed = np.stack([rng.normal(1800, 300, n_players), rng.uniform(300, 600, n_players)], 1)
    want = table.copy()
    out = c_oracle.rate_sequential(c_oracle.make_params(1000.0, 10.0, env_draw), ids, meta, want, -1, [-1, 0],
                                   seed=seed, quad=True, want_outputs=True)
    r = engine.Rater(engine.Env(draw_probability=env_draw), 0)
    lanes = int(rng.choice([0, 1, 2, 4, 8, 16, 32]))
    for path in ("rounds", "stream"):
        tab_d, seed_d = _dev(table), _dev(seed)
        pt = engine.PlayerTable(tab_d, -1, [-1, 0], seed_d)
        outs = r.alloc_outputs(n, slots)
        r.reset_errors()
        ids_d, meta_d = _dev(ids), _dev(meta)
        if path == "rounds":
            r.rate(pt, r.schedule(ids_d, meta_d, n_players, pt.mode_cols), outs, lanes)
        else:
            r.rate_stream(pt, ids_d, meta_d, r.link_pred(ids_d, meta_d, n_players, pt.mode_cols), outs, lanes)
        assert r.error_count() == out["errors"]
        got = tab_d.cpu().numpy()
        assert np.array_equal(np.isnan(got), np.isnan(want)), path
        m = ~np.isnan(want)
        assert not m.any() or np.abs(got[m] - want[m]).max() < TOL, path
        st = outs["status"].cpu().numpy()
        assert np.array_equal(st, out["status"]), path
        rated = st == 0
        if rated.any():
            assert np.abs(outs["quality"].cpu().numpy()[rated] - out["quality"][rated]).max() < 1e-12
            oq, wq = outs["out_queue"].cpu().numpy()[rated], out["out_queue"][rated]
            assert np.array_equal(np.isnan(oq), np.isnan(wq)) and np.nanmax(np.abs(oq - wq)) < TOL
    r.reset_errors()
    r.close()
