"""The rater port (oracle/rater_port.py) against (a) the golden fixtures made
by the unmodified reference and (b), when /root/reference is present (build
container only), the unmodified reference itself -- both bit-for-bit, since
port and reference run on the same arithmetic shim."""
import os
import sys

import numpy as np
import pytest

from oracle import rater_port as rp
from tests import helpers as H

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _same(a, b):
    return np.array_equal(a, b, equal_nan=True)


def test_port_reproduces_golden_mixed_small_prefix():
    g = np.load(os.path.join(GOLD, "mixed_small.npz"))
    out = H.run_objects(rp.rate_match, g["ids"], g["meta"], g["table_in"], g["tier"], g["ranked"], g["blitz"])
    assert _same(out["table"], g["table_out"])
    assert _same(out["quality"], g["quality"])
    assert _same(out["out_shared"], g["out_shared"])
    assert _same(out["out_delta"], g["out_delta"])
    assert _same(out["out_queue"], g["out_queue"])
    assert _same(H.seed_table(g["tier"], g["ranked"], g["blitz"]), g["seed"])


@pytest.mark.skipif(not os.path.exists("/root/reference/rater.py"), reason="reference tree not on this box")
def test_port_equals_unmodified_reference_live():
    sys.path.insert(0, "/root/reference")
    try:
        import logging
        import rater as ref_rater
        ref_rater.logger.setLevel(logging.ERROR)
    finally:
        sys.path.remove("/root/reference")
    assert ref_rater.vst_points == rp.vst_points
    assert ref_rater.env.beta == rp.env.beta and ref_rater.env.tau == rp.env.tau
    rng = np.random.default_rng(7)
    n_players = 40
    tier, ranked, blitz = H.random_seed_fields(rng, n_players)
    ids, meta = H.small_ladder(rng, 120, n_players, slots=10, modes=(0, 1, 2, 3, 4, 5), p_draw=0.1,
                               p_afk=0.1, p_invalid=0.05, p_unsupported=0.05, ragged=True)
    table = np.full((n_players, H.N_COLS, 2), np.nan)
    table[::4, H.SHARED_COL] = (1700.0, 220.0)
    a = H.run_objects(ref_rater.rate_match, ids, meta, table, tier, ranked, blitz)
    b = H.run_objects(rp.rate_match, ids, meta, table, tier, ranked, blitz)
    for k in ("table", "quality", "out_shared", "out_delta", "out_queue"):
        assert _same(a[k], b[k]), k
    for ma, mb in zip(a["matches"], b["matches"]):
        fa = [p.participant_items[0].any_afk for p in ma.participants]
        fb = [p.participant_items[0].any_afk for p in mb.participants]
        assert fa == fb
