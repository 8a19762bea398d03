"""Generates the golden fixtures in this directory by running the UNMODIFIED
reference /root/reference/rater.py (on the trueskill-API shim in oracle/shim,
mpmath backend at 50 digits -- the pinned trueskill==0.4.4 package is not
installable here) over synthetic ladders in created_at order, the way
worker.py:176-192 drives it.

Run in the build container only (it needs /root/reference):
    python tests/golden/make_golden.py
Outputs (committed): mixed_small.npz, config1.npz
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "shim"))
sys.path.insert(0, "/root/reference")

import logging  # noqa: E402
import rater as ref_rater  # noqa: E402  the unmodified reference module
ref_rater.logger.setLevel(logging.ERROR)

from oracle import c_oracle as co  # noqa: E402
from tests import helpers as H  # noqa: E402


def mixed_small():
    rng = np.random.default_rng(0xA11CE0AA)
    n_players, n = 90, 500
    tier, ranked, blitz = H.random_seed_fields(rng, n_players)
    ids6, meta6 = H.small_ladder(rng, n // 2, n_players, slots=6, modes=(0, 1, 2, 3), p_draw=0.05,
                                 p_afk=0.05, p_invalid=0.02, p_unsupported=0.02, ragged=True)
    ids10, meta10 = H.small_ladder(rng, n // 2, n_players, slots=10, modes=(4, 5), p_draw=0.05,
                                   p_afk=0.05, ragged=True)
    # one stream with 10 slots per match; 3v3 matches leave slots empty
    ids = np.full((n, 10), H.EMPTY, dtype=np.uint32)
    meta = np.empty(n, dtype=np.uint32)
    pick = rng.permutation(n)
    for k, m in enumerate(pick):
        if k < n // 2:
            ids[m, 0:3], ids[m, 5:8], meta[m] = ids6[k, 0:3], ids6[k, 3:6], meta6[k]
        else:
            ids[m], meta[m] = ids10[k - n // 2], meta10[k - n // 2]
    table = np.full((n_players, H.N_COLS, 2), np.nan)
    # a third of the players come in with stored ratings
    for i in range(0, n_players, 3):
        table[i, H.SHARED_COL] = (1500 + 20 * i, 150 + i)
        if i % 2 == 0:
            table[i, H.MODE_COLS[1]] = (1400 + 25 * i, 120 + i)
    out = H.run_objects(ref_rater.rate_match, ids, meta, table, tier, ranked, blitz)
    np.savez_compressed(os.path.join(HERE, "mixed_small.npz"), ids=ids, meta=meta, table_in=table,
                        tier=tier, ranked=ranked, blitz=blitz, seed=H.seed_table(tier, ranked, blitz),
                        table_out=out["table"], quality=out["quality"], out_shared=out["out_shared"],
                        out_delta=out["out_delta"], out_queue=out["out_queue"])


def config1():
    """BASELINE.json configs[0]: 10k 3v3 ranked matches over 5k players."""
    n, n_players, seed = 10_000, 5_000, 0xA11CE001
    gen = co.make_gen(seed, n_players, slots=6, mode=1)
    ids, meta = co.gen_matches(gen, 0, n)
    rng = np.random.default_rng(seed)
    tier = rng.integers(0, 30, n_players)
    none = np.full(n_players, -1)
    table = np.full((n_players, H.N_COLS, 2), np.nan)
    t0 = time.time()
    out = H.run_objects(ref_rater.rate_match, ids, meta, table, tier, none, none)
    dt = time.time() - t0
    print("config1: %d matches in %.1f s = %.0f matches/s (1 core)" % (n, dt, n / dt))
    np.savez_compressed(os.path.join(HERE, "config1.npz"), gen_seed=seed, tier=tier,
                        ids_head=ids[:64], meta_head=meta[:64],
                        table_out=out["table"][:, [H.SHARED_COL, H.MODE_COLS[1]]],
                        quality=out["quality"], ref_matches_per_s=n / dt)


if __name__ == "__main__":
    mixed_small()
    config1()
