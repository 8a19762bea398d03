"""Latency of one reference-sized batch (worker.py:18, BATCHSIZE = 500 matches): the columnar
call with host buffers (all of rate_match's outputs) against the faithful 8-pair table."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from analyzer_b200 import engine

dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
for n, u in ((500, 2500), (5000, 20000), (50000, 100000)):
    gen = engine.make_gen(0xA11CE0AA, u, slots=6, mode=1)
    ids_d, meta_d = engine.gen_matches(gen, 0, n, dev)
    ids = ids_d.cpu().numpy().view(np.uint32)
    meta = meta_d.cpu().numpy().view(np.uint32).copy()
    meta = (meta & ~np.uint32(0xF0)) | ((np.arange(n, dtype=np.uint32) % 3) << 4)      # casual / ranked / blitz
    table = np.full((u, 8, 2), np.nan)
    seeds = np.stack([np.full(u, 1800.0), np.full(u, 500.0)], 1)
    for path in ("auto", "rounds", "stream"):
        best = 1e9
        for it in range(12):
            t0 = time.perf_counter()
            res = r.rate_batch(ids, meta, table, seed=seeds, path=path)
            best = min(best, time.perf_counter() - t0)
        print("%6d matches, %6d players, path=%-6s: %.3f ms per batch (%.3g matches/s), rounds=%s" % (
            n, u, path, best * 1e3, n / best, res["n_rounds"]), flush=True)

# ---- the object API on the same batch size: worker.py's own call pattern (one rate_match per
# match, worker.py:192) and the batch call (rate_matches), with its three parts ---------------
import logging
from analyzer_b200 import rater


class Row(object):
    def __init__(self, **kw):
        self.__dict__.update(kw)


QUEUES = ["trueskill_" + m for m in ("casual", "ranked", "blitz", "br", "5v5_casual", "5v5_ranked")]


def graphs(n, u, seed):
    rng = np.random.default_rng(seed)

    def player(i):
        p = Row(api_id=str(i), skill_tier=i % 30, rank_points_ranked=None, rank_points_blitz=None,
                trueskill_mu=None, trueskill_sigma=None)
        for c in QUEUES:
            setattr(p, c + "_mu", None)
            setattr(p, c + "_sigma", None)
        return p

    players = [player(i) for i in range(u)]
    out = []
    for m in range(n):
        who = rng.choice(u, size=6, replace=False)
        rosters = []
        for t in range(2):
            parts = []
            for i in who[3 * t:3 * t + 3]:
                items = Row(any_afk=None)
                for c in QUEUES:
                    setattr(items, c + "_mu", None)
                    setattr(items, c + "_sigma", None)
                parts.append(Row(player=[players[i]], participant_items=[items], went_afk=0, trueskill_mu=None,
                                 trueskill_sigma=None, trueskill_delta=None))
            rosters.append(Row(winner=(t == m % 2), participants=parts))
        out.append(Row(api_id=str(m), game_mode=("casual", "ranked", "blitz")[m % 3], rosters=rosters,
                       participants=rosters[0].participants + rosters[1].participants, trueskill_quality=None))
    return out


rater.logger.setLevel(logging.WARNING)
rater.rate_matches(graphs(500, 2500, 2))            # warm-up
for rep in range(2):
    ms = graphs(500, 2500, 10 + rep)
    t0 = time.perf_counter()
    b = rater.ingest(ms)
    t1 = time.perf_counter()
    res = rater._rater(0).rate_batch(b.ids, b.meta, b.table, seed=b.seed, shared_col=0, mode_cols=(1, 2, 3, 4, 5, 6),
                                     want_outputs=True)
    t2 = time.perf_counter()
    rater._write_back(b, res)
    t3 = time.perf_counter()
    print("rate_matches(500 match graphs, %d players): %.2f ms = ingest %.2f + rate_batch %.2f + write-back %.2f" % (
        len(b.players), (t3 - t0) * 1e3, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3))
ms = graphs(300, 2500, 3)
t0 = time.perf_counter()
for m in ms:
    rater.rate_match(m)
dt = time.perf_counter() - t0
print("rate_match per match (worker.py unchanged), 300 calls: %.3f ms per call (%.3g matches/s)" % (dt / 300 * 1e3, 300 / dt))

# the same 500-match batch as a row stream (what the query of worker.py:176-191 yields, one row per
# participant): numpy ingest, GPU call, numpy updates -- no per-object Python
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import helpers as H
for rep in range(3):
    ms = graphs(500, 2500, 20 + rep)
    for m in ms:                      # helpers.rows_from_matches reads the port's attribute names
        for r_ in m.rosters:
            for p_ in r_.participants:
                for c in ("trueskill_5v5_casual", "trueskill_5v5_ranked"):
                    for o in (p_.player[0], p_.participant_items[0]):
                        if not hasattr(o, c + "_mu"):
                            setattr(o, c + "_mu", None); setattr(o, c + "_sigma", None)
    rows, _ = H.rows_from_matches(ms)
    t0 = time.perf_counter()
    b = rater.ingest_rows(rows)
    t1 = time.perf_counter()
    res = rater._rater(0).rate_batch(b.ids, b.meta, b.table, seed=b.seed, shared_col=0, mode_cols=(1, 2, 3, 4, 5, 6),
                                     want_outputs=True)
    t2 = time.perf_counter()
    up = rater.updates(b, res)
    t3 = time.perf_counter()
    print("rate_rows(500 matches as %d participant rows): %.2f ms = ingest_rows %.2f + rate_batch %.2f + updates %.2f" % (
        b.n_rows, (t3 - t0) * 1e3, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3))
