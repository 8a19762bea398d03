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
