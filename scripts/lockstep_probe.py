"""Where does the per-level time of the streaming kernel go?  G disjoint groups of 6
players play round-robin: every level has G matches, and every match depends on exactly
one earlier match (the previous one of its own group) -- no max over six different
predecessors.  Compare the time per level with config 2's (same width, same depth)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from analyzer_b200 import engine

dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
for G in (1, 64, 1024, 11264, 45056):
    levels = max(200, min(888, 10_000_000 // G)) if G > 1 else 20000
    n = G * levels
    m = np.arange(n, dtype=np.int64)
    g = (m % G).astype(np.uint32)
    ids = (g[:, None] * 6 + np.arange(6, dtype=np.uint32)[None, :]).astype(np.uint32)
    meta = np.full(n, 1 | (1 << 4), dtype=np.uint32)
    u = G * 6
    ids_d = torch.from_numpy(ids.view(np.int32)).to(dev)
    meta_d = torch.from_numpy(meta.view(np.int32)).to(dev)
    tab = engine.gen_table(1, 0, u, dev)
    pt = engine.PlayerTable.ladder(tab, mode=1)
    pred = r.link_pred(ids_d, meta_d, u, pt.mode_cols)
    best = 1e9
    for it in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r.rate_stream(pt, ids_d, meta_d, pred)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("groups=%6d levels=%5d matches=%8d: %8.3f ms  %.3f us per level  %.3g matches/s" % (
        G, levels, n, best, best * 1e3 / levels, n / best * 1e3), flush=True)
print("errors", r.error_count())
