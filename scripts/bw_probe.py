"""Bandwidth-bound probe: few dependencies (players >> matches), table far larger
than L2.  Times the two rating kernels alone (schedule / links precomputed)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u = int(float(sys.argv[1])), int(float(sys.argv[2]))
lanes = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [1, 2]
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE002, u, slots=6, mode=1)
ids, meta = engine.gen_matches(gen, 0, n, dev)
tab = engine.gen_table(1, 0, u, dev)
pt = engine.PlayerTable.ladder(tab, mode=1)
sch = r.schedule(ids, meta, u, pt.mode_cols)
pred = r.link_pred(ids, meta, u, pt.mode_cols)
print("rounds", sch.n_rounds)
ev = lambda: torch.cuda.Event(enable_timing=True)
for L_ in lanes:
    for path in ("rounds", "stream"):
        best = 1e9
        for it in range(3):
            e0, e1 = ev(), ev()
            e0.record()
            if path == "rounds":
                r.rate(pt, sch, None, L_)
            else:
                r.rate_stream(pt, ids, meta, pred, None, L_)   # includes building / reading back the versioned rows
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print("lanes=%d %-6s %8.3f ms  %.3g matches/s  %.0f GB/s algorithmic (220 B/match)" % (L_, path, best, n / best * 1e3, 220 * n / best / 1e6), flush=True)
