"""Per-hop latency of the two rating paths on a pure chain (every match shares
player 0, so R = N and nothing overlaps): time / N = latency of one dependent
hop.  Run under gpurun."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from analyzer_b200 import engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
ids = np.zeros((n, 6), dtype=np.uint32)
ids[:, 1:] = np.arange(1, 1 + 5 * n, dtype=np.uint32).reshape(n, 5)
meta = np.full(n, 1 | (1 << 4), dtype=np.uint32)
u = 1 + 5 * n
ids_d = torch.from_numpy(ids.view(np.int32)).to(dev)
meta_d = torch.from_numpy(meta.view(np.int32)).to(dev)
tab = engine.gen_table(1, 0, u, dev)
pt = engine.PlayerTable.ladder(tab, mode=1)
sch = r.schedule(ids_d, meta_d, u, pt.mode_cols)
pred = r.link_pred(ids_d, meta_d, u, pt.mode_cols)
ev = lambda: torch.cuda.Event(enable_timing=True)
for lanes in (1, 2, 8, 32):
    for path in ("rounds", "stream"):
        best = 1e9
        for it in range(3):
            e0, e1 = ev(), ev()
            e0.record()
            if path == "rounds":
                r.rate(pt, sch, None, lanes)
            else:
                r.rate_stream(pt, ids_d, meta_d, pred, None, lanes)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print("lanes=%2d %-6s %8.3f ms  -> %.3f us per hop" % (lanes, path, best, best * 1e3 / n))
