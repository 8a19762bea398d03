"""Reference-faithful layout (K = 2: shared + queue pair per player, quality, delta)
at config-2 size through the round path and the streaming path."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u = 10_000_000, 1_000_000
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE002, u, slots=6, mode=1)
ids, meta = engine.gen_matches(gen, 0, n, dev)
one = engine.gen_table(0xA11CE002, 0, u, dev)                      # [U, 1, 2]
tab0 = torch.full((u, 2, 2), float("nan"), dtype=torch.float64, device=dev)
tab0[:, 0] = one[:, 0]                                               # shared stored, queue NULL
tab = tab0.clone()
pt = engine.PlayerTable(tab, 0, [-1, 1])
sch = r.schedule(ids, meta, u, pt.mode_cols)
ev = lambda: torch.cuda.Event(enable_timing=True)
for lanes in (2, 4):
    for outs in (None, r.alloc_outputs(n, 6)):
        best = 1e9
        for it in range(3):
            tab.copy_(tab0)
            e0, e1 = ev(), ev()
            e0.record()
            r.rate(pt, sch, outs, lanes)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print("K=2 rounds lanes=%d outputs=%s: %.3f ms  %.3g matches/s  (412 B/match: %.0f GB/s)" % (
            lanes, outs is not None, best, n / best * 1e3, 412 * n / best / 1e6), flush=True)
pred = r.link_pred(ids, meta, u, pt.mode_cols)
for lanes in (2, 4):
    for outs in (None, r.alloc_outputs(n, 6)):
        best = 1e9
        for it in range(3):
            tab.copy_(tab0)
            e0, e1 = ev(), ev()
            e0.record()
            r.rate_stream(pt, ids, meta, pred, outs, lanes)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print("K=2 stream lanes=%d outputs=%s: %.3f ms  %.3g matches/s  (412 B/match: %.0f GB/s)" % (
            lanes, outs is not None, best, n / best * 1e3, 412 * n / best / 1e6), flush=True)
print("errors", r.error_count())
