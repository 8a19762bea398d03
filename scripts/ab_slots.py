"""Player-indexed rows (ts_link_pred + ts_rate_stream) against slot-indexed rows (ts_link_succ +
ts_rate_slots) on one ladder: link and rate phase of each, best of 3, tables compared bit for bit.
Usage: ab_slots.py N U SLOTS [zipf] [p_draw] [window]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
from analyzer_b200 import engine, _lib as L

n, u, slots = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
zipf = len(sys.argv) > 4 and sys.argv[4] == "zipf"
p_draw = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
window = int(float(sys.argv[6])) if len(sys.argv) > 6 else None
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(draw_probability=p_draw), 0)
gen = engine.make_gen(0xA11CE003, u, slots=slots, mode=1, p_draw=p_draw, zipf=zipf)
ids, meta = engine.gen_matches(gen, 0, n, dev)
tab0 = engine.gen_table(0xA11CE003, 0, u, dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
res = {}
for path in ("stream", "slots"):
    tab = tab0.clone()
    pt = engine.PlayerTable.ladder(tab, mode=1)
    best = None
    for it in range(4):
        tab.copy_(tab0)
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        if path == "stream":
            pred = r.link_pred_windows(ids, meta, u, pt.mode_cols)
            e1.record()
            r.rate_stream(pt, ids, meta, pred)
            del pred
        else:
            r.rate_slots(pt, ids, meta, window=window)
            e1.record()
        e2.record()
        torch.cuda.synchronize()
        t = (e0.elapsed_time(e1), e1.elapsed_time(e2))
        if it and (best is None or sum(t) < sum(best)):
            best = t
    res[path] = tab
    print("%-6s n=%g u=%g slots=%d %s: %s total %.3f ms  (%.3g matches/s)  errors %d" % (
        path, n, u, slots, "zipf" if zipf else "uniform",
        "link %.3f + rate %.3f" % best if path == "stream" else "link + rate interleaved per window", sum(best), n / sum(best) * 1e3,
        r.error_count()), flush=True)
print("tables bit-equal:", torch.equal(res["stream"], res["slots"]))
