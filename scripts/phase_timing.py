"""Run each phase of both paths once with a sync and a printed timing, so a
stall shows up by name.  Usage: phase_timing.py N U SLOTS [zipf] [p_draw]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u, slots = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
zipf = len(sys.argv) > 4 and sys.argv[4] == "zipf"
p_draw = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(draw_probability=p_draw), 0)


def phase(name, fn):
    torch.cuda.synchronize()
    t = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    print("%-14s %10.3f ms" % (name, (time.perf_counter() - t) * 1e3), flush=True)
    return out


gen = engine.make_gen(0xA11CE003, u, slots=slots, mode=1, p_draw=p_draw, zipf=zipf)
ids, meta = phase("gen", lambda: engine.gen_matches(gen, 0, n, dev))
tab = phase("gen_table", lambda: engine.gen_table(0xA11CE003, 0, u, dev))
tab2 = tab.clone()
pt, pt2 = engine.PlayerTable.ladder(tab, mode=1), engine.PlayerTable.ladder(tab2, mode=1)
pred = phase("link_pred", lambda: r.link_pred(ids, meta, u, pt.mode_cols))
phase("rate_stream", lambda: r.rate_stream(pt, ids, meta, pred))
if os.environ.get("SKIP_ROUNDS") != "1":
    sch = phase("schedule", lambda: r.schedule(ids, meta, u, pt2.mode_cols))
    print("rounds", sch.n_rounds, flush=True)
    phase("rate_rounds", lambda: r.rate(pt2, sch))
    print("max diff", float((tab - tab2).abs().max()), flush=True)
print("errors", r.error_count())
