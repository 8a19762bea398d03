"""A/B of the streaming kernel's memory options on one ladder: poll loads hoisted or row by
row (TSRATE_STREAM_HOIST) x next tile's rows prefetched into L2 or not (TSRATE_STREAM_PREFETCH),
link once, rate phase best of 3.  Usage: ab_stream.py N U SLOTS [zipf] [p_draw] [ladder|pair]"""
import itertools, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u, slots = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
zipf = len(sys.argv) > 4 and sys.argv[4] == "zipf"
p_draw = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
layout = sys.argv[6] if len(sys.argv) > 6 else "ladder"
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(draw_probability=p_draw), 0)
gen = engine.make_gen(0xA11CE003, u, slots=slots, mode=1, p_draw=p_draw, zipf=zipf)
ids, meta = engine.gen_matches(gen, 0, n, dev)
base = engine.gen_table(0xA11CE003, 0, u, dev)
if layout == "pair":
    tab0 = torch.full((u, 2, 2), float("nan"), dtype=torch.float64, device=dev)
    tab0[:, 0] = base.reshape(u, 2)
    tab = tab0.clone()
    pt = engine.PlayerTable(tab, 0, [-1, 1])
else:
    tab0 = base
    tab = tab0.clone()
    pt = engine.PlayerTable.ladder(tab, mode=1)
pred = r.link_pred_windows(ids, meta, u, pt.mode_cols)
ev = lambda: torch.cuda.Event(enable_timing=True)
ref = None
for hoist, pf in itertools.product(("0", "1"), ("0", "1")):
    if layout == "pair" and hoist == "1":
        continue
    os.environ["TSRATE_STREAM_HOIST"], os.environ["TSRATE_STREAM_PREFETCH"] = hoist, pf
    best = 1e9
    for it in range(3):
        tab.copy_(tab0)
        e0, e1 = ev(), ev()
        e0.record()
        r.rate_stream(pt, ids, meta, pred)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    if ref is None:
        ref = tab.clone()
    same = torch.equal(torch.nan_to_num(tab, nan=-1.0), torch.nan_to_num(ref, nan=-1.0))
    print("%s n=%g u=%g slots=%d %s hoist=%s prefetch=%s rate %.3f ms  same_bits=%s errors=%d" % (
        layout, n, u, slots, "zipf" if zipf else "uniform", hoist, pf, best, same, r.error_count()), flush=True)
