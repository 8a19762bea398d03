"""One link + one streaming rate of a big ladder (table far beyond L2), for ncu.
Usage: prof_big.py N U SLOTS [zipf] [p_draw] [ladder|pair]"""
import os, sys
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
    tab = torch.full((u, 2, 2), float("nan"), dtype=torch.float64, device=dev)
    tab[:, 0] = base.reshape(u, 2)
    pt = engine.PlayerTable(tab, 0, [-1, 1])
else:
    pt = engine.PlayerTable.ladder(base, mode=1)
pred = r.link_pred_windows(ids, meta, u, pt.mode_cols)
r.rate_stream(pt, ids, meta, pred)
torch.cuda.synchronize()
print("ok", layout, n, u, slots, "errors", r.error_count())
