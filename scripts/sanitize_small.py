"""Small end-to-end run of every kernel for compute-sanitizer (memcheck / racecheck)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from analyzer_b200 import engine

dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(draw_probability=0.02), 0)
n, u = 6000, 400
for slots in (6, 10):
    gen = engine.make_gen(7, u, slots=slots, mode=1, p_draw=0.05, p_afk=0.05, zipf=(slots == 10))
    ids, meta = engine.gen_matches(gen, 0, n, dev)
    a = engine.gen_table(7, 0, u, dev)
    b = a.clone()
    pa, pb = engine.PlayerTable.ladder(a, mode=1), engine.PlayerTable.ladder(b, mode=1)
    sch = r.schedule(ids, meta, u, pa.mode_cols)
    r.rate(pa, sch, r.alloc_outputs(n, slots))
    r.rate_stream_windows(pb, ids, meta, 2500)
    print(slots, "rounds", sch.n_rounds, "max diff", float((a - b).abs().max()), "errors", r.error_count())
# faithful layout with NULLs and seeds
ft = engine.PlayerTable.faithful(u, dev, seed=torch.full((u, 2), 1500.0, dtype=torch.float64, device=dev))
ft.seed[:, 1] = 400.0
gen = engine.make_gen(9, u, slots=6, mode=2)
ids, meta = engine.gen_matches(gen, 0, n, dev)
sch = r.schedule(ids, meta, u, ft.mode_cols)
r.rate(ft, sch, r.alloc_outputs(n, 6))
print("faithful ok", r.error_count())
