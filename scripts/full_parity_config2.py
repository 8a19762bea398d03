"""BASELINE.json configs[1] at full size (10M 3v3 matches, 1M players): final table
of both GPU paths against the sequential __float128 oracle, and the schedule
against the sequential schedule oracle.  CPU side takes about a minute."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from analyzer_b200 import engine
from oracle import c_oracle as co

n, u, seed = 10_000_000, 1_000_000, 0xA11CE002
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(seed, u, slots=6, mode=1)
ids_d, meta_d = engine.gen_matches(gen, 0, n, dev)
tab_s = engine.gen_table(seed, 0, u, dev)
tab_r = tab_s.clone()
want = tab_s.cpu().numpy()
ps, pr = engine.PlayerTable.ladder(tab_s, mode=1), engine.PlayerTable.ladder(tab_r, mode=1)
r.rate_stream(ps, ids_d, meta_d, r.link_pred(ids_d, meta_d, u, ps.mode_cols))
sch = r.schedule(ids_d, meta_d, u, pr.mode_cols)
r.rate(pr, sch)
assert r.error_count() == 0
ids, meta = ids_d.cpu().numpy().view(np.uint32), meta_d.cpu().numpy().view(np.uint32)
t0 = time.time()
rnd, order, offs, R = co.schedule(ids, meta, pr.mode_cols, u)
print("schedule oracle %.1f s, R = %d" % (time.time() - t0, R), flush=True)
print("round  bit-exact:", np.array_equal(sch.round.cpu().numpy().view(np.uint32), rnd))
print("order  bit-exact:", np.array_equal(sch.order.cpu().numpy().view(np.uint32), order))
print("offsets bit-exact:", np.array_equal(sch.round_offsets.cpu().numpy().view(np.uint32)[:R + 2], offs), flush=True)
t0 = time.time()
out = co.rate_sequential(co.make_params(), ids, meta, want, -1, pr.mode_cols[:6], quad=True)
print("float128 sequential oracle %.1f s, errors %d" % (time.time() - t0, out["errors"]), flush=True)
print("stream path max |delta| = %.3g" % np.abs(tab_s.cpu().numpy() - want).max())
print("rounds path max |delta| = %.3g" % np.abs(tab_r.cpu().numpy() - want).max())
