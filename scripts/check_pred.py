"""Device-side sanity check of ts_link_pred at scale: every pred is an earlier
match that contains the same player, and first appearances have TS_NONE."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u, slots = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
zipf = len(sys.argv) > 4 and sys.argv[4] == "zipf"
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE003, u, slots=slots, mode=1, p_draw=0.02, zipf=zipf)
ids, meta = engine.gen_matches(gen, 0, n, dev)
cols = [-1, 0]
pred = r.link_pred(ids, meta, u, cols)
torch.cuda.synchronize()
p64 = pred.to(torch.int64) & 0xFFFFFFFF
m_idx = torch.arange(n, device=dev).unsqueeze(1).expand(n, slots)
has = p64 != 0xFFFFFFFF
print("slots with pred:", int(has.sum()), "of", n * slots, flush=True)
print("pred < m everywhere:", bool((p64[has] < m_idx[has]).all()), flush=True)
bad_ge = int((p64[has] >= m_idx[has]).sum())
print("pred >= m count:", bad_ge, flush=True)
# the pred match must contain the player
chunk = 2_000_000
bad = 0
for lo in range(0, n, chunk):
    hi = min(n, lo + chunk)
    pp = p64[lo:hi].clamp(max=n - 1)
    rows = ids[pp.reshape(-1)].reshape(hi - lo, slots, slots)          # ids of the pred match of each slot
    same = (rows == ids[lo:hi].unsqueeze(2)).any(dim=2)
    bad += int((~same & has[lo:hi]).sum())
print("pred match lacks the player:", bad, flush=True)
# first appearance <=> NONE: count of NONE must equal number of distinct players seen
print("NONE count", int((~has).sum()), "distinct players", int(torch.unique(ids).numel()), flush=True)
