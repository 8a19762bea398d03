"""Config 2 on HBM-resident inputs: one shot vs windows vs windows with the
link of window k+1 overlapped with the rating of window k."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u = 10_000_000, 1_000_000
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE002, u, slots=6, mode=1)
ids, meta = engine.gen_matches(gen, 0, n, dev)
tab0 = engine.gen_table(0xA11CE002, 0, u, dev)
tab = tab0.clone()
pt = engine.PlayerTable.ladder(tab, mode=1)
ref = None
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for nwin, overlap in ((1, False), (2, False), (4, False), (8, False)):
    best = 1e9
    for it in range(4):
        tab.copy_(tab0); flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r.rate_stream_windows(pt, ids, meta, -(-n // nwin))
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    if ref is None:
        ref = tab.clone()
    print("windows=%2d overlap=%-5s %7.3f ms  %.3g matches/s  equal=%s" % (nwin, overlap, best, n / best * 1e3, torch.equal(tab, ref)), flush=True)
print("errors", r.error_count())
