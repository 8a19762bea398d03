"""Link phase only (sort + pred) over sub-windows with carried chain tails."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u = 10_000_000, 1_000_000
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE002, u, slots=6, mode=1)
ids, meta = engine.gen_matches(gen, 0, n, dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ref = None
for nwin in (1, 2, 3, 4, 6, 8):
    best = 1e9
    for it in range(4):
        flush.fill_(1)
        last = torch.full((u,), -1, dtype=torch.int32, device=dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        preds = []
        w = -(-n // nwin)
        for lo in range(0, n, w):
            hi = min(n, lo + w)
            preds.append(r.link_pred(ids[lo:hi], meta[lo:hi], u, [-1, 0], seq_base=lo, last=last))
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    pred = torch.cat(preds)
    if ref is None:
        ref = pred
    print("link windows=%d %7.3f ms equal=%s" % (nwin, best, torch.equal(pred, ref)), flush=True)
