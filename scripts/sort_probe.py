"""Link phase alone (radix sort by player + predecessor links), best of 5, CUDA events.
Usage: sort_probe.py N U SLOTS [window]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u, slots = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
ids, meta = engine.gen_matches(engine.make_gen(0xA11CE002, u, slots=slots, mode=1), 0, n, dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
best = 1e9
for it in range(6):
    flush.fill_(1)
    e0, e1 = ev(), ev()
    e0.record()
    pred = r.link_pred_windows(ids, meta, u, [-1, 0])
    e1.record()
    torch.cuda.synchronize()
    if it:
        best = min(best, e0.elapsed_time(e1))
print("%s link n=%g u=%g slots=%d: %.3f ms (%.1f G slot entries/s) checksum %d" % (
    os.path.basename(os.environ.get("TSRATE_LIB", "default")), n, u, slots, best, n * slots / best / 1e6, int(pred.long().sum())), flush=True)
