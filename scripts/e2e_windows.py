"""End-to-end time of Rater.rate_batch (pinned host buffers in, table out) against the number
of pipeline windows, with the pure copy time of the same bytes for comparison."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine

n, u = 10_000_000, 1_000_000
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE002, u, slots=6, mode=1)
ids_d, meta_d = engine.gen_matches(gen, 0, n, dev)
import numpy as np
ids_h = ids_d.cpu().pin_memory()
ids_pk = engine.pack_ids(ids_d.cpu().numpy().view(np.uint32), u, pin=True)
meta_h = meta_d.cpu().to(torch.uint8).pin_memory()
tab_h = engine.gen_table(0xA11CE002, 0, u, dev).cpu().pin_memory()
out_h = torch.empty_like(tab_h).pin_memory()
buf = torch.empty_like(ids_d)
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    buf.copy_(ids_h, non_blocking=True); torch.cuda.synchronize()
    print("H2D of ids alone (240 MB): %.3f ms" % ((time.perf_counter() - t0) * 1e3))
for form, ids_in in (("ids packed at %d bits" % ids_pk.bits, ids_pk), ("ids as 32-bit words", ids_h)):
  print(form)
  for windows in (1, 2, 3, 4, 6, 8, 12, 16):
    best = 1e9
    for it in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r.rate_batch(ids_in, meta_h, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, table_out=out_h,
                     windows=windows)
        best = min(best, time.perf_counter() - t0)
    print("windows=%2d: %.3f ms  %.3g matches/s" % (windows, best * 1e3, n / best), flush=True)

print("uneven windows (fractions of the batch), packed ids")
for fr in ([1, 3, 3, 3], [1, 2, 3, 4], [1, 2, 2, 2, 3], [0.5, 1.5, 2.5, 2.5, 3], [1, 3, 6], [1, 4, 5], [2, 3, 5], [0.5, 1, 2, 3, 3.5]):
    best = 1e9
    for it in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r.rate_batch(ids_pk, meta_h, tab_h, shared_col=-1, mode_cols=[-1, 0], want_outputs=False, table_out=out_h, windows=fr)
        best = min(best, time.perf_counter() - t0)
    print("windows=%s: %.3f ms  %.3g matches/s" % (fr, best * 1e3, n / best), flush=True)
