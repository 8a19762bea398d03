"""Timeline of the pipelined ingest of Rater.rate_batch: when each window's copy lands and when
its link and rate finish, on the device clock (events), for a given number of equal windows."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine
from analyzer_b200 import _lib as L

n, u = 10_000_000, 1_000_000
windows = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE002, u, slots=6, mode=1)
ids_s, meta_s = engine.gen_matches(gen, 0, n, dev)
ids_h = ids_s.cpu().pin_memory()
meta_h = meta_s.cpu().to(torch.uint8).pin_memory()
tab_h = engine.gen_table(0xA11CE002, 0, u, dev).cpu().pin_memory()
out_h = torch.empty_like(tab_h).pin_memory()
del ids_s, meta_s
ev = lambda: torch.cuda.Event(enable_timing=True)
for rep in range(3):
    cur = torch.cuda.current_stream(dev)
    side = torch.cuda.Stream(dev)
    torch.cuda.synchronize()
    t_host0 = time.perf_counter()
    e_start = ev(); e_start.record(cur)
    ids_d = torch.empty((n, 6), dtype=torch.int32, device=dev)
    meta_d = torch.empty((n,), dtype=torch.int32, device=dev)
    meta_in = torch.empty((n,), dtype=torch.uint8, device=dev)
    tab_d = tab_h.to(dev, non_blocking=True)
    ptab = engine.PlayerTable(tab_d, -1, [-1, 0])
    last = torch.full((u,), -1, dtype=torch.int32, device=dev)
    bounds = [(w * n) // windows for w in range(windows + 1)]
    side.wait_stream(cur)
    ready, t_copy, t_link, t_rate = [], [], [], []
    with torch.cuda.stream(side):
        for w in range(windows):
            lo, hi = bounds[w], bounds[w + 1]
            ids_d[lo:hi].copy_(ids_h[lo:hi], non_blocking=True)
            meta_in[lo:hi].copy_(meta_h[lo:hi], non_blocking=True)
            e = ev(); e.record(side); ready.append(e)
    t_host1 = time.perf_counter()
    for w in range(windows):
        lo, hi = bounds[w], bounds[w + 1]
        cur.wait_event(ready[w])
        r.widen_meta(meta_in[lo:hi], meta_d[lo:hi])
        flags = (L.STREAM_LOAD if w == 0 else 0) | (L.STREAM_STORE if w == windows - 1 else 0)
        pred = r.link_pred(ids_d[lo:hi], meta_d[lo:hi], u, ptab.mode_cols, seq_base=lo, last=last)
        e = ev(); e.record(cur); t_link.append(e)
        r.rate_stream(ptab, ids_d[lo:hi], meta_d[lo:hi], pred, None, 0, seq_base=lo, flags=flags)
        e = ev(); e.record(cur); t_rate.append(e)
    t_host2 = time.perf_counter()
    out_h.copy_(tab_d, non_blocking=True)
    e_end = ev(); e_end.record(cur)
    torch.cuda.synchronize()
    t_host3 = time.perf_counter()
    if rep == 2:
        print("windows=%d  host: setup+copies issued %.2f ms, all launches issued %.2f ms, done %.2f ms" % (
            windows, (t_host1 - t_host0) * 1e3, (t_host2 - t_host0) * 1e3, (t_host3 - t_host0) * 1e3))
        for w in range(windows):
            print("  window %d: copy landed %.2f  link done %.2f  rate done %.2f ms" % (
                w, e_start.elapsed_time(ready[w]), e_start.elapsed_time(t_link[w]), e_start.elapsed_time(t_rate[w])))
        print("  table back on the host: %.2f ms" % e_start.elapsed_time(e_end))
