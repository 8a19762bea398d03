"""Slot-indexed form, phase by phase: ts_link_succ and ts_rate_slots per window (CUDA events).
Usage: slots_phases.py N U SLOTS WINDOW"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
from analyzer_b200 import engine, _lib as L
from analyzer_b200.engine import _ptr, _stream_ptr

n, u, slots, window = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3]), int(float(sys.argv[4]))
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
ids, meta = engine.gen_matches(engine.make_gen(0xA11CE003, u, slots=slots, mode=1), 0, n, dev)
tab0 = engine.gen_table(0xA11CE003, 0, u, dev)
tab = tab0.clone()
pt = engine.PlayerTable.ladder(tab, mode=1)
desc = pt.desc()
lib = L.lib()
rows = torch.zeros(window * slots * 32, dtype=torch.uint8, device=dev)
succ = torch.empty(window * slots, dtype=torch.int32, device=dev)
ws = torch.empty(L.check(lib.ts_link_workspace_bytes(window, slots)), dtype=torch.uint8, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
epoch = 0
for rep in range(3):
    tab.copy_(tab0)
    tl = tr = 0.0
    evs = []
    for lo in range(0, n, window):
        hi = min(n, lo + window)
        epoch += 1
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        L.check(lib.ts_link_succ(r._ctx, _ptr(ids[lo:hi]), _ptr(meta[lo:hi]), hi - lo, slots, C.byref(desc), epoch, _ptr(succ),
                                 _ptr(rows), _ptr(ws), ws.numel(), _stream_ptr()))
        e1.record()
        L.check(lib.ts_rate_slots(r._ctx, C.byref(desc), _ptr(ids[lo:hi]), _ptr(meta[lo:hi]), _ptr(succ), hi - lo, slots, epoch,
                                  None, 0, _ptr(rows), _stream_ptr()))
        e2.record()
        evs.append((e0, e1, e2))
    torch.cuda.synchronize()
    tl = sum(a.elapsed_time(b) for a, b, c in evs)
    tr = sum(b.elapsed_time(c) for a, b, c in evs)
    print("window %g: %d windows, link %.3f ms, rate %.3f ms, total %.3f ms, errors %d" % (window, len(evs), tl, tr, tl + tr, r.error_count()), flush=True)
