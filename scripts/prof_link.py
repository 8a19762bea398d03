"""The link phases once, for ncu: the one-GPU sort + pred of a ladder (one window of it), then the
sharded link of the same stream with both ranks of a 2-way ladder played on this GPU.
Usage: prof_link.py N U SLOTS"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from analyzer_b200 import engine, sharded

n, u, slots = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
dev = torch.device("cuda", 0)
r = engine.Rater(engine.Env(), 0)
gen = engine.make_gen(0xA11CE003, u, slots=slots, mode=1)
ids, meta = engine.gen_matches(gen, 0, n, dev)
pred = r.link_pred(ids, meta, u, [-1, 0])
torch.cuda.synchronize()
lad = sharded.ShardedLadder(r, u, n, slots, 0, 2, emulate=True)
lad.set_stream(ids, meta)
lad.set_table(engine.gen_table(1, 0, u, dev).reshape(u, 2))
lad.reset()
lad.link(n, 1)
lad.rate(n, 1)
torch.cuda.synchronize()
print("ok", lad.status(), r.error_count())
