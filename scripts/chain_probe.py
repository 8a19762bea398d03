"""A ladder that is one long chain (a handful of players, every match waits for the one before):
the one-GPU stream kernel against the emulated sharded ladder, both delivery modes.  What it showed
(B200, 320k matches over 16 players): 6.2 us per match on one GPU -- thousands of warps poll the same
16 rows -- and 4.0 / 4.3 us emulating two ranks; such a ladder of millions of matches runs into the
watchdog (TSRATE_STREAM_MAX_SPINS) on two real GPUs.  Usage: chain_probe.py n_matches n_players world"""
import sys, time, torch
sys.path.insert(0, "/root/repo")
from analyzer_b200 import engine, sharded
import os
dev = torch.device("cuda", 0)
n, u, slots = int(sys.argv[1]), int(sys.argv[2]), 6
world = int(sys.argv[3])
r = engine.Rater(engine.Env(), 0)
ids, meta = engine.gen_matches(engine.make_gen(7, u, slots=slots, mode=1), 0, n, dev)
full = engine.gen_table(7, 0, u, dev).reshape(u, 2)
single = full.clone().reshape(u, 1, 2)
pt = engine.PlayerTable.ladder(single, mode=1)
pred = r.link_pred(ids, meta, u, pt.mode_cols)
torch.cuda.synchronize(); t0 = time.perf_counter()
r.rate_stream(pt, ids, meta, pred)
torch.cuda.synchronize(); t1 = time.perf_counter()
print("one gpu rate ms", (t1 - t0) * 1e3, "per match us", (t1 - t0) * 1e6 / n, flush=True)
for direct in ("1", "0"):
    os.environ["TSRATE_SHARD_DIRECT"] = direct
    lad = sharded.ShardedLadder(r, u, n, slots, 0, world, emulate=True)
    lad.set_stream(ids, meta); lad.set_table(full)
    lad.reset(); lad.barrier(); lad.link(n, 1); lad.check(); lad.barrier()
    t0 = time.perf_counter()
    lad.rate(n, 1); lad.barrier()
    t1 = time.perf_counter()
    try:
        lad.check()
        got = sharded.merge_tables([lad.table[g] for g in range(world)])
        print("emulate", world, "direct", direct, "rate ms", (t1 - t0) * 1e3, "equal", torch.equal(got, single.reshape(u, 2)), flush=True)
    except Exception as e:
        print("emulate", world, "direct", direct, "rate ms", (t1 - t0) * 1e3, "ERROR", e, flush=True)
    lad.close()
