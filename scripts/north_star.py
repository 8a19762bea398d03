"""BASELINE.json configs[3] and configs[4] at the sizes they state, on 8 B200s (torchrun, one
process per GPU):

  (a) 8 independent ladders, 1B matches / 100M players in all: every rank rates its own ladder of
      125M 3v3 matches over 12.5M players (link + streaming rate, best of 3) -- no data-path
      collective; aggregate = 1e9 matches / slowest rank's time;
  (b) ONE ladder of 1B matches over 100M players, table hash-sharded over the 8 GPUs, rated in two
      windows of 500M matches (a window holds fewer than 2^32 slot entries); checked bit for bit
      against the same ladder rated by one GPU alone in windows (every rank does that itself for its
      own shard of the table).

Prints one JSON object; writes nothing else.  ~2 minutes."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
from analyzer_b200 import engine, sharded, _lib as L

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
peak, _ = bench.hbm_peak()
out = {"n_gpus": world}
ev = lambda: torch.cuda.Event(enable_timing=True)


def barrier():
    torch.cuda.synchronize()
    dist.barrier()


# ---- (a) one ladder per GPU --------------------------------------------------------------------
raters = {}
def rater_of(p_draw):
    if p_draw not in raters:
        raters[p_draw] = engine.Rater(engine.Env(draw_probability=p_draw), local)
    return raters[p_draw]
SKIP_A = os.environ.get("NORTH_STAR_SKIP_A", "0") == "1"     # (b) only
bench.SEED = bench.SEED + rank            # a different ladder on every GPU
blk = None if SKIP_A else bench.config_block(rater_of, engine, "125M 3v3 matches over 12.5M players", 125_000_000, 12_500_000, 6, False, 0.0,
                         "ladder", 0, peak, reps=3, n_prefix=100_000, want_rounds=False)
if not SKIP_A:
    t = torch.tensor([blk["link_ms"] + blk["rate_ms"], blk["link_ms"], blk["rate_ms"], blk["check_max_abs_diff"]],
                     dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, link_ms, rate_ms, diff = (float(x) for x in t.cpu())
    out["config4"] = {"workload": "%d independent ladders, one per GPU: %gM matches / %gM players in all" % (
                          world, 125 * world, 12.5 * world), "ms_slowest_rank": ms, "link_ms": link_ms, "rate_ms": rate_ms,
                      "matches_per_s": 125e6 * world / (ms * 1e-3), "check_max_abs_diff": diff,
                      "check": "first 100k matches of every rank's ladder vs the sequential __float128 oracle"}
bench.SEED = bench.SEED - rank
r = rater_of(0.0)
r._ws = r._vt = None
torch.cuda.empty_cache()
barrier()

# ---- (b) one ladder over all GPUs, two windows ----------------------------------------------------
N, U, P = 1_000_000_000, 100_000_000, 6
WIN = 500_000_000
seed = bench.SEED + 0x300
gen = engine.make_gen(seed, U, slots=P, mode=1)
table0 = engine.gen_table(seed, 0, U, dev).reshape(U, 2)

# one GPU alone, in windows of 100M matches (the strong-scaling baseline and the bit-for-bit reference)
single = table0.clone().reshape(U, 1, 2)
pt = engine.PlayerTable.ladder(single, mode=1)
last = torch.full((U,), -1, dtype=torch.int32, device=dev)
w1 = 100_000_000
torch.cuda.synchronize()
t0 = time.perf_counter()
for lo in range(0, N, w1):
    ids, meta = engine.gen_matches(gen, lo, w1, dev)
    flags = (L.STREAM_LOAD if lo == 0 else 0) | (L.STREAM_STORE if lo + w1 >= N else 0)
    for a in range(0, w1, 10_000_000):         # link in pieces whose scatter stays in L2
        pass
    pred = torch.empty_like(ids)
    step = 10_000_000
    for a in range(0, w1, step):
        r.link_pred(ids[a:a + step], meta[a:a + step], U, pt.mode_cols, seq_base=lo + a, last=last, out=pred[a:a + step])
    r.rate_stream(pt, ids, meta, pred, None, 0, seq_base=lo, flags=flags)
    del ids, meta, pred
torch.cuda.synchronize()
one_gpu_s = time.perf_counter() - t0           # includes generating the stream (0.1 s per 100M)
assert r.error_count() == 0
want = single.reshape(U, 2)[rank::world].clone()
del single, pt, last
r._ws = r._vt = None
torch.cuda.empty_cache()
barrier()

ladder = sharded.ShardedLadder(r, U, WIN, P, rank, world, n_total=N)
ladder.set_table(table0)
del table0
torch.cuda.empty_cache()
link_ms, rate_ms, wall = 0.0, 0.0, 0.0
for lo in range(0, N, WIN):
    ids, meta = engine.gen_matches(gen, lo, WIN, dev)
    ladder.set_stream(ids, meta)
    del ids, meta
    torch.cuda.empty_cache()
    barrier()
    t0 = time.perf_counter()
    e0, e1, e2, e3 = ev(), ev(), ev(), ev()
    e0.record()
    if lo == 0:
        ladder.reset()
    ladder.barrier(host=False)
    ladder.link(WIN, 1, seq_base=lo)
    e1.record()
    ladder.barrier(host=False)
    e2.record()
    ladder.rate(WIN, 1, lo)
    e3.record()
    ladder.barrier()
    wall += time.perf_counter() - t0
    ladder.check()
    link_ms += e0.elapsed_time(e1)
    rate_ms += e2.elapsed_time(e3)
equal = torch.equal(ladder.table[rank], want)
maxdiff = float((ladder.table[rank] - want).abs().max())
t = torch.tensor([wall * 1e3, link_ms, rate_ms, 0.0 if equal else 1.0, maxdiff, one_gpu_s], dtype=torch.float64, device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
wall_ms, link_ms, rate_ms, unequal, maxdiff, one_gpu_s = (float(x) for x in t.cpu())
assert r.error_count() == 0
out["config5"] = {"workload": "ONE ladder, 1B 3v3 matches over 100M players, table hash-sharded over %d GPUs, 2 windows of 500M" % world,
                  "ms": wall_ms, "link_ms": link_ms, "rate_ms": rate_ms, "matches_per_s": N / (wall_ms * 1e-3),
                  "one_gpu_s_including_stream_generation": one_gpu_s, "one_gpu_matches_per_s": N / one_gpu_s,
                  "speedup_vs_one_gpu": one_gpu_s / (wall_ms * 1e-3),
                  "bit_equal_to_one_gpu": unequal == 0.0, "max_abs_diff_vs_one_gpu": maxdiff,
                  "hbm_per_gpu_gb": torch.cuda.max_memory_allocated(dev) / 1e9 + ladder.layouts[rank][1] / 1e9,
                  "timing": "wall clock of reset | link | rate | with host barriers, summed over the windows, max over ranks "
                            "(stream generation and slicing excluded)"}
ladder.close()
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
