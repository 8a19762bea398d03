"""The sharded ladder (ts_shard_link / ts_shard_rate, rows pushed to the rank that reads
them).  On one GPU every rank of a 1/2/4/8-way ladder is played by one launch
(emulate=True: same kernels, same index arithmetic, every pointer local); with 2 GPUs
the real thing runs, one process per GPU over NVLink peer memory.  The tables must equal
the single-GPU stream path BIT FOR BIT (both use the same explicitly rounded pass) and
stay within 1e-9 of the sequential float128 oracle."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

CASES = [(6, False, 0.0), (10, True, 0.02)]


def _single(engine, r, ids, meta, full, u, outputs=None):
    single = full.clone().reshape(u, 1, 2)
    pt = engine.PlayerTable.ladder(single, mode=1)
    pred = r.link_pred(ids, meta, u, pt.mode_cols)
    r.rate_stream(pt, ids, meta, pred, outputs)
    assert r.error_count() == 0
    return single.reshape(u, 2), pred


@pytest.mark.parametrize("world,direct,exchange", [(1, "1", True), (2, "1", True), (2, "0", False), (4, "0", True), (4, "1", False),
                                                   (8, "1", False), (8, "0", True), (16, "0", True)])
@pytest.mark.parametrize("slots,zipf,p_draw", CASES)
def test_emulated_ranks_equal_single_gpu_and_oracle(c_oracle, monkeypatch, world, direct, exchange, slots, zipf, p_draw):
    """direct = "1": rows stored straight into the receiver's mailbox; "0": as messages through the
    rings and the receiver's router blocks (what larger ladders on more GPUs use).
    exchange: the links' (player, slot) pairs are sent to the owners by the ranks that hold the
    matches (ts_shard_count / ts_shard_send); False: every rank reads the whole stream."""
    from analyzer_b200 import engine, sharded
    monkeypatch.setenv("TSRATE_SHARD_DIRECT", direct)
    co = c_oracle
    dev = torch.device("cuda", 0)
    n, u, seed = 60_011, 3_001, 0xA11CE005
    r = engine.Rater(engine.Env(draw_probability=p_draw), 0)
    gen = engine.make_gen(seed, u, slots=slots, mode=1, p_draw=p_draw, p_afk=0.02, zipf=zipf)
    ids, meta = engine.gen_matches(gen, 0, n, dev)
    full = engine.gen_table(seed, 0, u, dev).reshape(u, 2)
    outs1 = r.alloc_outputs(n, slots)
    want, pred = _single(engine, r, ids, meta, full, u, outs1)
    ladder = sharded.ShardedLadder(r, u, n, slots, 0, world, emulate=True, exchange=exchange)
    ladder.set_stream(ids, meta)
    ladder.set_table(full)
    outs = r.alloc_outputs(n, slots)
    ladder.run(n, mode=1, outputs=outs)
    assert r.error_count() == 0
    got = sharded.merge_tables([ladder.table[g] for g in range(world)])
    assert torch.equal(got, want), float((got - want).abs().max())
    for k in ("quality", "out_queue", "status"):
        a, b = outs[k], outs1[k]
        assert torch.equal(torch.nan_to_num(a.double(), nan=-7.0), torch.nan_to_num(b.double(), nan=-7.0)), k
    # the links: succ is the inverse of the single-GPU pred on every slot that takes part
    pred_h = pred.cpu().numpy().view(np.uint32)
    succ_full = np.full((n, slots), 0xFFFFFFFF, dtype=np.uint32)
    for g in range(world):
        nl = sharded.local_matches(n, g, world)
        gm = np.array([sharded.global_index(lm, g, world) for lm in range(nl)], dtype=np.int64)
        succ_full[gm] = ladder.succ[g][:nl].cpu().numpy().view(np.uint32)
    ids_h = ids.cpu().numpy().view(np.uint32)
    meta_h = meta.cpu().numpy().view(np.uint32)
    live = ((meta_h & 12) == 0)
    last = {}
    want_succ = np.full((n, slots), 0xFFFFFFFF, dtype=np.uint32)
    for m in range(n):
        if not live[m]:
            continue
        for s in range(slots):
            p = int(ids_h[m, s])
            if p in last:
                pm, ps = last[p]
                want_succ[pm, ps] = m
                assert pred_h[m, s] == pm
            last[p] = (m, s)
    used = live[:, None] & np.ones((1, slots), dtype=bool)
    assert np.array_equal(succ_full[used], want_succ[used])
    # float128 sequential oracle
    tab = full.cpu().numpy().reshape(u, 1, 2).copy()
    co.rate_sequential(co.make_params(draw_probability=p_draw), ids_h, meta_h, tab, -1, [-1, 0], quad=True)
    assert float(np.abs(got.cpu().numpy() - tab.reshape(u, 2)).max()) < 1e-9
    ladder.close()
    r.close()


@pytest.mark.parametrize("direct", ["0", "1"])
def test_emulated_windows_equal_one_shot(monkeypatch, direct):
    """A ladder rated window after window (the owners' table shards carry the state) equals one shot."""
    from analyzer_b200 import engine, sharded
    monkeypatch.setenv("TSRATE_SHARD_DIRECT", direct)
    dev = torch.device("cuda", 0)
    world, slots = 4, 6
    n, u, seed = 40_000, 2_003, 0xA11CE006
    r = engine.Rater(engine.Env(), 0)
    ids, meta = engine.gen_matches(engine.make_gen(seed, u, slots=slots, mode=1, p_afk=0.01), 0, n, dev)
    full = engine.gen_table(seed, 0, u, dev).reshape(u, 2)
    want, _ = _single(engine, r, ids, meta, full, u)
    win = 32 * world * 100                      # windows start at multiples of 32 * world
    ladder = sharded.ShardedLadder(r, u, win, slots, 0, world, emulate=True, n_total=n)    # rings sized for the whole run
    ladder.set_table(full)
    for lo in range(0, n, win):
        hi = min(n, lo + win)
        ladder.set_stream(ids[lo:hi], meta[lo:hi])
        ladder.run(hi - lo, mode=1, seq_base=lo, first=(lo == 0))
    got = sharded.merge_tables([ladder.table[g] for g in range(world)])
    assert torch.equal(got, want)
    assert r.error_count() == 0
    ladder.close()


def test_duplicate_player_and_bad_id_are_reported():
    from analyzer_b200 import _lib, engine, sharded
    dev = torch.device("cuda", 0)
    world, slots, n, u = 2, 6, 500, 400
    r = engine.Rater(engine.Env(), 0)
    ids, meta = engine.gen_matches(engine.make_gen(5, u, slots=slots, mode=1), 0, n, dev)
    full = engine.gen_table(5, 0, u, dev).reshape(u, 2)
    ladder = sharded.ShardedLadder(r, u, n, slots, 0, world, emulate=True)
    bad = ids.clone()
    bad[100, 3] = bad[100, 0]                   # the same player twice in one match
    ladder.set_stream(bad, meta)
    ladder.set_table(full)
    with pytest.raises(_lib.TsError, match="twice"):
        ladder.run(n, mode=1)
    # an id beyond the table is nobody's: the match is refused (ValueError status), everyone else goes on
    bad = ids.clone()
    bad[100, 3] = u + 5
    ladder.set_stream(bad, meta)
    ladder.set_table(full)
    r.reset_errors()
    outs = r.alloc_outputs(n, slots)
    ladder.run(n, mode=1, outputs=outs)
    assert int(outs["status"][100]) == _lib.ST_VALUE_ERROR
    fp, val = r.error_counts()
    assert (fp, val) == (0, 1)
    single = full.clone().reshape(u, 1, 2)
    pt = engine.PlayerTable.ladder(single, mode=1)
    r.reset_errors()
    r.rate_stream(pt, bad, meta, r.link_pred(bad, meta, u, pt.mode_cols))
    assert r.error_counts() == (0, 1)
    got = sharded.merge_tables([ladder.table[g] for g in range(world)])
    assert torch.equal(got, single.reshape(u, 2))
    ladder.close()


@pytest.mark.parametrize("exchange", [True, False])
def test_windows_smaller_than_a_round_of_blocks(exchange):
    """Windows of fewer than 32 * G matches: some ranks hold no match at all (they still own players,
    link them and take their rows back), blocks are ragged."""
    from analyzer_b200 import engine, sharded
    dev = torch.device("cuda", 0)
    slots, u = 6, 101
    r = engine.Rater(engine.Env(), 0)
    for world in (2, 4, 8):
        for n in (1, 31, 33, 40, 257):
            ids, meta = engine.gen_matches(engine.make_gen(100 + n, u, slots=slots, mode=1, p_afk=0.05), 0, n, dev)
            full = engine.gen_table(100 + n, 0, u, dev).reshape(u, 2)
            want, _ = _single(engine, r, ids, meta, full, u)
            ladder = sharded.ShardedLadder(r, u, n, slots, 0, world, emulate=True, exchange=exchange)
            ladder.set_stream(ids, meta)
            ladder.set_table(full)
            ladder.run(n, mode=1)
            got = sharded.merge_tables([ladder.table[g] for g in range(world)])
            assert torch.equal(got, want), (world, n)
            assert r.error_count() == 0
            ladder.close()


@pytest.mark.parametrize("exchange", [True, False])
def test_skewed_ownership_overflows_loudly_and_fits_with_more_room(exchange):
    """Every player owned by rank 0 (all ids even, 2 ranks): rank 0 is sent twice its share of the slots.
    The default capacity (1.25 shares) must report TS_SHARD_OVERFLOW, not rate with missing links; a
    ladder laid out for a window twice as large has the room and equals one GPU bit for bit."""
    from analyzer_b200 import _lib, engine, sharded
    dev = torch.device("cuda", 0)
    world, slots, n, u = 2, 6, 40_000, 3_000
    r = engine.Rater(engine.Env(), 0)
    ids, meta = engine.gen_matches(engine.make_gen(11, u // 2, slots=slots, mode=1), 0, n, dev)
    ids = ids * 2
    full = engine.gen_table(11, 0, u, dev).reshape(u, 2)
    want, _ = _single(engine, r, ids, meta, full, u)
    tight = sharded.ShardedLadder(r, u, n, slots, 0, world, emulate=True, exchange=exchange)
    tight.set_stream(ids, meta)
    tight.set_table(full)
    with pytest.raises(_lib.TsError, match="too small"):
        tight.run(n, mode=1)
    tight.close()
    r.reset_errors()
    roomy = sharded.ShardedLadder(r, u, 2 * n, slots, 0, world, emulate=True, exchange=exchange)
    roomy.set_stream(ids, meta)
    roomy.set_table(full)
    if exchange:
        roomy.run(n, mode=1)
    else:           # the whole-stream form sizes its workspace per call
        roomy.reset()
        roomy.barrier()
        roomy.link(n, 1, owned_cap=n * slots)
        roomy.check()
        roomy.rate(n, 1)
        roomy.barrier()
        roomy.check()
    got = sharded.merge_tables([roomy.table[g] for g in range(world)])
    assert torch.equal(got, want)
    assert r.error_count() == 0
    roomy.close()


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    from analyzer_b200 import engine, sharded
    n, u, seed = 200_000, 3_001, 0xA11CE005
    ok = True
    for slots, zipf, p_draw in CASES:
        r = engine.Rater(engine.Env(draw_probability=p_draw), rank)
        gen = engine.make_gen(seed, u, slots=slots, mode=1, p_draw=p_draw, p_afk=0.02, zipf=zipf)
        ids, meta = engine.gen_matches(gen, 0, n, dev)
        full = engine.gen_table(seed, 0, u, dev).reshape(u, 2)
        want, _ = _single(engine, r, ids, meta, full, u)
        for exchange in (True, False):      # links by pairs sent to the owners / by every rank reading the whole stream
            ladder = sharded.ShardedLadder(r, u, n, slots, rank, world, exchange=exchange)
            ladder.set_stream(ids, meta)
            for direct in ("1", "0"):       # rows straight into the peer's mailbox / as ring messages
                os.environ["TSRATE_SHARD_DIRECT"] = direct
                ladder.set_table(full)
                ladder.run(n, mode=1)
                errs = r.error_count()
                mine = ladder.table[rank]
                same = torch.equal(mine, want[rank::world]) and errs == 0        # bit for bit
                if not same:
                    diff = (mine - want[rank::world]).abs()
                    print("sharded mismatch: rank", rank, "slots", slots, "direct", direct, "exchange", exchange, "errs", errs,
                          "rows differing", int((diff > 0).any(dim=1).sum()), "max", float(diff.max()), flush=True)
                ok = ok and same
                dist.barrier()
            os.environ.pop("TSRATE_SHARD_DIRECT", None)
            ladder.close()
        r.close()
    results[rank] = ok
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_sharded_ladder_two_gpus_equals_single_gpu():
    world = 2
    with mp.Manager() as mgr:
        results = mgr.dict()
        mp.spawn(_worker, args=(world, 29541, results), nprocs=world, join=True)
        assert dict(results) == {0: True, 1: True}
