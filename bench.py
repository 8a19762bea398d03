#!/usr/bin/env python
"""bench.py -- matches rated/sec (fp64 TrueSkill) on B200, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A step is one pass of the hot path over one batch: schedule (radix sort, chain
link, rounds, round bucketing) + rate (all rounds) of the
workload -- BASELINE.json configs[1]: 10M synthetic 3v3 ranked matches over 1M
players, single (mu, sigma) ladder -- on inputs already resident in HBM.  With
N GPUs every rank rates its own independent ladder of that size (BASELINE.json
configs[3]: ladders shard by game mode x region; no data-path collective), so
scaling is weak and `value` is the sum over ranks / max-over-ranks time.

`e2e` is the same workload through the reference-facing call
(analyzer_b200.engine.Rater.rate_batch) with pinned HOST buffers: H2D of ids,
meta and the player table, schedule, rate, D2H of the table, every step.

`--impl reference` times the reference's CPU implementation of the path on the
host cores: oracle/rater_port.py (the restated rater.py policy) on the
trueskill-API shim at 50 digits -- the pinned trueskill==0.4.4 is not
installable (no network) and /root/reference cannot travel to the GPU box.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "matches rated/sec (fp64 TrueSkill)"
UNIT = "matches/s"
SEED = 0xA11CE002
FALLBACK_HBM_GBS = 6650.0     # /opt/skills/guides/B200_PROFILING.md, used only without MEASURED_PEAKS.json


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--matches", type=int, default=10_000_000)
    ap.add_argument("--players", type=int, default=1_000_000)
    ap.add_argument("--slots", type=int, default=6)
    ap.add_argument("--lanes", type=int, default=0, help="lanes per match of the rating kernel (0 = default)")
    ap.add_argument("--path", default="stream", choices=["rounds", "stream"],
                    help="headline path. stream: ts_link_pred + ts_rate_stream (sort, link, dataflow rating in stream "
                         "order); rounds: ts_schedule + ts_rate (sort, link, rounds by dataflow, round buckets, one "
                         "grid barrier per round).  Both are timed every run; the other one is reported under 'paths'.")
    ap.add_argument("--sharded", action="store_true",
                    help="ONE ladder of --matches over all ranks: player table hash-sharded in NVLink peer memory "
                         "(BASELINE.json configs[4]); strong scaling.  Default without it: one ladder per rank.")
    ap.add_argument("--zipf", action="store_true")
    ap.add_argument("--p-draw", type=float, default=0.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--host-numa", dest="numa", default="local", choices=["none", "local", "interleave"],
                    help="memory policy for the pinned host buffers of the e2e leg when several ranks run (see numa_policy)")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs block (configs 3 and 4 at full size) at N = 1")
    ap.add_argument("--no-sharded", action="store_true", help="skip the one-ladder-over-all-ranks block at N > 1")
    ap.add_argument("--sharded-matches", type=int, default=100_000_000)
    ap.add_argument("--sharded-players", type=int, default=10_000_000)
    ap.add_argument("--sharded-deadline", type=int, default=300)
    ap.add_argument("--emulate", type=int, default=0, help="with --sharded on one GPU: play this many ranks in one process")
    ap.add_argument("--shard-exchange", type=int, default=1,
                    help="sharded ladder links: 1 = pairs sent to their owners (ts_shard_count / ts_shard_send), 0 = every rank reads the whole stream")
    return ap.parse_args()


def workload_name(a):
    return "%s synthetic %s matches over %s players, single ladder (K=1 pair per player)%s%s" % (
        _si(a.matches), "3v3 ranked" if a.slots == 6 else "5v5 ranked", _si(a.players),
        ", Zipf activity" if a.zipf else "", ", %.0f%% draws" % (100 * a.p_draw) if a.p_draw else "")


def _si(n):
    for div, suf in ((10 ** 9, "B"), (10 ** 6, "M"), (10 ** 3, "k")):
        if n >= div and n % (div // 10 if div >= 10 else 1) == 0:
            return ("%g" % (n / div)) + suf
    return str(n)


def rate_bytes_per_match(slots, k=1):
    """record (4P ids + 4 meta) + gather and scatter of P pairs of 16 B, K pairs each."""
    return (4 * slots + 4) + 2 * slots * 16 * k


def sched_bytes_per_match(slots, n_players):
    bits = max(1, (n_players).bit_length())
    return slots * 16 * ((bits + 7) // 8) + 4 * slots + 8


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------
# CPU side (oracle/ is executed only here: cpu_baseline leg and --impl reference)
# ---------------------------------------------------------------------------
def cpu_reference_rate(ids, meta, table, seconds, draw_probability=0.0):
    """Single-rate port (one env.rate per match, like the K=1 GPU workload) on a
    prefix of the stream sized to run for about `seconds`; 1 core, because the
    reference is one process with one thread (worker.py:219-221)."""
    import numpy as np
    from oracle import rater_port as rp
    env = rp.make_env(draw_probability=draw_probability)
    probe = min(64, len(ids))
    ratings = {int(i): (float(table[int(i), 0, 0]), float(table[int(i), 0, 1])) for i in np.unique(ids[:probe])}
    t0 = time.perf_counter()
    rp.rate_ladder(ids[:probe], meta[:probe], ratings, env)
    per = (time.perf_counter() - t0) / probe
    n = int(max(probe, min(len(ids), seconds / per)))
    ratings = {int(i): (float(table[int(i), 0, 0]), float(table[int(i), 0, 1])) for i in np.unique(ids[:n])}
    t0 = time.perf_counter()
    rp.rate_ladder(ids[:n], meta[:n], ratings, env)
    dt = time.perf_counter() - t0
    return n / dt, n, dt


def cpu_rate_match_rate(ids, meta, seconds):
    """The full rate_match (shared + queue rate() + quality) of the port, for context."""
    import numpy as np
    from oracle import rater_port as rp
    n = min(len(ids), 256)
    players = {int(i): rp.new_player(skill_tier=int(i) % 30) for i in np.unique(ids[:n])}
    matches = rp.matches_from_columns(ids[:n], meta[:n], players)
    t0 = time.perf_counter()
    done = 0
    for m in matches:
        rp.rate_match(m)
        done += 1
        if time.perf_counter() - t0 > seconds:
            break
    return done / (time.perf_counter() - t0), done


def faithful_rate(rater, engine, ids_d, meta_d, table0, u, lanes):
    """The reference's own semantics on the same stream (K = 2: shared pair + queue pair per
    player, quality, two rate() per match -- rater.py:108-169), HBM-resident: every player
    starts with a stored shared rating and a NULL queue rating."""
    import torch
    dev = ids_d.device
    n, p = ids_d.shape
    mode = int((int(meta_d[0]) >> 4) & 15)
    cols = [-1] * 6
    cols[mode] = 1
    tab0 = torch.full((u, 2, 2), float("nan"), dtype=torch.float64, device=dev)
    tab0[:, 0] = table0.reshape(u, 2)
    tab = tab0.clone()
    pt = engine.PlayerTable(tab, 0, cols)
    out = {}
    for path in ("stream", "rounds"):
        best = None
        for it in range(3):
            tab.copy_(tab0)
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            if path == "stream":
                pred = rater.link_pred_windows(ids_d, meta_d, u, pt.mode_cols)
                e1.record()
                rater.rate_stream(pt, ids_d, meta_d, pred, None, lanes)
            else:
                sch = rater.schedule(ids_d, meta_d, u, pt.mode_cols)
                e1.record()
                rater.rate(pt, sch, None, lanes)
            e2.record()
            torch.cuda.synchronize()
            t = (e0.elapsed_time(e1), e1.elapsed_time(e2))
            if best is None or sum(t) < sum(best):
                best = t
        out[path] = {"schedule_ms": best[0], "rate_ms": best[1], "matches_per_s": n / (sum(best) * 1e-3)}
        out[path + "_table"] = tab.clone()
    diff = float((out.pop("stream_table") - out.pop("rounds_table")).abs().max())
    assert diff < 1e-9, diff
    assert rater.error_count() == 0
    out["algorithmic_bytes_per_match"] = rate_bytes_per_match(p, 2)
    out["note"] = "shared + queue pair per player (K = 2), best of 3, no per-participant outputs; both paths agree"
    if p == 6:
        # what worker.py:176-192 hands over: ONE chronological stream whose game mode varies from
        # match to match (here 40% ranked, 30% casual, 20% blitz, 10% br), against the 8-pair table
        del tab, tab0, pt
        g = torch.Generator(device=dev)
        g.manual_seed(7)
        x = torch.rand(n, device=dev, generator=g)
        modes = torch.where(x < 0.4, 1, torch.where(x < 0.7, 0, torch.where(x < 0.9, 2, 3))).to(torch.int32)
        meta_mix = (meta_d & ~0xF0) | (modes << 4)
        tab0 = torch.full((u, 8, 2), float("nan"), dtype=torch.float64, device=dev)
        tab0[:, 0] = table0.reshape(u, 2)
        tab = tab0.clone()
        pt = engine.PlayerTable(tab, 0, [1, 2, 3, 4, 5, 6])
        best = None
        for it in range(3):
            tab.copy_(tab0)
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            pred = rater.link_pred_windows(ids_d, meta_mix, u, pt.mode_cols)
            e1.record()
            rater.rate_stream(pt, ids_d, meta_mix, pred, None, lanes)
            e2.record()
            torch.cuda.synchronize()
            t = (e0.elapsed_time(e1), e1.elapsed_time(e2))
            if best is None or sum(t) < sum(best):
                best = t
        assert rater.error_count() == 0
        out["mixed_modes"] = {"schedule_ms": best[0], "rate_ms": best[1], "matches_per_s": n / (sum(best) * 1e-3),
                              "note": "stream path, game mode varying from match to match (4 queues), 8-pair table"}
    return out


class _Row(object):
    """ORM-like row: attributes only (the shape worker_test.py:6-63 fakes)."""

    def __init__(self, **kw):
        self.__dict__.update(kw)


_QUEUE_COLS = ["trueskill_" + m for m in ("casual", "ranked", "blitz", "br", "5v5_casual", "5v5_ranked")]


def _object_graphs(ids, meta, n_players):
    """Match / roster / participant / player graphs of a 'ranked' stream; players are shared
    between matches like in a SQLAlchemy session's identity map (worker.py:176-192)."""
    def player(i):
        p = _Row(api_id=str(i), skill_tier=i % 30, rank_points_ranked=None, rank_points_blitz=None,
                 trueskill_mu=None, trueskill_sigma=None)
        for c in _QUEUE_COLS:
            setattr(p, c + "_mu", None)
            setattr(p, c + "_sigma", None)
        return p

    players = [player(i) for i in range(n_players)]
    half = ids.shape[1] // 2
    out = []
    for m in range(len(ids)):
        rosters = []
        for t in range(2):
            parts = []
            for s in range(t * half, (t + 1) * half):
                items = _Row(api_id="", any_afk=False)
                for c in _QUEUE_COLS:
                    setattr(items, c + "_mu", None)
                    setattr(items, c + "_sigma", None)
                parts.append(_Row(api_id="", went_afk=0, participant_items=[items], player=[players[int(ids[m, s])]],
                                  trueskill_mu=None, trueskill_sigma=None, trueskill_delta=None))
            rosters.append(_Row(api_id="", winner=bool(int(meta[m]) & (1 << t)), participants=parts))
        out.append(_Row(api_id=str(m), game_mode="ranked", rosters=rosters,
                        participants=[p for r in rosters for p in r.participants], trueskill_quality=None))
    return out


def object_api_rate(ids_d, meta_d, n=5000):
    """The drop-in object API: analyzer_b200.rater.rate_matches on ORM-like object
    graphs (the shape worker.py:176-192 hands to rate_match), against the port's
    rate_match loop on the same kind of graphs.  Ingest and write-back are Python."""
    import logging
    import numpy as np
    from analyzer_b200 import rater as gpu_rater
    gpu_rater.logger.setLevel(logging.ERROR)
    ids = ids_d[:n].cpu().numpy().view(np.uint32) % 3000        # 3000 players: chains inside the batch
    meta = meta_d[:n].cpu().numpy().view(np.uint32)
    for row in ids:                                              # keep players distinct inside a match
        seen = set()
        for k in range(len(row)):
            while int(row[k]) in seen:
                row[k] = (int(row[k]) + 1) % 3000
            seen.add(int(row[k]))
    gpu_rater.rate_matches(_object_graphs(ids[:64], meta[:64], 3000))      # warm-up (context, first launches)
    g = _object_graphs(ids, meta, 3000)
    t0 = time.perf_counter()
    gpu_rater.rate_matches(g)
    dt_gpu = time.perf_counter() - t0
    # reference loop on the same graphs (the oracle is used here only as the CPU baseline / checker)
    from oracle import rater_port as rp
    g2 = _object_graphs(ids[:400], meta[:400], 3000)
    t0 = time.perf_counter()
    for m in g2:
        rp.rate_match(m)
    dt_cpu = (time.perf_counter() - t0) / len(g2)
    err = max(abs(a.participants[0].trueskill_mu - b.participants[0].trueskill_mu)
              for a, b in zip(g[:400], g2) if b.participants[0].trueskill_mu is not None)
    # the same batch as a ROW STREAM (what the query of worker.py:176-191 yields, one row per participant):
    # numpy ingest, one GPU call, numpy update rows -- no per-object Python (rater.ingest_rows / updates)
    p = ids.shape[1]
    rows_n = n * p
    nan = np.full(rows_n, np.nan)
    rows = dict(match=np.repeat(np.arange(n), p), game_mode=np.full(rows_n, 1), n_rosters=np.full(rows_n, 2),
                roster=np.tile(np.repeat([0, 1], p // 2), n), went_afk=np.zeros(rows_n, dtype=np.int64),
                winner=((meta[:, None] >> np.repeat([0, 1], p // 2)[None, :]) & 1).astype(bool).reshape(-1),
                player=ids.reshape(-1).astype(np.int64), skill_tier=(ids.reshape(-1) % 30).astype(np.int64),
                rank_points_ranked=nan, rank_points_blitz=nan, trueskill_mu=nan, trueskill_sigma=nan,
                queue=np.full((rows_n, 6, 2), np.nan))
    gpu_rater.rate_rows({k: v[:64 * p] for k, v in rows.items()})          # warm-up
    t0 = time.perf_counter()
    b, up = gpu_rater.rate_rows(rows)
    dt_rows = time.perf_counter() - t0
    first = next(i for i in range(n) if g[i].participants[0].trueskill_mu is not None)
    err_rows = abs(g[first].participants[0].trueskill_mu - up["part_mu"][first * p])
    small = {k: v[:500 * p] for k, v in rows.items()}
    t0 = time.perf_counter()
    gpu_rater.rate_rows(small)
    dt_small = time.perf_counter() - t0
    return {"rate_matches_per_s": n / dt_gpu, "batch": n, "reference_loop_per_s": 1.0 / dt_cpu,
            "note": "faithful shared + queue path (rounds), %d-match batch of object graphs, 3000 players; "
                    "Python ingest + write-back included" % n, "check_max_abs_diff": err,
            "rate_rows_per_s": n / dt_rows, "rate_rows_500_match_batch_ms": dt_small * 1e3,
            "rate_rows_note": "the same batch as participant rows: ingest_rows + one GPU call + updates (numpy, no per-object "
                              "Python); 500-match batch = the reference worker's BATCHSIZE", "rate_rows_check_abs_diff": err_rows}


def cpu_c_baselines(ids, meta, table, n_players):
    """the closed form in C, fp64, 1 core sequential and
    round-parallel OpenMP on all host cores (reported, not the target)."""
    import numpy as np
    from oracle import c_oracle as co
    prm = co.make_params()
    cols = [-1, 0]
    t = table.copy()
    t0 = time.perf_counter()
    co.rate_sequential(prm, ids, meta, t, -1, cols, quad=False)
    seq = len(ids) / (time.perf_counter() - t0)
    t0 = time.perf_counter()
    rnd, order, offs, R = co.schedule(ids, meta, cols, n_players)
    sched_s = time.perf_counter() - t0
    t = table.copy()
    t0 = time.perf_counter()
    co.rate_rounds_omp(prm, ids, meta, t, -1, cols, order, offs, R)
    omp = len(ids) / (time.perf_counter() - t0)
    return dict(c_fp64_1core=seq, c_fp64_omp=omp, omp_threads=co.max_threads(),
                c_schedule_matches_per_s=len(ids) / sched_s, sample_matches=len(ids), rounds=R)


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import numpy as np
    from oracle import c_oracle as co
    co.build()
    gen = co.make_gen(SEED, a.players, slots=a.slots, mode=1, p_draw=a.p_draw, zipf=a.zipf)
    budget = 90.0 / max(1, a.steps + a.warmup)          # whole run ~1.5 min of CPU work
    sample = max(64, int(budget * 300))
    ids, meta = co.gen_matches(gen, 0, sample)
    table = co.gen_table(SEED, 0, a.players)
    times = []
    n_done = 0
    for step in range(a.warmup + a.steps):
        rate, n_done, dt = cpu_reference_rate(ids, meta, table, budget, a.p_draw)
        if step >= a.warmup:
            times.append(dt / n_done)
    per_match = sum(times) / len(times)
    value = 1.0 / per_match
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * per_match * n_done,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "mpmath-50-digit",
        "data": "synthetic", "config": {"workload": workload_name(a), "seed": hex(SEED)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": "first %d matches of the workload's stream per step; oracle/rater_port.py "
                                   "rate_ladder (one env.rate per match) on the trueskill-API shim, mpmath 50 "
                                   "digits; 1 core because the reference is single-process/single-thread "
                                   "(worker.py:219-221)" % n_done},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------
# clocks during the timed region (NVML, sampled from a thread)
# ---------------------------------------------------------------------------
class ClockSampler(object):
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                self.samples.append((mhz, util))
                for b, name in self.REASONS.items():
                    if bits & b and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.nv:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join()
        mhz = sorted(m for m, _ in self.samples)
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(mhz)}


# ---------------------------------------------------------------------------
# BASELINE.json configs[2..3] on one GPU, each with a float128-oracle prefix check
# ---------------------------------------------------------------------------
def oracle_prefix_check(rater, engine, gen_args, env_draw, u, layout, n_prefix, lanes, sharded_ladder=None):
    """Rate the first n_prefix matches of a ladder through the SAME GPU path and compare the
    table with the sequential __float128 oracle on the host (a prefix is an exact check:
    the first k matches of a stream do not depend on the rest).  Returns max |diff|."""
    import numpy as np
    import torch
    from oracle import c_oracle as co
    dev = rater.device
    seed = gen_args["seed"]
    gen = engine.make_gen(seed, u, slots=gen_args["slots"], mode=1, p_draw=gen_args["p_draw"], zipf=gen_args["zipf"])
    ids_d, meta_d = engine.gen_matches(gen, 0, n_prefix, dev)
    ids = ids_d.cpu().numpy().view(np.uint32)
    meta = meta_d.cpu().numpy().view(np.uint32)
    touched = np.unique(ids)
    tab0 = engine.gen_table(seed, 0, u, dev)
    prm = co.make_params(draw_probability=env_draw)
    if layout == "ladder":
        want = tab0.cpu().numpy()
        co.rate_sequential(prm, ids, meta, want, -1, [-1, 0], quad=True)
        if sharded_ladder is not None:
            from analyzer_b200 import sharded
            lad = sharded_ladder
            lad.set_stream(ids_d, meta_d)
            lad.set_table(tab0.reshape(u, 2))
            lad.run(n_prefix, mode=1, lanes=lanes)
            diff = 0.0
            want_t = torch.from_numpy(want.reshape(u, 2)).to(dev)
            for g in lad.played:
                diff = max(diff, float((lad.table[g] - want_t[g::lad.world]).abs().max()))
            return diff
        pt = engine.PlayerTable.ladder(tab0, mode=1)
        rater.rate_stream(pt, ids_d, meta_d, rater.link_pred(ids_d, meta_d, u, pt.mode_cols), None, lanes)
        got = tab0.cpu().numpy()
    else:       # "pair": shared + queue pair, every player starts with a stored shared rating and a NULL queue rating
        t0 = torch.full((u, 2, 2), float("nan"), dtype=torch.float64, device=dev)
        t0[:, 0] = tab0.reshape(u, 2)
        want = t0.cpu().numpy()
        co.rate_sequential(prm, ids, meta, want, 0, [-1, 1], quad=True)
        pt = engine.PlayerTable(t0, 0, [-1, 1])
        rater.rate_stream(pt, ids_d, meta_d, rater.link_pred(ids_d, meta_d, u, pt.mode_cols), None, lanes)
        got = t0.cpu().numpy()
    assert rater.error_count() == 0
    d = np.abs(got[touched] - want[touched])
    return float(np.nanmax(d))


def big_traffic(kernel, n, p, k):
    """dram bytes per launch of `kernel` at this size from the committed ncu summary, or None."""
    tp = os.path.join(ROOT, "profiles", "r2_big_traffic.json")
    try:
        for row in json.load(open(tp)):
            if row["kernel"] == kernel and row["matches"] == n and row["slots"] == p and row["pairs"] == k:
                return row["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


def config_block(rater_of, engine, name, n, u, p, zipf, p_draw, layout, lanes, peak, reps=3, n_prefix=200_000,
                 want_rounds=True):
    """One BASELINE.json configuration on this GPU, HBM-resident: stream path (link + rate),
    best of `reps`, roofline of the rating kernel on the algorithmic bytes, oracle prefix check."""
    import torch
    rater = rater_of(p_draw)
    dev = rater.device
    seed = SEED + 0x100
    gen = engine.make_gen(seed, u, slots=p, mode=1, p_draw=p_draw, zipf=zipf)
    ids_d, meta_d = engine.gen_matches(gen, 0, n, dev)
    base = engine.gen_table(seed, 0, u, dev)
    k = 1 if layout == "ladder" else 2
    if layout == "ladder":
        tab0 = base
        make_pt = lambda t: engine.PlayerTable.ladder(t, mode=1)
    else:
        tab0 = torch.full((u, 2, 2), float("nan"), dtype=torch.float64, device=dev)
        tab0[:, 0] = base.reshape(u, 2)
        make_pt = lambda t: engine.PlayerTable(t, 0, [-1, 1])
    tab = tab0.clone()
    pt = make_pt(tab)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    best = None
    pred = None
    for it in range(1 + reps):
        tab.copy_(tab0)
        pred = None
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        pred = rater.link_pred_windows(ids_d, meta_d, u, pt.mode_cols)
        e1.record()
        rater.rate_stream(pt, ids_d, meta_d, pred, None, lanes)
        e2.record()
        torch.cuda.synchronize()
        t = (e0.elapsed_time(e1), e1.elapsed_time(e2))
        if it > 0 and (best is None or sum(t) < sum(best)):
            best = t
    assert rater.error_count() == 0
    del pred
    b = rate_bytes_per_match(p, k)
    achieved = b * n / (best[1] * 1e-3) / 1e9
    kernel = "k_rate_stream" if layout == "ladder" else "k_rate_stream_pair"
    traffic = big_traffic(kernel, n, p, k)
    out = {"workload": name, "path": "stream", "link_ms": best[0], "rate_ms": best[1],
           "matches_per_s": n / (sum(best) * 1e-3),
           "roofline": {"kernel": kernel, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": traffic, "algorithmic_bytes_per_match": b,
                        "traffic_over_algorithmic": (traffic / (b * n)) if traffic else None},
           "l2": "table %d MB, versioned rows %d MB (L2 is 126 MB)" % ((u * 16 * k) >> 20, (u * 32 * k) >> 20)}
    if layout == "ladder":
        # the slot-indexed form (ts_link_succ + ts_rate_slots: rows live where they are read), link and rate
        # interleaved window by window; same bits as the stream path
        try:
            ref = tab.clone()
            t_best = None
            for it in range(3):
                tab.copy_(tab0)
                e0, e1 = ev(), ev()
                e0.record()
                rater.rate_slots(pt, ids_d, meta_d, None, lanes)
                e1.record()
                torch.cuda.synchronize()
                if it and (t_best is None or e0.elapsed_time(e1) < t_best):
                    t_best = e0.elapsed_time(e1)
            assert rater.error_count() == 0
            out["slots_path"] = {"total_ms": t_best, "matches_per_s": n / (t_best * 1e-3),
                                 "bit_equal_to_stream_path": bool(torch.equal(tab, ref)),
                                 "note": "rows indexed by the consuming slot (sequential reads, scattered writes); link + rate "
                                         "per window of ~60 Mi slot entries"}
            del ref
            rater._slot_rows = rater._slot_succ = None
        except Exception as exc:
            out["slots_path"] = {"error": repr(exc)}
    if want_rounds:
        try:
            tab.copy_(tab0)
            e0, e1, e2 = ev(), ev(), ev()
            e0.record()
            sch = rater.schedule(ids_d, meta_d, u, pt.mode_cols)
            e1.record()
            rater.rate(pt, sch, None, lanes)
            e2.record()
            torch.cuda.synchronize()
            out["rounds"] = sch.n_rounds
            out["rounds_path"] = {"schedule_ms": e0.elapsed_time(e1), "rate_ms": e1.elapsed_time(e2),
                                  "roofline_frac_rate": b * n / (e1.elapsed_time(e2) * 1e-3) / 1e9 / peak}
            del sch
        except Exception as exc:
            out["rounds_path"] = {"error": repr(exc)}
    del ids_d, meta_d, tab, tab0, base, pt
    rater._ws = None
    torch.cuda.empty_cache()
    out["check_max_abs_diff"] = oracle_prefix_check(
        rater, engine, {"seed": seed, "slots": p, "p_draw": p_draw, "zipf": zipf}, p_draw, u, layout, n_prefix, lanes)
    out["check"] = "first %d matches through the same path vs the sequential __float128 oracle" % n_prefix
    assert out["check_max_abs_diff"] < 1e-9, out
    rater._ws = rater._vt = None
    torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------
# BASELINE.json configs[4]: ONE ladder over all ranks (strong scaling)
# ---------------------------------------------------------------------------
def sharded_block(a, rater, engine, dist, rank, world, dev, n, u, p, steps=5, warmup=2, emulate=0, n_prefix=200_000):
    """Rows pushed to the rank that rates the player's next match (analyzer_b200/sharded.py).
    Every rank also rates the whole ladder alone (the one-GPU stream path) for the
    strong-scaling baseline and the bit-for-bit check of its table shard."""
    import torch
    from analyzer_b200 import sharded
    G = emulate if emulate else world
    seed = SEED + 0x200
    gen = engine.make_gen(seed, u, slots=p, mode=1, p_draw=a.p_draw, zipf=a.zipf)
    ids_d, meta_d = engine.gen_matches(gen, 0, n, dev)
    full0 = engine.gen_table(seed, 0, u, dev).reshape(u, 2)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    # one GPU alone, same ladder
    one = None
    single = full0.clone().reshape(u, 1, 2)
    pt = engine.PlayerTable.ladder(single, mode=1)
    for it in range(3):
        single.copy_(full0.reshape(u, 1, 2))
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        pred = rater.link_pred_windows(ids_d, meta_d, u, pt.mode_cols)
        e1.record()
        rater.rate_stream(pt, ids_d, meta_d, pred, None, a.lanes)
        e2.record()
        torch.cuda.synchronize()
        t = (e0.elapsed_time(e1), e1.elapsed_time(e2))
        if it > 0 and (one is None or sum(t) < sum(one)):
            one = t
        del pred
    assert rater.error_count() == 0
    want = single.reshape(u, 2)
    rater._ws = rater._vt = None
    torch.cuda.empty_cache()

    ladder = sharded.ShardedLadder(rater, u, n, p, 0 if emulate else rank, G, emulate=bool(emulate), exchange=bool(a.shard_exchange))
    prefix_diff = oracle_prefix_check(rater, engine, {"seed": seed, "slots": p, "p_draw": a.p_draw, "zipf": a.zipf},
                                      a.p_draw, u, "ladder", min(n_prefix, n), a.lanes, sharded_ladder=ladder)
    ladder.set_stream(ids_d, meta_d)
    del ids_d, meta_d
    torch.cuda.empty_cache()
    walls, links, rates = [], [], []
    for it in range(warmup + steps):
        ladder.set_table(full0)
        barrier()
        t0 = time.perf_counter()
        e0, e1, e2, e3 = ev(), ev(), ev(), ev()
        e0.record()
        ladder.reset()
        ladder.barrier(host=False)     # the GPUs wait for each other on their streams; the host queues on
        ladder.link(n, 1)
        e1.record()
        ladder.barrier(host=False)
        e2.record()
        ladder.rate(n, 1, 0, None, a.lanes)
        e3.record()
        ladder.barrier()               # host: everything done on every rank
        dt = time.perf_counter() - t0
        ladder.check()
        if it >= warmup:
            walls.append(dt * 1e3)
            links.append(e0.elapsed_time(e1))
            rates.append(e2.elapsed_time(e3))
    equal = all(torch.equal(ladder.table[g], want[g::G]) for g in ladder.played)
    maxdiff = max(float((ladder.table[g] - want[g::G]).abs().max()) for g in ladder.played)
    t = torch.tensor([sum(walls) / len(walls), sum(links) / len(links), sum(rates) / len(rates), one[0], one[1],
                      0.0 if equal else 1.0, maxdiff, prefix_diff], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_wall, ms_link, ms_rate, one_link, one_rate, unequal, maxdiff, prefix_diff = (float(x) for x in t.cpu())
    ladder_exchange = ladder.exchange
    ladder.close()
    assert rater.error_count() == 0
    remote = (G - 1) / G
    rate_nv = remote * p * 32 * n / G                       # rows pushed to other GPUs, per GPU
    if ladder_exchange:      # (player, slot) pairs and block counts stored to peers + succ stored to peers
        link_nv = remote * (n * p * 8) / G + remote * (n / 32) * 2 * G + remote * (n * p * 8) / G
    else:                    # id stream read from peers + succ stored to peers
        link_nv = remote * (n * p * 4 + n * 4) + remote * (n * p * 8) / G
    out = {"workload": "%s synthetic %s matches over %s players, ONE ladder, table hash-sharded over %d GPUs%s" % (
               _si(n), "3v3" if p == 6 else "5v5", _si(u), G, " (all ranks played by one GPU)" if emulate else ""),
           "n_gpus": G, "scaling": "strong", "ms_per_step": ms_wall, "link_ms": ms_link, "rate_ms": ms_rate,
           "matches_per_s": n / (ms_wall * 1e-3),
           "timing": "ms_per_step = wall clock from before the mailbox reset to after the last (host) barrier, the "
                     "barriers between the phases included (one-word NCCL all-reduces on the streams), max over ranks; "
                     "link_ms / rate_ms = CUDA events around each rank's own launches, max over ranks",
           "one_gpu": {"link_ms": one_link, "rate_ms": one_rate, "matches_per_s": n / ((one_link + one_rate) * 1e-3)},
           "speedup_vs_one_gpu": (one_link + one_rate) / ms_wall,
           "strong_scaling_efficiency": (one_link + one_rate) / ms_wall / G,
           "nvlink": {"rate_bytes_per_gpu": rate_nv, "rate_gbs_per_gpu": rate_nv / (ms_rate * 1e-3) / 1e9,
                      "link_bytes_per_gpu": link_nv, "link_gbs_per_gpu": link_nv / (ms_link * 1e-3) / 1e9,
                      "peak_gbs_per_direction": 770.0,
                      "note": "rows pushed to other GPUs (32 B each); links: " +
                              ("(player, slot) pairs and per-block counts stored to the owners, succ messages stored to the raters; "
                               if ladder_exchange else "id stream read from peers + succ messages stored to peers; ") +
                              "770 GB/s = measured peer copy per direction (B200_PROFILING.md)"},
           "link_path": "exchange (every rank reads its own slice and sends each pair to its owner)" if ladder_exchange
                        else "every rank reads the whole id stream",
           "bit_equal_to_one_gpu": unequal == 0.0, "max_abs_diff_vs_one_gpu": maxdiff,
           "check_max_abs_diff": prefix_diff,
           "check": "first %d matches through the sharded path vs the sequential __float128 oracle; full ladder "
                    "compared bit for bit with the one-GPU stream path" % min(n_prefix, n)}
    assert unequal == 0.0 and prefix_diff < 1e-9, out
    return out


def numa_policy(local, mode):
    """Memory policy for this process's pinned host buffers (set before they are allocated): 'local' =
    prefer the NUMA node of this rank's GPU, 'interleave' = spread pages over all nodes.  With several
    ranks copying at once the host side of the H2D copies is what bounds e2e.  Returns what was done."""
    import ctypes
    import glob
    import torch
    try:
        nodes = sorted(int(p.rsplit("node", 1)[1]) for p in glob.glob("/sys/devices/system/node/node[0-9]*"))
        pr = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        try:
            gpu_node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
        except Exception:
            gpu_node = -1
        info = {"nodes": nodes, "gpu_node": gpu_node, "mode": mode}
        if mode == "none" or len(nodes) < 2:
            return info
        libc = ctypes.CDLL("libc.so.6", use_errno=True)
        if mode == "local" and gpu_node >= 0:
            mask, pol = ctypes.c_ulong(1 << gpu_node), 1          # MPOL_PREFERRED
        else:
            mask, pol = ctypes.c_ulong(sum(1 << n for n in nodes)), 3     # MPOL_INTERLEAVE
        rc = libc.syscall(238, pol, ctypes.byref(mask), ctypes.c_ulong(64))   # set_mempolicy (x86_64)
        info["rc"] = rc if rc == 0 else -ctypes.get_errno()
        return info
    except Exception as exc:
        return {"error": repr(exc)}


# ---------------------------------------------------------------------------
def run_b200(a):
    import numpy as np
    import torch
    from analyzer_b200 import engine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the rating path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    n, u, p = a.matches, a.players, a.slots
    seed = SEED + rank                      # an independent ladder per GPU
    rater = engine.Rater(engine.Env(draw_probability=a.p_draw), local)
    gen = engine.make_gen(seed, u, slots=p, mode=1, p_draw=a.p_draw, zipf=a.zipf)
    ids_d, meta_d = engine.gen_matches(gen, 0, n, dev)
    table0 = engine.gen_table(seed, 0, u, dev)
    table = table0.clone()
    ptab = engine.PlayerTable.ladder(table, mode=1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    ev = lambda: torch.cuda.Event(enable_timing=True)
    timings = {"stream": ([], []), "rounds": ([], [])}
    last_sch = [None]
    final = {}

    def step(path, timed):
        table.copy_(table0)
        flush.fill_(1)                       # evict ids/table from L2 between steps
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        if path == "stream":
            pred = rater.link_pred_windows(ids_d, meta_d, u, ptab.mode_cols)
            e1.record()
            rater.rate_stream(ptab, ids_d, meta_d, pred, None, a.lanes)
        else:
            sch = rater.schedule(ids_d, meta_d, u, ptab.mode_cols)
            e1.record()
            rater.rate(ptab, sch, None, a.lanes)
            last_sch[0] = sch              # its round count is read after the timed region: nothing in a step waits for the GPU
        e2.record()
        if timed:
            torch.cuda.synchronize()
            timings[path][0].append(e0.elapsed_time(e1))
            timings[path][1].append(e1.elapsed_time(e2))

    other = "rounds" if a.path == "stream" else "stream"
    for path in (other, a.path):
        for _ in range(a.warmup):
            step(path, False)
    # the secondary path first (outside the contract's timed region, device-timed all the same) ...
    barrier()
    for _ in range(a.steps):
        step(other, True)
    final[other] = table.clone()
    # ... then the headline path inside the contract's timed region
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    t_wall = time.perf_counter()
    for _ in range(a.steps):
        step(a.path, True)
    barrier()
    t_wall = time.perf_counter() - t_wall
    clocks = sampler.stop()
    final[a.path] = table.clone()
    n_rounds = last_sch[0].n_rounds if last_sch[0] is not None else 0
    last_sch[0] = None
    assert rater.error_count() == 0
    # both paths rate the same ladder: their tables must agree to rounding
    assert float((final["stream"] - final["rounds"]).abs().max()) < 1e-9, "stream and rounds paths disagree"

    def mean_ms(path):
        sc, rt = timings[path]
        return sum(sc) / len(sc), sum(rt) / len(rt)

    ms_sched, ms_rate = mean_ms(a.path)
    ms_sched2, ms_rate2 = mean_ms(other)
    ms_step = ms_sched + ms_rate
    if dist is not None:
        t = torch.tensor([ms_step, ms_sched, ms_rate, ms_sched2, ms_rate2], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_step, ms_sched, ms_rate, ms_sched2, ms_rate2 = (float(x) for x in t.cpu())

    # ---- e2e: the reference-facing call with pinned host buffers -----------
    e2e = None
    if not a.skip_e2e:
        numa = numa_policy(local, a.numa) if world > 1 else None
        # the id column at the width it needs (ceil(log2(U + 1)) bits per id, widened on the device) ...
        ids_h = engine.pack_ids(ids_d.cpu().numpy().view(np.uint32), u, pin=True)
        meta_h = meta_d.cpu().to(torch.uint8).pin_memory()    # the meta word uses 8 bits: shipped as bytes
        tab_h = table0.cpu().pin_memory()
        out_h = torch.empty_like(tab_h).pin_memory()
        k_e2e = max(1, min(a.steps, 5))
        res = None
        for it in range(1 + k_e2e):          # 1 warm-up
            if it == 1:
                barrier()
                t0 = time.perf_counter()
            res = rater.rate_batch(ids_h, meta_h, tab_h, shared_col=-1, mode_cols=ptab.mode_cols[:6],
                                   want_outputs=False, lanes_per_match=a.lanes, table_out=out_h, path=a.path)
        barrier()
        dt = (time.perf_counter() - t0) / k_e2e
        if dist is not None:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.cpu()[0])
        # the e2e result must be the same table the device-resident path produced
        assert torch.equal(out_h, final[a.path].cpu()), "e2e table differs from the HBM-resident run"
        h2d = ids_h.nbytes + meta_h.numel() * meta_h.element_size() + tab_h.numel() * 8
        e2e = {"value": n * world / dt, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": out_h.numel() * 8 + 8 + 16, "ms_per_step": dt * 1e3, "steps": k_e2e,
               "numa": numa,
               "ingest": "ids packed at %d bits each (ts_unpack_ids), meta as bytes (ts_widen_meta), table float64; "
                         "copies pipelined with link + rate over windows" % ids_h.bits}

    peak, peak_src = hbm_peak()
    b_rate = rate_bytes_per_match(p, 1)
    achieved = b_rate * n / (ms_rate * 1e-3) / 1e9
    kernel = "k_rate_stream" if a.path == "stream" else "k_rate"
    traffic = None
    tp = os.path.join(ROOT, "profiles", "rate_kernel_traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp)).get(kernel, {})
            if tj.get("matches") == n and tj.get("slots") == p:
                traffic = tj.get("dram_bytes_per_launch")
        except Exception:
            pass
    b_sched = sched_bytes_per_match(p, u)
    passes_p = (max(1, u.bit_length()) + 7) // 8
    passes_r = (max(1, n_rounds.bit_length()) + 7) // 8
    pred_passes = -(-(n * p) // (16 << 20))
    # kernels of this library per step (memsets / copies are not kernels): the sort is
    # 1 histogram + 1 scan + one single-sweep kernel per pass
    launches = {"stream": (2 + passes_p) + pred_passes + 3,                 # sort, pred passes, rows load, rate_stream, rows store
                "rounds": (2 + passes_p) + pred_passes + 1 + (2 + passes_r) + 1 + 1 + 1}   # sort, pred passes, levels, sort, bounds, permute, rate
    line = {
        "metric": METRIC, "value": n * world / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "seed": hex(SEED), "rounds": n_rounds, "path": a.path,
                   "per_gpu": "one independent ladder per GPU, no data-path collective" if world > 1 else "single GPU",
                   "l2": "256 MiB flush between steps; ids+meta (%d MB) exceed the 126 MB L2; the %d MB player "
                         "table %s" % ((n * p * 4 + n * 4) >> 20, (u * 16) >> 20,
                                       "is L2-resident by nature" if u * 32 < (100 << 20) else
                                       "(versioned rows: twice that) does not fit L2 either"),
                   "lanes_per_match": a.lanes or (2 if p == 6 else 8)},
        "phases": {"schedule_ms": ms_sched, "rate_ms": ms_rate,
                   "rate_only_matches_per_s": n * world / (ms_rate * 1e-3),
                   "schedule_only_matches_per_s": n * world / (ms_sched * 1e-3)},
        "paths": {a.path: {"schedule_ms": ms_sched, "rate_ms": ms_rate, "matches_per_s": n * world / (ms_step * 1e-3)},
                  other: {"schedule_ms": ms_sched2, "rate_ms": ms_rate2,
                          "matches_per_s": n * world / ((ms_sched2 + ms_rate2) * 1e-3)},
                  "note": "stream = sort + link + dataflow rating in stream order; rounds = sort + link + rounds by "
                          "dataflow + round buckets + one grid barrier per round; same ratings (checked)"},
        "roofline": {"kernel": kernel, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_match": b_rate,
                     "note": "one launch per step; duration = CUDA events around ts_rate on its stream"},
        "roofline_schedule": {"bound": "hbm", "achieved": b_sched * n / (ms_sched * 1e-3) / 1e9, "peak": peak,
                              "unit": "GB/s", "frac": b_sched * n / (ms_sched * 1e-3) / 1e9 / peak,
                              "algorithmic_bytes_per_match": b_sched},
        "gpu_launches": launches[a.path] * a.steps,
        "clocks": clocks,
        "wall_s_timed_region": t_wall,
    }
    if e2e:
        line["e2e"] = e2e
    if world > 1 and not a.no_sharded:
        # BASELINE.json configs[4] (strong scaling): ONE ladder over all ranks, on the same process group.
        # A deadline guards the contract: whatever happens in there, rank 0 prints its line and every rank exits 0.
        del ids_d, meta_d, table, table0, flush, final
        rater._ws = rater._vt = None
        torch.cuda.empty_cache()

        def give_up():
            if rank == 0:
                line["sharded"] = {"error": "no result within %d s" % a.sharded_deadline}
                print(json.dumps(line), flush=True)
            os._exit(0)

        timer = threading.Timer(a.sharded_deadline, give_up)
        timer.daemon = True
        timer.start()
        try:
            line["sharded"] = sharded_block(a, rater, engine, dist, rank, world, dev, a.sharded_matches,
                                            a.sharded_players, p)
        except Exception as exc:
            line["sharded"] = {"error": repr(exc)}
        timer.cancel()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0
    if world == 1 and p in (6, 10):
        try:
            line["faithful"] = faithful_rate(rater, engine, ids_d, meta_d, table0, u, a.lanes)
        except Exception as exc:
            line["faithful"] = {"error": repr(exc)}
    if world == 1 and not a.no_cpu_baseline:
        try:
            line["object_api"] = object_api_rate(ids_d, meta_d)
        except Exception as exc:
            line["object_api"] = {"error": repr(exc)}
    if world == 1 and not a.no_cpu_baseline:
        sample = min(n, 200_000)
        ids = ids_d[:sample].cpu().numpy().view(np.uint32)
        meta = meta_d[:sample].cpu().numpy().view(np.uint32)
        tab = table0.cpu().numpy()
        value, n_done, dt = cpu_reference_rate(ids, meta, tab, a.cpu_seconds, a.p_draw)
        full, n_full = cpu_rate_match_rate(ids, meta, 3.0)
        cb = {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
              "sample": "first %d matches of the same stream (%.1f s); oracle/rater_port.py rate_ladder (one "
                        "env.rate per match, like the GPU workload) on the restated trueskill API at 50 digits; "
                        "the pinned trueskill==0.4.4 is not installable here" % (n_done, dt),
              "rate_match_value": full,
              "rate_match_note": "full rate_match (2 x rate + quality) on %d matches" % n_full}
        try:
            cb.update(cpu_c_baselines(ids, meta, tab, u))
        except Exception as exc:   # the C oracle is optional context, never the product
            cb["c_error"] = repr(exc)
        line["cpu_baseline"] = cb
    if world == 1 and not a.no_configs:
        # the other single-GPU configurations BASELINE.json names, at their stated sizes
        del ids_d, meta_d, table, table0, flush, final
        rater._ws = rater._vt = None
        torch.cuda.empty_cache()
        raters = {a.p_draw: rater}

        def rater_of(p_draw):
            if p_draw not in raters:
                raters[p_draw] = engine.Rater(engine.Env(draw_probability=p_draw), local)
            return raters[p_draw]

        cfgs = {}
        for key, kw in (
                ("config3", dict(name="100M 5v5 matches over 10M players, Zipf activity, 2% draws (BASELINE.json configs[2])",
                                 n=100_000_000, u=10_000_000, p=10, zipf=True, p_draw=0.02, layout="ladder")),
                ("config4_ladder", dict(name="one of the 8 ladders of BASELINE.json configs[3]: 125M 3v3 matches over 12.5M "
                                             "players, K=1 pair per player", n=125_000_000, u=12_500_000, p=6, zipf=False,
                                        p_draw=0.0, layout="ladder")),
                ("config4_ladder_k2", dict(name="the same ladder with the reference's shared + queue pair per player (K=2: one "
                                                "32-byte sector of table per player)", n=125_000_000, u=12_500_000, p=6,
                                           zipf=False, p_draw=0.0, layout="pair", want_rounds=False))):
            try:
                cfgs[key] = config_block(rater_of, engine, lanes=a.lanes, peak=peak, **kw)
            except Exception as exc:
                cfgs[key] = {"error": repr(exc)}
                torch.cuda.empty_cache()
        line["configs"] = cfgs
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


def run_sharded(a):
    """--sharded: only the one-ladder-over-all-ranks block (sharded_block), at --matches / --players;
    on one GPU --emulate G plays G ranks in one process."""
    import torch
    from analyzer_b200 import engine
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    rater = engine.Rater(engine.Env(draw_probability=a.p_draw), local)
    out = sharded_block(a, rater, engine, dist, rank, world, dev, a.matches, a.players, a.slots, steps=a.steps,
                        warmup=a.warmup, emulate=a.emulate if world == 1 else 0)
    if rank == 0:
        line = {"metric": METRIC, "value": out["matches_per_s"], "unit": UNIT, "n_gpus": world, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": out["ms_per_step"], "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": out["workload"], "seed": hex(SEED), "path": "stream-sharded"}, "sharded": out}
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        sys.exit(run_reference(args))
    sys.exit(run_sharded(args) if args.sharded else run_b200(args))
