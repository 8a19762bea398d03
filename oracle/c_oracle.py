"""ORACLE -- test infrastructure, NOT product code.

ctypes front end of oracle/ts_oracle.c (sequential closed-form rater in
__float128 / fp64, the schedule oracle and the host twin of the synthetic
generator).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
leg may import this.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libtsoracle.so")
MAXP = 16
EMPTY = 0xFFFFFFFF

ST_OK, ST_SKIP_MODE, ST_AFK, ST_INVALID, ST_FP_ERROR, ST_VALUE_ERROR = range(6)


class Params(C.Structure):
    _fields_ = [("beta", C.c_double), ("tau", C.c_double),
                ("draw_margin", C.c_double * (MAXP + 1)), ("draw_is_zero", C.c_int32)]


class Gen(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_players", C.c_uint64), ("slots", C.c_int32),
                ("mode", C.c_int32), ("draw_thr", C.c_uint32), ("afk_thr", C.c_uint32),
                ("zipf", C.c_int32), ("zipf_a", C.c_double), ("zipf_b", C.c_double),
                ("zipf_q", C.c_double)]


def build(force=False):
    src = os.path.join(_HERE, "ts_oracle.c")
    if (not force and os.path.exists(_SO)
            and os.path.getmtime(_SO) >= os.path.getmtime(src)):
        return _SO
    subprocess.check_call(["make", "-C", _HERE, "-B", "libtsoracle.so"],
                          stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.tso_rate_sequential.restype = C.c_int64
        _lib.tso_rate_rounds_omp.restype = C.c_int64
        _lib.tso_schedule.restype = C.c_int64
        _lib.tso_max_threads.restype = C.c_int
    return _lib


def draw_margins(beta, draw_probability):
    """calc_draw_margin of the trueskill package for n = 0..MAXP players, at
    50 digits through the shim's ppf; exactly 0 when draw_probability == 0
    (the analytic limit)."""
    out = [0.0] * (MAXP + 1)
    if draw_probability == 0:
        return out
    import sys
    shim = os.path.join(_HERE, "shim")
    if shim not in sys.path:
        sys.path.insert(0, shim)
    import mpmath
    import trueskill
    mpmath.mp.dps = 50
    env = trueskill.TrueSkill(backend="mpmath", beta=beta, draw_probability=draw_probability)
    for n in range(1, MAXP + 1):
        out[n] = float(trueskill.calc_draw_margin(draw_probability, n, env))
    return out


def make_params(beta=1000.0, tau=10.0, draw_probability=0.0):
    p = Params()
    p.beta, p.tau = beta, tau
    for i, v in enumerate(draw_margins(beta, draw_probability)):
        p.draw_margin[i] = v
    p.draw_is_zero = 1 if draw_probability == 0 else 0
    return p


def _ptr(a, ctype):
    return a.ctypes.data_as(C.POINTER(ctype)) if a is not None else None


def _mode_cols(mode_cols):
    a = np.full(16, -1, dtype=np.int8)
    a[:len(mode_cols)] = mode_cols
    return a


def rate_sequential(params, ids, meta, table, shared_col, mode_cols, seed=None, quad=True,
                    want_outputs=False):
    """table: float64 [U, C, 2], updated in place.  Returns dict of outputs."""
    n, p = ids.shape
    assert ids.dtype == np.uint32 and meta.dtype == np.uint32 and table.dtype == np.float64
    assert ids.flags.c_contiguous and table.flags.c_contiguous
    cols = table.shape[1]
    mc = _mode_cols(mode_cols)
    status = np.zeros(n, dtype=np.uint8)
    out = {"status": status}
    q = sh = de = qu = None
    if want_outputs:
        q = np.full(n, np.nan)
        sh = np.full((n, p, 2), np.nan)
        de = np.full((n, p), np.nan)
        qu = np.full((n, p, 2), np.nan)
        out.update(quality=q, out_shared=sh, out_delta=de, out_queue=qu)
    bad = lib().tso_rate_sequential(
        C.byref(params), C.c_int64(n), C.c_int(p), _ptr(ids, C.c_uint32), _ptr(meta, C.c_uint32),
        _ptr(table, C.c_double), C.c_int(cols), C.c_int(shared_col), _ptr(mc, C.c_int8),
        _ptr(seed, C.c_double), C.c_int(1 if quad else 0),
        _ptr(q, C.c_double), _ptr(sh, C.c_double), _ptr(de, C.c_double), _ptr(qu, C.c_double),
        _ptr(status, C.c_uint8))
    out["errors"] = int(bad)
    return out


def schedule(ids, meta, mode_cols, n_players):
    n, p = ids.shape
    mc = _mode_cols(mode_cols)
    rnd = np.zeros(n, dtype=np.uint32)
    order = np.zeros(n, dtype=np.uint32)
    cap = n + 2
    offs = np.zeros(cap, dtype=np.uint32)
    r = lib().tso_schedule(C.c_int64(n), C.c_int(p), _ptr(ids, C.c_uint32), _ptr(meta, C.c_uint32),
                           _ptr(mc, C.c_int8), C.c_int64(n_players), _ptr(rnd, C.c_uint32),
                           _ptr(order, C.c_uint32), _ptr(offs, C.c_uint32), C.c_int64(cap))
    assert r >= 0
    return rnd, order, offs[:r + 2].copy(), int(r)


def rate_rounds_omp(params, ids, meta, table, shared_col, mode_cols, order, offsets, n_rounds,
                    seed=None, threads=0):
    n, p = ids.shape
    mc = _mode_cols(mode_cols)
    return int(lib().tso_rate_rounds_omp(
        C.byref(params), C.c_int64(n), C.c_int(p), _ptr(ids, C.c_uint32), _ptr(meta, C.c_uint32),
        _ptr(table, C.c_double), C.c_int(table.shape[1]), C.c_int(shared_col), _ptr(mc, C.c_int8),
        _ptr(seed, C.c_double), _ptr(order, C.c_uint32), _ptr(offsets, C.c_uint32),
        C.c_uint32(n_rounds), C.c_int(threads)))


def max_threads():
    return int(lib().tso_max_threads())


def make_gen(seed, n_players, slots=6, mode=1, p_draw=0.0, p_afk=0.0, zipf=False, zipf_q=1000.0):
    g = Gen()
    g.seed, g.n_players, g.slots, g.mode = seed, n_players, slots, mode
    g.draw_thr = min(int(p_draw * 2 ** 32), 2 ** 32 - 1)
    g.afk_thr = min(int(p_afk * 2 ** 32), 2 ** 32 - 1)
    g.zipf = 1 if zipf else 0
    g.zipf_q = zipf_q
    g.zipf_a = zipf_q ** 0.2
    g.zipf_b = (n_players + zipf_q) ** 0.2 - zipf_q ** 0.2
    return g


def gen_matches(gen, first, count):
    ids = np.empty((count, gen.slots), dtype=np.uint32)
    meta = np.empty(count, dtype=np.uint32)
    lib().tso_gen_matches(C.byref(gen), C.c_int64(first), C.c_int64(count),
                          _ptr(ids, C.c_uint32), _ptr(meta, C.c_uint32))
    return ids, meta


def gen_table(seed, first, count):
    t = np.empty((count, 1, 2), dtype=np.float64)
    lib().tso_gen_table(C.c_uint64(seed), C.c_int64(first), C.c_int64(count), _ptr(t, C.c_double))
    return t
