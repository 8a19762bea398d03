"""ctypes binding of libtsrate.so (include/tsrate.h).  There is no CPU
fallback: if the CUDA library is missing or cannot be loaded this module
raises, and so does everything built on it."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
# TSRATE_LIB: another build of the same library (tuning variants, scripts/build_variants.sh)
LIB_PATH = os.environ.get("TSRATE_LIB") or os.path.join(_HERE, "libtsrate.so")
SOURCES = ["api.cu", "gen.cu", "ingest.cu", "sched.cu", "rate.cu", "stream.cu", "stream_pair.cu", "slots.cu", "shard.cu", "static_sched.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--extended-lambda", "-shared", "-Xcompiler", "-fPIC"]

MAX_SLOTS = 16
EMPTY_SLOT = 0xFFFFFFFF
META_WIN_A, META_WIN_B, META_AFK, META_INVALID = 1, 2, 4, 8
META_MODE_SHIFT = 4
MODE_UNSUPPORTED = 15
STREAM_LOAD, STREAM_STORE = 1, 2
ST_OK, ST_SKIP_MODE, ST_AFK, ST_INVALID, ST_FP_ERROR, ST_VALUE_ERROR = range(6)

EXPORTS = ["ts_ctx_create", "ts_ctx_destroy", "ts_last_error", "ts_sm_count",
           "ts_schedule_workspace_bytes", "ts_schedule", "ts_rate", "ts_error_count",
           "ts_reset_errors", "ts_gen_matches", "ts_gen_table",
           "ts_schedule_static", "ts_schedule_static_max_slot_entries",
           "ts_link_workspace_bytes", "ts_link_pred", "ts_rate_stream_workspace_bytes", "ts_rate_stream",
           "ts_stream_stalled", "ts_rate_stream_table_workspace_bytes",
           "ts_seed_table", "ts_widen_meta", "ts_unpack_ids", "ts_error_counts", "ts_slot_rows_bytes", "ts_link_succ", "ts_rate_slots", "ts_ingest_rows",
           "ts_shard_alloc", "ts_shard_open", "ts_shard_close", "ts_shard_free", "ts_shard_local_matches",
           "ts_shard_link_workspace_bytes", "ts_shard_inbox_cap", "ts_shard_reset", "ts_shard_link", "ts_shard_rate", "ts_shard_status",
           "ts_shard_exchange_workspace_bytes", "ts_shard_count", "ts_shard_send"]
SHARD_BLOCK, SHARD_MAX = 32, 16
SHARD_STALLED, SHARD_OVERFLOW, SHARD_DUPLICATE = 1, 2, 4


class TsParams(C.Structure):
    _fields_ = [("beta", C.c_double), ("tau", C.c_double),
                ("draw_margin", C.c_double * (MAX_SLOTS + 1)),
                ("draw_is_zero", C.c_int32), ("reserved", C.c_int32)]


class TsTable(C.Structure):
    _fields_ = [("data", C.c_void_p), ("seed", C.c_void_p), ("n_players", C.c_int64),
                ("row_pairs", C.c_int32), ("shared_col", C.c_int32), ("mode_col", C.c_int8 * 16)]


class TsSchedInfo(C.Structure):
    _fields_ = [("n_rounds", C.c_uint32), ("n_gated", C.c_uint32), ("overflow", C.c_uint32),
                ("reserved", C.c_uint32)]


class TsOutputs(C.Structure):
    _fields_ = [("quality", C.c_void_p), ("out_shared", C.c_void_p), ("out_delta", C.c_void_p),
                ("out_queue", C.c_void_p), ("status", C.c_void_p)]


class TsShardView(C.Structure):
    _fields_ = [("rank", C.c_int32), ("n_shards", C.c_int32),
                ("ids", C.c_void_p * 16), ("meta", C.c_void_p * 16), ("succ", C.c_void_p * 16),
                ("rows", C.c_void_p * 16), ("table", C.c_void_p * 16), ("inbox", C.c_void_p * 16),
                ("inbox_count", C.c_void_p * 16), ("inbox_cap", C.c_int64), ("ring", C.c_void_p * 16),
                ("ring_final", C.c_void_p * 16), ("ring_tail", C.c_void_p * 16), ("ring_cap", C.c_int64), ("ring_epoch", C.c_int64),
                ("link_counts", C.c_void_p * 16), ("in_keys", C.c_void_p * 16), ("in_vals", C.c_void_p * 16),
                ("in_count", C.c_void_p * 16), ("in_cap", C.c_int64), ("link_blocks_cap", C.c_int64)]


class TsRowColumns(C.Structure):
    _fields_ = [("n_rows", C.c_int64), ("match", C.c_void_p), ("player", C.c_void_p), ("game_mode", C.c_void_p),
                ("n_rosters", C.c_void_p), ("roster", C.c_void_p), ("skill_tier", C.c_void_p), ("winner", C.c_void_p),
                ("went_afk", C.c_void_p), ("rank_points_ranked", C.c_void_p), ("rank_points_blitz", C.c_void_p),
                ("trueskill_mu", C.c_void_p), ("trueskill_sigma", C.c_void_p), ("queue", C.c_void_p)]


class TsGen(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_players", C.c_uint64), ("slots", C.c_int32),
                ("mode", C.c_int32), ("draw_thr", C.c_uint32), ("afk_thr", C.c_uint32),
                ("zipf", C.c_int32), ("zipf_a", C.c_double), ("zipf_b", C.c_double),
                ("zipf_q", C.c_double)]


def sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def build(force=False, verbose=False):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    srcs = sources()
    deps = srcs + [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "vw.cuh"), os.path.join(CSRC, "vtable.inc"), os.path.join(CSRC, "stream_common.cuh"),
                   os.path.join(os.path.dirname(_HERE), "include", "tsrate.h")]
    if not force and os.path.exists(LIB_PATH) and \
            all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    cmd = ["nvcc"] + NVCC_FLAGS + ["-o", LIB_PATH] + srcs
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd, cwd=CSRC)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "analyzer_b200: %s is missing -- run `python -c 'import __graft_entry__ as g; "
                "g.build()'`; there is no CPU fallback for the rating path" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.ts_last_error.restype = C.c_char_p
        L.ts_schedule_workspace_bytes.restype = C.c_int64
        L.ts_schedule_workspace_bytes.argtypes = [C.c_int64, C.c_int32]
        L.ts_error_count.restype = C.c_int64
        L.ts_error_count.argtypes = [C.c_void_p, C.c_void_p]
        L.ts_reset_errors.argtypes = [C.c_void_p, C.c_void_p]
        L.ts_ctx_create.argtypes = [C.POINTER(TsParams), C.c_int, C.POINTER(C.c_void_p)]
        L.ts_ctx_destroy.argtypes = [C.c_void_p]
        L.ts_ctx_destroy.restype = None
        L.ts_sm_count.argtypes = [C.c_void_p]
        L.ts_schedule.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                  C.POINTER(C.c_int8), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                  C.POINTER(C.c_uint32), C.c_void_p]
        L.ts_schedule_static.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                         C.POINTER(C.c_int8), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p]
        L.ts_schedule_static_max_slot_entries.restype = C.c_int64
        L.ts_rate.argtypes = [C.c_void_p, C.POINTER(TsTable), C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.POINTER(TsOutputs),
                              C.c_int32, C.c_void_p]
        L.ts_link_workspace_bytes.restype = C.c_int64
        L.ts_link_workspace_bytes.argtypes = [C.c_int64, C.c_int32]
        L.ts_link_pred.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                   C.POINTER(C.c_int8), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                   C.c_void_p]
        L.ts_rate_stream_workspace_bytes.restype = C.c_int64
        L.ts_rate_stream_workspace_bytes.argtypes = [C.c_int64]
        L.ts_rate_stream.argtypes = [C.c_void_p, C.POINTER(TsTable), C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int64, C.c_int32, C.c_int64, C.c_int32, C.POINTER(TsOutputs), C.c_int32,
                                     C.c_void_p, C.c_int64, C.c_void_p]
        L.ts_rate_stream_table_workspace_bytes.restype = C.c_int64
        L.ts_rate_stream_table_workspace_bytes.argtypes = [C.POINTER(TsTable)]
        L.ts_shard_alloc.argtypes = [C.c_int, C.c_int64, C.POINTER(C.c_void_p), C.c_void_p]
        L.ts_shard_open.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
        L.ts_shard_close.argtypes = [C.c_void_p]
        L.ts_shard_free.argtypes = [C.c_void_p]
        L.ts_shard_local_matches.restype = C.c_int64
        L.ts_shard_local_matches.argtypes = [C.c_int64, C.c_int32, C.c_int32]
        L.ts_shard_inbox_cap.restype = C.c_int64
        L.ts_shard_inbox_cap.argtypes = [C.c_int64, C.c_int32, C.c_int32]
        L.ts_shard_link_workspace_bytes.restype = C.c_int64
        L.ts_shard_link_workspace_bytes.argtypes = [C.c_int64, C.c_int64]
        L.ts_shard_reset.argtypes = [C.c_void_p, C.POINTER(TsShardView), C.c_int64, C.c_void_p]
        L.ts_shard_link.argtypes = [C.c_void_p, C.POINTER(TsShardView), C.c_int64, C.c_int32, C.c_int64, C.c_int8,
                                    C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
        L.ts_shard_rate.argtypes = [C.c_void_p, C.POINTER(TsShardView), C.c_int64, C.c_int32, C.c_int64, C.c_int8,
                                    C.c_int64, C.POINTER(TsOutputs), C.c_int32, C.c_int32, C.c_void_p]
        L.ts_shard_status.argtypes = [C.c_void_p, C.c_void_p]
        L.ts_shard_exchange_workspace_bytes.restype = C.c_int64
        L.ts_shard_exchange_workspace_bytes.argtypes = [C.c_int64, C.c_int32]
        L.ts_shard_count.argtypes = [C.c_void_p, C.POINTER(TsShardView), C.c_int64, C.c_int32, C.c_int64, C.c_int8, C.c_void_p]
        L.ts_shard_send.argtypes = [C.c_void_p, C.POINTER(TsShardView), C.c_int64, C.c_int32, C.c_int64, C.c_int8,
                                    C.c_void_p, C.c_int64, C.c_void_p]
        L.ts_error_counts.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_void_p]
        L.ts_slot_rows_bytes.restype = C.c_int64
        L.ts_slot_rows_bytes.argtypes = [C.c_int64, C.c_int32]
        L.ts_link_succ.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.POINTER(TsTable), C.c_uint32,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.ts_rate_slots.argtypes = [C.c_void_p, C.POINTER(TsTable), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                    C.c_uint32, C.POINTER(TsOutputs), C.c_int32, C.c_void_p, C.c_void_p]
        L.ts_ingest_rows.argtypes = [C.POINTER(TsRowColumns), C.c_double, C.POINTER(C.c_double)] + [C.c_void_p] * 9 + \
            [C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
        L.ts_stream_stalled.argtypes = [C.c_void_p, C.c_void_p]
        L.ts_seed_table.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double,
                                    C.POINTER(C.c_double), C.c_void_p, C.c_void_p, C.c_void_p]
        L.ts_widen_meta.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.ts_unpack_ids.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]
        L.ts_gen_matches.argtypes = [C.POINTER(TsGen), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                     C.c_void_p]
        L.ts_gen_table.argtypes = [C.c_uint64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


class TsError(RuntimeError):
    pass


def check(rc):
    if rc < 0:
        raise TsError("libtsrate error %d: %s" % (rc, lib().ts_last_error().decode()))
    return rc
