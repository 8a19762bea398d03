"""ts_schedule_static -- the scheduler entry point that allocates nothing.

Some hosts of this library cannot allocate device memory at all: a plugin
living inside a service that owns the device allocator, or a process whose HBM
another framework has already reserved.  For them libtsrate.so carries its
scheduler scratch as a fixed 2 MiB array inside the library image, and
ts_schedule_static runs the normal scheduler on it.  The caller still passes
inputs and outputs.

This test calls it through the C ABI, twice in a row on the same array (a
larger batch first, then a smaller and different one, so stale contents of the
shared scratch would show), and demands results bit-identical to ts_schedule
run with a caller-provided workspace and to the sequential schedule oracle.  It
also checks that the driver's free-memory figure does not move across a call,
that a batch of exactly the stated maximum fits, and that one beyond it is
refused with a clear error rather than overrunning.  The
file name sorts last on purpose: run after everything else, or alone in a
fresh process, it must pass either way.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


def _static_schedule(lib, L, ids_d, meta_d, n_players, cols, wait=True):
    n, p = ids_d.shape
    dev = ids_d.device
    i32 = lambda *shape: torch.empty(shape, dtype=torch.int32, device=dev)
    out = dict(round=i32(n), order=i32(n), offsets=i32(n + 2), ids_sorted=i32(n, p), meta_sorted=i32(n),
               info=torch.zeros(4, dtype=torch.int32, device=dev))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]            # cudaMemGetInfo: what the driver has left
    n_rounds = C.c_uint32(0)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.ts_schedule_static(
        dev.index or 0, ptr(ids_d), ptr(meta_d), n, p, n_players, (C.c_int8 * 16)(*cols),
        ptr(out["round"]), ptr(out["order"]), ptr(out["offsets"]), n + 2, ptr(out["ids_sorted"]),
        ptr(out["meta_sorted"]), ptr(out["info"]), C.byref(n_rounds) if wait else None,
        C.c_void_p(torch.cuda.current_stream().cuda_stream))
    L.check(rc)
    torch.cuda.synchronize()
    out["free_delta"] = free_before - torch.cuda.mem_get_info(dev)[0]
    out["n_rounds"] = int(n_rounds.value) if wait else int(_u32(out["info"])[0])
    return out


def test_static_scratch_schedule_matches_ts_schedule_twice_in_a_row(c_oracle):
    from analyzer_b200 import _lib as L
    from analyzer_b200 import engine
    lib = L.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    cols = [-1, 0] + [-1] * 14
    rater = engine.Rater(engine.Env(), 0)
    assert lib.ts_schedule_static_max_slot_entries() == 65536
    for call, (n, n_players, slots, wait, kw) in enumerate([
            (6500, 900, 10, True, dict(zipf=True, p_draw=0.02, p_afk=0.05)),    # close to the 65 536-entry limit
            (3000, 2000, 6, True, dict(p_afk=0.1)),                             # smaller and different: stale scratch would show
            (8192, 5000, 8, False, dict(p_afk=0.02))]):                         # exactly 65 536 entries; the call that only enqueues
        gen = engine.make_gen(0xA11CE0D0 + call, n_players, slots=slots, mode=1, **kw)
        ids_d, meta_d = engine.gen_matches(gen, 0, n, dev)
        _static_schedule(lib, L, ids_d, meta_d, n_players, cols, wait)     # loads the kernels this shape needs (lazy module loading)
        got = _static_schedule(lib, L, ids_d, meta_d, n_players, cols, wait)
        assert got["free_delta"] == 0                                      # the call itself takes no device memory
        want = rater.schedule(ids_d, meta_d, n_players, cols)       # the existing entry point, caller-provided workspace
        R = want.n_rounds
        assert got["n_rounds"] == R and R > 3
        assert np.array_equal(_u32(got["round"]), _u32(want.round))
        assert np.array_equal(_u32(got["order"]), _u32(want.order))
        assert np.array_equal(_u32(got["offsets"])[:R + 2], _u32(want.round_offsets)[:R + 2])
        assert np.array_equal(_u32(got["ids_sorted"]), _u32(want.ids_sorted))
        assert np.array_equal(_u32(got["meta_sorted"]), _u32(want.meta_sorted))
        assert np.array_equal(_u32(got["info"]), _u32(want.info))
        # and against the sequential schedule oracle, so that the two CUDA entry points cannot be wrong together
        rnd, order, offs, R0 = c_oracle.schedule(_u32(ids_d), _u32(meta_d), cols[:6], n_players)
        assert R0 == R and np.array_equal(_u32(got["round"]), rnd) and np.array_equal(_u32(got["order"]), order)
        assert np.array_equal(_u32(got["offsets"])[:R + 2], offs)
    # one entry too many: refused, nothing launched
    n = 65536 // 6 + 1
    ids_d = torch.zeros((n, 6), dtype=torch.int32, device=dev)
    meta_d = torch.zeros((n,), dtype=torch.int32, device=dev)
    with pytest.raises(L.TsError, match="exceeds the built-in scratch"):
        _static_schedule(lib, L, ids_d, meta_d, 10, cols)
    rater.close()
