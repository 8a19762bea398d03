// Shared device/host helpers of libtsrate (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/tsrate.h"

#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ < 1000
#error "libtsrate is written for sm_100a (B200) only"
#endif

struct ts_ctx {
    int device;
    int sm_count;
    ts_params params;
    // device control words (the only memory the library owns)
    unsigned long long *d_ctrl;  // TS_CTRL_BYTES of control words, layout below
};

// Control words (device): word 0 barrier counter, 2 matches that ended in TS_ST_FP_ERROR,
// 3 queue tail, 4 tail snapshot, 8..10 rotating per-level words of the frontier sweep;
// sticky until ts_reset_errors: 5 stall flag of the streaming kernels, 12 matches that
// ended in TS_ST_VALUE_ERROR, 13 TS_SHARD_* bits of the sharded ladder.
enum { CTRL_ARRIVE = 0, CTRL_ERRORS = 2, CTRL_TAIL = 3, CTRL_STALLED = 5, CTRL_LEVEL = 8, CTRL_VALUE_ERRORS = 12,
       CTRL_SHARD = 13, CTRL_WORDS = 16 };
constexpr size_t TS_CTRL_BYTES = CTRL_WORDS * 8;

void ts_set_error(const char *fmt, ...);
int ts_reset_barrier(unsigned long long *d_ctrl, cudaStream_t stream);   // arrive + release copies := 0
int ts_cuda_fail(cudaError_t e, const char *what);
#define TS_CUDA(call)                                             \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return ts_cuda_fail(e__, #call);  \
    } while (0)

// ---------------------------------------------------------------------------
// Grid-wide barrier for cooperatively launched persistent kernels: every block
// adds 1 to a monotone counter with a fire-and-forget reduction (no return
// trip) after fencing its writes, then polls the counter itself until the whole
// grid has arrived.  One L2 round trip less than an arrive/release pair: on
// B200 a coherent L2 round trip costs ~0.6 us and the barrier sits on the
// critical path of every round.
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_add_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ void grid_barrier(unsigned long long *ctrl, unsigned long long &generation)
{
    __syncthreads();
    generation += 1;
    if (threadIdx.x == 0) {
        __threadfence();                                   // this block's writes are visible ...
        red_add_u64(&ctrl[CTRL_ARRIVE], 1ull);             // ... before its arrival is
        const unsigned long long target = generation * (unsigned long long)gridDim.x;
        while (ld_relaxed_u64(&ctrl[CTRL_ARRIVE]) < target) {
        }
        __threadfence();
    }
    __syncthreads();
}

__device__ __forceinline__ bool meta_is_gated(uint32_t meta, const int8_t *mode_col)
{
    uint32_t mode = (meta >> TS_META_MODE_SHIFT) & 15u;
    bool unsupported = mode >= 6u || mode_col[mode] < 0;
    return unsupported || (meta & (TS_META_AFK | TS_META_INVALID)) != 0u;
}

struct ModeCols {
    int8_t col[16];
};
