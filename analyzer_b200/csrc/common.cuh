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
    unsigned long long *d_ctrl;  // [0] barrier arrive, [1] barrier release, [2] error count, [3..7] scratch
};

enum { CTRL_ARRIVE = 0, CTRL_RELEASE = 1, CTRL_ERRORS = 2, CTRL_TAIL = 3, CTRL_WORDS = 16 };

void ts_set_error(const char *fmt, ...);
int ts_cuda_fail(cudaError_t e, const char *what);
#define TS_CUDA(call)                                             \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return ts_cuda_fail(e__, #call);  \
    } while (0)

// ---------------------------------------------------------------------------
// Grid-wide barrier for cooperatively launched persistent kernels.
// Arrive on a monotone counter; the block that arrives last may run a serial
// step (the lambda) before it publishes the release word every block spins on.
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <class Serial>
__device__ __forceinline__ void grid_barrier(unsigned long long *ctrl, unsigned long long &generation,
                                             Serial serial)
{
    __syncthreads();
    generation += 1;
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned long long arrived = atomicAdd(&ctrl[CTRL_ARRIVE], 1ull) + 1ull;
        if (arrived == generation * (unsigned long long)gridDim.x) {
            serial();
            __threadfence();
            st_release_u64(&ctrl[CTRL_RELEASE], generation);
        } else {
            while (ld_acquire_u64(&ctrl[CTRL_RELEASE]) < generation) {
            }
        }
    }
    __syncthreads();
}

struct NoSerial {
    __device__ __forceinline__ void operator()() const {}
};

__device__ __forceinline__ bool meta_is_gated(uint32_t meta, const int8_t *mode_col)
{
    uint32_t mode = (meta >> TS_META_MODE_SHIFT) & 15u;
    bool unsupported = mode >= 6u || mode_col[mode] < 0;
    return unsupported || (meta & (TS_META_AFK | TS_META_INVALID)) != 0u;
}

struct ModeCols {
    int8_t col[16];
};
