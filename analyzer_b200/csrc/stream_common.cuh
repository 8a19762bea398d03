// Pieces shared by the streaming rating kernels (stream.cu: one pair per player;
// stream_pair.cu: shared + queue pair): self-validating row words, cp.async staging,
// group shuffles, tuning knobs.
#pragma once
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "vw.cuh"

namespace tsstream {

// SYS = rows may live in a peer GPU's memory: system scope (peer accesses bypass the local L2)
template <bool SYS = false>
__device__ __forceinline__ void ld_row(const unsigned long long *row, unsigned long long (&w)[4])
{
    if (SYS) {
        asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[0]), "=l"(w[1]) : "l"(row) : "memory");
        asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[2]), "=l"(w[3]) : "l"(row + 2) : "memory");
    } else {
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[0]), "=l"(w[1]) : "l"(row) : "memory");
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[2]), "=l"(w[3]) : "l"(row + 2) : "memory");
    }
}
template <bool SYS = false>
__device__ __forceinline__ void st_row(unsigned long long *row, double mu, double sigma, uint32_t ver)
{
    const unsigned long long v = (unsigned long long)ver << 32;
    const unsigned long long m = (unsigned long long)__double_as_longlong(mu);
    const unsigned long long s = (unsigned long long)__double_as_longlong(sigma);
    unsigned long long w0 = (m & 0xFFFFFFFFull) | v, w1 = (m >> 32) | v;
    unsigned long long w2 = (s & 0xFFFFFFFFull) | v, w3 = (s >> 32) | v;
    if (SYS) {
        asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(row), "l"(w0), "l"(w1) : "memory");
        asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(row + 2), "l"(w2), "l"(w3) : "memory");
    } else {
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(row), "l"(w0), "l"(w1) : "memory");
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(row + 2), "l"(w2), "l"(w3) : "memory");
    }
}
__device__ __forceinline__ bool row_has_version(const unsigned long long (&w)[4], uint32_t ver)
{
    return (uint32_t)(w[0] >> 32) == ver && (uint32_t)(w[1] >> 32) == ver && (uint32_t)(w[2] >> 32) == ver &&
           (uint32_t)(w[3] >> 32) == ver;
}
__device__ __forceinline__ double row_mu(const unsigned long long (&w)[4])
{
    return __longlong_as_double((long long)((w[0] & 0xFFFFFFFFull) | (w[1] << 32)));
}
__device__ __forceinline__ double row_sigma(const unsigned long long (&w)[4])
{
    return __longlong_as_double((long long)((w[2] & 0xFFFFFFFFull) | (w[3] << 32)));
}

__device__ __forceinline__ void cp_async_8(void *smem, const void *gmem)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_4(void *smem, const void *gmem)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// bit g * L of a lane mask: the first lane of every group of L lanes
template <int L>
constexpr uint32_t kGroupBase = L == 1 ? 0xFFFFFFFFu : L == 2 ? 0x55555555u : L == 4 ? 0x11111111u
                              : L == 8 ? 0x01010101u : L == 16 ? 0x00010001u : 0x00000001u;

// From the ballot of the lanes that have everything they wait for: bit g * L set where all L
// lanes of match group g are ready.  Executed by the whole warp.
template <int L>
__device__ __forceinline__ uint32_t complete_groups(bool ready)
{
    uint32_t c = __ballot_sync(0xFFFFFFFFu, ready);
#pragma unroll
    for (int o = 1; o < L; o <<= 1) c &= c >> o;
    return c & kGroupBase<L>;
}

// sum over the L lanes of every group; executed by the whole converged warp
template <int L>
__device__ __forceinline__ double group_sum(double x)
{
#pragma unroll
    for (int o = L / 2; o > 0; o >>= 1) x += __shfl_xor_sync(0xFFFFFFFFu, x, o);
    return x;
}
template <int L>
__device__ __forceinline__ uint32_t group_sum_u32(uint32_t x)
{
#pragma unroll
    for (int o = L / 2; o > 0; o >>= 1) x += __shfl_xor_sync(0xFFFFFFFFu, x, o);
    return x;
}


inline int stream_sleep_ns()
{
    static int ns = -1;
    if (ns < 0) {
        const char *e = getenv("TSRATE_STREAM_SLEEP_NS");
        ns = e ? atoi(e) : 20;
        if (ns < 0) ns = 0;
    }
    return ns;
}

inline int stream_max_spins()
{
    const char *e = getenv("TSRATE_STREAM_MAX_SPINS");     // read every call: tests lower it
    long v = e ? atol(e) : (1l << 24);
    return (int)(v < 1024 ? 1024 : v);
}

inline int stream_blocks_per_sm_cap()
{
    static int cap = -1;
    if (cap < 0) {
        const char *e = getenv("TSRATE_STREAM_BLOCKS_PER_SM");
        cap = e ? atoi(e) : 4;
        if (cap < 1) cap = 1;
    }
    return cap;
}


}  // namespace tsstream
