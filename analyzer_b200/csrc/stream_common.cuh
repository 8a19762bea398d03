// Pieces shared by the streaming rating kernels (stream.cu: one pair per player;
// stream_pair.cu: shared + queue pair): self-validating row words, cp.async staging,
// group shuffles, tuning knobs.
#pragma once
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "vw.cuh"

namespace tsstream {

// A row is 32 bytes and 32-byte aligned: one 256-bit access (LDG.256 / STG.256, new with sm_100)
// instead of two 128-bit ones -- half the L2 requests per poll, and half the NVLink packets when
// the row goes to another GPU.  The elements of a vector access are performed as separate 8-byte
// accesses in no particular order, which is all the protocol needs: every word validates itself.
// SYS = the row may live in, or be written from, another GPU: system scope.
template <bool SYS = false>
__device__ __forceinline__ void ld_row(const unsigned long long *row, unsigned long long (&w)[4])
{
    if (SYS)
        asm volatile("ld.relaxed.sys.global.v4.u64 {%0, %1, %2, %3}, [%4];"
                     : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(row) : "memory");
    else
        asm volatile("ld.relaxed.gpu.global.v4.u64 {%0, %1, %2, %3}, [%4];"
                     : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(row) : "memory");
}
template <bool SYS = false>
__device__ __forceinline__ void st_row(unsigned long long *row, double mu, double sigma, uint32_t ver)
{
    const unsigned long long v = (unsigned long long)ver << 32;
    const unsigned long long m = (unsigned long long)__double_as_longlong(mu);
    const unsigned long long s = (unsigned long long)__double_as_longlong(sigma);
    unsigned long long w0 = (m & 0xFFFFFFFFull) | v, w1 = (m >> 32) | v;
    unsigned long long w2 = (s & 0xFFFFFFFFull) | v, w3 = (s >> 32) | v;
    if (SYS)
        asm volatile("st.relaxed.sys.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(row), "l"(w0), "l"(w1), "l"(w2), "l"(w3) : "memory");
    else
        asm volatile("st.relaxed.gpu.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(row), "l"(w0), "l"(w1), "l"(w2), "l"(w3) : "memory");
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

// Match records (ids, links, meta) and the per-match outputs are touched once: they travel with
// an L2 evict-first policy so that, on ladders whose rows do not fit L2, the 50-odd bytes a match
// streams through do not push the rows out that its successors will poll.
__device__ __forceinline__ unsigned long long stream_policy()
{
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void cp_async_8(void *smem, const void *gmem, unsigned long long pol)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;" ::"r"(s), "l"(gmem), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_4(void *smem, const void *gmem, unsigned long long pol)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2;" ::"r"(s), "l"(gmem), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_pair_stream(double *p, double mu, double sigma, unsigned long long pol)
{
    asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p), "d"(mu), "d"(sigma), "l"(pol) : "memory");
}
__device__ __forceinline__ void prefetch_row_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
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


// per-warp staging buffers: a warp owns W = 32 / L consecutive matches at a time; `link` holds
// pred (single GPU: the version to wait for) or succ (sharded ladder: where the new row goes).
// Three buffers: the tile being rated, the next one (its records are there one tile early, so
// that its players' rows can be prefetched into L2 while this tile waits and computes), and the
// one after that, in flight.
// (Two buffers, and no prefetch, where three would not fit the 48 KB of static shared memory:
// one lane per 5v5 match.)
template <int P, int W>
struct alignas(16) WarpStage {
    static constexpr int kDepth = ((2 * W * P + W) * 4 * 3 * 8 <= 48 * 1024) ? 3 : 2;
    uint32_t ids[kDepth][W * P];
    uint32_t link[kDepth][W * P];
    uint32_t meta[kDepth][W];
};

template <int P, int W>
__device__ __forceinline__ void stage_warp_tile(WarpStage<P, W> &st, int b, const uint32_t *ids, const uint32_t *link,
                                                const uint32_t *meta, uint32_t base, uint32_t cnt, int wl,
                                                unsigned long long pol)
{
    const uint32_t *src_i = ids + (size_t)base * P, *src_p = link + (size_t)base * P;
    for (uint32_t i = wl; i < cnt * (P / 2); i += 32) {
        cp_async_8(&st.ids[b][2 * i], src_i + 2 * i, pol);
        cp_async_8(&st.link[b][2 * i], src_p + 2 * i, pol);
    }
    for (uint32_t i = wl; i < cnt; i += 32) cp_async_4(&st.meta[b][i], meta + base + i, pol);
}

// rows of the players of a staged tile -> L2 (row_words 64-bit words per player)
template <int P, int W>
__device__ __forceinline__ void prefetch_tile_rows(const WarpStage<P, W> &st, int b, const unsigned long long *rows,
                                                   uint32_t row_words, uint32_t cnt, uint32_t n_players, int wl)
{
    for (uint32_t i = wl; i < cnt * P; i += 32) {
        const uint32_t pid = st.ids[b][i];
        if (pid < n_players) prefetch_row_l2(rows + (size_t)pid * row_words);
    }
}

// One poll of the rows a lane still misses (rows of 4 words per player at `rows`): a row is taken
// when all four of its words carry the version the lane waits for.  HOIST: the loads of all
// missing rows are issued before any is checked -- as many DRAM / L2 round trips in flight as
// the lane has slots; otherwise row by row (measured faster when every row sits in L2).
template <int SPL, bool HOIST, bool SYS>
__device__ __forceinline__ bool poll_rows(const unsigned long long *rows, const uint32_t (&pid)[SPL],
                                          const uint32_t (&want)[SPL], bool (&have)[SPL], double (&q_mu)[SPL],
                                          double (&q_sg)[SPL])
{
    bool ready = true;
    if constexpr (HOIST) {
        unsigned long long w[SPL][4];
#pragma unroll
        for (int k = 0; k < SPL; ++k)
            if (!have[k]) ld_row<SYS>(rows + (size_t)pid[k] * 4, w[k]);
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            if (!have[k] && row_has_version(w[k], want[k])) {
                have[k] = true;
                q_mu[k] = row_mu(w[k]);
                q_sg[k] = row_sigma(w[k]);
            }
            ready &= have[k];
        }
    } else {
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            if (have[k]) continue;
            unsigned long long w[4];
            ld_row<SYS>(rows + (size_t)pid[k] * 4, w);
            if (row_has_version(w, want[k])) {
                have[k] = true;
                q_mu[k] = row_mu(w);
                q_sg[k] = row_sigma(w);
            }
            ready &= have[k];
        }
    }
    return ready;
}

// What a pass needs of the environment (rater.py:30-37), in registers.
struct PassEnv {
    double tau2, beta2, tb2, margin_full;
    const double *margins;       // draw margin by number of players (ragged teams)
    const double *seed;          // get_trueskill_seed per player, or nullptr
    int draw_is_zero;            // 0: draw margin in use; 1: draw_probability = 0, analytic limit; 2: such a draw fails
};

// One rating pass of the ladder layout (one (mu, sigma) pair per player; rater.py:132-169 with
// a single rate()), shared by the single-GPU stream kernel and the sharded one so that both give
// the same bits: every product-sum is written with an explicit rounding (no contraction is left
// to the compiler, which decides it per template instantiation).  Executed by the whole converged
// warp; lanes of groups that are not complete compute on placeholders.  L lanes share a match;
// lane `lane` of the group owns slots lane, lane + L, ...  q_* are the rows as read; mu2 / sg2
// the rows to publish (unchanged where the reference would raise).  Returns TS_ST_*.
template <int P, int L, int SPL>
__device__ __forceinline__ int rate_pass(const PassEnv &env, int lane, int first, bool draw_eff,
                                         const uint32_t (&pid)[SPL], const bool (&used)[SPL], bool bad_id,
                                         const double (&q_mu)[SPL], const double (&q_sg)[SPL],
                                         double (&mu2)[SPL], double (&sg2)[SPL], double &quality_d, double &quality_ss,
                                         int &n_out)
{
    constexpr int HALF = P / 2;
    double sgn[SPL], e_mu[SPL], e_sg[SPL];     // effective pairs; q_mu / q_sg stay the rows as read
    double d_q = 0.0, c2_q = 0.0, ss_q = 0.0;
    uint32_t cnts = bad_id ? 65536u : 0u;
    bool any_null = false;
#pragma unroll
    for (int k = 0; k < SPL; ++k) {
        e_mu[k] = q_mu[k]; e_sg[k] = q_sg[k];
        any_null |= used[k] && isnan(q_mu[k]);
    }
    if (__any_sync(0xFFFFFFFFu, any_null)) {       // first touch: seed (rater.py:118-119)
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            if (!(used[k] && isnan(q_mu[k]))) continue;
            double2 sd = make_double2(q_mu[k], q_sg[k]);
            if (env.seed) sd = __ldg(reinterpret_cast<const double2 *>(env.seed) + pid[k]);
            if (isnan(sd.x)) cnts |= 65536u;       // neither rating nor seed
            else { e_mu[k] = sd.x; e_sg[k] = sd.y; }
        }
    }
#pragma unroll
    for (int k = 0; k < SPL; ++k) {
        const int s = lane + k * L;
        sgn[k] = ((s >= HALF ? 1 : 0) == first) ? 1.0 : -1.0;
        const bool u = used[k];
        d_q = __fma_rn(sgn[k], u ? e_mu[k] : 0.0, d_q);
        c2_q = __dadd_rn(c2_q, u ? __fma_rn(e_sg[k], e_sg[k], env.tb2) : 0.0);
        ss_q = u ? __fma_rn(e_sg[k], e_sg[k], ss_q) : ss_q;
        cnts += (u ? (sgn[k] > 0.0 ? 1u : 256u) : 0u) + ((u && e_sg[k] == 0.0) ? 65536u : 0u);
    }
    d_q = group_sum<L>(d_q);
    c2_q = group_sum<L>(c2_q);
    ss_q = group_sum<L>(ss_q);
    cnts = group_sum_u32<L>(cnts);
    const int n0 = cnts & 255u, n1 = (cnts >> 8) & 255u, n = n0 + n1;
    int status = TS_ST_OK;
    if ((cnts >> 16) != 0u || n0 == 0 || n1 == 0) status = TS_ST_VALUE_ERROR;
    const double margin = n == P ? env.margin_full : env.margins[n];   // ragged teams are rare: no indexed load on the hot path
    double v_q = 0.0, w_q = 0.0;
    const double rc_q = rsqrt(c2_q);
    if (status == TS_ST_OK &&
        !tsvw::trunc_vw(__dmul_rn(d_q, rc_q), __dmul_rn(margin, rc_q), draw_eff, env.draw_is_zero, v_q, w_q))
        status = TS_ST_FP_ERROR;
    const bool ok = status == TS_ST_OK;
#pragma unroll
    for (int k = 0; k < SPL; ++k) {
        // a failed match writes nothing in the reference; here it republishes the rows it
        // read so that successors can go on
        const double si = __fma_rn(e_sg[k], e_sg[k], env.tau2);
        const double g = __dmul_rn(si, rc_q);
        mu2[k] = ok ? __fma_rn(__dmul_rn(sgn[k], g), v_q, e_mu[k]) : q_mu[k];
        sg2[k] = ok ? __dsqrt_rn(__dmul_rn(si, __fma_rn(-__dmul_rn(g, rc_q), w_q, 1.0))) : q_sg[k];
    }
    quality_d = d_q;
    quality_ss = ss_q;
    n_out = n;
    return status;
}

// match.trueskill_quality (rater.py:141) of a rated match, from the pass's team sums
__device__ __forceinline__ double match_quality(double d_q, double ss_q, int n, double beta2)
{
    const double nb2 = __dmul_rn((double)n, beta2);
    const double den = __dadd_rn(nb2, ss_q);
    return __dmul_rn(exp(__ddiv_rn(__dmul_rn(__dmul_rn(-0.5, d_q), d_q), den)), __dsqrt_rn(__ddiv_rn(nb2, den)));
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

// TSRATE_STREAM_PREFETCH=1: the rows of the next tile's players are prefetched into L2 one tile early.
// Off by default -- measured (scripts/ab_stream.py): where the rows do not fit L2 the kernel is bound by
// DRAM accesses and a prefetched row is often evicted again before it is polled (125M 3v3 matches over
// 12.5M players: rate 29.7 -> 36.9 ms); it gains 2 % on the deep Zipf chains of config 3 and nothing
// when the rows sit in L2.
inline int stream_prefetch_mode()
{
    const char *e = getenv("TSRATE_STREAM_PREFETCH");
    return e ? atoi(e) : 0;
}
// TSRATE_STREAM_HOIST=0: a poll checks row by row instead of issuing the loads of all its missing rows
// first.  With one 256-bit load per row the hoisted form is the faster one everywhere measured
// (config 2: 2.20 -> 2.13 ms; 125M / 12.5M: 29.7 -> 29.2 ms).
inline int stream_hoist_mode()
{
    const char *e = getenv("TSRATE_STREAM_HOIST");
    return e ? atoi(e) : 1;
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
