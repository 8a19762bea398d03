// fp64 TrueSkill rating kernel for sm_100a.
//
// Takes over rater.py:108-169 (gather the players' pairs, quality, rate(shared),
// trueskill_delta, rate(queue), write-back) for every match of a batch, walking
// the conflict-free rounds produced by ts_schedule.  The arithmetic is the
// two-team closed form of the trueskill package's factor graph (SURVEY.md
// Appendix A): team sums of mu and sigma^2 + tau^2 + beta^2, the truncated
// Gaussian corrections v and w, then one fused update per player.
//
// Mapping: L lanes of a warp share one match (L = 1, 2, 4, 8, 16 or 32; 32 is
// the literal "one warp per match" of the north star).  Each lane owns the slots
// lane, lane+L, ... of the match record, accumulates its partial team sums in
// registers, and the group combines them with xor-shuffles of width L.  The
// scalar v/w chain is then evaluated redundantly by the lanes of the group, so
// small L amortises it over more matches per warp (the fp64 pipe, not HBM, is
// what a wide L saturates -- see DESIGN.md).
//
// One persistent cooperative grid runs all rounds; rounds are separated by a
// grid barrier because round r+1 reads what round r wrote.  Player rows are
// read with ld.global.cg / written with st.global.cg (L2 is the coherence
// point across SMs; L1 is not coherent), as 16-byte vectors.  Match records of
// a tile are staged through shared memory with coalesced loads.
#include <stdlib.h>

#include "common.cuh"
#include "vw.cuh"

namespace {

constexpr int kBlock = 256;

struct RateArgs {
    ts_params prm;
    double *table;
    const double *seed;
    uint32_t n_players;           // ids at or beyond it are refused (TS_ST_VALUE_ERROR), never dereferenced
    int32_t row_pairs;
    int32_t shared_col;
    ModeCols modes;
    const uint32_t *ids_sorted;
    const uint32_t *meta_sorted;
    const uint32_t *order;
    const uint32_t *round_offsets;
    const ts_sched_info *info;
    ts_outputs out;
    unsigned long long *ctrl;
};

__device__ __forceinline__ double2 ld_pair(const double *p)
{
    return __ldcg(reinterpret_cast<const double2 *>(p));
}
__device__ __forceinline__ void st_pair(double *p, double mu, double sigma)
{
    __stcg(reinterpret_cast<double2 *>(p), make_double2(mu, sigma));
}

template <int L>
__device__ __forceinline__ double group_sum(double x)
{
#pragma unroll
    for (int o = L / 2; o > 0; o >>= 1) x += __shfl_xor_sync(0xFFFFFFFFu, x, o, L);
    return x;
}
template <int L>
__device__ __forceinline__ uint32_t group_sum_u32(uint32_t x)
{
#pragma unroll
    for (int o = L / 2; o > 0; o >>= 1) x += __shfl_xor_sync(0xFFFFFFFFu, x, o, L);
    return x;
}

using tsvw::trunc_vw;

// ---- asynchronous staging of match records (cp.async = LDGSTS) -------------
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

template <int P, int G>
struct RecordStage {
    uint32_t ids[2][G * P];
    uint32_t meta[2][G];
    uint32_t order[2][G];
};

// Start copying the records of tile `tile` of the round [lo, hi) into buffer b.
// Records do not depend on the player table, so this runs ahead of the grid
// barrier that orders the table accesses of consecutive rounds.
template <int P, int G>
__device__ __forceinline__ void stage_tile(RecordStage<P, G> &st, int b, const RateArgs &a, uint32_t lo,
                                           uint32_t hi, uint32_t tile, bool want_order)
{
    const uint32_t base = lo + tile * G;
    const uint32_t cnt = min((uint32_t)G, hi - base);
    const uint32_t *src = a.ids_sorted + (size_t)base * P;   // base * P * 4 bytes: 8-byte aligned (P even)
    for (uint32_t i = threadIdx.x; i < cnt * (P / 2); i += blockDim.x) cp_async_8(&st.ids[b][2 * i], src + 2 * i);
    for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) {
        cp_async_4(&st.meta[b][i], a.meta_sorted + base + i);
        if (want_order) cp_async_4(&st.order[b][i], a.order + base + i);
    }
}

template <int P, int L, bool SHARED>
__global__ void __launch_bounds__(kBlock) k_rate(const RateArgs a)
{
    constexpr int G = kBlock / L;               // matches per tile
    constexpr int SPL = (P + L - 1) / L;        // slots per lane
    constexpr int HALF = P / 2;
    __shared__ __align__(16) RecordStage<P, G> stage;

    const int tid = threadIdx.x;
    const int lane = tid % L;
    const int grp = tid / L;
    const double tau2 = a.prm.tau * a.prm.tau;
    const double beta2 = a.prm.beta * a.prm.beta;
    const double tb2 = tau2 + beta2;
    const double margin_full = a.prm.draw_margin[P];
    const uint32_t n_rounds = __ldcg(&a.info->n_rounds);
    if (__ldcg(&a.info->overflow) != 0u) {      // round_offsets was too small for this schedule: rate nothing, say so
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&a.ctrl[CTRL_VALUE_ERRORS], 1ull);
        return;
    }
    const bool want_order = a.out.quality || a.out.out_shared || a.out.out_delta || a.out.out_queue || a.out.status;
    unsigned long long generation = 0;

    // bucket 0: gated matches (rater.py:83-85, 102-106) -- outputs only; nothing
    // here touches the table, so no barrier is needed before round 1
    if (a.out.status != nullptr || a.out.quality != nullptr) {
        const uint32_t lo = __ldg(a.round_offsets), hi = __ldg(a.round_offsets + 1);
        for (uint32_t j = lo + blockIdx.x * kBlock + tid; j < hi; j += gridDim.x * kBlock) {
            uint32_t meta = __ldg(a.meta_sorted + j), m = __ldg(a.order + j);
            uint32_t mode = (meta >> TS_META_MODE_SHIFT) & 15u;
            bool skip = mode >= 6u || a.modes.col[mode] < 0;
            uint8_t st = skip ? TS_ST_SKIP_MODE : ((meta & TS_META_INVALID) ? TS_ST_INVALID : TS_ST_AFK);
            if (a.out.status) a.out.status[m] = st;
            if (!skip && a.out.quality) a.out.quality[m] = 0.0;
        }
    }
    if (n_rounds == 0) return;

    // round bounds travel one round ahead in registers: lo/hi of the current
    // round and hi of the next, so nothing on the post-barrier path waits for them
    uint32_t lo = __ldg(a.round_offsets + 1), hi = __ldg(a.round_offsets + 2);
    uint32_t hi_next = n_rounds >= 2 ? __ldg(a.round_offsets + 3) : hi;
    int pf = 0;   // staging buffer that holds / will hold the next tile of this block
    if (blockIdx.x < (hi - lo + G - 1) / G) stage_tile<P, G>(stage, pf, a, lo, hi, blockIdx.x, want_order);

    for (uint32_t r = 1; r <= n_rounds; ++r) {
        const uint32_t n_tiles = (hi - lo + G - 1) / G;
        for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            cp_async_wait_all();
            __syncthreads();               // tile's records landed; everyone is done with the other buffer
            const int cur = pf;
            pf ^= 1;
            if (tile + gridDim.x < n_tiles) stage_tile<P, G>(stage, pf, a, lo, hi, tile + gridDim.x, want_order);
            const uint32_t base = lo + tile * G;
            const uint32_t cnt = min((uint32_t)G, hi - base);
            const bool live = (uint32_t)grp < cnt;
            const uint32_t meta = live ? stage.meta[cur][grp] : 0u;
            const uint32_t mode = (meta >> TS_META_MODE_SHIFT) & 15u;
            const int qcol = live ? a.modes.col[mode & 7u] : 0;
            // winner first, stable (rater.py:144: rank = int(not winner))
            const int rank0 = (meta & TS_META_WIN_A) ? 0 : 1, rank1 = (meta & TS_META_WIN_B) ? 0 : 1;
            const int first = rank1 < rank0 ? 1 : 0;
            const bool is_draw = rank0 == rank1;

            uint32_t pid[SPL];
            double s_mu[SPL], s_sg[SPL], q_mu[SPL], q_sg[SPL], sgn[SPL];
            bool used[SPL], had[SPL];
            double d_s = 0.0, c2_s = 0.0, d_q = 0.0, c2_q = 0.0, ss_q = 0.0;
            uint32_t cnts = 0;   // n_first | n_second << 8 | bad << 16
            bool bad_id = false;

            // issue every gather of this lane before consuming any of them
            double2 raw_s[SPL], raw_q[SPL];
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const int s = lane + k * L;
                pid[k] = (live && s < P) ? stage.ids[cur][grp * P + s] : TS_EMPTY_SLOT;
                used[k] = pid[k] != TS_EMPTY_SLOT;
                if (used[k] && pid[k] >= a.n_players) { used[k] = false; bad_id = true; }   // corrupt id: no table access
                raw_s[k] = make_double2(0.0, 1.0);
                raw_q[k] = make_double2(0.0, 1.0);
                if (used[k]) {
                    const double *row = a.table + (size_t)pid[k] * a.row_pairs * 2;
                    if (SHARED) raw_s[k] = ld_pair(row + a.shared_col * 2);
                    raw_q[k] = ld_pair(row + qcol * 2);
                }
            }
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const int s = lane + k * L;
                had[k] = false;
                sgn[k] = ((s >= HALF ? 1 : 0) == first) ? 1.0 : -1.0;
                s_mu[k] = 0.0; s_sg[k] = 1.0; q_mu[k] = 0.0; q_sg[k] = 1.0;
                if (!used[k]) continue;
                bool bad = false;
                double2 sd = make_double2(0.0, 1.0);
                if (SHARED) {
                    sd = raw_s[k];
                    had[k] = !isnan(sd.x);
                    if (!had[k]) {            // rater.py:118-119: first touch -> seed
                        if (a.seed) sd = __ldg(reinterpret_cast<const double2 *>(a.seed) + pid[k]);
                        bad |= isnan(sd.x);
                    }
                }
                double2 qd = raw_q[k];
                if (isnan(qd.x)) {            // rater.py:127-130: fall back to the shared pair
                    if (SHARED) qd = sd;
                    else {
                        if (a.seed) qd = __ldg(reinterpret_cast<const double2 *>(a.seed) + pid[k]);
                        bad |= isnan(qd.x);
                    }
                }
                bad |= (qd.y == 0.0) || (SHARED && sd.y == 0.0);   // Rating(sigma=0): ValueError
                s_mu[k] = sd.x; s_sg[k] = sd.y; q_mu[k] = qd.x; q_sg[k] = qd.y;
                if (SHARED) {
                    d_s = fma(sgn[k], sd.x, d_s);
                    c2_s += fma(sd.y, sd.y, tb2);
                }
                d_q = fma(sgn[k], qd.x, d_q);
                c2_q += fma(qd.y, qd.y, tb2);
                ss_q = fma(qd.y, qd.y, ss_q);
                cnts += (sgn[k] > 0.0 ? 1u : 256u) + (bad ? 65536u : 0u);
            }
            cnts += bad_id ? 65536u : 0u;
            if (L > 1) {
                if (SHARED) { d_s = group_sum<L>(d_s); c2_s = group_sum<L>(c2_s); }
                d_q = group_sum<L>(d_q);
                c2_q = group_sum<L>(c2_q);
                ss_q = group_sum<L>(ss_q);
                cnts = group_sum_u32<L>(cnts);
            }
            const int n0 = cnts & 255u, n1 = (cnts >> 8) & 255u, n = n0 + n1;
            int status = TS_ST_OK;
            if ((cnts >> 16) != 0u || n0 == 0 || n1 == 0) status = TS_ST_VALUE_ERROR;
            const double margin = n == P ? margin_full : a.prm.draw_margin[n];   // ragged teams are rare

            double v_s = 0.0, w_s = 0.0, rc_s = 0.0, v_q, w_q;
            if (SHARED) {
                rc_s = rsqrt(c2_s);
                if (!trunc_vw(d_s * rc_s, margin * rc_s, is_draw, a.prm.draw_is_zero, v_s, w_s))
                    status = status ? status : TS_ST_FP_ERROR;
            }
            const double rc_q = rsqrt(c2_q);
            if (!trunc_vw(d_q * rc_q, margin * rc_q, is_draw, a.prm.draw_is_zero, v_q, w_q))
                status = status ? status : TS_ST_FP_ERROR;

            const uint32_t m = (live && want_order) ? stage.order[cur][grp] : 0u;
            if (live && lane == 0) {
                if (a.out.status) a.out.status[m] = (uint8_t)status;
                if (status != TS_ST_OK) atomicAdd(&a.ctrl[status == TS_ST_VALUE_ERROR ? CTRL_VALUE_ERRORS : CTRL_ERRORS], 1ull);
                else if (a.out.quality) {
                    // rater.py:141 on the queue-specific pre-update ratings (no tau)
                    const double nb2 = (double)n * beta2;
                    const double den = nb2 + ss_q;
                    a.out.quality[m] = exp(-0.5 * d_q * d_q / den) * sqrt(nb2 / den);
                }
            }
            if (status != TS_ST_OK) continue;   // the reference raises: nothing is written
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                if (!used[k]) continue;
                const int s = lane + k * L;
                double *row = a.table + (size_t)pid[k] * a.row_pairs * 2;
                const size_t o = (size_t)m * P + s;
                if (SHARED) {
                    const double si = fma(s_sg[k], s_sg[k], tau2);
                    const double g = si * rc_s;
                    const double mu2 = fma(sgn[k] * g, v_s, s_mu[k]);
                    const double sg2 = sqrt(si * (1.0 - (g * rc_s) * w_s));
                    st_pair(row + a.shared_col * 2, mu2, sg2);
                    if (a.out.out_shared) st_pair(a.out.out_shared + o * 2, mu2, sg2);
                    if (a.out.out_delta)   // rater.py:150-153
                        a.out.out_delta[o] = had[k] ? (mu2 - sg2) - (s_mu[k] - s_sg[k]) : 0.0;
                }
                const double si = fma(q_sg[k], q_sg[k], tau2);
                const double g = si * rc_q;
                const double mu2 = fma(sgn[k] * g, v_q, q_mu[k]);
                const double sg2 = sqrt(si * (1.0 - (g * rc_q) * w_q));
                st_pair(row + qcol * 2, mu2, sg2);
                if (a.out.out_queue) st_pair(a.out.out_queue + o * 2, mu2, sg2);
            }
        }
        if (r == n_rounds) break;
        // run ahead of the barrier: bounds and first tile of the next round
        lo = hi;
        hi = hi_next;
        if (r + 2 <= n_rounds) hi_next = __ldg(a.round_offsets + r + 3);
        if (blockIdx.x < (hi - lo + G - 1) / G) stage_tile<P, G>(stage, pf, a, lo, hi, blockIdx.x, want_order);
        grid_barrier(a.ctrl, generation);
    }
}

// Fewer resident blocks make the per-round grid barrier cheaper (fewer arrivals
// on one L2 word); more blocks give more gathers in flight for fat rounds.
// TSRATE_BLOCKS_PER_SM overrides the default for experiments.
int rate_blocks_per_sm_cap()
{
    static int cap = -1;
    if (cap < 0) {
        const char *e = getenv("TSRATE_BLOCKS_PER_SM");
        cap = e ? atoi(e) : 1;   // measured on config 2 at L = 2: 1 block/SM 3.67 ms, 2 blocks/SM 4.26 ms
        if (cap < 1) cap = 1;
    }
    return cap;
}

template <int P, int L, bool SHARED>
int launch_rate(ts_ctx *ctx, const RateArgs &args, cudaStream_t stream)
{
    int per_sm = 0;
    TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_rate<P, L, SHARED>, kBlock, 0));
    if (per_sm < 1) { ts_set_error("rate kernel does not fit on an SM"); return TS_E_CUDA; }
    per_sm = min(per_sm, rate_blocks_per_sm_cap());
    dim3 grid(ctx->sm_count * per_sm), block(kBlock);
    void *kargs[] = {(void *)&args};
    if (int rc = ts_reset_barrier(ctx->d_ctrl, stream)) return rc;
    TS_CUDA(cudaLaunchCooperativeKernel((void *)k_rate<P, L, SHARED>, grid, block, kargs, 0, stream));
    return 0;
}

template <int P, bool SHARED>
int dispatch_l(ts_ctx *ctx, const RateArgs &a, int lanes, cudaStream_t s)
{
    switch (lanes) {
    case 1: return launch_rate<P, 1, SHARED>(ctx, a, s);
    case 2: return launch_rate<P, 2, SHARED>(ctx, a, s);
    case 4: return launch_rate<P, 4, SHARED>(ctx, a, s);
    case 8: return launch_rate<P, 8, SHARED>(ctx, a, s);
    case 16: return launch_rate<P, 16, SHARED>(ctx, a, s);
    case 32: return launch_rate<P, 32, SHARED>(ctx, a, s);
    }
    ts_set_error("lanes_per_match must be 0,1,2,4,8,16 or 32 (got %d)", lanes);
    return TS_E_ARG;
}

}  // namespace

extern "C" int ts_rate(ts_ctx *ctx, const ts_table *table, const uint32_t *ids_sorted,
                       const uint32_t *meta_sorted, const uint32_t *order,
                       const uint32_t *round_offsets, const ts_sched_info *info, int64_t n_matches,
                       int32_t slots, const ts_outputs *outputs, int32_t lanes_per_match, void *stream)
{
    if (!ctx || !table || !table->data || !ids_sorted || !meta_sorted || !order || !round_offsets || !info) {
        ts_set_error("ts_rate: null argument");
        return TS_E_ARG;
    }
    if (n_matches < 0 || n_matches >= (int64_t)1 << 32) {
        ts_set_error("ts_rate: n_matches %lld outside [0, 2^32): rate the ladder in windows", (long long)n_matches);
        return TS_E_ARG;
    }
    if (slots != 6 && slots != 10) {
        ts_set_error("ts_rate: slots must be 6 (3v3) or 10 (5v5); pad smaller teams with TS_EMPTY_SLOT (got %d)", slots);
        return TS_E_UNSUPPORTED;
    }
    if (table->row_pairs < 1 || table->shared_col >= table->row_pairs) {
        ts_set_error("ts_rate: bad table layout");
        return TS_E_ARG;
    }
    for (int i = 0; i < 6; ++i)
        if (table->mode_col[i] >= table->row_pairs) {
            ts_set_error("ts_rate: mode_col[%d] outside the row", i);
            return TS_E_ARG;
        }
    if (n_matches == 0) return 0;
    TS_CUDA(cudaSetDevice(ctx->device));
    RateArgs a;
    a.prm = ctx->params;
    a.table = table->data;
    a.seed = table->seed;
    a.n_players = (uint32_t)(table->n_players > 0xFFFFFFFFll ? 0xFFFFFFFFll : table->n_players);
    a.row_pairs = table->row_pairs;
    a.shared_col = table->shared_col;
    for (int i = 0; i < 16; ++i) a.modes.col[i] = i < 6 ? table->mode_col[i] : (int8_t)-1;
    a.ids_sorted = ids_sorted;
    a.meta_sorted = meta_sorted;
    a.order = order;
    a.round_offsets = round_offsets;
    a.info = info;
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
    cudaStream_t s = (cudaStream_t)stream;
    const bool shared = table->shared_col >= 0;
    int lanes = lanes_per_match;
    if (lanes == 0) lanes = slots == 6 ? 2 : 8;   // 3v3: one lane per team; 5v5: two lanes hold three slots at most
    if (slots == 6) return shared ? dispatch_l<6, true>(ctx, a, lanes, s) : dispatch_l<6, false>(ctx, a, lanes, s);
    return shared ? dispatch_l<10, true>(ctx, a, lanes, s) : dispatch_l<10, false>(ctx, a, lanes, s);
}
